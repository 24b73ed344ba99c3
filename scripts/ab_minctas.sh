# A/B of the register cap of the TMA kernels (HIPE_MIN_CTAS) on one B200; tuning aid.
set -x
cd "$(dirname "$0")/.."
python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/ab_pytest.log 2>&1; tail -2 gpurun_out/ab_pytest.log
for m in 3 2; do
  (cd hand-integral-pose-estimation_b200 && HIPE_NVCC_EXTRA="-DHIPE_MIN_CTAS=$m" python -m hipe_b200._build --force) > /dev/null 2>&1
  for cfg in "f32 64 64" "bf16 128 64" "f32 32 96" "f32 96 56" "bf16 32 96"; do
    set -- $cfg
    HIPE_NO_BUILD=1 python scripts/sweep.py --dtype $1 --B $2 --dim $3 --clusters 0 --variants 1 --ops fwd,bwd,fused 2>&1 | tail -1 | sed "s/^/minctas=$m /" >> gpurun_out/ab_minctas.log
  done
done
HIPE_NO_BUILD=1 python scripts/sweep.py --dtype f32 --B 96 --dim 56 --ops resident >> gpurun_out/ab_minctas.log 2>&1
cat gpurun_out/ab_minctas.log
