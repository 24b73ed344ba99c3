set -x
python bench.py --dtype bf16 --batch 512 --steps 30 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/bench_bf16_b512.log 2>&1
python bench.py --fused --steps 50 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/bench_fused.log 2>&1
python bench.py --dim 96 --batch 32 --steps 30 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/bench_96.log 2>&1
python scripts/sweep.py --dtype bf16 --B 128 --clusters 0 --variants 1 --ops resident,fwd,bwd,fused > gpurun_out/sweep_bf16.log 2>&1
python scripts/sweep.py --dtype f32 --B 64 --clusters 0 --variants 1 --ops resident,fwd,bwd,fused > gpurun_out/sweep_f32.log 2>&1
python scripts/sweep.py --dtype f32 --B 32 --dim 96 --clusters 0 --variants 1 --ops resident,fwd,bwd,fused > gpurun_out/sweep_f32_96.log 2>&1
tail -n 3 gpurun_out/bench_bf16_b512.log gpurun_out/bench_fused.log gpurun_out/bench_96.log gpurun_out/sweep_*.log
