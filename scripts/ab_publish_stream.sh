timeout 600 python -m pytest tests/test_gpu_sharded.py -x -q -m gpu > gpurun_out/r2_pytest_sharded.log 2>&1; tail -3 gpurun_out/r2_pytest_sharded.log
for m in side main side main; do
HIPE_PUBLISH_STREAM=$m timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-e2e --no-cpu-baseline --configs C3 > gpurun_out/r2_bench_n2_$m.json 2> gpurun_out/r2_bench_n2_$m.err
python - <<PY
import json
b=json.loads(open('gpurun_out/r2_bench_n2_$m.json').read().strip().splitlines()[-1])
print('$m', b['ms_per_step'], b['ms_per_step_min'], b['value'], b['host_enqueue_ms_per_step'], {k:round(v['ms'],4) for k,v in b['kernels'].items()}, b['sharded_parity'], b['configs']['C3']['ms_per_step'])
PY
done
