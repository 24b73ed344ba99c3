# publish on the side stream vs in stream order, alternating, N ranks (N = $1, default 2)
N=${1:-2}
for m in side main side main; do
HIPE_PUBLISH_STREAM=$m timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-e2e --no-cpu-baseline --configs C3 > gpurun_out/r2_bench_n${N}_$m.json 2> gpurun_out/r2_bench_n${N}_$m.err
python - <<PY
import json
b=json.loads(open('gpurun_out/r2_bench_n${N}_$m.json').read().strip().splitlines()[-1])
print('$m', 'N=$N', round(b['ms_per_step'],4), round(b['ms_per_step_min'],4), round(b['ms_per_step_max'],4), round(b['value']), round(b['host_enqueue_ms_per_step'],3), {k:round(v['ms'],4) for k,v in b['kernels'].items()}, b['sharded_parity']['max'], 'C3', round(b['configs']['C3']['ms_per_step'],4))
PY
done
