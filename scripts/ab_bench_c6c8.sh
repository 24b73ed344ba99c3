# the driver's bench line (headline part) with the streamed plan's cluster size: library default vs C=8 / lead 21
cd "$(dirname "$0")/.."
for v in new old new old new old; do
  if [ $v = old ]; then export HIPE_CLUSTER_STREAM=1 HIPE_CLUSTER_C=8 HIPE_CLUSTER_LEAD=21 HIPE_CLUSTER_RSLOTS=8; else unset HIPE_CLUSTER_STREAM HIPE_CLUSTER_C HIPE_CLUSTER_LEAD HIPE_CLUSTER_RSLOTS; fi
  python bench.py --no-configs --no-e2e --no-cpu-baseline --no-alt 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', round(b['ms_per_step'],4), round(b['ms_per_step_min'],4), round(b['ms_per_step_max'],4), round(b['kernels']['joint_loss_cluster']['ms'],4), round(b['roofline']['frac'],3))"
done
