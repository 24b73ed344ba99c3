# A/B of the resident launch's ring split (reduce slots, write slots; 13 x 16 KB in total); tuning aid.
cd "$(dirname "$0")/.."
for rw in "7 6" "8 5" "9 4" "6 7" "5 8" "10 3"; do
  set -- $rw
  (cd hand-integral-pose-estimation_b200 && HIPE_NVCC_EXTRA="-DHIPE_RES_RSLOTS=$1 -DHIPE_RES_WSLOTS=$2" python -m hipe_b200._build --force) > /dev/null 2>&1
  for lag in 3 4; do
  HIPE_NO_BUILD=1 HIPE_RESIDENT_LAG=$lag python scripts/sweep.py --dtype f32 --B 64 --ops resident 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('rslots $1 wslots $2 lag $lag', d['ms'], d['min_ms'], d['GBs_2V'])"
  done
done
