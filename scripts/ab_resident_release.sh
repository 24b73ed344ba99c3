# A/B of when the resident launch's consumers release a ring slot (HIPE_RES_RELEASE 0/1/2); tuning aid.
cd "$(dirname "$0")/.."
for m in 0 1 2; do
  (cd hand-integral-pose-estimation_b200 && HIPE_NVCC_EXTRA="-DHIPE_RES_RELEASE=$m" python -m hipe_b200._build --force) > /dev/null 2>&1
  HIPE_NO_BUILD=1 python scripts/sweep.py --dtype f32 --B 64 --ops resident 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('release mode $m', d['ms'], d['min_ms'], d['GBs_2V'])"
done
