# A/B of how the resident launch's partials cross CTAs (HIPE_RES_SELFVALID 0/1, poll sleep); tuning aid.
# Same box, same call: only numbers printed by one run of this script are comparable.
cd "$(dirname "$0")/.."
for cfg in "-DHIPE_RES_SELFVALID=0" "-DHIPE_RES_SELFVALID=1" "-DHIPE_RES_SELFVALID=1 -DHIPE_RES_POLL_NS=32" "-DHIPE_RES_SELFVALID=0" "-DHIPE_RES_SELFVALID=1"; do
  (cd hand-integral-pose-estimation_b200 && HIPE_NVCC_EXTRA="$cfg" python -m hipe_b200._build --force) > /dev/null 2>&1
  HIPE_NO_BUILD=1 python scripts/sweep.py --dtype f32 --B 64 --ops resident 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$cfg', d['ms'], d['min_ms'], d['GBs_2V'])"
done
(cd hand-integral-pose-estimation_b200 && python -m hipe_b200._build --force) > /dev/null 2>&1
