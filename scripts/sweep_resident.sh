# Sweep of the resident launch's knobs (lag, item size) on one B200; tuning aid.
cd "$(dirname "$0")/.."
for st in 4 8 16; do for lag in 2 3 4 6 8 12; do
  HIPE_RESIDENT_ITEM_STAGES=$st HIPE_RESIDENT_LAG=$lag python scripts/sweep.py --dtype f32 --B 64 --ops resident 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('stages $st lag $lag', d['ms'], d['min_ms'], d['GBs_2V'])"
done; done
