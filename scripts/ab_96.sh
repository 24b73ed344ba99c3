# 96^3: the L2-window launch (hipe_joint_loss_resident) against the stream-fused one, same box
set -x
python scripts/sweep.py --B 32 --dim 96 --ops fused --variants 1 --clusters 8
for lag in 2 3 4 6; do HIPE_RESIDENT_LAG=$lag python scripts/sweep.py --B 32 --dim 96 --ops resident; done
HIPE_RESIDENT_ITEM_STAGES=6 python scripts/sweep.py --B 32 --dim 96 --ops resident
HIPE_RESIDENT_ITEM_STAGES=4 python scripts/sweep.py --B 32 --dim 96 --ops resident
python scripts/sweep.py --B 32 --dim 96 --dtype bf16 --ops resident
python scripts/sweep.py --B 32 --dim 96 --dtype bf16 --ops fused --variants 1 --clusters 2
