# Wait profile of the cluster-resident launch: instrumented build, runs, normal build restored.
set -x
cd "$(dirname "$0")/.."
HIPE_NVCC_EXTRA=-DHIPE_CLU_PROF python hand-integral-pose-estimation_b200/hipe_b200/_build.py --force > /dev/null || exit 1
python scripts/diag_cluster_waits.py --dtype f32 --B 64
HIPE_CLUSTER_C=8 python scripts/diag_cluster_waits.py --dtype f32 --B 64
python scripts/diag_cluster_waits.py --dtype bf16 --B 128
HIPE_CLUSTER_C=3 python scripts/diag_cluster_waits.py --dtype bf16 --B 128
python hand-integral-pose-estimation_b200/hipe_b200/_build.py --force > /dev/null
