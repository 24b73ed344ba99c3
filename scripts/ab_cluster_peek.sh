# early barrier peek on/off, same box (two builds)
cd "$(dirname "$0")/.."
for v in 0 1 0 1; do
  HIPE_NVCC_EXTRA=-DHIPE_CLU_PEEK=$v python hand-integral-pose-estimation_b200/hipe_b200/_build.py --force > /dev/null || exit 1
  echo "== HIPE_CLU_PEEK=$v"
  python scripts/ab_cluster.py peek
done
python hand-integral-pose-estimation_b200/hipe_b200/_build.py --force > /dev/null
