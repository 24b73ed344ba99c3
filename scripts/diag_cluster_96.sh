# DRAM traffic of the streamed cluster launch against its lead (what the L2 must hold): four-metric ncu
# passes; then the wait profile of the instrumented build
cd "$(dirname "$0")/.."
M="--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_op_read_hit_rate.pct --clock-control none -k regex:k_cluster -s 2 -c 1 --csv"
run() {  # label, env..., -- prof_one args
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" ncu $M python scripts/prof_one.py --op cluster "$@" 2>/dev/null | grep -E '"(gpu__time_duration.sum|dram__bytes_read.sum|dram__bytes_write.sum|lts__t_sector_op_read_hit_rate.pct)"' | awk -F'","' -v l="$label" '{gsub(/"/,"",$NF); printf "%s %s %s %s\n", l, $(NF-2), $(NF-1), $NF}'
}
for L in 25 31 36 44; do run "f32_96 C=16 LEAD=$L" HIPE_CLUSTER_LEAD=$L -- --dtype f32 --dim 96 --B 32; done
run "f32_64 STREAM C=8" HIPE_CLUSTER_STREAM=1 HIPE_CLUSTER_C=8 -- --dtype f32 --dim 64 --B 64
HIPE_NVCC_EXTRA=-DHIPE_CLU_PROF python hand-integral-pose-estimation_b200/hipe_b200/_build.py --force > /dev/null || exit 1
python scripts/diag_cluster_waits.py --dtype f32 --B 32 --dim 96
HIPE_CLUSTER_STREAM=1 HIPE_CLUSTER_C=8 python scripts/diag_cluster_waits.py --dtype f32 --B 64 --dim 64
