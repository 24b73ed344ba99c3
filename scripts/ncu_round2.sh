# Round-2 ncu evidence for profiles/: the launch list of the default bench command (headline part),
# then one --set full capture per dominant kernel and shape.  One GPU.  bash scripts/ncu_round2.sh r2
set -x
tag=${1:-r2}
cd "$(dirname "$0")/.."
B="python bench.py --steps 10 --warmup 3 --blocks 3 --no-e2e --no-cpu-baseline --configs C3,C4"
$B > gpurun_out/${tag}_bench_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^k_|hipe' -c 600 --csv --log-file gpurun_out/${tag}_launches.csv \
    $B > gpurun_out/${tag}_ncu_launches.log 2>&1
cap() {  # name, kernel regex, prof_one args...
  name=$1; kre=$2; shift 2
  python scripts/prof_one.py "$@" > /dev/null 2>&1 || return
  ncu --set full --clock-control none --import-source on -k regex:$kre -s 2 -c 1 -f -o gpurun_out/${tag}_$name \
      python scripts/prof_one.py "$@" > gpurun_out/${tag}_$name.log 2>&1
}
cap cluster_f32_64_B64 k_cluster --op cluster --dtype f32 --B 64
cap cluster_bf16_64_B512 k_cluster --op cluster --dtype bf16 --B 512
cap cluster_f32_56_B32 k_cluster --op cluster --dtype f32 --dim 56 --B 32
cap cluster_f32_96_B32 k_cluster --op cluster --dtype f32 --dim 96 --B 32
# the resident form of the headline shape (the default before the streamed form), for comparison
export HIPE_CLUSTER_STREAM=0
cap clusterres_f32_64_B64 k_cluster --op cluster --dtype f32 --B 64
unset HIPE_CLUSTER_STREAM
cap fused_f32_96_B32 k_stream --op fused --dtype f32 --dim 96 --B 32
cap fwd_f32_64_B32 k_stream --op fwd --dtype f32 --B 32
ls -la gpurun_out/${tag}_*.ncu-rep
