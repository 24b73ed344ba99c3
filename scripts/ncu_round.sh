# ncu evidence for profiles/: launch list of the default bench command, then one --set full
# capture per kernel of interest.  One GPU.  Usage: bash scripts/ncu_round.sh r1b
set -x
tag=${1:-r1b}
cd "$(dirname "$0")/.."
python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_bench_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^k_|hipe' -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1
cap() {  # name, kernel regex, prof_one args...
  name=$1; kre=$2; shift 2
  python scripts/prof_one.py "$@" > /dev/null 2>&1 || return
  ncu --set full --clock-control none --import-source on -k regex:$kre -s 2 -c 1 -f -o gpurun_out/${tag}_$name \
      python scripts/prof_one.py "$@" > gpurun_out/${tag}_$name.log 2>&1
}
cap resident_f32 k_resident --op resident --dtype f32
cap fwd_bf16 k_stream --op fwd --dtype bf16 --B 128
cap fwd_f32 k_stream --op fwd --dtype f32
cap bwd_f32 k_stream --op bwd --dtype f32
ls -la gpurun_out/*.ncu-rep
