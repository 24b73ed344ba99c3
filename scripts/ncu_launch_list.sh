set -x
cd /root/repo
python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1b_bench_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^k_|hipe' -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r1b_ncu_launches.log 2>&1
tail -3 gpurun_out/launches.csv | cut -c1-300
python -m pytest tests -m gpu -q 2>&1 | tail -3
python __graft_entry__.py smoke 2>&1 | tail -1
