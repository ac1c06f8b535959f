set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_v2.json 2> gpurun_out/bench_v2.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_v2.json 2> gpurun_out/bench_l1_16_v2.err
mkdir -p /tmp/prof
ncu --set full --clock-control none --import-source on -k regex:msm_accumulate -s 15 -c 5 -o /tmp/prof/acc $B > gpurun_out/ncu_acc.log 2>&1
ncu -i /tmp/prof/acc.ncu-rep --page raw --csv > gpurun_out/acc_raw.csv
ncu --set full --clock-control none --import-source on -k regex:ntt_pass -s 54 -c 6 -o /tmp/prof/ntt $B > gpurun_out/ncu_ntt.log 2>&1
ncu -i /tmp/prof/ntt.ncu-rep --page raw --csv > gpurun_out/ntt_raw.csv
ls -la /tmp/prof
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_l1_16.csv python bench.py --workload l1-16sigs --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ncu_l1_16.log 2>&1
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > gpurun_out/tests_v2.log 2>&1
cat gpurun_out/tests_v2.log
