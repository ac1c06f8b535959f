set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/tests_final.log 2>&1
cat gpurun_out/tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; tail -2 gpurun_out/smoke_final.log
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_final.json 2> gpurun_out/bench_l1_16_final.err
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final.csv $B > gpurun_out/ncu_final.log 2>&1
mkdir -p /tmp/prof
ncu --set full --clock-control none --import-source on -k regex:msm_accumulate -s 15 -c 5 -o /tmp/prof/acc $B > gpurun_out/ncu_acc.log 2>&1
ncu -i /tmp/prof/acc.ncu-rep --page raw --csv > gpurun_out/acc_raw_final.csv
ncu --set full --clock-control none --import-source on -k regex:ntt_pass -s 54 -c 6 -o /tmp/prof/ntt $B > gpurun_out/ncu_ntt.log 2>&1
ncu -i /tmp/prof/ntt.ncu-rep --page raw --csv > gpurun_out/ntt_raw_final.csv
python bench_cli.py > gpurun_out/cli_wall_final.json 2> gpurun_out/cli_wall_final.err
ls -la gpurun_out | tail -15
