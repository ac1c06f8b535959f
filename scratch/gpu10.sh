set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_v6.log 2>&1
cat gpurun_out/tests_v6.log
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_v6.json 2> gpurun_out/bench_v6.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_v6.json 2> gpurun_out/bench_l1_16_v6.err
timeout 300 python bench_sweep.py --min-log 21 --max-log 24 --max-log-g2 20 --out gpurun_out/sweep_v6.json > gpurun_out/sweep_v6.log 2> gpurun_out/sweep_v6.err
tail -12 gpurun_out/sweep_v6.err
