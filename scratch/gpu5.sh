set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_v4.log 2>&1
cat gpurun_out/tests_v4.log
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_v4.json 2> gpurun_out/bench_v4.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_v4.json 2> gpurun_out/bench_l1_16_v4.err
python bench_cli.py > gpurun_out/cli_wall.json 2> gpurun_out/cli_wall.err
cat gpurun_out/cli_wall.json
timeout 600 python bench_sweep.py --out gpurun_out/sweep.json > gpurun_out/sweep.log 2> gpurun_out/sweep.err
tail -42 gpurun_out/sweep.err
