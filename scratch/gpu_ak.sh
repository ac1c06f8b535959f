set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_ak.log 2>&1
cat gpurun_out/tests_ak.log
timeout 300 python bench_sweep.py --ntt-only --apply-key --min-log 21 --max-log 21 --out gpurun_out/sweep_ak.json > gpurun_out/sweep_ak.log 2>&1
tail -6 gpurun_out/sweep_ak.log
