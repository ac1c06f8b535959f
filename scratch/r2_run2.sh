#!/bin/bash
# round 2, run 2: sync-free counting-sort MSM pipeline
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_tests.log 2>&1
tail -5 gpurun_out/r2_tests.log
( timeout 300 python bench.py --steps 10 --warmup 3 --per-config none --no-cpu-baseline ) > gpurun_out/r2_bench.log 2>&1
( G16_W_TABLE=1 timeout 300 python bench.py --steps 10 --warmup 3 --per-config none --no-cpu-baseline ) > gpurun_out/r2_bench_wtable.log 2>&1
( timeout 300 python bench.py --steps 5 --warmup 3 --workload l1-16sigs --per-config none --no-cpu-baseline ) > gpurun_out/r2_bench_l16.log 2>&1
( G16_W_TABLE=1 timeout 300 python bench.py --steps 5 --warmup 3 --workload l1-16sigs --per-config none --no-cpu-baseline ) > gpurun_out/r2_bench_l16_wtable.log 2>&1
tail -c 400 gpurun_out/r2_bench.log
