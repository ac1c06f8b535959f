#!/bin/bash
# round 2, run 1: whole GPU suite + both bench arms on the round-1 kernels (baseline for this round's changes)
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
nproc > gpurun_out/r1_nproc.txt; free -g >> gpurun_out/r1_nproc.txt
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r1_tests.log 2>&1
( time python bench.py --impl reference --steps 5 --warmup 1 ) > gpurun_out/r1_ref.log 2>&1
( time python bench.py --steps 10 --warmup 3 ) > gpurun_out/r1_bench.log 2>&1
tail -3 gpurun_out/r1_tests.log; tail -c 600 gpurun_out/r1_ref.log; tail -c 300 gpurun_out/r1_bench.log
