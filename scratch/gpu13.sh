set -x
(G16_PAIR_MERGE=1 G16_PAIR_MERGE_MIN=0 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "msm or prove" 2>&1 | tail -6) > gpurun_out/tests_merge2.log 2>&1
cat gpurun_out/tests_merge2.log
B="python bench.py --steps 10 --no-cpu-baseline"
$B > gpurun_out/ab3_base.json 2>/dev/null
G16_PAIR_MERGE=1 $B > gpurun_out/ab3_merge.json 2>gpurun_out/ab3_merge.err
B="python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline"
$B > gpurun_out/ab324_base.json 2>/dev/null
G16_PAIR_MERGE=1 $B > gpurun_out/ab324_merge.json 2>gpurun_out/ab324_merge.err
