set -x
(G16_PAIR_MERGE=1 G16_PAIR_MERGE_MIN=0 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/tests_merge.log 2>&1
cat gpurun_out/tests_merge.log
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "msm" 2>&1 | tail -5) > gpurun_out/tests_v8.log 2>&1
cat gpurun_out/tests_v8.log
B="python bench.py --steps 10 --no-cpu-baseline"
$B > gpurun_out/ab2_base.json 2>/dev/null
G16_ACC_AFTER_NTT=1 $B > gpurun_out/ab2_gate.json 2>/dev/null
G16_PAIR_MERGE=1 $B > gpurun_out/ab2_merge.json 2>gpurun_out/ab2_merge.err
G16_PAIR_MERGE=1 G16_ACC_AFTER_NTT=1 $B > gpurun_out/ab2_merge_gate.json 2>/dev/null
B="python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline"
$B > gpurun_out/ab224_base.json 2>/dev/null
G16_ACC_AFTER_NTT=1 $B > gpurun_out/ab224_gate.json 2>/dev/null
G16_PAIR_MERGE=1 $B > gpurun_out/ab224_merge.json 2>gpurun_out/ab224_merge.err
G16_PAIR_MERGE=1 G16_ACC_AFTER_NTT=1 $B > gpurun_out/ab224_merge_gate.json 2>/dev/null
tail -3 gpurun_out/ab2_merge.err gpurun_out/ab224_merge.err
