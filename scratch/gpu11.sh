set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_v7.log 2>&1
cat gpurun_out/tests_v7.log
B="python bench.py --steps 10 --no-cpu-baseline --no-verify"
$B > gpurun_out/ab_base.json 2>/dev/null
G16_ACC_MINB=4 $B > gpurun_out/ab_minb4.json 2>/dev/null
$B --msm-c 17 > gpurun_out/ab_c17.json 2>/dev/null
$B --msm-c 15 > gpurun_out/ab_c15.json 2>/dev/null
B="python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline --no-verify"
$B > gpurun_out/ab24_base.json 2>/dev/null
G16_ACC_MINB=4 $B > gpurun_out/ab24_minb4.json 2>/dev/null
$B --msm-c 20 > gpurun_out/ab24_c20.json 2>/dev/null
$B --msm-c 18 > gpurun_out/ab24_c18.json 2>/dev/null
for f in gpurun_out/ab*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print(round(d['ms_per_step'],2), {k:(v['c'],round(v['accumulate_ms'],2)) for k,v in d['msm'].items()}, round(d['roofline']['frac'],3))
"; done
