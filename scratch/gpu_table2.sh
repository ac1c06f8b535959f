set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/tests_t2.log 2>&1
cat gpurun_out/tests_t2.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_t2.log 2>&1; tail -2 gpurun_out/smoke_t2.log
python bench.py > gpurun_out/bench_t2.json 2> gpurun_out/bench_t2.err
G16_H_TABLE=0 python bench.py --no-cpu-baseline > gpurun_out/bench_t2_notable.json 2> gpurun_out/bench_t2_notable.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_t2.json 2> gpurun_out/bench_l1_16_t2.err
G16_H_TABLE=0 python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_t2_notable.json 2> gpurun_out/bench_l1_16_t2_notable.err
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
mkdir -p /tmp/prof
ncu --set full --clock-control none --import-source on -k regex:msm_accumulate -s 15 -c 5 -o /tmp/prof/acc $B > gpurun_out/ncu_acc_t2.log 2>&1
ncu -i /tmp/prof/acc.ncu-rep --page raw --csv > gpurun_out/acc_raw_t2.csv
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_t2.csv $B > gpurun_out/ncu_launch_t2.log 2>&1
for f in bench_t2 bench_t2_notable bench_l1_16_t2 bench_l1_16_t2_notable; do python - <<PY
import json
d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
print("$f", round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3), d["roofline"].get("ms_per_launch"), d["roofline"].get("frac"), d.get("verified"), d["trace_last_step_ms"]["H"])
PY
done
ls -la gpurun_out | tail -20
