set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_tail.log 2>&1
cat gpurun_out/tests_tail.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_tail.log 2>&1; tail -2 gpurun_out/smoke_tail.log
python bench.py > gpurun_out/bench_tail.json 2> gpurun_out/bench_tail.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_tail.json 2> gpurun_out/bench_l1_tail.err
for f in bench_tail bench_l1_tail; do python - <<PY
import json
d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
print("$f", round(d["ms_per_step"],3), d["e2e"]["ms_per_step"], (d.get("e2e_pipelined") or {}).get("ms_per_step"), d["roofline"]["frac"], d.get("verified"), d["trace_last_step_ms"]["H"])
PY
done
