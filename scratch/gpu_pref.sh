set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_pref.log 2>&1
cat gpurun_out/tests_pref.log
python bench.py --no-cpu-baseline > gpurun_out/bench_pref.json 2> gpurun_out/bench_pref.err
tail -3 gpurun_out/bench_pref.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_pref.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],3), d["e2e"]["ms_per_step"], d["e2e_pipelined"], d.get("verified"))
PY
