set -x
(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "table or prove" 2>&1 | tail -8) > gpurun_out/tests_table.log 2>&1
cat gpurun_out/tests_table.log
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_table_off.json 2> gpurun_out/bench_table_off.err
G16_H_TABLE=1 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_table_on.json 2> gpurun_out/bench_table_on.err
G16_H_TABLE=1 python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_table_on.json 2> gpurun_out/bench_l1_table_on.err
for f in off on; do python - <<PY
import json
d=json.loads(open("gpurun_out/bench_table_$f.json").read().strip().splitlines()[-1])
print("$f", d["ms_per_step"], d["e2e"]["value"], d["roofline"].get("ms_per_launch"), d.get("verified"), d.get("open_ms"))
PY
done
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_l1_table_on.json").read().strip().splitlines()[-1])
print("l1 on", d["ms_per_step"], d["roofline"].get("ms_per_launch"), d.get("verified"))
PY
tail -3 gpurun_out/bench_table_on.err
