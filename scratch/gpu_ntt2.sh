set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/tests_ntt.log 2>&1
cat gpurun_out/tests_ntt.log
python bench.py --no-cpu-baseline > gpurun_out/bench_ntt_r4.json 2> gpurun_out/bench_ntt_r4.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_ntt_r4.json 2> gpurun_out/bench_l1_ntt_r4.err
timeout 200 python bench_sweep.py --ntt-only --min-log 20 --max-log 26 --out gpurun_out/sweep_ntt_r4.json > gpurun_out/sweep_ntt_r4.log 2>&1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
mkdir -p /tmp/prof
ncu --set full --clock-control none --import-source on -k regex:ntt_pass -s 54 -c 6 -o /tmp/prof/ntt $B > gpurun_out/ncu_ntt3.log 2>&1
ncu -i /tmp/prof/ntt.ncu-rep --page raw --csv > gpurun_out/ntt_raw3.csv
for f in bench_ntt_r4 bench_l1_ntt_r4; do python - <<PY
import json
d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
print("$f", round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3), d["phases_ms"]["ntt_pipeline"], d.get("verified"), d["trace_last_step_ms"]["H"])
PY
done
python - <<PY
import json
d=json.load(open("gpurun_out/sweep_ntt_r4.json"))
print([(l["log_n"], round(l["ms"],3), round(l["frac_of_modmul_peak"],3)) for l in d["lines"]])
PY
