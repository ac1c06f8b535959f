set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/tests_final.log 2>&1
cat gpurun_out/tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; tail -2 gpurun_out/smoke_final.log
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_final.json 2> gpurun_out/bench_l1_16_final.err
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final.csv $B > gpurun_out/ncu_final.log 2>&1
python bench_cli.py > gpurun_out/cli_wall_final.json 2> gpurun_out/cli_wall_final.err
(time timeout 600 python bench.py --workload c2-layer-two --steps 3 --no-cpu-baseline) > gpurun_out/bench_l2_final.json 2> gpurun_out/bench_l2_final.err
tail -4 gpurun_out/bench_l2_final.err
for f in bench_final bench_l1_16_final bench_l2_final; do python - <<PY
import json
d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
print("$f", round(d["ms_per_step"],3), round(d["e2e"]["ms_per_step"],3), d["roofline"].get("ms_per_launch"), d["roofline"].get("frac"), d.get("verified"), d["msm"]["H"]["c"], d.get("zkey_open_s"))
PY
done
cat gpurun_out/cli_wall_final.json | head -30
