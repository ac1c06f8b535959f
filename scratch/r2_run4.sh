#!/bin/bash
# round 2, run 4: G2 pair kernel correctness + gate experiments (fixed ordering)
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q ) > gpurun_out/r4_tests.log 2>&1
tail -4 gpurun_out/r4_tests.log
run() { # name, env...
  name=$1; shift
  for wl in c2-layer-one l1-16sigs; do
    env "$@" timeout 300 python bench.py --steps 8 --warmup 3 --workload $wl --per-config none --no-cpu-baseline --no-verify > gpurun_out/r4_${name}_$wl.log 2>&1
    python - <<PY
import json
for l in open("gpurun_out/r4_${name}_$wl.log"):
    if l.startswith("{"):
        d=json.loads(l); print("$name $wl ms %.3f e2e %.3f bit_exact %s B2acc %.2f ntt_in_step %.2f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["bit_exact"], d["msm"]["B2"]["accumulate_ms"], d["roofline_ntt"]["ms_in_step"]))
PY
  done
}
run pair G16_G2_PAIR=1
run nopair G16_G2_PAIR=0
run pair_gate1 G16_G2_PAIR=1 G16_GATE=1
run pair_gate3 G16_G2_PAIR=1 G16_GATE=3
run pair_gate11 G16_G2_PAIR=1 G16_GATE=11
run pair_gate3_h2 G16_G2_PAIR=1 G16_GATE=3 G16_HPRIO=2
run pair_h1 G16_G2_PAIR=1 G16_HPRIO=1
