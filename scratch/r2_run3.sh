#!/bin/bash
# round 2, run 3: Merkle tests + scheduling knobs (G16_GATE x G16_HPRIO)
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_gpu_merkle.py -x -q -s ) > gpurun_out/r3_merkle.log 2>&1
tail -15 gpurun_out/r3_merkle.log
for combo in "0 0" "1 0" "3 0" "0 2" "1 2" "3 2" "4 0" "4 2"; do
  set -- $combo
  for wl in c2-layer-one l1-16sigs; do
    G16_GATE=$1 G16_HPRIO=$2 timeout 300 python bench.py --steps 8 --warmup 3 --workload $wl --per-config none --no-cpu-baseline --no-verify > gpurun_out/r3_g$1_h$2_$wl.log 2>&1
    python - <<PY
import json
for l in open("gpurun_out/r3_g$1_h$2_$wl.log"):
    if l.startswith("{"):
        d=json.loads(l); print("gate $1 hprio $2 $wl ms %.3f e2e %.3f bit_exact %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["bit_exact"]))
PY
  done
done
