set -x
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-verify"
for v in 0 32 64 128; do
  if [ $v = 0 ]; then $B > gpurun_out/l2_$v.json 2> gpurun_out/l2_$v.err; else G16_L2_FETCH=$v $B > gpurun_out/l2_$v.json 2> gpurun_out/l2_$v.err; fi
  python - <<PY
import json
d=json.loads(open("gpurun_out/l2_$v.json").read().strip().splitlines()[-1])
print("L2_FETCH=$v", round(d["ms_per_step"],3), d["roofline"]["ms_per_launch"], d["phases_ms"]["ntt_pipeline"])
PY
done
B1="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
for v in 0 32 64; do
  if [ $v = 0 ]; then ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:msm_accumulate -s 19 -c 1 --csv --log-file gpurun_out/l2_ncu_$v.csv $B1 > /dev/null 2>&1
  else G16_L2_FETCH=$v ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:msm_accumulate -s 19 -c 1 --csv --log-file gpurun_out/l2_ncu_$v.csv $B1 > /dev/null 2>&1; fi
  grep -v "^==" gpurun_out/l2_ncu_$v.csv | cut -d, -f5,13- | tail -3
done
