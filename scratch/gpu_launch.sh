set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
$B > gpurun_out/plain.json 2> gpurun_out/plain.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final.csv $B > gpurun_out/ncu_final.log 2>&1
tail -2 gpurun_out/ncu_final.log | cut -c1-200
