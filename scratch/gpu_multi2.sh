set -x
nvidia-smi -L | head -4
python bench.py --steps 3 --no-cpu-baseline --no-verify > /dev/null 2> gpurun_out/multi_warm.err
N=2
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$T --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n${N}_replica.json 2> gpurun_out/bench_n${N}_replica.err
$T --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --mode shard --no-cpu-baseline > gpurun_out/bench_n${N}_shard.json 2> gpurun_out/bench_n${N}_shard.err
$T --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --mode shard-rep --no-cpu-baseline > gpurun_out/bench_n${N}_shardrep.json 2> gpurun_out/bench_n${N}_shardrep.err
for f in replica shard shardrep; do python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n2_$f.json").read().strip().splitlines()[-1])
    print("$f", d["n_gpus"], round(d["ms_per_step"],3), d["value"], d["e2e"]["ms_per_step"], d.get("verified"), d["scaling"])
except Exception as e:
    print("$f FAILED", e)
PY
tail -3 gpurun_out/bench_n2_$f.err
done
