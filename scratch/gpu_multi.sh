# usage: gpu_multi.sh N [N2 ...]   -- replica and shard benches on N GPUs of one node
set -x
nvidia-smi -L | head -8
python bench.py --steps 5 --no-cpu-baseline --no-verify > /dev/null 2> gpurun_out/multi_warm.err   # makes the fixture cache
python bench.py --workload l1-16sigs --steps 2 --no-cpu-baseline --no-verify > /dev/null 2>> gpurun_out/multi_warm.err
for N in "$@"; do
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$T --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_replica.json 2> gpurun_out/bench_n${N}_replica.err
$T --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --mode shard > gpurun_out/bench_n${N}_shard.json 2> gpurun_out/bench_n${N}_shard.err
$T --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --workload l1-16sigs > gpurun_out/bench_n${N}_replica_l1_16.json 2> gpurun_out/bench_n${N}_replica_l1_16.err
$T --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 --workload l1-16sigs --mode shard > gpurun_out/bench_n${N}_shard_l1_16.json 2> gpurun_out/bench_n${N}_shard_l1_16.err
done
wc -c gpurun_out/bench_n*_*.json
