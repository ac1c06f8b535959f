set -x
N=8
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$T --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --mode shard > gpurun_out/bench_n${N}_sharddist.json 2> gpurun_out/bench_n${N}_sharddist.err
$T --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 --workload l1-16sigs --mode shard > gpurun_out/bench_n${N}_sharddist_l1_16.json 2> gpurun_out/bench_n${N}_sharddist_l1_16.err
N=4
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$T --master-port 29516 bench.py --gpus $N --steps 5 --warmup 3 --workload l1-16sigs --mode shard > gpurun_out/bench_n${N}_sharddist_l1_16.json 2> gpurun_out/bench_n${N}_sharddist_l1_16.err
tail -n 4 gpurun_out/bench_n*_sharddist*.err
wc -c gpurun_out/bench_n*_sharddist*.json
