set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_v5.log 2>&1
cat gpurun_out/tests_v5.log
python bench.py --steps 5 --no-cpu-baseline --no-verify > /dev/null 2> gpurun_out/multi_warm.err
python bench.py --workload l1-16sigs --steps 2 --no-cpu-baseline --no-verify > /dev/null 2>> gpurun_out/multi_warm.err
N=2
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$T --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --mode shard > gpurun_out/bench_n${N}_sharddist.json 2> gpurun_out/bench_n${N}_sharddist.err
$T --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 --workload l1-16sigs --mode shard > gpurun_out/bench_n${N}_sharddist_l1_16.json 2> gpurun_out/bench_n${N}_sharddist_l1_16.err
tail -n 5 gpurun_out/bench_n${N}_sharddist.err gpurun_out/bench_n${N}_sharddist_l1_16.err
wc -c gpurun_out/bench_n2_sharddist*.json
