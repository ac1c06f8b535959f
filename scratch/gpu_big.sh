set -x
C=/tmp/g16_bench_cache; P=zk-proof-of-assets_b200/prover
python bench.py --steps 3 --no-cpu-baseline > gpurun_out/bench_chk.json 2> gpurun_out/bench_chk.err
for i in 1 2; do G16_VERBOSE=1 $P $C/c2-layer-one.zkey $C/c2-layer-one.wtns /tmp/p.json /tmp/u.json; done 2> gpurun_out/cli_verbose.log
cat gpurun_out/cli_verbose.log
(time python bench.py --workload c2-layer-two --steps 3 --no-cpu-baseline) > gpurun_out/bench_l2_final.json 2> gpurun_out/bench_l2_final.err
tail -4 gpurun_out/bench_l2_final.err
rm -f $C/c2-layer-two.*
df -h /tmp | tail -1; free -g | head -2
(time timeout 1200 python bench.py --workload c2-layer-three --steps 2 --no-cpu-baseline) > gpurun_out/bench_l3_final.json 2> gpurun_out/bench_l3_final.err
tail -6 gpurun_out/bench_l3_final.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
