set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/tests_v3.log 2>&1
cat gpurun_out/tests_v3.log
python bench.py --steps 10 --no-cpu-baseline > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_v3.json 2> gpurun_out/bench_l1_16_v3.err
# CLI wall (GNU time style) with and without the resident-key service, page cache warm
C=/tmp/g16_bench_cache; P=zk-proof-of-assets_b200/prover
for W in c2-layer-one l1-16sigs; do
  for i in 1 2 3; do /usr/bin/time -f "cli_inprocess $W %e s %M KB" $P $C/$W.zkey $C/$W.wtns /tmp/p.json /tmp/u.json; done
done 2> gpurun_out/cli_wall.log
$P --serve --socket /tmp/s.sock 2> gpurun_out/service.log &
sleep 1
for W in c2-layer-one l1-16sigs; do
  for i in 1 2 3 4; do G16_VERBOSE=1 G16_SERVICE=/tmp/s.sock /usr/bin/time -f "cli_service $W %e s %M KB" $P $C/$W.zkey $C/$W.wtns /tmp/p.json /tmp/u.json; done
done 2>> gpurun_out/cli_wall.log
$P --stats --socket /tmp/s.sock >> gpurun_out/cli_wall.log
$P --shutdown --socket /tmp/s.sock
wait
cat gpurun_out/cli_wall.log
