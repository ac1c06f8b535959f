set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/tests_final.log 2>&1
cat gpurun_out/tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; tail -2 gpurun_out/smoke_final.log
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err
python bench.py --workload l1-16sigs --steps 3 --no-cpu-baseline > gpurun_out/bench_l1_16_final.json 2> gpurun_out/bench_l1_16_final.err
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final.csv $B > gpurun_out/ncu_final.log 2>&1





python bench_cli.py > gpurun_out/cli_wall_final.json 2> gpurun_out/cli_wall_final.err
ls -la gpurun_out | tail -15
C=/tmp/g16_bench_cache; P=zk-proof-of-assets_b200/prover
for i in 1 2; do G16_VERBOSE=1 $P $C/c2-layer-one.zkey $C/c2-layer-one.wtns /tmp/p.json /tmp/u.json; done 2> gpurun_out/cli_verbose.log
cat gpurun_out/cli_verbose.log
