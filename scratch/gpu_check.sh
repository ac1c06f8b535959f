set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/tests_chk.log 2>&1
cat gpurun_out/tests_chk.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/bench_chk2.json 2> gpurun_out/bench_chk2.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_chk2.json 2>/dev/null
tail -c 900 gpurun_out/bench_chk2.json
