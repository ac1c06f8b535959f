"""experiment: K proofs in flight on ONE GPU (K handles of the same key, K host threads) -> aggregate throughput"""
import sys, time, threading, json
sys.path.insert(0, "/root/repo")
import bench
pkg = bench.load_pkg()
name = sys.argv[1] if len(sys.argv) > 1 else "c2-layer-one"
zkey, wtns = bench.fixture(name)
ncons = bench.WORKLOADS[name][2]
for K in (1, 2, 3):
    hs = [pkg.ZKey(zkey, device=0) for _ in range(K)]
    for h in hs:
        h.load_witness(wtns)
        for _ in range(3):
            h.prove_loaded()
    steps = 12
    def run(h):
        for _ in range(steps):
            h.prove_loaded()
    ts = [threading.Thread(target=run, args=(h,)) for h in hs]
    t0 = time.perf_counter()
    for t in ts: t.start()
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    print(json.dumps({"workload": name, "in_flight": K, "ms_per_proof_aggregate": dt / (K * steps) * 1e3,
                      "constraints_per_s": ncons * K * steps / dt}), flush=True)
    for h in hs: h.close()
