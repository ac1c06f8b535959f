#!/usr/bin/env python3
"""bench_cli.py -- wall time of the reference's own boundary: `prover <zkey> <wtns> <proof.json> <public.json>`
(scripts/g16_prove.sh:246-252), from exec to exit, as GNU time reports it in the reference's STATS lines
(scripts/lib/cmd_executor.sh:6,17).  Two arms per workload, page cache warm, zkey on local disk:
  in-process : every call reads, parses and uploads the key (what rapidsnark does too)
  service    : `prover --serve` holds the key on the GPU; the same command line is a thin client (G16_SERVICE=...)
One JSON line per workload.  Uses the fixtures bench.py caches under G16_BENCH_CACHE.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (fixture cache + workload table)

EXE = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")


def timed(cmd, env):
    t = time.perf_counter()
    p = subprocess.run(cmd, env=env, capture_output=True, text=True)
    dt = time.perf_counter() - t
    if p.returncode != 0:
        raise SystemExit("prover failed: " + p.stderr)
    return dt, p.stderr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="c2-layer-one,l1-16sigs")
    ap.add_argument("--runs", type=int, default=5)
    ap.add_argument("--devices", default="0")
    args = ap.parse_args()
    sock = "/tmp/g16_bench_cli.%d.sock" % os.getpid()
    for name in args.workloads.split(","):
        bench.fixture(name, load=False)
        zp, wp = os.path.join(bench.CACHE_DIR, name + ".zkey"), os.path.join(bench.CACHE_DIR, name + ".wtns")
        out = ["/tmp/g16_cli_proof.json", "/tmp/g16_cli_public.json"]
        env = dict(os.environ)
        env.pop("G16_SERVICE", None)
        env.pop("CUDA_DEVICE_MAX_CONNECTIONS", None)     # importing bench set it for THIS process; the command line decides for itself
        env.pop("G16_CONNECTIONS_DEFAULTED", None)
        cold = [timed([EXE, zp, wp] + out, env)[0] for _ in range(args.runs)]
        ref_proof = open(out[0]).read()
        srv = subprocess.Popen([EXE, "--serve", "--socket", sock, "--devices", args.devices], stderr=subprocess.DEVNULL)
        try:
            for _ in range(100):
                if subprocess.run([EXE, "--ping", "--socket", sock], capture_output=True).returncode == 0:
                    break
                time.sleep(0.05)
            senv = dict(env, G16_SERVICE=sock, G16_VERBOSE="1")
            first, _ = timed([EXE, zp, wp] + out, senv)
            warm, notes = [], []
            for _ in range(args.runs):
                dt, err = timed([EXE, zp, wp] + out, senv)
                warm.append(dt)
                notes.append(err.strip())
            ok = json.loads(open(out[0]).read()).keys() == json.loads(ref_proof).keys()
        finally:
            subprocess.run([EXE, "--shutdown", "--socket", sock], capture_output=True)
            srv.wait(timeout=60)
        nv, npub, ncons, desc = bench.WORKLOADS[name]
        print(json.dumps({
            "metric": "prover_cli_wall_s", "workload": name, "description": desc, "n_constraints": ncons,
            "zkey_bytes": os.path.getsize(zp), "runs": args.runs,
            "in_process_s": {"median": statistics.median(cold), "min": min(cold), "all": cold},
            "service_first_call_s": first,
            "service_resident_s": {"median": statistics.median(warm), "min": min(warm), "all": warm},
            "service_reply_last": notes[-1] if notes else "", "outputs_well_formed": bool(ok)}), flush=True)


if __name__ == "__main__":
    main()
