#!/usr/bin/env python3
"""bench.py -- Groth16 prove throughput of the B200 prover (and of the CPU reference arm) on one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one Groth16 prove (the thing scripts/g16_prove.sh:246-260 of the reference execs a prover for) of one
witness against a resident proving key.  Metric = constraints proven per second (BASELINE.json: "Groth16 prove s &
constraints/s per layer"); `ms_per_step` is the prove time, `msm_mpts_per_s` the MSM point rate inside it.
  value : witness already in HBM (g16_witness_load done), timed g16_prove_loaded, CUDA-side work + host finalisation
  e2e   : the public call ZKey.prove(wtns bytes in pinned host memory) -> JSON texts in host memory, per step
N > 1   : one process per GPU, every rank proves its own independent batch proof (the reference's GNU-parallel batch
          loop, scripts/full_workflow.sh:552): weak scaling, no data-path collective; barrier + max over ranks.
          --mode shard: ONE proof over all ranks instead (strong scaling; NCCL for slices and 768-byte partials).
Inputs are synthetic (SYNTH-1 circuits shaped like the reference's layers, known-trapdoor setup): the reference commits
no .zkey/.wtns and none can be generated offline.  The reference arm (--impl reference) times the CPU oracle port
(oracle/c/oracle.c, all host threads) because neither rapidsnark nor snarkjs can be installed here.
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# name -> (n_vars, n_public, n_constraints, description); sizes from SURVEY.md section 6a / 8(d)
WORKLOADS = {
    "c2-layer-one": (2083343, 1, 2094670, "configs[1] tests/4_sigs_2_batches_12_height layer_one batch (2 sigs), domain 2^21"),
    "c1-layer-one": (1632790, 1, 1641360, "configs[0] tests/1_sigs_1_batches_5_height layer_one (1 sig), domain 2^21"),
    "l1-16sigs": (8651416, 1, 8702867, "configs[2] layer_one batch-ECDSA 16 sigs, domain 2^24"),
    "c2-layer-two": (21356921, 2, 21481888, "configs[1] layer_two (2 sigs, h=12), domain 2^25"),
    "c1-layer-three": (28128132, 13, 28328996, "configs[3] layer_three recursive verifier (1 batch), domain 2^25"),
    "c2-layer-three": (52367163, 13, 52722297, "configs[1] layer_three (2 batches), domain 2^26"),
    "tiny": (30000, 1, 32000, "smoke-sized circuit, domain 2^15"),
}
DEFAULT_WORKLOAD = "c2-layer-one"
CIRCUIT_SEED, TRAPDOOR_SEED = 0xB2000001, 0xB2000002
CACHE_DIR = os.environ.get("G16_BENCH_CACHE", "/tmp/g16_bench_cache")


def load_pkg():
    if "g16b200" in sys.modules:
        return sys.modules["g16b200"]
    spec = importlib.util.spec_from_file_location("g16b200", os.path.join(ROOT, "zk-proof-of-assets_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["g16b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def fixture(name, rank=0):
    """(zkey bytes, wtns bytes) of a workload; generated once per box (GPU fixed-base multiplications through
    libg16synth.so, which is bench support and not part of either timed arm) and cached under CACHE_DIR."""
    nv, npub, ncons, _ = WORKLOADS[name]
    os.makedirs(CACHE_DIR, exist_ok=True)
    zp, wp = os.path.join(CACHE_DIR, name + ".zkey"), os.path.join(CACHE_DIR, name + ".wtns")
    if not (os.path.exists(zp) and os.path.exists(wp)):
        lock = zp + ".lock"
        try:
            fd = os.open(lock, os.O_CREAT | os.O_EXCL | os.O_WRONLY)
        except FileExistsError:
            fd = None
        if fd is not None:
            try:
                pkg = load_pkg()
                sy = pkg.Synth(nv, npub, ncons, CIRCUIT_SEED, TRAPDOOR_SEED)
                zk, wt = sy.write()
                vk = json.dumps(sy.vkey()).encode()
                sy.close()
                for path, data in ((zp[:-5] + ".vkey.json", vk), (wp, wt), (zp, zk)):
                    with open(path + ".tmp", "wb") as f:
                        f.write(data)
                    os.replace(path + ".tmp", path)
            finally:
                os.close(fd)
                os.unlink(lock)
        else:
            while os.path.exists(lock):
                time.sleep(0.5)
    with open(zp, "rb") as f:
        zk = f.read()
    with open(wp, "rb") as f:
        wt = f.read()
    return zk, wt


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                power.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for k, nm in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---- work model (DESIGN.md section "Kernels and rooflines") ---------------------------------------------------
MODMUL_OPS = 136          # 32x32->64 multiply-adds per 254-bit Montgomery product (128 wide + 8 quotient digits)
G1_MADD_MODMULS = 10      # XYZZ mixed addition: 8M + 2S over Fq
G2_MADD_MODMULS = 28      # the same over Fq2: 8 x 3 (Karatsuba) + 2 x 2
NTT_BYTES_PER_POINT = 384  # H pipeline, strict: 6 transforms x (32 B read + 32 B write) per domain point


H_TABLE = {"auto": 0, "on": 1, "off": -1}


def measured_traffic(workload):
    """dram bytes from the committed ncu --set full capture of this workload (profiles/traffic.json), or {}"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload, {})
    except (OSError, ValueError):
        return {}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except OSError:
        return {}


def dist_setup(n_gpus):
    """-> (rank, world, local_rank, dist module or None)"""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        return rank, world, local, dist
    return 0, 1, 0, None


def run_b200(args):
    import torch
    rank, world, local, dist = dist_setup(args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the prover has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pkg = load_pkg()
    nv, npub, ncons, desc = WORKLOADS[args.workload]
    if rank == 0:
        fixture(args.workload)          # one rank generates, the others read the cache
    if dist:
        dist.barrier()
    zkey, wtns = fixture(args.workload, rank)

    t_open = time.perf_counter()
    shard = args.mode in ("shard", "shard-rep") and world > 1
    dist_ntt = args.mode == "shard"     # coset transforms spread over the ranks too (one all-to-all per polynomial)
    if shard:
        zk = pkg.ZKey(zkey, device=local, msm_c=args.msm_c, shard_rank=rank, shard_count=world, h_table=H_TABLE[args.h_table])
    else:
        zk = pkg.ZKey(zkey, device=local, msm_c=args.msm_c, h_table=H_TABLE[args.h_table])
    t_open = time.perf_counter() - t_open
    zkey_len = len(zkey)
    del zkey
    pinned = torch.empty(len(wtns), dtype=torch.uint8).pin_memory()
    pinned.copy_(torch.frombuffer(bytearray(wtns), dtype=torch.uint8))

    def sync_all():
        torch.cuda.synchronize(dev)
        if dist:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # sharded mode: ONE proof over all ranks (base-range shards of sections 5-9, the H pipeline replicated, one
    # NCCL all-gather of 768-byte partials, every rank finalises); pinned r,s so all ranks agree
    R_PIN, S_PIN = 0x1234567 << 200, 0x7654321 << 190

    def step_loaded():
        if shard and dist_ntt:
            return pkg.prove_sharded_distributed(zk, None, r=R_PIN, s=S_PIN)
        if shard:
            return zk.combine(pkg.all_gather_partials(zk.prove_partial()), r=R_PIN, s=S_PIN)
        return zk.prove_loaded()

    def step_e2e():
        if shard and dist_ntt:
            return pkg.prove_sharded_distributed(zk, pinned, r=R_PIN, s=S_PIN)
        if shard:
            return pkg.prove_sharded(zk, pinned, r=R_PIN, s=S_PIN)
        return zk.prove(pinned)

    # ---- resident arm: witness in HBM before the timed region -----------------------------------------------------
    zk.load_witness(pinned)
    for _ in range(args.warmup):
        proof, public = step_loaded()
    clocks = ClockSampler(local)
    phase = {}
    acc_ms = {k: 0.0 for k in ("A", "B1", "B2", "C", "H")}
    sync_all()
    if rank == 0:
        clocks.start()
    launches0 = pkg.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    w0 = time.perf_counter()
    for _ in range(args.steps):
        proof, public = step_loaded()
        for k, v in zk.timings().items():
            phase[k] = phase.get(k, 0.0) + v
        for k, v in zk.accumulate_ms().items():
            acc_ms[k] += v
    e1.record()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - w0
    dev_ms = e0.elapsed_time(e1)
    launches = pkg.launch_count() - launches0
    clk = clocks.stop() if rank == 0 else None
    resident_s = max_over_ranks(max(dev_ms / 1e3, wall))
    stats = zk.msm_stats()

    # ---- end-to-end arm: pinned host bytes in, JSON texts out, every step ------------------------------------------
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    sync_all()
    w0 = time.perf_counter()
    for _ in range(args.steps):
        proof, public = step_e2e()
    torch.cuda.synchronize(dev)
    e2e_s = max_over_ranks(time.perf_counter() - w0)
    d2h = len(proof) + len(public)

    # ---- the same through ZKey.prove_stream: the next step's witness is copied (pinned host -> device) while this step
    # computes (g16_witness_prefetch / g16_witness_commit); every step still has its own H2D copy and its own JSON out
    pipe_s = None
    if not shard:
        for _ in zk.prove_stream([pinned] * max(2, args.warmup // 2)):
            pass
        sync_all()
        w0 = time.perf_counter()
        for proof, public in zk.prove_stream([pinned] * args.steps):
            pass
        torch.cuda.synchronize(dev)
        pipe_s = max_over_ranks(time.perf_counter() - w0)

    # ---- self-check of the last proof (untimed): it must verify ------------------------------------------------------
    ok = None
    if rank == 0 and args.verify:
        # g16_verify: the native host verifier (csrc/verify*.hpp), itself pinned on the reference's committed
        # (vkey, public, proof) triples in tests/test_verify_native.py
        with open(os.path.join(CACHE_DIR, args.workload + ".vkey.json")) as f:
            ok = bool(pkg.verify(f.read(), public, proof))

    steps = args.steps
    jobs = 1 if shard else world     # proofs finished per step over the whole job
    g1_adds = sum(stats[k]["additions"] for k in ("A", "B1", "C", "H"))
    g2_adds = stats["B2"]["additions"]
    msm_points = 3 * nv - npub - 1 + zk.domain_size + nv
    msm_ms = sum(phase[k] for k in ("msm_A", "msm_B1", "msm_B2", "msm_C", "msm_H")) / steps
    acc_total_ms = sum(acc_ms.values()) / steps
    int_ops = MODMUL_OPS * (G1_MADD_MODMULS * g1_adds + G2_MADD_MODMULS * g2_adds)
    out = None
    if rank == 0:
        peaks = measured_peaks()
        traffic = measured_traffic(args.workload)
        # the H section's fixed-base window table (opts.h_table): one bucket set of c+4-bit windows; told apart from the
        # plain launch by the window width the run reports
        plain_c = max(6, min(20, zk.domain_size.bit_length() - 1 - 5))     # msm_pick_c (csrc/msm.cuh)
        h_tabled = not args.msm_c and stats["H"]["c"] > plain_c
        if not h_tabled and "msm_accumulate_H_plain" in traffic:
            traffic = dict(traffic, msm_accumulate_H=traffic["msm_accumulate_H_plain"])
            traffic.pop("msm_accumulate", None)
        acc_alg_bytes = 68 * g1_adds + 132 * g2_adds      # one base record + one 4-byte entry per mixed addition
        h_ms = acc_ms["H"] / steps
        h_ops = MODMUL_OPS * G1_MADD_MODMULS * stats["H"]["additions"]
        imad_peak = pkg.bench_int_pipe(1)            # wide multiply-adds per second, measured on this GPU now
        modmul_peak = pkg.bench_int_pipe(2)          # dependent-chain Montgomery products per second
        hbm_peak = peaks.get("hbm_gbs")
        ntt_gbs = NTT_BYTES_PER_POINT * zk.domain_size / (phase["ntt_pipeline"] / steps * 1e-3) / 1e9
        out = {
            "metric": "groth16_prove_constraints_per_s", "value": ncons * steps * jobs / resident_s,
            "unit": "constraints/s", "n_gpus": world, "steps": steps, "warmup": args.warmup,
            "ms_per_step": resident_s / steps * 1e3, "higher_is_better": True, "scaling": "strong" if shard else "weak",
            "vs_baseline": None, "dtype": "u32 limbs (256-bit Montgomery Fr/Fq/Fq2)", "data": "synthetic",
            "config": {"workload": args.workload, "description": desc, "n_vars": nv, "n_public": npub,
                       "n_constraints": ncons, "domain_size": zk.domain_size, "n_coefs": zk.n_coefs,
                       "h_window_table": h_tabled,
                       "zkey_bytes": zkey_len, "proofs_per_step": jobs,
                       "parallelism": ("base-range shards x%d, %s, NCCL all-gather of 768 B partials"
                                       % (world, "A/B/C transforms on ranks 0-%d + all-to-all of slices" % (min(world, 3) - 1)
                                          if dist_ntt else "transforms replicated")) if shard else "replica-per-gpu",
                       "l2": "inputs exceed L2 (bases + scalars > 1 GB touched per step); no flush",
                       "timing": "cuda events on the default stream around synchronous prove calls, max(event, host clock), max over ranks"},
            "prove_s": resident_s / steps,
            "msm_mpts_per_s": msm_points / (msm_ms * 1e-3) / 1e6 if msm_ms else None,
            "phases_ms": {k: v / steps for k, v in phase.items()},
            "msm": {k: dict(v, accumulate_ms=acc_ms[k] / steps) for k, v in stats.items()},
            "trace_last_step_ms": zk.trace(),
            "zkey_open_s": t_open,
            "clocks": clk,
            "e2e_pipelined": ({"value": ncons * steps * jobs / pipe_s, "unit": "constraints/s",
                               "ms_per_step": pipe_s / steps * 1e3, "h2d_bytes_per_step": len(wtns),
                               "api": "ZKey.prove_stream: witness i+1 copied while proof i computes"} if pipe_s else None),
            "e2e": {"value": ncons * steps * jobs / e2e_s, "unit": "constraints/s", "ms_per_step": e2e_s / steps * 1e3,
                    "h2d_bytes_per_step": len(wtns), "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            # the dominant kernel = the bucket accumulation of the H MSM (one launch per step, the largest MSM at every
            # config); CUDA events on its own stream around that launch.  The four witness-MSM launches overlap the
            # transforms and each other, so their event times are not kernel durations: they are listed apart.
            "roofline": {"kernel": "msm_accumulate_kernel<G1>, H MSM launch (bucket accumulation; dominant kernel)",
                         "bound": "int", "achieved": h_ops / (h_ms * 1e-3) / 1e12 if h_ms else None,
                         "peak": imad_peak / 1e12, "unit": "T wide-IMAD/s",
                         "frac": (h_ops / (h_ms * 1e-3)) / imad_peak if h_ms else None,
                         "units_per_launch": stats["H"]["additions"], "ops_per_unit": MODMUL_OPS * G1_MADD_MODMULS,
                         "ms_per_launch": h_ms,
                         "peak_source": "g16_bench_int_pipe(1): IMAD.WIDE.U32 carry chains, measured in this run "
                                        "(MEASURED_PEAKS.json has no integer figure)",
                         "modmul_peak_per_s": modmul_peak, "share_of_step": h_ms / (resident_s / steps * 1e3),
                         "algorithmic_bytes": 68 * stats["H"]["additions"],
                         "traffic": traffic.get("msm_accumulate_H", {}).get("bytes_per_launch"),
                         "traffic_source": traffic.get("msm_accumulate_H", {}).get("source")},
            "roofline_all_accumulate": {"kernel": "all five msm_accumulate launches of a step (overlapped with other streams)",
                                        "achieved": int_ops / (acc_total_ms * 1e-3) / 1e12 if acc_total_ms else None,
                                        "frac": (int_ops / (acc_total_ms * 1e-3)) / imad_peak if acc_total_ms else None,
                                        "sum_of_event_ms": acc_total_ms, "algorithmic_bytes": acc_alg_bytes,
                                        "traffic": traffic.get("msm_accumulate", {}).get("bytes_per_step")},
            "roofline_ntt": {"kernel": "ntt_pass_kernel x passes + r1cs/join (H pipeline)", "bound": "hbm",
                             "achieved": ntt_gbs, "peak": hbm_peak, "unit": "GB/s",
                             "frac": ntt_gbs / hbm_peak if hbm_peak else None,
                             "traffic": (traffic["ntt_pass"]["bytes_per_launch"] * traffic["ntt_pass"]["launches_per_step"])
                             if "ntt_pass" in traffic else None,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if hbm_peak else "unavailable"},
            "verified": ok,
        }
    zk.close()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args.workload, budget_s=args.cpu_budget)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


# ---- CPU arm: the oracle port (oracle/c/oracle.c) on the host cores ---------------------------------------------
def sample_workload(name, per_step_budget_s):
    """A bounded sample of `name` for the CPU: the workload itself if one prove fits the per-step budget, else a
    SYNTH-1 circuit of the same shape (same wire classes, same coefficients per constraint) with fewer constraints.
    The metric is a rate (constraints/s), so the sample's rate is reported as is."""
    from oracle import cport
    nv, npub, ncons, _ = WORKLOADS[name]
    # calibrate on 2^17 constraints (about a second)
    cal = "cpu-cal-131000"
    WORKLOADS[cal] = (125000, npub, 131000, "calibration sample")
    zk, wt = fixture(cal)
    t = time.perf_counter()
    cport.prove_points(zk, wt, 1, 2)
    rate = 131000 / (time.perf_counter() - t)
    fit = int(rate * per_step_budget_s)
    if fit >= ncons:
        return name, "1 full prove of the workload per step"
    lg = max(17, min(fit, ncons).bit_length() - 1)
    sc = (1 << lg) - 72                       # just under a power of two, like the reference's circuits
    sv = int(sc * nv / ncons)
    sname = "%s-sample-%d" % (name, sc)
    WORKLOADS[sname] = (sv, npub, sc, "SYNTH-1 sample of %s: %d constraints, %d wires" % (name, sc, sv))
    return sname, "1 prove of a %d-constraint / %d-wire SYNTH-1 circuit of the same shape per step (workload has %d)" % (sc, sv, ncons)


def cpu_prove_loop(sname, steps, warmup):
    from oracle import cport
    zk, wt = fixture(sname)
    for _ in range(warmup):
        cport.prove_points(zk, wt, 1, 2)
    t = time.perf_counter()
    for _ in range(steps):
        pts = cport.prove_points(zk, wt, 1, 2)
    return time.perf_counter() - t, pts, cport.last_timings(), cport.threads()


def cpu_baseline(name, budget_s=25.0):
    """cpu_baseline leg of the default run: ~budget_s of CPU work on a bounded sample, rank 0, N=1."""
    sname, what = sample_workload(name, budget_s)
    dt, _, phases, threads = cpu_prove_loop(sname, 1, 0)
    return {"value": WORKLOADS[sname][2] / dt, "unit": "constraints/s", "cores": threads, "kind": "port",
            "sample": what, "seconds": dt, "phases_s": phases,
            "note": "oracle/c/oracle.c (C restatement of the snarkjs/rapidsnark algorithm, pthreads); rapidsnark and "
                    "snarkjs cannot be installed offline"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nv, npub, ncons, desc = WORKLOADS[args.workload]
    per_step = max(2.0, args.ref_budget / max(1, args.steps + args.warmup))
    sname, what = sample_workload(args.workload, per_step)
    dt, pts, phases, threads = cpu_prove_loop(sname, args.steps, args.warmup)
    sc = WORKLOADS[sname][2]
    v = sc * args.steps / dt
    print(json.dumps({
        "impl": "reference", "metric": "groth16_prove_constraints_per_s", "value": v, "unit": "constraints/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 limbs (256-bit Montgomery)",
        "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "n_vars": nv, "n_public": npub, "n_constraints": ncons},
        "cpu_baseline": {"value": v, "unit": "constraints/s", "cores": threads, "kind": "port", "sample": what,
                         "phases_s": phases},
        "e2e": {"value": v, "unit": "constraints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="replica", choices=["replica", "shard", "shard-rep"],
                    help="N > 1: replica = one independent proof per GPU (weak scaling, default); shard = one proof "
                         "over all GPUs (strong scaling): sections 5-9 split by base range, the A/B/C coset transforms on "
                         "ranks 0-2 with an all-to-all of their slices, one all-gather of partial sums; shard-rep = the "
                         "same with the transforms repeated on every rank (no slice exchange)")
    ap.add_argument("--msm-c", type=int, default=0, help="Pippenger window bits (0 = automatic)")
    ap.add_argument("--h-table", default="auto", choices=sorted(H_TABLE),
                    help="fixed-base window table for the H section (auto = on when it fits in a quarter of free HBM)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", dest="verify", action="store_false")
    ap.add_argument("--cpu-budget", type=float, default=25.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--ref-budget", type=float, default=150.0, help="seconds for the whole --impl reference run")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
