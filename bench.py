#!/usr/bin/env python3
"""bench.py -- Groth16 prove throughput of the B200 prover (and of the CPU reference arm) on one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one Groth16 prove (the thing scripts/g16_prove.sh:246-260 of the reference execs a prover for) of one
witness against a resident proving key.  Metric = constraints proven per second (BASELINE.json: "Groth16 prove s &
constraints/s per layer"); `ms_per_step` is the prove time, `msm_mpts_per_s` the MSM point rate inside it.
  value      : witness already in HBM (g16_witness_load done), timed g16_prove_loaded, CUDA-side work + host finalisation
  e2e        : the public call ZKey.prove(wtns bytes in pinned host memory) -> JSON texts in host memory, per step
  per_config : N = 1 only; the larger BASELINE configurations (l1-16sigs 2^24, c1-layer-three 2^25) measured the same
               way in the same run with fewer steps (--per-config none to skip)
N > 1        : one process per GPU, every rank proves its own independent batch proof (the reference's GNU-parallel
               batch loop, scripts/full_workflow.sh:552): weak scaling, no data-path collective; barrier + max over
               ranks.  The same run then measures ONE proof over all ranks (base-range shards, NCCL for slices and
               768-byte partials) and reports it under "strong".  --mode shard|shard-rep makes that the headline instead.
Inputs are synthetic (SYNTH-1 circuits shaped like the reference's layers, known-trapdoor setup): the reference commits
no .zkey/.wtns and none can be generated offline.  The reference arm (--impl reference) times the CPU oracle port
(oracle/c/oracle.c, all host threads) on the SAME workload, because neither rapidsnark nor snarkjs can be installed
here; it makes its fixture on the host cores too (oracle/c/synth1.h), so that arm loads none of the repository's CUDA
libraries and never touches the GPU.
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
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # before torch creates the CUDA context (see csrc/prover.cu)

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
DEFAULT_PER_CONFIG = "l1-16sigs,c1-layer-three"
CIRCUIT_SEED, TRAPDOOR_SEED = 0xB2000001, 0xB2000002
CACHE_DIR = os.environ.get("G16_BENCH_CACHE", "/tmp/g16_bench_cache")
# pinned blinding factors: the bit-exact self-check of every run, and the sharded modes (all ranks must agree)
R_PIN, S_PIN = 0x1234567 << 200, 0x7654321 << 190


def load_pkg():
    if "g16b200" in sys.modules:
        return sys.modules["g16b200"]
    spec = importlib.util.spec_from_file_location("g16b200", os.path.join(ROOT, "zk-proof-of-assets_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["g16b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def proof_object(pts):
    """closed-form (pi_a, pi_b, pi_c) affine ints -> the dict a rapidsnark-style proof.json parses to"""
    a, b, c = pts
    return {"pi_a": [str(a[0]), str(a[1]), "1"],
            "pi_b": [[str(b[0][0]), str(b[0][1])], [str(b[1][0]), str(b[1][1])], ["1", "0"]],
            "pi_c": [str(c[0]), str(c[1]), "1"], "protocol": "groth16"}


def fixture_paths(name):
    base = os.path.join(CACHE_DIR, name)
    return base + ".zkey", base + ".wtns", base + ".vkey.json", base + ".expected.json"


def fixture(name, maker="gpu", load=True):
    """(zkey bytes, wtns bytes) of a workload, generated once per box and cached under CACHE_DIR together with its
    verification key and the closed-form proof for (R_PIN, S_PIN).  maker = "gpu": fixed-base multiplications through
    libg16synth.so (bench support, not part of either timed arm); "cpu": oracle/c/synth1.h on the host cores (what the
    reference arm uses: no repository CUDA library, no GPU).  Both write the same bytes."""
    nv, npub, ncons, _ = WORKLOADS[name]
    os.makedirs(CACHE_DIR, exist_ok=True)
    zp, wp, vp, ep = fixture_paths(name)
    if not all(os.path.exists(p) for p in (zp, wp, vp, ep)):
        lock = zp + ".lock"
        try:
            fd = os.open(lock, os.O_CREAT | os.O_EXCL | os.O_WRONLY)
        except FileExistsError:
            fd = None
        if fd is not None:
            try:
                if maker == "cpu":
                    from oracle import cport
                    sy = cport.Synth(nv, npub, ncons, CIRCUIT_SEED, TRAPDOOR_SEED)
                else:
                    sy = load_pkg().Synth(nv, npub, ncons, CIRCUIT_SEED, TRAPDOOR_SEED)
                zk, wt = sy.write()
                vk = json.dumps(sy.vkey()).encode()
                ex = json.dumps({"r": str(R_PIN), "s": str(S_PIN), "maker": maker,
                                 "proof": proof_object(sy.expected(R_PIN, S_PIN))}).encode()
                if hasattr(sy, "close"):
                    sy.close()
                for path, data in ((vp, vk), (ep, ex), (wp, wt), (zp, zk)):
                    with open(path + ".tmp", "wb") as f:
                        f.write(data)
                    os.replace(path + ".tmp", path)
            finally:
                os.close(fd)
                os.unlink(lock)
        else:
            while os.path.exists(lock):
                time.sleep(0.5)
    if not load:
        return None, None
    with open(zp, "rb") as f:
        zk = f.read()
    with open(wp, "rb") as f:
        wt = f.read()
    return zk, wt


def fixture_expected(name):
    with open(fixture_paths(name)[3]) as f:
        return json.load(f)


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


# ---- work model -------------------------------------------------------
MODMUL_OPS = 136          # 32x32->64 multiply-adds per 254-bit Montgomery product (128 wide + 8 quotient digits)
G1_MADD_MODMULS = 10      # XYZZ mixed addition: 8M + 2S over Fq
G2_MADD_MODMULS = 28      # the same over Fq2: 8 x 3 (Karatsuba) + 2 x 2
G1_ADD_MODMULS = 14       # full XYZZ addition (bucket reduction)
G2_ADD_MODMULS = 42
NTT_BYTES_PER_POINT = 384  # H pipeline, strict: 6 transforms x (32 B read + 32 B write) per domain point

H_TABLE = {"auto": 0, "on": 1, "off": -1}


def step_work(n, logn, n_coefs, stats):
    """Algorithmic modmuls of one prove, by phase: accumulations = one mixed addition per non-zero
    digit; transforms = 6 x (n/2) log2 n butterflies + 3n coset factors; reductions = 2 full additions per bucket;
    r1cs + join = one product per coefficient, n for C = A.B, 2n for the join and de-Montgomery."""
    acc = sum(G1_MADD_MODMULS * stats[k]["additions"] for k in ("A", "B1", "C", "H")) + G2_MADD_MODMULS * stats["B2"]["additions"]
    ntt = 6 * (n // 2) * logn + 3 * n
    red = 0
    for k in ("A", "B1", "B2", "C", "H"):
        windows = 1 if stats[k].get("tabled") else stats[k]["W"]
        buckets = windows << (stats[k]["c"] - 1)
        red += 2 * buckets * (G2_ADD_MODMULS if k == "B2" else G1_ADD_MODMULS)
    other = n_coefs + n + 2 * n
    return {"accumulate": acc, "transforms": ntt, "reductions": red, "r1cs_join": other}


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


def dist_setup():
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


class Ctx:
    """what every measurement needs: torch, the package, rank layout, collectives"""

    def __init__(self, args):
        import torch
        self.torch = torch
        self.rank, self.world, self.local, self.dist = dist_setup()
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the prover has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.pkg = load_pkg()
        self.args = args
        self._peak = None

    def sync_all(self):
        self.torch.cuda.synchronize(self.dev)
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, x):
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def pinned(self, data):
        t = self.torch.empty(len(data), dtype=self.torch.uint8).pin_memory()
        t.copy_(self.torch.frombuffer(bytearray(data), dtype=self.torch.uint8))
        return t

    def int_peak(self):
        """wide multiply-adds per second of this GPU, measured now: the best of the carry-chain microbenchmarks (one
        chain per thread, four independent chains per thread) and of dependent Montgomery products x 136"""
        if self._peak is None:
            one = self.pkg.bench_int_pipe(1)
            four = self.pkg.bench_int_pipe(3)
            modmul = self.pkg.bench_int_pipe(2)
            self._peak = {"peak": max(one, four, modmul * MODMUL_OPS), "one_chain": one, "four_chains": four,
                          "modmul_per_s": modmul, "modmul_x136": modmul * MODMUL_OPS}
        return self._peak

    def fixture(self, name):
        if self.rank == 0:
            fixture(name, load=False)          # one rank generates, the others read the cache
        if self.dist:
            self.dist.barrier()
        return fixture(name)


def timed_steps(cx, fn, steps):
    """K calls of fn bracketed by barrier + synchronize; device events and host clock, max over ranks -> seconds"""
    torch = cx.torch
    cx.sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    w0 = time.perf_counter()
    out = None
    for _ in range(steps):
        out = fn()
    e1.record()
    torch.cuda.synchronize(cx.dev)
    wall = time.perf_counter() - w0
    return cx.max_over_ranks(max(e0.elapsed_time(e1) / 1e3, wall)), out


def self_check(cx, zk, name, pinned, shard_step=None):
    """untimed: (i) the proof under pinned r,s equals the closed-form known-trapdoor proof cached with the fixture, bit
    for bit; (ii) a proof under fresh randomness passes the native pairing verifier"""
    pkg = cx.pkg
    exp = fixture_expected(name)
    if shard_step:
        proof, public = shard_step()
    else:
        proof, public = zk.prove(pinned, r=R_PIN, s=S_PIN)
    bit_exact = json.loads(proof) == exp["proof"]
    ok = None
    if cx.args.verify and not shard_step:
        p2, u2 = zk.prove(pinned)
        with open(fixture_paths(name)[2]) as f:
            ok = bool(pkg.verify(f.read(), u2, p2))
    return bit_exact, ok, exp.get("maker")


def measure(cx, name, steps, warmup, full=True):
    """resident + end-to-end arms of one workload on every rank (replicas) -> result dict on rank 0"""
    torch, pkg, args = cx.torch, cx.pkg, cx.args
    nv, npub, ncons, desc = WORKLOADS[name]
    zkey, wtns = cx.fixture(name)
    t_open = time.perf_counter()
    zk = pkg.ZKey(zkey, device=cx.local, msm_c=args.msm_c, h_table=H_TABLE[args.h_table])
    t_open = time.perf_counter() - t_open
    zkey_len, wtns_len = len(zkey), len(wtns)
    del zkey
    pinned = cx.pinned(wtns)
    del wtns

    # ---- resident arm: witness in HBM before the timed region ---------------------------------------------------
    zk.load_witness(pinned)
    for _ in range(warmup):
        zk.prove_loaded()
    clocks = ClockSampler(cx.local)
    phase = {}
    acc_ms = {k: 0.0 for k in ("A", "B1", "B2", "C", "H")}
    if cx.rank == 0:
        clocks.start()
    launches0 = pkg.launch_count()

    def step_loaded():
        out = zk.prove_loaded()
        for k, v in zk.timings().items():
            phase[k] = phase.get(k, 0.0) + v
        for k, v in zk.accumulate_ms().items():
            acc_ms[k] += v
        return out

    resident_s, _ = timed_steps(cx, step_loaded, steps)
    launches = pkg.launch_count() - launches0
    clk = clocks.stop() if cx.rank == 0 else None
    stats = zk.msm_stats()
    trace = zk.trace()

    # ---- end-to-end arm: pinned host bytes in, JSON texts out, every step -----------------------------------------
    for _ in range(max(1, warmup // 2)):
        zk.prove(pinned)
    e2e_s, (proof, public) = timed_steps(cx, lambda: zk.prove(pinned), steps)
    d2h = len(proof) + len(public)

    # ---- the same through ZKey.prove_stream: witness i+1 is copied while proof i computes ---------------------------
    pipe_s = None
    if full:
        for _ in zk.prove_stream([pinned] * max(2, warmup // 2)):
            pass

        def stream_all():
            for out in zk.prove_stream([pinned] * steps):
                pass
            return out
        pipe_s, _ = timed_steps(cx, stream_all, 1)

    # ---- the H pipeline (r1cs + 6 transforms + join) timed alone, for its own roofline -----------------------------
    ntt_alone_ms = None
    if cx.rank == 0:
        zk.load_witness(pinned)
        zk.h_pipeline_dev()
        t = []
        for _ in range(5):
            zk.h_pipeline_dev()
            t.append(zk.timings()["ntt_pipeline"])
        ntt_alone_ms = sorted(t)[len(t) // 2]

    bit_exact = ok = maker = None
    if cx.rank == 0:
        bit_exact, ok, maker = self_check(cx, zk, name, pinned)

    out = None
    if cx.rank == 0:
        n, logn = zk.domain_size, zk.domain_size.bit_length() - 1
        plain_c = max(6, min(20, logn - 5))                                # msm_pick_c (csrc/msm.cuh)
        h_tabled = not args.msm_c and stats["H"]["c"] > plain_c
        stats["H"]["tabled"] = h_tabled
        ms_step = resident_s / steps * 1e3
        peak = cx.int_peak()
        imad_peak = peak["peak"]
        work = step_work(n, logn, zk.n_coefs, stats)
        total_ops = MODMUL_OPS * sum(work.values())
        h_ms = acc_ms["H"] / steps
        h_ops = MODMUL_OPS * G1_MADD_MODMULS * stats["H"]["additions"]
        acc_total_ms = sum(acc_ms.values()) / steps
        msm_points = 3 * nv - npub - 1 + n + nv
        msm_ms = sum(phase[k] for k in ("msm_A", "msm_B1", "msm_B2", "msm_C", "msm_H")) / steps
        traffic = measured_traffic(name)
        if not h_tabled and "msm_accumulate_H_plain" in traffic:
            traffic = dict(traffic, msm_accumulate_H=traffic["msm_accumulate_H_plain"])
        hbm_peak = measured_peaks().get("hbm_gbs")
        ntt_ops = MODMUL_OPS * work["transforms"]
        ntt_in_step_ms = phase["ntt_pipeline"] / steps
        out = {
            "metric": "groth16_prove_constraints_per_s", "value": ncons * steps * cx.world / resident_s,
            "unit": "constraints/s", "n_gpus": cx.world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 limbs (256-bit Montgomery Fr/Fq/Fq2)", "data": "synthetic",
            "config": {"workload": name, "description": desc, "n_vars": nv, "n_public": npub,
                       "n_constraints": ncons, "domain_size": n, "n_coefs": zk.n_coefs, "h_window_table": h_tabled,
                       "zkey_bytes": zkey_len, "proofs_per_step": cx.world, "parallelism": "replica-per-gpu",
                       "l2": "inputs exceed L2 (bases + scalars > 1 GB touched per step); no flush",
                       "timing": "cuda events on the default stream around synchronous prove calls, max(event, host clock), max over ranks"},
            "prove_s": resident_s / steps,
            "msm_mpts_per_s": msm_points / (msm_ms * 1e-3) / 1e6 if msm_ms else None,
            "phases_ms": {k: v / steps for k, v in phase.items()},
            "msm": {k: dict(v, accumulate_ms=acc_ms[k] / steps) for k, v in stats.items()},
            "trace_last_step_ms": trace,
            "zkey_open_s": t_open,
            "clocks": clk,
            "e2e_pipelined": ({"value": ncons * steps * cx.world / pipe_s, "unit": "constraints/s",
                               "ms_per_step": pipe_s / steps * 1e3, "h2d_bytes_per_step": wtns_len,
                               "api": "ZKey.prove_stream: witness i+1 copied while proof i computes"} if pipe_s else None),
            "e2e": {"value": ncons * steps * cx.world / e2e_s, "unit": "constraints/s", "ms_per_step": e2e_s / steps * 1e3,
                    "h2d_bytes_per_step": wtns_len, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            # the dominant kernel = the bucket accumulation of the H MSM (one launch per step, the largest MSM at every
            # config); CUDA events on its own stream around that launch
            "roofline": {"kernel": "msm_accumulate_kernel<G1>, H MSM launch (bucket accumulation; dominant kernel)",
                         "bound": "int", "achieved": h_ops / (h_ms * 1e-3) / 1e12 if h_ms else None,
                         "peak": imad_peak / 1e12, "unit": "T wide-IMAD/s",
                         "frac": (h_ops / (h_ms * 1e-3)) / imad_peak if h_ms else None,
                         "units_per_launch": stats["H"]["additions"], "ops_per_unit": MODMUL_OPS * G1_MADD_MODMULS,
                         "ms_per_launch": h_ms,
                         "peak_source": "max of g16_bench_int_pipe ops 1 (one carry chain per thread), 3 (four independent "
                                        "chains) and 2 x 136 (dependent Montgomery products), measured in this run "
                                        "(MEASURED_PEAKS.json has no integer figure)",
                         "peak_candidates_T": {k: v / 1e12 for k, v in peak.items() if k not in ("peak", "modmul_per_s")},
                         "modmul_peak_per_s": peak["modmul_per_s"], "share_of_step": h_ms / ms_step,
                         "algorithmic_bytes": 68 * stats["H"]["additions"],
                         "traffic": traffic.get("msm_accumulate_H", {}).get("bytes_per_launch"),
                         "traffic_source": traffic.get("msm_accumulate_H", {}).get("source")},
            # every algorithmic wide multiply-add of the step over the step time: how well the whole proof uses the pipe
            "roofline_step": {"bound": "int", "achieved": total_ops / (ms_step * 1e-3) / 1e12, "peak": imad_peak / 1e12,
                              "unit": "T wide-IMAD/s", "frac": total_ops / (ms_step * 1e-3) / imad_peak,
                              "modmuls": work, "ops_per_modmul": MODMUL_OPS},
            "roofline_all_accumulate": {"kernel": "all five msm_accumulate launches of a step (overlapped with other streams)",
                                        "achieved": MODMUL_OPS * work["accumulate"] / (acc_total_ms * 1e-3) / 1e12 if acc_total_ms else None,
                                        "frac": MODMUL_OPS * work["accumulate"] / (acc_total_ms * 1e-3) / imad_peak if acc_total_ms else None,
                                        "sum_of_event_ms": acc_total_ms,
                                        "algorithmic_bytes": 68 * sum(stats[k]["additions"] for k in ("A", "B1", "C", "H")) + 132 * stats["B2"]["additions"],
                                        "traffic": traffic.get("msm_accumulate", {}).get("bytes_per_step")},
            # the six transforms: integer-bound at 256 bits (7 wide multiply-adds per byte against a machine balance of
            # 1.4), so the integer peak is the yardstick; the HBM figure SURVEY 8(d) asked for is kept beside it
            "roofline_ntt": {"kernel": "ntt_pass_kernel x passes (6 transforms of the H pipeline, timed alone)", "bound": "int",
                             "achieved": ntt_ops / (ntt_alone_ms * 1e-3) / 1e12 if ntt_alone_ms else None,
                             "peak": imad_peak / 1e12, "unit": "T wide-IMAD/s",
                             "frac": ntt_ops / (ntt_alone_ms * 1e-3) / imad_peak if ntt_alone_ms else None,
                             "ms_alone": ntt_alone_ms, "ms_in_step": ntt_in_step_ms,
                             "hbm": {"achieved_gbs": NTT_BYTES_PER_POINT * n / (ntt_alone_ms * 1e-3) / 1e9 if ntt_alone_ms else None,
                                     "peak_gbs": hbm_peak,
                                     "frac": NTT_BYTES_PER_POINT * n / (ntt_alone_ms * 1e-3) / 1e9 / hbm_peak if hbm_peak and ntt_alone_ms else None,
                                     "traffic": (traffic["ntt_pass"]["bytes_per_launch"] * traffic["ntt_pass"]["launches_per_step"])
                                     if "ntt_pass" in traffic else None}},
            "bit_exact": bit_exact, "bit_exact_against": "closed-form known-trapdoor proof for pinned r,s (fixture made by the %s generator)" % maker,
            "verified": ok,
        }
    zk.close()
    del zk
    torch.cuda.empty_cache()
    return out, pinned


def measure_strong(cx, name, steps, warmup, mode):
    """ONE proof over all ranks: base-range shards of sections 5-9; mode "shard" spreads the coset transforms over the
    ranks too (all-to-all of slices), "shard-rep" repeats them on every rank; 768-byte partials all-gathered."""
    pkg, args = cx.pkg, cx.args
    nv, npub, ncons, desc = WORKLOADS[name]
    zkey, wtns = cx.fixture(name)
    zk = pkg.ZKey(zkey, device=cx.local, msm_c=args.msm_c, shard_rank=cx.rank, shard_count=cx.world, h_table=H_TABLE[args.h_table])
    del zkey
    pinned = cx.pinned(wtns)
    dist_ntt = mode == "shard"

    def step_loaded():
        if dist_ntt:
            return pkg.prove_sharded_distributed(zk, None, r=R_PIN, s=S_PIN)
        return zk.combine(pkg.all_gather_partials(zk.prove_partial()), r=R_PIN, s=S_PIN)

    def step_e2e():
        if dist_ntt:
            return pkg.prove_sharded_distributed(zk, pinned, r=R_PIN, s=S_PIN)
        return pkg.prove_sharded(zk, pinned, r=R_PIN, s=S_PIN)

    zk.load_witness(pinned)
    for _ in range(warmup):
        step_loaded()
    phase = {}

    def timed():
        out = step_loaded()
        for k, v in zk.timings().items():
            phase[k] = phase.get(k, 0.0) + v
        return out
    resident_s, _ = timed_steps(cx, timed, steps)
    for _ in range(max(1, warmup // 2)):
        step_e2e()
    e2e_s, (proof, public) = timed_steps(cx, step_e2e, steps)
    out = None
    if cx.rank == 0:
        bit_exact = json.loads(proof) == fixture_expected(name)["proof"]
        ph = {k: v / steps for k, v in phase.items()}
        bound = max((("transform", ph.get("r1cs_eval", 0) + ph.get("ntt_pipeline", 0)), ("H accumulate", ph.get("msm_H", 0)),
                     ("witness MSMs", max(ph.get(k, 0) for k in ("msm_A", "msm_B1", "msm_B2", "msm_C")))), key=lambda kv: kv[1])
        out = {"mode": mode, "workload": name, "n_gpus": cx.world, "scaling": "strong", "ms_per_proof": resident_s / steps * 1e3,
               "value": ncons * steps / resident_s, "unit": "constraints/s",
               "e2e_ms_per_proof": e2e_s / steps * 1e3, "phases_ms_rank0": ph, "bounded_by_rank0": bound[0], "bit_exact": bit_exact,
               "parallelism": "base-range shards x%d, %s, NCCL all-gather of 768 B partials"
                              % (cx.world, "coset transforms spread over the ranks + exchange of slices" if dist_ntt else "transforms replicated")}
    zk.close()
    return out


def run_b200(args):
    cx = Ctx(args)
    name = args.workload
    if args.mode != "replica" and cx.world > 1:
        out = measure_strong(cx, name, args.steps, args.warmup, args.mode)
        if cx.rank == 0:
            nv, npub, ncons, desc = WORKLOADS[name]
            line = {"metric": "groth16_prove_constraints_per_s", "value": out["value"], "unit": "constraints/s", "n_gpus": cx.world,
                    "steps": args.steps, "warmup": args.warmup, "ms_per_step": out["ms_per_proof"], "higher_is_better": True,
                    "scaling": "strong", "vs_baseline": None, "dtype": "u32 limbs (256-bit Montgomery Fr/Fq/Fq2)", "data": "synthetic",
                    "config": {"workload": name, "description": desc, "n_vars": nv, "n_public": npub, "n_constraints": ncons,
                               "parallelism": out["parallelism"]}, "strong": out}
            print(json.dumps(line), flush=True)
    else:
        out, pinned = measure(cx, name, args.steps, args.warmup)
        if cx.world > 1 and args.strong:
            # the same workload as ONE proof over all ranks, both transform placements
            strong = [measure_strong(cx, name, max(3, args.steps // 2), 3, m) for m in ("shard-rep", "shard")]
            if cx.rank == 0:
                best = min(strong, key=lambda s: s["ms_per_proof"])
                out["strong"] = dict(best, speedup_vs_one_gpu_replica=out["ms_per_step"] / best["ms_per_proof"], modes=strong)
        if cx.rank == 0 and cx.world == 1:
            if not args.no_cpu_baseline:
                out["cpu_baseline"] = cpu_baseline(name, budget_s=args.cpu_budget)
            per = [w for w in args.per_config.split(",") if w and w != "none" and w != name]
            out["per_config"] = []
            for w in per:
                t0 = time.perf_counter()
                try:
                    r, _ = measure(cx, w, max(3, args.steps // 3), 3, full=False)
                    keep = ("value", "unit", "ms_per_step", "steps", "config", "e2e", "roofline", "roofline_step", "roofline_ntt",
                            "phases_ms", "trace_last_step_ms", "bit_exact", "verified", "zkey_open_s", "msm_mpts_per_s", "gpu_launches")
                    r = {k: r[k] for k in keep}
                    r["wall_s_including_fixture"] = time.perf_counter() - t0
                    out["per_config"].append(r)
                except Exception as e:           # a config that does not fit this box is reported, not fatal
                    out["per_config"].append({"config": {"workload": w}, "error": repr(e)})
        if cx.rank == 0:
            print(json.dumps(out), flush=True)
    if cx.dist:
        cx.dist.barrier()
        cx.dist.destroy_process_group()


# ---- CPU arm: the oracle port (oracle/c/oracle.c) on the host cores ---------------------------------------------
def sample_workload(name, fit_constraints):
    """A bounded sample of `name` for the CPU: a SYNTH-1 circuit of the same shape (same wire classes, same coefficients
    per constraint) with about fit_constraints constraints.  The metric is a rate (constraints/s)."""
    nv, npub, ncons, _ = WORKLOADS[name]
    lg = max(17, min(fit_constraints, ncons).bit_length() - 1)
    sc = (1 << lg) - 72                       # just under a power of two, like the reference's circuits
    sv = int(sc * nv / ncons)
    sname = "%s-sample-%d" % (name, sc)
    WORKLOADS[sname] = (sv, npub, sc, "SYNTH-1 sample of %s: %d constraints, %d wires" % (name, sc, sv))
    return sname, "1 prove of a %d-constraint / %d-wire SYNTH-1 circuit of the same shape per step (workload has %d)" % (sc, sv, ncons)


def cpu_plan(name, steps, warmup, budget_s, maker):
    """-> (workload to prove, description, calibrated seconds per prove, fixture bytes).  The full workload whenever
    (steps + warmup) proves of it fit the budget; calibrated on the SECOND prove of a 2^17 circuit (the first call of
    a process may build the library)."""
    from oracle import cport
    cport.lib()
    nv, npub, ncons, _ = WORKLOADS[name]
    cal = "cpu-cal-131000"
    WORKLOADS[cal] = (125000, npub, 131000, "calibration sample")
    zk, wt = fixture(cal, maker)
    cport.prove_points(zk, wt, 1, 2)
    t = time.perf_counter()
    cport.prove_points(zk, wt, 1, 2)
    rate = 131000 / (time.perf_counter() - t)
    # CPU Pippenger gets faster per constraint as n grows, so this estimate errs towards the full workload
    if ncons / rate * (steps + warmup) <= budget_s:
        return name, "1 full prove of the workload per step (%d constraints)" % ncons
    fit = int(rate * budget_s / max(1, steps + warmup))
    return sample_workload(name, fit)


def cpu_prove_loop(sname, steps, warmup, maker):
    from oracle import cport
    zk, wt = fixture(sname, maker)
    pts = None
    for _ in range(warmup):
        pts = cport.prove_points(zk, wt, R_PIN, S_PIN)
    t = time.perf_counter()
    for _ in range(steps):
        pts = cport.prove_points(zk, wt, R_PIN, S_PIN)
    dt = time.perf_counter() - t
    exp = fixture_expected(sname)["proof"]
    return dt, proof_object(pts) == exp, cport.last_timings(), cport.threads()


def cpu_baseline(name, budget_s=30.0):
    """cpu_baseline leg of the default run: the oracle port on the host cores, rank 0, N=1: one prove of the full
    workload when it fits the budget (c2-layer-one: about 3 s on 16 cores), else of a bounded sample of it."""
    sname, what = cpu_plan(name, 2, 0, budget_s, "gpu")
    dt, exact, phases, threads = cpu_prove_loop(sname, 1, 1 if sname == name else 0, "gpu")
    return {"value": WORKLOADS[sname][2] / dt, "unit": "constraints/s", "cores": threads, "kind": "port",
            "sample": what, "seconds": dt, "phases_s": phases, "bit_exact": exact,
            "note": "oracle/c/oracle.c (C restatement of the snarkjs/rapidsnark algorithm, pthreads, plain unsigned __int128 field "
                    "arithmetic); rapidsnark and snarkjs cannot be installed offline, so every GPU/CPU ratio here is against "
                    "this port, not against the reference's own provers"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nv, npub, ncons, desc = WORKLOADS[args.workload]
    sname, what = cpu_plan(args.workload, args.steps, args.warmup, args.ref_budget, "cpu")
    dt, exact, phases, threads = cpu_prove_loop(sname, args.steps, args.warmup, "cpu")
    sc = WORKLOADS[sname][2]
    v = sc * args.steps / dt
    print(json.dumps({
        "impl": "reference", "metric": "groth16_prove_constraints_per_s", "value": v, "unit": "constraints/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 limbs (256-bit Montgomery)",
        "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "n_vars": nv, "n_public": npub, "n_constraints": ncons,
                   "same_workload_as_gpu_arm": sname == args.workload},
        "cpu_baseline": {"value": v, "unit": "constraints/s", "cores": threads, "kind": "port", "sample": what,
                         "phases_s": phases, "bit_exact": exact,
                         "note": "oracle/c/oracle.c on all host threads; fixture made on the host cores by oracle/c/synth1.h; "
                                 "no repository CUDA library is loaded and the GPU is not used in this arm"},
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
    ap.add_argument("--per-config", default=DEFAULT_PER_CONFIG,
                    help="N = 1: comma-separated larger workloads also measured in this run (fewer steps), or 'none'")
    ap.add_argument("--mode", default="replica", choices=["replica", "shard", "shard-rep"],
                    help="N > 1: replica = one independent proof per GPU (weak scaling, default; the run also reports ONE "
                         "proof over all GPUs under 'strong'); shard / shard-rep = make the one-proof-over-all-GPUs "
                         "measurement the headline (coset transforms spread over the ranks / repeated on every rank)")
    ap.add_argument("--no-strong", dest="strong", action="store_false", help="N > 1: skip the strong-scaling block")
    ap.add_argument("--msm-c", type=int, default=0, help="Pippenger window bits (0 = automatic)")
    ap.add_argument("--h-table", default="auto", choices=sorted(H_TABLE),
                    help="fixed-base window table for the H section (auto = on when it fits in a quarter of free HBM)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", dest="verify", action="store_false")
    ap.add_argument("--cpu-budget", type=float, default=30.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--ref-budget", type=float, default=300.0, help="seconds for the whole --impl reference run")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
