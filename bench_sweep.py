#!/usr/bin/env python3
"""bench_sweep.py -- BASELINE.json configs[4]: standalone BN254 G1/G2 MSM and coset-NTT sweep, 2^20 .. 2^26.

    python bench_sweep.py [--min-log 20] [--max-log 26] [--max-log-g2 24] [--iters 5] [--out profiles/rNN_sweep.json]

Per size and group: bases P_i = k_i * G (k_i < 2^31, made on the GPU by libg16synth's fixed-base tables, bench support),
scalars (a) uniform in [0, r) and (b) witness-like (55 % in {0,1}, 35 % < 2^64, 10 % full width -- the SYNTH-1 mix).
Every timed MSM result is checked against the closed form (sum k_i s_i) * G computed on the host with integer
arithmetic, so a number is only printed for a correct result.  Timing: device-resident bases and scalars, whole
operator call (digits + sort + accumulate + reductions + host Horner + affine result in host memory), median of
`iters` runs after 3 warm-ups, host clock around the synchronous C-ABI call.
NTT line: to_odd_coset (iNTT -> x inc^i -> NTT, rows a5-a7) on three n-point polynomials, as one proof does.
One JSON line per measurement; a summary table goes to stderr.
"""
import argparse
import ctypes
import importlib.util
import json
import os
import statistics
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
MODMUL_OPS, G1_MADD, G2_MADD = 136, 10, 28


def load_pkg():
    spec = importlib.util.spec_from_file_location("g16b200", os.path.join(ROOT, "zk-proof-of-assets_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["g16b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def make_scalars(n, kind, rng):
    """n x 4 uint64 limbs (LE), standard form, < r"""
    s = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64)
    s[:, 3] &= np.uint64((1 << 60) - 1)          # < 2^252 < r
    if kind == "witness":
        u = rng.random(n)
        bits = u < 0.55
        small = (u >= 0.55) & (u < 0.90)
        s[bits, 1:] = 0
        s[bits, 0] &= np.uint64(1)
        s[small, 1:] = 0
    return s


def dot_mod_r(k, s):
    """sum k_i * s_i mod r with k < 2^31 (uint64 array) and s = n x 4 uint64 limbs: 16-bit pieces, 2^16-row chunks"""
    n = len(k)
    h = s.view(np.uint16).reshape(n, 16)
    tot = [0] * 16
    for lo in range(0, n, 1 << 16):
        part = (h[lo:lo + (1 << 16)].astype(np.uint64) * k[lo:lo + (1 << 16), None]).sum(axis=0)
        for j in range(16):
            tot[j] += int(part[j])
    return sum(t << (16 * j) for j, t in enumerate(tot)) % R


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-log", type=int, default=20)
    ap.add_argument("--max-log", type=int, default=26)
    ap.add_argument("--max-log-g2", type=int, default=24)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--ntt-only", action="store_true", help="skip the MSM sweeps")
    ap.add_argument("--apply-key", action="store_true", help="with --ntt-only: still time the batch apply-key operator")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench_sweep.py: no CUDA device; the operators have no CPU fallback")
    pkg = load_pkg()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(0xB2000001)
    lines = []

    def emit(d):
        lines.append(d)
        print(json.dumps(d), flush=True)

    # the integer peak bench.py uses: the best of one carry chain, four independent chains, dependent Montgomery products x 136
    imad_peak = max(pkg.bench_int_pipe(1), pkg.bench_int_pipe(3), pkg.bench_int_pipe(2) * MODMUL_OPS)
    for is_g2 in (() if args.ntt_only else (False, True)):
        top = min(args.max_log, args.max_log_g2) if is_g2 else args.max_log
        if top < args.min_log:
            continue
        nmax = 1 << top
        rec = 128 if is_g2 else 64
        k = rng.integers(1, 1 << 31, size=nmax, dtype=np.uint64)
        d_bases = torch.empty(nmax * rec, dtype=torch.uint8, device=dev)
        t0 = time.perf_counter()
        step = 1 << 21
        for lo in range(0, nmax, step):           # bases in slices: bounded host copies
            kb = np.zeros((min(step, nmax - lo), 4), dtype=np.uint64)
            kb[:, 0] = k[lo:lo + step]
            pts = pkg.fixed_base(is_g2, kb.tobytes(), len(kb))
            d_bases[lo * rec:(lo + len(kb)) * rec].copy_(torch.frombuffer(bytearray(pts), dtype=torch.uint8))
        gen_s = time.perf_counter() - t0
        for lg in range(args.min_log, top + 1):
            n = 1 << lg
            msm = pkg.Msm(is_g2, n)
            for kind in ("uniform", "witness"):
                s = make_scalars(n, kind, rng)
                d_s = torch.from_numpy(s.view(np.uint8).reshape(-1)).to(dev)
                # closed form through the fixed-base tables (independent of the MSM code): one record e * G
                want = pkg.fixed_base(is_g2, dot_mod_r(k[:n], s).to_bytes(32, "little"), 1)
                ts = []
                for it in range(3 + args.iters):
                    torch.cuda.synchronize(dev)
                    t = time.perf_counter()
                    got = msm.run(d_bases.data_ptr(), d_s.data_ptr(), n)
                    dt = time.perf_counter() - t
                    if it >= 3:
                        ts.append(dt)
                ok = got == want
                st = msm.stats()
                med = statistics.median(ts)
                ops = MODMUL_OPS * (G2_MADD if is_g2 else G1_MADD) * st["additions"]
                emit({"op": "msm_g2" if is_g2 else "msm_g1", "log_n": lg, "n": n, "scalars": kind, "ms": med * 1e3,
                      "mpts_per_s": n / med / 1e6, "additions": st["additions"], "c": st["c"], "W": st["W"],
                      "tasks": st["tasks"], "imad_T_per_s_whole_call": ops / med / 1e12,
                      "frac_of_imad_peak_whole_call": ops / med / imad_peak, "correct": bool(ok)})
                if not ok:
                    raise SystemExit("MSM result differs from the closed form at 2^%d (%s)" % (lg, kind))
                del d_s
            msm.close()
        del d_bases
        torch.cuda.empty_cache()
        print("# %s bases for 2^%d made in %.1f s" % ("G2" if is_g2 else "G1", top, gen_s), file=sys.stderr)

    hbm = None
    try:
        hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
    except OSError:
        pass
    modmul_peak = pkg.bench_int_pipe(2)
    for lg in range(args.min_log, args.max_log + 1):
        n = 1 << lg
        a = torch.from_numpy(make_scalars(3 * n, "uniform", rng).view(np.uint8).reshape(-1)).to(dev)
        ts = []
        for it in range(3 + args.iters):
            torch.cuda.synchronize(dev)
            t = time.perf_counter()
            pkg.coset_ntt_dev(a.data_ptr(), lg, 3)
            dt = time.perf_counter() - t
            if it >= 3:
                ts.append(dt)
        med = statistics.median(ts)
        modmuls = 3 * (2 * (n // 2) * lg + n)      # two transforms of (n/2) log n butterflies + the coset/scale factor
        emit({"op": "coset_ntt_x3", "log_n": lg, "n": n, "ms": med * 1e3, "mpts_per_s": 3 * n / med / 1e6,
              "strict_bytes": 3 * 128 * n, "gb_per_s_strict": 3 * 128 * n / med / 1e9,
              "frac_of_hbm_peak_strict": (3 * 128 * n / med / 1e9 / hbm) if hbm else None,
              "modmul_per_s": modmuls / med, "frac_of_modmul_peak": modmuls / med / modmul_peak})
        del a
    # batch apply key (operator level): out[i] = (first * inc^i) * P_i on device-resident records
    if not args.ntt_only or args.apply_key:
        for is_g2, lg in ((False, min(args.max_log, 21)), (True, min(args.max_log, 19))):
            n = 1 << lg
            rec = 128 if is_g2 else 64
            kb = np.zeros((n, 4), dtype=np.uint64)
            kb[:, 0] = rng.integers(1, 1 << 31, size=n, dtype=np.uint64)
            pts = pkg.fixed_base(is_g2, kb.tobytes(), n)
            d_p = torch.frombuffer(bytearray(pts), dtype=torch.uint8).to(dev)
            d_o = torch.empty_like(d_p)
            first = int(rng.integers(1, 1 << 62)) << 190 | 12345
            ts = []
            for it in range(1 + 3):
                torch.cuda.synchronize(dev)
                t = time.perf_counter()
                pkg.batch_apply_key_dev(is_g2, d_p.data_ptr(), n, first, 1, d_o.data_ptr())
                dt = time.perf_counter() - t
                if it >= 1:
                    ts.append(dt)
            med = statistics.median(ts)
            # spot check against the fixed-base generator: (first * k_0) * G
            want = pkg.fixed_base(is_g2, (first * int(kb[0, 0]) % R).to_bytes(32, "little"), 1)
            ok = bytes(d_o[:rec].cpu().numpy().tobytes()) == want
            popcnt = bin(first % R).count("1")
            mm = (254 * 9 + popcnt * 10 + 400) * (3 if is_g2 else 1)
            emit({"op": "apply_key_g2" if is_g2 else "apply_key_g1", "log_n": lg, "n": n, "ms": med * 1e3,
                  "mpts_per_s": n / med / 1e6, "modmul_per_point_model": mm, "modmul_per_s": mm * n / med,
                  "frac_of_modmul_peak": mm * n / med / modmul_peak, "correct": bool(ok)})
            if not ok:
                raise SystemExit("apply key result differs from the closed form")
            del d_p, d_o
    print("\nop | log n | scalars | ms | Mpts/s | frac of int peak", file=sys.stderr)
    for d in lines:
        print("%s | %d | %s | %.3f | %.1f | %.2f" % (d["op"], d["log_n"], d.get("scalars", "-"), d["ms"], d["mpts_per_s"],
                                                  d.get("frac_of_imad_peak_whole_call", d.get("frac_of_modmul_peak"))),
              file=sys.stderr)
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"imad_peak_per_s": imad_peak, "modmul_peak_per_s": modmul_peak, "hbm_gbs": hbm, "lines": lines}, f, indent=1)


if __name__ == "__main__":
    main()
