"""Parity of the CUDA path against the oracle, through the C ABI (bit-exact: all integer work)."""
import json
import os
import random
import subprocess

import pytest

from conftest import ROOT
from oracle import binfile, bn254 as bn, factory, groth16 as g16, verify as V

pytestmark = pytest.mark.gpu

Q, R = bn.Q, bn.R
MQ, MR = bn.MONT_R_Q, bn.MONT_R_R


def le(x):
    return int(x).to_bytes(32, "little")


def ints(b):
    return [int.from_bytes(b[i:i + 32], "little") for i in range(0, len(b), 32)]


EDGE = lambda p: [0, 1, 2, p - 1, p - 2, (1 << 256) % p, (1 << 255) % p, (p + 1) // 2, 0xFFFFFFFF, 1 << 32, (1 << 64) - 1,
                  (1 << 224) - 1]


# ---- field (row a17) ------------------------------------------------------------------------------------
@pytest.mark.parametrize("field,p", [(0, Q), (1, R)])
def test_field_ops(pkg, field, p):
    rng = random.Random(1234 + field)
    e = EDGE(p)
    a = [x for x in e for _ in e] + [rng.randrange(p) for _ in range(4000)]
    b = [y for _ in e for y in e] + [rng.randrange(p) for _ in range(4000)]
    A, B = b"".join(map(le, a)), b"".join(map(le, b))
    rinv = bn.inv_mod(1 << 256, p)
    assert ints(pkg.field_op(field, 0, A, B)) == [x * y * rinv % p for x, y in zip(a, b)]
    assert ints(pkg.field_op(field, 1, A, B)) == [(x + y) % p for x, y in zip(a, b)]
    assert ints(pkg.field_op(field, 2, A, B)) == [(x - y) % p for x, y in zip(a, b)]
    assert ints(pkg.field_op(field, 3, A)) == [(x << 256) % p for x in a]
    assert ints(pkg.field_op(field, 4, A)) == [x * rinv % p for x in a]
    assert ints(pkg.field_op(field, 6, A)) == [x * x * rinv % p for x in a]      # the dedicated squaring (kept as an option: field.cuh)
    sub = a[:200]
    got = ints(pkg.field_op(field, 5, b"".join(map(le, sub))))
    # Montgomery inverse: inv(aR) = a^-1 R, i.e. got * x == R^2
    assert all((g * x - (1 << 512)) % p == 0 if x else g == 0 for g, x in zip(got, sub))


def test_field_mul_accepts_unreduced_second_operand(pkg):
    """The witness is multiplied raw (row a4): b may be any 256-bit value."""
    rng = random.Random(5)
    a = [rng.randrange(R) for _ in range(500)]
    b = [rng.randrange(1 << 256) for _ in range(496)] + [(1 << 256) - 1, R, R + 1, 2 * R]
    rinv = bn.inv_mod(1 << 256, R)
    got = ints(pkg.field_op(1, 0, b"".join(map(le, a)), b"".join(map(le, b))))
    assert got == [x * y * rinv % R for x, y in zip(a, b)]


# ---- curve (row a17) ------------------------------------------------------------------------------------
def rec_g1(p):
    return binfile.enc_g1(p)


def rec_g2(p):
    return binfile.enc_g2(p)


def test_g1_group_law(pkg):
    G = bn.G1
    ks = [1, 2, 3, 5, 2, R - 1, R - 2, 77, 1, 12345678901234567890]
    pts = [G.mul(bn.G1_GEN, k) for k in ks]
    raw = b"".join(rec_g1(p) for p in pts)
    want = G.mul(bn.G1_GEN, sum(ks) % R)
    assert binfile.dec_g1(pkg.curve_op(False, 0, raw, len(pts))) == want
    assert binfile.dec_g1(pkg.curve_op(False, 1, raw, len(pts))) == want
    out = pkg.curve_op(False, 2, raw, len(pts))
    assert [binfile.dec_g1(out[64 * i:64 * i + 64]) for i in range(len(pts))] == [G.mul(p, 2) for p in pts]
    # P + (-P) = identity, identity records are skipped, P + P through the mixed-add doubling branch
    p = pts[3]
    raw = rec_g1(p) + rec_g1(G.neg(p)) + rec_g1(None) + rec_g1(pts[4]) + rec_g1(pts[4])
    assert binfile.dec_g1(pkg.curve_op(False, 0, raw, 5)) == G.mul(bn.G1_GEN, 4)
    assert binfile.dec_g1(pkg.curve_op(False, 1, raw, 5)) == G.mul(bn.G1_GEN, 4)
    assert pkg.curve_op(False, 1, rec_g1(p) + rec_g1(G.neg(p)), 2) == bytes(64)


def test_g2_group_law(pkg):
    G = bn.G2
    ks = [1, 2, 3, 5, 2, R - 1, 77, 31337]
    pts = [G.mul(bn.G2_GEN, k) for k in ks]
    raw = b"".join(rec_g2(p) for p in pts)
    want = G.mul(bn.G2_GEN, sum(ks) % R)
    assert binfile.dec_g2(pkg.curve_op(True, 0, raw, len(pts))) == want
    assert binfile.dec_g2(pkg.curve_op(True, 1, raw, len(pts))) == want
    out = pkg.curve_op(True, 2, raw, len(pts))
    assert [binfile.dec_g2(out[128 * i:128 * i + 128]) for i in range(len(pts))] == [G.mul(p, 2) for p in pts]


def test_fixed_base_generator_matches_oracle(pkg):
    rng = random.Random(9)
    ks = [0, 1, 2, R - 1, 255, 256, 1 << 128] + [rng.randrange(R) for _ in range(20)]
    sc = b"".join(map(le, ks))
    o1 = pkg.fixed_base(False, sc, len(ks))
    o2 = pkg.fixed_base(True, sc, len(ks))
    for i, k in enumerate(ks):
        assert binfile.dec_g1(o1[64 * i:64 * i + 64]) == bn.G1.mul(bn.G1_GEN, k)
        assert binfile.dec_g2(o2[128 * i:128 * i + 128]) == bn.G2.mul(bn.G2_GEN, k)


# ---- MSM (rows a9-a13) -------------------------------------------------------------------------------------
def _msm_case(pkg, is_g2, n, seed, scalar_kind="uniform", zero_bases=0.0, msm_c=0):
    """bases k_i * G from the GPU fixed-base generator (itself checked above) -> expected = (sum k_i s_i) G"""
    rng = random.Random(seed)
    ks = [rng.randrange(1, R) for _ in range(n)]
    for i in range(n):
        if rng.random() < zero_bases:
            ks[i] = 0
    if scalar_kind == "uniform":
        ss = [rng.randrange(R) for _ in range(n)]
    elif scalar_kind == "witness":  # mostly bits / small values, like a circom witness
        ss = [rng.choice([0, 1, 1, 0, rng.randrange(1 << 64), rng.randrange(R), 1, 0, 2, R - 1]) for _ in range(n)]
    elif scalar_kind == "ones":
        ss = [1] * n
    else:
        ss = [0] * n
    bases = pkg.fixed_base(is_g2, b"".join(map(le, ks)), n)
    got = pkg.msm_host(is_g2, bases, b"".join(map(le, ss)), n, msm_c=msm_c)
    e = sum(k * s for k, s in zip(ks, ss)) % R
    if is_g2:
        assert binfile.dec_g2(got) == bn.G2.mul(bn.G2_GEN, e)
    else:
        assert binfile.dec_g1(got) == bn.G1.mul(bn.G1_GEN, e)


@pytest.mark.parametrize("n", [1, 2, 31, 33, 1000, 5000])
def test_msm_g1_uniform(pkg, n):
    _msm_case(pkg, False, n, 100 + n)


@pytest.mark.parametrize("kind", ["witness", "ones", "zeros"])
def test_msm_g1_scalar_distributions(pkg, kind):
    _msm_case(pkg, False, 3000, 7, scalar_kind=kind, zero_bases=0.3)


@pytest.mark.parametrize("c", [2, 4, 7, 11, 13, 16, 19])
def test_msm_g1_window_sizes(pkg, c):
    _msm_case(pkg, False, 2000, 40 + c, msm_c=c)


def test_msm_g1_large(pkg):
    _msm_case(pkg, False, 1 << 17, 21, scalar_kind="witness", zero_bases=0.2)
    _msm_case(pkg, False, (1 << 16) + 123, 22)


@pytest.mark.parametrize("n,kind", [(1, "uniform"), (257, "uniform"), (3000, "witness"), (20000, "uniform")])
def test_msm_g2(pkg, n, kind):
    _msm_case(pkg, True, n, 300 + n, scalar_kind=kind, zero_bases=0.35)


def test_msm_against_oracle_pippenger_inputs(pkg):
    """same points, oracle's own msm (not the closed form)"""
    rng = random.Random(77)
    pts = [bn.G1.mul(bn.G1_GEN, rng.randrange(1, R)) for _ in range(40)] + [None, None]
    ss = [rng.randrange(R) for _ in range(42)]
    got = pkg.msm_host(False, b"".join(rec_g1(p) for p in pts), b"".join(map(le, ss)), 42)
    assert binfile.dec_g1(got) == bn.G1.msm(pts, ss)
    assert pkg.msm_host(False, b"", b"", 0) == bytes(64)  # empty input -> infinity


def test_msm_repeated_and_opposite_bases(pkg):
    """the same base many times in one bucket, and P next to -P: the doubling / cancellation cases of the group law
    (XYZZ mixed addition, and the batched-affine pair round when G16_PAIR_MERGE is on)"""
    rng = random.Random(99)
    ks = [5, 5, R - 5, 5, 7, 7, 7, 7, R - 7, 11, R - 11, 11, 13] * 40
    rng.shuffle(ks)
    for ss in ([1] * len(ks), [3] * len(ks), [rng.choice([1, 2, R - 1]) for _ in ks]):
        for is_g2 in (False, True):
            bases = pkg.fixed_base(is_g2, b"".join(map(le, ks)), len(ks))
            got = pkg.msm_host(is_g2, bases, b"".join(map(le, ss)), len(ks), msm_c=4)
            e = sum(k * s for k, s in zip(ks, ss)) % R
            if is_g2:
                assert binfile.dec_g2(got) == bn.G2.mul(bn.G2_GEN, e)
            else:
                assert binfile.dec_g1(got) == bn.G1.mul(bn.G1_GEN, e)


MERGE_WORKER = r'''
import os, sys, random
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
from conftest import load_pkg
from oracle import bn254 as bn, binfile
pkg = load_pkg()
R = bn.R
le = lambda x: int(x).to_bytes(32, "little")
rng = random.Random(5)
cases = [([5, 5, R - 5, 5, 7, 7, 7, 7, R - 7, 11, R - 11, 11, 13] * 40, None, 4),
         ([rng.randrange(1, R) for _ in range(6000)], "uniform", 0),
         ([rng.randrange(1, R) for _ in range(6001)], "witness", 9)]
for ks, kind, c in cases:
    if kind is None:
        ss = [rng.choice([1, 2, R - 1]) for _ in ks]
    elif kind == "uniform":
        ss = [rng.randrange(R) for _ in ks]
    else:
        ss = [rng.choice([0, 1, 1, rng.randrange(1 << 64), rng.randrange(R), R - 1]) for _ in ks]
    e = sum(k * s for k, s in zip(ks, ss)) % R
    for is_g2 in (False, True):
        bases = pkg.fixed_base(is_g2, b"".join(map(le, ks)), len(ks))
        got = pkg.msm_host(is_g2, bases, b"".join(map(le, ss)), len(ks), msm_c=c)
        want = bn.G2.mul(bn.G2_GEN, e) if is_g2 else bn.G1.mul(bn.G1_GEN, e)
        assert (binfile.dec_g2(got) if is_g2 else binfile.dec_g1(got)) == want, (kind, is_g2)
print("merge ok")
'''


PROVE_WORKER = r'''
import json
from oracle import groth16 as g16
sy = pkg.Synth(30000, 1, 32000, 11, 12)
zkey, wtns = sy.write()
r, s = 0x1234567 << 200, 0x7654321 << 190
with pkg.ZKey(zkey, h_table=1) as zk:
    proof, public = zk.prove(wtns, r=r, s=s)
    assert json.loads(proof) == g16.proof_object(*sy.expected(r, s))
print("prove ok")
'''


@pytest.mark.parametrize("env", [{"G16_G2_PAIR": "1"}, {"G16_DENSE_MIN": "0"}, {"G16_W_TABLE": "1"}, {"G16_CHAIN": "1"},
                                 {"G16_SERIAL": "1"}, {"G16_GATE": "3", "G16_HPRIO": "1"}])
def test_optional_kernel_paths_and_schedules(tmp_path, env):
    """the paths that are selected by size or kept as options (each knob is read once per process, hence the subprocess):
    the lane-pair G2 accumulation, the fixed-slot digits + radix sort of a dense MSM (forced at every size), window tables
    for the witness sections, and the alternative stream schedules -- MSM sums equal the closed form, the proof is bit-exact"""
    import sys
    (tmp_path / "w.py").write_text(MERGE_WORKER.replace('print("merge ok")', "") + PROVE_WORKER)
    p = subprocess.run([sys.executable, str(tmp_path / "w.py"), ROOT], env=dict(os.environ, **env), capture_output=True, text=True, timeout=900)
    assert p.returncode == 0 and "prove ok" in p.stdout, p.stdout + p.stderr


def test_msm_linearity_at_scale(pkg):
    """size-independent property at 2^20: MSM(P, s) + MSM(P, t) == MSM(P, s + t)"""
    import numpy as np
    n = 1 << 20
    rng = np.random.default_rng(5)
    ks = rng.integers(1, 1 << 62, size=n, dtype=np.uint64)
    kb = np.zeros((n, 4), dtype=np.uint64)
    kb[:, 0] = ks
    bases = pkg.fixed_base(False, kb.tobytes(), n)
    s = np.zeros((n, 4), dtype=np.uint64)
    t = np.zeros((n, 4), dtype=np.uint64)
    s[:, :3] = rng.integers(0, 1 << 63, size=(n, 3), dtype=np.uint64)
    t[:, :3] = rng.integers(0, 1 << 63, size=(n, 3), dtype=np.uint64)
    u = s + t  # limbs < 2^64, no carries
    ps = binfile.dec_g1(pkg.msm_host(False, bases, s.tobytes(), n))
    pt = binfile.dec_g1(pkg.msm_host(False, bases, t.tobytes(), n))
    pu = binfile.dec_g1(pkg.msm_host(False, bases, u.tobytes(), n))
    assert bn.G1.add(ps, pt) == pu
    # closed form: sum k_i u_i
    e = 0
    uu = [int(u[:, j].astype(object) @ ks.astype(object)) for j in range(3)]
    e = (uu[0] + (uu[1] << 64) + (uu[2] << 128)) % R
    assert pu == bn.G1.mul(bn.G1_GEN, e)


# ---- NTT pipeline (rows a5-a8) ---------------------------------------------------------------------------
@pytest.mark.parametrize("logn", [1, 2, 5, 10, 11, 13])
def test_coset_ntt_matches_oracle(pkg, logn):
    n = 1 << logn
    rng = random.Random(logn)
    a = [rng.randrange(R) for _ in range(n)]
    if n >= 4:
        a[0], a[1], a[2] = 0, 1, R - 1
    wn = g16.root_of_unity(logn)
    inc = g16.root_of_unity(logn + 1)
    coef = g16.intt(a, wn)
    t = 1
    for i in range(n):
        coef[i] = coef[i] * t % R
        t = t * inc % R
    want = g16.ntt(coef, wn)
    got = ints(pkg.coset_ntt_host(b"".join(le(x * MR % R) for x in a), logn))
    assert got == [x * MR % R for x in want]


def test_coset_ntt_polynomial_property_at_scale(pkg):
    """2^20: evaluations of a low-degree polynomial on the domain -> its evaluations on the odd coset"""
    import numpy as np
    logn = 20
    n = 1 << logn
    w = g16.root_of_unity(logn)
    inc = g16.root_of_unity(logn + 1)
    c = [3, 5, 7, 11]  # f(x) = 3 + 5x + 7x^2 + 11x^3
    f = lambda x: (c[0] + c[1] * x + c[2] * x * x + c[3] * x * x * x) % R
    xs = [1] * n
    for i in range(1, n):
        xs[i] = xs[i - 1] * w % R
    data = b"".join(le(f(x) * MR % R) for x in xs)
    got = ints(pkg.coset_ntt_host(data, logn))
    for i in [0, 1, 2, 12345, n // 2, n - 1]:
        assert got[i] == f(inc * xs[i] % R) * MR % R


# ---- full prove (rows a1-a16) ------------------------------------------------------------------------------
def test_prove_bit_exact_vs_oracle(pkg, small_fixture):
    f = small_fixture
    r, s = 0xDEADBEEF << 200 | 17, (R - 5)
    with pkg.ZKey(bytes(f["zkey_bytes"])) as zk:
        assert (zk.n_vars, zk.n_public, zk.domain_size) == (85, 2, 128)
        for style, name in ((pkg.STYLE_RAPIDSNARK, "rapidsnark"), (pkg.STYLE_SNARKJS, "snarkjs")):
            proof, public = zk.prove(bytes(f["wtns_bytes"]), r=r, s=s, style=style)
            want_proof, want_public = g16.prove(f["zk"], f["w"], r, s, style=name)
            assert proof == want_proof and public == want_public
        # H scalars on their own (rows a4-a8)
        assert ints(zk.h_scalars()) == g16.h_scalars(f["zk"], f["w"])
        # fresh randomness: must verify, and differ between calls
        p1, u1 = zk.prove(bytes(f["wtns_bytes"]))
        p2, _ = zk.prove(bytes(f["wtns_bytes"]))
        assert p1 != p2
        assert V.verify(f["vkey"], json.loads(u1), json.loads(p1))
        assert V.verify(f["vkey"], json.loads(u1), json.loads(p2))


def test_prove_tiny_handmade_circuit(pkg):
    cons = [([(2, 1)], [(2, 1)], [(3, 1)]), ([(3, 1)], [(2, 1)], [(4, 1)]),
            ([(4, 1), (2, 1), (0, 5)], [(0, 1)], [(1, 1)])]
    circ = factory.ListCircuit(5, 1, cons, [1, 35, 3, 9, 27])
    zk, vkey = factory.setup(circ, factory.Trapdoor(77))
    with pkg.ZKey(binfile.write_zkey(zk)) as h:
        proof, public = h.prove(binfile.write_wtns(circ.witness()), r=1111, s=2222)
    assert (proof, public) == g16.prove(zk, circ.witness(), 1111, 2222)
    assert V.verify(vkey, ["35"], json.loads(proof))


def test_prove_sharded_equals_whole(pkg, small_fixture):
    """Base-range sharding: 3 shards' partial sums combine to the same proof."""
    f = small_fixture
    r, s = 123456789, 987654321
    want = g16.prove(f["zk"], f["w"], r, s)
    partials = b""
    handles = [pkg.ZKey(bytes(f["zkey_bytes"]), shard_rank=k, shard_count=3) for k in range(3)]
    for h in handles:
        h.load_witness(bytes(f["wtns_bytes"]))
        partials += h.prove_partial()
    assert handles[0].combine(partials, r=r, s=s) == want
    with pytest.raises(pkg.G16Error):
        handles[1].prove_loaded(r=r, s=s)
    for h in handles:
        h.close()


def test_error_conventions(pkg, small_fixture):
    f = small_fixture
    with pkg.ZKey(bytes(f["zkey_bytes"])) as zk:
        with pytest.raises(pkg.G16Error) as e:
            zk.prove(binfile.write_wtns(f["w"][:-1]))
        assert e.value.code == pkg.G16_ERR_WITNESS_LEN and "Invalid witness length. Circuit: 85, witness: 84" in e.value.message
        with pytest.raises(pkg.G16Error) as e:
            zk.prove(binfile.write_wtns(f["w"], prime=Q))
        assert e.value.code == pkg.G16_ERR_CURVE and "Curve of the witness does not match" in e.value.message
        with pytest.raises(pkg.G16Error) as e:
            zk.prove(b"wtnX" + bytes(f["wtns_bytes"])[4:])
        assert e.value.code == pkg.G16_ERR_FORMAT
        with pytest.raises(pkg.G16Error) as e:
            zk.prove(bytes(f["wtns_bytes"])[:-7])
        assert e.value.code == pkg.G16_ERR_FORMAT
    with pytest.raises(pkg.G16Error) as e:
        pkg.ZKey(bytes(f["zkey_bytes"])[:5000])
    assert e.value.code == pkg.G16_ERR_FORMAT
    with pytest.raises(pkg.G16Error) as e:
        pkg.ZKey(bytes(f["wtns_bytes"]))
    assert e.value.code == pkg.G16_ERR_FORMAT


def test_cli_and_js_api_mirror(pkg, small_fixture, tmp_path):
    f = small_fixture
    zp, wp = tmp_path / "c_final.zkey", tmp_path / "witness.wtns"
    zp.write_bytes(bytes(f["zkey_bytes"]))
    wp.write_bytes(bytes(f["wtns_bytes"]))
    exe = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")
    env = dict(os.environ, G16_R="1111", G16_S="2222")
    p = subprocess.run([exe, str(zp), str(wp), str(tmp_path / "proof.json"), str(tmp_path / "public.json")], env=env,
                       capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    want = g16.prove(f["zk"], f["w"], 1111, 2222)
    assert ((tmp_path / "proof.json").read_text(), (tmp_path / "public.json").read_text()) == want
    # snarkjs JS API shape
    proof, signals = pkg.groth16.prove(str(zp), str(wp), r=1111, s=2222)
    assert proof == g16.proof_object(*g16.prove_points(f["zk"], f["w"], 1111, 2222), style="snarkjs")
    assert signals == [str(x) for x in f["w"][1:3]]
    assert V.verify(f["vkey"], signals, proof)
    # failure = non-zero exit + message (scripts/lib/error_handling.sh:16-41)
    bad = tmp_path / "bad.wtns"
    bad.write_bytes(binfile.write_wtns(f["w"][:-2]))
    p = subprocess.run([exe, str(zp), str(bad), str(tmp_path / "p2.json"), str(tmp_path / "u2.json")], capture_output=True, text=True)
    assert p.returncode != 0 and "Invalid witness length" in p.stderr
    assert not (tmp_path / "p2.json").exists()


# ---- SYNTH-1 generator twins + closed form at scale -----------------------------------------------------------
def test_cuda_fixture_generator_matches_python_factory(pkg, small_fixture):
    f = small_fixture
    sy = pkg.Synth(85, 2, 100, 0xB2000001, 0xB2000002)
    zkey, wtns = sy.write()
    assert bytes(wtns) == f["wtns_bytes"]
    assert bytes(zkey) == f["zkey_bytes"]
    assert sy.vkey() == {k: v for k, v in f["vkey"].items()}
    pa, pb, pc = factory.closed_form_proof(f["circ"], f["w"], f["td"], 5, 6)
    assert sy.expected(5, 6) == (pa, pb, pc)


@pytest.mark.parametrize("n_vars,n_public,n_cons", [(30000, 1, 32000), (200000, 13, 262130)])
def test_prove_closed_form_at_scale(pkg, n_vars, n_public, n_cons):
    """Known trapdoor => the proof for pinned r,s is three scalar multiplications, whatever the size."""
    sy = pkg.Synth(n_vars, n_public, n_cons, 11, 12)
    zkey, wtns = sy.write()
    r, s = 0xABCDEF0123456789 << 64, 0x1F2E3D4C5B6A7988 << 128
    with pkg.ZKey(zkey) as zk:
        proof, public = zk.prove(wtns, r=r, s=s)
        pa, pb, pc = sy.expected(r, s)
        assert json.loads(proof) == g16.proof_object(pa, pb, pc)
        p2, u2 = zk.prove(wtns)
        assert V.verify(sy.vkey(), json.loads(u2), json.loads(p2))
        st = zk.msm_stats()
        assert st["H"]["additions"] > 0 and st["B2"]["additions"] > 0


@pytest.mark.parametrize("msm_c", [0, 6, 9, 13])
def test_prove_with_h_window_table_bit_exact(pkg, small_fixture, msm_c):
    """opts.h_table = 1: the H MSM runs over the fixed-base window table (rows 2^(c w) * H_i, one bucket set);
    the proof bytes must not change.  msm_c sweeps the table's row count (43, 29, 20 rows)."""
    f = small_fixture
    r, s = (R - 1), 0x1234567 << 190
    want = g16.prove(f["zk"], f["w"], r, s)
    with pkg.ZKey(bytes(f["zkey_bytes"]), msm_c=msm_c, h_table=1) as zk:
        assert zk.prove(bytes(f["wtns_bytes"]), r=r, s=s) == want
        assert zk.prove(bytes(f["wtns_bytes"]), r=r, s=s) == want      # the table is reused, not rebuilt
    # sharded: each shard tabulates its own slice of section 9
    partials = b""
    handles = [pkg.ZKey(bytes(f["zkey_bytes"]), shard_rank=k, shard_count=2, msm_c=msm_c, h_table=1) for k in range(2)]
    for h in handles:
        h.load_witness(bytes(f["wtns_bytes"]))
        partials += h.prove_partial()
    assert handles[0].combine(partials, r=r, s=s) == want
    for h in handles:
        h.close()


def test_prove_closed_form_at_scale_with_h_window_table(pkg):
    sy = pkg.Synth(200000, 13, 262130, 11, 12)
    zkey, wtns = sy.write()
    r, s = 0xABCDEF0123456789 << 64, 0x1F2E3D4C5B6A7988 << 128
    with pkg.ZKey(zkey, h_table=1) as zk, pkg.ZKey(zkey, h_table=-1) as plain:
        proof, public = zk.prove(wtns, r=r, s=s)
        assert json.loads(proof) == g16.proof_object(*sy.expected(r, s))
        assert plain.prove(wtns, r=r, s=s) == (proof, public)
        a, b = zk.msm_stats()["H"], plain.msm_stats()["H"]
        assert a["additions"] < b["additions"] and a["c"] > b["c"]      # wider windows: fewer mixed additions


def test_infinity_records_in_every_section(pkg, small_fixture):
    """all-zero (infinity) records in sections 5-9, the zkey convention for wires a matrix never touches: skipped by the
    digit kernels and by the H window table; the proof must equal the oracle's on the same edited key."""
    import copy
    f = small_fixture
    zk = copy.deepcopy(f["zk"])
    rng = random.Random(31)
    for sec in (zk.A, zk.B1, zk.C, zk.H):
        for i in rng.sample(range(len(sec)), max(1, len(sec) // 5)):
            sec[i] = None
    for i in rng.sample(range(len(zk.B2)), len(zk.B2) // 5):
        zk.B2[i] = None
        zk.B1[i] = None
    zk.H[0] = None
    zk.H[-1] = None
    r, s = 77, 88
    want = g16.prove(zk, f["w"], r, s)
    raw = binfile.write_zkey(zk)
    for h_table in (1, -1):
        with pkg.ZKey(raw, h_table=h_table) as h:
            assert h.prove(bytes(f["wtns_bytes"]), r=r, s=s) == want


def test_prove_stream_double_buffered_witness(pkg, small_fixture):
    """g16_witness_prefetch / g16_witness_commit: the next witness is copied while the current proof computes; each
    proof must be the oracle's proof for ITS witness (also for a witness that does not satisfy the circuit: the prover,
    like the reference's, does not check)."""
    f = small_fixture
    rng = random.Random(5)
    w2 = list(f["w"])
    for i in rng.sample(range(1, len(w2)), 20):
        w2[i] = rng.randrange(R)
    ws = [f["w"], w2, f["w"], w2, w2]
    r, s = 31337, 271828
    want = [g16.prove(f["zk"], w, r, s) for w in ws]
    with pkg.ZKey(bytes(f["zkey_bytes"])) as zk:
        got = list(zk.prove_stream([binfile.write_wtns(w) for w in ws], r=r, s=s))
        assert got == want
        assert want[0] != want[1]
        assert list(zk.prove_stream([])) == []
        with pytest.raises(pkg.G16Error):
            zk.commit_witness()                       # nothing prefetched
        with pytest.raises(pkg.G16Error) as e:
            zk.prefetch_witness(binfile.write_wtns(f["w"][:-1]))
        assert e.value.code == pkg.G16_ERR_WITNESS_LEN
        # a failed prefetch leaves the loaded witness in place
        assert zk.prove_loaded(r=r, s=s) == want[-1]


# ---- batch apply key (operator level) ---------------------------------------------------
@pytest.mark.parametrize("is_g2", [False, True])
def test_batch_apply_key_matches_oracle_scalar_multiplication(pkg, is_g2):
    """out[i] = (first * inc^i) * P_i, ffjavascript G.batchApplyKey; inc = 1 is the `zkey contribute` case (every point of
    sections 8 and 9 times 1/delta').  Oracle: double-and-add in Python."""
    rng = random.Random(2024 + is_g2)
    G, gen, rec, dec = (bn.G2, bn.G2_GEN, rec_g2, binfile.dec_g2) if is_g2 else (bn.G1, bn.G1_GEN, rec_g1, binfile.dec_g1)
    size = 128 if is_g2 else 64
    n = 12 if is_g2 else 40
    pts = [G.mul(gen, rng.randrange(1, R)) for _ in range(n)]
    pts[3] = None                                   # infinity record stays infinity
    raw = b"".join(rec(p) for p in pts)
    for first, inc in ((rng.randrange(1, R), 1), (rng.randrange(1, R), rng.randrange(2, R)), (1, 1), (R - 1, 1), (0, 5), (7, 0)):
        out = pkg.batch_apply_key(is_g2, raw, n, first, inc)
        for i, p in enumerate(pts):
            k = first * pow(inc, i, R) % R
            want = None if p is None or k == 0 else G.mul(p, k)
            assert dec(out[size * i:size * i + size]) == want, (first, inc, i)
    assert pkg.batch_apply_key(is_g2, b"", 0, 5, 1) == b""
    with pytest.raises(pkg.G16Error):
        pkg.batch_apply_key(is_g2, raw, n, R, 1)     # scalar not reduced


def test_batch_apply_key_is_the_contribute_update_of_sections_8_and_9(pkg, small_fixture):
    """A `zkey contribute` with secret d turns delta into d*delta and sections 8 (C) and 9 (H) into (1/d) times
    themselves (snarkjs zkey_contribute.js); a key updated that way with this operator must still produce proofs
    that verify against the vkey with the updated delta."""
    import copy
    f = small_fixture
    d = 0x1234567890ABCDEF1234567
    dinv = pow(d, -1, R)
    zk = copy.deepcopy(f["zk"])
    c_raw = b"".join(rec_g1(p) for p in zk.C)
    h_raw = b"".join(rec_g1(p) for p in zk.H)
    c_new = pkg.batch_apply_key(False, c_raw, len(zk.C), dinv, 1)
    h_new = pkg.batch_apply_key(False, h_raw, len(zk.H), dinv, 1)
    zk.C = [binfile.dec_g1(c_new[64 * i:64 * i + 64]) for i in range(len(zk.C))]
    zk.H = [binfile.dec_g1(h_new[64 * i:64 * i + 64]) for i in range(len(zk.H))]
    zk.delta1 = bn.G1.mul(zk.delta1, d)
    zk.delta2 = bn.G2.mul(zk.delta2, d)
    vkey = copy.deepcopy(f["vkey"])
    x, y = zk.delta2
    vkey["vk_delta_2"] = [[str(x[0]), str(x[1])], [str(y[0]), str(y[1])], ["1", "0"]]
    with pkg.ZKey(binfile.write_zkey(zk)) as h:
        proof, public = h.prove(bytes(f["wtns_bytes"]))
    assert V.verify(vkey, json.loads(public), json.loads(proof))
    assert not V.verify(f["vkey"], json.loads(public), json.loads(proof))     # the old delta no longer fits


@pytest.mark.parametrize("n_vars,n_public,n_cons,shards", [(85, 2, 100, 1), (85, 2, 100, 3), (30000, 1, 32000, 1), (200000, 13, 262130, 5)])
def test_split_transform_halves_recombine_bit_exact(pkg, n_vars, n_public, n_cons, shards):
    """g16_h_transform_half + g16_prove_partial_joined_halves: the six half transforms (coefficients split by parity, half
    domain, recombined inside the join) give the same proof as the whole transforms, for every shard's H range --
    including ranges that straddle the middle of the domain.  2^15 and 2^18 points: one- and two-pass half transforms."""
    import torch
    sy = pkg.Synth(n_vars, n_public, n_cons, 11, 12)
    zkey, wtns = sy.write()
    r, s = 0xABCDEF0123456789 << 64, 0x1F2E3D4C5B6A7988 << 128
    handles = [pkg.ZKey(zkey, shard_rank=k, shard_count=shards) for k in range(shards)]
    n = handles[0].domain_size
    halves = []
    for poly in range(3):                     # any handle can transform: the polynomial buffers are whole on every rank
        for parity in range(2):
            h = handles[(2 * poly + parity) % shards]
            h.load_witness(wtns)
            h.h_transform_half(poly, parity)
            ptr, _, _ = h.h_buffer(poly)
            buf = torch.as_tensor(pkg._DevMem(ptr, n * 32), device="cuda").clone()
            assert torch.equal(buf[:n * 16], buf[n * 16:])      # the half, twice
            halves.append(buf)
    want, got = b"", b""
    for h in handles:
        h.load_witness(wtns)
        want += h.prove_partial()
        _, lo, cnt = h.h_buffer(0)
        got += h.prove_partial_joined_halves([b.data_ptr() + lo * 32 for b in halves])
    # partial sums are projective (their representation depends on the order of additions): compare the proofs
    assert handles[0].combine(got, r=r, s=s) == handles[0].combine(want, r=r, s=s)
    proof, _ = handles[0].combine(got, r=r, s=s)
    assert json.loads(proof) == g16.proof_object(*sy.expected(r, s))
    for h in handles:
        h.close()


SHARD_WORKER = r'''
import os, sys, json
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
from conftest import load_pkg
pkg = load_pkg()
dist.init_process_group("gloo")          # two ranks share cuda:0 in the test, so the gather runs over gloo
rank, world = dist.get_rank(), dist.get_world_size()
zkey = open(sys.argv[2], "rb").read(); wtns = open(sys.argv[3], "rb").read()
zk = pkg.ZKey(zkey, device=0, shard_rank=rank, shard_count=world)
proof, public = pkg.prove_sharded(zk, wtns, r=424242, s=171717)
open(sys.argv[4] + ".%d" % rank, "w").write(proof + "\n" + public)
# same proof with the three coset transforms spread over the ranks and their slices exchanged
p2, u2 = pkg.prove_sharded_distributed(zk, wtns, r=424242, s=171717)
open(sys.argv[4] + ".dist.%d" % rank, "w").write(p2 + "\n" + u2)
zk.close()
dist.barrier(); dist.destroy_process_group()
'''


@pytest.mark.parametrize("world", [2, 3, 6, 7])
def test_prove_sharded_ranks(pkg, small_fixture, tmp_path, world):
    """One proof over several ranks (base-range shards, 768-byte partial gather; then also with the transforms
    spread over the ranks and an all-to-all of their slices) == the whole-key proof, on every rank.  From six ranks on
    each transform is split in halves over two ranks (7: an H range that straddles the middle of the domain)."""
    import socket
    import sys
    f = small_fixture
    (tmp_path / "k.zkey").write_bytes(bytes(f["zkey_bytes"]))
    (tmp_path / "w.wtns").write_bytes(bytes(f["wtns_bytes"]))
    (tmp_path / "worker.py").write_text(SHARD_WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(tmp_path / "worker.py"), ROOT, str(tmp_path / "k.zkey"),
                                       str(tmp_path / "w.wtns"), str(tmp_path / "out")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out = p.communicate(timeout=300)[0]
        assert p.returncode == 0, out
    want = "\n".join(g16.prove(f["zk"], f["w"], 424242, 171717))
    for rank in range(world):
        assert (tmp_path / ("out.%d" % rank)).read_text() == want
        assert (tmp_path / ("out.dist.%d" % rank)).read_text() == want
