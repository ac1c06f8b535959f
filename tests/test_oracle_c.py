"""The C oracle (CPU baseline / mid-size checker) against the Python oracle, which is pinned on the reference's vectors."""
import json
import random

from oracle import binfile, bn254 as bn, cport, factory, groth16 as g16

R, Q = bn.R, bn.Q


def le(x):
    return int(x).to_bytes(32, "little")


def ints(b):
    return [int.from_bytes(b[i:i + 32], "little") for i in range(0, len(b), 32)]


def test_c_field_ops():
    rng = random.Random(3)
    for field, p in ((0, Q), (1, R)):
        a = [0, 1, p - 1, (1 << 256) % p] + [rng.randrange(p) for _ in range(300)]
        b = [p - 1, 0, p - 1, 5] + [rng.randrange(p) for _ in range(300)]
        A, B = b"".join(map(le, a)), b"".join(map(le, b))
        rinv = bn.inv_mod(1 << 256, p)
        assert ints(cport.field_op(field, 0, A, B)) == [x * y * rinv % p for x, y in zip(a, b)]
        assert ints(cport.field_op(field, 1, A, B)) == [(x + y) % p for x, y in zip(a, b)]
        assert ints(cport.field_op(field, 2, A, B)) == [(x - y) % p for x, y in zip(a, b)]
        assert ints(cport.field_op(field, 3, A)) == [(x << 256) % p for x in a]
        got = ints(cport.field_op(field, 5, A))
        assert all((g * x - (1 << 512)) % p == 0 if x else g == 0 for g, x in zip(got, a))


def test_c_msm_matches_python(small_fixture):
    zk, w = small_fixture["zk"], small_fixture["w"]
    sc = b"".join(map(le, w))
    g1 = b"".join(binfile.enc_g1(p) for p in zk.A)
    assert binfile.dec_g1(cport.msm(False, g1, sc, len(w))) == bn.G1.msm(zk.A, w)
    g2 = b"".join(binfile.enc_g2(p) for p in zk.B2)
    assert binfile.dec_g2(cport.msm(True, g2, sc, len(w))) == bn.G2.msm(zk.B2, w)
    assert cport.msm(False, b"", b"", 0) == bytes(64)
    # many points, several threads: bases k_i G from the fixed-base table, closed-form sum
    rng = random.Random(11)
    fb = bn.FixedBase(bn.G1)
    ks = [rng.randrange(1, 1 << 40) for _ in range(700)] + [0] * 20
    ss = [rng.randrange(R) for _ in range(720)]
    bases = b"".join(binfile.enc_g1(fb.mul(k) if k else None) for k in ks)
    want = bn.G1.mul(bn.G1_GEN, sum(k * s for k, s in zip(ks, ss)) % R)
    assert binfile.dec_g1(cport.msm(False, bases, b"".join(map(le, ss)), 720)) == want


def test_c_coset_ntt_matches_python():
    for logn in (1, 3, 9, 12):
        n = 1 << logn
        rng = random.Random(logn)
        a = [rng.randrange(R) for _ in range(n)]
        wn, inc = g16.root_of_unity(logn), g16.root_of_unity(logn + 1)
        coef = g16.intt(a, wn)
        t = 1
        for i in range(n):
            coef[i] = coef[i] * t % R
            t = t * inc % R
        want = g16.ntt(coef, wn)
        got = ints(cport.coset_ntt(b"".join(le(x * bn.MONT_R_R % R) for x in a), logn))
        assert got == [x * bn.MONT_R_R % R for x in want]


def test_c_prove_matches_python_and_closed_form(small_fixture):
    f = small_fixture
    r, s = 0xFEEDFACE << 180, R - 12345
    assert ints(cport.h_scalars(f["zkey_bytes"], f["wtns_bytes"])) == g16.h_scalars(f["zk"], f["w"])
    for style in ("rapidsnark", "snarkjs"):
        assert cport.prove(f["zkey_bytes"], f["wtns_bytes"], r, s, style) == g16.prove(f["zk"], f["w"], r, s, style)
    pa, pb, pc = factory.closed_form_proof(f["circ"], f["w"], f["td"], r, s)
    assert cport.prove_points(f["zkey_bytes"], f["wtns_bytes"], r, s) == (pa, pb, pc)
    t = cport.last_timings()
    assert t["total"] > 0


def test_c_prove_error_codes(small_fixture):
    import pytest
    f = small_fixture
    with pytest.raises(ValueError, match="code 3"):
        cport.prove_points(f["zkey_bytes"], binfile.write_wtns(f["w"][:-1]), 1, 2)
    with pytest.raises(ValueError, match="code 4"):
        cport.prove_points(f["zkey_bytes"], binfile.write_wtns(f["w"], prime=Q), 1, 2)
    with pytest.raises(ValueError, match="code 6"):
        cport.prove_points(f["zkey_bytes"][:3000], f["wtns_bytes"], 1, 2)


def test_c_fixture_generator_matches_python_factory():
    """oracle/c/synth1.h (host-core SYNTH-1 generator the reference arm of bench.py uses) writes the bytes of the
    Python factory, and its closed-form proof equals the factory's."""
    for nv, npub, nc in ((85, 2, 100), (300, 1, 310), (1500, 13, 1600)):
        circ = factory.SynthCircuit(n_vars=nv, n_public=npub, n_constraints=nc, seed=0xB2000001)
        td = factory.Trapdoor(0xB2000002)
        zk, vkey = factory.setup(circ, td)
        w = circ.witness()
        sy = cport.Synth(nv, npub, nc, 0xB2000001, 0xB2000002)
        z2, w2 = sy.write()
        assert bytes(z2) == binfile.write_zkey(zk)
        assert bytes(w2) == binfile.write_wtns(w)
        assert sy.vkey() == vkey
        r, s = 0x1234567 << 200, 0x7654321 << 190
        want = factory.closed_form_proof(circ, w, td, r % R, s % R)
        assert sy.expected(r % R, s % R) == want
        assert cport.prove_points(bytes(z2), bytes(w2), r, s) == want


def test_cport_host_tag_is_stable_across_processes():
    """the marker that decides whether liboracle.so is rebuilt must not depend on Python's per-process hash salt"""
    import subprocess
    import sys
    code = "import sys; sys.path.insert(0, %r); from oracle import cport; cport.lib(); print(open(cport._LIB + '.host').read())"
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    a = subprocess.run([sys.executable, "-c", code % root], capture_output=True, text=True, check=True).stdout
    before = os.path.getmtime(cport._LIB)
    b = subprocess.run([sys.executable, "-c", code % root], capture_output=True, text=True, check=True,
                       env=dict(os.environ, PYTHONHASHSEED="12345")).stdout
    assert a == b and len(a.strip()) == 64
    assert os.path.getmtime(cport._LIB) == before      # second process did not rebuild
