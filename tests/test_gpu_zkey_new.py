"""`zkey new` on the GPU (g16_zkey_new, csrc/zkey_new.cu) through the C ABI and the CLI, against the oracle
(oracle/zkey_new.py: byte for byte, circuit hash included) and against the known-trapdoor closed form; the new key
then goes through the rest of the reference's setup and prove steps on the same library."""
import json
import os
import subprocess

import pytest

from oracle import binfile, bn254 as bn, factory, verify as V, zkey_new as ZN

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sections(data):
    _, secs = binfile.read_binfile(data, b"zkey")
    return {t: bytes(data[v[0][0]:v[0][0] + v[0][1]]) for t, v in secs.items()}


@pytest.fixture(scope="module")
def small(tmp_path_factory):
    d = tmp_path_factory.mktemp("zkey_new")
    circ = factory.SynthCircuit(n_vars=85, n_public=2, n_constraints=100, seed=0xB2000001)
    td = factory.Trapdoor(0xB2000002)
    td.gamma = td.delta = 1
    (d / "c.r1cs").write_bytes(ZN.write_r1cs(circ, n_pub_out=1))
    for power in (7, 8):
        (d / ("pot%d.ptau" % power)).write_bytes(ZN.write_ptau(td.tau, td.alpha, td.beta, power))
    return {"dir": d, "circ": circ, "td": td}


@pytest.mark.parametrize("power", [8, 7])
def test_zkey_new_matches_the_oracle_byte_for_byte(pkg, small, power):
    d = small["dir"]
    out = d / ("new%d.zkey" % power)
    cs = pkg.zkey_new(d / "c.r1cs", d / ("pot%d.ptau" % power), out)
    want = ZN.zkey_new((d / "c.r1cs").read_bytes(), (d / ("pot%d.ptau" % power)).read_bytes())
    got = out.read_bytes()
    a, b = sections(got), sections(want)
    for t in range(1, 11):
        assert a[t] == b[t], t
    assert got == want and cs == b[10][:64]


def test_new_key_goes_through_contribute_export_prove_verify(pkg, small):
    """the reference's sequence (scripts/g16_setup.sh:239-295, g16_prove.sh, g16_verify.sh) on one library"""
    d, circ = small["dir"], small["circ"]
    pkg.zkey_new(d / "c.r1cs", d / "pot8.ptau", d / "k0.zkey")
    pkg.zkey_contribute(d / "k0.zkey", d / "k1.zkey", name="first", entropy="abc")
    pkg.zkey_beacon(d / "k1.zkey", d / "k2.zkey", "0102030405060708090a0b0c0d0e0f101112131415161718191a1b1c1d1e1f", 4, name="final")
    for key in ("k0.zkey", "k2.zkey"):
        vkey = json.loads(pkg.zkey_export_vkey(d / key))
        with pkg.ZKey((d / key).read_bytes()) as zk:
            proof, public = zk.prove(binfile.write_wtns(circ.witness()))
        assert V.verify(vkey, json.loads(public), json.loads(proof)), key
        assert pkg.verify(json.dumps(vkey), public, proof)


def test_zkey_new_at_2_to_the_10_equals_the_closed_form(pkg, tmp_path):
    """1000 constraints, every coefficient class (+-1, 31-bit, full width), rows of more than one chunk (wire 0)"""
    circ = factory.SynthCircuit(n_vars=800, n_public=5, n_constraints=1000, seed=77)
    td = factory.Trapdoor(78)
    td.gamma = td.delta = 1
    zk, _ = factory.setup(circ, td)
    want = sections(binfile.write_zkey(zk))
    (tmp_path / "c.r1cs").write_bytes(ZN.write_r1cs(circ))
    (tmp_path / "pot.ptau").write_bytes(ZN.write_ptau(td.tau, td.alpha, td.beta, 11))
    cs = pkg.zkey_new(tmp_path / "c.r1cs", tmp_path / "pot.ptau", tmp_path / "k.zkey")
    got = sections((tmp_path / "k.zkey").read_bytes())
    for t in range(1, 10):
        assert got[t] == want[t], t
    fb1 = bn.FixedBase(bn.G1)
    n = circ.domain_size
    mono, t = [], 1
    for _ in range(2 * n - 1):
        mono.append(fb1.mul(t))
        t = t * td.tau % bn.R
    assert cs == ZN.circuit_hash(zk.alpha1, zk.beta1, zk.beta2, zk.ic, mono, n, zk.C, zk.A, zk.B1, zk.B2)
    assert got[10] == cs + bytes(4)


def test_zkey_new_refusals_and_cli(pkg, small, tmp_path):
    d = small["dir"]
    td = small["td"]
    (tmp_path / "pot6.ptau").write_bytes(ZN.write_ptau(td.tau, td.alpha, td.beta, 6))
    with pytest.raises(pkg.G16Error, match="circuit too big for this power of tau ceremony"):
        pkg.zkey_new(d / "c.r1cs", tmp_path / "pot6.ptau", tmp_path / "x.zkey")
    assert not (tmp_path / "x.zkey").exists()
    # a ceremony file that was not prepared for phase 2
    raw = (d / "pot8.ptau").read_bytes()
    _, secs = binfile.read_binfile(raw, b"ptau")
    plain = binfile.write_binfile(b"ptau", 1, [(t, raw[v[0][0]:v[0][0] + v[0][1]]) for t, v in sorted(secs.items()) if t < 12])
    (tmp_path / "plain.ptau").write_bytes(plain)
    with pytest.raises(pkg.G16Error, match="not prepared for phase 2") as e:
        pkg.zkey_new(d / "c.r1cs", tmp_path / "plain.ptau", tmp_path / "x.zkey")
    assert e.value.code == pkg.G16_ERR_FORMAT
    with pytest.raises(pkg.G16Error):
        pkg.zkey_new(d / "pot8.ptau", d / "pot8.ptau", tmp_path / "x.zkey")            # not an .r1cs
    bad = bytearray((d / "c.r1cs").read_bytes())
    bad = bytes(bad[:-200])                                                             # truncated constraints
    (tmp_path / "bad.r1cs").write_bytes(bad)
    with pytest.raises(pkg.G16Error):
        pkg.zkey_new(tmp_path / "bad.r1cs", d / "pot8.ptau", tmp_path / "x.zkey")
    # the command line, snarkjs' argument order
    exe = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")
    r = subprocess.run([exe, "--zkey-new", str(d / "c.r1cs"), str(d / "pot8.ptau"), str(tmp_path / "cli.zkey")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Circuit Hash" in r.stdout
    assert (tmp_path / "cli.zkey").read_bytes() == ZN.zkey_new((d / "c.r1cs").read_bytes(), raw)
    r = subprocess.run([exe, "--zkey-new", str(d / "c.r1cs")], capture_output=True, text=True)
    assert r.returncode == 2


@pytest.mark.parametrize("n_vars,n_public,n_cons", [(30000, 1, 32000), (2083343, 1, 2094670)])
def test_zkey_new_at_scale_agrees_with_the_fixture_generator_and_proves(pkg, tmp_path, n_vars, n_public, n_cons):
    """2^15, and BASELINE's c2-layer-one shape (2^21): sections 4-7 of the new key equal the known-trapdoor generator's
    (they do not depend on gamma, delta), and a proof made with the new key verifies against its exported vkey -- which
    pins IC, C and H (a wrong point in any of them breaks the pairing equation)."""
    import time
    sy = pkg.Synth(n_vars, n_public, n_cons, 11, 12)
    sy.write_setup_inputs(tmp_path / "c.r1cs", tmp_path / "pot.ptau")
    t0 = time.perf_counter()
    pkg.zkey_new(tmp_path / "c.r1cs", tmp_path / "pot.ptau", tmp_path / "new.zkey")
    dt = time.perf_counter() - t0
    ref, wtns = sy.write()
    ref = bytes(ref)
    got = (tmp_path / "new.zkey").read_bytes()
    _, sa = binfile.read_binfile(got, b"zkey")
    _, sb = binfile.read_binfile(ref, b"zkey")
    for t in (1, 4, 5, 6, 7):
        (pa, la), (pb, lb) = sa[t][0], sb[t][0]
        assert la == lb and got[pa:pa + la] == ref[pb:pb + lb], t
    for t in (3, 8, 9):
        assert sa[t][0][1] == sb[t][0][1]
    vkey = pkg.zkey_export_vkey(tmp_path / "new.zkey")
    with pkg.ZKey(got) as zk:
        proof, public = zk.prove(wtns)
    assert pkg.verify(vkey, public, proof)
    # the rest of the reference's setup on the new key: contribute, then `zkey verify` of the result against circuit + ceremony
    del got, ref
    pkg.zkey_contribute(tmp_path / "new.zkey", tmp_path / "c1.zkey", name="first", entropy="e")
    t0 = time.perf_counter()
    assert pkg.zkey_verify(tmp_path / "c.r1cs", tmp_path / "pot.ptau", tmp_path / "c1.zkey")
    dv = time.perf_counter() - t0
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "zkey_new_%d.json" % n_cons), "w") as f:
            json.dump({"n_vars": n_vars, "n_public": n_public, "n_constraints": n_cons, "zkey_new_wall_s": dt,
                       "zkey_verify_wall_s": dv, "zkey_bytes": os.path.getsize(tmp_path / "new.zkey")}, f)


def test_zkey_verify_agrees_with_the_oracle(pkg, small, tmp_path):
    """`zkey verify` (scripts/g16_verify.sh:190-202): the fresh key, a chain of this library's contributions and a chain made
    by the oracle all pass, in the library and in the oracle; broken keys are refused with the check that failed"""
    import random
    from oracle import setup as S
    d = small["dir"]
    r1cs, ptau = d / "c.r1cs", d / "pot8.ptau"
    pkg.zkey_new(r1cs, ptau, tmp_path / "k0.zkey")
    assert pkg.zkey_verify(r1cs, ptau, tmp_path / "k0.zkey")
    pkg.zkey_contribute(tmp_path / "k0.zkey", tmp_path / "k1.zkey", name="first", entropy="abc")
    pkg.zkey_beacon(tmp_path / "k1.zkey", tmp_path / "k2.zkey", "0102030405060708090a0b0c0d0e0f101112131415161718191a1b1c1d1e1f", 4, name="final")
    assert pkg.zkey_verify(r1cs, ptau, tmp_path / "k2.zkey")
    init = (tmp_path / "k0.zkey").read_bytes()
    k2 = (tmp_path / "k2.zkey").read_bytes()
    assert S.verify_from_init(init, k2, random.Random(1)) is None            # the oracle accepts the library's chain
    ok, _ = S.contribute(init, "oracle", b"oracle entropy")
    (tmp_path / "o1.zkey").write_bytes(ok)
    assert pkg.zkey_verify(r1cs, ptau, tmp_path / "o1.zkey")                  # and the library the oracle's
    secs = sections(k2)

    def rebuilt(changes, name):
        out = bytearray(b"zkey" + (1).to_bytes(4, "little") + (10).to_bytes(4, "little"))
        for t in range(1, 11):
            body = changes.get(t, secs[t])
            out += t.to_bytes(4, "little") + len(body).to_bytes(8, "little") + body
        (tmp_path / name).write_bytes(bytes(out))
        return tmp_path / name
    s8 = bytearray(secs[8])
    s8[0:64], s8[64:128] = s8[64:128], s8[0:64]
    with pytest.raises(pkg.G16Error, match=r"section 8 \(C\) is not the initial section divided by delta"):
        pkg.zkey_verify(r1cs, ptau, rebuilt({8: bytes(s8)}, "b8.zkey"))
    s9 = bytearray(secs[9])
    s9[64:128] = s9[0:64]
    with pytest.raises(pkg.G16Error, match=r"section 9 \(H\)"):
        pkg.zkey_verify(r1cs, ptau, rebuilt({9: bytes(s9)}, "b9.zkey"))
    s5 = bytearray(secs[5])
    s5[0:64] = s5[64:128]
    with pytest.raises(pkg.G16Error, match=r"section 5 \(A\) differs"):
        pkg.zkey_verify(r1cs, ptau, rebuilt({5: bytes(s5)}, "b5.zkey"))
    s10 = bytearray(secs[10])
    s10[-1] ^= 1
    with pytest.raises(pkg.G16Error, match="not the beacon it claims to be"):
        pkg.zkey_verify(r1cs, ptau, rebuilt({10: bytes(s10)}, "b10.zkey"))
    s10 = bytearray(secs[10])
    s10[68 + 64 + 5] ^= 1                              # first record's g1_s: no longer what the transcript hashed
    with pytest.raises(pkg.G16Error, match="contribution 1"):
        pkg.zkey_verify(r1cs, ptau, rebuilt({10: bytes(s10)}, "b10b.zkey"))
    other = tmp_path / "other.zkey"
    pkg.zkey_contribute(tmp_path / "k0.zkey", other, name="first", entropy="another")
    so = sections(other.read_bytes())
    with pytest.raises(pkg.G16Error, match="delta"):
        pkg.zkey_verify(r1cs, ptau, rebuilt({10: sections((tmp_path / "k1.zkey").read_bytes())[10], 2: so[2], 8: so[8], 9: so[9]}, "bd.zkey"))
    exe = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")
    r = subprocess.run([exe, "--zkey-verify", str(r1cs), str(ptau), str(tmp_path / "k2.zkey")], capture_output=True, text=True)
    assert r.returncode == 0 and "ZKey Ok!" in r.stdout, r.stderr
    r = subprocess.run([exe, "--zkey-verify", str(r1cs), str(ptau), str(tmp_path / "b8.zkey")], capture_output=True, text=True)
    assert r.returncode == 1 and "section 8" in r.stderr
