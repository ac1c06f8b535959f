"""`wtns check` on the GPU (g16_wtns_check, csrc/wtns_check.cu) through the C ABI and the CLI against the oracle
(oracle/wtns_check.py): same verdict, same first failing constraint."""
import os
import subprocess

import pytest

from oracle import binfile, bn254 as bn, factory, wtns_check as WC, zkey_new as ZN

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")


@pytest.mark.parametrize("n_vars,n_public,n_cons", [(85, 2, 100), (800, 5, 1000)])
def test_wtns_check_matches_the_oracle(pkg, tmp_path, n_vars, n_public, n_cons):
    circ = factory.SynthCircuit(n_vars=n_vars, n_public=n_public, n_constraints=n_cons, seed=0xB2000001)
    r1cs = ZN.write_r1cs(circ, n_pub_out=1)
    (tmp_path / "c.r1cs").write_bytes(r1cs)
    w = circ.witness()
    (tmp_path / "w.wtns").write_bytes(binfile.write_wtns(w))
    assert pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "w.wtns") is None
    for wire in (0, 1, n_public + 3, n_vars // 2, n_vars - 1):
        bad = list(w)
        bad[wire] = (bad[wire] + 5) % bn.R
        if wire == 0:
            continue                                  # wire 0 must stay the constant 1: refused below
        raw = binfile.write_wtns(bad)
        (tmp_path / "bad.wtns").write_bytes(raw)
        want = WC.wtns_check(r1cs, raw)
        assert want is not None
        assert pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "bad.wtns") == want
    # several broken wires: still the lowest failing constraint
    bad = list(w)
    for wire in (n_vars - 1, n_vars // 3, n_public + 1):
        bad[wire] = (bad[wire] * 3 + 7) % bn.R
    raw = binfile.write_wtns(bad)
    (tmp_path / "bad.wtns").write_bytes(raw)
    assert pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "bad.wtns") == WC.wtns_check(r1cs, raw)


def test_wtns_check_refusals_and_cli(pkg, tmp_path):
    circ = factory.SynthCircuit(n_vars=85, n_public=2, n_constraints=100, seed=0xB2000001)
    (tmp_path / "c.r1cs").write_bytes(ZN.write_r1cs(circ))
    w = circ.witness()
    (tmp_path / "w.wtns").write_bytes(binfile.write_wtns(w))
    (tmp_path / "short.wtns").write_bytes(binfile.write_wtns(w[:-1]))
    with pytest.raises(pkg.G16Error, match="Invalid witness length. Circuit: 85, witness: 84") as e:
        pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "short.wtns")
    assert e.value.code == pkg.G16_ERR_WITNESS_LEN
    w0 = list(w)
    w0[0] = 2
    (tmp_path / "w0.wtns").write_bytes(binfile.write_wtns(w0))
    with pytest.raises(pkg.G16Error, match="wire 0"):
        pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "w0.wtns")
    with pytest.raises(pkg.G16Error):
        pkg.wtns_check(tmp_path / "w.wtns", tmp_path / "w.wtns")              # not an .r1cs
    r = subprocess.run([EXE, "--wtns-check", str(tmp_path / "c.r1cs"), str(tmp_path / "w.wtns")], capture_output=True, text=True)
    assert r.returncode == 0 and "WITNESS IS CORRECT" in r.stdout, r.stderr
    bad = list(w)
    bad[40] = (bad[40] + 2) % bn.R
    (tmp_path / "bad.wtns").write_bytes(binfile.write_wtns(bad))
    r = subprocess.run([EXE, "--wtns-check", str(tmp_path / "c.r1cs"), str(tmp_path / "bad.wtns")], capture_output=True, text=True)
    assert r.returncode == 1 and "WITNESS CHECKING FAILED" in r.stdout
    assert "constraint %d " % WC.wtns_check((tmp_path / "c.r1cs").read_bytes(), (tmp_path / "bad.wtns").read_bytes()) in r.stdout


def test_wtns_check_at_baseline_size(pkg, tmp_path):
    """c2-layer-one's shape (2.09 M constraints): the generator's witness passes; one changed wire is caught at a constraint
    that reads it"""
    import time
    sy = pkg.Synth(2083343, 1, 2094670, 11, 12)
    sy.write_setup_inputs(tmp_path / "c.r1cs", tmp_path / "pot.ptau")
    os.remove(tmp_path / "pot.ptau")
    _, wtns = sy.write(want_zkey=False)
    wtns = bytearray(wtns)
    (tmp_path / "w.wtns").write_bytes(wtns)
    t0 = time.perf_counter()
    assert pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "w.wtns") is None
    dt = time.perf_counter() - t0
    wire = 2 + 20 * 61728 + 12                            # a "small" wire (class by (i - nPublic - 1) % 20): not a bit, whose 0 <-> 1 flip stays valid
    at = len(wtns) - (2083343 - wire) * 32
    wtns[at + 1] ^= 1
    (tmp_path / "bad.wtns").write_bytes(wtns)
    bad = pkg.wtns_check(tmp_path / "c.r1cs", tmp_path / "bad.wtns")
    assert bad is not None and bad >= wire - 1            # SYNTH-1: wire i is defined by constraint i - 1, read only later
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "wtns_check_2094670.txt"), "w") as f:
            f.write("wtns check, 2094670 constraints, file to verdict: %.3f s\n" % dt)
