"""`zkey contribute` / `zkey beacon` on the GPU through the C ABI and the CLI: byte-for-byte
against the oracle (oracle/setup.py), and the contract that matters to the workflow (scripts/g16_setup.sh:254-295):
the contributed key proves, and its proofs verify under the verification key exported from it and not under the old one."""
import json
import os
import subprocess
import time

import pytest

from oracle import cport, setup as S, verify as V   # checker only

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BEACON = "0102030405060708090a0b0c0d0e0f101112231415161718221a1b1c1d1e1f"   # scripts/g16_setup.sh:275


def test_contribute_and_beacon_match_oracle_and_keep_proving(pkg, tmp_path):
    sy = cport.Synth(300, 2, 320, 21, 22)
    zk0, wt = (bytes(x) for x in sy.write())
    p0, p1, p2 = (str(tmp_path / n) for n in ("c_0.zkey", "c_1.zkey", "c_final.zkey"))
    open(p0, "wb").write(zk0)
    h1 = pkg.zkey_contribute(p0, p1, name="First contributor", entropy="random text for entropy")
    want1, wh1 = S.contribute(zk0, "First contributor", b"random text for entropy")
    got1 = open(p1, "rb").read()
    assert got1 == want1 and h1 == wh1
    h2 = pkg.zkey_beacon(p1, p2, BEACON, 10, name="Final Beacon phase2")
    want2, wh2 = S.contribute(want1, "Final Beacon phase2", S.beacon_seed(bytes.fromhex(BEACON), 10), beacon=(bytes.fromhex(BEACON), 10))
    got2 = open(p2, "rb").read()
    assert got2 == want2 and h2 == wh2
    cs, recs = S.parse_mpc(S.parse_sections(got2)[10])
    assert cs == bytes(64) and len(recs) == 2 and recs[1]["params"][:2] == bytes([1, 19]) and recs[1]["params"][21:23] == bytes([2, 10])
    vk0, vk2 = json.loads(pkg.zkey_export_vkey(p0)), json.loads(pkg.zkey_export_vkey(p2))
    assert vk0["vk_delta_2"] != vk2["vk_delta_2"] and vk0["vk_alpha_1"] == vk2["vk_alpha_1"] and vk0["IC"] == vk2["IC"]
    with pkg.ZKey(p2) as zk:
        proof, public = zk.prove(wt)
    assert V.verify(vk2, json.loads(public), json.loads(proof)) and pkg.verify(vk2, public, proof)
    assert not V.verify(vk0, json.loads(public), json.loads(proof))
    # fresh entropy (getrandom): two contributions to the same key differ, both keep proving
    pa, pb = str(tmp_path / "a.zkey"), str(tmp_path / "b.zkey")
    assert pkg.zkey_contribute(p0, pa) != pkg.zkey_contribute(p0, pb)
    assert open(pa, "rb").read() != open(pb, "rb").read()


def test_setup_cli(pkg, tmp_path):
    sy = cport.Synth(85, 1, 100, 5, 6)
    zk0, wt = (bytes(x) for x in sy.write())
    p0, p1, vk = (str(tmp_path / n) for n in ("k_0.zkey", "k_final.zkey", "vkey.json"))
    open(p0, "wb").write(zk0)
    open(tmp_path / "w.wtns", "wb").write(wt)
    exe = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")
    r = subprocess.run([exe, "--zkey-contribute", p0, p1, "--name=First contributor", "-e=random text for entropy"], capture_output=True, text=True)
    assert r.returncode == 0 and "Contribution Hash" in r.stdout, r.stderr
    assert open(p1, "rb").read() == S.contribute(zk0, "First contributor", b"random text for entropy")[0]
    r = subprocess.run([exe, "--zkey-export-vkey", p1, vk], capture_output=True, text=True)
    assert r.returncode == 0 and open(vk).read() == S.export_vkey(open(p1, "rb").read())
    r = subprocess.run([exe, p1, str(tmp_path / "w.wtns"), str(tmp_path / "proof.json"), str(tmp_path / "public.json")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe, "--verify", vk, str(tmp_path / "public.json"), str(tmp_path / "proof.json")], capture_output=True, text=True)
    assert r.returncode == 0 and "OK" in r.stdout
    r = subprocess.run([exe, "--zkey-contribute", str(tmp_path / "missing.zkey"), p1], capture_output=True, text=True)
    assert r.returncode != 0 and "cannot open" in r.stderr


def test_contribute_at_baseline_size(pkg, tmp_path):
    """c2-layer-one (1.08 GB key, 2 M + 2 M points in sections 8 and 9): the contributed key proves and verifies"""
    import bench
    zkey, wtns = bench.fixture("c2-layer-one")
    zp = bench.fixture_paths("c2-layer-one")[0]
    out = str(tmp_path / "contributed.zkey")
    t = time.perf_counter()
    pkg.zkey_contribute(zp, out, name="First contributor", entropy="random text for entropy")
    dt = time.perf_counter() - t
    print("zkey contribute, c2-layer-one: %.2f s wall" % dt)
    vk = json.loads(pkg.zkey_export_vkey(out))
    with pkg.ZKey(out) as zk:
        proof, public = zk.prove(wtns)
    assert V.verify(vk, json.loads(public), json.loads(proof))
    with open(bench.fixture_paths("c2-layer-one")[2]) as f:
        assert not V.verify(json.load(f), json.loads(public), json.loads(proof))
    os.unlink(out)
