"""Phase-2 ceremony bookkeeping and vkey export (file level): host-only parts.
The hashes against hashlib, the vkey layout against the reference's committed verification keys, and the native
exporter against the oracle (oracle/setup.py) on a SYNTH-1 key made on the host cores."""
import base64
import hashlib
import json
import os

import pytest

from oracle import bn254 as bn, cport, setup as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_native_hashes_match_hashlib(pkg):
    for n in (0, 1, 63, 64, 127, 128, 129, 255, 256, 1000):
        data = bytes((i * 7 + n) & 255 for i in range(n))
        assert pkg.test_hash(0, data) == hashlib.blake2b(data, digest_size=64).digest()
        assert pkg.test_hash(1, data) == hashlib.sha256(data).digest()


def test_vkey_layout_is_the_references(pkg):
    """json.dumps(indent=1) of a parsed reference vkey reproduces the committed bytes: the layout the exporters write;
    and vk_alphabeta_12 of those keys is e(alpha, beta) in the tower basis the oracle's exporter uses"""
    with open(os.path.join(ROOT, "tests", "golden", "vkey_layout.json")) as f:
        files = json.load(f)["files"]
    for item in files:
        raw = base64.b64decode(item["b64"]).decode()
        vk = json.loads(raw)
        assert json.dumps(vk, indent=1) == raw.rstrip("\n")
        assert list(vk) == ["protocol", "curve", "nPublic", "vk_alpha_1", "vk_beta_2", "vk_gamma_2", "vk_delta_2", "vk_alphabeta_12", "IC"]
        alpha = (int(vk["vk_alpha_1"][0]), int(vk["vk_alpha_1"][1]))
        beta = tuple((int(c[0]), int(c[1])) for c in vk["vk_beta_2"][:2])
        ab = bn.pairing_snarkjs(beta, alpha)
        tower = [[[str((ab[2 * i + j] + 9 * ab[2 * i + j + 6]) % bn.Q), str(ab[2 * i + j + 6] % bn.Q)] for i in range(3)] for j in range(2)]
        assert tower == vk["vk_alphabeta_12"]


@pytest.mark.parametrize("n_public", [1, 13])
def test_native_vkey_export_matches_oracle(pkg, tmp_path, n_public):
    sy = cport.Synth(200, n_public, 210, 3, 4)
    zk, _ = sy.write()
    (tmp_path / "k.zkey").write_bytes(zk)
    text = pkg.zkey_export_vkey(str(tmp_path / "k.zkey"))
    assert text == S.export_vkey(bytes(zk))
    vk = json.loads(text)
    assert {k: vk[k] for k in sy.vkey()} == sy.vkey()
    (tmp_path / "bad.zkey").write_bytes(bytes(zk)[:1000])
    with pytest.raises(pkg.G16Error):
        pkg.zkey_export_vkey(str(tmp_path / "bad.zkey"))


def test_hash_to_g2_matches_oracle_and_lands_in_the_subgroup(pkg):
    for tr in (bytes(range(64)), bytes(64), b"\xff" * 64):
        p = S.hash_to_g2(tr)
        assert pkg.test_hash(2, tr) == S._g2_lem(p)
        assert bn.G2.to_affine(bn.G2.jmul(bn.G2.to_jac(p), bn.R)) is None     # order r
        x, y = p
        assert bn.f2_mul(y, y) == bn.f2_add(bn.f2_mul(bn.f2_mul(x, x), x), bn.B2)   # on the twist
