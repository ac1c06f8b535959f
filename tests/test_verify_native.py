"""Native Groth16 verifier (g16_verify / `prover --verify`) pinned on the reference's own
vectors: every (vkey, public, proof) triple the reference commits must verify, tampered copies must not, and the
verdicts must agree with the Python oracle's pairing check.  Host code only: runs without a GPU."""
import copy
import json
import os
import subprocess

import pytest

from conftest import ROOT, load_pkg

EXE = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")


@pytest.fixture(scope="module")
def triples(golden):
    return golden["triples"]


def test_every_reference_triple_verifies(triples):
    pkg = load_pkg()
    assert len(triples) == 7
    for t in triples:
        assert pkg.verify(t["vkey"], t["public"], t["proof"]) is True, t["name"]


def test_tampered_inputs_are_rejected_like_the_oracle(triples):
    from oracle import verify as V   # checker: the pinned Python pairing
    pkg = load_pkg()
    t = triples[0]
    cases = []
    bad = copy.deepcopy(t)
    bad["public"][0] = str(int(bad["public"][0]) + 1)
    cases.append(bad)                                   # wrong public signal
    bad = copy.deepcopy(t)
    bad["proof"]["pi_a"], bad["proof"]["pi_c"] = bad["proof"]["pi_c"], bad["proof"]["pi_a"]
    cases.append(bad)                                   # valid curve points, wrong proof
    bad = copy.deepcopy(t)
    bad["proof"]["pi_b"] = triples[1]["proof"]["pi_b"]
    cases.append(bad)                                   # pi_b of another proof
    bad = copy.deepcopy(t)
    bad["proof"]["pi_c"][1] = str(int(bad["proof"]["pi_c"][1]) + 1)
    cases.append(bad)                                   # pi_c not on the curve
    bad = copy.deepcopy(t)
    bad["vkey"]["vk_delta_2"] = t["vkey"]["vk_gamma_2"]
    cases.append(bad)                                   # wrong key
    bad = copy.deepcopy(t)
    bad["public"] = bad["public"] + ["1"]
    cases.append(bad)                                   # public signal count != len(IC) - 1
    for c in cases:
        got = pkg.verify(c["vkey"], c["public"], c["proof"])
        assert got is False
        assert got == bool(V.verify(c["vkey"], c["public"], c["proof"]))


def test_other_triples_cross_checked_against_other_keys(triples):
    pkg = load_pkg()
    a, b = triples[0], triples[2]      # layer one and layer two: different circuits
    assert pkg.verify(b["vkey"], a["public"], a["proof"]) is False or len(a["public"]) != len(b["public"])
    assert pkg.verify(a["vkey"], a["public"], triples[1]["proof"]) is False


def test_malformed_json_is_an_error_not_a_verdict(triples):
    pkg = load_pkg()
    t = triples[0]
    with pytest.raises(pkg.G16Error) as e:
        pkg.verify("{not json", t["public"], t["proof"])
    assert e.value.code == pkg.G16_ERR_FORMAT
    with pytest.raises(pkg.G16Error):
        pkg.verify({"protocol": "groth16"}, t["public"], t["proof"])          # missing keys
    bad = copy.deepcopy(t["proof"])
    bad["pi_a"][0] = "12x"
    with pytest.raises(pkg.G16Error):
        pkg.verify(t["vkey"], t["public"], bad)


def test_snarkjs_and_rapidsnark_text_layouts_and_cli(triples, tmp_path):
    """byte layouts of both provers (indentation, key order, extra `curve` key) parse alike; CLI exit-status contract"""
    pkg = load_pkg()
    t = triples[5]                      # snarkjs-style fixture (experiments/scripts/groth16_input_prep)
    assert pkg.verify(json.dumps(t["vkey"], indent=1), json.dumps(t["public"], indent=1), json.dumps(t["proof"], indent=1))
    assert pkg.verify(json.dumps(t["vkey"], separators=(",", ":")), json.dumps(t["public"], separators=(",", ":")),
                      json.dumps(t["proof"], separators=(",", ":")))
    for k in ("vkey", "public", "proof"):
        (tmp_path / (k + ".json")).write_text(json.dumps(t[k]))
    p = subprocess.run([EXE, "--verify", str(tmp_path / "vkey.json"), str(tmp_path / "public.json"), str(tmp_path / "proof.json")],
                       capture_output=True, text=True)
    assert p.returncode == 0 and "OK!" in p.stdout
    pub = list(t["public"])
    pub[-1] = str(int(pub[-1]) ^ 1)
    (tmp_path / "bad.json").write_text(json.dumps(pub))
    p = subprocess.run([EXE, "--verify", str(tmp_path / "vkey.json"), str(tmp_path / "bad.json"), str(tmp_path / "proof.json")],
                       capture_output=True, text=True)
    assert p.returncode == 1 and "Invalid proof" in p.stderr
    p = subprocess.run([EXE, "--verify", str(tmp_path / "nope.json"), str(tmp_path / "bad.json"), str(tmp_path / "proof.json")],
                       capture_output=True, text=True)
    assert p.returncode != 0 and "cannot open" in p.stderr


def test_verifies_proofs_of_the_oracle_prover(small_fixture):
    """a proof made by the Python oracle on a SYNTH-1 circuit (fresh randomness) passes the native verifier"""
    from oracle import groth16 as g16
    pkg = load_pkg()
    f = small_fixture
    proof, public = g16.prove(f["zk"], f["w"], 987654321, 123456789)
    assert pkg.verify(f["vkey"], public, proof) is True
    pub = json.loads(public)
    pub[0] = str(int(pub[0]) + 1)
    assert pkg.verify(f["vkey"], pub, proof) is False


def test_sanitize_is_byte_identical_to_the_reference_files(golden, tmp_path):
    """scripts/sanitize_groth16_proof.py outputs committed by the reference (the two cases that also commit the vkey):
    same dict, and the same bytes Python's json.dump wrote (e(-alpha, beta) limbs included)."""
    pkg = load_pkg()
    from oracle import verify as V
    n = 0
    for it in golden["sanitized"]:
        if it["vkey"] is None:
            continue
        text = pkg.sanitize(it["vkey"], it["public"], it["proof"])
        assert json.loads(text) == it["sanitized"], it["name"]
        assert text == json.dumps(it["sanitized"])          # the layout of json.dump(fullCircomInput, outfile)
        assert json.loads(text) == V.sanitize(it["vkey"], it["public"], it["proof"])
        n += 1
    assert n == 2
    it = [x for x in golden["sanitized"] if x["vkey"] is not None][0]
    for k in ("vkey", "public", "proof"):
        (tmp_path / (k + ".json")).write_text(json.dumps(it[k]))
    p = subprocess.run([EXE, "--sanitize", str(tmp_path / "vkey.json"), str(tmp_path / "public.json"),
                        str(tmp_path / "proof.json"), str(tmp_path / "sanitized_proof.json")], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert (tmp_path / "sanitized_proof.json").read_text() == json.dumps(it["sanitized"])


def test_sanitize_of_every_reference_triple_matches_the_oracle(triples):
    from oracle import verify as V
    pkg = load_pkg()
    for t in triples[:4]:
        assert json.loads(pkg.sanitize(t["vkey"], t["public"], t["proof"])) == V.sanitize(t["vkey"], t["public"], t["proof"])
