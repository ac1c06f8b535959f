"""The prove restatement against its two independent checks: the pinned verifier and the closed-form
known-trapdoor proof; plus container codecs."""
import json

import pytest

from oracle import binfile, bn254 as bn, factory, groth16 as g16, verify as V


def test_prove_equals_closed_form_and_verifies(small_fixture):
    f = small_fixture
    r, s = 0x1234567890ABCDEF << 100, 987654321987654321
    proof, public = g16.prove(f["zk"], f["w"], r, s)
    pa, pb, pc = factory.closed_form_proof(f["circ"], f["w"], f["td"], r, s)
    assert json.loads(proof) == g16.proof_object(pa, pb, pc)
    assert V.verify(f["vkey"], json.loads(public), json.loads(proof))
    assert json.loads(public) == [str(x) for x in f["w"][1:3]]


def test_prove_blinding_changes_proof_not_validity(small_fixture):
    f = small_fixture
    p1, u1 = g16.prove(f["zk"], f["w"], 1, 2)
    p2, u2 = g16.prove(f["zk"], f["w"], 3, 4)
    assert p1 != p2 and u1 == u2
    assert V.verify(f["vkey"], json.loads(u2), json.loads(p2))


def test_wrong_witness_does_not_verify(small_fixture):
    f = small_fixture
    w = list(f["w"])
    w[40] = (w[40] + 1) % bn.R
    p, u = g16.prove(f["zk"], w, 5, 6)
    assert not V.verify(f["vkey"], json.loads(u), json.loads(p))


def test_tiny_handmade_circuit():
    # x * x = y ; y * x = z ; (z + x + 5) * 1 = out   with x = 3 -> out = 35 (the circom tutorial cubic)
    cons = [([(2, 1)], [(2, 1)], [(3, 1)]), ([(3, 1)], [(2, 1)], [(4, 1)]),
            ([(4, 1), (2, 1), (0, 5)], [(0, 1)], [(1, 1)])]
    circ = factory.ListCircuit(5, 1, cons, [1, 35, 3, 9, 27])
    assert circ.check_witness(circ.witness())
    td = factory.Trapdoor(77)
    zk, vkey = factory.setup(circ, td)
    proof, public = g16.prove(zk, circ.witness(), 1111, 2222)
    assert json.loads(public) == ["35"]
    assert V.verify(vkey, ["35"], json.loads(proof))
    assert not V.verify(vkey, ["36"], json.loads(proof))
    pa, pb, pc = factory.closed_form_proof(circ, circ.witness(), td, 1111, 2222)
    assert json.loads(proof) == g16.proof_object(pa, pb, pc)


def test_binfile_roundtrip(small_fixture):
    f = small_fixture
    zk2 = binfile.read_zkey(f["zkey_bytes"])
    for k in ("n_vars", "n_public", "domain_size", "alpha1", "beta1", "beta2", "gamma2", "delta1", "delta2", "ic",
              "coefs", "A", "B1", "B2", "C", "H"):
        assert getattr(zk2, k) == getattr(f["zk"], k), k
    w2 = binfile.read_wtns(f["wtns_bytes"])
    assert w2["witness"] == f["w"] and w2["prime"] == bn.R and w2["n8"] == 32
    assert binfile.write_zkey(zk2) == f["zkey_bytes"]


def test_zero_points_and_zero_scalars_in_msm():
    pts = [bn.G1.mul(bn.G1_GEN, k) for k in (1, 2, 3)] + [None]
    assert bn.G1.msm(pts, [0, 0, 0, 5]) is None
    assert bn.G1.msm(pts, [1, 0, 2, 9]) == bn.G1.mul(bn.G1_GEN, 7)
    assert bn.G1.msm([], []) is None


def test_ntt_matches_naive_dft():
    n, w = 16, g16.root_of_unity(4)
    a = [(i * i * 7919 + 3) % bn.R for i in range(n)]
    want = [sum(a[j] * pow(w, i * j, bn.R) for j in range(n)) % bn.R for i in range(n)]
    assert g16.ntt(a, w) == want
    assert g16.intt(want, w) == a
