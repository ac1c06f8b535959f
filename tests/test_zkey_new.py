"""`zkey new` oracle (oracle/zkey_new.py, point arithmetic only) against the known-trapdoor factory: with gamma = delta = 1
the factory's closed form IS what `zkey new` computes, so every section except the circuit hash must agree byte for byte;
and the key must prove and verify.  CPU only."""
import json

import pytest

from oracle import binfile, bn254 as bn, factory, groth16 as g16, verify as V, zkey_new as ZN
from oracle import setup as S


@pytest.fixture(scope="module")
def case():
    circ = factory.SynthCircuit(n_vars=85, n_public=2, n_constraints=100, seed=0xB2000001)
    td = factory.Trapdoor(0xB2000002)
    td.gamma = td.delta = 1
    zk, vkey = factory.setup(circ, td)
    return {"circ": circ, "td": td, "zk": zk, "vkey": vkey, "want": binfile.write_zkey(zk), "r1cs": ZN.write_r1cs(circ, n_pub_out=1)}


def sections(data):
    _, secs = binfile.read_binfile(data, b"zkey")
    return {t: bytes(data[v[0][0]:v[0][0] + v[0][1]]) for t, v in secs.items()}


def test_r1cs_round_trip(case):
    cir = ZN.read_r1cs(case["r1cs"])
    circ = case["circ"]
    assert (cir["n_vars"], cir["n_public"], cir["n_constraints"]) == (circ.n_vars, circ.n_public, circ.n_constraints)
    for c in range(circ.n_constraints):
        A, B, C = circ.constraint(c)
        assert cir["constraints"][c] == ([(s, v % bn.R) for s, v in A], [(s, v % bn.R) for s, v in B], [(s, v % bn.R) for s, v in C])
    with pytest.raises(ValueError):
        ZN.read_r1cs(case["r1cs"][:-1])


def test_zkey_new_equals_the_closed_form_with_a_larger_ceremony(case):
    td = case["td"]
    got = ZN.zkey_new(case["r1cs"], ZN.write_ptau(td.tau, td.alpha, td.beta, 8))      # circuit domain 2^7
    a, b = sections(got), sections(case["want"])
    for t in range(1, 10):
        assert a[t] == b[t], t
    assert a[10][64:] == bytes(4) and a[10][:64] != bytes(64)
    # the vkey exported from the new key is the factory's
    exported = json.loads(S.export_vkey(got))
    exported.pop("vk_alphabeta_12")
    assert exported == case["vkey"]


def test_zkey_new_with_a_ceremony_of_the_circuits_own_power_proves_and_verifies(case):
    """power == circuit power: the ceremony's level power+1 lacks the top power of tau, so section 9 differs from the
    closed form -- and must not matter: h Z has degree 2n - 2"""
    td, circ = case["td"], case["circ"]
    got = ZN.zkey_new(case["r1cs"], ZN.write_ptau(td.tau, td.alpha, td.beta, 7))
    a, b = sections(got), sections(case["want"])
    for t in (1, 2, 3, 4, 5, 6, 7, 8):
        assert a[t] == b[t], t
    assert a[9] != b[9]
    zk = binfile.read_zkey(got)
    proof, public = g16.prove(zk, circ.witness(), 1234567, 7654321)
    assert V.verify(case["vkey"], json.loads(public), json.loads(proof))
    with pytest.raises(ValueError, match="circuit too big"):
        ZN.zkey_new(case["r1cs"], ZN.write_ptau(td.tau, td.alpha, td.beta, 6))
