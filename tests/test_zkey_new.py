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


def test_zkey_verify_oracle_accepts_a_chain_of_contributions_and_names_what_is_broken(case):
    """`zkey verify` (scripts/g16_verify.sh:190-202) = the initial key recomputed from circuit + ceremony, then the chain of
    contributions and the two delta-scaled sections checked against it"""
    import random
    td = case["td"]
    init = ZN.zkey_new(case["r1cs"], ZN.write_ptau(td.tau, td.alpha, td.beta, 8))
    rng = random.Random(5)
    assert S.verify_from_init(init, init, rng) is None
    k1, _ = S.contribute(init, "first", b"entropy one")
    bh = bytes(range(1, 33))
    k2, _ = S.contribute(k1, "beacon", S.beacon_seed(bh, 3), beacon=(bh, 3))
    assert S.verify_from_init(init, k1, rng) is None
    assert S.verify_from_init(init, k2, rng) is None
    secs = sections(k2)

    def rebuilt(changes):
        out = bytearray(b"zkey" + (1).to_bytes(4, "little") + (10).to_bytes(4, "little"))
        for t in range(1, 11):
            body = changes.get(t, secs[t])
            out += t.to_bytes(4, "little") + len(body).to_bytes(8, "little") + body
        return bytes(out)
    # a point of section 8 swapped for another valid point
    s8 = bytearray(secs[8])
    s8[0:64], s8[64:128] = s8[64:128], s8[0:64]
    assert "section 8" in S.verify_from_init(init, rebuilt({8: bytes(s8)}), rng)
    s5 = bytearray(secs[5])
    s5[0:64] = s5[64:128]
    assert "section 5" in S.verify_from_init(init, rebuilt({5: bytes(s5)}), rng)
    # a contribution whose delta does not follow: splice k1's record into a key whose sections were scaled by another secret
    other, _ = S.contribute(init, "first", b"another entropy")
    so = sections(other)
    assert "delta" in S.verify_from_init(init, rebuilt({10: sections(k1)[10], 2: so[2], 8: so[8], 9: so[9]}), rng)
    # the beacon record claims another beacon hash
    s10 = bytearray(secs[10])
    s10[-1] ^= 1
    assert "beacon" in S.verify_from_init(init, rebuilt({10: bytes(s10)}), rng)
