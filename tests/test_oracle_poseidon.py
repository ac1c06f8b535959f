"""The Poseidon / Merkle oracle (oracle/poseidon.py) against the reference's own artefacts (tests/golden/merkle_vectors.json,
copied from the reference tree by tests/golden/make_merkle_vectors.py)."""
import json
import os

import pytest

from oracle import poseidon as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def vec():
    with open(os.path.join(ROOT, "tests", "golden", "merkle_vectors.json")) as f:
        return json.load(f)


def test_grain_parameters_shape():
    consts, mds = P.params()
    assert len(consts) == (P.RF + P.RP) * P.T and len(mds) == 3 and all(len(r) == 3 for r in mds)
    assert all(0 <= c < P.R for c in consts)
    # circomlib's published Poseidon(2) known answer
    assert P.hash2(1, 2) == 0x115cc0f5e7d690413df64c6b9662e9cf2a3617f2743245519e19607a4417189a


def test_root_of_the_reference_anonymity_set(vec):
    v = vec["anon_set_10"]
    rows = P.parse_anon_set(v["csv"])
    assert len(rows) == v["leaves"]
    leaves = P.build_leaves(rows)
    assert len(leaves) == v["padded_leaves"] and leaves[len(rows):] == [0] * (len(leaves) - len(rows))
    levels = P.build_tree(leaves)
    assert str(levels[-1][0]) == v["root"]           # the value merkle_tree.rs logged and wrote to merkle_root.json
    root_txt, _ = P.merkle_files(rows, [])
    assert root_txt == '{\n  "__bigint__": "%s"\n}' % v["root"]


def test_merkle_paths_the_reference_commits(vec):
    v = vec["layer_two_input"]
    for a, b, el, idx in zip(v["leaf_addresses"], v["leaf_balances"], v["path_elements"], v["path_indices"]):
        leaf = P.hash2(int(a), int(b))
        assert str(P.root_from_proof(leaf, [int(e) for e in el], idx)) == v["merkle_root"]
    e = vec["experiments_merkle_inputs"]
    assert str(P.root_from_proof(int(e["leaf"]), [int(x) for x in e["pathElements"]], e["pathIndices"])) == e["root"]


def test_proofs_and_file_layout(vec):
    rows = P.parse_anon_set(vec["anon_set_10"]["csv"])
    owned = [rows[3], rows[7]]
    root_txt, proofs_txt = P.merkle_files(rows, owned)
    d = json.loads(proofs_txt)
    assert [int(x["address"]["__bigint__"]) for x in d["leaves"]] == [rows[3][0], rows[7][0]]
    assert d["path_indices"] == [[1, 1, 0, 0], [1, 1, 1, 0]]
    root = int(json.loads(root_txt)["__bigint__"])
    for leaf, el, idx in zip(d["leaves"], d["path_elements"], d["path_indices"]):
        assert P.root_from_proof(int(leaf["hash"]["__bigint__"]), [int(e["__bigint__"]) for e in el], idx) == root
    assert proofs_txt.startswith('{\n  "leaves": [\n    {\n      "address": {\n        "__bigint__": "')
    with pytest.raises(ValueError):
        P.merkle_files(rows, [rows[7], rows[3]])     # the reference scans forward only: out-of-order owners are an error
