"""Poseidon Merkle tree on the GPU through the C ABI, against the oracle (oracle/poseidon.py, pinned
on the reference's artefacts) and against those artefacts directly (tests/golden/merkle_vectors.json)."""
import json
import os
import random
import subprocess

import pytest

from oracle import poseidon as P   # checker only

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = P.R


@pytest.fixture(scope="module")
def vec():
    with open(os.path.join(ROOT, "tests", "golden", "merkle_vectors.json")) as f:
        return json.load(f)


def test_poseidon_hash_matches_oracle(pkg):
    rng = random.Random(5)
    pairs = [(0, 0), (1, 2), (R - 1, R - 1), (0, R - 1), (1 << 160, 10 ** 26)] + [(rng.randrange(R), rng.randrange(R)) for _ in range(300)]
    assert pkg.poseidon_hash2(pairs) == [P.hash2(a, b) for a, b in pairs]
    assert pkg.poseidon_hash2([]) == []
    with pytest.raises(pkg.G16Error):
        pkg.poseidon_hash2([(R, 1)])              # light_poseidon rejects inputs that are not field elements


def test_reference_anonymity_set_root_and_committed_paths(pkg, vec):
    v = vec["anon_set_10"]
    rows = P.parse_anon_set(v["csv"])
    with pkg.MerkleTree(rows) as t:
        assert (t.n_rows, t.n_leaves, t.levels) == (v["leaves"], v["padded_leaves"], 5)   # "height 5" in the reference's log
        assert str(t.root()) == v["root"]
        levels = P.build_tree(P.build_leaves(rows))
        for lv, nodes in enumerate(levels):
            assert [t.node(lv, i) for i in range(len(nodes))] == nodes
        for i in range(t.n_leaves):
            assert t.proof(i) == P.proof(levels, i)
        assert t.find(levels[0][7]) == 7 and t.find(levels[0][7], 8) is None and t.find(12345) is None
        assert t.find(0) == 10                                   # first padding leaf
    l2 = vec["layer_two_input"]
    leaf = pkg.poseidon_hash2([(int(l2["leaf_addresses"][0]), int(l2["leaf_balances"][0]))])[0]
    cur = leaf
    for e, i in zip(l2["path_elements"][0], l2["path_indices"][0]):
        cur = pkg.poseidon_hash2([(int(e), cur) if i else (cur, int(e))])[0]
    assert str(cur) == l2["merkle_root"]
    ex = vec["experiments_merkle_inputs"]
    cur = int(ex["leaf"])
    for e, i in zip(ex["pathElements"], ex["pathIndices"]):
        cur = pkg.poseidon_hash2([(int(e), cur) if i else (cur, int(e))])[0]
    assert str(cur) == ex["root"]


@pytest.mark.parametrize("n", [1, 2, 3, 1000, 4097])
def test_ragged_sizes_match_oracle(pkg, n):
    rng = random.Random(n)
    rows = [(rng.randrange(1 << 160), rng.randrange(10 ** 24)) for _ in range(n)]
    leaves = P.build_leaves(rows)
    levels = P.build_tree(leaves)
    with pkg.MerkleTree(rows) as t:
        assert t.n_leaves == len(leaves) and t.levels == len(levels)
        assert t.root() == levels[-1][0]
        for i in {0, n - 1, len(leaves) - 1, n // 2}:
            assert t.proof(i) == P.proof(levels, i)


def test_merkle_tree_cli_writes_the_reference_files(pkg, vec, tmp_path):
    rows = P.parse_anon_set(vec["anon_set_10"]["csv"])
    (tmp_path / "anon.csv").write_text(vec["anon_set_10"]["csv"])
    owned = [rows[3], rows[7]]
    poa = {"accountAttestations": [{"signature": {"r": {"__bigint__": "1"}}, "accountData": {"address": {"__bigint__": str(a)}, "balance": {"__bigint__": str(b)}}}
                                   for a, b in owned]}
    (tmp_path / "poa.json").write_text(json.dumps(poa))
    exe = os.path.join(ROOT, "zk-proof-of-assets_b200", "prover")
    p = subprocess.run([exe, "--merkle-tree", "--anon-set", str(tmp_path / "anon.csv"), "--poa-input-data", str(tmp_path / "poa.json"),
                        "--output-dir", str(tmp_path)], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert "Root hash %s written to file" % vec["anon_set_10"]["root"] in p.stdout
    want_root, want_proofs = P.merkle_files(rows, owned)
    assert (tmp_path / "merkle_root.json").read_text() == want_root
    assert (tmp_path / "merkle_proofs.json").read_text() == want_proofs
    # an owner that is not in the set (or out of order) fails like the reference's panic
    (tmp_path / "poa.json").write_text(json.dumps({"accountAttestations": list(reversed(poa["accountAttestations"]))}))
    p = subprocess.run([exe, "--merkle-tree", "-a", str(tmp_path / "anon.csv"), "-p", str(tmp_path / "poa.json"), "-o", str(tmp_path)],
                       capture_output=True, text=True)
    assert p.returncode != 0 and "does not exist in the anonymity set" in p.stderr
    # library call, no owners: root only
    pkg.merkle_files(str(tmp_path / "anon.csv"), None, str(tmp_path))
    assert (tmp_path / "merkle_proofs.json").read_text() == P.merkle_files(rows, [])[1]


def test_ten_million_leaves_in_seconds(pkg):
    """the size the reference quotes 2.5 h for (scripts/README.md:913-915); checked through properties the domain offers:
    sampled leaves and sampled parent/child triples against the oracle, and a proof that folds to the root"""
    import numpy as np
    n = 10_000_000
    rng = np.random.default_rng(7)
    rows = np.zeros((n, 16), dtype=np.uint32)
    rows[:, 0:5] = rng.integers(0, 1 << 32, size=(n, 5), dtype=np.uint64).astype(np.uint32)     # 160-bit addresses
    rows[:, 8:10] = rng.integers(0, 1 << 32, size=(n, 2), dtype=np.uint64).astype(np.uint32)    # 64-bit balances
    with pkg.MerkleTree(rows) as t:
        assert t.n_leaves == 1 << 24 and t.levels == 25
        assert t.build_ms < 5000
        def val(i, lo, hi):
            return int.from_bytes(rows[i, lo:hi].tobytes(), "little")
        for i in (0, 1, n // 3, n - 1):
            assert t.node(0, i) == P.hash2(val(i, 0, 8), val(i, 8, 16))
        assert t.node(0, n) == 0 and t.node(0, (1 << 24) - 1) == 0
        for lv, i in ((1, 0), (1, n // 2 - 1), (7, 1234), (23, 1)):
            assert t.node(lv, i) == P.hash2(t.node(lv - 1, 2 * i), t.node(lv - 1, 2 * i + 1))
        el, ix = t.proof(n - 1)
        assert P.root_from_proof(t.node(0, n - 1), el, ix) == t.root()
        print("10M-leaf tree: %.1f ms on the GPU" % t.build_ms)
