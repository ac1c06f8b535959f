"""Poseidon (circom parameters) and the anonymity-set Merkle tree -- oracle for SURVEY.md 8(f) rank 4.

ORACLE CODE: TEST INFRASTRUCTURE.  CPU restatement (Python big-int) of what the reference's
scripts/merkle_tree.rs computes:
  hasher            light_poseidon Poseidon::<Fr>::new_circom(2)  (merkle_tree.rs:150-181): t = 3, 8 full + 57 partial
                    rounds, x^5, state [0, in0, in1], output state[0] -- the circomlib `Poseidon(2)` the circuits use
                    (circuits/merkle.circom:39)
  leaves            hash(address, balance), then zero leaves up to the next power of two (merkle_tree.rs:206-269)
  tree / proofs     rs_merkle from_leaves / proof: parent = hash(left, right); proof hashes bottom-up
                    (merkle_tree.rs:290-396); path_indices = index bits, low first
  files             merkle_root.json {"__bigint__": ..}, merkle_proofs.json {leaves, path_elements, path_indices},
                    serde_json pretty layout (scripts/README.md:633-673)
The round constants and the MDS matrix are not in the reference tree (light-poseidon / circomlib hold them): they are
regenerated here with the Poseidon paper's Grain-LFSR procedure (generate_parameters_grain.sage 1 0 254 3 8 57) and
PINNED on the reference's own artefacts by tests/test_oracle_poseidon.py:
  * the root the reference logged for tests/1_sigs_1_batches_5_height/anonymity_set_10.csv
    (tests/1_sigs_1_batches_5_height/logs/merkle_tree.log:9),
  * the Merkle path in tests/1_sigs_1_batches_5_height/layer_two_input.json (leaf, 4 siblings, root),
  * the 24-level path in experiments/scripts/merkle_inputs.json.
"""
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
T, RF, RP = 3, 8, 57


def grain_parameters(t=T, rf=RF, rp=RP, n=254, prime=R):
    """-> (round constants [(rf+rp)*t], MDS matrix t x t) of the Poseidon reference generator for a prime field (field
    tag 1) and the x^5 S-box (tag 0): 80-bit Grain LFSR, 160 warm-up clocks, self-shrinking output."""
    state = []
    for value, width in ((1, 2), (0, 4), (n, 12), (t, 12), (rf, 10), (rp, 10)):
        state += [int(b) for b in bin(value)[2:].zfill(width)]
    state += [1] * 30

    def clock():
        b = state[62] ^ state[51] ^ state[38] ^ state[23] ^ state[13] ^ state[0]
        state.pop(0)
        state.append(b)
        return b

    for _ in range(160):
        clock()

    def bit():
        while True:
            keep = clock()
            b = clock()
            if keep:
                return b

    def draw():
        v = 0
        for _ in range(n):
            v = (v << 1) | bit()
        return v

    consts = []
    while len(consts) < (rf + rp) * t:
        v = draw()
        if v < prime:                      # rejection sampling
            consts.append(v)
    while True:                            # Cauchy matrix 1 / (x_i + y_j) from 2t distinct field elements
        vals = [draw() % prime for _ in range(2 * t)]
        if len(set(vals)) != 2 * t:
            continue
        xs, ys = vals[:t], vals[t:]
        if any((x + y) % prime == 0 for x in xs for y in ys):
            continue
        return consts, [[pow((x + y) % prime, prime - 2, prime) for y in ys] for x in xs]


_PARAMS = None


def params():
    global _PARAMS
    if _PARAMS is None:
        _PARAMS = grain_parameters()
    return _PARAMS


def hash2(a, b):
    """circomlib Poseidon(2) / light_poseidon new_circom(2)"""
    consts, mds = params()
    st = [0, a % R, b % R]
    for r in range(RF + RP):
        st = [(st[i] + consts[r * T + i]) % R for i in range(T)]
        if r < RF // 2 or r >= RF // 2 + RP:
            st = [pow(x, 5, R) for x in st]
        else:
            st[0] = pow(st[0], 5, R)
        st = [sum(mds[i][j] * st[j] for j in range(T)) % R for i in range(T)]
    return st[0]


def build_leaves(rows):
    """rows = [(address int, balance int)] -> leaf hashes padded with zero leaves to a power of two
    (merkle_tree.rs:206-269; a single row stays a single leaf: 2^ceil(log2 1) = 1)"""
    leaves = [hash2(a, b) for a, b in rows]
    size = 1
    while size < len(leaves):
        size *= 2
    return leaves + [0] * (size - len(leaves))


def build_tree(leaves):
    """-> levels, levels[0] = leaves, levels[-1] = [root]"""
    levels = [list(leaves)]
    while len(levels[-1]) > 1:
        cur = levels[-1]
        levels.append([hash2(cur[i], cur[i + 1]) for i in range(0, len(cur), 2)])
    return levels


def proof(levels, index):
    """-> (path_elements bottom-up, path_indices): sibling per level; index bit 1 = the node is a right child"""
    elements, indices = [], []
    for lvl in levels[:-1]:
        elements.append(lvl[index ^ 1])
        indices.append(index & 1)
        index >>= 1
    return elements, indices


def root_from_proof(leaf, elements, indices):
    """circuits/merkle.circom:24-47 MerkleProofVerify"""
    cur = leaf
    for e, i in zip(elements, indices):
        cur = hash2(e, cur) if i else hash2(cur, e)
    return cur


def parse_anon_set(text):
    """csv with headings address,eth_balance; address hex with 0x, balance decimal (merkle_tree.rs:232-247)"""
    rows = []
    lines = [ln.strip() for ln in text.splitlines() if ln.strip()]
    for ln in lines[1:]:
        addr, bal = ln.split(",")[:2]
        rows.append((int(addr[2:], 16), int(bal)))
    return rows


def _pretty(obj, depth=0):
    """serde_json::to_writer_pretty: 2-space indent, no trailing newline, empty containers as [] / {}"""
    pad, pad1 = "  " * depth, "  " * (depth + 1)
    if isinstance(obj, dict):
        if not obj:
            return "{}"
        items = ['%s"%s": %s' % (pad1, k, _pretty(v, depth + 1)) for k, v in obj.items()]
        return "{\n" + ",\n".join(items) + "\n" + pad + "}"
    if isinstance(obj, list):
        if not obj:
            return "[]"
        return "[\n" + ",\n".join(pad1 + _pretty(v, depth + 1) for v in obj) + "\n" + pad + "]"
    if isinstance(obj, str):
        return '"%s"' % obj
    return str(obj)


def merkle_files(rows, owned):
    """rows = the anonymity set, owned = [(address, balance)] of the prover's accounts, in anonymity-set order
    (merkle_tree.rs:315-335 scans forward only) -> (merkle_root.json text, merkle_proofs.json text)"""
    leaves = build_leaves(rows)
    levels = build_tree(leaves)
    big = lambda v: {"__bigint__": str(v)}          # noqa: E731
    out = {"leaves": [], "path_elements": [], "path_indices": []}
    at = 0
    for a, b in owned:
        h = hash2(a, b)
        while leaves[at] != h:
            if at == len(leaves) - 1:
                raise ValueError("Owned leaf %d does not exist in the anonymity set" % a)
            at += 1
        el, idx = proof(levels, at)
        out["leaves"].append({"address": big(a), "balance": big(b), "hash": big(h)})
        out["path_elements"].append([big(e) for e in el])
        out["path_indices"].append(idx)
    return _pretty(big(levels[-1][0])), _pretty(out)
