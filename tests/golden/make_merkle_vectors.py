#!/usr/bin/env python3
"""Collects the reference's own Merkle / Poseidon artefacts into tests/golden/merkle_vectors.json (run in the build
container, where /root/reference exists; the GPU box only sees the committed JSON).  Nothing is computed here: every
value is copied from a file the reference commits."""
import json
import os
import re

REF = os.environ.get("REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "merkle_vectors.json")


def main():
    t = os.path.join(REF, "tests", "1_sigs_1_batches_5_height")
    with open(os.path.join(t, "anonymity_set_10.csv")) as f:
        anon = f.read()
    with open(os.path.join(t, "logs", "merkle_tree.log")) as f:
        log = f.read()
    root = re.search(r"Root hash (\d+) written", log).group(1)
    leaves = re.search(r"Done creating (\d+) leaves", log).group(1)
    padded = re.search(r"after adding padding nodes\): (\d+)", log).group(1)
    with open(os.path.join(t, "layer_two_input.json")) as f:
        l2 = json.load(f)
    with open(os.path.join(REF, "experiments", "scripts", "merkle_inputs.json")) as f:
        ex = json.load(f)
    out = {
        "_source": "silversixpence-crypto/zk-proof-of-assets: tests/1_sigs_1_batches_5_height/{anonymity_set_10.csv,"
                   "logs/merkle_tree.log,layer_two_input.json}, experiments/scripts/merkle_inputs.json",
        "anon_set_10": {"csv": anon, "root": root, "leaves": int(leaves), "padded_leaves": int(padded)},
        "layer_two_input": {k: l2[k] for k in ("leaf_addresses", "leaf_balances", "merkle_root", "path_elements", "path_indices")},
        "experiments_merkle_inputs": {"leaf": str(ex["leaf"]), "root": str(ex["root"]),
                                      "pathElements": [str(x) for x in ex["pathElements"]],
                                      "pathIndices": [int(x) for x in ex["pathIndices"]]},
    }
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print("wrote", OUT)


if __name__ == "__main__":
    main()
