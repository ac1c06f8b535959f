#!/usr/bin/env python3
"""Collect the reference's committed Groth16 artefacts into one fixture file.

Run in the build container (where /root/reference exists):
    python tests/golden/make_reference_vectors.py
Writes tests/golden/reference_vectors.json.  /root/reference does not exist on
the GPU box, so tests read only the committed fixture.

What is collected:
  * (vkey, public, proof) triples that must satisfy the Groth16 pairing equation;
  * sanitized_proof.json files: 6x43-bit limb forms of the proof, of the vkey
    parts (gamma2, delta2, IC) and of e(-alpha1, beta2) ("negalfa1xbeta2");
  * the exact bytes of proof.json/public.json in rapidsnark and snarkjs styles
    (output-format vectors, row a16).
"""
import base64
import hashlib
import json
import os

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors.json")

TRIPLES = [
    # name, vkey, public, proof
    ("c2_layer_one_batch_0", "tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json",
     "tests/4_sigs_2_batches_12_height/layer_one/batch_0/public.json",
     "tests/4_sigs_2_batches_12_height/layer_one/batch_0/proof.json"),
    ("c2_layer_one_batch_1", "tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json",
     "tests/4_sigs_2_batches_12_height/layer_one/batch_1/public.json",
     "tests/4_sigs_2_batches_12_height/layer_one/batch_1/proof.json"),
    ("c2_layer_two_batch_0", "tests/4_sigs_2_batches_12_height/layer_two/layer_two_vkey.json",
     "tests/4_sigs_2_batches_12_height/layer_two/batch_0/public.json",
     "tests/4_sigs_2_batches_12_height/layer_two/batch_0/proof.json"),
    ("c2_layer_two_batch_1", "tests/4_sigs_2_batches_12_height/layer_two/layer_two_vkey.json",
     "tests/4_sigs_2_batches_12_height/layer_two/batch_1/public.json",
     "tests/4_sigs_2_batches_12_height/layer_two/batch_1/proof.json"),
    ("c2_layer_three", "tests/4_sigs_2_batches_12_height/layer_three/layer_three_vkey.json",
     "tests/4_sigs_2_batches_12_height/layer_three/public.json",
     "tests/4_sigs_2_batches_12_height/layer_three/proof.json"),
    ("experiments_groth16_input_prep", "experiments/scripts/groth16_input_prep/vkey.json",
     "experiments/scripts/groth16_input_prep/public.json",
     "experiments/scripts/groth16_input_prep/proof.json"),
    ("circom_pairing_fixture_0", "git_modules/circom-pairing/python/fixtures/vkey-0.json",
     "git_modules/circom-pairing/python/fixtures/public-0.json",
     "git_modules/circom-pairing/python/fixtures/proof-0.json"),
]

SANITIZED = [
    # name, sanitized_proof.json, proof.json, public.json, vkey (or None)
    ("c1_layer_one_batch_0", "tests/1_sigs_1_batches_5_height/layer_one/batch_0", None),
    ("c1_layer_two_batch_0", "tests/1_sigs_1_batches_5_height/layer_two/batch_0", None),
    ("c2_layer_one_batch_0", "tests/4_sigs_2_batches_12_height/layer_one/batch_0",
     "tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json"),
    ("c2_layer_two_batch_1", "tests/4_sigs_2_batches_12_height/layer_two/batch_1",
     "tests/4_sigs_2_batches_12_height/layer_two/layer_two_vkey.json"),
]

FORMAT_FILES = [
    ("rapidsnark", "tests/1_sigs_1_batches_5_height/layer_one/batch_0/proof.json",
     "tests/1_sigs_1_batches_5_height/layer_one/batch_0/public.json"),
    ("rapidsnark_13_public", "tests/1_sigs_1_batches_5_height/layer_three/proof.json",
     "tests/1_sigs_1_batches_5_height/layer_three/public.json"),
    ("snarkjs", "experiments/scripts/groth16_input_prep/proof.json",
     "experiments/scripts/groth16_input_prep/public.json"),
]


def _raw(path):
    with open(os.path.join(REF, path), "rb") as f:
        return f.read()


def _json(path):
    return json.loads(_raw(path))


def main():
    out = {"source": "silversixpence-crypto/zk-proof-of-assets @ 954ab723442bb0324c4654ba734b9a4a0c616464",
           "triples": [], "sanitized": [], "formats": []}
    for name, vk, pub, prf in TRIPLES:
        out["triples"].append({"name": name, "paths": [vk, pub, prf],
                               "vkey": _json(vk), "public": _json(pub), "proof": _json(prf)})
    for name, d, vk in SANITIZED:
        out["sanitized"].append({"name": name, "dir": d,
                                 "sanitized": _json(d + "/sanitized_proof.json"),
                                 "proof": _json(d + "/proof.json"),
                                 "public": _json(d + "/public.json"),
                                 "vkey": _json(vk) if vk else None})
    for style, prf, pub in FORMAT_FILES:
        pb, ub = _raw(prf), _raw(pub)
        out["formats"].append({"style": style, "paths": [prf, pub],
                               "proof_b64": base64.b64encode(pb).decode(),
                               "public_b64": base64.b64encode(ub).decode(),
                               "proof_sha256": hashlib.sha256(pb).hexdigest(),
                               "public_sha256": hashlib.sha256(ub).hexdigest()})
    with open(OUT, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
