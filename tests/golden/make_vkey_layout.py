#!/usr/bin/env python3
"""Copies the exact bytes of two verification keys the reference commits into tests/golden/vkey_layout.json (run in the
build container, where /root/reference exists): the layout `snarkjs zkey export verificationkey` writes, which
g16_zkey_export_vkey and oracle/setup.py export_vkey must reproduce."""
import base64
import json
import os

REF = os.environ.get("REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "vkey_layout.json")
FILES = ["tests/4_sigs_2_batches_12_height/layer_one/layer_one_vkey.json",
         "tests/4_sigs_2_batches_12_height/layer_three/layer_three_vkey.json"]

out = {"_source": "silversixpence-crypto/zk-proof-of-assets: " + ", ".join(FILES), "files": []}
for rel in FILES:
    with open(os.path.join(REF, rel), "rb") as f:
        out["files"].append({"path": rel, "b64": base64.b64encode(f.read()).decode()})
with open(OUT, "w") as f:
    json.dump(out, f, indent=1)
    f.write("\n")
print("wrote", OUT)
