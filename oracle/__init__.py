"""CPU oracle for the Groth16/BN254 `groth16 prove` hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(`zk-proof-of-assets_b200/`) may import, link or execute anything from this
directory; only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py` do, and there only as the checker or the
timed CPU baseline.

Provenance: the reference (silversixpence-crypto/zk-proof-of-assets) holds no
prover source.  The arithmetic lives in third-party packages that are absent
from /root/reference: snarkjs 0.7.2 -> ffjavascript 0.2.62 -> wasmcurves 0.2.2
(pnpm-lock.yaml:2231-2244, :1406-1411, :2459-2462) and iden3/rapidsnark at an
unpinned HEAD (scripts/machine_initialization.sh:301-314).  This oracle restates
their published algorithm and anchors on the
reference's own call site (scripts/g16_prove.sh:246-260) and committed
artefacts:

  * pinned: pairing / verifier / point decoding / JSON layout against the
    reference's committed (vkey, public, proof) triples and the e(-alpha,beta)
    known-answer limbs in every sanitized_proof.json
    (tests/golden/reference_vectors.json, made by
    tests/golden/make_reference_vectors.py);
  * PROVER BITS: PARITY UNPINNED.  The reference pins no prover output (proofs
    are randomised, no test fixes r,s).  The prove restatement is pinned only
    indirectly: every proof it makes must satisfy the pinned verifier, and must
    equal the closed-form known-trapdoor proof.
"""
