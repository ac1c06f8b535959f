"""`snarkjs wtns check <circuit.r1cs> <witness.wtns>` -- oracle for the witness check of scripts/g16_verify.sh:205-209.

ORACLE CODE: TEST INFRASTRUCTURE.  Python big-int restatement of snarkjs 0.7.2 src/wtns_check.js (not under /root/reference,
pnpm-lock.yaml:2231-2244): for every constraint <A, w> * <B, w> == <C, w> mod r; the first failing constraint stops the
check.  PARITY: the arithmetic is the R1CS definition itself (pinned on nothing but that); the wording of snarkjs' log lines
is not restated here.
"""
from . import bn254 as bn
from .binfile import read_wtns
from .zkey_new import read_r1cs

R = bn.R


def wtns_check(r1cs_bytes, wtns_bytes):
    """-> None when every constraint holds, else the lowest index of a constraint that fails"""
    cir = read_r1cs(r1cs_bytes)
    wt = read_wtns(wtns_bytes)
    if wt["prime"] != R:
        raise ValueError("Curve of the witness does not match the curve of the circuit")
    w = wt["witness"]
    if len(w) != cir["n_vars"]:
        raise ValueError("Invalid witness length. Circuit: %d, witness: %d" % (cir["n_vars"], len(w)))
    for c, (A, B, C) in enumerate(cir["constraints"]):
        va = sum(co * w[s] for s, co in A) % R
        vb = sum(co * w[s] for s, co in B) % R
        vc = sum(co * w[s] for s, co in C) % R
        if va * vb % R != vc:
            return c
    return None
