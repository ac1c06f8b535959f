"""`snarkjs zkey new <circuit.r1cs> <pot_final.ptau> <circuit_0000.zkey>` -- oracle for SURVEY.md 8(f) rank 2 (the first setup step).

ORACLE CODE: TEST INFRASTRUCTURE.  CPU restatement (Python big-int group arithmetic, no trapdoor) of what the reference
runs in scripts/g16_setup.sh:239-254 (`snarkjs groth16 setup` = zkey new) on the files scripts/g16_setup.sh:196-237
names.  snarkjs 0.7.2 is not under /root/reference (pnpm-lock.yaml:2231-2244) and cannot run here, so this follows its
published algorithm (src/zkey_new.js, r1csfile 0.0.47 readR1cs, src/powersoftau_utils.js, powersoftau_preparephase2.js):
  .r1cs v1 : sec1 { u32 n8 | prime | u32 nWires | nPubOut | nPubIn | nPrvIn | u64 nLabels | u32 nConstraints }
             sec2 { per constraint: A, B, C as  u32 n | n x (u32 wire | coef[n8] LE standard form) }   sec3 { nWires x u64 }
  .ptau v1 : sec1 { u32 n8 | q | u32 power | u32 ceremonyPower }   2 tauG1 (2^(power+1)-1)  3 tauG2 (2^power)
             4 alphaTauG1  5 betaTauG1 (2^power)  6 betaG2 (1)  7 contributions
             prepared for phase 2:  12 tauG1 / 13 tauG2 / 14 alphaTauG1 / 15 betaTauG1 in LAGRANGE form, levels p = 0 .. power
             one after the other (level p starts at point 2^p - 1); section 12 also holds level power+1, made from the
             2^(power+1)-1 powers that exist (the missing top power counts as zero).  Points: affine Montgomery LE.
  zkey new : domain 2^k >= nConstraints + nPublic + 1;  gamma = delta = 1 (vk_gamma_2 = delta2 = G2, delta1 = G1)
             A_s  = sum_{A[c][s]} coef * tauG1_L[c]  (+ tauG1_L[nConstraints + s] for s <= nPublic: the public-input rows)
             B1_s / B2_s = sum_{B[c][s]} coef * tauG1_L[c] / tauG2_L[c]
             IC_s (s <= nPublic), C_s (others) = sum_A coef * betaTauG1_L[c] + sum_B coef * alphaTauG1_L[c] + sum_C coef * tauG1_L[c]
                                                 (+ betaTauG1_L[nConstraints + s] for the public-input rows)
             H_i  = level-(k+1) tauG1_L[2 i + 1]  (odd positions);   section 4 = A and B coefficients x R^2 + public rows
  section 10 = csHash[64] | u32 0.
PARITY: everything except csHash is checked against the known-trapdoor factory (oracle/factory.py `setup` with gamma =
delta = 1 is, in the exponent, exactly what zkey new computes from points) and, end to end, by proving with the new key
and verifying against its exported vkey.  csHash (Blake2b-512 of the key's points in uncompressed big-endian form, in
the order alpha1 beta1 beta2 gamma2 delta1 delta2 | IC | tau^(i+n) - tau^i, i < n-1 | C | A | B1 | B2, each list led
by its big-endian u32 length) is restated from the published source without a file to pin it on: UNPINNED.
"""
import hashlib
import struct

from . import bn254 as bn
from . import groth16 as g16
from .binfile import enc_g1, enc_g2, dec_g1, dec_g2, read_binfile, write_binfile, _le
from .factory import lagrange_at

R, Q = bn.R, bn.Q


# ---- .r1cs ---------------------------------------------------------------------------------------------------------
def write_r1cs(circ, n_pub_out=0):
    """the circuit (factory.SynthCircuit / ListCircuit) as an .r1cs image; n_pub_out of the public wires are outputs"""
    n_pub = circ.n_public
    sec1 = (struct.pack("<I", 32) + _le(R) + struct.pack("<IIII", circ.n_vars, n_pub_out, n_pub - n_pub_out,
                                                         circ.n_vars - 1 - n_pub)
            + struct.pack("<Q", circ.n_vars) + struct.pack("<I", circ.n_constraints))
    body = []
    for c in range(circ.n_constraints):
        for lc in circ.constraint(c):
            body.append(struct.pack("<I", len(lc)))
            for wire, coef in lc:
                body.append(struct.pack("<I", wire) + _le(coef % R))
    sec3 = b"".join(struct.pack("<Q", i) for i in range(circ.n_vars))
    return write_binfile(b"r1cs", 1, [(1, sec1), (2, b"".join(body)), (3, sec3)])


def read_r1cs(data):
    """-> dict(n_vars, n_public, n_constraints, constraints=[(A, B, C)] of (wire, coef) lists)"""
    _, secs = read_binfile(data, b"r1cs")
    p, _ = secs[1][0]
    n8 = struct.unpack_from("<I", data, p)[0]
    if n8 != 32 or int.from_bytes(data[p + 4:p + 36], "little") != R:
        raise ValueError("r1cs: the circuit's field is not the curve's scalar field")
    n_wires, n_out, n_pub_in, n_prv = struct.unpack_from("<IIII", data, p + 36)
    n_cons = struct.unpack_from("<I", data, p + 36 + 16 + 8)[0]
    pos, size = secs[2][0]
    end = pos + size
    cons = []
    for _ in range(n_cons):
        lcs = []
        for _ in range(3):
            n = struct.unpack_from("<I", data, pos)[0]
            pos += 4
            lc = []
            for _ in range(n):
                wire = struct.unpack_from("<I", data, pos)[0]
                lc.append((wire, int.from_bytes(data[pos + 4:pos + 36], "little")))
                pos += 36
            lcs.append(lc)
        cons.append(tuple(lcs))
    if pos != end:
        raise ValueError("r1cs: constraints section has the wrong size")
    return {"n_vars": n_wires, "n_public": n_out + n_pub_in, "n_constraints": n_cons, "constraints": cons}


# ---- .ptau (known tau, alpha, beta: a test fixture; real ceremonies publish the same layout) -----------------------
def lagrange_levels(tau, power, truncated_top):
    """scalars of the Lagrange sections: levels 0 .. power (+ level power+1 without the top power of tau)"""
    out = []
    for p in range(power + 1):
        out.append([1] if p == 0 else lagrange_at(tau, p))
    if truncated_top:
        p = power + 1
        n = 1 << p
        full = lagrange_at(tau, p)
        w = g16.root_of_unity(p)
        top = pow(tau, n - 1, R) * bn.inv_mod(n, R) % R
        # L_j(tau) = 1/n sum_k w^(-jk) tau^k: dropping k = n-1 removes w^(-j(n-1)) tau^(n-1) / n = w^j tau^(n-1) / n
        out.append([(full[j] - pow(w, j, R) * top) % R for j in range(n)])
    return out


def write_ptau(tau, alpha, beta, power):
    fb1, fb2 = bn.FixedBase(bn.G1), bn.FixedBase(bn.G2)
    n = 1 << power
    tp = [1] * (2 * n - 1)
    for i in range(1, 2 * n - 1):
        tp[i] = tp[i - 1] * tau % R
    sec1 = struct.pack("<I", 32) + _le(Q) + struct.pack("<II", power, power)
    sec2 = b"".join(enc_g1(fb1.mul(t)) for t in tp)
    sec3 = b"".join(enc_g2(fb2.mul(t)) for t in tp[:n])
    sec4 = b"".join(enc_g1(fb1.mul(alpha * t % R)) for t in tp[:n])
    sec5 = b"".join(enc_g1(fb1.mul(beta * t % R)) for t in tp[:n])
    sec6 = enc_g2(fb2.mul(beta))
    sec7 = struct.pack("<I", 0)
    lv = lagrange_levels(tau, power, True)
    flat = [s for level in lv for s in level]
    flat_p = [s for level in lv[:power + 1] for s in level]
    g1 = lambda k: enc_g1(fb1.mul(k)) if k else bytes(64)        # noqa: E731
    sec12 = b"".join(g1(s) for s in flat)
    sec13 = b"".join(enc_g2(fb2.mul(s)) if s else bytes(128) for s in flat_p)
    sec14 = b"".join(g1(alpha * s % R) for s in flat_p)
    sec15 = b"".join(g1(beta * s % R) for s in flat_p)
    return write_binfile(b"ptau", 1, [(1, sec1), (2, sec2), (3, sec3), (4, sec4), (5, sec5), (6, sec6), (7, sec7),
                                      (12, sec12), (13, sec13), (14, sec14), (15, sec15)])


# ---- zkey new ------------------------------------------------------------------------------------------------------
def _points(data, secs, typ, first, count, size, dec):
    pos, total = secs[typ][0]
    if (first + count) * size > total:
        raise ValueError("ptau: section %d is too short for this circuit (powers of tau too small)" % typ)
    return [dec(bytes(data[pos + (first + i) * size:pos + (first + i + 1) * size])) for i in range(count)]


def _be(x):
    return int(x).to_bytes(32, "big")


def _u_g1(p):          # ffjavascript toRprUncompressed: x | y big-endian, infinity = 0x40 then zeros
    return bytes([0x40]) + bytes(63) if p is None else _be(p[0]) + _be(p[1])


def _u_g2(p):
    return bytes([0x40]) + bytes(127) if p is None else _be(p[0][1]) + _be(p[0][0]) + _be(p[1][1]) + _be(p[1][0])


def _lincomb(curve, terms):
    """sum coef * P over (coef, P); small negative coefficients (r - k) as -(k P), like a signed digit"""
    acc = None
    for coef, p in terms:
        coef %= R
        if coef == 0 or p is None:
            continue
        t = curve.mul(p, coef) if coef <= R // 2 else curve.neg(curve.mul(p, R - coef))
        acc = curve.add(acc, t)
    return acc


def circuit_hash(alpha1, beta1, beta2, IC, mono, n, Cp, A_pts, B1_pts, B2_pts):
    """csHash of a fresh key (see the module header; mono = tauG1 in monomial form, at least 2n - 1 points)"""
    h = hashlib.blake2b(digest_size=64)
    for b in (_u_g1(alpha1), _u_g1(beta1), _u_g2(beta2), _u_g2(bn.G2_GEN), _u_g1(bn.G1_GEN), _u_g2(bn.G2_GEN)):
        h.update(b)

    def feed(points, enc):
        h.update(struct.pack(">I", len(points)))
        for p in points:
            h.update(enc(p))
    feed(IC, _u_g1)
    feed([bn.G1.add(mono[i + n], bn.G1.neg(mono[i])) for i in range(n - 1)], _u_g1)
    feed(Cp, _u_g1)
    feed(A_pts, _u_g1)
    feed(B1_pts, _u_g1)
    feed(B2_pts, _u_g2)
    return h.digest()


def zkey_new(r1cs_bytes, ptau_bytes):
    """-> zkey bytes (sections 1-10) of the circuit under the ceremony's tau, alpha, beta, with gamma = delta = 1"""
    cir = read_r1cs(r1cs_bytes)
    _, ps = read_binfile(ptau_bytes, b"ptau")
    hp, _ = ps[1][0]
    if struct.unpack_from("<I", ptau_bytes, hp)[0] != 32 or int.from_bytes(ptau_bytes[hp + 4:hp + 36], "little") != Q:
        raise ValueError("ptau: not a BN254 ceremony")
    power = struct.unpack_from("<I", ptau_bytes, hp + 36)[0]
    if 12 not in ps or 13 not in ps or 14 not in ps or 15 not in ps:
        raise ValueError("ptau: not prepared for phase 2 (sections 12-15 missing)")
    n_vars, n_pub, n_cons = cir["n_vars"], cir["n_public"], cir["n_constraints"]
    need = n_cons + n_pub + 1
    k = max(1, (need - 1).bit_length())
    if k > power:
        raise ValueError("circuit too big for this power of tau ceremony. %d > 2**%d" % (need, power))
    n = 1 << k
    first = n - 1                                     # level k starts here
    tau1 = _points(ptau_bytes, ps, 12, first, n, 64, dec_g1)
    tau2 = _points(ptau_bytes, ps, 13, first, n, 128, dec_g2)
    alpha_l = _points(ptau_bytes, ps, 14, first, n, 64, dec_g1)
    beta_l = _points(ptau_bytes, ps, 15, first, n, 64, dec_g1)
    top = _points(ptau_bytes, ps, 12, 2 * n - 1, 2 * n, 64, dec_g1)     # level k+1
    alpha1 = _points(ptau_bytes, ps, 4, 0, 1, 64, dec_g1)[0]
    beta1 = _points(ptau_bytes, ps, 5, 0, 1, 64, dec_g1)[0]
    beta2 = _points(ptau_bytes, ps, 6, 0, 1, 128, dec_g2)[0]
    ta = [[] for _ in range(n_vars)]
    tb1 = [[] for _ in range(n_vars)]
    tb2 = [[] for _ in range(n_vars)]
    tc = [[] for _ in range(n_vars)]
    coefs = []
    for c, (A, B, C) in enumerate(cir["constraints"]):
        for s, co in A:
            ta[s].append((co, tau1[c]))
            tc[s].append((co, beta_l[c]))
            coefs.append((0, c, s, co))
        for s, co in B:
            tb1[s].append((co, tau1[c]))
            tb2[s].append((co, tau2[c]))
            tc[s].append((co, alpha_l[c]))
            coefs.append((1, c, s, co))
        for s, co in C:
            tc[s].append((co, tau1[c]))
    for s in range(n_pub + 1):
        ta[s].append((1, tau1[n_cons + s]))
        tc[s].append((1, beta_l[n_cons + s]))
        coefs.append((0, n_cons + s, s, 1))
    A_pts = [_lincomb(bn.G1, t) for t in ta]
    B1_pts = [_lincomb(bn.G1, t) for t in tb1]
    B2_pts = [_lincomb(bn.G2, t) for t in tb2]
    M_pts = [_lincomb(bn.G1, t) for t in tc]
    IC, Cp = M_pts[:n_pub + 1], M_pts[n_pub + 1:]
    H = [top[2 * i + 1] for i in range(n)]
    mono = _points(ptau_bytes, ps, 2, 0, 2 * n - 1, 64, dec_g1)
    cs = circuit_hash(alpha1, beta1, beta2, IC, mono, n, Cp, A_pts, B1_pts, B2_pts)
    sec2 = (struct.pack("<I", 32) + _le(Q) + struct.pack("<I", 32) + _le(R) + struct.pack("<III", n_vars, n_pub, n)
            + enc_g1(alpha1) + enc_g1(beta1) + enc_g2(beta2) + enc_g2(bn.G2_GEN) + enc_g1(bn.G1_GEN) + enc_g2(bn.G2_GEN))
    r2 = bn.MONT_R_R * bn.MONT_R_R % R
    sec4 = struct.pack("<I", len(coefs)) + b"".join(struct.pack("<III", m, c, s) + _le(v % R * r2 % R) for m, c, s, v in coefs)
    g1s = lambda pts: b"".join(enc_g1(p) for p in pts)          # noqa: E731
    sec10 = cs + struct.pack("<I", 0)
    return write_binfile(b"zkey", 1, [(1, struct.pack("<I", 1)), (2, sec2), (3, g1s(IC)), (4, sec4), (5, g1s(A_pts)),
                                      (6, g1s(B1_pts)), (7, b"".join(enc_g2(p) for p in B2_pts)), (8, g1s(Cp)),
                                      (9, g1s(H)), (10, sec10)])
