"""Phase-2 ceremony steps and verification-key export -- oracle for SURVEY.md 8(f) rank 2 (file-level part).

ORACLE CODE: TEST INFRASTRUCTURE.  CPU restatement (Python big-int + hashlib) of what the reference runs in
scripts/g16_setup.sh:254-283:
    snarkjs zkey contribute <in.zkey> <out.zkey> --name=.. -e=..          (:262-266)
    snarkjs zkey beacon <in.zkey> <out.zkey> <beaconHash> <numIterationsExp> -n=..   (:272-277)
    snarkjs zkey export verificationkey <zkey> <vkey.json>                (:287-295)
snarkjs 0.7.2 is not under /root/reference (pnpm-lock.yaml:2231-2244) and cannot run here, so this follows its published
algorithm (src/zkey_contribute.js, zkey_beacon.js, zkey_utils.js writeMPCParams, zkey_export_verificationkey.js):
  * delta' <- RNG;  delta1, delta2 <- delta' * delta;  sections 8 (C) and 9 (H) <- (1/delta') * points; 3-7 copied
  * section 10 = csHash[64] | u32 n | n x { deltaAfter G1 | g1_s G1 | g1_sx G1 | g2_spx G2 | transcript[64] | u32 len | params }
    (points in the zkey's own little-endian Montgomery records; params = 1,len,name  [2,numIterationsExp,3,len,beaconHash])
  * transcript = Blake2b-512(csHash | hashPubKey(previous contributions) | g1_s | g1_sx), points hashed as big-endian
    standard-form coordinates (Fq2 as c1 | c0); proof of knowledge g2_spx = delta' * hashToG2(transcript)
PARITY UNPINNED where snarkjs draws from its ChaCha generator (the reference commits no zkey to pin those bytes on): the
value of delta', the point g1_s and the map hashToG2 are THIS repository's own derivations (Blake2b, below).  Any delta'
gives a valid contribution and the updated key proves and verifies; what cannot be claimed without node is that
`snarkjs zkey verify` accepts the proof-of-knowledge record.  The vkey export IS pinned: key order, number formats and
the vk_alphabeta_12 tower basis on the reference's committed vkey files (tests/golden/reference_vectors.json).
"""
import hashlib
import json

from . import bn254 as bn
from .binfile import read_binfile

R, Q = bn.R, bn.Q
TAG = b"g16b200/phase2/"


def parse_sections(data):
    """{section type: payload bytes} of a zkey image"""
    _, secs = read_binfile(data, b"zkey")
    return {t: bytes(data[v[0][0]:v[0][0] + v[0][1]]) for t, v in secs.items()}


G2_COFACTOR = 2 * Q - R


def blake(*parts):
    h = hashlib.blake2b(digest_size=64)
    for p in parts:
        h.update(p)
    return h.digest()


def be32(x):
    return int(x).to_bytes(32, "big")


def hash_g1(p):
    return be32(p[0]) + be32(p[1])


def hash_g2(p):
    return be32(p[0][1]) + be32(p[0][0]) + be32(p[1][1]) + be32(p[1][0])


def f2_sqrt(a):
    """a root of a in Fq2 = Fq[u]/(u^2+1) by the complex method, or None; q = 3 mod 4"""
    a0, a1 = a
    leg = lambda x: pow(x, (Q - 1) // 2, Q)      # noqa: E731
    rt = lambda x: pow(x, (Q + 1) // 4, Q)       # noqa: E731
    if a1 == 0:
        if a0 == 0:
            return (0, 0)
        return (rt(a0), 0) if leg(a0) == 1 else (0, rt(Q - a0))
    norm = (a0 * a0 + a1 * a1) % Q
    if leg(norm) != 1:
        return None
    s = rt(norm)
    half = (Q + 1) // 2
    d = (a0 + s) * half % Q
    if leg(d) != 1:
        d = (a0 - s) * half % Q
        if leg(d) != 1:
            return None
    x0 = rt(d)
    x1 = a1 * bn.inv_mod(2 * x0 % Q, Q) % Q
    return (x0, x1)


def f2_is_negative(y):
    """the larger of the two roots: compare (c1, c0) with the negation"""
    c0, c1 = y
    if c1 != 0:
        return c1 > (Q - 1) // 2
    return c0 > (Q - 1) // 2


def hash_to_g2(transcript):
    ctr = 0
    while True:
        c = ctr.to_bytes(4, "little")
        x = (int.from_bytes(blake(TAG + b"g2/c0", transcript, c), "little") % Q,
             int.from_bytes(blake(TAG + b"g2/c1", transcript, c), "little") % Q)
        rhs = bn.f2_add(bn.f2_mul(bn.f2_mul(x, x), x), bn.B2)
        y = f2_sqrt(rhs)
        ctr += 1
        if y is None:
            continue
        if f2_is_negative(y):
            y = ((Q - y[0]) % Q, (Q - y[1]) % Q)
        # cofactor clearing: the point is NOT in the order-r subgroup yet, so the scalar must not be reduced mod r
        # (Curve.mul does that): use the Jacobian ladder directly
        p = bn.G2.to_affine(bn.G2.jmul(bn.G2.to_jac((x, y)), G2_COFACTOR))
        if p is not None:
            return p


def parse_mpc(sec10):
    """-> (csHash, [contribution dicts with raw bytes])"""
    cs = sec10[:64]
    n = int.from_bytes(sec10[64:68], "little")
    pos = 68
    out = []
    for _ in range(n):
        rec = sec10[pos:pos + 64 * 3 + 128 + 64]
        plen = int.from_bytes(sec10[pos + 384:pos + 388], "little")
        params = sec10[pos + 388:pos + 388 + plen]
        out.append({"points": rec[:320], "transcript": rec[320:384], "params": params})
        pos += 388 + plen
    return cs, out


def _g1_from_lem(b):
    x, y = int.from_bytes(b[:32], "little"), int.from_bytes(b[32:64], "little")
    if x == 0 and y == 0:
        return None
    ri = bn.inv_mod(1 << 256, Q)
    return (x * ri % Q, y * ri % Q)


def _g2_from_lem(b):
    v = [int.from_bytes(b[32 * i:32 * i + 32], "little") for i in range(4)]
    if not any(v):
        return None
    ri = bn.inv_mod(1 << 256, Q)
    v = [x * ri % Q for x in v]
    return ((v[0], v[1]), (v[2], v[3]))


def _g1_lem(p):
    if p is None:
        return bytes(64)
    return b"".join(((c << 256) % Q).to_bytes(32, "little") for c in p)


def _g2_lem(p):
    if p is None:
        return bytes(128)
    return b"".join(((c << 256) % Q).to_bytes(32, "little") for c in (p[0][0], p[0][1], p[1][0], p[1][1]))


def hash_pubkey(c):
    """bytes hashPubKey feeds the hasher for one stored contribution"""
    pts = c["points"]
    return (hash_g1(_g1_from_lem(pts[0:64])) + hash_g1(_g1_from_lem(pts[64:128])) + hash_g1(_g1_from_lem(pts[128:192])) +
            hash_g2(_g2_from_lem(pts[192:320])) + c["transcript"])


def derive(seed):
    d = int.from_bytes(blake(TAG + b"delta", seed), "little") % R or 1
    s = int.from_bytes(blake(TAG + b"s", seed), "little") % R or 1
    return d, s


def beacon_seed(beacon_hash, num_iterations_exp):
    cur = bytes(beacon_hash)
    for _ in range(1 << num_iterations_exp):
        cur = hashlib.sha256(cur).digest()
    return cur


def contribute(zkey_bytes, name, seed, beacon=None):
    """-> (new zkey bytes, contribution hash).  seed: bytes the secrets are derived from; beacon = (hash bytes, exp)
    marks the record as a beacon contribution (the caller passes seed = beacon_seed(*beacon))."""
    secs = parse_sections(zkey_bytes)
    h = bytearray(secs[2])
    o = 4 + 32 + 4 + 32 + 12
    delta1 = _g1_from_lem(h[o + 64 + 64 + 128 + 128:o + 64 + 64 + 128 + 128 + 64])
    delta2 = _g2_from_lem(h[o + 64 + 64 + 128 + 128 + 64:o + 64 + 64 + 128 + 128 + 64 + 128])
    cs, contribs = parse_mpc(secs[10])
    d, s = derive(seed)
    g1_s = bn.G1.mul(bn.G1_GEN, s)
    g1_sx = bn.G1.mul(g1_s, d)
    transcript = blake(cs, *[hash_pubkey(c) for c in contribs], hash_g1(g1_s), hash_g1(g1_sx))
    g2_sp = hash_to_g2(transcript)
    g2_spx = bn.G2.mul(g2_sp, d)
    nd1, nd2 = bn.G1.mul(delta1, d), bn.G2.mul(delta2, d)
    h[o + 384:o + 448] = _g1_lem(nd1)
    h[o + 448:o + 576] = _g2_lem(nd2)
    dinv = bn.inv_mod(d, R)

    def scale(sec):
        out = bytearray()
        for i in range(0, len(sec), 64):
            out += _g1_lem(bn.G1.mul(_g1_from_lem(sec[i:i + 64]), dinv))
        return bytes(out)

    params = b""
    if name:
        nb = name.encode()[:64]
        params += bytes([1, len(nb)]) + nb
    if beacon is not None:
        params += bytes([2, beacon[1], 3, len(beacon[0])]) + bytes(beacon[0])
    new = {"points": _g1_lem(nd1) + _g1_lem(g1_s) + _g1_lem(g1_sx) + _g2_lem(g2_spx), "transcript": transcript, "params": params}
    sec10 = cs + (len(contribs) + 1).to_bytes(4, "little")
    for c in contribs + [new]:
        sec10 += c["points"] + c["transcript"] + len(c["params"]).to_bytes(4, "little") + c["params"]
    new_secs = dict(secs)
    new_secs[2], new_secs[8], new_secs[9], new_secs[10] = bytes(h), scale(secs[8]), scale(secs[9]), sec10
    out = bytearray(b"zkey" + (1).to_bytes(4, "little") + (10).to_bytes(4, "little"))
    for t in range(1, 11):
        out += t.to_bytes(4, "little") + len(new_secs[t]).to_bytes(8, "little") + new_secs[t]
    return bytes(out), blake(hash_pubkey(new))


def export_vkey(zkey_bytes):
    """snarkjs zkey export verificationkey -> the text of verification_key.json (JSON.stringify(vKey, null, 1))"""
    secs = parse_sections(zkey_bytes)
    h = secs[2]
    o = 4 + 32 + 4 + 32
    npub = int.from_bytes(h[o + 4:o + 8], "little")
    o += 12
    alpha1 = _g1_from_lem(h[o:o + 64])
    beta2 = _g2_from_lem(h[o + 128:o + 256])
    gamma2 = _g2_from_lem(h[o + 256:o + 384])
    delta2 = _g2_from_lem(h[o + 448:o + 576])
    g1j = lambda p: [str(p[0]), str(p[1]), "1"] if p else ["0", "1", "0"]                                       # noqa: E731
    g2j = lambda p: [[str(p[0][0]), str(p[0][1])], [str(p[1][0]), str(p[1][1])], ["1", "0"]]                     # noqa: E731
    ab = bn.pairing_snarkjs(beta2, alpha1)
    tower = [[[str((ab[2 * i + j] + 9 * ab[2 * i + j + 6]) % Q), str(ab[2 * i + j + 6] % Q)] for i in range(3)] for j in range(2)]
    vk = {"protocol": "groth16", "curve": "bn128", "nPublic": npub, "vk_alpha_1": g1j(alpha1), "vk_beta_2": g2j(beta2),
          "vk_gamma_2": g2j(gamma2), "vk_delta_2": g2j(delta2), "vk_alphabeta_12": tower,
          "IC": [g1j(_g1_from_lem(secs[3][64 * i:64 * i + 64])) for i in range(npub + 1)]}
    return json.dumps(vk, indent=1)


# ---- zkey verify (scripts/g16_verify.sh:190-202: `snarkjs zkey verify <r1cs> <ptau> <zkey>`) ------------------------------
def same_ratio(a1, a2, b1, b2):
    """e(a1, b2) == e(a2, b1): (a1, a2) in G1 and (b1, b2) in G2 differ by the same factor"""
    return bn.pairing(b2, a1) == bn.pairing(b1, a2)


def parse_params(params):
    """-> (name | None, (beacon hash, exp) | None) of a contribution's parameter bytes"""
    name, exp, bh, pos = None, None, None, 0
    while pos < len(params):
        t = params[pos]
        if t == 1:
            n = params[pos + 1]
            name = params[pos + 2:pos + 2 + n].decode()
            pos += 2 + n
        elif t == 2:
            exp = params[pos + 1]
            pos += 2
        elif t == 3:
            n = params[pos + 1]
            bh = bytes(params[pos + 2:pos + 2 + n])
            pos += 2 + n
        else:
            raise ValueError("zkey: unknown contribution parameter %d" % t)
    return name, ((bh, exp) if bh is not None and exp is not None else None)


def verify_from_init(init_zkey, zkey, rng):
    """snarkjs zkey_verify_frominit.js on THIS repository's transcript conventions: `init_zkey` is what `zkey new` makes from the
    circuit and the ceremony file (oracle/zkey_new.py), `zkey` the key under test.  -> None when the key is the initial key
    plus a valid chain of contributions, else a string naming the first check that fails.  rng: random.Random for the
    coefficients of the random linear combinations over sections 8 and 9."""
    si, sz = parse_sections(init_zkey), parse_sections(zkey)
    o = 4 + 32 + 4 + 32 + 12
    if sz[1] != si[1]:
        return "not a groth16 key"
    if sz[2][:o + 384] != si[2][:o + 384]:
        return "header differs from the initial key (sizes, alpha, beta or gamma)"
    for t, what in ((3, "IC"), (4, "coefficients"), (5, "A"), (6, "B1"), (7, "B2")):
        if sz[t] != si[t]:
            return "section %d (%s) differs from the initial key" % (t, what)
    cs_i, _ = parse_mpc(si[10])
    cs, contribs = parse_mpc(sz[10])
    if cs != cs_i:
        return "circuit hash differs from the initial key"
    cur = _g1_from_lem(si[2][o + 384:o + 448])
    seen = []
    for i, c in enumerate(contribs):
        pts = c["points"]
        after, g1_s, g1_sx = (_g1_from_lem(pts[64 * k:64 * k + 64]) for k in range(3))
        g2_spx = _g2_from_lem(pts[192:320])
        if None in (after, g1_s, g1_sx, g2_spx):
            return "contribution %d: point at infinity" % (i + 1)
        transcript = blake(cs, *[hash_pubkey(p) for p in seen], hash_g1(g1_s), hash_g1(g1_sx))
        if transcript != c["transcript"]:
            return "contribution %d: transcript mismatch" % (i + 1)
        g2_sp = hash_to_g2(transcript)
        if not same_ratio(g1_s, g1_sx, g2_sp, g2_spx):
            return "contribution %d: invalid proof of knowledge" % (i + 1)
        if not same_ratio(cur, after, g2_sp, g2_spx):
            return "contribution %d: delta does not follow from the previous one" % (i + 1)
        _, beacon = parse_params(c["params"])
        if beacon is not None:
            d, s = derive(beacon_seed(*beacon))
            want_s = bn.G1.mul(bn.G1_GEN, s)
            if g1_s != want_s or g1_sx != bn.G1.mul(want_s, d):
                return "contribution %d: not the beacon it claims to be" % (i + 1)
        cur = after
        seen.append(c)
    delta1 = _g1_from_lem(sz[2][o + 384:o + 448])
    delta2 = _g2_from_lem(sz[2][o + 448:o + 576])
    if delta1 != cur:
        return "delta in the header is not the last contribution's"
    if not same_ratio(bn.G1_GEN, delta1, bn.G2_GEN, delta2):
        return "delta1 and delta2 differ"
    for t, what in ((8, "C"), (9, "H")):
        if len(sz[t]) != len(si[t]):
            return "section %d (%s) has the wrong size" % (t, what)
        n = len(sz[t]) // 64
        r = [rng.getrandbits(128) | 1 for _ in range(n)]
        a1 = bn.G1.msm([_g1_from_lem(si[t][64 * k:64 * k + 64]) for k in range(n)], r)
        a2 = bn.G1.msm([_g1_from_lem(sz[t][64 * k:64 * k + 64]) for k in range(n)], r)
        if (a1 is None) != (a2 is None) or (a1 is not None and not same_ratio(a1, a2, delta2, bn.G2_GEN)):
            return "section %d (%s) is not the initial section divided by delta" % (t, what)
    return None
