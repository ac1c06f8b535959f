"""iden3 binfile container + .zkey / .wtns readers and writers (oracle; test infrastructure).

Restates rows a1-a3 of SURVEY.md section 8(a) (upstream homes: @iden3/binfileutils 0.0.11,
snarkjs 0.7.2 src/zkey_utils.js + src/wtns_utils.js; call site scripts/g16_prove.sh:246-260):

  binfile : magic[4] | u32 version | u32 nSections | { u32 type | u64 size | payload }*   (little-endian)
  .wtns v2: sec1 { u32 n8 | prime | u32 nWitness }  sec2 { nWitness x n8 bytes, standard form }
  .zkey v1: sec1 { u32 protocol=1 }  sec2 header  sec3 IC  sec4 coefs  sec5 A  sec6 B1  sec7 B2  sec8 C  sec9 H
            points affine, Montgomery (x 2^256), LE; G2 = x.c0 x.c1 y.c0 y.c1; infinity = all zero.
"""
import struct

from . import bn254 as bn

RINV_Q = bn.inv_mod(bn.MONT_R_Q, bn.Q)
RINV_R = bn.inv_mod(bn.MONT_R_R, bn.R)


def read_binfile(data, magic):
    if data[:4] != magic:
        raise ValueError("bad magic %r (want %r)" % (data[:4], magic))
    version, nsec = struct.unpack_from("<II", data, 4)
    pos = 12
    sections = {}
    for _ in range(nsec):
        typ, size = struct.unpack_from("<IQ", data, pos)
        pos += 12
        sections.setdefault(typ, []).append((pos, size))
        pos += size
    if pos != len(data):
        raise ValueError("trailing or missing bytes in binfile")
    return version, sections


def write_binfile(magic, version, sections):
    out = [magic, struct.pack("<II", version, len(sections))]
    for typ, payload in sections:
        out.append(struct.pack("<IQ", typ, len(payload)))
        out.append(payload)
    return b"".join(out)


def _le(x, n=32):
    return int(x).to_bytes(n, "little")


def _from_le(b):
    return int.from_bytes(b, "little")


# ---- point codecs (Montgomery affine, LE) ---------------------------------
def enc_g1(p):
    if p is None:
        return bytes(64)
    return _le(p[0] * bn.MONT_R_Q % bn.Q) + _le(p[1] * bn.MONT_R_Q % bn.Q)


def dec_g1(b):
    if b == bytes(64):
        return None
    return (_from_le(b[:32]) * RINV_Q % bn.Q, _from_le(b[32:64]) * RINV_Q % bn.Q)


def enc_g2(p):
    if p is None:
        return bytes(128)
    (x0, x1), (y0, y1) = p
    return b"".join(_le(v * bn.MONT_R_Q % bn.Q) for v in (x0, x1, y0, y1))


def dec_g2(b):
    if b == bytes(128):
        return None
    v = [_from_le(b[i * 32:(i + 1) * 32]) * RINV_Q % bn.Q for i in range(4)]
    return ((v[0], v[1]), (v[2], v[3]))


# ---- wtns -----------------------------------------------------------------
def write_wtns(witness, prime=bn.R):
    sec1 = struct.pack("<I", 32) + _le(prime) + struct.pack("<I", len(witness))
    sec2 = b"".join(_le(w % prime) for w in witness)
    return write_binfile(b"wtns", 2, [(1, sec1), (2, sec2)])


def read_wtns(data):
    version, secs = read_binfile(data, b"wtns")
    if version != 2:
        raise ValueError("wtns version %d" % version)
    p1, _ = secs[1][0]
    n8 = struct.unpack_from("<I", data, p1)[0]
    prime = _from_le(data[p1 + 4:p1 + 4 + n8])
    nw = struct.unpack_from("<I", data, p1 + 4 + n8)[0]
    p2, s2 = secs[2][0]
    if s2 != nw * n8:
        raise ValueError("wtns section 2 size mismatch")
    w = [_from_le(data[p2 + i * n8:p2 + (i + 1) * n8]) for i in range(nw)]
    return {"n8": n8, "prime": prime, "witness": w}


# ---- zkey -----------------------------------------------------------------
class ZKey:
    """Parsed Groth16 proving key (all values out of Montgomery form, as Python ints)."""

    def __init__(self):
        self.n_vars = self.n_public = self.domain_size = 0
        self.alpha1 = self.beta1 = self.beta2 = self.gamma2 = self.delta1 = self.delta2 = None
        self.ic = []
        self.coefs = []  # (m, c, s, coef) with coef in standard form
        self.A = []
        self.B1 = []
        self.B2 = []
        self.C = []
        self.H = []


def write_zkey(zk):
    sec1 = struct.pack("<I", 1)
    sec2 = (struct.pack("<I", 32) + _le(bn.Q) + struct.pack("<I", 32) + _le(bn.R)
            + struct.pack("<III", zk.n_vars, zk.n_public, zk.domain_size)
            + enc_g1(zk.alpha1) + enc_g1(zk.beta1) + enc_g2(zk.beta2) + enc_g2(zk.gamma2)
            + enc_g1(zk.delta1) + enc_g2(zk.delta2))
    sec3 = b"".join(enc_g1(p) for p in zk.ic)
    r2 = bn.MONT_R_R * bn.MONT_R_R % bn.R
    sec4 = struct.pack("<I", len(zk.coefs)) + b"".join(
        struct.pack("<III", m, c, s) + _le(v % bn.R * r2 % bn.R) for (m, c, s, v) in zk.coefs)
    sec5 = b"".join(enc_g1(p) for p in zk.A)
    sec6 = b"".join(enc_g1(p) for p in zk.B1)
    sec7 = b"".join(enc_g2(p) for p in zk.B2)
    sec8 = b"".join(enc_g1(p) for p in zk.C)
    sec9 = b"".join(enc_g1(p) for p in zk.H)
    # sec 10 (contributions): u32 count = 0 after a 64-byte circuit hash, ignored by prove
    sec10 = bytes(64) + struct.pack("<I", 0)
    return write_binfile(b"zkey", 1, [(1, sec1), (2, sec2), (3, sec3), (4, sec4), (5, sec5),
                                      (6, sec6), (7, sec7), (8, sec8), (9, sec9), (10, sec10)])


def read_zkey(data):
    version, secs = read_binfile(data, b"zkey")
    if version != 1:
        raise ValueError("zkey version %d" % version)
    p, _ = secs[1][0]
    if struct.unpack_from("<I", data, p)[0] != 1:
        raise ValueError("zkey protocol is not groth16")
    zk = ZKey()
    p, _ = secs[2][0]
    n8q = struct.unpack_from("<I", data, p)[0]
    q = _from_le(data[p + 4:p + 4 + n8q])
    p += 4 + n8q
    n8r = struct.unpack_from("<I", data, p)[0]
    r = _from_le(data[p + 4:p + 4 + n8r])
    p += 4 + n8r
    if q != bn.Q or r != bn.R:
        raise ValueError("zkey is not over bn128")
    zk.n_vars, zk.n_public, zk.domain_size = struct.unpack_from("<III", data, p)
    p += 12
    zk.alpha1 = dec_g1(data[p:p + 64]); p += 64
    zk.beta1 = dec_g1(data[p:p + 64]); p += 64
    zk.beta2 = dec_g2(data[p:p + 128]); p += 128
    zk.gamma2 = dec_g2(data[p:p + 128]); p += 128
    zk.delta1 = dec_g1(data[p:p + 64]); p += 64
    zk.delta2 = dec_g2(data[p:p + 128]); p += 128

    def pts(sec, size, dec):
        p0, sz = secs[sec][0]
        assert sz % size == 0
        return [dec(data[p0 + i * size:p0 + (i + 1) * size]) for i in range(sz // size)]

    zk.ic = pts(3, 64, dec_g1)
    p0, _ = secs[4][0]
    ncoef = struct.unpack_from("<I", data, p0)[0]
    r2inv = bn.inv_mod(bn.MONT_R_R * bn.MONT_R_R, bn.R)
    for i in range(ncoef):
        o = p0 + 4 + i * 44
        m, c, s = struct.unpack_from("<III", data, o)
        zk.coefs.append((m, c, s, _from_le(data[o + 12:o + 44]) * r2inv % bn.R))
    zk.A = pts(5, 64, dec_g1)
    zk.B1 = pts(6, 64, dec_g1)
    zk.B2 = pts(7, 128, dec_g2)
    zk.C = pts(8, 64, dec_g1)
    zk.H = pts(9, 64, dec_g1)
    return zk
