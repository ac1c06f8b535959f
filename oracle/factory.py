"""Known-trapdoor fixture factory (oracle; test infrastructure).

The reference commits no .zkey/.wtns/.r1cs and none can be produced
offline, so parity fixtures are synthetic: a deterministic R1CS ("SYNTH-1", below) plus
a trusted setup whose trapdoor (tau, alpha, beta, gamma, delta) WE choose.  With the
trapdoor known every proof element has a closed form in the exponent
(SURVEY.md section 8(c)(i)), so a prover's output is checkable bit-exactly with three
scalar multiplications, independent of any MSM/NTT code.

The zkey this emits follows snarkjs `zkey new` conventions:
  A_i = a_i(tau) G1, B1_i = b_i(tau) G1, B2_i = b_i(tau) G2,
  C_i = (beta a_i + alpha b_i + c_i)(tau)/delta G1 (private i), IC_i = (...)/gamma G1 (public i),
  H_j = L^(2n)_{2j+1}(tau)/delta G1  (odd Lagrange basis of the 2n domain),
  section 4 = A and B coefficients (x R^2) plus the rows c = nConstraints+i, s = i, coef 1 for i <= nPublic.

SYNTH-1 circuit (same spec is implemented in oracle/c/oracle.c and in the CUDA fixture
generator csrc/synth.cu; all three must agree bit for bit):
  wire 0 = 1; wires 1..nPublic are public; every wire i >= 1 is defined by constraint i-1;
  class of private wire i: u = i-(nPublic+1), u%20 in 0..10 bit | 11..17 small | 18 F1 | 19 F2;
  public wires are class F2.  Constraints nVars-1 .. nConstraints-1 are linear fillers.
"""
from . import bn254 as bn
from . import groth16 as g16
from .binfile import ZKey

R = bn.R
M64 = (1 << 64) - 1


def mix64(x):
    x = (x + 0x9E3779B97F4A7C15) & M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def h64(seed, tag, i, k):
    return mix64(mix64(mix64(seed ^ (tag * 0xD6E8FEB86659FD93 & M64)) + i & M64) + k & M64)


def hfield(seed, tag, i, k):
    """253-bit field element from four 64-bit hashes (no modular reduction needed)."""
    v = 0
    for j in range(4):
        v |= h64(seed, tag, i, 4 * k + j) << (64 * j)
    return v & ((1 << 253) - 1)


def hsmall(seed, tag, i, k):
    return (h64(seed, tag, i, k) & 0x7FFFFFFF) | 1


TAG_BIT, TAG_SMALL, TAG_F1, TAG_F2, TAG_FILL, TAG_TRAP = 1, 2, 3, 4, 5, 100


class Trapdoor:
    def __init__(self, seed):
        self.tau = hfield(seed, TAG_TRAP, 0, 0) | 2
        self.alpha = hfield(seed, TAG_TRAP, 1, 0) | 2
        self.beta = hfield(seed, TAG_TRAP, 2, 0) | 2
        self.gamma = hfield(seed, TAG_TRAP, 3, 0) | 2
        self.delta = hfield(seed, TAG_TRAP, 4, 0) | 2


class SynthCircuit:
    def __init__(self, n_vars, n_public, n_constraints, seed):
        assert n_vars >= n_public + 21 and n_constraints >= n_vars - 1
        self.n_vars, self.n_public, self.n_constraints, self.seed = n_vars, n_public, n_constraints, seed
        self.nblocks = (n_vars - n_public - 1) // 20
        need = n_constraints + n_public + 1
        self.log_n = max(1, (need - 1).bit_length())
        self.domain_size = 1 << self.log_n
        self._w = {}

    # -- wire selection -------------------------------------------------------
    def pick(self, lo, hi, h):
        q = (h & 0xFFFFFFFF) % self.nblocks
        cls = lo + (h >> 32) % (hi - lo)
        return self.n_public + 1 + 20 * q + cls

    def wire_class(self, i):
        if i == 0:
            return "one"
        if i <= self.n_public:
            return "F2"
        u = (i - self.n_public - 1) % 20
        return "bit" if u < 11 else "small" if u < 18 else "F1" if u == 18 else "F2"

    # -- constraint c -> (A, B, C) as lists of (wire, coef) -------------------
    def constraint(self, c):
        sd = self.seed
        if c < self.n_vars - 1:
            i = c + 1
            k = self.wire_class(i)
            if k == "bit":
                return [(i, 1)], [(i, 1), (0, R - 1)], []
            if k == "small":
                b0 = self.pick(0, 11, h64(sd, TAG_SMALL, i, 0))
                b1 = self.pick(0, 11, h64(sd, TAG_SMALL, i, 1))
                c0, c1, c2 = (hsmall(sd, TAG_SMALL, i, 2 + j) for j in range(3))
                return [(b0, c0), (b1, c1)], [(0, c2)], [(i, 1)]
            if k == "F1":
                s0 = self.pick(0, 18, h64(sd, TAG_F1, i, 0))
                s1 = self.pick(11, 18, h64(sd, TAG_F1, i, 1))
                t0 = self.pick(11, 18, h64(sd, TAG_F1, i, 2))
                a0, a1, b0 = (hfield(sd, TAG_F1, i, 1 + j) for j in range(3))
                return [(s0, a0), (s1, a1)], [(t0, b0)], [(i, 1)]
            s0 = self.pick(18, 19, h64(sd, TAG_F2, i, 0))
            s1 = self.pick(0, 19, h64(sd, TAG_F2, i, 1))
            t0 = self.pick(18, 19, h64(sd, TAG_F2, i, 2))
            t1 = self.pick(0, 11, h64(sd, TAG_F2, i, 3))
            a0, b0 = hfield(sd, TAG_F2, i, 1), hfield(sd, TAG_F2, i, 2)
            a1, b1 = hsmall(sd, TAG_F2, i, 12), hsmall(sd, TAG_F2, i, 13)
            return [(s0, a0), (s1, a1)], [(t0, b0), (t1, b1)], [(i, 1)]
        s0 = self.pick(0, 20, h64(sd, TAG_FILL, c, 0))
        s1 = self.pick(0, 20, h64(sd, TAG_FILL, c, 1))
        a0, a1 = hfield(sd, TAG_FILL, c, 1), hsmall(sd, TAG_FILL, c, 8)
        return [(s0, a0), (s1, a1)], [(0, 1)], [(s0, a0), (s1, a1)]

    # -- witness ------------------------------------------------------------------
    def value(self, i):
        if i in self._w:
            return self._w[i]
        k = self.wire_class(i)
        if k == "one":
            v = 1
        elif k == "bit":
            v = h64(self.seed, TAG_BIT, i, 0) & 1
        else:
            A, B, _ = self.constraint(i - 1)
            va = sum(co * self.value(s) for s, co in A) % R
            vb = sum(co * self.value(s) for s, co in B) % R
            v = va * vb % R
        self._w[i] = v
        return v

    def witness(self):
        return [self.value(i) for i in range(self.n_vars)]

    def coef_records(self):
        """zkey section-4 records (m, c, s, coef), constraint order, A entries then B entries,
        then the public-input rows."""
        out = []
        for c in range(self.n_constraints):
            A, B, _ = self.constraint(c)
            out += [(0, c, s, co) for s, co in A]
            out += [(1, c, s, co) for s, co in B]
        out += [(0, self.n_constraints + i, i, 1) for i in range(self.n_public + 1)]
        return out

    def check_witness(self, w):
        for c in range(self.n_constraints):
            A, B, C = self.constraint(c)
            va = sum(co * w[s] for s, co in A) % R
            vb = sum(co * w[s] for s, co in B) % R
            vc = sum(co * w[s] for s, co in C) % R
            if va * vb % R != vc:
                return False
        return True


class ListCircuit:
    """Explicit R1CS for hand-made tiny cases: constraints = [(A, B, C)] of (wire, coef) lists."""

    def __init__(self, n_vars, n_public, constraints, witness):
        self.n_vars, self.n_public, self._cons, self._wit = n_vars, n_public, constraints, witness
        self.n_constraints = len(constraints)
        need = self.n_constraints + n_public + 1
        self.log_n = max(1, (need - 1).bit_length())
        self.domain_size = 1 << self.log_n

    def constraint(self, c):
        return self._cons[c]

    def witness(self):
        return list(self._wit)

    coef_records = SynthCircuit.coef_records
    check_witness = SynthCircuit.check_witness


def lagrange_at(tau, log_n):
    """[L_c(tau)] for the 2^log_n domain: L_c(tau) = w^c (tau^n - 1) / (n (tau - w^c))."""
    n = 1 << log_n
    w = g16.root_of_unity(log_n)
    zt = (pow(tau, n, R) - 1) % R
    pw = [1] * n
    for c in range(1, n):
        pw[c] = pw[c - 1] * w % R
    den = [(n * (tau - x)) % R for x in pw]
    # batch inversion
    pref = [1] * (n + 1)
    for c in range(n):
        pref[c + 1] = pref[c] * den[c] % R
    inv = bn.inv_mod(pref[n], R)
    out = [0] * n
    for c in range(n - 1, -1, -1):
        out[c] = pw[c] * zt % R * (inv * pref[c] % R) % R
        inv = inv * den[c] % R
    return out


def qap_at_tau(circ, tau):
    """a_i(tau), b_i(tau), c_i(tau) for every wire (including the public-input rows of A)."""
    L = lagrange_at(tau, circ.log_n)
    a = [0] * circ.n_vars
    b = [0] * circ.n_vars
    cc = [0] * circ.n_vars
    for c in range(circ.n_constraints):
        A, B, C = circ.constraint(c)
        for s, co in A:
            a[s] = (a[s] + co * L[c]) % R
        for s, co in B:
            b[s] = (b[s] + co * L[c]) % R
        for s, co in C:
            cc[s] = (cc[s] + co * L[c]) % R
    for i in range(circ.n_public + 1):
        a[i] = (a[i] + L[circ.n_constraints + i]) % R
    return a, b, cc


def setup(circ, td):
    """-> (ZKey, vkey dict) for the circuit under trapdoor td."""
    fb1 = bn.FixedBase(bn.G1)
    fb2 = bn.FixedBase(bn.G2)
    a, b, cc = qap_at_tau(circ, td.tau)
    dinv = bn.inv_mod(td.delta, R)
    ginv = bn.inv_mod(td.gamma, R)
    zk = ZKey()
    zk.n_vars, zk.n_public, zk.domain_size = circ.n_vars, circ.n_public, circ.domain_size
    zk.alpha1 = fb1.mul(td.alpha)
    zk.beta1 = fb1.mul(td.beta)
    zk.beta2 = fb2.mul(td.beta)
    zk.gamma2 = fb2.mul(td.gamma)
    zk.delta1 = fb1.mul(td.delta)
    zk.delta2 = fb2.mul(td.delta)
    mixed = [(td.beta * a[i] + td.alpha * b[i] + cc[i]) % R for i in range(circ.n_vars)]
    zk.ic = [fb1.mul(mixed[i] * ginv % R) for i in range(circ.n_public + 1)]
    zk.coefs = circ.coef_records()
    zk.A = [fb1.mul(v) for v in a]
    zk.B1 = [fb1.mul(v) for v in b]
    zk.B2 = [fb2.mul(v) for v in b]
    zk.C = [fb1.mul(mixed[i] * dinv % R) for i in range(circ.n_public + 1, circ.n_vars)]
    L2 = lagrange_at(td.tau, circ.log_n + 1)
    zk.H = [fb1.mul(L2[2 * j + 1] * dinv % R) for j in range(circ.domain_size)]
    vkey = {
        "protocol": "groth16", "curve": "bn128", "nPublic": circ.n_public,
        "vk_alpha_1": [str(zk.alpha1[0]), str(zk.alpha1[1]), "1"],
        "vk_beta_2": _g2j(zk.beta2), "vk_gamma_2": _g2j(zk.gamma2), "vk_delta_2": _g2j(zk.delta2),
        "IC": [[str(p[0]), str(p[1]), "1"] if p else ["0", "1", "0"] for p in zk.ic],
    }
    return zk, vkey


def _g2j(p):
    return [[str(p[0][0]), str(p[0][1])], [str(p[1][0]), str(p[1][1])], ["1", "0"]]


def closed_form_exponents(circ, w, td, r, s):
    """Discrete logs of (pi_a, pi_b, pi_c) w.r.t. the generators, from the trapdoor alone."""
    a, b, cc = qap_at_tau(circ, td.tau)
    At = sum(x * y for x, y in zip(w, a)) % R
    Bt = sum(x * y for x, y in zip(w, b)) % R
    # C(tau) as the prover sees it: interpolation of A_T[c] * B_T[c] over the domain
    L = lagrange_at(td.tau, circ.log_n)
    Ct = 0
    for c in range(circ.n_constraints):
        A, B, _ = circ.constraint(c)
        va = sum(co * w[i] for i, co in A) % R
        vb = sum(co * w[i] for i, co in B) % R
        Ct = (Ct + va * vb % R * L[c]) % R
    dinv = bn.inv_mod(td.delta, R)
    priv = sum(w[i] * (td.beta * a[i] + td.alpha * b[i] + cc[i]) for i in range(circ.n_public + 1, circ.n_vars)) % R
    ea = (td.alpha + At + r * td.delta) % R
    eb = (td.beta + Bt + s * td.delta) % R
    ec = ((priv + At * Bt - Ct) * dinv + s * ea + r * eb - r * s % R * td.delta) % R
    return ea, eb, ec


def closed_form_proof(circ, w, td, r, s):
    ea, eb, ec = closed_form_exponents(circ, w, td, r, s)
    return bn.G1.mul(bn.G1_GEN, ea), bn.G2.mul(bn.G2_GEN, eb), bn.G1.mul(bn.G1_GEN, ec)
