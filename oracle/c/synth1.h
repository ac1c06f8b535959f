/*
 * synth1.h -- SYNTH-1 fixture generator on the host cores (included at the end of oracle.c).
 *
 * ORACLE CODE: TEST INFRASTRUCTURE.  Third implementation of the known-trapdoor fixture factory, bit for bit equal to
 * oracle/factory.py (Python big-int, the specification) and to zk-proof-of-assets_b200/csrc/synth.cu (CUDA, bench
 * support).  It exists so that `bench.py --impl reference` and the CPU tests can make the .zkey/.wtns of a BASELINE
 * workload without loading any of the repository's CUDA libraries and without touching the GPU.  The reference step it
 * stands in for is `snarkjs zkey new` + the circom witness generator (scripts/g16_setup.sh:240-283,
 * scripts/g16_prove.sh:228-239 of silversixpence-crypto/zk-proof-of-assets); see oracle/factory.py for the circuit spec.
 */

static inline uint64_t sy_mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
static inline uint64_t sy_h64(uint64_t seed, uint64_t tag, uint64_t i, uint64_t k) {
    return sy_mix64(sy_mix64(sy_mix64(seed ^ (tag * 0xD6E8FEB86659FD93ull)) + i) + k);
}
static inline fe sy_hfield(uint64_t seed, uint64_t tag, uint64_t i, uint64_t k) { /* 253 bits, standard form */
    fe v;
    for (int j = 0; j < 4; j++) v.v[j] = sy_h64(seed, tag, i, 4 * k + (uint64_t)j);
    v.v[3] &= (1ull << 61) - 1;
    return v;
}
static inline fe sy_small(uint64_t x) { fe v = {{x, 0, 0, 0}}; return v; }
static inline uint64_t sy_hsmall(uint64_t seed, uint64_t tag, uint64_t i, uint64_t k) {
    return (sy_h64(seed, tag, i, k) & 0x7FFFFFFFull) | 1;
}

enum { SY_BIT = 1, SY_SMALL = 2, SY_F1 = 3, SY_F2 = 4, SY_FILL = 5, SY_TRAP = 100 };
enum { SYC_ONE, SYC_BIT, SYC_SMALL, SYC_F1, SYC_F2 };

typedef struct { uint32_t wire; fe coef; } sy_term; /* coef in standard form */
typedef struct { sy_term A[2], B[2], C[2]; int nA, nB, nC; } sy_cons;
typedef struct {
    uint32_t n_vars, n_public, n_constraints, nblocks, domain;
    int log_n;
    uint64_t seed;
} sy_circ;

static inline uint32_t sy_pick(const sy_circ* c, uint32_t lo, uint32_t hi, uint64_t h) {
    uint32_t q = (uint32_t)((h & 0xFFFFFFFFull) % c->nblocks);
    uint32_t cls = lo + (uint32_t)((h >> 32) % (hi - lo));
    return c->n_public + 1 + 20 * q + cls;
}
static inline int sy_class(const sy_circ* c, uint32_t i) {
    if (i == 0) return SYC_ONE;
    if (i <= c->n_public) return SYC_F2;
    uint32_t u = (i - c->n_public - 1) % 20;
    return u < 11 ? SYC_BIT : u < 18 ? SYC_SMALL : u == 18 ? SYC_F1 : SYC_F2;
}
static void sy_constraint(const sy_circ* ci, uint32_t c, sy_cons* k) {
    const uint64_t sd = ci->seed;
    k->nA = k->nB = k->nC = 0;
#define T_(arr, cnt, w, co) do { k->arr[k->cnt].wire = (w); k->arr[k->cnt].coef = (co); k->cnt++; } while (0)
    if (c < ci->n_vars - 1) {
        uint32_t i = c + 1;
        int cl = sy_class(ci, i);
        if (cl == SYC_BIT) {
            fe m1;
            memcpy(m1.v, FR.p, 32);
            m1.v[0] -= 1;
            T_(A, nA, i, sy_small(1));
            T_(B, nB, i, sy_small(1));
            T_(B, nB, 0, m1);
        } else if (cl == SYC_SMALL) {
            uint32_t b0 = sy_pick(ci, 0, 11, sy_h64(sd, SY_SMALL, i, 0));
            uint32_t b1 = sy_pick(ci, 0, 11, sy_h64(sd, SY_SMALL, i, 1));
            T_(A, nA, b0, sy_small(sy_hsmall(sd, SY_SMALL, i, 2)));
            T_(A, nA, b1, sy_small(sy_hsmall(sd, SY_SMALL, i, 3)));
            T_(B, nB, 0, sy_small(sy_hsmall(sd, SY_SMALL, i, 4)));
            T_(C, nC, i, sy_small(1));
        } else if (cl == SYC_F1) {
            uint32_t s0 = sy_pick(ci, 0, 18, sy_h64(sd, SY_F1, i, 0));
            uint32_t s1 = sy_pick(ci, 11, 18, sy_h64(sd, SY_F1, i, 1));
            uint32_t t0 = sy_pick(ci, 11, 18, sy_h64(sd, SY_F1, i, 2));
            T_(A, nA, s0, sy_hfield(sd, SY_F1, i, 1));
            T_(A, nA, s1, sy_hfield(sd, SY_F1, i, 2));
            T_(B, nB, t0, sy_hfield(sd, SY_F1, i, 3));
            T_(C, nC, i, sy_small(1));
        } else {
            uint32_t s0 = sy_pick(ci, 18, 19, sy_h64(sd, SY_F2, i, 0));
            uint32_t s1 = sy_pick(ci, 0, 19, sy_h64(sd, SY_F2, i, 1));
            uint32_t t0 = sy_pick(ci, 18, 19, sy_h64(sd, SY_F2, i, 2));
            uint32_t t1 = sy_pick(ci, 0, 11, sy_h64(sd, SY_F2, i, 3));
            T_(A, nA, s0, sy_hfield(sd, SY_F2, i, 1));
            T_(A, nA, s1, sy_small(sy_hsmall(sd, SY_F2, i, 12)));
            T_(B, nB, t0, sy_hfield(sd, SY_F2, i, 2));
            T_(B, nB, t1, sy_small(sy_hsmall(sd, SY_F2, i, 13)));
            T_(C, nC, i, sy_small(1));
        }
        return;
    }
    uint32_t s0 = sy_pick(ci, 0, 20, sy_h64(sd, SY_FILL, c, 0));
    uint32_t s1 = sy_pick(ci, 0, 20, sy_h64(sd, SY_FILL, c, 1));
    fe a0 = sy_hfield(sd, SY_FILL, c, 1), a1 = sy_small(sy_hsmall(sd, SY_FILL, c, 8));
    T_(A, nA, s0, a0);
    T_(A, nA, s1, a1);
    T_(B, nB, 0, sy_small(1));
    T_(C, nC, s0, a0);
    T_(C, nC, s1, a1);
#undef T_
}

static int sy_circ_init(sy_circ* ci, uint32_t nv, uint32_t npub, uint32_t ncons, uint64_t seed) {
    if (nv < npub + 21 || ncons + 1 < nv) return -1;
    ci->n_vars = nv; ci->n_public = npub; ci->n_constraints = ncons; ci->seed = seed;
    ci->nblocks = (nv - npub - 1) / 20;
    uint64_t need = (uint64_t)ncons + npub + 1;
    ci->log_n = 1;
    while (((uint64_t)1 << ci->log_n) < need) ci->log_n++;
    if (ci->log_n > 27) return -1;
    ci->domain = 1u << ci->log_n;
    return 0;
}

/* ---- [L_c(tau)] on the 2^log_n domain: w^c (tau^n - 1) / (n (tau - w^c)), every step over all cores -------------- */
typedef struct { fe *pw, *den, *pref, *out; fe tau, nn, zt; size_t n; } sy_lag;
static void sy_lag_den_range(size_t lo, size_t hi, void* p) {
    sy_lag* x = (sy_lag*)p;
    for (size_t c = lo; c < hi; c++) {
        fe t;
        fr_sub(&t, &x->tau, &x->pw[c]);
        fr_mul(&x->den[c], &x->nn, &t);
    }
}
static void sy_lag_inv_range(size_t lo, size_t hi, void* p) { /* one batch inversion per slice */
    sy_lag* x = (sy_lag*)p;
    if (lo >= hi) return;
    fe acc, inv;
    fr_one(&acc);
    for (size_t c = lo; c < hi; c++) {
        x->pref[c] = acc;
        fr_mul(&acc, &acc, &x->den[c]);
    }
    fr_inv(&inv, &acc);
    for (size_t c = hi; c-- > lo;) {
        fe di, t;
        fr_mul(&di, &inv, &x->pref[c]);
        fr_mul(&inv, &inv, &x->den[c]);
        fr_mul(&t, &x->pw[c], &x->zt);
        fr_mul(&x->out[c], &t, &di);
    }
}
static fe* sy_lagrange_at(const fe* tau, int log_n) {
    size_t n = (size_t)1 << log_n;
    sy_lag x;
    x.n = n;
    x.tau = *tau;
    x.pw = (fe*)malloc(n * sizeof(fe));
    x.den = (fe*)malloc(n * sizeof(fe));
    x.pref = (fe*)malloc(n * sizeof(fe));
    x.out = (fe*)malloc(n * sizeof(fe));
    fe w, tn = *tau, one;
    fr_root_of_unity(&w, log_n);
    make_twiddles(x.pw, n, &w);
    for (int i = 0; i < log_n; i++) fr_sqr(&tn, &tn);
    fr_one(&one);
    fr_sub(&x.zt, &tn, &one);
    fe nn = {{(uint64_t)n, 0, 0, 0}};
    f_to_mont(&x.nn, &nn, &FR);
    par_range(n, 1024, sy_lag_den_range, &x);
    par_range(n, 1024, sy_lag_inv_range, &x);
    free(x.pw); free(x.den); free(x.pref);
    return x.out;
}

typedef struct {
    sy_circ ci;
    fe tau, alpha, beta, gamma, delta; /* Montgomery */
    fe *w, *a, *b, *c, *hs;            /* Montgomery */
    fe At, Bt, Ct, priv;
    int level;
} sy_state;

static void sy_witness_range(size_t lo, size_t hi, void* p) {
    sy_state* s = (sy_state*)p;
    sy_cons k;
    for (size_t i = lo ? lo : 1; i < hi; i++) {
        if (sy_class(&s->ci, (uint32_t)i) != s->level) continue;
        if (s->level == SYC_BIT) {
            if (sy_h64(s->ci.seed, SY_BIT, i, 0) & 1) fr_one(&s->w[i]);
            else memset(&s->w[i], 0, sizeof(fe));
            continue;
        }
        sy_constraint(&s->ci, (uint32_t)i - 1, &k);
        fe va, vb, co, t;
        memset(&va, 0, sizeof(va));
        memset(&vb, 0, sizeof(vb));
        for (int j = 0; j < k.nA; j++) { f_to_mont(&co, &k.A[j].coef, &FR); fr_mul(&t, &co, &s->w[k.A[j].wire]); fr_add(&va, &va, &t); }
        for (int j = 0; j < k.nB; j++) { f_to_mont(&co, &k.B[j].coef, &FR); fr_mul(&t, &co, &s->w[k.B[j].wire]); fr_add(&vb, &vb, &t); }
        fr_mul(&s->w[i], &va, &vb);
    }
}

static void sy_trap(fe* out, uint64_t seed, uint64_t i) {
    fe v = sy_hfield(seed, SY_TRAP, i, 0);
    v.v[0] |= 2;
    f_to_mont(out, &v, &FR);
}

static void sy_compute(sy_state* s, uint64_t td_seed) {
    sy_circ* ci = &s->ci;
    uint32_t nv = ci->n_vars;
    sy_trap(&s->tau, td_seed, 0);
    sy_trap(&s->alpha, td_seed, 1);
    sy_trap(&s->beta, td_seed, 2);
    sy_trap(&s->gamma, td_seed, 3);
    sy_trap(&s->delta, td_seed, 4);
    s->w = (fe*)calloc(nv, sizeof(fe));
    s->a = (fe*)calloc(nv, sizeof(fe));
    s->b = (fe*)calloc(nv, sizeof(fe));
    s->c = (fe*)calloc(nv, sizeof(fe));
    fr_one(&s->w[0]);
    /* witness by dependency level of the wire classes (a wire only reads wires of lower classes) */
    for (s->level = SYC_BIT; s->level <= SYC_F2; s->level++) par_range(nv, 4096, sy_witness_range, s);
    fe* L = sy_lagrange_at(&s->tau, ci->log_n);
    sy_cons k;
    memset(&s->Ct, 0, sizeof(fe));
    for (uint32_t cc = 0; cc < ci->n_constraints; cc++) {
        sy_constraint(ci, cc, &k);
        fe va, vb, co, t;
        memset(&va, 0, sizeof(va));
        memset(&vb, 0, sizeof(vb));
        for (int j = 0; j < k.nA; j++) {
            f_to_mont(&co, &k.A[j].coef, &FR);
            fr_mul(&t, &co, &L[cc]); fr_add(&s->a[k.A[j].wire], &s->a[k.A[j].wire], &t);
            fr_mul(&t, &co, &s->w[k.A[j].wire]); fr_add(&va, &va, &t);
        }
        for (int j = 0; j < k.nB; j++) {
            f_to_mont(&co, &k.B[j].coef, &FR);
            fr_mul(&t, &co, &L[cc]); fr_add(&s->b[k.B[j].wire], &s->b[k.B[j].wire], &t);
            fr_mul(&t, &co, &s->w[k.B[j].wire]); fr_add(&vb, &vb, &t);
        }
        for (int j = 0; j < k.nC; j++) {
            f_to_mont(&co, &k.C[j].coef, &FR);
            fr_mul(&t, &co, &L[cc]); fr_add(&s->c[k.C[j].wire], &s->c[k.C[j].wire], &t);
        }
        fr_mul(&t, &va, &vb);
        fr_mul(&t, &t, &L[cc]);
        fr_add(&s->Ct, &s->Ct, &t);
    }
    for (uint32_t i = 0; i <= ci->n_public; i++) fr_add(&s->a[i], &s->a[i], &L[ci->n_constraints + i]);
    free(L);
    memset(&s->At, 0, sizeof(fe));
    memset(&s->Bt, 0, sizeof(fe));
    memset(&s->priv, 0, sizeof(fe));
    for (uint32_t i = 0; i < nv; i++) {
        fe t, m;
        fr_mul(&t, &s->w[i], &s->a[i]); fr_add(&s->At, &s->At, &t);
        fr_mul(&t, &s->w[i], &s->b[i]); fr_add(&s->Bt, &s->Bt, &t);
        if (i > ci->n_public) {
            fr_mul(&m, &s->beta, &s->a[i]);
            fr_mul(&t, &s->alpha, &s->b[i]); fr_add(&m, &m, &t);
            fr_add(&m, &m, &s->c[i]);
            fr_mul(&t, &s->w[i], &m); fr_add(&s->priv, &s->priv, &t);
        }
    }
    /* H_j = L^(2n)_{2j+1}(tau) / delta */
    fe* L2 = sy_lagrange_at(&s->tau, ci->log_n + 1);
    fe dinv;
    fr_inv(&dinv, &s->delta);
    s->hs = (fe*)malloc((size_t)ci->domain * sizeof(fe));
    for (uint32_t j = 0; j < ci->domain; j++) fr_mul(&s->hs[j], &L2[2 * (size_t)j + 1], &dinv);
    free(L2);
}
static void sy_free(sy_state* s) { free(s->w); free(s->a); free(s->b); free(s->c); free(s->hs); }

static uint64_t sy_count_coefs(const sy_circ* ci) {
    uint64_t n = 0;
    sy_cons k;
    for (uint32_t c = 0; c < ci->n_constraints; c++) {
        sy_constraint(ci, c, &k);
        n += (uint64_t)(k.nA + k.nB);
    }
    return n + ci->n_public + 1;
}

typedef struct { uint64_t ncoef, off[11], size[11], total; } sy_layout;
static void sy_layout_of(const sy_circ* ci, sy_layout* l) {
    uint64_t nv = ci->n_vars, np = ci->n_public, n = ci->domain;
    l->ncoef = sy_count_coefs(ci);
    l->size[1] = 4;
    l->size[2] = 4 + 32 + 4 + 32 + 12 + 64 + 64 + 128 + 128 + 64 + 128;
    l->size[3] = (np + 1) * 64;
    l->size[4] = 4 + l->ncoef * 44;
    l->size[5] = nv * 64;
    l->size[6] = nv * 64;
    l->size[7] = nv * 128;
    l->size[8] = (nv - np - 1) * 64;
    l->size[9] = n * 64;
    l->size[10] = 68;
    uint64_t pos = 12;
    for (int s = 1; s <= 10; s++) {
        l->off[s] = pos + 12;
        pos += 12 + l->size[s];
    }
    l->total = pos;
}

static void sy_std_scalars(uint8_t* out, const fe* mont, size_t n) {
    for (size_t i = 0; i < n; i++) {
        fe v;
        f_from_mont(&v, &mont[i], &FR);
        memcpy(out + 32 * i, v.v, 32);
    }
}
static void sy_g1_generator(g1_aff* g) {
    fq_one(&g->x);
    fq_add(&g->y, &g->x, &g->x);
    g->inf = 0;
}
static void sy_g2_generator(g2_aff* g) {
    static const uint64_t c[4][4] = {{0x46debd5cd992f6edull, 0x674322d4f75edaddull, 0x426a00665e5c4479ull, 0x1800deef121f1e76ull},
                                     {0x97e485b7aef312c2ull, 0xf1aa493335a9e712ull, 0x7260bfb731fb5d25ull, 0x198e9393920d483aull},
                                     {0x4ce6cc0166fa7daaull, 0xe3d1e7690c43d37bull, 0x4aab71808dcb408full, 0x12c85ea5db8c6debull},
                                     {0x55acdadcd122975bull, 0xbc4b313370b38ef3ull, 0xec9e99ad690c3395ull, 0x090689d0585ff075ull}};
    fe m[4];
    for (int i = 0; i < 4; i++) { memcpy(m[i].v, c[i], 32); f_to_mont(&m[i], &m[i], &FQ); }
    g->x.c0 = m[0]; g->x.c1 = m[1]; g->y.c0 = m[2]; g->y.c1 = m[3];
    g->inf = 0;
}

static inline void sy_w32(uint8_t** p, uint32_t v) { memcpy(*p, &v, 4); *p += 4; }
static inline void sy_w64(uint8_t** p, uint64_t v) { memcpy(*p, &v, 8); *p += 8; }
static inline void sy_w256(uint8_t** p, const uint64_t v[4]) { memcpy(*p, v, 32); *p += 32; }

typedef struct { uint8_t* dst; const fe* src; } sy_std_ctx;
static void sy_std_range(size_t lo, size_t hi, void* p) {
    sy_std_ctx* c = (sy_std_ctx*)p;
    sy_std_scalars(c->dst + 32 * lo, c->src + lo, hi - lo);
}
/* Montgomery Fr scalars -> k*G records at dst */
static void sy_emit_g1(const g1_xy* t1, const fe* mont, size_t n, uint8_t* dst) {
    uint8_t* sd = (uint8_t*)malloc(n * 32 + 32);
    sy_std_ctx c = {sd, mont};
    par_range(n, 4096, sy_std_range, &c);
    g1_fb_mul(t1, sd, n, dst);
    free(sd);
}
static void sy_emit_g2(const g2_xy* t2, const fe* mont, size_t n, uint8_t* dst) {
    uint8_t* sd = (uint8_t*)malloc(n * 32 + 32);
    sy_std_ctx c = {sd, mont};
    par_range(n, 4096, sy_std_range, &c);
    g2_fb_mul(t2, sd, n, dst);
    free(sd);
}

static void sy_mixed(fe* out, const sy_state* s, uint32_t i) {
    fe t;
    fr_mul(out, &s->beta, &s->a[i]);
    fr_mul(&t, &s->alpha, &s->b[i]);
    fr_add(out, out, &t);
    fr_add(out, out, &s->c[i]);
}

EXPORT int orc_synth_sizes(uint32_t nv, uint32_t npub, uint32_t ncons, uint64_t* zkey_bytes, uint64_t* wtns_bytes,
                           uint32_t* domain_size) {
    sy_circ ci;
    if (sy_circ_init(&ci, nv, npub, ncons, 0)) return 1;
    sy_layout l;
    sy_layout_of(&ci, &l);
    if (zkey_bytes) *zkey_bytes = l.total;
    if (wtns_bytes) *wtns_bytes = 12 + 12 + 40 + 12 + (uint64_t)nv * 32;
    if (domain_size) *domain_size = ci.domain;
    return 0;
}

/* Fills caller buffers (any may be NULL) with: the bytes of the .zkey, the bytes of the .wtns, the verification key
 * points in standard form (alpha1 64 | beta2 128 | gamma2 128 | delta2 128 | IC (nPublic+1) x 64: the layout of
 * g16_synth_vkey) and the four exponent sums the closed-form proof needs (A(tau), B(tau), C(tau), private sum;
 * 4 x 32 B LE standard form). */
EXPORT int orc_synth_write(uint32_t nv, uint32_t npub, uint32_t ncons, uint64_t circuit_seed, uint64_t trapdoor_seed,
                           uint8_t* zkey_out, uint8_t* wtns_out, uint8_t* vk_out, uint8_t* sums_out) {
    sy_state s;
    memset(&s, 0, sizeof(s));
    if (sy_circ_init(&s.ci, nv, npub, ncons, circuit_seed)) return 1;
    sy_compute(&s, trapdoor_seed);
    sy_circ* ci = &s.ci;
    if (wtns_out) {
        uint8_t* p = wtns_out;
        memcpy(p, "wtns", 4); p += 4;
        sy_w32(&p, 2); sy_w32(&p, 2);
        sy_w32(&p, 1); sy_w64(&p, 40);
        sy_w32(&p, 32); sy_w256(&p, FR.p); sy_w32(&p, nv);
        sy_w32(&p, 2); sy_w64(&p, (uint64_t)nv * 32);
        sy_std_ctx c = {p, s.w};
        par_range(nv, 4096, sy_std_range, &c);
    }
    if (sums_out) {
        fe v[4] = {s.At, s.Bt, s.Ct, s.priv};
        sy_std_scalars(sums_out, v, 4);
    }
    if (zkey_out || vk_out) {
        g1_aff gen1;
        g2_aff gen2;
        sy_g1_generator(&gen1);
        sy_g2_generator(&gen2);
        g1_xy* t1 = g1_fb_table(&gen1);
        g2_xy* t2 = g2_fb_table(&gen2);
        fe dinv, ginv;
        fr_inv(&dinv, &s.delta);
        fr_inv(&ginv, &s.gamma);
        uint8_t g1h[3 * 64], g2h[3 * 128];
        fe h1[3] = {s.alpha, s.beta, s.delta}, h2[3] = {s.beta, s.gamma, s.delta};
        sy_emit_g1(t1, h1, 3, g1h);
        sy_emit_g2(t2, h2, 3, g2h);
        fe* sc = (fe*)malloc((size_t)nv * sizeof(fe));
        uint8_t* ic = (uint8_t*)malloc((size_t)(npub + 1) * 64);
        for (uint32_t i = 0; i <= npub; i++) { sy_mixed(&sc[i], &s, i); fr_mul(&sc[i], &sc[i], &ginv); }
        sy_emit_g1(t1, sc, npub + 1, ic);
        if (vk_out) { /* Montgomery records -> standard form */
            uint8_t* rec = (uint8_t*)malloc(64 + 384 + (size_t)(npub + 1) * 64);
            memcpy(rec, g1h, 64);
            memcpy(rec + 64, g2h, 384);
            memcpy(rec + 448, ic, (size_t)(npub + 1) * 64);
            size_t words = (64 + 384 + (size_t)(npub + 1) * 64) / 32;
            for (size_t i = 0; i < words; i++) {
                fe v;
                memcpy(v.v, rec + 32 * i, 32);
                f_from_mont(&v, &v, &FQ);
                memcpy(vk_out + 32 * i, v.v, 32);
            }
            free(rec);
        }
        if (zkey_out) {
            sy_layout l;
            sy_layout_of(ci, &l);
            uint8_t* p = zkey_out;
            memcpy(p, "zkey", 4); p += 4;
            sy_w32(&p, 1); sy_w32(&p, 10);
            for (int sec = 1; sec <= 10; sec++) {
                uint8_t* h = zkey_out + l.off[sec] - 12;
                sy_w32(&h, (uint32_t)sec);
                sy_w64(&h, l.size[sec]);
            }
            p = zkey_out + l.off[1];
            sy_w32(&p, 1);
            p = zkey_out + l.off[2];
            sy_w32(&p, 32); sy_w256(&p, FQ.p);
            sy_w32(&p, 32); sy_w256(&p, FR.p);
            sy_w32(&p, nv); sy_w32(&p, npub); sy_w32(&p, ci->domain);
            memcpy(p, g1h, 64); p += 64;          /* alpha1 */
            memcpy(p, g1h + 64, 64); p += 64;     /* beta1 */
            memcpy(p, g2h, 128); p += 128;        /* beta2 */
            memcpy(p, g2h + 128, 128); p += 128;  /* gamma2 */
            memcpy(p, g1h + 128, 64); p += 64;    /* delta1 */
            memcpy(p, g2h + 256, 128); p += 128;  /* delta2 */
            memcpy(zkey_out + l.off[3], ic, (size_t)(npub + 1) * 64);
            p = zkey_out + l.off[4];
            sy_w32(&p, (uint32_t)l.ncoef);
            sy_cons k;
            for (uint32_t c = 0; c <= ci->n_constraints + npub; c++) {
                if (c < ci->n_constraints) sy_constraint(ci, c, &k);
                else { k.nA = 1; k.nB = 0; k.A[0].wire = c - ci->n_constraints; k.A[0].coef = sy_small(1); }
                for (int m = 0; m < 2; m++) {
                    const sy_term* t = m ? k.B : k.A;
                    for (int j = 0; j < (m ? k.nB : k.nA); j++) {
                        fe v;
                        f_to_mont(&v, &t[j].coef, &FR);
                        f_to_mont(&v, &v, &FR); /* coef * R^2 */
                        sy_w32(&p, (uint32_t)m); sy_w32(&p, c); sy_w32(&p, t[j].wire); sy_w256(&p, v.v);
                    }
                }
            }
            sy_emit_g1(t1, s.a, nv, zkey_out + l.off[5]);
            sy_emit_g1(t1, s.b, nv, zkey_out + l.off[6]);
            sy_emit_g2(t2, s.b, nv, zkey_out + l.off[7]);
            for (uint32_t i = npub + 1; i < nv; i++) { sy_mixed(&sc[i - npub - 1], &s, i); fr_mul(&sc[i - npub - 1], &sc[i - npub - 1], &dinv); }
            sy_emit_g1(t1, sc, nv - npub - 1, zkey_out + l.off[8]);
            sy_emit_g1(t1, s.hs, ci->domain, zkey_out + l.off[9]);
            memset(zkey_out + l.off[10], 0, 68);
        }
        free(sc); free(ic); free(t1); free(t2);
    }
    sy_free(&s);
    return 0;
}
