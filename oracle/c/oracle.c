/*
 * oracle.c -- CPU restatement (plain C + pthreads) of the Groth16 `prove` hot path.
 *
 * ORACLE CODE: TEST INFRASTRUCTURE.  Used only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs, as the checker or the timed CPU baseline.  The product (zk-proof-of-assets_b200/) never
 * links or calls it.
 *
 * The reference (silversixpence-crypto/zk-proof-of-assets) holds no prover source: scripts/g16_prove.sh:246-260 execs
 * rapidsnark (unpinned HEAD, scripts/machine_initialization.sh:301-314) or snarkjs 0.7.2 -> ffjavascript 0.2.62 ->
 * wasmcurves 0.2.2 (pnpm-lock.yaml:2231-2244).  Neither can be built here (no node, nasm, gmp, libsodium; no network),
 * so this file restates their published algorithm, function by function:
 *   binfile sections        @iden3/binfileutils readBinFile / rapidsnark BinFileUtils          (row a1)
 *   zkey + wtns headers     snarkjs zkey_utils.js readHeader, wtns_utils.js readHeader         (rows a2, a3)
 *   build_abc               snarkjs groth16_prove.js buildABC1                                  (row a4)
 *   ntt / coset             ffjavascript engine_fft.js, engine_applykey.js                      (rows a5-a7)
 *   join                    groth16_prove.js joinABC + batchFromMontgomery                      (row a8)
 *   msm                     ffjavascript engine_multiexp.js / rapidsnark multiexp.cpp           (rows a9-a13)
 *   final combination       groth16_prove.js (pi_a, pi_b, pi_c)                                 (row a15)
 * PARITY: pinned against the Python oracle, oracle/bn254.py + groth16.py (which is pinned against the reference's committed vkey/proof triples) by
 * tests/test_oracle_c.py; prover bits themselves are unpinned in the reference (see oracle/__init__.py).
 *
 * Arithmetic: 4 x 64-bit Montgomery (R = 2^256) with unsigned __int128, like rapidsnark's fr.asm/fq.asm.
 * Threads: pthreads over contiguous index ranges (this image's gcc has no libgomp); rapidsnark uses OpenMP the same way.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

typedef unsigned __int128 u128;
typedef struct { uint64_t v[4]; } fe;

typedef struct {
    uint64_t p[4], r1[4], r2[4], inv;
} fparams;

/* ---- threads: fn(lo, hi, ctx) over T contiguous slices of [0, n) ---------------------------------------------- */
static int g_threads = 0;
static int n_threads(void) {
    if (g_threads <= 0) {
        const char* e = getenv("ORC_THREADS");
        g_threads = e ? atoi(e) : (int)sysconf(_SC_NPROCESSORS_ONLN);
        if (g_threads <= 0) g_threads = 1;
        if (g_threads > 256) g_threads = 256;
    }
    return g_threads;
}
typedef void (*range_fn)(size_t lo, size_t hi, void* ctx);
typedef struct { range_fn fn; size_t lo, hi; void* ctx; } par_job;
static void* par_entry(void* p) {
    par_job* j = (par_job*)p;
    j->fn(j->lo, j->hi, j->ctx);
    return NULL;
}
static void par_range(size_t n, size_t min_per_thread, range_fn fn, void* ctx) {
    int T = n_threads();
    if (min_per_thread && (size_t)T > n / min_per_thread + 1) T = (int)(n / min_per_thread + 1);
    if (T <= 1 || n == 0) { fn(0, n, ctx); return; }
    pthread_t th[256];
    par_job job[256];
    for (int t = 0; t < T; t++) {
        job[t].fn = fn; job[t].ctx = ctx;
        job[t].lo = n * (size_t)t / (size_t)T;
        job[t].hi = n * (size_t)(t + 1) / (size_t)T;
        if (t) pthread_create(&th[t], NULL, par_entry, &job[t]);
    }
    par_entry(&job[0]);
    for (int t = 1; t < T; t++) pthread_join(th[t], NULL);
}

static const fparams FQ = {{0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull, 0x30644e72e131a029ull},
                           {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull, 0x0e0a77c19a07df2full},
                           {0xf32cfc5b538afa89ull, 0xb5e71911d44501fbull, 0x47ab1eff0a417ff6ull, 0x06d89f71cab8351full},
                           0x87d20782e4866389ull};
static const fparams FR = {{0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull},
                           {0xac96341c4ffffffbull, 0x36fc76959f60cd29ull, 0x666ea36f7879462eull, 0x0e0a77c19a07df2full},
                           {0x1bb8e645ae216da7ull, 0x53fe3ab1e35c59e3ull, 0x8c49833d53bb8085ull, 0x0216d0b17f4e44a5ull},
                           0xc2e1f593efffffffull};

/* ---- generic prime-field kernels ------------------------------------------------------------------------- */
static inline int ge(const uint64_t a[4], const uint64_t b[4]) {
    for (int i = 3; i >= 0; i--) {
        if (a[i] > b[i]) return 1;
        if (a[i] < b[i]) return 0;
    }
    return 1;
}
static inline void sub_raw(uint64_t r[4], const uint64_t a[4], const uint64_t b[4]) {
    u128 bw = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a[i] - b[i] - bw;
        r[i] = (uint64_t)d;
        bw = (d >> 64) & 1;
    }
}
static inline void f_add(fe* r, const fe* a, const fe* b, const fparams* P) {
    u128 c = 0;
    uint64_t t[4];
    for (int i = 0; i < 4; i++) {
        c += (u128)a->v[i] + b->v[i];
        t[i] = (uint64_t)c;
        c >>= 64;
    }
    if (ge(t, P->p)) sub_raw(r->v, t, P->p);
    else memcpy(r->v, t, 32);
}
static inline void f_sub(fe* r, const fe* a, const fe* b, const fparams* P) {
    uint64_t t[4];
    u128 bw = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a->v[i] - b->v[i] - bw;
        t[i] = (uint64_t)d;
        bw = (d >> 64) & 1;
    }
    if (bw) {
        u128 c = 0;
        for (int i = 0; i < 4; i++) {
            c += (u128)t[i] + P->p[i];
            t[i] = (uint64_t)c;
            c >>= 64;
        }
    }
    memcpy(r->v, t, 32);
}
/* Montgomery product a*b/R mod p (CIOS); a < p, b < 2^256 */
static inline void f_mul(fe* r, const fe* a, const fe* b, const fparams* P) {
    uint64_t t0 = 0, t1 = 0, t2 = 0, t3 = 0, t4 = 0;
    for (int i = 0; i < 4; i++) {
        u128 c;
        uint64_t bi = b->v[i], carry, t5;
        c = (u128)a->v[0] * bi + t0; t0 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)a->v[1] * bi + t1 + carry; t1 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)a->v[2] * bi + t2 + carry; t2 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)a->v[3] * bi + t3 + carry; t3 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)t4 + carry; t4 = (uint64_t)c; t5 = (uint64_t)(c >> 64);
        uint64_t m = t0 * P->inv;
        c = (u128)m * P->p[0] + t0; carry = (uint64_t)(c >> 64);
        c = (u128)m * P->p[1] + t1 + carry; t0 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)m * P->p[2] + t2 + carry; t1 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)m * P->p[3] + t3 + carry; t2 = (uint64_t)c; carry = (uint64_t)(c >> 64);
        c = (u128)t4 + carry; t3 = (uint64_t)c; t4 = t5 + (uint64_t)(c >> 64);
    }
    uint64_t t[4] = {t0, t1, t2, t3};
    if (t4 || ge(t, P->p)) sub_raw(r->v, t, P->p);
    else memcpy(r->v, t, 32);
}
static inline int f_is_zero(const fe* a) { return (a->v[0] | a->v[1] | a->v[2] | a->v[3]) == 0; }
static inline void f_neg(fe* r, const fe* a, const fparams* P) {
    if (f_is_zero(a)) { *r = *a; return; }
    sub_raw(r->v, P->p, a->v);
}
static void f_pow(fe* r, const fe* a, const uint64_t e[4], const fparams* P) {
    fe acc;
    memcpy(acc.v, P->r1, 32);
    for (int i = 255; i >= 0; i--) {
        f_mul(&acc, &acc, &acc, P);
        if ((e[i >> 6] >> (i & 63)) & 1) f_mul(&acc, &acc, a, P);
    }
    *r = acc;
}
static void f_inv(fe* r, const fe* a, const fparams* P) { /* Fermat; inv(0) = 0 */
    uint64_t e[4];
    memcpy(e, P->p, 32);
    e[0] -= 2;
    f_pow(r, a, e, P);
}
static inline void f_to_mont(fe* r, const fe* a, const fparams* P) {
    fe r2;
    memcpy(r2.v, P->r2, 32);
    f_mul(r, &r2, a, P);
}
static inline void f_from_mont(fe* r, const fe* a, const fparams* P) {
    fe one = {{1, 0, 0, 0}};
    f_mul(r, a, &one, P);
}

/* ---- Fq, Fr, Fq2 named wrappers (what curve_tmpl.h binds to) ------------------------------------------------ */
#define WRAP(name, P)                                                                                          \
    static inline void name##_add(fe* r, const fe* a, const fe* b) { f_add(r, a, b, &P); }                     \
    static inline void name##_sub(fe* r, const fe* a, const fe* b) { f_sub(r, a, b, &P); }                     \
    static inline void name##_mul(fe* r, const fe* a, const fe* b) { f_mul(r, a, b, &P); }                     \
    static inline void name##_sqr(fe* r, const fe* a) { f_mul(r, a, a, &P); }                                  \
    static inline void name##_neg(fe* r, const fe* a) { f_neg(r, a, &P); }                                     \
    static inline void name##_inv(fe* r, const fe* a) { f_inv(r, a, &P); }                                     \
    static inline int name##_is_zero(const fe* a) { return f_is_zero(a); }                                     \
    static inline void name##_one(fe* r) { memcpy(r->v, P.r1, 32); }
WRAP(fq, FQ)
WRAP(fr, FR)

typedef struct { fe c0, c1; } fe2; /* Fq[u]/(u^2+1) */
static inline void fq2_add(fe2* r, const fe2* a, const fe2* b) { fq_add(&r->c0, &a->c0, &b->c0); fq_add(&r->c1, &a->c1, &b->c1); }
static inline void fq2_sub(fe2* r, const fe2* a, const fe2* b) { fq_sub(&r->c0, &a->c0, &b->c0); fq_sub(&r->c1, &a->c1, &b->c1); }
static inline void fq2_neg(fe2* r, const fe2* a) { fq_neg(&r->c0, &a->c0); fq_neg(&r->c1, &a->c1); }
static inline void fq2_mul(fe2* r, const fe2* a, const fe2* b) {
    fe t0, t1, s, sa, sb;
    fq_mul(&t0, &a->c0, &b->c0);
    fq_mul(&t1, &a->c1, &b->c1);
    fq_add(&sa, &a->c0, &a->c1);
    fq_add(&sb, &b->c0, &b->c1);
    fq_mul(&s, &sa, &sb);
    fq_sub(&r->c0, &t0, &t1);
    fq_sub(&s, &s, &t0);
    fq_sub(&r->c1, &s, &t1);
}
static inline void fq2_sqr(fe2* r, const fe2* a) {
    fe t, sa, sb;
    fq_mul(&t, &a->c0, &a->c1);
    fq_add(&sa, &a->c0, &a->c1);
    fq_sub(&sb, &a->c0, &a->c1);
    fq_mul(&r->c0, &sa, &sb);
    fq_add(&r->c1, &t, &t);
}
static inline void fq2_inv(fe2* r, const fe2* a) {
    fe d, t;
    fq_sqr(&d, &a->c0);
    fq_sqr(&t, &a->c1);
    fq_add(&d, &d, &t);
    fq_inv(&d, &d);
    fq_mul(&r->c0, &a->c0, &d);
    fq_mul(&t, &a->c1, &d);
    fq_neg(&r->c1, &t);
}
static inline int fq2_is_zero(const fe2* a) { return fq_is_zero(&a->c0) && fq_is_zero(&a->c1); }
static inline void fq2_one(fe2* r) { fq_one(&r->c0); memset(&r->c1, 0, 32); }

/* ---- G1 / G2 ------------------------------------------------------------------------------------------------ */
#define FE fe
#define FN(n) fq_##n
#define CN(n) g1_##n
#include "curve_tmpl.h"
#undef FE
#undef FN
#undef CN
#define FE fe2
#define FN(n) fq2_##n
#define CN(n) g2_##n
#include "curve_tmpl.h"
#undef FE
#undef FN
#undef CN

/* ---- NTT over Fr (rows a5-a7) ------------------------------------------------------------------------------- */
static void fr_root_of_unity(fe* w, int log) { /* ffjavascript: w[28] = 5^((r-1)/2^28), w[k] = w[k+1]^2 */
    uint64_t e[4];
    memcpy(e, FR.p, 32);
    e[0] -= 1;
    for (int i = 0; i < 4; i++) e[i] = (e[i] >> 28) | (i < 3 ? e[i + 1] << 36 : 0);
    fe five = {{5, 0, 0, 0}};
    f_to_mont(&five, &five, &FR);
    f_pow(w, &five, e, &FR);
    for (int i = 28; i > log; i--) fr_sqr(w, w);
}

/* in-place radix-2 DIT, natural order in and out; tw[k] = w^k, k < n/2 */
typedef struct { fe* a; const fe* tw; int logn; size_t half, step; } ntt_ctx;
static void ntt_bitrev_range(size_t lo, size_t hi, void* p) {
    ntt_ctx* c = (ntt_ctx*)p;
    for (size_t i = lo; i < hi; i++) {
        size_t j = 0;
        for (int b = 0; b < c->logn; b++) j |= ((i >> b) & 1) << (c->logn - 1 - b);
        if (i < j) { fe t = c->a[i]; c->a[i] = c->a[j]; c->a[j] = t; }
    }
}
static void ntt_stage_range(size_t lo, size_t hi, void* p) {
    ntt_ctx* c = (ntt_ctx*)p;
    fe* a = c->a;
    for (size_t k = lo; k < hi; k++) {
        size_t grp = k / c->half, j = k % c->half;
        size_t i0 = grp * 2 * c->half + j, i1 = i0 + c->half;
        fe v, u = a[i0];
        fr_mul(&v, &a[i1], &c->tw[j * c->step]);
        fr_add(&a[i0], &u, &v);
        fr_sub(&a[i1], &u, &v);
    }
}
static void ntt_inplace(fe* a, int logn, const fe* tw) {
    size_t n = (size_t)1 << logn;
    ntt_ctx c = {a, tw, logn, 0, 0};
    par_range(n, 4096, ntt_bitrev_range, &c);
    for (int s = 1; s <= logn; s++) {
        c.half = (size_t)1 << (s - 1);
        c.step = n >> s;
        par_range(n / 2, 4096, ntt_stage_range, &c);
    }
}

typedef struct { fe* tw; const fe* lead; size_t B; } tw_ctx;
static void tw_range(size_t lo, size_t hi, void* p) {
    tw_ctx* c = (tw_ctx*)p;
    for (size_t b = lo > 1 ? lo : 1; b < hi; b++)
        for (size_t i = 0; i < c->B; i++) fr_mul(&c->tw[b * c->B + i], &c->lead[b], &c->tw[i]);
}
static void make_twiddles(fe* tw, size_t half, const fe* w) {
    if (half == 0) return;
    fr_one(&tw[0]);
    /* blocked so that it parallelises: tw[b*B + i] = w^(b*B) * w^i */
    size_t B = 1024;
    if (half <= B) {
        for (size_t i = 1; i < half; i++) fr_mul(&tw[i], &tw[i - 1], w);
        return;
    }
    for (size_t i = 1; i < B; i++) fr_mul(&tw[i], &tw[i - 1], w);
    fe wB;
    fr_mul(&wB, &tw[B - 1], w);
    size_t nb = half / B;
    fe* lead = (fe*)malloc(nb * sizeof(fe));
    fr_one(&lead[0]);
    for (size_t b = 1; b < nb; b++) fr_mul(&lead[b], &lead[b - 1], &wB);
    tw_ctx c = {tw, lead, B};
    par_range(nb, 1, tw_range, &c);
    free(lead);
}

typedef struct {
    int logn;
    fe *tw, *twi; /* forward and inverse twiddles */
    fe inc, ninv;
} ntt_domain;

static void domain_init(ntt_domain* d, int logn) {
    size_t n = (size_t)1 << logn, half = n / 2 ? n / 2 : 1;
    d->logn = logn;
    d->tw = (fe*)malloc(half * sizeof(fe));
    d->twi = (fe*)malloc(half * sizeof(fe));
    fe w, wi;
    fr_root_of_unity(&w, logn);
    fr_inv(&wi, &w);
    make_twiddles(d->tw, n / 2 ? n / 2 : 1, &w);
    make_twiddles(d->twi, n / 2 ? n / 2 : 1, &wi);
    if (logn < 28) fr_root_of_unity(&d->inc, logn + 1);
    else { fe s = {{25, 0, 0, 0}}; f_to_mont(&d->inc, &s, &FR); } /* Fr.shift = nqr^2 */
    fe nn = {{(uint64_t)n, 0, 0, 0}};
    f_to_mont(&nn, &nn, &FR);
    fr_inv(&d->ninv, &nn);
}
static void domain_free(ntt_domain* d) { free(d->tw); free(d->twi); }

/* evaluations on the domain -> evaluations on the odd coset (rows a5, a6, a7), Montgomery in and out */
typedef struct { fe* a; const ntt_domain* d; } coset_ctx;
static void coset_scale_range(size_t lo, size_t hi, void* p) { /* a[i] *= inc^i / n   (Fr.batchApplyKey) */
    coset_ctx* c = (coset_ctx*)p;
    uint64_t e[4] = {lo, 0, 0, 0};
    fe cur;
    f_pow(&cur, &c->d->inc, e, &FR);
    fr_mul(&cur, &cur, &c->d->ninv);
    for (size_t i = lo; i < hi; i++) {
        fr_mul(&c->a[i], &c->a[i], &cur);
        fr_mul(&cur, &cur, &c->d->inc);
    }
}
static void coset_transform(fe* a, const ntt_domain* d) {
    size_t n = (size_t)1 << d->logn;
    ntt_inplace(a, d->logn, d->twi);
    coset_ctx c = {a, d};
    par_range(n, 4096, coset_scale_range, &c);
    ntt_inplace(a, d->logn, d->tw);
}

/* ---- binfile (row a1) ----------------------------------------------------------------------------------------- */
typedef struct { const uint8_t* p; uint64_t size; } section;
static int parse_binfile(const uint8_t* data, size_t len, const char* magic, uint32_t max_version, section secs[16]) {
    memset(secs, 0, 16 * sizeof(section));
    if (len < 12 || memcmp(data, magic, 4) != 0) return -1;
    uint32_t ver, nsec;
    memcpy(&ver, data + 4, 4);
    memcpy(&nsec, data + 8, 4);
    if (ver > max_version) return -1;
    size_t pos = 12;
    for (uint32_t i = 0; i < nsec; i++) {
        if (pos + 12 > len) return -1;
        uint32_t typ;
        uint64_t size;
        memcpy(&typ, data + pos, 4);
        memcpy(&size, data + pos + 4, 8);
        pos += 12;
        if (size > len - pos) return -1;
        if (typ < 16 && !secs[typ].p) { secs[typ].p = data + pos; secs[typ].size = size; }
        pos += size;
    }
    return 0;
}

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static double g_times[8]; /* [0] abc [1] ntt+join [2] msm A [3] B1 [4] B2 [5] C [6] H [7] total */

typedef struct { const uint8_t *recs, *w; fe *prod, *A, *B, *C, *out; uint32_t n, n_vars; volatile int bad; } abc_ctx;
static void abc_prod_range(size_t lo, size_t hi, void* p) {
    abc_ctx* x = (abc_ctx*)p;
    for (size_t i = lo; i < hi; i++) {
        const uint8_t* r = x->recs + i * 44;
        uint32_t m, c, s;
        memcpy(&m, r, 4); memcpy(&c, r + 4, 4); memcpy(&s, r + 8, 4);
        if (m > 1 || c >= x->n || s >= x->n_vars) { x->bad = 1; continue; }
        fe coef, ws;
        memcpy(coef.v, r + 12, 32);
        memcpy(ws.v, x->w + (size_t)s * 32, 32);
        fr_mul(&x->prod[i], &coef, &ws);
    }
}
static void abc_c_range(size_t lo, size_t hi, void* p) {
    abc_ctx* x = (abc_ctx*)p;
    for (size_t i = lo; i < hi; i++) fr_mul(&x->C[i], &x->A[i], &x->B[i]);
}
static void abc_join_range(size_t lo, size_t hi, void* p) { /* joinABC + batchFromMontgomery */
    abc_ctx* x = (abc_ctx*)p;
    for (size_t i = lo; i < hi; i++) {
        fe t;
        fr_mul(&t, &x->A[i], &x->B[i]);
        fr_sub(&t, &t, &x->C[i]);
        f_from_mont(&x->out[i], &t, &FR);
    }
}

/* rows a4-a8: H scalars (standard form) into out[n] */
static int h_scalars(const section* zs, const uint8_t* w, uint32_t n_vars, uint32_t n, fe* out) {
    int logn = 0;
    while ((1u << logn) < n) logn++;
    uint32_t ncoef;
    memcpy(&ncoef, zs[4].p, 4);
    if (zs[4].size != 4 + (uint64_t)ncoef * 44) return -1;
    const uint8_t* recs = zs[4].p + 4;
    double t0 = now_s();
    fe* A = (fe*)calloc(n, sizeof(fe));
    fe* B = (fe*)calloc(n, sizeof(fe));
    fe* C = (fe*)malloc((size_t)n * sizeof(fe));
    fe* prod = (fe*)malloc((size_t)(ncoef ? ncoef : 1) * sizeof(fe));
    /* buildABC1: X_m[c] += coef (x) w[s]; the products in parallel, the scatter-add in file order */
    abc_ctx ac = {recs, w, prod, A, B, C, out, n, n_vars, 0};
    par_range(ncoef, 4096, abc_prod_range, &ac);
    if (ac.bad) { free(A); free(B); free(C); free(prod); return -1; }
    for (uint32_t i = 0; i < ncoef; i++) {
        const uint8_t* r = recs + (size_t)i * 44;
        uint32_t m, c;
        memcpy(&m, r, 4); memcpy(&c, r + 4, 4);
        fe* X = m ? B : A;
        fr_add(&X[c], &X[c], &prod[i]);
    }
    free(prod);
    par_range(n, 4096, abc_c_range, &ac);
    g_times[0] = now_s() - t0;
    t0 = now_s();
    ntt_domain d;
    domain_init(&d, logn);
    coset_transform(A, &d);
    coset_transform(B, &d);
    coset_transform(C, &d);
    domain_free(&d);
    par_range(n, 4096, abc_join_range, &ac);
    g_times[1] = now_s() - t0;
    free(A); free(B); free(C);
    return 0;
}

static void dec_g1(g1_aff* a, const uint8_t* p) {
    memcpy(a->x.v, p, 32);
    memcpy(a->y.v, p + 32, 32);
    a->inf = fq_is_zero(&a->x) && fq_is_zero(&a->y);
}
static void dec_g2(g2_aff* a, const uint8_t* p) {
    memcpy(&a->x, p, 64);
    memcpy(&a->y, p + 64, 64);
    a->inf = fq2_is_zero(&a->x) && fq2_is_zero(&a->y);
}

/* ---- exported API ------------------------------------------------------------------------------------------------- */
#define EXPORT __attribute__((visibility("default")))

EXPORT int orc_threads(void) { return n_threads(); }
EXPORT void orc_set_threads(int t) { if (t > 0) g_threads = t > 256 ? 256 : t; }
EXPORT void orc_last_timings(double out[8]) { memcpy(out, g_times, sizeof(g_times)); }

/* field op on n elements (parity of the restatement itself against oracle/bn254.py) */
EXPORT void orc_field_op(int field, int op, const uint8_t* a, const uint8_t* b, uint8_t* out, size_t n) {
    const fparams* P = field ? &FR : &FQ;
    for (size_t i = 0; i < n; i++) {
        fe x, y, r;
        memcpy(x.v, a + 32 * i, 32);
        if (b) memcpy(y.v, b + 32 * i, 32);
        switch (op) {
            case 0: f_mul(&r, &x, &y, P); break;
            case 1: f_add(&r, &x, &y, P); break;
            case 2: f_sub(&r, &x, &y, P); break;
            case 3: f_to_mont(&r, &x, P); break;
            case 4: f_from_mont(&r, &x, P); break;
            default: f_inv(&r, &x, P); break;
        }
        memcpy(out + 32 * i, r.v, 32);
    }
}

/* MSM over zkey records; out = affine Montgomery record (all-zero = infinity) */
EXPORT int orc_msm_g1(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t out[64]) {
    g1_jac r;
    g1_aff a;
    g1_msm(&r, bases, scalars, n);
    g1_to_affine(&a, &r);
    memset(out, 0, 64);
    if (!a.inf) { memcpy(out, a.x.v, 32); memcpy(out + 32, a.y.v, 32); }
    return 0;
}
EXPORT int orc_msm_g2(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t out[128]) {
    g2_jac r;
    g2_aff a;
    g2_msm(&r, bases, scalars, n);
    g2_to_affine(&a, &r);
    memset(out, 0, 128);
    if (!a.inf) { memcpy(out, &a.x, 64); memcpy(out + 64, &a.y, 64); }
    return 0;
}

/* odd-coset transform of 2^logn Montgomery Fr elements, in place */
EXPORT int orc_coset_ntt(uint8_t* data, int logn) {
    ntt_domain d;
    domain_init(&d, logn);
    coset_transform((fe*)data, &d);
    domain_free(&d);
    return 0;
}

/* rows a4-a8 from file bytes; out = domain_size x 32 B standard form */
EXPORT int orc_h_scalars(const uint8_t* zkey, size_t zlen, const uint8_t* wtns, size_t wlen, uint8_t* out) {
    section zs[16], ws[16];
    if (parse_binfile(zkey, zlen, "zkey", 1, zs) || parse_binfile(wtns, wlen, "wtns", 2, ws)) return 6;
    if (!zs[2].p || !zs[4].p || !ws[2].p) return 6;
    uint32_t nv, n;
    memcpy(&nv, zs[2].p + 72, 4);
    memcpy(&n, zs[2].p + 80, 4);
    if (ws[2].size != (uint64_t)nv * 32) return 3;
    return h_scalars(zs, ws[2].p, nv, n, (fe*)out) ? 6 : 0;
}

/* The whole prove step from file bytes (what `prover zkey wtns proof public` computes).  r, s: 32 B LE standard form.
 * out: pi_a x,y | pi_b x.c0,x.c1,y.c0,y.c1 | pi_c x,y   (8 x 32 B LE standard form; infinity = zeros).
 * Returns 0, 3 (witness length), 4 (curve), 6 (format) -- the codes of include/g16_b200.h. */
EXPORT int orc_prove(const uint8_t* zkey, size_t zlen, const uint8_t* wtns, size_t wlen, const uint8_t r_[32],
                     const uint8_t s_[32], uint8_t out[256]) {
    double t_all = now_s();
    section zs[16], ws[16];
    if (parse_binfile(zkey, zlen, "zkey", 1, zs) || parse_binfile(wtns, wlen, "wtns", 2, ws)) return 6;
    for (int i = 1; i <= 9; i++) if (!zs[i].p) return 6;
    if (!ws[1].p || !ws[2].p) return 6;
    uint32_t proto;
    memcpy(&proto, zs[1].p, 4);
    if (proto != 1) return 6;
    const uint8_t* h = zs[2].p;
    uint32_t n8q, n8r, nv, npub, n;
    memcpy(&n8q, h, 4);
    memcpy(&n8r, h + 36, 4);
    if (n8q != 32 || n8r != 32 || memcmp(h + 4, FQ.p, 32) || memcmp(h + 40, FR.p, 32)) return 4;
    memcpy(&nv, h + 72, 4); memcpy(&npub, h + 76, 4); memcpy(&n, h + 80, 4);
    h += 84;
    uint32_t n8w, nw;
    memcpy(&n8w, ws[1].p, 4);
    if (n8w != 32 || memcmp(ws[1].p + 4, FR.p, 32)) return 4;
    memcpy(&nw, ws[1].p + 36, 4);
    if (nw != nv) return 3;
    if (ws[2].size != (uint64_t)nv * 32) return 6;
    if (zs[5].size != (uint64_t)nv * 64 || zs[6].size != (uint64_t)nv * 64 || zs[7].size != (uint64_t)nv * 128 ||
        zs[8].size != (uint64_t)(nv - npub - 1) * 64 || zs[9].size != (uint64_t)n * 64) return 6;
    g1_aff alpha1, beta1, delta1;
    g2_aff beta2, delta2;
    dec_g1(&alpha1, h); dec_g1(&beta1, h + 64); dec_g2(&beta2, h + 128);
    dec_g1(&delta1, h + 384); dec_g2(&delta2, h + 448);
    const uint8_t* w = ws[2].p;

    fe* P = (fe*)malloc((size_t)n * sizeof(fe));
    if (h_scalars(zs, w, nv, n, P)) { free(P); return 6; }
    g1_jac a, b1, c, hh;
    g2_jac b2;
    double t0 = now_s();
    g1_msm(&a, zs[5].p, w, nv); g_times[2] = now_s() - t0; t0 = now_s();
    g1_msm(&b1, zs[6].p, w, nv); g_times[3] = now_s() - t0; t0 = now_s();
    g2_msm(&b2, zs[7].p, w, nv); g_times[4] = now_s() - t0; t0 = now_s();
    g1_msm(&c, zs[8].p, w + (size_t)(npub + 1) * 32, nv - npub - 1); g_times[5] = now_s() - t0; t0 = now_s();
    g1_msm(&hh, zs[9].p, (const uint8_t*)P, n); g_times[6] = now_s() - t0;
    free(P);

    /* row a15 */
    uint64_t r[4], s[4];
    memcpy(r, r_, 32);
    memcpy(s, s_, 32);
    while (ge(r, FR.p)) sub_raw(r, r, FR.p);
    while (ge(s, FR.p)) sub_raw(s, s, FR.p);
    fe rm, sm, rs;
    memcpy(rm.v, r, 32); memcpy(sm.v, s, 32);
    f_to_mont(&rm, &rm, &FR); f_to_mont(&sm, &sm, &FR);
    fr_mul(&rs, &rm, &sm);
    f_from_mont(&rs, &rs, &FR);
    g1_jac d1, t, pi_a, pb1, pi_c;
    g2_jac d2, t2, pi_b;
    g1_from_affine(&d1, &delta1);
    g2_from_affine(&d2, &delta2);
    g1_from_affine(&t, &alpha1); g1_add(&pi_a, &a, &t);
    g1_mul_scalar(&t, &d1, r); g1_add(&pi_a, &pi_a, &t);
    g2_from_affine(&t2, &beta2); g2_add(&pi_b, &b2, &t2);
    g2_mul_scalar(&t2, &d2, s); g2_add(&pi_b, &pi_b, &t2);
    g1_from_affine(&t, &beta1); g1_add(&pb1, &b1, &t);
    g1_mul_scalar(&t, &d1, s); g1_add(&pb1, &pb1, &t);
    g1_add(&pi_c, &c, &hh);
    g1_mul_scalar(&t, &pi_a, s); g1_add(&pi_c, &pi_c, &t);
    g1_mul_scalar(&t, &pb1, r); g1_add(&pi_c, &pi_c, &t);
    g1_mul_scalar(&t, &d1, rs.v); g1_neg(&t, &t); g1_add(&pi_c, &pi_c, &t);

    g1_aff A, C;
    g2_aff B;
    g1_to_affine(&A, &pi_a); g2_to_affine(&B, &pi_b); g1_to_affine(&C, &pi_c);
    memset(out, 0, 256);
    fe o;
    if (!A.inf) { f_from_mont(&o, &A.x, &FQ); memcpy(out, o.v, 32); f_from_mont(&o, &A.y, &FQ); memcpy(out + 32, o.v, 32); }
    if (!B.inf) {
        f_from_mont(&o, &B.x.c0, &FQ); memcpy(out + 64, o.v, 32);
        f_from_mont(&o, &B.x.c1, &FQ); memcpy(out + 96, o.v, 32);
        f_from_mont(&o, &B.y.c0, &FQ); memcpy(out + 128, o.v, 32);
        f_from_mont(&o, &B.y.c1, &FQ); memcpy(out + 160, o.v, 32);
    }
    if (!C.inf) { f_from_mont(&o, &C.x, &FQ); memcpy(out + 192, o.v, 32); f_from_mont(&o, &C.y, &FQ); memcpy(out + 224, o.v, 32); }
    g_times[7] = now_s() - t_all;
    return 0;
}

/* ---- SYNTH-1 fixture generator (orc_synth_sizes, orc_synth_write) ------------------------------------------------ */
#include "synth1.h"
