/* Jacobian short-Weierstrass (a = 0) group law over a field given by macros; included twice (Fq -> G1, Fq2 -> G2).
 * ORACLE CODE - test infrastructure.  Restates wasmcurves 0.2.2 build_curve_jacobian_a0.
 * Required macros: FE (element type), FN(name) (field function prefix), CN(name) (curve symbol prefix). */

typedef struct { FE x, y; int inf; } CN(aff);
typedef struct { FE x, y, z; } CN(jac); /* z == 0 -> identity */

static inline void CN(set_inf)(CN(jac) * p) { memset(p, 0, sizeof(*p)); }
static inline int CN(is_inf)(const CN(jac) * p) { return FN(is_zero)(&p->z); }

static void CN(dbl)(CN(jac) * r, const CN(jac) * p) {
    if (CN(is_inf)(p)) { *r = *p; return; }
    FE a, b, c, d, e, f, t;
    FN(sqr)(&a, &p->x);
    FN(sqr)(&b, &p->y);
    FN(sqr)(&c, &b);
    FN(add)(&t, &p->x, &b); FN(sqr)(&t, &t); FN(sub)(&t, &t, &a); FN(sub)(&t, &t, &c); FN(add)(&d, &t, &t);
    FN(add)(&e, &a, &a); FN(add)(&e, &e, &a);
    FN(sqr)(&f, &e);
    FE x3, y3, z3;
    FN(sub)(&x3, &f, &d); FN(sub)(&x3, &x3, &d);
    FN(mul)(&z3, &p->y, &p->z); FN(add)(&z3, &z3, &z3);
    FN(sub)(&t, &d, &x3); FN(mul)(&y3, &e, &t);
    FN(add)(&c, &c, &c); FN(add)(&c, &c, &c); FN(add)(&c, &c, &c);
    FN(sub)(&y3, &y3, &c);
    r->x = x3; r->y = y3; r->z = z3;
}

/* r = p + q (q affine, not infinity) */
static void CN(madd)(CN(jac) * r, const CN(jac) * p, const FE* qx, const FE* qy) {
    if (CN(is_inf)(p)) { r->x = *qx; r->y = *qy; FN(one)(&r->z); return; }
    FE z1z1, u2, s2, h, hh, i, j, rr, v, t;
    FN(sqr)(&z1z1, &p->z);
    FN(mul)(&u2, qx, &z1z1);
    FN(mul)(&s2, qy, &p->z); FN(mul)(&s2, &s2, &z1z1);
    FN(sub)(&h, &u2, &p->x);
    FN(sub)(&rr, &s2, &p->y);
    if (FN(is_zero)(&h)) {
        if (FN(is_zero)(&rr)) { CN(jac) q; q.x = *qx; q.y = *qy; FN(one)(&q.z); CN(dbl)(r, &q); }
        else CN(set_inf)(r);
        return;
    }
    FN(sqr)(&hh, &h);
    FN(add)(&i, &hh, &hh); FN(add)(&i, &i, &i);
    FN(mul)(&j, &h, &i);
    FN(add)(&rr, &rr, &rr);
    FN(mul)(&v, &p->x, &i);
    FE x3, y3, z3;
    FN(sqr)(&x3, &rr); FN(sub)(&x3, &x3, &j); FN(sub)(&x3, &x3, &v); FN(sub)(&x3, &x3, &v);
    FN(sub)(&t, &v, &x3); FN(mul)(&y3, &rr, &t);
    FN(mul)(&t, &p->y, &j); FN(add)(&t, &t, &t); FN(sub)(&y3, &y3, &t);
    FN(add)(&z3, &p->z, &h); FN(sqr)(&z3, &z3); FN(sub)(&z3, &z3, &z1z1); FN(sub)(&z3, &z3, &hh);
    r->x = x3; r->y = y3; r->z = z3;
}

static void CN(add)(CN(jac) * r, const CN(jac) * p, const CN(jac) * q) {
    if (CN(is_inf)(q)) { *r = *p; return; }
    if (CN(is_inf)(p)) { *r = *q; return; }
    FE z1z1, z2z2, u1, u2, s1, s2, h, i, j, rr, v, t;
    FN(sqr)(&z1z1, &p->z);
    FN(sqr)(&z2z2, &q->z);
    FN(mul)(&u1, &p->x, &z2z2);
    FN(mul)(&u2, &q->x, &z1z1);
    FN(mul)(&s1, &p->y, &q->z); FN(mul)(&s1, &s1, &z2z2);
    FN(mul)(&s2, &q->y, &p->z); FN(mul)(&s2, &s2, &z1z1);
    FN(sub)(&h, &u2, &u1);
    FN(sub)(&rr, &s2, &s1);
    if (FN(is_zero)(&h)) {
        if (FN(is_zero)(&rr)) CN(dbl)(r, p);
        else CN(set_inf)(r);
        return;
    }
    FN(add)(&i, &h, &h); FN(sqr)(&i, &i);
    FN(mul)(&j, &h, &i);
    FN(add)(&rr, &rr, &rr);
    FN(mul)(&v, &u1, &i);
    FE x3, y3, z3;
    FN(sqr)(&x3, &rr); FN(sub)(&x3, &x3, &j); FN(sub)(&x3, &x3, &v); FN(sub)(&x3, &x3, &v);
    FN(sub)(&t, &v, &x3); FN(mul)(&y3, &rr, &t);
    FN(mul)(&t, &s1, &j); FN(add)(&t, &t, &t); FN(sub)(&y3, &y3, &t);
    FN(add)(&z3, &p->z, &q->z); FN(sqr)(&z3, &z3); FN(sub)(&z3, &z3, &z1z1); FN(sub)(&z3, &z3, &z2z2);
    FN(mul)(&z3, &z3, &h);
    r->x = x3; r->y = y3; r->z = z3;
}

static void CN(neg)(CN(jac) * r, const CN(jac) * p) { *r = *p; FN(neg)(&r->y, &p->y); }

static void CN(to_affine)(CN(aff) * r, const CN(jac) * p) {
    if (CN(is_inf)(p)) { memset(r, 0, sizeof(*r)); r->inf = 1; return; }
    FE zi, zi2, zi3;
    FN(inv)(&zi, &p->z);
    FN(sqr)(&zi2, &zi);
    FN(mul)(&zi3, &zi2, &zi);
    FN(mul)(&r->x, &p->x, &zi2);
    FN(mul)(&r->y, &p->y, &zi3);
    r->inf = 0;
}

static void CN(from_affine)(CN(jac) * r, const CN(aff) * a) {
    if (a->inf) { CN(set_inf)(r); return; }
    r->x = a->x; r->y = a->y; FN(one)(&r->z);
}

/* k * p, k = 4 little-endian 64-bit limbs (standard form) */
static void CN(mul_scalar)(CN(jac) * r, const CN(jac) * p, const uint64_t k[4]) {
    CN(jac) acc;
    CN(set_inf)(&acc);
    for (int i = 255; i >= 0; i--) {
        CN(dbl)(&acc, &acc);
        if ((k[i >> 6] >> (i & 63)) & 1) CN(add)(&acc, &acc, p);
    }
    *r = acc;
}

static inline uint32_t CN(window)(const uint64_t* s, int bit, int c) {
    int limb = bit >> 6, sh = bit & 63;
    if (limb >= 4) return 0;
    uint64_t v = s[limb] >> sh;
    if (sh + c > 64 && limb + 1 < 4) v |= s[limb + 1] << (64 - sh);
    return (uint32_t)(v & ((1u << c) - 1));
}

/* Pippenger over one contiguous chunk, single thread (rapidsnark multiexp.cpp / ffjavascript engine_multiexp.js
 * compute the same sum with their own chunking; the sum does not depend on it).  bases: records of REC bytes
 * (affine, Montgomery, all-zero = infinity); scalars: 32 B LE standard form. */
static void CN(pippenger)(CN(jac) * out, const uint8_t* bases, const uint8_t* scalars, size_t n) {
    CN(set_inf)(out);
    if (n == 0) return;
    int c = 1;
    {
        double best = 1e300;
        for (int k = 1; k <= 18; k++) {
            double cost = ((double)n + 2.0 * (double)(1u << k)) * (double)((254 + k - 1) / k);
            if (cost < best) { best = cost; c = k; }
        }
    }
    int W = (254 + c - 1) / c;
    size_t nb = (size_t)1 << c;
    CN(jac)* bucket = (CN(jac)*)malloc(nb * sizeof(CN(jac)));
    CN(jac) total;
    CN(set_inf)(&total);
    for (int w = W - 1; w >= 0; w--) {
        for (int i = 0; i < c; i++) CN(dbl)(&total, &total);
        memset(bucket, 0, nb * sizeof(CN(jac)));
        for (size_t i = 0; i < n; i++) {
            uint64_t s[4];
            memcpy(s, scalars + 32 * i, 32);
            uint32_t d = CN(window)(s, w * c, c);
            if (!d) continue;
            FE q[2]; /* records sit at arbitrary byte offsets of the file: copy, do not cast */
            memcpy(q, bases + sizeof(FE) * 2 * i, sizeof(q));
            if (FN(is_zero)(&q[0]) && FN(is_zero)(&q[1])) continue;
            CN(madd)(&bucket[d], &bucket[d], &q[0], &q[1]);
        }
        CN(jac) run, acc;
        CN(set_inf)(&run);
        CN(set_inf)(&acc);
        for (size_t b = nb - 1; b >= 1; b--) {
            CN(add)(&run, &run, &bucket[b]);
            CN(add)(&acc, &acc, &run);
        }
        CN(add)(&total, &total, &acc);
    }
    free(bucket);
    *out = total;
}

/* all cores: one Pippenger per thread over a contiguous slice, then the partial sums are added */
typedef struct { const uint8_t *bases, *scalars; size_t n; int T; CN(jac) * part; } CN(msm_ctx);
static void CN(msm_range)(size_t lo, size_t hi, void* p) {
    CN(msm_ctx)* c = (CN(msm_ctx)*)p;
    for (size_t t = lo; t < hi; t++) {
        size_t a = c->n * t / (size_t)c->T, b = c->n * (t + 1) / (size_t)c->T;
        CN(pippenger)(&c->part[t], c->bases + a * 2 * sizeof(FE), c->scalars + a * 32, b - a);
    }
}
static void CN(msm)(CN(jac) * out, const uint8_t* bases, const uint8_t* scalars, size_t n) {
    int T = n_threads();
    if ((size_t)T > n / 256 + 1) T = (int)(n / 256 + 1);
    CN(jac)* part = (CN(jac)*)malloc((size_t)T * sizeof(CN(jac)));
    CN(msm_ctx) c = {bases, scalars, n, T, part};
    par_range((size_t)T, 0, CN(msm_range), &c);
    CN(set_inf)(out);
    for (int t = 0; t < T; t++) CN(add)(out, out, &part[t]);
    free(part);
}

/* ---- fixed-base scalar multiplication k*G (used by the SYNTH-1 fixture generator in synth1.h only) ----------------
 * FB_W windows of FB_C bits; table[w][d-1] = d * 2^(FB_C w) * G in affine form, so k*G is at most FB_W mixed
 * additions.  Results are brought back to affine in batches with one inversion each (Montgomery's trick). */
#ifndef FB_C
#define FB_C 13
#define FB_W 20
#define FB_ROW ((1u << FB_C) - 1)
#endif
typedef struct { FE x, y; } CN(xy);

/* out[i] = affine form of in[i] (all-zero for the identity); scratch holds n elements */
static void CN(batch_to_affine)(CN(xy) * out, const CN(jac) * in, size_t n, FE* scratch) {
    FE acc, inv;
    FN(one)(&acc);
    for (size_t i = 0; i < n; i++) {
        if (CN(is_inf)(&in[i])) continue;
        scratch[i] = acc;
        FN(mul)(&acc, &acc, &in[i].z);
    }
    FN(inv)(&inv, &acc);
    for (size_t i = n; i-- > 0;) {
        if (CN(is_inf)(&in[i])) { memset(&out[i], 0, sizeof(out[i])); continue; }
        FE zi, zi2, zi3;
        FN(mul)(&zi, &inv, &scratch[i]);
        FN(mul)(&inv, &inv, &in[i].z);
        FN(sqr)(&zi2, &zi);
        FN(mul)(&zi3, &zi2, &zi);
        FN(mul)(&out[i].x, &in[i].x, &zi2);
        FN(mul)(&out[i].y, &in[i].y, &zi3);
    }
}

typedef struct { CN(xy) * table; CN(aff) gen; } CN(fb_ctx);
static void CN(fb_table_range)(size_t lo, size_t hi, void* p) {
    CN(fb_ctx)* c = (CN(fb_ctx)*)p;
    CN(jac)* row = (CN(jac)*)malloc(FB_ROW * sizeof(CN(jac)));
    FE* scratch = (FE*)malloc(FB_ROW * sizeof(FE));
    for (size_t w = lo; w < hi; w++) {
        CN(jac) base;
        CN(from_affine)(&base, &c->gen);
        for (size_t i = 0; i < FB_C * w; i++) CN(dbl)(&base, &base);
        row[0] = base;
        for (size_t d = 1; d < FB_ROW; d++) CN(add)(&row[d], &row[d - 1], &base);
        CN(batch_to_affine)(c->table + w * FB_ROW, row, FB_ROW, scratch);
    }
    free(row);
    free(scratch);
}
static CN(xy) * CN(fb_table)(const CN(aff) * gen) {
    CN(fb_ctx) c;
    c.table = (CN(xy)*)malloc((size_t)FB_W * FB_ROW * sizeof(CN(xy)));
    c.gen = *gen;
    par_range(FB_W, 0, CN(fb_table_range), &c);
    return c.table;
}

/* out[i] = scalars[i] * G as a zkey record (affine, Montgomery, all-zero = infinity); scalars: 32 B LE standard form */
typedef struct { const CN(xy) * table; const uint8_t* scalars; uint8_t* out; } CN(fbm_ctx);
static void CN(fb_mul_range)(size_t lo, size_t hi, void* p) {
    CN(fbm_ctx)* c = (CN(fbm_ctx)*)p;
    enum { BATCH = 1024 };
    CN(jac)* acc = (CN(jac)*)malloc(BATCH * sizeof(CN(jac)));
    CN(xy)* aff = (CN(xy)*)malloc(BATCH * sizeof(CN(xy)));
    FE* scratch = (FE*)malloc(BATCH * sizeof(FE));
    for (size_t b = lo; b < hi; b += BATCH) {
        size_t m = hi - b < BATCH ? hi - b : BATCH;
        for (size_t i = 0; i < m; i++) {
            uint64_t s[4];
            memcpy(s, c->scalars + 32 * (b + i), 32);
            CN(set_inf)(&acc[i]);
            for (int w = 0; w < FB_W; w++) {
                uint32_t d = CN(window)(s, w * FB_C, FB_C);
                if (!d) continue;
                const CN(xy)* q = c->table + (size_t)w * FB_ROW + (d - 1);
                CN(madd)(&acc[i], &acc[i], &q->x, &q->y);
            }
        }
        CN(batch_to_affine)(aff, acc, m, scratch);
        memcpy(c->out + sizeof(CN(xy)) * b, aff, m * sizeof(CN(xy)));
    }
    free(acc);
    free(aff);
    free(scratch);
}
static void CN(fb_mul)(const CN(xy) * table, const uint8_t* scalars, size_t n, uint8_t* out) {
    CN(fbm_ctx) c = {table, scalars, out};
    par_range(n, 256, CN(fb_mul_range), &c);
}
