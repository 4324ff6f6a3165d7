/*
 * torus.c — CPU restatement of the 2^64-torus / f64-FFT programmable bootstrap.  See torus.h for what it
 * restates, what pins it and what does not.  TEST INFRASTRUCTURE ONLY.
 *
 * Build: -ffp-contract=off is REQUIRED (oracle/Makefile): every floating-point operation written below must
 * be one individually rounded IEEE-754 binary64 operation, because the CUDA kernels are compared with this
 * file word for word.  fma() calls are the only fused operations.
 */
#include "torus.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

struct tor_ctx {
    orc_params p;
    uint32_t M;       /* N / 2 complex points */
    tor_cplx *tw;     /* [M] merged-twist twiddles, heap order */
    tor_cplx *bsk_f;  /* [n][2][ell][2][M] */
    uint64_t *ksk;    /* [kN][ell_ks][n+1] */
    int has_bsk, has_ksk;
};

/* ------------------------------------------------------------------------------------------ pieces */

/* tfhe-rs core_crypto/commons/math/decomposition/decomposer.rs: closest_representable — round x to the
 * nearest multiple of 2^(64 - ell*logB) (ties away from zero on the unsigned value, wrapping at 2^64) */
uint64_t tor_closest_representable(uint64_t x, uint32_t logB, uint32_t ell) {
    const uint32_t rep = logB * ell;
    if (rep >= 64) return x;
    const uint32_t non_rep = 64 - rep;
    const uint64_t rounding = (x >> (non_rep - 1)) & 1u;
    return ((x >> non_rep) + rounding) << non_rep;
}

/* tfhe-rs .../decomposition/iter.rs: SignedDecompositionIter + decompose_one_level — balanced digits in
 * [-B/2, B/2], least significant level first; sum_i digit_i * 2^(64 - (ell - i) logB) == closest_representable(x) */
void tor_decompose(uint64_t x, uint32_t logB, uint32_t ell, int64_t *digits) {
    const uint32_t rep = logB * ell;
    uint64_t state = rep >= 64 ? x : tor_closest_representable(x, logB, ell) >> (64 - rep);
    const uint64_t mask = (1ull << logB) - 1;
    for (uint32_t i = 0; i < ell; i++) {
        const uint64_t res = state & mask;
        state >>= logB;
        uint64_t carry = ((res - 1ull) | state) & res;
        carry >>= (logB - 1);
        state += carry;
        digits[i] = (int64_t)res - (int64_t)(carry << logB);
    }
}

/* crypto/lwe.rs:28-34 (x read as a 64-bit integer, i.e. x / 2^64 of the torus): round(x * 2N / 2^64) in [0, 2N] */
uint32_t tor_mod_switch(uint64_t x, uint32_t N) {
    uint32_t logN = 0;
    while ((1u << logN) < N) logN++;
    uint64_t a = x >> (64 - logN - 2);
    return (uint32_t)((a >> 1) + (a & 1));
}

/* multiply by X^shift in Z[X]/(X^N+1), shift in [0, 2N]   (vtfhe/glwe_poly.rs:127-142, crypto/poly.rs:152-184) */
void tor_rotate(uint32_t N, const uint64_t *in, uint32_t shift, uint64_t *out) {
    for (uint32_t i = 0; i < N; i++) {
        const uint32_t src = (i + 2 * N - (shift % (2 * N))) % (2 * N);
        out[i] = (src & N) ? (uint64_t)0 - in[src & (N - 1)] : in[src];
    }
}

uint64_t tor_delta(uint32_t space) {
    uint32_t l = 0;
    while ((1u << l) < space) l++;
    return 1ull << (64 - l);
}

/* get_testv (crypto/mod.rs:17-27): p blocks of N/p copies of lut[i]*delta, then times X^-(block/2) */
void tor_testv_lut(uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out) {
    const uint32_t bs = N / p, sh = bs / 2;
    for (uint32_t i = 0; i < N; i++) {
        const uint32_t src = i + sh;
        out[i] = src < N ? lut[src / bs] * delta : (uint64_t)0 - lut[(src - N) / bs] * delta;
    }
}

/* ------------------------------------------------------------------------------------------ FFT
 * A real polynomial p of degree < N is folded into M = N/2 complex coefficients z_j = p_j + i p_{j+M}; then
 * p(w^(4k+1)) = sum_j z_j w^((4k+1) j) with w = exp(i pi / N), because X^M = i at every such point.  Those M
 * values determine p (the other N/2 roots of X^N + 1 are their conjugates), and the negacyclic product is the
 * pointwise product.  Evaluating at all roots of Y^M = i is done like the reference's NTT evaluates at the roots
 * of X^N = -1 (crypto/poly.rs:9-34): a Cooley-Tukey tree in which node v splits Y^len - c into Y^(len/2) -+ sqrt(c),
 * twiddle sqrt(c), natural order in, bit-reversed order out, the twist merged into the twiddles.
 * With c = w^e: root e = N/2; node twiddle w^(e/2); left child e/2, right child e/2 + N (w^N = -1), hence
 * tw[2v+1] = i * tw[2v] exactly — defined that way here, so the kernel may load even nodes only. */
void tor_twiddles(uint32_t N, tor_cplx *tw) {
    const uint32_t M = N / 2;
    uint32_t *c = (uint32_t *)malloc(sizeof(uint32_t) * M);
    tw[0].re = tw[0].im = 0.0;
    c[1] = N / 2;
    for (uint32_t v = 1; v < M; v++) {
        const uint32_t e = c[v] / 2;
        if (v == 1 || (v & 1) == 0) {
            const double ang = M_PI * (double)e / (double)N;
            tw[v].re = cos(ang);
            tw[v].im = sin(ang);
        } else {
            tw[v].re = -tw[v - 1].im;
            tw[v].im = tw[v - 1].re;
        }
        if (2 * v < M) {
            c[2 * v] = e;
            c[2 * v + 1] = e + N;
        }
    }
    free(c);
}

/* (u, v) -> (u + w v, u - w v) as six fused multiply-adds (the second output is 2u - (u + w v)) */
static inline void ct_bfly(tor_cplx *u, tor_cplx *v, double wr, double wi) {
    const double tr = fma(-wi, v->im, u->re);
    const double nr = fma(wr, v->re, tr);
    const double ti = fma(wi, v->re, u->im);
    const double ni = fma(wr, v->im, ti);
    const double br = fma(2.0, u->re, -nr);
    const double bi = fma(2.0, u->im, -ni);
    u->re = nr;
    u->im = ni;
    v->re = br;
    v->im = bi;
}
/* (u, v) -> (u + v, (u - v) conj(w)): four adds, two multiplies, two fused multiply-adds */
static inline void gs_bfly(tor_cplx *u, tor_cplx *v, double wr, double wi) {
    const double sr = u->re + v->re, si = u->im + v->im;
    const double dr = u->re - v->re, di = u->im - v->im;
    const double m0 = di * wi, m1 = dr * wi;
    u->re = sr;
    u->im = si;
    v->re = fma(dr, wr, m0);
    v->im = fma(di, wr, -m1);
}

void tor_fft_forward(uint32_t N, const tor_cplx *tw, tor_cplx *z) {
    const uint32_t M = N / 2;
    for (uint32_t m = 1; m < M; m <<= 1) {
        const uint32_t t = M / (2 * m);
        for (uint32_t i = 0; i < m; i++) {
            const double wr = tw[m + i].re, wi = tw[m + i].im;
            for (uint32_t j = 2 * i * t; j < 2 * i * t + t; j++) ct_bfly(&z[j], &z[j + t], wr, wi);
        }
    }
}
void tor_fft_inverse(uint32_t N, const tor_cplx *tw, tor_cplx *z) {
    const uint32_t M = N / 2;
    for (uint32_t m = M / 2; m >= 1; m >>= 1) {
        const uint32_t t = M / (2 * m);
        for (uint32_t i = 0; i < m; i++) {
            const double wr = tw[m + i].re, wi = tw[m + i].im;
            for (uint32_t j = 2 * i * t; j < 2 * i * t + t; j++) gs_bfly(&z[j], &z[j + t], wr, wi);
        }
    }
}

/* fractional part of y to 51 bits, as a torus element.  1.5 * 2^52 puts the integer being rounded into the low
 * mantissa bits of the sum (valid for |y| < 2^51; the products of a CMUX stay below 2^34). */
uint64_t tor_from_double(double y) {
    const double MAGIC = 6755399441055744.0;
    const double t = y + MAGIC;
    const double r = t - MAGIC;
    const double f = y - r;
    const double h = fma(f, 2251799813685248.0, MAGIC);
    int64_t v;
    memcpy(&v, &h, 8);
    v -= 0x4338000000000000LL;
    return (uint64_t)v << 13;
}

/* ------------------------------------------------------------------------------------------ context */
static uint32_t ilog2u(uint32_t x) {
    uint32_t l = 0;
    while ((1u << l) < x) l++;
    return l;
}

tor_ctx *tor_ctx_create(const orc_params *p) {
    if (!p || p->k != 1 || p->N < 8 || (p->N & (p->N - 1)) || p->ell < 1 || p->logB < 1 || p->ell * p->logB > 64 ||
        p->ell_ks < 1 || p->logB_ks < 1 || p->ell_ks * p->logB_ks > 64 || p->n < 1 ||
        (p->order != TOR_ORDER_LWEKS_PBS && p->order != TOR_ORDER_PBS_LWEKS))
        return NULL;
    tor_ctx *c = (tor_ctx *)calloc(1, sizeof(tor_ctx));
    c->p = *p;
    c->M = p->N / 2;
    c->tw = (tor_cplx *)malloc(sizeof(tor_cplx) * c->M);
    tor_twiddles(p->N, c->tw);
    return c;
}
void tor_ctx_destroy(tor_ctx *c) {
    if (!c) return;
    free(c->tw);
    free(c->bsk_f);
    free(c->ksk);
    free(c);
}
const tor_cplx *tor_ctx_bsk_fourier(const tor_ctx *c) { return c->bsk_f; }
const tor_cplx *tor_ctx_twiddles(const tor_ctx *c) { return c->tw; }
size_t tor_ct_len(const tor_ctx *c) {
    return c->p.order == TOR_ORDER_LWEKS_PBS ? (size_t)c->p.k * c->p.N + 1 : (size_t)c->p.n + 1;
}

static size_t ggsw_polys(const tor_ctx *c) { return (size_t)2 * c->p.ell * 2; }

/* Key polynomial -> Fourier row.  Coefficients are read as signed 64-bit integers, converted to binary64
 * (round to nearest) and scaled by 2^-64 / M: torus units, and the 1/M of the unnormalised inverse transform.
 * The mask GLev (j < k) is negated so that the external product is one sum (ggsw_ct.rs:104-110 subtracts it). */
void tor_ggsw_to_fourier(const tor_ctx *c, const uint64_t *ggsw, tor_cplx *out) {
    const uint32_t N = c->p.N, M = c->M;
    const double scale = ldexp(1.0, -64 - (int)ilog2u(M));
    for (uint32_t j = 0; j < 2; j++)
        for (uint32_t lvl = 0; lvl < c->p.ell; lvl++)
            for (uint32_t r = 0; r < 2; r++) {
                const size_t pi = ((size_t)j * c->p.ell + lvl) * 2 + r;
                const uint64_t *src = ggsw + pi * N;
                tor_cplx *z = out + pi * M;
                for (uint32_t q = 0; q < M; q++) {
                    z[q].re = (double)(int64_t)src[q] * scale;
                    z[q].im = (double)(int64_t)src[q + M] * scale;
                }
                tor_fft_forward(N, c->tw, z);
                if (j < c->p.k)
                    for (uint32_t q = 0; q < M; q++) {
                        z[q].re = -z[q].re;
                        z[q].im = -z[q].im;
                    }
            }
}

int tor_ctx_set_bsk(tor_ctx *c, const uint64_t *bsk) {
    if (!c || !bsk) return -1;
    const size_t per = ggsw_polys(c);
    free(c->bsk_f);
    c->bsk_f = (tor_cplx *)malloc(sizeof(tor_cplx) * c->p.n * per * c->M);
    if (!c->bsk_f) return -2;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)c->p.n; i++)
        tor_ggsw_to_fourier(c, bsk + (size_t)i * per * c->p.N, c->bsk_f + (size_t)i * per * c->M);
    c->has_bsk = 1;
    return 0;
}
int tor_ctx_set_lwe_ksk(tor_ctx *c, const uint64_t *ksk) {
    if (!c || !ksk) return -1;
    const size_t words = (size_t)c->p.k * c->p.N * c->p.ell_ks * (c->p.n + 1);
    free(c->ksk);
    c->ksk = (uint64_t *)malloc(sizeof(uint64_t) * words);
    if (!c->ksk) return -2;
    memcpy(c->ksk, ksk, sizeof(uint64_t) * words);
    c->has_ksk = 1;
    return 0;
}

/* ------------------------------------------------------------------------------------------ hot path */

/* GGSW (Fourier rows) (x) GLWE -> GLWE        vtfhe/ggsw_ct.rs:98-112, glev_ct.rs:92-110 on the torus:
 * for input polynomial j and level lvl (both ascending): D = FFT(digit polynomial), out_r += D (.) K[j][lvl][r],
 * each complex multiply-add as four fused multiply-adds in the order re*re, -im*im, re*im, im*re, from +0.0 */
void tor_external_product(const tor_ctx *c, const tor_cplx *ggsw_f, const uint64_t *glwe_in, uint64_t *glwe_out) {
    const uint32_t N = c->p.N, M = c->M, ell = c->p.ell, logB = c->p.logB;
    tor_cplx *out = (tor_cplx *)calloc((size_t)2 * M, sizeof(tor_cplx));
    tor_cplx *z = (tor_cplx *)malloc(sizeof(tor_cplx) * M);
    int64_t *dig = (int64_t *)malloc(sizeof(int64_t) * (size_t)N * ell);
    for (uint32_t j = 0; j < 2; j++) {
        for (uint32_t q = 0; q < N; q++) tor_decompose(glwe_in[(size_t)j * N + q], logB, ell, dig + (size_t)q * ell);
        for (uint32_t lvl = 0; lvl < ell; lvl++) {
            for (uint32_t q = 0; q < M; q++) {
                z[q].re = (double)dig[(size_t)q * ell + lvl];
                z[q].im = (double)dig[(size_t)(q + M) * ell + lvl];
            }
            tor_fft_forward(N, c->tw, z);
            for (uint32_t r = 0; r < 2; r++) {
                const tor_cplx *K = ggsw_f + (((size_t)j * ell + lvl) * 2 + r) * M;
                tor_cplx *o = out + (size_t)r * M;
                for (uint32_t q = 0; q < M; q++) {
                    double re = o[q].re, im = o[q].im;
                    re = fma(z[q].re, K[q].re, re);
                    re = fma(-z[q].im, K[q].im, re);
                    im = fma(z[q].re, K[q].im, im);
                    im = fma(z[q].im, K[q].re, im);
                    o[q].re = re;
                    o[q].im = im;
                }
            }
        }
    }
    for (uint32_t r = 0; r < 2; r++) {
        tor_cplx *o = out + (size_t)r * M;
        tor_fft_inverse(N, c->tw, o);
        for (uint32_t q = 0; q < M; q++) {
            glwe_out[(size_t)r * N + q] = tor_from_double(o[q].re);
            glwe_out[(size_t)r * N + q + M] = tor_from_double(o[q].im);
        }
    }
    free(dig);
    free(z);
    free(out);
}

/* acc <- acc + GGSW (x) (X^ms(mask) acc - acc)            vtfhe/mod.rs:119-136 */
void tor_cmux_step(const tor_ctx *c, const tor_cplx *ggsw_f, uint64_t mask_element, uint64_t *acc) {
    const uint32_t N = c->p.N;
    const uint32_t sh = tor_mod_switch(mask_element, N);
    uint64_t *d = (uint64_t *)malloc(sizeof(uint64_t) * 2 * N), *e = (uint64_t *)malloc(sizeof(uint64_t) * 2 * N);
    for (uint32_t j = 0; j < 2; j++) {
        tor_rotate(N, acc + (size_t)j * N, sh, d + (size_t)j * N);
        for (uint32_t q = 0; q < N; q++) d[(size_t)j * N + q] -= acc[(size_t)j * N + q];
    }
    tor_external_product(c, ggsw_f, d, e);
    for (uint32_t q = 0; q < 2 * N; q++) acc[q] += e[q];
    free(d);
    free(e);
}

/* acc0 = X^ms(-b) (0, testv); `steps` CMUX iterations      ivc_based_vpbs.rs:109-124,277-353 */
void tor_blind_rotate_steps(const tor_ctx *c, const uint64_t *lwe, const uint64_t *testv, uint32_t steps, uint64_t *acc) {
    const uint32_t N = c->p.N, n = c->p.n;
    memset(acc, 0, sizeof(uint64_t) * N);
    tor_rotate(N, testv, tor_mod_switch((uint64_t)0 - lwe[n], N), acc + N);
    const size_t per = ggsw_polys(c) * c->M;
    for (uint32_t i = 0; i < steps && i < n; i++) tor_cmux_step(c, c->bsk_f + (size_t)i * per, lwe[i], acc);
}

/* crypto/glwe.rs:96-106: a'_0 = a_0, a'_j = -a_{N-j}, body = b_0 */
void tor_sample_extract(uint32_t N, const uint64_t *glwe, uint64_t *lwe) {
    lwe[0] = glwe[0];
    for (uint32_t j = 1; j < N; j++) lwe[j] = (uint64_t)0 - glwe[N - j];
    lwe[N] = glwe[N];
}

/* vtfhe/mod.rs:31-54, lev_ct.rs:63-76: (0, .., 0, b) + sum_i sum_j digit_j(a_i) KSK[i][j] */
void tor_lwe_keyswitch(const tor_ctx *c, const uint64_t *big, uint64_t *small) {
    const uint32_t kN = c->p.k * c->p.N, n = c->p.n, ell = c->p.ell_ks, logB = c->p.logB_ks;
    memset(small, 0, sizeof(uint64_t) * n);
    small[n] = big[kN];
    int64_t dig[64];
    for (uint32_t i = 0; i < kN; i++) {
        tor_decompose(big[i], logB, ell, dig);
        for (uint32_t j = 0; j < ell; j++) {
            if (!dig[j]) continue;
            const uint64_t d = (uint64_t)dig[j];
            const uint64_t *row = c->ksk + ((size_t)i * ell + j) * (n + 1);
            for (uint32_t x = 0; x <= n; x++) small[x] += d * row[x];
        }
    }
}

int tor_pbs(const tor_ctx *c, const uint64_t *in, const uint64_t *testv, uint64_t *out) {
    if (!c->has_bsk || !c->has_ksk) return -1;
    const uint32_t N = c->p.N, n = c->p.n;
    uint64_t *acc = (uint64_t *)malloc(sizeof(uint64_t) * 2 * N);
    uint64_t *tmp = (uint64_t *)malloc(sizeof(uint64_t) * (N + 1 > n + 1 ? N + 1 : n + 1));
    if (c->p.order == TOR_ORDER_LWEKS_PBS) {
        tor_lwe_keyswitch(c, in, tmp);
        tor_blind_rotate_steps(c, tmp, testv, n, acc);
        tor_sample_extract(N, acc, out);
    } else {
        tor_blind_rotate_steps(c, in, testv, n, acc);
        tor_sample_extract(N, acc, tmp);
        tor_lwe_keyswitch(c, tmp, out);
    }
    free(tmp);
    free(acc);
    return 0;
}

int tor_pbs_batch(const tor_ctx *c, size_t B, const uint64_t *in, const uint64_t *testvs, const uint32_t *lut_idx, uint64_t *out) {
    if (!c->has_bsk || !c->has_ksk) return -1;
    const size_t w = tor_ct_len(c);
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t b = 0; b < (int64_t)B; b++)
        tor_pbs(c, in + (size_t)b * w, testvs + (size_t)(lut_idx ? lut_idx[b] : 0) * c->p.N, out + (size_t)b * w);
    return orc_max_threads();
}

/* ------------------------------------------------------------------------------------------ fixtures */
int64_t tor_rng_tuniform(orc_rng *r, uint32_t b) {
    const uint64_t u = orc_rng_u64(r) & ((1ull << (b + 2)) - 1);
    return (int64_t)((u + 1) >> 1) - ((int64_t)1 << b);
}
static uint64_t noise(orc_rng *r, int noise_log2) { return noise_log2 < 0 ? 0 : (uint64_t)tor_rng_tuniform(r, (uint32_t)noise_log2); }

void tor_lwe_encrypt(orc_rng *r, uint32_t n, const uint64_t *s, uint64_t m, int noise_log2, uint64_t *ct) {
    uint64_t b = m + noise(r, noise_log2);
    for (uint32_t i = 0; i < n; i++) {
        ct[i] = orc_rng_u64(r);
        if (s[i]) b += ct[i] * s[i];
    }
    ct[n] = b;
}
uint64_t tor_lwe_decrypt(uint32_t n, const uint64_t *s, const uint64_t *ct) {
    uint64_t b = ct[n];
    for (uint32_t i = 0; i < n; i++) b -= ct[i] * s[i];
    return b;
}
void tor_lwe_encrypt_batch(uint64_t seed, uint32_t n, const uint64_t *s, size_t B, const uint64_t *msgs, int noise_log2, uint64_t *cts) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < (int64_t)B; b++) {
        orc_rng r;
        orc_rng_seed(&r, seed * 0x9E3779B97F4A7C15ull + (uint64_t)b);
        tor_lwe_encrypt(&r, n, s, msgs[b], noise_log2, cts + (size_t)b * (n + 1));
    }
}
void tor_lwe_decrypt_batch(uint32_t n, const uint64_t *s, size_t B, const uint64_t *cts, uint64_t *noisy) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < (int64_t)B; b++) noisy[b] = tor_lwe_decrypt(n, s, cts + (size_t)b * (n + 1));
}
/* main.rs:61-64: round(noisy / delta) mod space */
uint64_t tor_decode(uint64_t noisy, uint64_t delta, uint32_t space) { return ((noisy + delta / 2) / delta) % space; }
double tor_centered_diff(uint64_t a, uint64_t b) { return (double)(int64_t)(a - b) * 0x1p-64; }

/* a(X) * s(X) mod X^N + 1 for a binary s: sum of the rotations X^i a over the set bits */
static void mul_binary(uint32_t N, const uint64_t *a, const uint64_t *s, uint64_t *out) {
    memset(out, 0, sizeof(uint64_t) * N);
    for (uint32_t i = 0; i < N; i++) {
        if (!s[i]) continue;
        for (uint32_t q = 0; q < N - i; q++) out[q + i] += a[q] * s[i];
        for (uint32_t q = N - i; q < N; q++) out[q + i - N] -= a[q] * s[i];
    }
}
void tor_glwe_decrypt(uint32_t N, const uint64_t *s, const uint64_t *ct, uint64_t *m) {
    mul_binary(N, ct, s, m);
    for (uint32_t q = 0; q < N; q++) m[q] = ct[N + q] - m[q];
}

/* GGSW of the scalar m (crypto/ggsw.rs:26-36, glev.rs:26-38 on the torus): GLev j < k encrypts m s_j, GLev k
 * encrypts m, level i scaled by 2^(64 - (ell - i) logB) */
void tor_ggsw_encrypt(const tor_ctx *c, orc_rng *r, const uint64_t *s, uint64_t m, int noise_log2, uint64_t *ggsw) {
    const uint32_t N = c->p.N, ell = c->p.ell, logB = c->p.logB;
    uint64_t *as = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (uint32_t j = 0; j < 2; j++)
        for (uint32_t lvl = 0; lvl < ell; lvl++) {
            uint64_t *a = ggsw + (((size_t)j * ell + lvl) * 2) * N, *b = a + N;
            const uint64_t scale = (uint64_t)1 << (64 - (ell - lvl) * logB);
            for (uint32_t q = 0; q < N; q++) a[q] = orc_rng_u64(r);
            mul_binary(N, a, s, as);
            for (uint32_t q = 0; q < N; q++) {
                uint64_t pt = j < c->p.k ? m * scale * s[q] : (q == 0 ? m * scale : 0);
                b[q] = as[q] + noise(r, noise_log2) + pt;
            }
        }
    free(as);
}
void tor_bsk_gen(const tor_ctx *c, orc_rng *r, const uint64_t *s_lwe, const uint64_t *s_glwe, int noise_log2, uint64_t *bsk) {
    const size_t per = ggsw_polys(c) * c->p.N;
    uint64_t *seeds = (uint64_t *)malloc(sizeof(uint64_t) * c->p.n);
    for (uint32_t i = 0; i < c->p.n; i++) seeds[i] = orc_rng_u64(r);
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t i = 0; i < (int64_t)c->p.n; i++) {
        orc_rng ri;
        orc_rng_seed(&ri, seeds[i]);
        tor_ggsw_encrypt(c, &ri, s_glwe, s_lwe[i], noise_log2, bsk + (size_t)i * per);
    }
    free(seeds);
}
/* row (i, j) = LWE_{s_small}( -s_big[i] * 2^(64 - (ell_ks - j) logB_ks) )          vtfhe/mod.rs:52 */
void tor_lwe_ksk_gen(const tor_ctx *c, orc_rng *r, const uint64_t *s_big, const uint64_t *s_small, int noise_log2, uint64_t *ksk) {
    const uint32_t kN = c->p.k * c->p.N, n = c->p.n, ell = c->p.ell_ks, logB = c->p.logB_ks;
    uint64_t *seeds = (uint64_t *)malloc(sizeof(uint64_t) * kN);
    for (uint32_t i = 0; i < kN; i++) seeds[i] = orc_rng_u64(r);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)kN; i++) {
        orc_rng ri;
        orc_rng_seed(&ri, seeds[i]);
        for (uint32_t j = 0; j < ell; j++) {
            const uint64_t scale = (uint64_t)1 << (64 - (ell - j) * logB);
            tor_lwe_encrypt(&ri, n, s_small, (uint64_t)0 - s_big[i] * scale, noise_log2, ksk + ((size_t)i * ell + j) * (n + 1));
        }
    }
    free(seeds);
}
