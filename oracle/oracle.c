/*
 * oracle.c — CPU restatement of the reference's in-tree exact-NTT TFHE PBS.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Each function cites the reference
 * lines it restates; paths are relative to
 * /root/reference/research/verifiable-fhe/src .
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;
#define Q ORC_Q
#define EPS 0xFFFFFFFFULL /* 2^64 mod q */

/* ------------------------------------------------------------------ field */
/* Goldilocks field, the reference's F = GoldilocksField (plonky2 0.2.0, Cargo.toml:7). */

/* branch-free: the conditions are data-dependent coin flips, a branch would mispredict */
static inline uint64_t gl_add(uint64_t a, uint64_t b) {
    uint64_t s = a + b;
    uint64_t c = (uint64_t)(s < a) | (uint64_t)(s >= Q);
    return s - (Q & (0 - c));
}
static inline uint64_t gl_sub(uint64_t a, uint64_t b) {
    uint64_t d = a - b;
    return d + (Q & (0 - (uint64_t)(a < b)));
}
static inline uint64_t gl_neg(uint64_t a) { return a ? Q - a : 0; }

static inline uint64_t reduce128(u128 x) {
    uint64_t lo = (uint64_t)x, hi = (uint64_t)(x >> 64);
    uint64_t hh = hi >> 32, hl = hi & EPS;
    uint64_t t0 = lo - hh;
    t0 -= EPS & (0 - (uint64_t)(lo < hh)); /* borrowed 2^64 == EPS */
    uint64_t t1 = hl * EPS;
    uint64_t s = t0 + t1;
    s += EPS & (0 - (uint64_t)(s < t0));
    s -= Q & (0 - (uint64_t)(s >= Q));
    return s;
}
static inline uint64_t gl_mul(uint64_t a, uint64_t b) { return reduce128((u128)a * b); }

/* exported (non-inlinable under -fPIC) entry points for the tests */
uint64_t orc_add(uint64_t a, uint64_t b) { return gl_add(a, b); }
uint64_t orc_sub(uint64_t a, uint64_t b) { return gl_sub(a, b); }
uint64_t orc_neg(uint64_t a) { return gl_neg(a); }
uint64_t orc_mul(uint64_t a, uint64_t b) { return gl_mul(a, b); }
uint64_t orc_mul_slow(uint64_t a, uint64_t b) { return (uint64_t)(((u128)a * b) % Q); }

uint64_t orc_pow(uint64_t a, uint64_t e) {
    uint64_t r = 1;
    while (e) {
        if (e & 1) r = gl_mul(r, a);
        a = gl_mul(a, a);
        e >>= 1;
    }
    return r;
}
uint64_t orc_inv(uint64_t a) { return orc_pow(a, Q - 2); }

static uint32_t ilog2(uint32_t x) {
    uint32_t l = 0;
    while ((1u << l) < x) l++;
    return l;
}
static uint32_t bitrev(uint32_t x, uint32_t bits) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}

/* -------------------------------------------------------------------- NTT */
/* ntt/gen_param_file.sage:3-10 (psi = g^((q-1)/2N), g = 7 the smallest primitive root),
 * :86-88,111-112 (tables stored bit-reversed), NINV = (1/2)^logN. */
int orc_ntt_tables(uint32_t N, uint64_t *roots, uint64_t *invroots, uint64_t *ninv) {
    if (N < 2 || (N & (N - 1))) return -1;
    uint32_t logN = ilog2(N);
    uint64_t psi = orc_pow(7, (Q - 1) / (2ULL * N));
    uint64_t psi_inv = orc_inv(psi);
    uint64_t *pw = (uint64_t *)malloc(sizeof(uint64_t) * N * 2);
    if (!pw) return -2;
    uint64_t *pwi = pw + N;
    pw[0] = pwi[0] = 1;
    for (uint32_t j = 1; j < N; j++) {
        pw[j] = gl_mul(pw[j - 1], psi);
        pwi[j] = gl_mul(pwi[j - 1], psi_inv);
    }
    for (uint32_t j = 0; j < N; j++) {
        roots[j] = pw[bitrev(j, logN)];
        invroots[j] = pwi[bitrev(j, logN)];
    }
    free(pw);
    *ninv = orc_pow(orc_inv(2), logN);
    return 0;
}

/* crypto/poly.rs:9-34: Cooley-Tukey, natural in -> bit-reversed out, twiddle ROOTS[m+i]. */
void orc_ntt_forward(uint32_t N, const uint64_t *roots, uint64_t *a) {
    for (uint32_t m = 1; m < N; m <<= 1) {
        uint32_t t = N / (2 * m);
        for (uint32_t i = 0; i < m; i++) {
            uint32_t j1 = 2 * i * t, j2 = j1 + t;
            uint64_t s = roots[m + i];
            for (uint32_t j = j1; j < j2; j++) {
                uint64_t u = a[j];
                uint64_t v = gl_mul(a[j + t], s);
                a[j] = gl_add(u, v);
                a[j + t] = gl_sub(u, v);
            }
        }
    }
}

/* crypto/poly.rs:36-64: Gentleman-Sande, bit-reversed in -> natural out, then * NINV. */
void orc_ntt_backward(uint32_t N, const uint64_t *invroots, uint64_t ninv, uint64_t *a) {
    for (uint32_t m = N / 2; m >= 1; m >>= 1) {
        uint32_t t = N / (2 * m);
        uint32_t j1 = 0;
        for (uint32_t i = 0; i < m; i++) {
            uint32_t j2 = j1 + t;
            uint64_t s = invroots[m + i];
            for (uint32_t j = j1; j < j2; j++) {
                uint64_t u = a[j], v = a[j + t];
                a[j] = gl_add(u, v);
                a[j + t] = gl_mul(gl_sub(u, v), s);
            }
            j1 += 2 * t;
        }
    }
    for (uint32_t j = 0; j < N; j++) a[j] = gl_mul(a[j], ninv);
}

/* ----------------------------------------------------------------- pieces */

uint32_t orc_num_limbs(uint32_t logB) { return (64 + logB - 1) / logB; } /* ceil_div_usize(F::BITS, LOGB), glev.rs:28 */

/* vtfhe/glwe_poly.rs:28-50.  bits = split_le(x, L*logB); sgn = top bit of that split
 * (bit 63 when L*logB == 64, a constant 0 when L*logB > 64 since x < 2^64);
 * x' = sgn ? -x : x; limb k_i = bits [i*logB, (i+1)*logB) of x';
 * digit_i = k_i + carry_{i-1} - MSB(k_i)*B with carry_i = MSB(k_i) (of the RAW limb);
 * final digit negated when sgn. */
void orc_decompose_signed(uint64_t x, uint32_t logB, int64_t *digits) {
    uint32_t L = orc_num_limbs(logB);
    uint32_t total = L * logB;
    int sgn = (total == 64) ? (int)(x >> 63) : 0;
    uint64_t xc = sgn ? (Q - x) : x; /* sgn => x >= 2^63 > 0 */
    int64_t carry = 0;
    for (uint32_t i = 0; i < L; i++) {
        uint32_t sh = i * logB;
        uint64_t kk = sh >= 64 ? 0 : (xc >> sh);
        if (logB < 64) kk &= ((1ULL << logB) - 1);
        int64_t msb = (int64_t)((kk >> (logB - 1)) & 1);
        int64_t d = (int64_t)kk + carry - (int64_t)((uint64_t)msb << logB);
        carry = msb;
        digits[i] = sgn ? -d : d;
    }
}

void orc_decompose(uint64_t x, uint32_t logB, uint64_t *digits) {
    int64_t d[64];
    uint32_t L = orc_num_limbs(logB);
    orc_decompose_signed(x, logB, d);
    for (uint32_t i = 0; i < L; i++) digits[i] = d[i] >= 0 ? (uint64_t)d[i] : Q - (uint64_t)(-d[i]);
}

/* crypto/lwe.rs:28-34 (called with p = N, ivc_based_vpbs.rs:277) and its circuit twin
 * vtfhe/mod.rs:80-107: keep the top log2(N)+2 bits of the canonical u64, round on the last. */
uint32_t orc_mod_switch(uint64_t x, uint32_t N) {
    uint64_t s = x >> (64 - ilog2(N) - 2);
    uint64_t carry = s & 1;
    s >>= 1;
    return (uint32_t)(s + carry);
}

/* multiply by X^shift in Z_q[X]/(X^N+1): vtfhe/glwe_poly.rs:127-142 composed as in
 * vtfhe/mod.rs:80-107 (rounding shift by 1, then the bits up to weight N);
 * plain twin crypto/poly.rs:152-184 (right_shift).  shift in [0, 2N]. */
void orc_rotate(uint32_t N, const uint64_t *in, uint32_t shift, uint64_t *out) {
    int negate = 0;
    shift %= 2 * N;
    if (shift >= N) {
        negate = 1;
        shift -= N;
    }
    for (uint32_t i = 0; i < N; i++) {
        uint64_t v = (i < shift) ? gl_neg(in[N - shift + i]) : in[i - shift];
        out[i] = negate ? gl_neg(v) : v;
    }
}

uint64_t orc_delta(uint32_t plaintext_space) { return Q >> ilog2(plaintext_space); } /* lwe.rs:46-48 */

/* crypto/mod.rs:17-27: p blocks of N/p copies of lut[i]*delta, left-shifted by half a block
 * (left_shift = multiply by X^-s, crypto/poly.rs:152-166). */
void orc_testv_lut(uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out) {
    uint32_t bs = N / p, sh = bs / 2;
    uint64_t *tmp = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (uint32_t i = 0; i < N; i++) tmp[i] = gl_mul(lut[i / bs] % Q, delta);
    for (uint32_t i = 0; i < N; i++) out[i] = (i < N - sh) ? tmp[i + sh] : gl_neg(tmp[i - N + sh]);
    free(tmp);
}
void orc_testv(uint32_t N, uint32_t p, uint64_t delta, uint64_t *out) {
    uint64_t *lut = (uint64_t *)malloc(sizeof(uint64_t) * p);
    for (uint32_t i = 0; i < p; i++) lut[i] = i;
    orc_testv_lut(N, p, delta, lut, out);
    free(lut);
}

/* crypto/glwe.rs:96-106: per mask poly a_0, -a_{N-1}, ..., -a_1 ; body coeff 0 last. */
void orc_sample_extract(uint32_t N, uint32_t k, const uint64_t *glwe, uint64_t *lwe) {
    for (uint32_t j = 0; j < k; j++) {
        const uint64_t *a = glwe + (size_t)j * N;
        lwe[(size_t)j * N] = a[0];
        for (uint32_t i = 1; i < N; i++) lwe[(size_t)j * N + i] = gl_neg(a[N - i]);
    }
    lwe[(size_t)k * N] = glwe[(size_t)k * N];
}
/* crypto/glwe.rs:108-113 */
void orc_partial_sample_extract(uint32_t N, uint32_t k, uint32_t nz, const uint64_t *glwe, uint64_t *lwe) {
    uint64_t *full = (uint64_t *)malloc(sizeof(uint64_t) * ((size_t)k * N + 1));
    orc_sample_extract(N, k, glwe, full);
    memcpy(lwe, full, sizeof(uint64_t) * nz);
    lwe[nz] = full[(size_t)k * N];
    free(full);
}

/* ---------------------------------------------------------------- context */
struct orc_ctx {
    orc_params p;
    uint32_t L, L_ks; /* limb counts */
    uint64_t *roots, *invroots, ninv;
    uint64_t *bsk_ntt;      /* [n] GGSW, NTT domain */
    uint64_t *glwe_ksk_ntt; /* one GGSW, NTT domain */
    uint64_t *lwe_ksk;      /* [kN][ell_ks][n+1] */
};

static size_t ggsw_len(const orc_params *p) { return (size_t)(p->k + 1) * p->ell * (p->k + 1) * p->N; }

orc_ctx *orc_ctx_create(const orc_params *p) {
    if (!p || p->N < 2 || (p->N & (p->N - 1)) || p->k < 1 || p->ell < 1 || p->logB < 1 || p->logB > 62) return NULL;
    if (p->ell > orc_num_limbs(p->logB)) return NULL;
    orc_ctx *c = (orc_ctx *)calloc(1, sizeof(orc_ctx));
    c->p = *p;
    c->L = orc_num_limbs(p->logB);
    c->L_ks = p->logB_ks ? orc_num_limbs(p->logB_ks) : 0;
    c->roots = (uint64_t *)malloc(sizeof(uint64_t) * p->N);
    c->invroots = (uint64_t *)malloc(sizeof(uint64_t) * p->N);
    if (orc_ntt_tables(p->N, c->roots, c->invroots, &c->ninv)) {
        orc_ctx_destroy(c);
        return NULL;
    }
    return c;
}
void orc_ctx_destroy(orc_ctx *c) {
    if (!c) return;
    free(c->roots);
    free(c->invroots);
    free(c->bsk_ntt);
    free(c->glwe_ksk_ntt);
    free(c->lwe_ksk);
    free(c);
}
const uint64_t *orc_ctx_roots(const orc_ctx *c) { return c->roots; }
const uint64_t *orc_ctx_invroots(const orc_ctx *c) { return c->invroots; }
uint64_t orc_ctx_ninv(const orc_ctx *c) { return c->ninv; }
const uint64_t *orc_ctx_bsk_ntt(const orc_ctx *c) { return c->bsk_ntt; }

static uint64_t *copy_maybe_ntt(const orc_ctx *c, const uint64_t *src, size_t n_polys, int ntt_domain) {
    uint32_t N = c->p.N;
    uint64_t *dst = (uint64_t *)malloc(sizeof(uint64_t) * n_polys * N);
    if (!dst) return NULL;
    memcpy(dst, src, sizeof(uint64_t) * n_polys * N);
    if (!ntt_domain) {
#pragma omp parallel for schedule(static)
        for (long i = 0; i < (long)n_polys; i++) orc_ntt_forward(N, c->roots, dst + (size_t)i * N);
    }
    return dst;
}
/* crypto/mod.rs:41-44: every polynomial of every GGSW forward-NTT'd. */
int orc_ctx_set_bsk(orc_ctx *c, const uint64_t *bsk, int ntt_domain) {
    free(c->bsk_ntt);
    c->bsk_ntt = copy_maybe_ntt(c, bsk, (size_t)c->p.n * ggsw_len(&c->p) / c->p.N, ntt_domain);
    return c->bsk_ntt ? 0 : -1;
}
int orc_ctx_set_glwe_ksk(orc_ctx *c, const uint64_t *ksk, int ntt_domain) {
    free(c->glwe_ksk_ntt);
    c->glwe_ksk_ntt = copy_maybe_ntt(c, ksk, ggsw_len(&c->p) / c->p.N, ntt_domain);
    return c->glwe_ksk_ntt ? 0 : -1;
}
int orc_ctx_set_lwe_ksk(orc_ctx *c, const uint64_t *ksk) {
    size_t len = (size_t)c->p.k * c->p.N * c->p.ell_ks * (c->p.n + 1);
    free(c->lwe_ksk);
    c->lwe_ksk = (uint64_t *)malloc(sizeof(uint64_t) * len);
    if (!c->lwe_ksk) return -1;
    memcpy(c->lwe_ksk, ksk, sizeof(uint64_t) * len);
    return 0;
}

/* --------------------------------------------------------------- hot path */

/* vtfhe/glev_ct.rs:92-110: decompose every coefficient, keep limbs L-ell..L-1, forward-NTT
 * each limb polynomial, out[r] += limb_hat[i] (.) glev[i][r].  `out` is accumulated into
 * (NTT domain) with sign `negate` so that ggsw_ct.rs:108-111 (last - sum of others) needs
 * no second buffer. */
static void glev_mul_acc(const orc_ctx *c, const uint64_t *glev_ntt, const uint64_t *poly, int negate, uint64_t *out,
                         uint64_t *scratch /* [ell*N] */) {
    const uint32_t N = c->p.N, K = c->p.k + 1, ell = c->p.ell, L = c->L, logB = c->p.logB;
    uint64_t digits[64];
    for (uint32_t x = 0; x < N; x++) {
        orc_decompose(poly[x], logB, digits);
        for (uint32_t i = 0; i < ell; i++) scratch[(size_t)i * N + x] = digits[L - ell + i];
    }
    for (uint32_t i = 0; i < ell; i++) {
        uint64_t *lh = scratch + (size_t)i * N;
        orc_ntt_forward(N, c->roots, lh);
        for (uint32_t r = 0; r < K; r++) {
            const uint64_t *row = glev_ntt + ((size_t)i * K + r) * N;
            uint64_t *o = out + (size_t)r * N;
            if (negate)
                for (uint32_t x = 0; x < N; x++) o[x] = gl_sub(o[x], gl_mul(lh[x], row[x]));
            else
                for (uint32_t x = 0; x < N; x++) o[x] = gl_add(o[x], gl_mul(lh[x], row[x]));
        }
    }
}

/* vtfhe/ggsw_ct.rs:98-112: glev_muls[j] = GLev_j.mul(glwe.polys[j]);
 * result = iNTT(glev_muls[K-1] - sum_{j<K-1} glev_muls[j]). */
void orc_external_product(const orc_ctx *c, const uint64_t *ggsw_ntt, const uint64_t *glwe_in, uint64_t *glwe_out) {
    const uint32_t N = c->p.N, K = c->p.k + 1, ell = c->p.ell;
    const size_t glev = (size_t)ell * K * N;
    uint64_t *acc = (uint64_t *)calloc((size_t)K * N + (size_t)ell * N, sizeof(uint64_t));
    uint64_t *scratch = acc + (size_t)K * N;
    for (uint32_t j = 0; j < K; j++)
        glev_mul_acc(c, ggsw_ntt + j * glev, glwe_in + (size_t)j * N, j != K - 1, acc, scratch);
    for (uint32_t r = 0; r < K; r++) {
        orc_ntt_backward(N, c->invroots, c->ninv, acc + (size_t)r * N);
        memcpy(glwe_out + (size_t)r * N, acc + (size_t)r * N, sizeof(uint64_t) * N);
    }
    free(acc);
}

/* vtfhe/mod.rs:119-136: shifted = rotate(glwe, mask); diff = shifted - glwe;
 * out = GGSW (x) diff + glwe.  The rotation amount is the mod-switched mask element
 * (rotate_poly reads the top bits of the canonical element, mod.rs:80-107). */
void orc_cmux_step(const orc_ctx *c, const uint64_t *ggsw_ntt, uint64_t mask_element, uint64_t *acc) {
    const uint32_t N = c->p.N, K = c->p.k + 1;
    uint32_t sh = orc_mod_switch(mask_element, N);
    uint64_t *diff = (uint64_t *)malloc(sizeof(uint64_t) * 2 * K * N);
    uint64_t *xp = diff + (size_t)K * N;
    for (uint32_t j = 0; j < K; j++) {
        orc_rotate(N, acc + (size_t)j * N, sh, diff + (size_t)j * N);
        for (uint32_t x = 0; x < N; x++) diff[(size_t)j * N + x] = gl_sub(diff[(size_t)j * N + x], acc[(size_t)j * N + x]);
    }
    orc_external_product(c, ggsw_ntt, diff, xp);
    for (size_t x = 0; x < (size_t)K * N; x++) acc[x] = gl_add(acc[x], xp[x]);
    free(diff);
}

/* ivc_based_vpbs.rs:109-124 (first_step: rotate the trivial (0,testv) by the NEGATED body,
 * :110-112) then :323-353 (n CMUX steps with bsk[x], ct[x]). */
void orc_blind_rotate_steps(const orc_ctx *c, const uint64_t *lwe, const uint64_t *testv, uint32_t steps,
                            uint64_t *acc) {
    const uint32_t N = c->p.N, K = c->p.k + 1, n = c->p.n;
    memset(acc, 0, sizeof(uint64_t) * (size_t)K * N);
    orc_rotate(N, testv, orc_mod_switch(gl_neg(lwe[n]), N), acc + (size_t)(K - 1) * N);
    if (steps > n) steps = n;
    for (uint32_t i = 0; i < steps; i++) orc_cmux_step(c, c->bsk_ntt + (size_t)i * ggsw_len(&c->p), lwe[i], acc);
}
void orc_blind_rotate(const orc_ctx *c, const uint64_t *lwe, const uint64_t *testv, uint64_t *acc) {
    orc_blind_rotate_steps(c, lwe, testv, c->p.n, acc);
}

/* vtfhe/mod.rs:31-54 with lev_ct.rs:63-76: out = (0,..,0,b) + sum_i sum_j digit_j(a_i) * KSK[i][j];
 * KSK rows encrypt -s_i (mod.rs:52).  The engine keeps the top ell_ks limbs, the same
 * convention as glev_ct.rs:97-99; when ell_ks == ceil(64/logB_ks) this is lev_ct.rs verbatim. */
void orc_lwe_keyswitch(const orc_ctx *c, const uint64_t *lwe_big, uint64_t *lwe_small) {
    const uint32_t n = c->p.n, kN = c->p.k * c->p.N, ell = c->p.ell_ks, L = c->L_ks, logB = c->p.logB_ks;
    uint64_t digits[64];
    memset(lwe_small, 0, sizeof(uint64_t) * (n + 1));
    lwe_small[n] = lwe_big[kN];
    for (uint32_t i = 0; i < kN; i++) {
        orc_decompose(lwe_big[i], logB, digits);
        for (uint32_t j = 0; j < ell; j++) {
            uint64_t d = digits[L - ell + j];
            if (!d) continue;
            const uint64_t *row = c->lwe_ksk + ((size_t)i * ell + j) * (n + 1);
            for (uint32_t x = 0; x <= n; x++) lwe_small[x] = gl_add(lwe_small[x], gl_mul(d, row[x]));
        }
    }
}

size_t orc_lwe_in_len(const orc_ctx *c) {
    return c->p.order == ORC_ORDER_LWEKS_PBS ? (size_t)c->p.k * c->p.N + 1 : (size_t)c->p.n + 1;
}

int orc_pbs(const orc_ctx *c, const uint64_t *lwe_in, const uint64_t *testv, uint64_t *lwe_out) {
    const uint32_t N = c->p.N, K = c->p.k + 1, n = c->p.n;
    if (!c->bsk_ntt) return -1;
    uint64_t *acc = (uint64_t *)malloc(sizeof(uint64_t) * (2 * (size_t)K * N + n + 1));
    uint64_t *acc2 = acc + (size_t)K * N;
    uint64_t *small = acc2 + (size_t)K * N;
    int rc = 0;
    if (c->p.order == ORC_ORDER_PBS_GLWEKS) {
        /* ivc_based_vpbs.rs:355-371: last IVC step = KSK (x) acc (no CMUX), then the LWE is read
         * off with the partial key's n coefficients (main.rs:59, glwe.rs:108-113). */
        if (!c->glwe_ksk_ntt) {
            rc = -2;
        } else {
            orc_blind_rotate(c, lwe_in, testv, acc);
            orc_external_product(c, c->glwe_ksk_ntt, acc, acc2);
            orc_partial_sample_extract(N, c->p.k, n, acc2, lwe_out);
        }
    } else {
        if (!c->lwe_ksk) {
            rc = -2;
        } else {
            orc_lwe_keyswitch(c, lwe_in, small);
            orc_blind_rotate(c, small, testv, acc);
            orc_sample_extract(N, c->p.k, acc, lwe_out);
        }
    }
    free(acc);
    return rc;
}

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int orc_pbs_batch(const orc_ctx *c, size_t B, const uint64_t *lwe_in, const uint64_t *testvs, const uint32_t *lut_idx,
                  uint64_t *lwe_out) {
    const size_t len = orc_lwe_in_len(c);
    int bad = 0;
#pragma omp parallel for schedule(dynamic, 1)
    for (long b = 0; b < (long)B; b++) {
        const uint64_t *tv = testvs + (size_t)(lut_idx ? lut_idx[b] : 0) * c->p.N;
        if (orc_pbs(c, lwe_in + (size_t)b * len, tv, lwe_out + (size_t)b * len)) bad = 1;
    }
    return bad ? -1 : orc_max_threads();
}

/* ------------------------------------------------- fixtures: rng and keys */
/* The reference draws from rand::thread_rng (unseeded); fixtures need determinism, so the
 * oracle uses xoshiro256** seeded through splitmix64.  Distributions follow the reference. */
static uint64_t splitmix64(uint64_t *x) {
    uint64_t z = (*x += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
void orc_rng_seed(orc_rng *r, uint64_t seed) {
    for (int i = 0; i < 4; i++) r->s[i] = splitmix64(&seed);
    r->have_spare = 0;
    r->spare = 0.0;
}
static inline uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
uint64_t orc_rng_u64(orc_rng *r) {
    uint64_t *s = r->s;
    uint64_t result = rotl64(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl64(s[3], 45);
    return result;
}
uint64_t orc_rng_field(orc_rng *r) { /* F::rand(): uniform in [0,q) */
    uint64_t v;
    do v = orc_rng_u64(r);
    while (v >= Q);
    return v;
}
static double rng_normal(orc_rng *r) {
    if (r->have_spare) {
        r->have_spare = 0;
        return r->spare;
    }
    double u1, u2;
    do u1 = (double)(orc_rng_u64(r) >> 11) * (1.0 / 9007199254740992.0);
    while (u1 <= 0.0);
    u2 = (double)(orc_rng_u64(r) >> 11) * (1.0 / 9007199254740992.0);
    double m = sqrt(-2.0 * log(u1)), a = 6.283185307179586476925 * u2;
    r->spare = m * sin(a);
    r->have_spare = 1;
    return m * cos(a);
}
/* crypto/lwe.rs:40-44: round(Normal(0, sigma*q)) as i64 -> field */
uint64_t orc_rng_error(orc_rng *r, double sigma) {
    if (sigma == 0.0) return 0;
    double e = nearbyint(rng_normal(r) * sigma * (double)Q);
    int64_t ei = (int64_t)e;
    return ei >= 0 ? (uint64_t)ei : Q - (uint64_t)(-ei);
}

void orc_lwe_keygen(orc_rng *r, uint32_t n, uint64_t *s) {
    for (uint32_t i = 0; i < n; i++) s[i] = orc_rng_u64(r) & 1;
}
void orc_glwe_keygen(orc_rng *r, uint32_t N, uint32_t k, uint64_t *s) {
    for (size_t i = 0; i < (size_t)k * N; i++) s[i] = orc_rng_u64(r) & 1;
}
/* crypto/glwe.rs:19-38: binary key whose flattened coefficients beyond nz are zero */
void orc_glwe_partial_key(orc_rng *r, uint32_t N, uint32_t k, uint32_t nz, uint64_t *s) {
    for (size_t i = 0; i < (size_t)k * N; i++) {
        uint64_t b = orc_rng_u64(r) & 1;
        s[i] = i < nz ? b : 0;
    }
}
void orc_lwe_encrypt(orc_rng *r, uint32_t n, const uint64_t *s, uint64_t m, double sigma, uint64_t *ct) {
    uint64_t body = 0;
    for (uint32_t i = 0; i < n; i++) {
        ct[i] = orc_rng_field(r);
        body = gl_add(body, gl_mul(s[i], ct[i]));
    }
    ct[n] = gl_add(gl_add(body, m), orc_rng_error(r, sigma));
}
uint64_t orc_lwe_decrypt(uint32_t n, const uint64_t *s, const uint64_t *ct) {
    uint64_t ip = 0;
    for (uint32_t i = 0; i < n; i++) ip = gl_add(ip, gl_mul(s[i], ct[i]));
    return gl_sub(ct[n], ip);
}
/* batched forms of the two functions above (test infrastructure for the >= 10^6-trial decryption gate,
 * SURVEY 8d): ciphertext b is lwe.rs:50-59 under its own generator seeded with seed + b */
void orc_lwe_encrypt_batch(uint64_t seed, uint32_t n, const uint64_t *s, size_t B, const uint64_t *msgs, double sigma,
                           uint64_t *cts) {
#pragma omp parallel for schedule(static)
    for (long b = 0; b < (long)B; b++) {
        orc_rng r;
        orc_rng_seed(&r, seed + (uint64_t)b);
        orc_lwe_encrypt(&r, n, s, msgs[b], sigma, cts + (size_t)b * (n + 1));
    }
}
void orc_lwe_decrypt_batch(uint32_t n, const uint64_t *s, size_t B, const uint64_t *cts, uint64_t *noisy) {
#pragma omp parallel for schedule(static)
    for (long b = 0; b < (long)B; b++) noisy[b] = orc_lwe_decrypt(n, s, cts + (size_t)b * (n + 1));
}

/* main.rs:61-64: round(noisy/delta) mod plaintext_space (integer rounding instead of f64) */
uint64_t orc_decode(uint64_t noisy, uint64_t delta, uint32_t plaintext_space) {
    u128 v = (u128)noisy + delta / 2;
    return (uint64_t)(v / delta) % plaintext_space;
}

static void poly_mul(const orc_ctx *c, const uint64_t *a, const uint64_t *b, uint64_t *out) {
    const uint32_t N = c->p.N;
    uint64_t *t = (uint64_t *)malloc(sizeof(uint64_t) * 2 * N);
    memcpy(t, a, sizeof(uint64_t) * N);
    memcpy(t + N, b, sizeof(uint64_t) * N);
    orc_ntt_forward(N, c->roots, t);
    orc_ntt_forward(N, c->roots, t + N);
    for (uint32_t i = 0; i < N; i++) out[i] = gl_mul(t[i], t[N + i]);
    orc_ntt_backward(N, c->invroots, c->ninv, out);
    free(t);
}
/* crypto/glwe.rs:49-57: body = sum a_j*s_j + e + m */
void orc_glwe_encrypt(const orc_ctx *c, orc_rng *r, const uint64_t *s, const uint64_t *m, double sigma, uint64_t *ct) {
    const uint32_t N = c->p.N, k = c->p.k;
    uint64_t *tmp = (uint64_t *)malloc(sizeof(uint64_t) * N);
    uint64_t *body = ct + (size_t)k * N;
    for (size_t i = 0; i < (size_t)k * N; i++) ct[i] = orc_rng_field(r);
    for (uint32_t i = 0; i < N; i++) body[i] = orc_rng_error(r, sigma);
    for (uint32_t j = 0; j < k; j++) {
        poly_mul(c, ct + (size_t)j * N, s + (size_t)j * N, tmp);
        for (uint32_t i = 0; i < N; i++) body[i] = gl_add(body[i], tmp[i]);
    }
    for (uint32_t i = 0; i < N; i++) body[i] = gl_add(body[i], m[i]);
    free(tmp);
}
void orc_glwe_decrypt(const orc_ctx *c, const uint64_t *s, const uint64_t *ct, uint64_t *m) {
    const uint32_t N = c->p.N, k = c->p.k;
    uint64_t *tmp = (uint64_t *)malloc(sizeof(uint64_t) * N);
    memcpy(m, ct + (size_t)k * N, sizeof(uint64_t) * N);
    for (uint32_t j = 0; j < k; j++) {
        poly_mul(c, ct + (size_t)j * N, s + (size_t)j * N, tmp);
        for (uint32_t i = 0; i < N; i++) m[i] = gl_sub(m[i], tmp[i]);
    }
    free(tmp);
}
/* crypto/glev.rs:26-38: level i encrypts m * B^(first+i), first = ceil(64/logB) - ell */
static void glev_encrypt(const orc_ctx *c, orc_rng *r, const uint64_t *s, const uint64_t *m, double sigma, uint64_t *glev) {
    const uint32_t N = c->p.N, K = c->p.k + 1, ell = c->p.ell;
    uint64_t base = orc_pow(2, c->p.logB);
    uint32_t first = c->L - ell;
    uint64_t *sm = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (uint32_t i = 0; i < ell; i++) {
        uint64_t f = orc_pow(base, first + i);
        for (uint32_t x = 0; x < N; x++) sm[x] = gl_mul(m[x], f);
        orc_glwe_encrypt(c, r, s, sm, sigma, glev + (size_t)i * K * N);
    }
    free(sm);
}
/* crypto/ggsw.rs:26-36: GLev(m*s_i) for i<k, then GLev(m) */
void orc_ggsw_encrypt(const orc_ctx *c, orc_rng *r, const uint64_t *s, const uint64_t *m, double sigma, uint64_t *ggsw) {
    const uint32_t N = c->p.N, K = c->p.k + 1, ell = c->p.ell;
    const size_t glev = (size_t)ell * K * N;
    uint64_t *ms = (uint64_t *)malloc(sizeof(uint64_t) * N);
    for (uint32_t i = 0; i < K; i++) {
        if (i < K - 1) {
            poly_mul(c, m, s + (size_t)i * N, ms);
            glev_encrypt(c, r, s, ms, sigma, ggsw + i * glev);
        } else {
            glev_encrypt(c, r, s, m, sigma, ggsw + i * glev);
        }
    }
    free(ms);
}
/* crypto/mod.rs:29-45 (without the final ntt_forward: standard domain out) */
void orc_bsk_gen(const orc_ctx *c, orc_rng *r, const uint64_t *s_lwe, const uint64_t *s_glwe, double sigma, uint64_t *bsk) {
    const uint32_t N = c->p.N;
    uint64_t *m = (uint64_t *)calloc(N, sizeof(uint64_t));
    for (uint32_t i = 0; i < c->p.n; i++) {
        m[0] = s_lwe[i];
        orc_ggsw_encrypt(c, r, s_glwe, m, sigma, bsk + (size_t)i * ggsw_len(&c->p));
    }
    free(m);
}
/* crypto/ggsw.rs:38-48 (standard domain out): GLev_{s_to}(s_from[i]) for i<k, GLev_{s_to}(1) */
void orc_glwe_ksk_gen(const orc_ctx *c, orc_rng *r, const uint64_t *s_to, const uint64_t *s_from, double sigma,
                      uint64_t *ksk) {
    const uint32_t N = c->p.N, K = c->p.k + 1, ell = c->p.ell;
    const size_t glev = (size_t)ell * K * N;
    uint64_t *one = (uint64_t *)calloc(N, sizeof(uint64_t));
    one[0] = 1;
    for (uint32_t i = 0; i < K; i++)
        glev_encrypt(c, r, s_to, i < K - 1 ? s_from + (size_t)i * N : one, sigma, ksk + i * glev);
    free(one);
}
/* vtfhe/mod.rs:52 ("the KSK encrypts -s_i"); levels as glev.rs:26-38 with (logB_ks, ell_ks) */
void orc_lwe_ksk_gen(const orc_ctx *c, orc_rng *r, const uint64_t *s_big, const uint64_t *s_small, double sigma,
                     uint64_t *ksk) {
    const uint32_t n = c->p.n, kN = c->p.k * c->p.N, ell = c->p.ell_ks;
    uint64_t base = orc_pow(2, c->p.logB_ks);
    uint32_t first = c->L_ks - ell;
    for (uint32_t i = 0; i < kN; i++)
        for (uint32_t j = 0; j < ell; j++) {
            uint64_t m = gl_mul(gl_neg(s_big[i]), orc_pow(base, first + j));
            orc_lwe_encrypt(r, n, s_small, m, sigma, ksk + ((size_t)i * ell + j) * (n + 1));
        }
}

static uint64_t abs_diff_mod(uint64_t a, uint64_t b) {
    uint64_t d = a > b ? a - b : b - a;
    return d < Q - d ? d : Q - d;
}
/* crypto/glwe.rs:65-78 */
double orc_glwe_max_error(const orc_ctx *c, const uint64_t *s, const uint64_t *ct, const uint64_t *m) {
    const uint32_t N = c->p.N;
    uint64_t *mb = (uint64_t *)malloc(sizeof(uint64_t) * N), mx = 0;
    orc_glwe_decrypt(c, s, ct, mb);
    for (uint32_t i = 0; i < N; i++) {
        uint64_t d = abs_diff_mod(m[i], mb[i]);
        if (d > mx) mx = d;
    }
    free(mb);
    return (double)mx / (double)Q;
}
/* crypto/glwe.rs:80-94 */
double orc_glwe_avg_error(const orc_ctx *c, const uint64_t *s, const uint64_t *ct, const uint64_t *m) {
    const uint32_t N = c->p.N;
    uint64_t *mb = (uint64_t *)malloc(sizeof(uint64_t) * N);
    double sum = 0;
    orc_glwe_decrypt(c, s, ct, mb);
    for (uint32_t i = 0; i < N; i++) sum += (double)abs_diff_mod(m[i], mb[i]);
    free(mb);
    return sum / ((double)Q * (double)N);
}
double orc_centered_diff(uint64_t a, uint64_t b) {
    uint64_t d = gl_sub(a, b);
    if (d > Q / 2) return -(double)(Q - d) / (double)Q;
    return (double)d / (double)Q;
}
