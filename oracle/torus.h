/*
 * torus.h — CPU restatement (plain C) of the programmable bootstrap on the 2^64 torus with a
 * double-precision negacyclic FFT ("arith = TORUS64_FFT").  TEST INFRASTRUCTURE ONLY, same rules as
 * oracle.h: nothing under tfhe_b200/ may include, link or call this.
 *
 * Why a second arithmetic.  The reference names its shortint parameter set
 * PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 (sdk/relayer/src/tfhe.ts:13-28,
 * relayer-sdk/scripts/generate-test-keys.ts:23-35) and takes it from tfhe-rs / node-tfhe 1.4.0-alpha.3
 * (sdk/relayer/package.json:69-72), a THIRD-PARTY dependency that is ABSENT from /root/reference.
 * That set is defined on the torus Z/2^64 with polynomial products by an f64 FFT.  This file restates
 * the published algorithm of that dependency (tfhe-rs core_crypto: SignedDecomposer, the "fold the real
 * polynomial of degree N into N/2 complex points and twist" negacyclic FFT, external product, CMUX blind
 * rotation, sample extraction, LWE key switch) on the structure of the in-tree PBS
 * (research/verifiable-fhe/src/vtfhe/ivc_based_vpbs.rs:109-124,277-371; layouts and sign conventions of
 * vtfhe/crypto/{ggsw,glev,glwe,lwe}.rs as in oracle.h), so that only the arithmetic differs between the
 * two oracles.
 *
 * PARITY STATUS: UNPINNED against tfhe-rs.  An f64 FFT is only defined up to rounding; two correct
 * implementations differ in the low ~26 bits of every accumulator word.  The north-star criterion for this
 * case is "decryptions bit-exact over >= 10^6 trials, output noise variance within the parameter set's
 * bound"; those are the tests.  What IS pinned bit-exactly is the CUDA engine against THIS restatement:
 * every floating-point operation below is an individually rounded IEEE-754 binary64 add / mul / fma in a
 * fixed order (compiled with -ffp-contract=off), and the CUDA kernels issue the same operations in the
 * same order, so the two produce identical 64-bit words.
 *
 * Conventions (identical to oracle.h unless stated):
 *   modulus 2^64 (wrapping uint64_t); LWE [n+1]; GLWE [(k+1)][N]; GGSW [(k+1)][ell][(k+1)][N] with GLev j < k
 *   encrypting +m*s_j and GLev k encrypting m (crypto/ggsw.rs:26-36), level i (least significant first)
 *   scaled by 2^(64 - (ell - i)*logB); external product = P_k - sum_{j<k} P_j (vtfhe/ggsw_ct.rs:98-112).
 *   (ell, logB) are tfhe-rs's (level, base_log): the decomposition keeps the TOP ell*logB bits.
 *   LWE KSK [k*N][ell_ks][n+1], row (i, j) encrypts -s_big[i] * 2^(64 - (ell_ks - j)*logB_ks) (vtfhe/mod.rs:52).
 */
#ifndef TFHE_ORACLE_TORUS_H
#define TFHE_ORACLE_TORUS_H
#include <stddef.h>
#include <stdint.h>

#include "oracle.h"

#ifdef __cplusplus
extern "C" {
#endif

#define TOR_ORDER_LWEKS_PBS 1 /* key switch kN -> n, mod switch, blind rotate, extract: ct [kN+1]   ("KS_PBS") */
#define TOR_ORDER_PBS_LWEKS 2 /* mod switch, blind rotate, extract, key switch kN -> n: ct [n+1]   (classic gate bootstrap) */

typedef struct {
    double re, im;
} tor_cplx;

/* ---- pieces ---- */
/* tfhe-rs SignedDecomposer{base_log, level}: closest representable, then balanced digits, least significant level first */
void tor_decompose(uint64_t x, uint32_t logB, uint32_t ell, int64_t *digits);
uint64_t tor_closest_representable(uint64_t x, uint32_t logB, uint32_t ell);
uint32_t tor_mod_switch(uint64_t x, uint32_t N);                                  /* crypto/lwe.rs:28-34 on 2^64 */
void tor_rotate(uint32_t N, const uint64_t *in, uint32_t shift, uint64_t *out);  /* X^shift * in, shift in [0,2N] */
uint64_t tor_delta(uint32_t plaintext_space);                                     /* 2^64 / space */
void tor_testv_lut(uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out); /* crypto/mod.rs:17-27 */

/* merged-twist twiddles for ring degree N: tw[v], v in [1, N/2), heap order (children 2v, 2v+1); tw[0] unused */
void tor_twiddles(uint32_t N, tor_cplx *tw);
/* forward: N/2 complex points z[j] = (p[j], p[j+N/2]), natural -> bit-reversed evaluation order, in place */
void tor_fft_forward(uint32_t N, const tor_cplx *tw, tor_cplx *z);
/* inverse of the above WITHOUT the 1/(N/2) factor (the engine folds it into the key) */
void tor_fft_inverse(uint32_t N, const tor_cplx *tw, tor_cplx *z);
/* y (a real in torus units) -> nearest multiple of 2^-51 of its fractional part, as a 64-bit torus element */
uint64_t tor_from_double(double y);

/* ---- context ---- */
typedef struct tor_ctx tor_ctx;
tor_ctx *tor_ctx_create(const orc_params *p); /* p->order is a TOR_ORDER_*; (ell, logB) in the tfhe-rs convention */
void tor_ctx_destroy(tor_ctx *c);
int tor_ctx_set_bsk(tor_ctx *c, const uint64_t *bsk_standard);   /* [n] GGSW, coefficient domain */
int tor_ctx_set_lwe_ksk(tor_ctx *c, const uint64_t *ksk);
const tor_cplx *tor_ctx_bsk_fourier(const tor_ctx *c);           /* [n][(k+1)][ell][(k+1)][N/2], scaled and sign-folded */
const tor_cplx *tor_ctx_twiddles(const tor_ctx *c);

/* ---- hot path ---- */
/* one GGSW in coefficient domain -> Fourier rows as the external product consumes them */
void tor_ggsw_to_fourier(const tor_ctx *c, const uint64_t *ggsw, tor_cplx *out);
void tor_external_product(const tor_ctx *c, const tor_cplx *ggsw_f, const uint64_t *glwe_in, uint64_t *glwe_out);
void tor_cmux_step(const tor_ctx *c, const tor_cplx *ggsw_f, uint64_t mask_element, uint64_t *acc);
void tor_blind_rotate_steps(const tor_ctx *c, const uint64_t *lwe_small, const uint64_t *testv, uint32_t steps, uint64_t *acc_out);
void tor_lwe_keyswitch(const tor_ctx *c, const uint64_t *lwe_big, uint64_t *lwe_small);
void tor_sample_extract(uint32_t N, const uint64_t *glwe, uint64_t *lwe);
int tor_pbs(const tor_ctx *c, const uint64_t *lwe_in, const uint64_t *testv, uint64_t *lwe_out);
int tor_pbs_batch(const tor_ctx *c, size_t B, const uint64_t *lwe_in, const uint64_t *testvs, const uint32_t *lut_idx, uint64_t *lwe_out);
size_t tor_ct_len(const tor_ctx *c);

/* ---- fixtures only ---- */
/* tfhe-rs TUniform(b): integers in [-2^b, 2^b], the two end points with half the probability of the others */
int64_t tor_rng_tuniform(orc_rng *r, uint32_t bound_log2);
/* noise spec: noise_log2 >= 0 selects TUniform(noise_log2); noise_log2 < 0 selects no noise */
void tor_lwe_encrypt(orc_rng *r, uint32_t n, const uint64_t *s, uint64_t m, int noise_log2, uint64_t *ct);
uint64_t tor_lwe_decrypt(uint32_t n, const uint64_t *s, const uint64_t *ct);
void tor_lwe_encrypt_batch(uint64_t seed, uint32_t n, const uint64_t *s, size_t B, const uint64_t *msgs, int noise_log2, uint64_t *cts);
void tor_lwe_decrypt_batch(uint32_t n, const uint64_t *s, size_t B, const uint64_t *cts, uint64_t *noisy);
uint64_t tor_decode(uint64_t noisy, uint64_t delta, uint32_t plaintext_space);
void tor_glwe_decrypt(uint32_t N, const uint64_t *s, const uint64_t *ct, uint64_t *m); /* k = 1, schoolbook negacyclic */
void tor_ggsw_encrypt(const tor_ctx *c, orc_rng *r, const uint64_t *s_glwe, uint64_t m_scalar, int noise_log2, uint64_t *ggsw);
void tor_bsk_gen(const tor_ctx *c, orc_rng *r, const uint64_t *s_lwe, const uint64_t *s_glwe, int noise_log2, uint64_t *bsk);
void tor_lwe_ksk_gen(const tor_ctx *c, orc_rng *r, const uint64_t *s_big, const uint64_t *s_small, int noise_log2, uint64_t *ksk);
double tor_centered_diff(uint64_t a, uint64_t b); /* signed (a - b) / 2^64 */

#ifdef __cplusplus
}
#endif
#endif
