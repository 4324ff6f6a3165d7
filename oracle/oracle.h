/*
 * oracle.h — CPU restatement (plain C) of the reference's in-tree exact-NTT TFHE
 * programmable bootstrap.  TEST INFRASTRUCTURE ONLY: nothing under tfhe_b200/
 * may include, link or call this; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Reference followed (paths relative to /root/reference/research/verifiable-fhe/src):
 *   vtfhe/crypto/poly.rs  lwe.rs  glwe.rs  glev.rs  ggsw.rs  mod.rs   (plain arithmetic)
 *   vtfhe/glwe_poly.rs:28-50 (decompose)  vtfhe/mod.rs:31-136 (LWE-KS, mod-switch, rotate, CMUX)
 *   vtfhe/glev_ct.rs:92-110  vtfhe/ggsw_ct.rs:98-112  vtfhe/lev_ct.rs:63-76
 *   vtfhe/ivc_based_vpbs.rs:109-124,277-371 (driver order)  main.rs:23-64 (parameters)
 *   ntt/gen_param_file.sage:3-10,86-116 (table derivation)
 *
 * Parity status: the NTT (tables + forward + backward) is PINNED bit-exactly by the
 * reference's golden vectors ntt/params_{8..2048}.rs (tests/golden/ntt_golden.npz).
 * Decomposition, key switch, mod switch and whole-PBS *values* are pinned by nothing in
 * the reference (its tests are decrypt-properties on fresh randomness); for those this
 * restatement is the pin and the reference's properties are re-checked on it.  The
 * reference's real arithmetic core (Go github.com/luxfi/tfhe, version unpinned, absent
 * from the tree) cannot be built here: parity against *it* is unpinned.
 *
 * All values are canonical Goldilocks elements (uint64_t < q, q = 2^64 - 2^32 + 1),
 * little-endian u64 arrays, layouts exactly as the reference's flatten():
 *   LWE  [n+1]  (mask..., body)              GLWE [(k+1)][N]  (mask polys..., body poly)
 *   GLev [ell][(k+1)][N]                     GGSW [(k+1)][ell][(k+1)][N]
 *   BSK  [n] GGSW                            LWE-KSK [k*N][ell_ks][n+1]
 */
#ifndef TFHE_ORACLE_H
#define TFHE_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_Q 0xFFFFFFFF00000001ULL

/* order of key switch and bootstrap inside one PBS */
#define ORC_ORDER_PBS_GLWEKS 0 /* in-tree: mod-switch, blind rotate, GLWE key switch (KSK (x) acc), partial extract */
#define ORC_ORDER_LWEKS_PBS 1  /* "KS_PBS": LWE key switch kN->n, mod-switch, blind rotate, full extract */

typedef struct {
    uint32_t n;       /* small LWE dimension                                     main.rs:27 */
    uint32_t N;       /* ring degree, power of two                                ntt/mod.rs:9 */
    uint32_t k;       /* GLWE dimension (reference's K = k+1)                     main.rs:26 */
    uint32_t ell;     /* gadget levels kept by the external product               main.rs:24 */
    uint32_t logB;    /* log2 gadget base                                         main.rs:23 */
    uint32_t ell_ks;  /* levels of the key switch (GLWE-KS: must equal ell)       */
    uint32_t logB_ks; /* log2 base of the key switch (GLWE-KS: must equal logB)   */
    uint32_t order;   /* ORC_ORDER_* */
} orc_params;

/* ---- field ---- */
uint64_t orc_add(uint64_t a, uint64_t b);
uint64_t orc_sub(uint64_t a, uint64_t b);
uint64_t orc_mul(uint64_t a, uint64_t b);
uint64_t orc_mul_slow(uint64_t a, uint64_t b); /* (u128)a*b % q, to cross-check orc_mul */
uint64_t orc_neg(uint64_t a);
uint64_t orc_pow(uint64_t a, uint64_t e);
uint64_t orc_inv(uint64_t a);

/* ---- NTT (A1/A2) ---- */
/* tables as gen_param_file.sage: psi = 7^((q-1)/2N), roots[j] = psi^bitrev(j), ninv = 2^-logN */
int orc_ntt_tables(uint32_t N, uint64_t *roots, uint64_t *invroots, uint64_t *ninv);
void orc_ntt_forward(uint32_t N, const uint64_t *roots, uint64_t *a);
void orc_ntt_backward(uint32_t N, const uint64_t *invroots, uint64_t ninv, uint64_t *a);

/* ---- pieces (A3, A6, A7, A9, A10) ---- */
uint32_t orc_num_limbs(uint32_t logB); /* ceil(64/logB) */
/* all L = ceil(64/logB) signed digits of x as field elements, LSB limb first (glwe_poly.rs:28-50) */
void orc_decompose(uint64_t x, uint32_t logB, uint64_t *digits);
/* the same digits as signed integers (|d| <= B/2, or <= 2^(64-(L-1)logB) for a short top limb) */
void orc_decompose_signed(uint64_t x, uint32_t logB, int64_t *digits);
uint32_t orc_mod_switch(uint64_t x, uint32_t N); /* lwe.rs:28-34 -> [0, 2N] */
void orc_rotate(uint32_t N, const uint64_t *in, uint32_t shift, uint64_t *out); /* X^shift * in, shift in [0,2N] */
uint64_t orc_delta(uint32_t plaintext_space);                                     /* lwe.rs:46-48 */
void orc_testv(uint32_t N, uint32_t p, uint64_t delta, uint64_t *out);           /* crypto/mod.rs:17-27 */
void orc_testv_lut(uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out);
void orc_sample_extract(uint32_t N, uint32_t k, const uint64_t *glwe, uint64_t *lwe);            /* glwe.rs:96-106 */
void orc_partial_sample_extract(uint32_t N, uint32_t k, uint32_t nz, const uint64_t *glwe, uint64_t *lwe); /* :108-113 */

/* ---- context: parameters + tables + evaluation keys ---- */
typedef struct orc_ctx orc_ctx;
orc_ctx *orc_ctx_create(const orc_params *p);
void orc_ctx_destroy(orc_ctx *c);
const uint64_t *orc_ctx_roots(const orc_ctx *c);
const uint64_t *orc_ctx_invroots(const orc_ctx *c);
uint64_t orc_ctx_ninv(const orc_ctx *c);
/* keys are handed over in the STANDARD (coefficient) domain or already in the
 * reference's NTT domain (ntt_domain != 0), layouts as in the header comment */
int orc_ctx_set_bsk(orc_ctx *c, const uint64_t *bsk, int ntt_domain);
int orc_ctx_set_glwe_ksk(orc_ctx *c, const uint64_t *ksk_ggsw, int ntt_domain);
int orc_ctx_set_lwe_ksk(orc_ctx *c, const uint64_t *ksk);
const uint64_t *orc_ctx_bsk_ntt(const orc_ctx *c);

/* ---- hot path (A4, A5, A8, A11, A12) ---- */
/* GGSW (NTT domain) (x) GLWE (coefficient domain) -> GLWE (coefficient domain)  ggsw_ct.rs:98-112 */
void orc_external_product(const orc_ctx *c, const uint64_t *ggsw_ntt, const uint64_t *glwe_in, uint64_t *glwe_out);
/* acc <- acc + GGSW (x) (X^ms(mask) * acc - acc)                                vtfhe/mod.rs:119-136 */
void orc_cmux_step(const orc_ctx *c, const uint64_t *ggsw_ntt, uint64_t mask_element, uint64_t *acc);
/* acc0 = X^ms(-b) * (0,..,0,testv); n CMUX steps with the context's BSK        ivc_based_vpbs.rs:109-124,277-353 */
void orc_blind_rotate(const orc_ctx *c, const uint64_t *lwe_small, const uint64_t *testv, uint64_t *acc_out);
/* same but stop after `steps` CMUX iterations (debug / per-iteration parity) */
void orc_blind_rotate_steps(const orc_ctx *c, const uint64_t *lwe_small, const uint64_t *testv, uint32_t steps,
                            uint64_t *acc_out);
/* LWE key switch big (k*N) -> small (n)                                          vtfhe/mod.rs:31-54, lev_ct.rs:63-76 */
void orc_lwe_keyswitch(const orc_ctx *c, const uint64_t *lwe_big, uint64_t *lwe_small);
/* one programmable bootstrap in the context's order.
 * order 0: lwe_in [n+1] -> lwe_out [n+1];  order 1: lwe_in [kN+1] -> lwe_out [kN+1] */
int orc_pbs(const orc_ctx *c, const uint64_t *lwe_in, const uint64_t *testv, uint64_t *lwe_out);
/* batch (OpenMP over ciphertexts): testvs [n_luts][N], lut_idx [B] (NULL = all 0). returns threads used */
int orc_pbs_batch(const orc_ctx *c, size_t B, const uint64_t *lwe_in, const uint64_t *testvs, const uint32_t *lut_idx,
                  uint64_t *lwe_out);
size_t orc_lwe_in_len(const orc_ctx *c);  /* words per input/output ciphertext of orc_pbs */
int orc_max_threads(void);

/* ---- fixtures only: RNG, key generation, encryption, decryption, noise probes (A13) ---- */
typedef struct {
    uint64_t s[4];
    int have_spare;
    double spare;
} orc_rng;
void orc_rng_seed(orc_rng *r, uint64_t seed);
uint64_t orc_rng_u64(orc_rng *r);
uint64_t orc_rng_field(orc_rng *r);                  /* uniform in [0,q) */
uint64_t orc_rng_error(orc_rng *r, double sigma);    /* round(N(0, sigma*q)) as field element  lwe.rs:40-44 */

void orc_lwe_keygen(orc_rng *r, uint32_t n, uint64_t *s);                                      /* lwe.rs:10-14 */
void orc_glwe_keygen(orc_rng *r, uint32_t N, uint32_t k, uint64_t *s);                         /* glwe.rs:15-17 */
void orc_glwe_partial_key(orc_rng *r, uint32_t N, uint32_t k, uint32_t nz, uint64_t *s);       /* glwe.rs:19-38 */
void orc_lwe_encrypt(orc_rng *r, uint32_t n, const uint64_t *s, uint64_t m, double sigma, uint64_t *ct); /* lwe.rs:50-59 */
uint64_t orc_lwe_decrypt(uint32_t n, const uint64_t *s, const uint64_t *ct);                   /* lwe.rs:62-69 */
void orc_lwe_encrypt_batch(uint64_t seed, uint32_t n, const uint64_t *s, size_t B, const uint64_t *msgs, double sigma,
                           uint64_t *cts);                                                        /* B x lwe.rs:50-59 */
void orc_lwe_decrypt_batch(uint32_t n, const uint64_t *s, size_t B, const uint64_t *cts, uint64_t *noisy); /* B x :62-69 */
uint64_t orc_decode(uint64_t noisy, uint64_t delta, uint32_t plaintext_space);                 /* main.rs:61-64 */
void orc_glwe_encrypt(const orc_ctx *c, orc_rng *r, const uint64_t *s, const uint64_t *m, double sigma, uint64_t *ct);
void orc_glwe_decrypt(const orc_ctx *c, const uint64_t *s, const uint64_t *ct, uint64_t *m);   /* glwe.rs:60-63 */
/* GGSW(m) in the standard domain, m a polynomial [N]                                           ggsw.rs:26-36, glev.rs:26-38 */
void orc_ggsw_encrypt(const orc_ctx *c, orc_rng *r, const uint64_t *s, const uint64_t *m, double sigma, uint64_t *ggsw);
/* BSK_i = GGSW(s_lwe[i]) (standard domain)                                                     crypto/mod.rs:29-45 */
void orc_bsk_gen(const orc_ctx *c, orc_rng *r, const uint64_t *s_lwe, const uint64_t *s_glwe, double sigma, uint64_t *bsk);
/* GLWE->GLWE KSK as one GGSW (standard domain)                                                 ggsw.rs:38-48 */
void orc_glwe_ksk_gen(const orc_ctx *c, orc_rng *r, const uint64_t *s_to, const uint64_t *s_from, double sigma,
                      uint64_t *ksk);
/* LWE KSK [kN][ell_ks][n+1], rows encrypt -s_big[i] * B^(first+j)                              vtfhe/mod.rs:52 */
void orc_lwe_ksk_gen(const orc_ctx *c, orc_rng *r, const uint64_t *s_big, const uint64_t *s_small, double sigma,
                     uint64_t *ksk);
double orc_glwe_max_error(const orc_ctx *c, const uint64_t *s, const uint64_t *ct, const uint64_t *m); /* glwe.rs:65-78 */
double orc_glwe_avg_error(const orc_ctx *c, const uint64_t *s, const uint64_t *ct, const uint64_t *m); /* glwe.rs:80-94 */
/* signed centred distance (a-b) mod q as double / q */
double orc_centered_diff(uint64_t a, uint64_t b);

#ifdef __cplusplus
}
#endif
#endif
