/*
 * lux_tfhe.h — C ABI of the B200-native TFHE evaluation engine (libluxtfhe_b200.so).
 *
 * This is the drop-in boundary for the reference's evaluation hot path.  Reference
 * interfaces each entry point replaces (paths relative to /root/reference):
 *
 *   packages/wasm/src/index.ts:14-23   LuxFHE.{add,sub,eq,lt}(ct1, ct2, bootstrapKey, secretKey)
 *   packages/wasm/src/node.ts:29-96    init()/getLuxFHE(); keys and ciphertexts as Uint8Array by value
 *   js/v2-sdk/src/luxd/client.ts:128-147  POST /evaluate {op, left, right?, bitWidth}; ops at :34-39
 *   research/verifiable-fhe/src/main.rs:19-67 + vtfhe/ivc_based_vpbs.rs:159-386   one PBS
 *   research/verifiable-fhe/src/vtfhe/crypto/{poly,lwe,glwe,ggsw,mod}.rs           the pieces
 *
 * Differences a caller sees: the bootstrap key is uploaded ONCE (lux_load_*) and stays resident
 * in HBM in the NTT domain instead of being passed by value on every call; every op has a batch
 * dimension; errors are return codes + lux_last_error() instead of JS exceptions.
 *
 * Conventions
 *   - every function returns LUX_OK (0) or a negative lux_status; nothing aborts the process
 *   - all values are canonical Goldilocks elements (uint64_t < q = 2^64 - 2^32 + 1), little-endian; engines created with
 *     arith = LUX_ARITH_TORUS64_FFT carry torus words (any uint64_t, arithmetic mod 2^64) instead
 *   - layouts are the reference's flatten() orders (crypto/ggsw.rs:62-64, glev.rs:52-54, glwe.rs:133-135):
 *       LWE  [n+1] = mask..., body          GLWE [(k+1)][N] = mask polys..., body poly
 *       GGSW [(k+1)][ell][(k+1)][N]         BSK [n] GGSW          LWE-KSK [k*N][ell_ks][n+1]
 *   - host pointers unless the function name ends in _device; the caller owns every buffer it
 *     passes; the engine owns device memory; no pointer outlives the call except handles
 *   - every entry point is blocking and thread-safe per engine: calls on one engine from several host threads are serialised
 *     inside the library (a per-engine lock held for the whole call); calls on different engines run concurrently.
 *     lux_last_error() returns the message of the calling thread's own last failure on that engine
 *   - no CPU fallback exists: without a CUDA device lux_engine_create fails with LUX_ERR_NO_DEVICE
 */
#ifndef LUX_TFHE_H
#define LUX_TFHE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LUX_Q 0xFFFFFFFF00000001ULL

typedef enum {
    LUX_OK = 0,
    LUX_ERR_INVALID = -1,     /* bad argument / unsupported parameter combination */
    LUX_ERR_NO_DEVICE = -2,   /* no usable CUDA device */
    LUX_ERR_CUDA = -3,        /* CUDA runtime error (message in lux_last_error) */
    LUX_ERR_NO_KEY = -4,      /* evaluation key needed by the call has not been loaded */
    LUX_ERR_OOM = -5,
    LUX_ERR_FORMAT = -6,      /* serialized container: bad magic, version, length or checksum */
    LUX_ERR_MISMATCH = -7     /* serialized container is valid but for another parameter set / arithmetic / kind */
} lux_status;

/* order of key switch and bootstrap inside one PBS */
#define LUX_ORDER_PBS_GLWEKS 0 /* in-tree order: mod-switch, blind rotate, GLWE key switch, partial extract
                                  (vtfhe/ivc_based_vpbs.rs:109-124,355-371).  ct length n+1 */
#define LUX_ORDER_LWEKS_PBS 1  /* "KS_PBS" (PARAM_MESSAGE_2_CARRY_2_KS_PBS_*, sdk/relayer/src/tfhe.ts:15):
                                  LWE key switch, mod-switch, blind rotate, full extract.  ct length k*N+1 */

#define LUX_ORDER_PBS_LWEKS 2  /* classic gate-bootstrap order (torus arithmetic only): mod-switch, blind rotate, full extract,
                                  LWE key switch.  ct length n+1 */

/* arithmetic of the ciphertext modulus and of the polynomial products */
#define LUX_ARITH_GOLDILOCKS_NTT 0 /* q = 2^64 - 2^32 + 1, exact NTT: the in-tree reference arithmetic
                                      (research/verifiable-fhe/src/vtfhe/crypto/poly.rs:9-64); bit-exact vs the reference */
#define LUX_ARITH_TORUS64_FFT 1    /* q = 2^64 (wrapping words), f64 negacyclic FFT: the arithmetic the reference's NAMED shortint
                                      set is defined on (PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128, sdk/relayer/src/tfhe.ts:13-28,
                                      from the out-of-tree tfhe-rs).  (ell, logB) and (ell_ks, logB_ks) are then tfhe-rs's
                                      (level, base_log): the decomposition keeps the TOP ell*logB bits.  Supported: N = 2048, k = 1,
                                      ell = 1, logB <= 31, logB_ks <= 7.  Results are bit-identical to oracle/torus.c (fixed rounding
                                      order), decrypt-exact vs any other f64-FFT implementation */

/* key domains accepted by lux_load_bsk / lux_load_glwe_ksk */
#define LUX_DOMAIN_STANDARD 0 /* coefficient domain (Ggsw::encrypt output, crypto/ggsw.rs:26-36) */
#define LUX_DOMAIN_NTT 1      /* the reference's NTT domain (compute_bsk output, crypto/mod.rs:41-44) */

typedef struct {
    uint32_t n;       /* small LWE dimension                       main.rs:27 */
    uint32_t N;       /* ring degree: 512, 1024, 2048 or 4096      ntt/mod.rs:9 */
    uint32_t k;       /* GLWE dimension; the engine supports k = 1 main.rs:26 (K = k+1) */
    uint32_t ell;     /* gadget levels of the external product     main.rs:24 */
    uint32_t logB;    /* log2 gadget base                          main.rs:23 */
    uint32_t ell_ks;  /* LWE key switch levels (order 1); for order 0 the GLWE-KS reuses (ell, logB) */
    uint32_t logB_ks; /* log2 base of the LWE key switch (<= 24) */
    uint32_t order;   /* LUX_ORDER_* */
    uint32_t arith;   /* LUX_ARITH_* (0 = the in-tree Goldilocks arithmetic) */
} lux_params;

typedef struct lux_engine lux_engine;

/* ---- lifecycle -------------------------------------------------------------------------------- */
const char *lux_version(void);                                   /* replaces LuxFHE.version(), index.ts:15 */
int lux_engine_create(const lux_params *p, int device, lux_engine **out);
void lux_engine_destroy(lux_engine *e);
/* message of the last failing call on this engine (or of the last failing create when e == NULL) */
const char *lux_last_error(const lux_engine *e);
int lux_engine_params(const lux_engine *e, lux_params *out);
size_t lux_ct_words(const lux_engine *e);                        /* words per ciphertext entering/leaving a PBS */
/* kernels launched by this engine since creation (bench.py's gpu_launches) */
uint64_t lux_launch_count(const lux_engine *e);
/* ciphertexts resident on the GPU at once in the blind-rotate launch chosen for a batch of B (the
 * bootstrap key is streamed from HBM/L2 once per such wave; bench.py's roofline arithmetic) */
size_t lux_wave_ciphertexts(const lux_engine *e, size_t B);
/* when enabled, PBS calls bracket their kernels with CUDA events on the engine's stream */
int lux_set_timing(lux_engine *e, int enabled);
/* device milliseconds of the last timed PBS call: out[0] = key switch kernel, out[1] = blind-rotate kernel */
int lux_last_timing(const lux_engine *e, float out[2]);
int lux_synchronize(lux_engine *e);

/* ---- evaluation keys: uploaded once, converted to the NTT domain on the device ------------------
 * replaces the per-call `bootstrapKey` argument of index.ts:19-22 */
int lux_load_bsk(lux_engine *e, const uint64_t *bsk, size_t n_words, int domain);          /* crypto/mod.rs:29-45 */
int lux_load_glwe_ksk(lux_engine *e, const uint64_t *ksk, size_t n_words, int domain);     /* crypto/ggsw.rs:38-48 */
int lux_load_lwe_ksk(lux_engine *e, const uint64_t *ksk, size_t n_words);                  /* vtfhe/mod.rs:31-54 */
/* multi-GPU replication: the converted keys of rank 0 are broadcast (NCCL, by the caller) as one
 * opaque device blob and adopted by the other ranks.  bytes = lux_key_blob_bytes(). */
size_t lux_key_blob_bytes(const lux_engine *e);
int lux_export_keys_device(lux_engine *e, void *dev_blob, size_t bytes);
int lux_import_keys_device(lux_engine *e, const void *dev_blob, size_t bytes);

/* ---- several GPUs of one box behind one handle --------------------------------------------------
 * the boundary is create(params, n_gpus).  The reference's host (a Node process: packages/wasm/src/node.ts:29-63
 * loads ONE evaluator; js/v2-sdk/src/luxd/client.ts:128-147 posts whole batches) has no collective library of its own, so
 * replication and sharding happen inside this library: one engine per GPU, keys converted on GPU 0 and copied device to
 * device to the others, every batch call slices its batch into contiguous shares (whole kernel waves where the batch is large
 * enough) run by one host thread per GPU.  n_gpus <= 0 takes every visible device.  Calls on one handle are serialised. */
typedef struct lux_multi lux_multi;
int lux_engine_create_multi(const lux_params *p, int n_gpus, lux_multi **out);
void lux_multi_destroy(lux_multi *m);
int lux_multi_gpus(const lux_multi *m);
lux_engine *lux_multi_engine(lux_multi *m, int i);                 /* engine of GPU i (owned by the handle) */
const char *lux_multi_last_error(const lux_multi *m);
int lux_multi_load_bsk(lux_multi *m, const uint64_t *bsk, size_t n_words, int domain);
int lux_multi_load_glwe_ksk(lux_multi *m, const uint64_t *ksk, size_t n_words, int domain);
int lux_multi_load_lwe_ksk(lux_multi *m, const uint64_t *ksk, size_t n_words);
int lux_multi_replicate_keys(lux_multi *m);                        /* after the last load: GPU 0's resident keys -> all GPUs */
int lux_multi_pbs_batch(lux_multi *m, const uint64_t *lwe_in, const uint64_t *testvs, uint32_t n_luts,
                        const uint32_t *lut_idx, size_t B, uint64_t *lwe_out);

/* ---- serialized keys and ciphertexts ------------------------------------------------------------
 * The reference moves keys and ciphertexts BY VALUE as opaque bytes (packages/wasm/src/node.ts:65-96 Uint8Array,
 * js/v2-sdk/src/luxd/client.ts:128-147,232-244 number[]).  The producers of those bytes (Go luxfi/tfhe, tfhe-rs safe_serialize)
 * are out of tree, so this is the engine's own container, not a converter: an 80-byte little-endian header (magic, version,
 * kind, lux_params, aux, count, payload length, CRC-32 of payload and of header) followed by the words of the layouts above.
 * aux = bit width for LUX_BLOB_RADIX (0 for LUX_BLOB_LWE), LUX_DOMAIN_* for keys.  Host-only, no GPU needed. */
typedef enum {
    LUX_BLOB_LWE = 1,       /* count LWE ciphertexts [count][ct_words] */
    LUX_BLOB_RADIX = 2,     /* count euintN integers [count][blocks][ct_words], aux = N */
    LUX_BLOB_BSK = 3,
    LUX_BLOB_GLWE_KSK = 4,
    LUX_BLOB_LWE_KSK = 5
} lux_blob_kind;
typedef struct {
    uint32_t version, kind;
    lux_params params;
    uint32_t aux;
    uint64_t count, payload_words;
} lux_blob_info;
size_t lux_serialized_bytes(const lux_params *p, uint32_t kind, uint32_t aux, uint64_t count);    /* 0 = invalid combination */
int lux_serialize(const lux_params *p, uint32_t kind, uint32_t aux, uint64_t count, const uint64_t *words, uint8_t *out, size_t cap);
int lux_blob_inspect(const uint8_t *blob, size_t bytes, lux_blob_info *info);                     /* header only */
/* checks header, payload CRC and, when given, that the container matches `expect` / `expect_kind` (LUX_ERR_MISMATCH
 * otherwise); copies the payload to words_out when it is not NULL */
int lux_deserialize(const uint8_t *blob, size_t bytes, const lux_params *expect, uint32_t expect_kind, uint64_t *words_out,
                    size_t cap_words, lux_blob_info *info);
/* a serialized evaluation key -> this engine; a key of another parameter set or arithmetic is refused */
int lux_load_key_blob(lux_engine *e, const uint8_t *blob, size_t bytes);

/* ---- programmable bootstrap, batched ------------------------------------------------------------
 * lwe_in  [B][ct_words]   testvs [n_luts][N] test-vector polynomials (crypto/mod.rs:17-27)
 * lut_idx [B] index into testvs per ciphertext, or NULL for all 0      lwe_out [B][ct_words]
 * replaces verified_pbs's FHE arithmetic (ivc_based_vpbs.rs:277-371) */
int lux_pbs_batch(lux_engine *e, const uint64_t *lwe_in, const uint64_t *testvs, uint32_t n_luts,
                  const uint32_t *lut_idx, size_t B, uint64_t *lwe_out);
/* same with every buffer already in device memory; stream = cudaStream_t or NULL for the engine's */
int lux_pbs_batch_device(lux_engine *e, const uint64_t *d_lwe_in, const uint64_t *d_testvs, uint32_t n_luts,
                         const uint32_t *d_lut_idx, size_t B, uint64_t *d_lwe_out, void *stream);
/* lux_pbs_batch_device that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * lux_pbs_batch_device keeps the key-switched ciphertexts and the key switch's digit matrix in engine-owned scratch that it
 * grows with cudaMalloc; this variant keeps them in one file-scope `static __device__` array of LUX_STATIC_SCRATCH_BYTES inside
 * the library (every GPU that loads the library holds it until exit) and calls no cudaMalloc / cudaFree / allocating
 * library routine.  Same arguments, same result words as lux_pbs_batch_device.  Limits:
 *  - LUX_ARITH_TORUS64_FFT engines only, with the bootstrap key and the LWE key switch key loaded;
 *  - the batch must fit the array: ceil(B/128)*128 * k*N*ell_ks bytes of digits (rounded up to 256) plus
 *    B * (n+1) words (order LWEKS_PBS) or B * (k*N+1) words (order PBS_GLWEKS); at the named S2 set that is B <= 1024.
 *    A larger batch is refused with LUX_ERR_INVALID and a message that states the bytes needed, the limit and the largest B;
 *  - every call on a GPU, from any engine, shares the array, and it keeps the last call's contents (nothing in it has to start
 *    at zero: every word the kernels read is written earlier in the same call).  Calls that overlap ON THE GPU (two streams,
 *    two engines on one device) race; calls on one engine are serialised on the host like every other entry point, so pass the
 *    same stream, or order the streams yourself. */
#define LUX_STATIC_SCRATCH_BYTES (16u << 20)
int lux_pbs_batch_device_static(lux_engine *e, const uint64_t *d_lwe_in, const uint64_t *d_testvs, uint32_t n_luts,
                                const uint32_t *d_lut_idx, size_t B, uint64_t *d_lwe_out, void *stream);

/* ---- boolean gates: ONE programmable bootstrap per gate (BASELINE config 1, "single boolean-gate (NAND)
 * programmable bootstrap"; parameter set main.rs:23-30).  A bit m is the LWE encryption of m * lux_delta(4)
 * (main.rs:48-51).  gate(a, b) = sign-bootstrap(c * (a + b) + k) + q/8 with the constant test vector q/8:
 * the linear part and the final offset run on the device inside the call.  a, b, out: [B][ct_words].
 * The reference's own gate layer (Go luxfi/tfhe) is out of tree; this is the textbook TFHE gate bootstrap
 * restated over the in-tree encoding. */
typedef enum {
    LUX_GATE_NAND = 0,
    LUX_GATE_AND = 1,
    LUX_GATE_OR = 2,
    LUX_GATE_NOR = 3,
    LUX_GATE_XOR = 4,
    LUX_GATE_XNOR = 5
} lux_gate;
int lux_gate_batch(lux_engine *e, int gate, const uint64_t *a, const uint64_t *b, size_t B, uint64_t *out);
int lux_gate_batch_device(lux_engine *e, int gate, const uint64_t *d_a, const uint64_t *d_b, size_t B, uint64_t *d_out,
                          void *stream);
/* NOT is linear (delta - x): no bootstrap, host-side helper */
int lux_gate_not(const lux_engine *e, const uint64_t *a, size_t B, uint64_t *out);

/* test vector for a p-entry lookup table with one padding bit: p blocks of N/p copies of
 * lut[i]*delta, pre-rotated by half a block (get_testv, crypto/mod.rs:17-27).  Host-side helper. */
int lux_make_testv(uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out);
uint64_t lux_delta(uint32_t plaintext_space);                    /* get_delta, crypto/lwe.rs:46-48 */
/* the same two helpers for a given LUX_ARITH_* (torus: delta = 2^64 / space, wrapping products) */
int lux_make_testv_arith(uint32_t arith, uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out);
uint64_t lux_delta_arith(uint32_t arith, uint32_t plaintext_space);

/* ---- the pieces, exposed for parity checks against the reference ------------------------------- */
/* negacyclic NTT in the reference's convention (crypto/poly.rs:9-64): forward natural -> bit-reversed,
 * backward bit-reversed -> natural including the 1/N.  N from the engine's parameters. */
int lux_ntt_forward(lux_engine *e, const uint64_t *in, size_t n_polys, uint64_t *out);
int lux_ntt_backward(lux_engine *e, const uint64_t *in, size_t n_polys, uint64_t *out);
/* standalone NTT for any power-of-two N in [8, 2048] (covers every golden vector ntt/params_*.rs) */
int lux_ntt_any(lux_engine *e, uint32_t N, int inverse, const uint64_t *in, size_t n_polys, uint64_t *out);
/* signed gadget digits (vtfhe/glwe_poly.rs:28-50): out [count][ceil(64/logB)] field elements, LSB limb first */
int lux_decompose(lux_engine *e, const uint64_t *in, size_t count, uint32_t logB, uint64_t *out);
/* mod-switch to [0, 2N] (crypto/lwe.rs:28-38) */
int lux_mod_switch(lux_engine *e, const uint64_t *in, size_t count, uint32_t *out);
/* LWE key switch k*N -> n (vtfhe/mod.rs:31-54): in [B][k*N+1] -> out [B][n+1] */
int lux_lwe_keyswitch(lux_engine *e, const uint64_t *in, size_t B, uint64_t *out);
/* blind rotation only: lwe [B][n+1] under the small key; runs `steps` CMUX iterations
 * (steps = n for the whole rotation) and returns the raw accumulators acc_out [B][(k+1)*N]
 * (blind_rotation_step, vtfhe/mod.rs:119-136) */
int lux_blind_rotate(lux_engine *e, const uint64_t *lwe, const uint64_t *testvs, uint32_t n_luts,
                     const uint32_t *lut_idx, size_t B, uint32_t steps, uint64_t *acc_out);
/* GGSW (x) GLWE for `count` independent pairs (vtfhe/ggsw_ct.rs:98-112); ggsw in `domain` */
int lux_external_product(lux_engine *e, const uint64_t *ggsw, int domain, const uint64_t *glwe_in, size_t count,
                         uint64_t *glwe_out);

/* ---- pieces of the torus / f64-FFT arithmetic (LUX_ARITH_TORUS64_FFT engines only), for parity with oracle/torus.c --- */
/* negacyclic transform of folded polynomials: in/out [n_polys][N/2] complex (re, im) pairs of doubles; forward takes
 * z[j] = (p[j], p[j+N/2]) in natural order and returns the evaluations in the engine's (bit-reversed) order; inverse is
 * unnormalised (the engine folds 1/(N/2) into the key) */
int lux_fft(lux_engine *e, const double *in, size_t n_polys, int inverse, double *out);
/* one-level SignedDecomposer digits (tfhe-rs decomposer.rs) of the top logB bits: out [count] signed */
int lux_torus_decompose(lux_engine *e, const uint64_t *in, size_t count, int64_t *out);
/* the resident Fourier-domain bootstrap key as the kernels read it: out [n][32][4][32] (re, im) pairs */
int lux_export_bsk_fourier(lux_engine *e, double *out, size_t n_doubles);

/* ---- radix integers (euintN) over the PBS: see lux_radix.h ------------------------------------ */

#ifdef __cplusplus
}
#endif
#endif
