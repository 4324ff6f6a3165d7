// engine.cu — C ABI (include/lux_tfhe.h) over the sm_100a kernels in tfhe_kernels.cuh.
// One engine per GPU; one process per GPU.  No CPU fallback: every entry point that computes
// launches CUDA kernels, and lux_engine_create fails without a device.
#include "engine.h"
#include "ks_mma.cuh"
#include "ks_tc5.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>

// templated kernels (blind rotate, batched NTT, external product) are instantiated per ring size in
// br_n{9,10,11,12}.cu (br_inst.cuh) so that they compile in parallel.
constexpr int KS_TB = 8, KS_TX = 128, KS_RC = 64;
__global__ void __launch_bounds__(KS_TX) lwe_keyswitch_kernel(const KSArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    lwe_keyswitch_body<KS_TB, KS_TX, KS_RC>(a, threadIdx.x, blockIdx.x, blockIdx.y, smem, blockIdx.z);
}
__global__ void ks_finish_kernel(const KSArgs a) { ks_finish_body(a, blockIdx.x * (size_t)blockDim.x + threadIdx.x); }

__global__ void ksk_colsum_kernel(const u64 *ksk, u64 *ksum, u32 rows, u32 cols) {
    ksk_colsum_body(ksk, ksum, rows, cols, blockIdx.x * blockDim.x + threadIdx.x);
}

__global__ void decompose_kernel(const u64 *in, u64 *out, size_t count, u32 logB, u32 L) {
    const bool centred = (L * logB == 64);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count * L; i += (size_t)gridDim.x * blockDim.x)
        out[i] = gl_digit(in[i / L], (u32)(i % L), logB, centred);
}

__global__ void mod_switch_kernel(const u64 *in, u32 *out, size_t count, u32 logN) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x)
        out[i] = gl_mod_switch(in[i], logN);
}


// field-arithmetic probe: a[] arbitrary u64 ("lazy"), b[] canonical; out[7][count] canonical residues of
// a+b, a-b, a*b, and both outputs of the two butterflies (u=a, v=b', w=b) where b' = a ^ b rotated.
__global__ void field_probe_kernel(const u64 *a, const u64 *b, u64 *out, size_t count) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        const u64 x = a[i], y = b[i];
        out[0 * count + i] = gl_canon(gl_add(x, y));
        out[1 * count + i] = gl_canon(gl_sub(x, y));
        out[2 * count + i] = gl_mul(x, y);
        u64 u = x, v = a[(i + 1) % count];
        gl_ct_butterfly(u, v, y);
        out[3 * count + i] = gl_canon(u);
        out[4 * count + i] = gl_canon(v);
        u = x;
        v = a[(i + 1) % count];
        gl_gs_butterfly(u, v, y);
        out[5 * count + i] = gl_canon(u);
        out[6 * count + i] = gl_canon(v);
    }
}

// integer-pipe micro-benchmarks (roofline denominators for the NTT/MAC work): each thread
// runs 8 independent chains of one instruction kind, no memory traffic.
//   kind 0: mad.wide.u32 (IMAD.WIDE.U32, fma pipe)   kind 1: 32-bit add (IADD3, alu pipe)
//   kind 2: Goldilocks butterfly (gl_ct_butterfly) in registers = the engine's own arithmetic kernel
template <int KIND>
__global__ void __launch_bounds__(256) int_peak_kernel(u64 *sink, u32 iters, u64 seed) {
    u64 x[8];
    const u64 t = seed + blockIdx.x * 256 + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = t * (2 * i + 3) + i;
    const u32 m = (u32)(t | 1);
    for (u32 it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (KIND == 0) {
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[i]) : "r"((u32)x[i]), "r"(m));
                } else if (KIND == 1) {
                    u32 lo = (u32)x[i], hi = (u32)(x[i] >> 32);
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(lo) : "r"(m));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(hi) : "r"(lo));
                    x[i] = ((u64)hi << 32) | lo;
                } else {
                    if (i & 1) gl_ct_butterfly(x[i - 1], x[i], (u64)m << 17 | r);
                }
            }
        }
    }
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc ^= x[i];
    if (acc == 0x1234567) sink[0] = acc;
}

// FP64 pipe peak: 16 independent DFMA chains per thread, register-only (the denominator of the torus arithmetic's roofline)
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *sink, u32 iters, double seed) {
    double x[16];
    const double a = 1.0 + 1e-9 * (double)threadIdx.x, b = seed * 1e-9;
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = seed + i + blockIdx.x;
    for (u32 it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 16; i++) x[i] = __fma_rn(x[i], a, b);
        }
    }
    double acc = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) acc += x[i];
    if (acc == 0.12345) sink[0] = acc;
}

// any power-of-two N in [8, 2048]: one CTA per polynomial, one thread per butterfly, textbook
// stage loop in shared memory (crypto/poly.rs:9-64).  Parity utility for the small golden vectors.
__global__ void ntt_any_kernel(const u64 *in, u64 *out, const u64 *tw, u64 scale, u32 N, u32 inverse) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64 *a = (u64 *)smem;
    const u32 h = threadIdx.x;
    const u64 *src = in + (size_t)blockIdx.x * N;
    a[h] = src[h];
    a[h + N / 2] = src[h + N / 2];
    __syncthreads();
    if (!inverse) {
        for (u32 m = 1; m < N; m <<= 1) {
            const u32 t = N / (2 * m), i = h / t, j = 2 * i * t + (h % t);
            u64 u = a[j], v = a[j + t];
            gl_ct_butterfly(u, v, tw[m + i]);
            a[j] = gl_canon(u);
            a[j + t] = gl_canon(v);
            __syncthreads();
        }
    } else {
        for (u32 m = N / 2; m >= 1; m >>= 1) {
            const u32 t = N / (2 * m), i = h / t, j = 2 * i * t + (h % t);
            u64 u = a[j], v = a[j + t];
            gl_gs_butterfly(u, v, tw[m + i]);
            a[j] = gl_canon(u);
            a[j + t] = gl_canon(v);
            __syncthreads();
        }
    }
    u64 *dst = out + (size_t)blockIdx.x * N;
    dst[h] = gl_mul(a[h], scale);
    dst[h + N / 2] = gl_mul(a[h + N / 2], scale);
}

// ------------------------------------------------------------------------------------------ helpers
static thread_local char g_create_error[512] = "";

static u32 ilog2u(u32 x) {
    u32 l = 0;
    while ((1u << l) < x) l++;
    return l;
}
static u32 bitrev_u(u32 x, u32 bits) {
    u32 r = 0;
    for (u32 i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
static u64 gl_pow_host(u64 a, u64 e) {
    u64 r = 1;
    while (e) {
        if (e & 1) r = gl_mul(r, a);
        a = gl_mul(a, a);
        e >>= 1;
    }
    return r;
}

// NTT tables exactly as the reference derives them (ntt/gen_param_file.sage:3-10,86-88,111-112):
// psi = 7^((q-1)/2N), ROOTS[j] = psi^bitrev(j), INVROOTS[j] = psi^-bitrev(j), NINV = 2^-logN.
void lux_host_ntt_tables(u32 N, std::vector<u64> &roots, std::vector<u64> &invroots, u64 &ninv) {
    const u32 logN = ilog2u(N);
    const u64 psi = gl_pow_host(7, (GL_Q - 1) / (2ULL * N));
    const u64 psi_inv = gl_pow_host(psi, GL_Q - 2);
    std::vector<u64> pw(N), pwi(N);
    pw[0] = pwi[0] = 1;
    for (u32 j = 1; j < N; j++) {
        pw[j] = gl_mul(pw[j - 1], psi);
        pwi[j] = gl_mul(pwi[j - 1], psi_inv);
    }
    roots.resize(N);
    invroots.resize(N);
    for (u32 j = 0; j < N; j++) {
        roots[j] = pw[bitrev_u(j, logN)];
        invroots[j] = pwi[bitrev_u(j, logN)];
    }
    ninv = gl_pow_host(gl_pow_host(2, GL_Q - 2), logN);
}

int lux_ensure_scratch(lux_engine *e, int slot, size_t bytes) {
    if (e->scratch_bytes[slot] >= bytes) return LUX_OK;
    if (e->scratch[slot]) cudaFree(e->scratch[slot]);
    e->scratch[slot] = nullptr;
    e->scratch_bytes[slot] = 0;
    size_t want = bytes + bytes / 4 + 4096;
    cudaError_t st = cudaMalloc(&e->scratch[slot], want);
    if (st != cudaSuccess) LUX_FAIL(e, LUX_ERR_OOM, "cudaMalloc(%zu) for scratch %d: %s", want, slot, cudaGetErrorString(st));
    e->scratch_bytes[slot] = want;
    return LUX_OK;
}

static int ensure_scratch(lux_engine *e, int slot, size_t bytes) { return lux_ensure_scratch(e, slot, bytes); }

static size_t ggsw_words(const lux_params &p) { return (size_t)(p.k + 1) * p.ell * (p.k + 1) * p.N; }

// ------------------------------------------------------------------------------------ kernel launchers
static int launch_br(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    int rc;
    switch (e->logN) {
        case 9: rc = lux_launch_br_9(e, a, s); break;
        case 10: rc = lux_launch_br_10(e, a, s); break;
        case 11: rc = lux_launch_br_11(e, a, s); break;
        case 12: rc = lux_launch_br_12(e, a, s); break;
        default: LUX_FAIL(e, LUX_ERR_INVALID, "unsupported N");
    }
    return rc;
}

// batched NTT of n_polys device polynomials, out may alias in
static int launch_ntt(lux_engine *e, const u64 *d_in, u64 *d_out, size_t n_polys, bool inverse, u64 scale, cudaStream_t s) {
    NttArgs a{d_in, d_out, e->d_tw_fwd, scale, (u32)n_polys, inverse ? 1u : 0u};
    switch (e->logN) {
        case 9: return lux_launch_ntt_9(e, a, s);
        case 10: return lux_launch_ntt_10(e, a, s);
        case 11: return lux_launch_ntt_11(e, a, s);
        case 12: return lux_launch_ntt_12(e, a, s);
    }
    LUX_FAIL(e, LUX_ERR_INVALID, "unsupported N");
}

// the int8 tensor-core form needs signed digits that fit a byte, exact int32 byte-plane sums and whole K tiles
static bool ks_mma_eligible(const lux_params &p) {
    const uint64_t K = (uint64_t)p.k * p.N * p.ell_ks;
    return p.order != LUX_ORDER_PBS_GLWEKS && p.logB_ks >= 1 && p.logB_ks <= 7 && K % ksmma::BK == 0 &&
           K * (1ull << (p.logB_ks - 1)) * 255ull < (1ull << 31);
}

// derive the K-major byte planes of the resident LWE key switch key (after a load or an import of the key blob)
static int derive_ksk_bytes(lux_engine *e) {
    if (!ks_mma_eligible(e->p)) return LUX_OK;
    const lux_params &p = e->p;
    const u32 K = p.k * p.N * p.ell_ks, cols = p.n + 1;
    e->ks_npad = (8 * cols + ksmma::BN - 1) / ksmma::BN * ksmma::BN;
    if (!e->d_kskT) {
        cudaError_t st = cudaMalloc(&e->d_kskT, (size_t)e->ks_npad * K);
        if (st != cudaSuccess) LUX_FAIL(e, LUX_ERR_OOM, "cudaMalloc(%zu) for key switch byte planes: %s", (size_t)e->ks_npad * K, cudaGetErrorString(st));
    }
    ksmma::ks_bytes_transpose_kernel<<<e->sm_count * 16, 256, 0, e->stream>>>(e->d_lksk, e->d_kskT, K, cols, e->ks_npad);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    // the tcgen05 form reads the same planes as 256 x 128 stage tiles in the tensor core's shared-memory image
    e->ks_npad5 = (8 * cols + kstc5::BN - 1) / kstc5::BN * kstc5::BN;
    if (!e->d_kskT5) {
        cudaError_t st = cudaMalloc(&e->d_kskT5, (size_t)e->ks_npad5 * K);
        if (st != cudaSuccess) LUX_FAIL(e, LUX_ERR_OOM, "cudaMalloc(%zu) for key switch stage tiles: %s", (size_t)e->ks_npad5 * K, cudaGetErrorString(st));
    }
    kstc5::ks_bytes_tiled_kernel<<<e->sm_count * 16, 256, 0, e->stream>>>(e->d_lksk, e->d_kskT5, K, cols, e->ks_npad5);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

static int launch_ks_mma(lux_engine *e, const u64 *d_in, u64 *d_out, size_t B, cudaStream_t s, bool tc5) {
    const lux_params &p = e->p;
    const u32 kN = p.k * p.N, K = kN * p.ell_ks;
    const u32 mpad = (u32)((B + ksmma::BM - 1) / ksmma::BM * ksmma::BM);
    const size_t need = (size_t)mpad * K;
    static_assert(ksmma::BM == kstc5::BM && ksmma::BK == kstc5::BK, "both forms pad the batch and K alike");
    if (e->ks_dig_bytes < need) {
        if (e->d_ks_dig) cudaFree(e->d_ks_dig);
        e->d_ks_dig = nullptr;
        e->ks_dig_bytes = 0;
        const size_t want = need + need / 4;
        cudaError_t st = cudaMalloc(&e->d_ks_dig, want);
        if (st != cudaSuccess) LUX_FAIL(e, LUX_ERR_OOM, "cudaMalloc(%zu) for key switch digits: %s", want, cudaGetErrorString(st));
        e->ks_dig_bytes = want;
    }
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(ksmma::lwe_keyswitch_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ksmma::SMEM_BYTES));
        CUDA_TRY(e, cudaFuncSetAttribute(kstc5::lwe_keyswitch_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kstc5::SMEM_BYTES));
        configured[e->device & 63] = true;
    }
    if (tc5) {
        kstc5::ks_digits_tiled_kernel<<<e->sm_count * 8, 256, 0, s>>>(d_in, e->d_ks_dig, (u32)B, mpad, kN, p.ell_ks, p.logB_ks, e->L_ks, lux_is_torus(e) ? 1u : 0u);
        e->launches++;
        CUDA_TRY(e, cudaGetLastError());
        kstc5::Args a{e->d_ks_dig, e->d_kskT5, d_in, d_out, (u32)B, p.n, kN, K, lux_is_torus(e) ? 1u : 0u, mpad / kstc5::BM, e->ks_npad5 / kstc5::BN};
        const u32 tiles = a.mtiles * a.ntiles;
        kstc5::lwe_keyswitch_tc5_kernel<<<std::min<u32>(tiles, (u32)e->sm_count), kstc5::THREADS, kstc5::SMEM_BYTES, s>>>(a);
        e->launches++;
        CUDA_TRY(e, cudaGetLastError());
        return LUX_OK;
    }
    ksmma::ks_digits_kernel<<<e->sm_count * 8, 256, 0, s>>>(d_in, e->d_ks_dig, (u32)B, mpad, kN, p.ell_ks, p.logB_ks, e->L_ks, lux_is_torus(e) ? 1u : 0u);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    ksmma::Args a{e->d_ks_dig, e->d_kskT, d_in, d_out, (u32)B, p.n, kN, K, lux_is_torus(e) ? 1u : 0u};
    dim3 grid(e->ks_npad / ksmma::BN, mpad / ksmma::BM);
    ksmma::lwe_keyswitch_mma_kernel<<<grid, ksmma::THREADS, ksmma::SMEM_BYTES, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

static int launch_ks(lux_engine *e, const u64 *d_in, u64 *d_out, size_t B, cudaStream_t s) {
    const lux_params &p = e->p;
    // the tensor-core forms are faster at every batch size (profiles/r01e_ks_probe.txt: 0.07 ms against 0.17 ms for one
    // ciphertext, 0.075 against 4.2 ms for 148), and of the two the tcgen05 / tensor-memory kernel (ks_tc5.cuh) again at every
    // batch size (profiles/r02e_ks_probe.txt: 0.039 against 0.075 ms for one ciphertext, 0.251 against 0.780 ms for 4,440);
    // the integer-pipe kernel below serves parameter sets they cannot take
    if (e->d_kskT && (e->ks_mode != 0 || lux_is_torus(e))) return launch_ks_mma(e, d_in, d_out, B, s, e->ks_mode != 1 && e->d_kskT5);
    if (lux_is_torus(e)) LUX_FAIL(e, LUX_ERR_INVALID, "torus arithmetic: key switch needs the tensor-core form (logB_ks <= 7)");
    KSArgs a{p.n, p.k * p.N, p.ell_ks, p.logB_ks, e->L_ks, (u32)B, e->d_lksk, e->d_ksum, d_in, d_out, 1, nullptr};
    const size_t smem = sizeof(u32) * KS_RC * p.ell_ks * KS_TB;
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(lwe_keyswitch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
        configured[e->device & 63] = true;
    }
    dim3 grid((p.n + 1 + KS_TX - 1) / KS_TX, (unsigned)((B + KS_TB - 1) / KS_TB));
    // small batches leave most SMs idle and make each CTA walk all k*N*ell_ks key rows one after the other:
    // cut the input coefficients into slices (grid z) and add the exact partial sums in a second tiny kernel
    const unsigned ctas = grid.x * grid.y, max_splits = (a.kN + KS_RC - 1) / KS_RC;
    unsigned splits = ctas >= (unsigned)e->sm_count ? 1u : std::min<unsigned>(max_splits, (2u * e->sm_count + ctas - 1) / ctas);
    if (e->ks_splits > 0) splits = std::min<unsigned>(max_splits, (unsigned)e->ks_splits);
    if (splits > 1) {
        int rc = ensure_scratch(e, 4, sizeof(u64) * splits * B * (p.n + 1));
        if (rc) return rc;
        a.splits = splits;
        a.partial = (u64 *)e->scratch[4];
        grid.z = splits;
    }
    lwe_keyswitch_kernel<<<grid, KS_TX, smem, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    if (splits > 1) {
        const size_t total = B * (size_t)(p.n + 1);
        ks_finish_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(a);
        e->launches++;
        CUDA_TRY(e, cudaGetLastError());
    }
    return LUX_OK;
}

// boolean gates (BASELINE config 1).  Bits are m * q/4 (main.rs:48-51: delta = q >> 2, one padding bit);
// gate(a, b) = sign-bootstrap(coef * (a + b) + konst) + q/8 with the constant test vector q/8:
//   x = coef*(a+b) + konst lands on +-q/8 or +-3q/8 (+-q/4 for XOR/XNOR), the bootstrap returns +q/8 on the
//   positive half-plane and -q/8 on the negative one, and the final +q/8 maps that to q/4 or 0.
// out[b][i] = coef * (a[b][i] + c[b][i]) (+ konst on the body), coef in {1, -1, 2, -2}
__global__ void gate_prepare_kernel(const u64 *a, const u64 *c, u64 *out, u32 w, size_t total, int coef, u64 konst) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        u64 s = gl_canon(gl_add(a[i], c[i]));
        if (coef == 2 || coef == -2) s = gl_canon(gl_add(s, s));
        if (coef < 0) s = gl_neg(s);
        if (i % w == w - 1) s = gl_canon(gl_add(s, konst));
        out[i] = s;
    }
}
__global__ void fill_kernel(u64 *p, u64 v, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

static BRArgs make_br_args(const lux_engine *e, const u64 *d_lwe, const u64 *d_testv, const u32 *d_idx, size_t B, u64 *d_out) {
    BRArgs a{};
    a.n = e->p.n;
    a.steps = e->p.n;
    a.ell = e->p.ell;
    a.logB = e->p.logB;
    a.L = e->L;
    a.B = (u32)B;
    a.bsk = e->d_bsk;
    a.ksk = nullptr;
    a.lwe = d_lwe;
    a.testv = d_testv;
    a.lut_idx = d_idx;
    a.lut_div = 1;
    a.out = d_out;
    a.tw_fwd = e->d_tw_fwd;
    return a;
}


static fft64::Args make_fft_args(const lux_engine *e, const u64 *d_lwe, const u64 *d_testv, const u32 *d_idx, u32 lut_div, size_t B, u64 *d_out) {
    fft64::Args a{};
    a.n = e->p.n;
    a.steps = e->p.n;
    a.logB = e->p.logB;
    a.B = (u32)B;
    a.out_mode = 0;
    a.lut_div = lut_div ? lut_div : 1;
    a.bsk = (const double2 *)e->d_bsk;
    a.lwe = d_lwe;
    a.testv = d_testv;
    a.lut_idx = d_idx;
    a.out = d_out;
    return a;
}

// PBS of the torus / f64-FFT arithmetic: order 1 = key switch, blind rotate + extract; order 2 = blind rotate + extract, key switch
static int torus_pbs_device(lux_engine *e, const u64 *d_in, const u64 *d_testvs, const u32 *d_idx, u32 lut_div, size_t B, u64 *d_out,
                            cudaStream_t s, u64 body_add) {
    if (!e->has_lksk) LUX_FAIL(e, LUX_ERR_NO_KEY, "LWE key switch key not loaded");
    CUDA_TRY(e, cudaSetDevice(e->device));
    const bool ks_first = e->p.order == LUX_ORDER_LWEKS_PBS;
    const size_t big = (size_t)e->p.k * e->p.N + 1, small = (size_t)e->p.n + 1;
    int rc = ensure_scratch(e, 0, sizeof(u64) * B * (ks_first ? small : big));
    if (rc) return rc;
    u64 *d_mid = (u64 *)e->scratch[0];
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev[0], s));
    if (ks_first && (rc = launch_ks(e, d_in, d_mid, B, s))) return rc;
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev[1], s));
    fft64::Args a = make_fft_args(e, ks_first ? d_mid : d_in, d_testvs, d_idx, lut_div, B, ks_first ? d_out : d_mid);
    a.body_add = body_add;
    if ((rc = lux_fft_launch_br(e, a, s))) return rc;
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev[2], s));
    if (!ks_first && (rc = launch_ks(e, d_mid, d_out, B, s))) return rc;
    if (e->timing) {
        CUDA_TRY(e, cudaEventRecord(e->ev[3], s));
        CUDA_TRY(e, cudaEventSynchronize(e->ev[3]));
        float ks_a = 0.f, ks_b = 0.f;
        CUDA_TRY(e, cudaEventElapsedTime(&ks_a, e->ev[0], e->ev[1]));
        CUDA_TRY(e, cudaEventElapsedTime(&e->last_ms[1], e->ev[1], e->ev[2]));
        CUDA_TRY(e, cudaEventElapsedTime(&ks_b, e->ev[2], e->ev[3]));
        e->last_ms[0] = ks_a + ks_b;
    }
    return LUX_OK;
}

// key switch + blind rotation with the raw accumulators out ([B][2N]) instead of one extracted sample: the radix layer reads
// several coefficients of an accumulator itself (dual tables, radix.cu).  Engines of order LUX_ORDER_LWEKS_PBS only.
int lux_pbs_device_glwe(lux_engine *e, const u64 *d_in, const u64 *d_testvs, const u32 *d_lut_idx, u32 lut_div, size_t B, u64 *d_glwe,
                        cudaStream_t s) {
    if (B == 0) return LUX_OK;
    if (B > 0x7fffffffu) LUX_FAIL(e, LUX_ERR_INVALID, "batch too large");
    if (e->p.order != LUX_ORDER_LWEKS_PBS) LUX_FAIL(e, LUX_ERR_INVALID, "accumulator output needs the key-switch-first order");
    if (!e->has_bsk) LUX_FAIL(e, LUX_ERR_NO_KEY, "bootstrap key not loaded");
    if (!e->has_lksk) LUX_FAIL(e, LUX_ERR_NO_KEY, "LWE key switch key not loaded");
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc = ensure_scratch(e, 0, sizeof(u64) * B * (e->p.n + 1));
    if (rc) return rc;
    u64 *d_small = (u64 *)e->scratch[0];
    if ((rc = launch_ks(e, d_in, d_small, B, s))) return rc;
    if (lux_is_torus(e)) {
        fft64::Args a = make_fft_args(e, d_small, d_testvs, d_lut_idx, lut_div, B, d_glwe);
        a.out_mode = 2;
        return lux_fft_launch_br(e, a, s);
    }
    BRArgs a = make_br_args(e, d_small, d_testvs, d_lut_idx, B, d_glwe);
    a.lut_div = lut_div ? lut_div : 1;
    a.out_mode = 2;
    return launch_br(e, a, s);
}

int lux_pbs_device_grouped(lux_engine *e, const u64 *d_in, const u64 *d_testvs, const u32 *d_lut_idx, u32 lut_div, size_t B,
                           u64 *d_out, cudaStream_t s, u64 body_add) {
    if (B == 0) return LUX_OK;
    if (B > 0x7fffffffu) LUX_FAIL(e, LUX_ERR_INVALID, "batch too large");
    if (!e->has_bsk) LUX_FAIL(e, LUX_ERR_NO_KEY, "bootstrap key not loaded");
    const bool ks_first = e->p.order == LUX_ORDER_LWEKS_PBS;
    if (lux_is_torus(e)) return torus_pbs_device(e, d_in, d_testvs, d_lut_idx, lut_div, B, d_out, s, body_add);
    if (ks_first && !e->has_lksk) LUX_FAIL(e, LUX_ERR_NO_KEY, "LWE key switch key not loaded");
    if (!ks_first && !e->has_gksk) LUX_FAIL(e, LUX_ERR_NO_KEY, "GLWE key switch key not loaded");
    CUDA_TRY(e, cudaSetDevice(e->device));
    const u64 *d_small = d_in;
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev[0], s));
    if (ks_first) {
        int rc = ensure_scratch(e, 0, sizeof(u64) * B * (e->p.n + 1));
        if (rc) return rc;
        rc = launch_ks(e, d_in, (u64 *)e->scratch[0], B, s);
        if (rc) return rc;
        d_small = (const u64 *)e->scratch[0];
    }
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev[1], s));
    BRArgs a = make_br_args(e, d_small, d_testvs, d_lut_idx, B, d_out);
    a.lut_div = lut_div ? lut_div : 1;
    a.body_add = body_add;
    if (ks_first) {
        a.out_mode = 0;
    } else {
        a.ksk = e->d_gksk;
        a.out_mode = 1;
        a.n_out = e->p.n;
    }
    int rc = launch_br(e, a, s);
    if (rc) return rc;
    if (e->timing) {
        CUDA_TRY(e, cudaEventRecord(e->ev[2], s));
        CUDA_TRY(e, cudaEventSynchronize(e->ev[2]));
        CUDA_TRY(e, cudaEventElapsedTime(&e->last_ms[0], e->ev[0], e->ev[1]));
        CUDA_TRY(e, cudaEventElapsedTime(&e->last_ms[1], e->ev[1], e->ev[2]));
    }
    return LUX_OK;
}

// --------------------------------------------------------------------------------------------- ABI
static thread_local char tl_err[512] = "";
static thread_local const lux_engine *tl_err_engine = nullptr;
void lux_note_error(const lux_engine *e) {
    tl_err_engine = e;
    snprintf(tl_err, sizeof(tl_err), "%s", e->err);
}

// ---- host-pointer pipeline state (lux_pbs_batch): two copy streams, per-slot events, pinned staging grown on demand
struct HostPipe {
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[2] = {}, ev_k[2] = {}, ev_out[2] = {};
    void *h_in[2] = {}, *h_out[2] = {};
    size_t cap = 0;
};
static int host_pipe_init(lux_engine *e) {
    if (e->pipe) return LUX_OK;
    HostPipe *hp = new (std::nothrow) HostPipe();
    if (!hp) LUX_FAIL(e, LUX_ERR_OOM, "out of host memory");
    e->pipe = hp;
    CUDA_TRY(e, cudaStreamCreateWithFlags(&hp->s_in, cudaStreamNonBlocking));
    CUDA_TRY(e, cudaStreamCreateWithFlags(&hp->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
        CUDA_TRY(e, cudaEventCreateWithFlags(&hp->ev_in[i], cudaEventDisableTiming));
        CUDA_TRY(e, cudaEventCreateWithFlags(&hp->ev_k[i], cudaEventDisableTiming));
        CUDA_TRY(e, cudaEventCreateWithFlags(&hp->ev_out[i], cudaEventDisableTiming));
    }
    return LUX_OK;
}
static int host_pipe_reserve(lux_engine *e, size_t bytes) {
    HostPipe *hp = e->pipe;
    if (hp->cap >= bytes) return LUX_OK;
    for (int i = 0; i < 2; i++) {
        if (hp->h_in[i]) cudaFreeHost(hp->h_in[i]);
        if (hp->h_out[i]) cudaFreeHost(hp->h_out[i]);
        hp->h_in[i] = hp->h_out[i] = nullptr;
    }
    hp->cap = 0;
    for (int i = 0; i < 2; i++) {
        if (cudaMallocHost(&hp->h_in[i], bytes) != cudaSuccess || cudaMallocHost(&hp->h_out[i], bytes) != cudaSuccess)
            LUX_FAIL(e, LUX_ERR_OOM, "cudaMallocHost(%zu) for the host staging buffers failed", bytes);
    }
    hp->cap = bytes;
    return LUX_OK;
}
static void host_pipe_destroy(lux_engine *e) {
    HostPipe *hp = e->pipe;
    if (!hp) return;
    for (int i = 0; i < 2; i++) {
        if (hp->h_in[i]) cudaFreeHost(hp->h_in[i]);
        if (hp->h_out[i]) cudaFreeHost(hp->h_out[i]);
        if (hp->ev_in[i]) cudaEventDestroy(hp->ev_in[i]);
        if (hp->ev_k[i]) cudaEventDestroy(hp->ev_k[i]);
        if (hp->ev_out[i]) cudaEventDestroy(hp->ev_out[i]);
    }
    if (hp->s_in) cudaStreamDestroy(hp->s_in);
    if (hp->s_out) cudaStreamDestroy(hp->s_out);
    delete hp;
    e->pipe = nullptr;
}
// page-locked (cudaHostAlloc / cudaHostRegister) caller memory is copied from directly
static bool host_is_pinned(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost;
}

extern "C" {

const char *lux_version(void) { return "tfhe_b200 0.2.0 (sm_100a; Goldilocks exact-NTT and 2^64-torus f64-FFT PBS)"; }

// the failing thread reads its own message (another thread's later failure on the same engine cannot overwrite it)
const char *lux_last_error(const lux_engine *e) {
    if (!e) return g_create_error;
    if (tl_err_engine == e && tl_err[0]) return tl_err;
    return e->err;
}

int lux_engine_create(const lux_params *p, int device, lux_engine **out) {
#define CREATE_FAIL(code, ...)                                          \
    do {                                                                \
        snprintf(g_create_error, sizeof(g_create_error), __VA_ARGS__);  \
        return (code);                                                  \
    } while (0)
    if (!p || !out) CREATE_FAIL(LUX_ERR_INVALID, "null argument");
    *out = nullptr;
    if (p->k != 1) CREATE_FAIL(LUX_ERR_INVALID, "only k = 1 is supported (got %u)", p->k);
    if (p->arith != LUX_ARITH_GOLDILOCKS_NTT && p->arith != LUX_ARITH_TORUS64_FFT) CREATE_FAIL(LUX_ERR_INVALID, "bad arith %u", p->arith);
    const bool torus = p->arith == LUX_ARITH_TORUS64_FFT;
    if (torus) {
        if (p->N != 2048) CREATE_FAIL(LUX_ERR_INVALID, "torus arithmetic: N must be 2048 (got %u)", p->N);
        if (p->ell != 1 || p->logB < 1 || p->logB > 31) CREATE_FAIL(LUX_ERR_INVALID, "torus arithmetic: ell = 1, logB <= 31 (got %u, %u)", p->ell, p->logB);
        if (p->order != LUX_ORDER_LWEKS_PBS && p->order != LUX_ORDER_PBS_LWEKS) CREATE_FAIL(LUX_ERR_INVALID, "torus arithmetic: order must be LUX_ORDER_LWEKS_PBS or LUX_ORDER_PBS_LWEKS");
        if (p->logB_ks < 1 || p->ell_ks < 1 || p->ell_ks * p->logB_ks > 64 || !ks_mma_eligible(*p))
            CREATE_FAIL(LUX_ERR_INVALID, "torus arithmetic: unsupported key-switch gadget (ell_ks=%u, logB_ks=%u)", p->ell_ks, p->logB_ks);
        if (p->n < 1) CREATE_FAIL(LUX_ERR_INVALID, "bad LWE dimension n=%u", p->n);
    } else {
    if (p->N != 512 && p->N != 1024 && p->N != 2048 && p->N != 4096)
        CREATE_FAIL(LUX_ERR_INVALID, "N must be 512, 1024, 2048 or 4096 (got %u)", p->N);
    if (p->logB < 1 || p->logB > 62 || p->ell < 1 || p->ell > (64 + p->logB - 1) / p->logB)
        CREATE_FAIL(LUX_ERR_INVALID, "bad gadget (ell=%u, logB=%u)", p->ell, p->logB);
    if (p->order != LUX_ORDER_PBS_GLWEKS && p->order != LUX_ORDER_LWEKS_PBS)
        CREATE_FAIL(LUX_ERR_INVALID, "bad order %u", p->order);
    if (p->order == LUX_ORDER_LWEKS_PBS &&
        (p->logB_ks < 1 || p->logB_ks > 24 || p->ell_ks < 1 || p->ell_ks > (64 + p->logB_ks - 1) / p->logB_ks))
        CREATE_FAIL(LUX_ERR_INVALID, "bad key-switch gadget (ell_ks=%u, logB_ks=%u)", p->ell_ks, p->logB_ks);
    if (p->n < 1 || (p->order == LUX_ORDER_PBS_GLWEKS && p->n > p->N))
        CREATE_FAIL(LUX_ERR_INVALID, "bad LWE dimension n=%u", p->n);
    }
    int ndev = 0;
    cudaError_t st = cudaGetDeviceCount(&ndev);
    if (st != cudaSuccess || ndev == 0)
        CREATE_FAIL(LUX_ERR_NO_DEVICE, "no CUDA device: %s (this engine has no CPU fallback)",
                    st != cudaSuccess ? cudaGetErrorString(st) : "device count is 0");
    if (device < 0 || device >= ndev) CREATE_FAIL(LUX_ERR_INVALID, "device %d out of range [0,%d)", device, ndev);
    if ((st = cudaSetDevice(device)) != cudaSuccess) CREATE_FAIL(LUX_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(st));
    cudaDeviceProp prop;
    if ((st = cudaGetDeviceProperties(&prop, device)) != cudaSuccess)
        CREATE_FAIL(LUX_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(st));
    if (prop.major != 10)
        CREATE_FAIL(LUX_ERR_NO_DEVICE, "device %d is sm_%d%d; this build carries sm_100a code only", device, prop.major, prop.minor);

    lux_engine *e = new (std::nothrow) lux_engine();
    if (!e) CREATE_FAIL(LUX_ERR_OOM, "out of host memory");
    e->p = *p;
    e->device = device;
    e->sm_count = prop.multiProcessorCount;
    e->logN = ilog2u(p->N);
    e->L = (64 + p->logB - 1) / p->logB;
    e->L_ks = p->logB_ks ? (64 + p->logB_ks - 1) / p->logB_ks : 0;
    std::vector<u64> roots, inv;
    lux_host_ntt_tables(p->N, roots, inv, e->ninv);
    // the kernels hard-code the twiddles of NTT stages 0..3 as powers of two (gl64.cuh gl_root_log2): refuse to run
    // if the generated table disagrees
#if defined(GL_POW2_TWIDDLES)
    for (u32 j = 0; j < 16 && j < p->N; j++)
        if (roots[j] != gl_pow_host(2, (u64)gl_root_log2((int)j))) {
            delete e;
            CREATE_FAIL(LUX_ERR_INVALID, "internal: ROOTS[%u] is not 2^%d", j, gl_root_log2((int)j));
        }
#endif
    bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaMalloc(&e->d_tw_fwd, sizeof(u64) * p->N * 2) == cudaSuccess;
    if (ok) {
        e->d_tw_inv = e->d_tw_fwd + p->N;
        ok = cudaMemcpy(e->d_tw_fwd, roots.data(), sizeof(u64) * p->N, cudaMemcpyHostToDevice) == cudaSuccess &&
             cudaMemcpy(e->d_tw_inv, inv.data(), sizeof(u64) * p->N, cudaMemcpyHostToDevice) == cudaSuccess;
    }
    // key blob: [header 8 words][bsk][glwe ksk][lwe ksk][ksum]
    e->bsk_words = (size_t)p->n * ggsw_words(*p);
    e->gksk_words = p->order == LUX_ORDER_PBS_GLWEKS ? ggsw_words(*p) : 0;
    e->lksk_words = p->order != LUX_ORDER_PBS_GLWEKS ? (size_t)p->k * p->N * p->ell_ks * (p->n + 1) : 0;
    e->ksum_words = p->order != LUX_ORDER_PBS_GLWEKS ? (size_t)p->n + 1 : 0;
    e->blob_words = 8 + e->bsk_words + e->gksk_words + e->lksk_words + e->ksum_words;
    ok = ok && cudaMalloc(&e->d_blob, sizeof(u64) * e->blob_words) == cudaSuccess;
    for (int i = 0; i < 4 && ok; i++) ok = cudaEventCreate(&e->ev[i]) == cudaSuccess;
    if (ok && torus) ok = lux_fft_init(e) == LUX_OK;
    if (!ok) {
        snprintf(g_create_error, sizeof(g_create_error), "device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        lux_engine_destroy(e);
        return LUX_ERR_OOM;
    }
    e->d_bsk = e->d_blob + 8;
    e->d_gksk = e->d_bsk + e->bsk_words;
    e->d_lksk = e->d_gksk + e->gksk_words;
    e->d_ksum = e->d_lksk + e->lksk_words;
    *out = e;
    return LUX_OK;
#undef CREATE_FAIL
}

void lux_engine_destroy(lux_engine *e) {
    if (!e) return;
    e->mtx.lock();   // a call still running on another thread finishes first; using the handle afterwards is the caller's bug
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    lux_radix_destroy(e);
    for (int i = 0; i < LUX_SCRATCH_SLOTS; i++)
        if (e->scratch[i]) cudaFree(e->scratch[i]);
    for (auto &kv : e->any_tables) cudaFree(kv.second);
    if (e->d_kskT) cudaFree(e->d_kskT);
    if (e->d_kskT5) cudaFree(e->d_kskT5);
    if (e->d_ks_dig) cudaFree(e->d_ks_dig);
    if (e->d_fft_twB) cudaFree(e->d_fft_twB);
    if (e->d_fft_twFull) cudaFree(e->d_fft_twFull);
    if (e->d_tw_fwd) cudaFree(e->d_tw_fwd);
    if (e->d_blob) cudaFree(e->d_blob);
    for (int i = 0; i < 4; i++)
        if (e->ev[i]) cudaEventDestroy(e->ev[i]);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->d_fft_trace) cudaFree(e->d_fft_trace);
    host_pipe_destroy(e);
    e->mtx.unlock();
    delete e;
}

int lux_engine_params(const lux_engine *e, lux_params *out) {
    if (!e || !out) return LUX_ERR_INVALID;
    *out = e->p;
    return LUX_OK;
}
size_t lux_ct_words(const lux_engine *e) {
    return e->p.order == LUX_ORDER_LWEKS_PBS ? (size_t)e->p.k * e->p.N + 1 : (size_t)e->p.n + 1;   // orders 0 and 2: small key
}
uint64_t lux_launch_count(const lux_engine *e) { return e ? e->launches : 0; }
size_t lux_wave_ciphertexts(const lux_engine *e, size_t B) {
    if (!e) return 0;
    if (lux_is_torus(e)) return lux_fft_wave(e, B);
    switch (e->logN) {
        case 9: return lux_br_wave_9(e, B);
        case 10: return lux_br_wave_10(e, B);
        case 11: return lux_br_wave_11(e, B);
        case 12: return lux_br_wave_12(e, B);
    }
    return 0;
}
int lux_set_timing(lux_engine *e, int enabled) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    e->timing = enabled != 0;
    return LUX_OK;
}
int lux_last_timing(const lux_engine *e, float out[2]) {
    if (!e || !out) return LUX_ERR_INVALID;
    out[0] = e->last_ms[0];
    out[1] = e->last_ms[1];
    return LUX_OK;
}
int lux_synchronize(lux_engine *e) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}
int lux_debug_force_teams(lux_engine *e, int teams) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    e->force_teams = teams;
    return LUX_OK;
}
int lux_debug_ks_splits(lux_engine *e, int splits) {
    LUX_LOCK(e);
    if (!e || splits < 0) return LUX_ERR_INVALID;
    e->ks_splits = splits;
    return LUX_OK;
}
int lux_debug_br_regs(lux_engine *e, int regs) {
    LUX_LOCK(e);
    if (!e || (regs != 4 && regs != 8 && regs != 16)) return LUX_ERR_INVALID;
    e->br_regs = regs;
    return LUX_OK;
}
int lux_debug_fft_variant(lux_engine *e, int variant) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    e->fft_variant = variant;
    return LUX_OK;
}
int lux_debug_fft_lat_mode(lux_engine *e, int mode) {
    if (!e || mode < -1 || mode > 1) return LUX_ERR_INVALID;
    e->fft_lat_mode = mode;
    return LUX_OK;
}
int lux_debug_fft_tail_split(lux_engine *e, int on) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    e->fft_tail_split = on != 0;
    return LUX_OK;
}
int lux_debug_fft_stagger(lux_engine *e, int ns) {
    LUX_LOCK(e);
    if (!e || ns < 0) return LUX_ERR_INVALID;
    e->fft_stagger_ns = ns;
    return LUX_OK;
}
// returns the phase-cycle counters accumulated since the previous call and clears them (first call: allocates, enables)
int lux_debug_fft_trace(lux_engine *e, unsigned long long out[16]) {
    LUX_LOCK(e);
    if (!e || !out) return LUX_ERR_INVALID;
    CUDA_TRY(e, cudaSetDevice(e->device));
    if (!e->d_fft_trace) {
        CUDA_TRY(e, cudaMalloc(&e->d_fft_trace, 16 * sizeof(u64)));
        CUDA_TRY(e, cudaMemset(e->d_fft_trace, 0, 16 * sizeof(u64)));
    }
    CUDA_TRY(e, cudaDeviceSynchronize());
    CUDA_TRY(e, cudaMemcpy(out, e->d_fft_trace, 16 * sizeof(u64), cudaMemcpyDeviceToHost));
    CUDA_TRY(e, cudaMemset(e->d_fft_trace, 0, 16 * sizeof(u64)));
    return LUX_OK;
}
int lux_debug_cluster_regs(lux_engine *e, int regs) {
    LUX_LOCK(e);
    if (!e || (regs != 0 && regs != 4 && regs != 8 && regs != 16)) return LUX_ERR_INVALID;
    e->cluster_regs = regs;
    return LUX_OK;
}
int lux_debug_generic_gadget(lux_engine *e, int on) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    e->generic_gadget = on != 0;
    return LUX_OK;
}
int lux_debug_ks_mode(lux_engine *e, int mode) {
    LUX_LOCK(e);
    if (!e || mode < -1 || mode > 2) return LUX_ERR_INVALID;
    e->ks_mode = mode;
    return LUX_OK;
}
int lux_debug_split_mode(lux_engine *e, int mode) {
    LUX_LOCK(e);
    if (!e || mode < -1 || mode > 2) return LUX_ERR_INVALID;
    e->split_mode = mode;
    return LUX_OK;
}

// ---- keys
static int load_ggsw_array(lux_engine *e, u64 *d_dst, const u64 *src, size_t words, size_t expect, int domain, const char *what) {
    if (!src) LUX_FAIL(e, LUX_ERR_INVALID, "%s: null pointer", what);
    if (words != expect) LUX_FAIL(e, LUX_ERR_INVALID, "%s: expected %zu words, got %zu", what, expect, words);
    if (domain != LUX_DOMAIN_STANDARD && domain != LUX_DOMAIN_NTT) LUX_FAIL(e, LUX_ERR_INVALID, "%s: bad domain %d", what, domain);
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaMemcpyAsync(d_dst, src, sizeof(u64) * words, cudaMemcpyHostToDevice, e->stream));
    if (domain == LUX_DOMAIN_STANDARD) {
        int rc = launch_ntt(e, d_dst, d_dst, words / e->p.N, false, 1, e->stream);  // crypto/mod.rs:41-44
        if (rc) return rc;
    }
    int rc = LUX_ERR_INVALID;   // reference NTT order -> resident key rows (fold 1/N and sign, lane-contiguous layout)
    switch (e->logN) {
        case 9: rc = lux_launch_keyfmt_9(e, d_dst, words / e->p.N, e->stream); break;
        case 10: rc = lux_launch_keyfmt_10(e, d_dst, words / e->p.N, e->stream); break;
        case 11: rc = lux_launch_keyfmt_11(e, d_dst, words / e->p.N, e->stream); break;
        case 12: rc = lux_launch_keyfmt_12(e, d_dst, words / e->p.N, e->stream); break;
    }
    if (rc) return rc;
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

// torus arithmetic: coefficient-domain key words -> Fourier rows in the ring layout (oracle/torus.c tor_ggsw_to_fourier)
static int load_bsk_torus(lux_engine *e, const uint64_t *bsk, size_t n_words, int domain) {
    if (!bsk) LUX_FAIL(e, LUX_ERR_INVALID, "lux_load_bsk: null pointer");
    if (n_words != e->bsk_words) LUX_FAIL(e, LUX_ERR_INVALID, "lux_load_bsk: expected %zu words, got %zu", e->bsk_words, n_words);
    if (domain != LUX_DOMAIN_STANDARD) LUX_FAIL(e, LUX_ERR_INVALID, "lux_load_bsk: the torus arithmetic takes coefficient-domain keys only");
    CUDA_TRY(e, cudaSetDevice(e->device));
    u64 *d_tmp = nullptr;
    cudaError_t st = cudaMalloc(&d_tmp, sizeof(u64) * n_words);
    if (st != cudaSuccess) LUX_FAIL(e, LUX_ERR_OOM, "cudaMalloc(%zu) for key upload: %s", sizeof(u64) * n_words, cudaGetErrorString(st));
    int rc = LUX_OK;
    if (cudaMemcpyAsync(d_tmp, bsk, sizeof(u64) * n_words, cudaMemcpyHostToDevice, e->stream) != cudaSuccess) rc = LUX_ERR_CUDA;
    if (rc == LUX_OK) rc = lux_fft_key_convert(e, d_tmp, e->d_bsk, n_words / e->p.N, e->stream);
    if (cudaStreamSynchronize(e->stream) != cudaSuccess && rc == LUX_OK) rc = LUX_ERR_CUDA;
    cudaFree(d_tmp);
    if (rc == LUX_ERR_CUDA) LUX_FAIL(e, LUX_ERR_CUDA, "key upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    return rc;
}

int lux_load_bsk(lux_engine *e, const uint64_t *bsk, size_t n_words, int domain) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (lux_is_torus(e)) {
        int rc = load_bsk_torus(e, bsk, n_words, domain);
        if (rc == LUX_OK) e->has_bsk = true;
        return rc;
    }
    int rc = load_ggsw_array(e, e->d_bsk, bsk, n_words, e->bsk_words, domain, "lux_load_bsk");
    if (rc == LUX_OK) e->has_bsk = true;
    return rc;
}
int lux_load_glwe_ksk(lux_engine *e, const uint64_t *ksk, size_t n_words, int domain) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (e->p.order != LUX_ORDER_PBS_GLWEKS) LUX_FAIL(e, LUX_ERR_INVALID, "GLWE key switch key only applies to LUX_ORDER_PBS_GLWEKS");
    int rc = load_ggsw_array(e, e->d_gksk, ksk, n_words, e->gksk_words, domain, "lux_load_glwe_ksk");
    if (rc == LUX_OK) e->has_gksk = true;
    return rc;
}
int lux_load_lwe_ksk(lux_engine *e, const uint64_t *ksk, size_t n_words) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (e->p.order == LUX_ORDER_PBS_GLWEKS) LUX_FAIL(e, LUX_ERR_INVALID, "LWE key switch key does not apply to LUX_ORDER_PBS_GLWEKS");
    if (!ksk) LUX_FAIL(e, LUX_ERR_INVALID, "lux_load_lwe_ksk: null pointer");
    if (n_words != e->lksk_words) LUX_FAIL(e, LUX_ERR_INVALID, "lux_load_lwe_ksk: expected %zu words, got %zu", e->lksk_words, n_words);
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaMemcpyAsync(e->d_lksk, ksk, sizeof(u64) * n_words, cudaMemcpyHostToDevice, e->stream));
    const u32 cols = e->p.n + 1, rows = e->p.k * e->p.N * e->p.ell_ks;
    if (!lux_is_torus(e)) {
        ksk_colsum_kernel<<<(cols + 127) / 128, 128, 0, e->stream>>>(e->d_lksk, e->d_ksum, rows, cols);
        e->launches++;
    }
    CUDA_TRY(e, cudaGetLastError());
    int rc = derive_ksk_bytes(e);
    if (rc) return rc;
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    e->has_lksk = true;
    return LUX_OK;
}

size_t lux_key_blob_bytes(const lux_engine *e) { return e ? sizeof(u64) * e->blob_words : 0; }

int lux_export_keys_device(lux_engine *e, void *dev_blob, size_t bytes) {
    LUX_LOCK(e);
    if (!e || !dev_blob) return LUX_ERR_INVALID;
    if (bytes != sizeof(u64) * e->blob_words) LUX_FAIL(e, LUX_ERR_INVALID, "key blob is %zu bytes, got %zu", sizeof(u64) * e->blob_words, bytes);
    CUDA_TRY(e, cudaSetDevice(e->device));
    u64 hdr[8] = {0x4c55584b45593031ULL, e->has_bsk, e->has_gksk, e->has_lksk, e->blob_words, e->p.arith, 0, 0};
    CUDA_TRY(e, cudaMemcpyAsync(e->d_blob, hdr, sizeof(hdr), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(dev_blob, e->d_blob, bytes, cudaMemcpyDeviceToDevice, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}
int lux_import_keys_device(lux_engine *e, const void *dev_blob, size_t bytes) {
    LUX_LOCK(e);
    if (!e || !dev_blob) return LUX_ERR_INVALID;
    if (bytes != sizeof(u64) * e->blob_words) LUX_FAIL(e, LUX_ERR_INVALID, "key blob is %zu bytes, got %zu", sizeof(u64) * e->blob_words, bytes);
    CUDA_TRY(e, cudaSetDevice(e->device));
    u64 hdr[8];
    CUDA_TRY(e, cudaMemcpy(hdr, dev_blob, sizeof(hdr), cudaMemcpyDeviceToHost));
    if (hdr[0] != 0x4c55584b45593031ULL || hdr[4] != e->blob_words || hdr[5] != e->p.arith) LUX_FAIL(e, LUX_ERR_INVALID, "key blob header mismatch");
    CUDA_TRY(e, cudaMemcpyAsync(e->d_blob, dev_blob, bytes, cudaMemcpyDeviceToDevice, e->stream));
    if (hdr[3] != 0) {
        int rc = derive_ksk_bytes(e);
        if (rc) return rc;
    }
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    e->has_bsk = hdr[1] != 0;
    e->has_gksk = hdr[2] != 0;
    e->has_lksk = hdr[3] != 0;
    return LUX_OK;
}

// ---- PBS
int lux_pbs_batch_device(lux_engine *e, const uint64_t *d_in, const uint64_t *d_testvs, uint32_t n_luts,
                         const uint32_t *d_lut_idx, size_t B, uint64_t *d_out, void *stream) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!d_in || !d_testvs || !d_out || n_luts == 0) LUX_FAIL(e, LUX_ERR_INVALID, "lux_pbs_batch_device: null/empty argument");
    return lux_pbs_device_grouped(e, d_in, d_testvs, d_lut_idx, 1, B, d_out, stream ? (cudaStream_t)stream : e->stream);
}

// ---- PBS that allocates nothing (include/lux_tfhe.h, lux_pbs_batch_device_static; no counterpart in the reference).
// The scratch of torus_pbs_device — the key switch's digit matrix and the ciphertexts
// between key switch and blind rotation — lives in this file-scope device array instead of cudaMalloc'ed engine scratch.
// In host code `g_static_scratch` is only a stand-in for the symbol: its device address is taken once per device, with that
// device current, by cudaGetSymbolAddress below, and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_static_scratch[LUX_STATIC_SCRATCH_BYTES];
static unsigned char *g_static_scratch_ptr[64] = {};
static std::mutex g_static_scratch_mtx;

static int static_scratch_ptr(lux_engine *e, unsigned char **out) {
    if (e->device < 0 || e->device >= 64) LUX_FAIL(e, LUX_ERR_INVALID, "device ordinal %d outside the static scratch table", e->device);
    std::lock_guard<std::mutex> lk(g_static_scratch_mtx);
    if (!g_static_scratch_ptr[e->device]) {
        void *p = nullptr;
        cudaError_t st = cudaGetSymbolAddress(&p, g_static_scratch);
        if (st != cudaSuccess || !p) LUX_FAIL(e, LUX_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on GPU %d: %s", e->device, cudaGetErrorString(st));
        CUDA_TRY(e, cudaFuncSetAttribute(ksmma::lwe_keyswitch_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ksmma::SMEM_BYTES));
        g_static_scratch_ptr[e->device] = (unsigned char *)p;
    }
    *out = g_static_scratch_ptr[e->device];
    return LUX_OK;
}

int lux_pbs_batch_device_static(lux_engine *e, const uint64_t *d_in, const uint64_t *d_testvs, uint32_t n_luts,
                                const uint32_t *d_lut_idx, size_t B, uint64_t *d_out, void *stream) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!d_in || !d_testvs || !d_out || n_luts == 0) LUX_FAIL(e, LUX_ERR_INVALID, "lux_pbs_batch_device_static: null/empty argument");
    if (!lux_is_torus(e)) LUX_FAIL(e, LUX_ERR_INVALID, "lux_pbs_batch_device_static: LUX_ARITH_TORUS64_FFT engines only");
    if (!e->has_bsk) LUX_FAIL(e, LUX_ERR_NO_KEY, "bootstrap key not loaded");
    if (!e->has_lksk || !e->d_kskT) LUX_FAIL(e, LUX_ERR_NO_KEY, "LWE key switch key not loaded");
    if (B == 0) return LUX_OK;
    const lux_params &p = e->p;
    const bool ks_first = p.order == LUX_ORDER_LWEKS_PBS;
    const size_t kN = (size_t)p.k * p.N, K = kN * p.ell_ks, mid_words = ks_first ? (size_t)p.n + 1 : kN + 1;
    auto dig_bytes = [&](size_t b) { return ((b + ksmma::BM - 1) / ksmma::BM * ksmma::BM * K + 255) / 256 * 256; };
    const size_t need = dig_bytes(B) + sizeof(u64) * B * mid_words;
    if (B > 0x7fffffffu || need > LUX_STATIC_SCRATCH_BYTES) {
        size_t bmax = 0;
        while (dig_bytes(bmax + 1) + sizeof(u64) * (bmax + 1) * mid_words <= LUX_STATIC_SCRATCH_BYTES) bmax++;
        LUX_FAIL(e, LUX_ERR_INVALID, "lux_pbs_batch_device_static: a batch of %zu needs %zu bytes of scratch, the static array holds %u "
                 "(largest batch at this parameter set: %zu); split the batch or call lux_pbs_batch_device", B, need,
                 (unsigned)LUX_STATIC_SCRATCH_BYTES, bmax);
    }
    CUDA_TRY(e, cudaSetDevice(e->device));
    unsigned char *base = nullptr;
    int rc = static_scratch_ptr(e, &base);
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : e->stream;
    int8_t *d_dig = (int8_t *)base;                       // [Mpad][K]; ks_digits_kernel writes every row, the padding rows as zero
    u64 *d_mid = (u64 *)(base + dig_bytes(B));            // [B][mid_words]; every word written before it is read
    const u32 mpad = (u32)((B + ksmma::BM - 1) / ksmma::BM * ksmma::BM);
    auto key_switch = [&](const u64 *src, u64 *dst) -> int {
        ksmma::ks_digits_kernel<<<e->sm_count * 8, 256, 0, s>>>(src, d_dig, (u32)B, mpad, (u32)kN, p.ell_ks, p.logB_ks, e->L_ks, 1u);
        e->launches++;
        CUDA_TRY(e, cudaGetLastError());
        ksmma::Args ka{d_dig, e->d_kskT, src, dst, (u32)B, p.n, (u32)kN, (u32)K, 1u};
        dim3 grid(e->ks_npad / ksmma::BN, mpad / ksmma::BM);
        ksmma::lwe_keyswitch_mma_kernel<<<grid, ksmma::THREADS, ksmma::SMEM_BYTES, s>>>(ka);
        e->launches++;
        CUDA_TRY(e, cudaGetLastError());
        return LUX_OK;
    };
    if (ks_first && (rc = key_switch(d_in, d_mid))) return rc;
    fft64::Args a = make_fft_args(e, ks_first ? d_mid : d_in, d_testvs, d_lut_idx, 1, B, ks_first ? d_out : d_mid);
    if ((rc = lux_fft_launch_br(e, a, s))) return rc;
    if (!ks_first && (rc = key_switch(d_mid, d_out))) return rc;
    return LUX_OK;
}

int lux_pbs_batch(lux_engine *e, const uint64_t *lwe_in, const uint64_t *testvs, uint32_t n_luts,
                  const uint32_t *lut_idx, size_t B, uint64_t *lwe_out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!lwe_in || !testvs || !lwe_out || n_luts == 0) LUX_FAIL(e, LUX_ERR_INVALID, "lux_pbs_batch: null/empty argument");
    if (B == 0) return LUX_OK;
    if (lut_idx)
        for (size_t b = 0; b < B; b++)
            if (lut_idx[b] >= n_luts) LUX_FAIL(e, LUX_ERR_INVALID, "lut_idx[%zu] = %u >= n_luts = %u", b, lut_idx[b], n_luts);
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t w = lux_ct_words(e), N = e->p.N;
    int rc;
    // Chunked, double-buffered pipeline: H2D of chunk c+1 and D2H of chunk c-1 run on their own streams while chunk c is
    // bootstrapped; pageable caller memory goes through pinned staging owned by the engine (the memcpy of the next chunk
    // overlaps the GPU), pinned caller memory is copied from / to directly.  A chunk is a whole number of kernel waves.
    if ((rc = host_pipe_init(e))) return rc;
    HostPipe &hp = *e->pipe;
    const size_t wave = lux_wave_ciphertexts(e, B);
    size_t chunk = B;
    if (B >= 4 * wave) chunk = 2 * wave;
    else if (B >= 2 * wave) chunk = wave;
    const size_t n_chunks = (B + chunk - 1) / chunk;
    const size_t chunk_bytes = sizeof(u64) * chunk * w;
    if ((rc = ensure_scratch(e, 1, 2 * chunk_bytes))) return rc;
    if ((rc = ensure_scratch(e, 2, 2 * chunk_bytes))) return rc;
    if ((rc = ensure_scratch(e, 3, sizeof(u64) * n_luts * N + sizeof(u32) * B))) return rc;
    const bool in_pinned = host_is_pinned(lwe_in), out_pinned = host_is_pinned(lwe_out);
    if ((!in_pinned || !out_pinned) && (rc = host_pipe_reserve(e, chunk_bytes))) return rc;
    u64 *d_tv = (u64 *)e->scratch[3];
    u32 *d_idx = lut_idx ? (u32 *)(d_tv + (size_t)n_luts * N) : nullptr;
    CUDA_TRY(e, cudaMemcpyAsync(d_tv, testvs, sizeof(u64) * n_luts * N, cudaMemcpyHostToDevice, e->stream));
    if (lut_idx) CUDA_TRY(e, cudaMemcpyAsync(d_idx, lut_idx, sizeof(u32) * B, cudaMemcpyHostToDevice, e->stream));
    auto retire = [&](size_t c) -> int {   // chunk c's results reach the caller's buffer
        const int slot = (int)(c & 1);
        CUDA_TRY(e, cudaEventSynchronize(hp.ev_out[slot]));
        if (!out_pinned) {
            const size_t c0 = c * chunk, nb = std::min(chunk, B - c0);
            memcpy(lwe_out + c0 * w, hp.h_out[slot], sizeof(u64) * nb * w);
        }
        return LUX_OK;
    };
    for (size_t c = 0; c < n_chunks; c++) {
        const int slot = (int)(c & 1);
        const size_t c0 = c * chunk, nb = std::min(chunk, B - c0), bytes = sizeof(u64) * nb * w;
        if (c >= 2 && (rc = retire(c - 2))) return rc;   // also frees this slot's device and staging buffers
        u64 *d_in = (u64 *)((unsigned char *)e->scratch[1] + slot * chunk_bytes);
        u64 *d_out = (u64 *)((unsigned char *)e->scratch[2] + slot * chunk_bytes);
        const uint64_t *src = lwe_in + c0 * w;
        if (!in_pinned) {
            memcpy(hp.h_in[slot], src, bytes);
            src = (const uint64_t *)hp.h_in[slot];
        }
        CUDA_TRY(e, cudaMemcpyAsync(d_in, src, bytes, cudaMemcpyHostToDevice, hp.s_in));
        CUDA_TRY(e, cudaEventRecord(hp.ev_in[slot], hp.s_in));
        CUDA_TRY(e, cudaStreamWaitEvent(e->stream, hp.ev_in[slot], 0));
        if ((rc = lux_pbs_device_grouped(e, d_in, d_tv, d_idx ? d_idx + c0 : nullptr, 1, nb, d_out, e->stream))) return rc;
        CUDA_TRY(e, cudaEventRecord(hp.ev_k[slot], e->stream));
        CUDA_TRY(e, cudaStreamWaitEvent(hp.s_out, hp.ev_k[slot], 0));
        uint64_t *dst = out_pinned ? lwe_out + c0 * w : (uint64_t *)hp.h_out[slot];
        CUDA_TRY(e, cudaMemcpyAsync(dst, d_out, bytes, cudaMemcpyDeviceToHost, hp.s_out));
        CUDA_TRY(e, cudaEventRecord(hp.ev_out[slot], hp.s_out));
    }
    for (size_t c = n_chunks >= 2 ? n_chunks - 2 : 0; c < n_chunks; c++)
        if ((rc = retire(c))) return rc;
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

// ---- boolean gates
static int gate_coeffs(int gate, int *coef, u64 *konst) {
    const u64 h = GL_Q >> 3;                       // q/8 = delta/2 (delta = q >> 2 is even)
    const u64 h3 = gl_canon(gl_add(gl_canon(gl_add(h, h)), h));
    switch (gate) {
        case LUX_GATE_NAND: *coef = -1; *konst = h3; return 1;                       // {3h, h, -h}
        case LUX_GATE_AND: *coef = 1; *konst = gl_neg(h3); return 1;                 // {-3h, -h, h}
        case LUX_GATE_OR: *coef = 1; *konst = gl_neg(h); return 1;                   // {-h, h, 3h}
        case LUX_GATE_NOR: *coef = -1; *konst = h; return 1;                         // {h, -h, -3h}
        case LUX_GATE_XOR: *coef = 2; *konst = gl_neg(gl_canon(gl_add(h, h))); return 1;   // {-2h, 2h, -2h}
        case LUX_GATE_XNOR: *coef = -2; *konst = gl_canon(gl_add(h, h)); return 1;         // {2h, -2h, 2h}
        default: return 0;
    }
}
int lux_gate_batch_device(lux_engine *e, int gate, const uint64_t *d_a, const uint64_t *d_b, size_t B, uint64_t *d_out,
                          void *stream) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!d_a || !d_b || !d_out) LUX_FAIL(e, LUX_ERR_INVALID, "lux_gate_batch_device: null argument");
    int coef;
    u64 konst;
    if (!gate_coeffs(gate, &coef, &konst)) LUX_FAIL(e, LUX_ERR_INVALID, "unknown gate %d", gate);
    if (B == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : e->stream;
    const size_t w = lux_ct_words(e), N = e->p.N, total = B * w;
    int rc;
    if ((rc = ensure_scratch(e, 6, sizeof(u64) * total))) return rc;
    if ((rc = ensure_scratch(e, 7, sizeof(u64) * N))) return rc;
    u64 *d_x = (u64 *)e->scratch[6], *d_tv = (u64 *)e->scratch[7];
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, (size_t)e->sm_count * 16);
    gate_prepare_kernel<<<grid, 256, 0, s>>>(d_a, d_b, d_x, (u32)w, total, coef, konst);
    fill_kernel<<<(unsigned)((N + 255) / 256), 256, 0, s>>>(d_tv, GL_Q >> 3, N);
    e->launches += 2;
    CUDA_TRY(e, cudaGetLastError());
    return lux_pbs_device_grouped(e, d_x, d_tv, nullptr, 1, B, d_out, s, GL_Q >> 3);
}
int lux_gate_batch(lux_engine *e, int gate, const uint64_t *a, const uint64_t *b, size_t B, uint64_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!a || !b || !out) LUX_FAIL(e, LUX_ERR_INVALID, "lux_gate_batch: null argument");
    if (B == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t bytes = sizeof(u64) * B * lux_ct_words(e);
    int rc;
    if ((rc = ensure_scratch(e, 1, bytes))) return rc;
    if ((rc = ensure_scratch(e, 2, bytes))) return rc;
    if ((rc = ensure_scratch(e, 3, bytes))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], a, bytes, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[3], b, bytes, cudaMemcpyHostToDevice, e->stream));
    if ((rc = lux_gate_batch_device(e, gate, (const u64 *)e->scratch[1], (const u64 *)e->scratch[3], B, (u64 *)e->scratch[2], nullptr)))
        return rc;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], bytes, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}
// NOT needs no bootstrap: delta - x  (1 -> 0, 0 -> 1), host-side helper on [B][ct_words] words
int lux_gate_not(const lux_engine *e, const uint64_t *a, size_t B, uint64_t *out) {
    if (!e || !a || !out) return LUX_ERR_INVALID;
    const size_t w = lux_ct_words(e);
    for (size_t i = 0; i < B * w; i++) {
        u64 v = gl_neg(a[i]);
        if (i % w == w - 1) v = gl_canon(gl_add(v, GL_Q >> 2));
        out[i] = v;
    }
    return LUX_OK;
}

// get_testv (crypto/mod.rs:17-27): blocks of lut[i]*delta, then left_shift(block/2) = times X^-(block/2)
int lux_make_testv(uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out) {
    if (!lut || !out || p == 0 || N == 0 || N % p) return LUX_ERR_INVALID;
    const u32 bs = N / p, sh = bs / 2;
    for (u32 i = 0; i < N; i++) {
        const u32 src = i + sh;
        if (src < N)
            out[i] = gl_mul(lut[src / bs] % GL_Q, delta % GL_Q);
        else
            out[i] = gl_neg(gl_mul(lut[(src - N) / bs] % GL_Q, delta % GL_Q));
    }
    return LUX_OK;
}
uint64_t lux_delta(uint32_t plaintext_space) { return GL_Q >> ilog2u(plaintext_space); }
uint64_t lux_delta_arith(uint32_t arith, uint32_t plaintext_space) {
    if (arith != LUX_ARITH_TORUS64_FFT) return lux_delta(plaintext_space);
    const u32 l = ilog2u(plaintext_space);
    return l == 0 ? 0 : 1ull << (64 - l);
}
int lux_make_testv_arith(uint32_t arith, uint32_t N, uint32_t p, uint64_t delta, const uint64_t *lut, uint64_t *out) {
    if (arith != LUX_ARITH_TORUS64_FFT) return lux_make_testv(N, p, delta, lut, out);
    if (!lut || !out || p == 0 || N == 0 || N % p) return LUX_ERR_INVALID;
    const u32 bs = N / p, sh = bs / 2;
    for (u32 i = 0; i < N; i++) {
        const u32 src = i + sh;
        out[i] = src < N ? lut[src / bs] * delta : (u64)0 - lut[(src - N) / bs] * delta;
    }
    return LUX_OK;
}

// ---- pieces
static int ntt_host(lux_engine *e, const u64 *in, size_t n_polys, u64 *out, bool inverse) {
    if (!e) return LUX_ERR_INVALID;
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (n_polys == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t bytes = sizeof(u64) * n_polys * e->p.N;
    int rc = ensure_scratch(e, 1, bytes);
    if (rc) return rc;
    u64 *d = (u64 *)e->scratch[1];
    CUDA_TRY(e, cudaMemcpyAsync(d, in, bytes, cudaMemcpyHostToDevice, e->stream));
    if ((rc = launch_ntt(e, d, d, n_polys, inverse, inverse ? e->ninv : 1, e->stream))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(out, d, bytes, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}
int lux_ntt_forward(lux_engine *e, const uint64_t *in, size_t n_polys, uint64_t *out) { return ntt_host(e, in, n_polys, out, false); }
int lux_ntt_backward(lux_engine *e, const uint64_t *in, size_t n_polys, uint64_t *out) { return ntt_host(e, in, n_polys, out, true); }

int lux_ntt_any(lux_engine *e, uint32_t N, int inverse, const uint64_t *in, size_t n_polys, uint64_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (N < 8 || N > 2048 || (N & (N - 1))) LUX_FAIL(e, LUX_ERR_INVALID, "lux_ntt_any: N must be a power of two in [8,2048]");
    if (n_polys == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    u64 ninv;
    {
        std::vector<u64> r, ri;
        lux_host_ntt_tables(N, r, ri, ninv);
        if (!e->any_tables.count(N)) {
            u64 *d = nullptr;
            CUDA_TRY(e, cudaMalloc(&d, sizeof(u64) * 2 * N));
            CUDA_TRY(e, cudaMemcpy(d, r.data(), sizeof(u64) * N, cudaMemcpyHostToDevice));
            CUDA_TRY(e, cudaMemcpy(d + N, ri.data(), sizeof(u64) * N, cudaMemcpyHostToDevice));
            e->any_tables[N] = d;
        }
    }
    const u64 *tw = e->any_tables[N] + (inverse ? N : 0);
    const size_t bytes = sizeof(u64) * n_polys * N;
    int rc = ensure_scratch(e, 1, bytes);
    if (rc) return rc;
    u64 *d = (u64 *)e->scratch[1];
    CUDA_TRY(e, cudaMemcpyAsync(d, in, bytes, cudaMemcpyHostToDevice, e->stream));
    ntt_any_kernel<<<(unsigned)n_polys, N / 2, sizeof(u64) * N, e->stream>>>(d, d, tw, inverse ? ninv : 1, N, inverse ? 1 : 0);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(out, d, bytes, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_decompose(lux_engine *e, const uint64_t *in, size_t count, uint32_t logB, uint64_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (logB < 1 || logB > 62) LUX_FAIL(e, LUX_ERR_INVALID, "logB out of range");
    if (count == 0) return LUX_OK;
    const u32 L = (64 + logB - 1) / logB;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * count))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u64) * count * L))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], in, sizeof(u64) * count, cudaMemcpyHostToDevice, e->stream));
    decompose_kernel<<<e->sm_count * 4, 256, 0, e->stream>>>((const u64 *)e->scratch[1], (u64 *)e->scratch[2], count, logB, L);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], sizeof(u64) * count * L, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_mod_switch(lux_engine *e, const uint64_t *in, size_t count, uint32_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (count == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * count))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u32) * count))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], in, sizeof(u64) * count, cudaMemcpyHostToDevice, e->stream));
    mod_switch_kernel<<<e->sm_count * 4, 256, 0, e->stream>>>((const u64 *)e->scratch[1], (u32 *)e->scratch[2], count, e->logN);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], sizeof(u32) * count, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_lwe_keyswitch(lux_engine *e, const uint64_t *in, size_t B, uint64_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (e->p.order == LUX_ORDER_PBS_GLWEKS) LUX_FAIL(e, LUX_ERR_INVALID, "engine has no LWE key switch (order 0)");
    if (!e->has_lksk) LUX_FAIL(e, LUX_ERR_NO_KEY, "LWE key switch key not loaded");
    if (B == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t wi = (size_t)e->p.k * e->p.N + 1, wo = (size_t)e->p.n + 1;
    int rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * B * wi))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u64) * B * wo))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], in, sizeof(u64) * B * wi, cudaMemcpyHostToDevice, e->stream));
    if ((rc = launch_ks(e, (const u64 *)e->scratch[1], (u64 *)e->scratch[2], B, e->stream))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], sizeof(u64) * B * wo, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_blind_rotate(lux_engine *e, const uint64_t *lwe, const uint64_t *testvs, uint32_t n_luts, const uint32_t *lut_idx,
                     size_t B, uint32_t steps, uint64_t *acc_out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!lwe || !testvs || !acc_out || n_luts == 0) LUX_FAIL(e, LUX_ERR_INVALID, "null/empty argument");
    if (!e->has_bsk) LUX_FAIL(e, LUX_ERR_NO_KEY, "bootstrap key not loaded");
    if (steps > e->p.n) LUX_FAIL(e, LUX_ERR_INVALID, "steps %u > n %u", steps, e->p.n);
    if (B == 0) return LUX_OK;
    if (lut_idx)
        for (size_t b = 0; b < B; b++)
            if (lut_idx[b] >= n_luts) LUX_FAIL(e, LUX_ERR_INVALID, "lut_idx out of range");
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t wi = (size_t)e->p.n + 1, N = e->p.N, wo = 2 * N;
    int rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * B * wi))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u64) * B * wo))) return rc;
    if ((rc = ensure_scratch(e, 3, sizeof(u64) * n_luts * N + sizeof(u32) * B))) return rc;
    u64 *d_tv = (u64 *)e->scratch[3];
    u32 *d_idx = lut_idx ? (u32 *)(d_tv + (size_t)n_luts * N) : nullptr;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], lwe, sizeof(u64) * B * wi, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_tv, testvs, sizeof(u64) * n_luts * N, cudaMemcpyHostToDevice, e->stream));
    if (lut_idx) CUDA_TRY(e, cudaMemcpyAsync(d_idx, lut_idx, sizeof(u32) * B, cudaMemcpyHostToDevice, e->stream));
    if (lux_is_torus(e)) {
        fft64::Args fa = make_fft_args(e, (const u64 *)e->scratch[1], d_tv, d_idx, 1, B, (u64 *)e->scratch[2]);
        fa.steps = steps;
        fa.out_mode = 2;
        if ((rc = lux_fft_launch_br(e, fa, e->stream))) return rc;
    } else {
    BRArgs a = make_br_args(e, (const u64 *)e->scratch[1], d_tv, d_idx, B, (u64 *)e->scratch[2]);
    a.steps = steps;
    a.out_mode = 2;
    if ((rc = launch_br(e, a, e->stream))) return rc;
    }
    CUDA_TRY(e, cudaMemcpyAsync(acc_out, e->scratch[2], sizeof(u64) * B * wo, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_fft(lux_engine *e, const double *in, size_t n_polys, int inverse, double *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!lux_is_torus(e)) LUX_FAIL(e, LUX_ERR_INVALID, "lux_fft: engine is not LUX_ARITH_TORUS64_FFT");
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (n_polys == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t bytes = sizeof(double) * n_polys * e->p.N;   // N/2 complex points
    int rc;
    if ((rc = ensure_scratch(e, 1, bytes))) return rc;
    if ((rc = ensure_scratch(e, 2, bytes))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], in, bytes, cudaMemcpyHostToDevice, e->stream));
    if ((rc = lux_fft_probe(e, e->scratch[1], e->scratch[2], n_polys, inverse, e->stream))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], bytes, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_torus_decompose(lux_engine *e, const uint64_t *in, size_t count, int64_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!lux_is_torus(e)) LUX_FAIL(e, LUX_ERR_INVALID, "lux_torus_decompose: engine is not LUX_ARITH_TORUS64_FFT");
    if (!in || !out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (count == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * count))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u64) * count))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], in, sizeof(u64) * count, cudaMemcpyHostToDevice, e->stream));
    if ((rc = lux_fft_digits(e, (const u64 *)e->scratch[1], (long long *)e->scratch[2], count, e->p.logB, e->stream))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], sizeof(u64) * count, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_export_bsk_fourier(lux_engine *e, double *out, size_t n_doubles) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!lux_is_torus(e)) LUX_FAIL(e, LUX_ERR_INVALID, "lux_export_bsk_fourier: engine is not LUX_ARITH_TORUS64_FFT");
    if (!e->has_bsk) LUX_FAIL(e, LUX_ERR_NO_KEY, "bootstrap key not loaded");
    if (!out || n_doubles != e->bsk_words) LUX_FAIL(e, LUX_ERR_INVALID, "lux_export_bsk_fourier: expected %zu doubles", e->bsk_words);
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaMemcpy(out, e->d_bsk, sizeof(double) * n_doubles, cudaMemcpyDeviceToHost));
    return LUX_OK;
}

int lux_external_product(lux_engine *e, const uint64_t *ggsw, int domain, const uint64_t *glwe_in, size_t count, uint64_t *glwe_out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (lux_is_torus(e)) LUX_FAIL(e, LUX_ERR_INVALID, "lux_external_product: not available for the torus arithmetic (use lux_blind_rotate with steps)");
    if (!ggsw || !glwe_in || !glwe_out) LUX_FAIL(e, LUX_ERR_INVALID, "null pointer");
    if (count == 0) return LUX_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const size_t gw = ggsw_words(e->p), N = e->p.N;
    int rc;
    if ((rc = ensure_scratch(e, 4, sizeof(u64) * count * gw))) return rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * count * 2 * N))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u64) * count * 2 * N))) return rc;
    if ((rc = load_ggsw_array(e, (u64 *)e->scratch[4], ggsw, count * gw, count * gw, domain, "lux_external_product"))) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(e->scratch[1], glwe_in, sizeof(u64) * count * 2 * N, cudaMemcpyHostToDevice, e->stream));
    BRArgs a = make_br_args(e, nullptr, nullptr, nullptr, count, (u64 *)e->scratch[2]);
    a.bsk = (const u64 *)e->scratch[4];
    switch (e->logN) {
        case 9: rc = lux_launch_xp_9(e, a, (const u64 *)e->scratch[1], gw, e->stream); break;
        case 10: rc = lux_launch_xp_10(e, a, (const u64 *)e->scratch[1], gw, e->stream); break;
        case 11: rc = lux_launch_xp_11(e, a, (const u64 *)e->scratch[1], gw, e->stream); break;
        case 12: rc = lux_launch_xp_12(e, a, (const u64 *)e->scratch[1], gw, e->stream); break;
        default: LUX_FAIL(e, LUX_ERR_INVALID, "unsupported N");
    }
    if (rc) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(glwe_out, e->scratch[2], sizeof(u64) * count * 2 * N, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

int lux_debug_field_probe(lux_engine *e, const uint64_t *a, const uint64_t *b, size_t count, uint64_t *out) {
    LUX_LOCK(e);
    if (!e) return LUX_ERR_INVALID;
    if (!a || !b || !out || count == 0) LUX_FAIL(e, LUX_ERR_INVALID, "null/empty argument");
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc;
    if ((rc = ensure_scratch(e, 1, sizeof(u64) * count * 2))) return rc;
    if ((rc = ensure_scratch(e, 2, sizeof(u64) * count * 7))) return rc;
    u64 *d_a = (u64 *)e->scratch[1], *d_b = d_a + count;
    CUDA_TRY(e, cudaMemcpyAsync(d_a, a, sizeof(u64) * count, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_b, b, sizeof(u64) * count, cudaMemcpyHostToDevice, e->stream));
    field_probe_kernel<<<e->sm_count, 256, 0, e->stream>>>(d_a, d_b, (u64 *)e->scratch[2], count);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(out, e->scratch[2], sizeof(u64) * count * 7, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return LUX_OK;
}

// out[0] = IMAD.WIDE.U32 per second, out[1] = 32-bit IADD per second, out[2] = Goldilocks butterflies per second
// (register-only, whole GPU), measured with CUDA events.
int lux_debug_int_peak(lux_engine *e, double out[3]) {
    LUX_LOCK(e);
    if (!e || !out) return LUX_ERR_INVALID;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc;
    if ((rc = ensure_scratch(e, 5, 64))) return rc;
    const u32 iters = 4096, grid = (u32)e->sm_count * 8;
    for (int kind = 0; kind < 3; kind++) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; rep++) {
            CUDA_TRY(e, cudaEventRecord(e->ev[0], e->stream));
            if (kind == 0) int_peak_kernel<0><<<grid, 256, 0, e->stream>>>((u64 *)e->scratch[5], iters, 1);
            if (kind == 1) int_peak_kernel<1><<<grid, 256, 0, e->stream>>>((u64 *)e->scratch[5], iters, 1);
            if (kind == 2) int_peak_kernel<2><<<grid, 256, 0, e->stream>>>((u64 *)e->scratch[5], iters, 1);
            e->launches++;
            CUDA_TRY(e, cudaEventRecord(e->ev[1], e->stream));
            CUDA_TRY(e, cudaEventSynchronize(e->ev[1]));
            float ms;
            CUDA_TRY(e, cudaEventElapsedTime(&ms, e->ev[0], e->ev[1]));
            if (rep && ms < best) best = ms;
        }
        const double per_thread = (double)iters * 8 * (kind == 0 ? 8 : kind == 1 ? 16 : 4);
        out[kind] = per_thread * grid * 256 / (best * 1e-3);
    }
    return LUX_OK;
}

// out[0] = DFMA lane-operations per second (register-only, whole GPU), measured with CUDA events.
int lux_debug_fp64_peak(lux_engine *e, double out[1]) {
    LUX_LOCK(e);
    if (!e || !out) return LUX_ERR_INVALID;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int rc;
    if ((rc = ensure_scratch(e, 5, 64))) return rc;
    const u32 iters = 2048, grid = (u32)e->sm_count * 8;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        CUDA_TRY(e, cudaEventRecord(e->ev[0], e->stream));
        fp64_peak_kernel<<<grid, 256, 0, e->stream>>>((double *)e->scratch[5], iters, 1.0);
        e->launches++;
        CUDA_TRY(e, cudaEventRecord(e->ev[1], e->stream));
        CUDA_TRY(e, cudaEventSynchronize(e->ev[1]));
        float ms;
        CUDA_TRY(e, cudaEventElapsedTime(&ms, e->ev[0], e->ev[1]));
        if (rep && ms < best) best = ms;
    }
    out[0] = (double)iters * 64 * grid * 256 / (best * 1e-3);
    return LUX_OK;
}

}  // extern "C"
