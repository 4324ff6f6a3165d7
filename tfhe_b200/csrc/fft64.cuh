// fft64.cuh — the blind rotation on the 2^64 torus with a double-precision negacyclic FFT (sm_100a).
//
// This is the second arithmetic of the engine (lux_params.arith = LUX_ARITH_TORUS64_FFT): the one the reference's
// named shortint set PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 (sdk/relayer/src/tfhe.ts:13-28) is defined on.
// The algorithm — SignedDecomposer digits, N/2-point complex transform with the negacyclic twist merged into the
// twiddles, external product, CMUX loop — and every rounding it performs are specified by oracle/torus.c, which
// this file reproduces word for word: each floating-point operation below is an explicit round-to-nearest
// intrinsic (__fma_rn, __dadd_rn, __dmul_rn) in the oracle's order, so nothing is contracted or re-associated.
//
// Why FP64 (profiles/r02_arith_ubench.txt, measured on B200): the FP64 pipe issues 64 lanes/SM/clk, a complex
// radix-2 butterfly is 6 DFMA -> 3.07 T butterflies/s against 0.545 T/s for the 64-bit Goldilocks butterfly and
// 4.05 T/s for a 31-bit Harvey butterfly (which needs twice the transforms and 2.2x the butterflies per transform).
//
// Work split ("warp per polynomial").  N = 2048 -> M = 1024 complex points = 32 x 32.  ONE WARP owns one polynomial
// of a ciphertext and keeps the whole transform in registers, 32 points per lane:
//   column layout  x[i] = z[lane + 32 i]      pass A: the 5 stages with strides 512..32 pair registers; the twiddle
//                                             of a butterfly depends on i only -> uniform over the warp, read from
//                                             the kernel-parameter constant bank
//   row layout     y[i] = z[32 lane + i]      pass B: strides 16..1; twiddles per lane, from a shared-memory table
//                                             that holds EVEN tree nodes only (tw[2v+1] = i tw[2v], exact)
// with ONE transpose through shared memory between the passes (padded rows: 33 x 16 B, conflict-free both ways).
// A transform therefore crosses shared memory once instead of once per 3-4 stages: shared-memory bandwidth
// (128 B/clk/SM), not the FP64 pipe, is what a 16-byte-per-point transform runs out of first.
// A team = 2 warps = one ciphertext (k = 1): warp r decomposes and transforms input polynomial r, the two warps
// swap their spectra through shared memory, and warp r accumulates, inverse-transforms and updates OUTPUT
// polynomial r.  Warp r only ever touches accumulator polynomial r, so the CMUX needs two 64-thread barriers (around
// the swap) and no other cross-warp synchronisation.
// The bootstrap key streams through a ring of shared-memory stages filled by a dedicated producer warp with
// cp.async.bulk (TMA) + mbarriers and is shared by the TEAMS ciphertexts of the CTA.
#pragma once
#include "tfhe_kernels.cuh"
#include <type_traits>

namespace fft64 {

constexpr double MAGIC = 6755399441055744.0;   // 1.5 * 2^52
constexpr double TWO51 = 2251799813685248.0;   // 2^51
constexpr int LOGN = 11, N = 1 << LOGN, M = N / 2;   // ring degree of the named set
constexpr int ROWP = 33;                       // padded row length of the transpose buffer (in points)
constexpr int BUF_POINTS = 31 * ROWP + 32;     // one warp's transpose / swap buffer (the last row needs no padding)
constexpr int TWB_ENTRIES = 512;               // pass-B table: 32 (m = 32) + 32 + 64 + 128 + 256 even nodes

struct Args {
    u32 n, steps, logB, B, out_mode, lut_div;
    const double2 *bsk;     // [n][32 rows i][4 polys (j, r)][32 lanes], Fourier domain, scaled, sign folded
    const u64 *lwe;         // [B][n+1] under the small key
    const u64 *testv;       // [n_luts][N]
    const u32 *lut_idx;     // [ceil(B / lut_div)] or nullptr
    u64 *out;               // out_mode 0: [B][N+1] sample extract; 2: [B][2N] raw accumulator
    const double2 *twB;     // [TWB_ENTRIES] pass-B table (global; copied to shared memory)
    u64 body_add;
    u32 stagger_ns;         // GROUPS > 1: start delay of group g is g * stagger_ns
    u64 *trace;             // phase-cycle counters (FFT64_TRACE_BUILD) or nullptr
    double2 twA[32];        // tree nodes 1..31: the warp-uniform twiddles of pass A
    u32 ct0;                // first ciphertext of THIS launch; B is one past its last.  A batch may be cut into a run of whole
                            // waves and a tail with fewer ciphertexts per CTA (br_fft.cu lux_fft_launch_br): only the kernels a
                            // tail can take (latency kernel, one to three ciphertexts per CTA) read it; the four-per-CTA kernel
                            // always starts at ciphertext 0 and is compiled exactly as before (last member for the same reason)
};

// ------------------------------------------------------------------------------------------ scalars
// oracle/torus.c tor_from_double
__device__ __forceinline__ u64 from_double(double y) {
    const double t = __dadd_rn(y, MAGIC);
    const double r = __dsub_rn(t, MAGIC);
    const double f = __dsub_rn(y, r);
    const double h = __fma_rn(f, TWO51, MAGIC);
    const long long v = __double_as_longlong(h) - 0x4338000000000000LL;
    return (u64)v << 13;
}
// exact int32 -> double without the conversion pipe: 2^52 + 2^31 + d sits in the mantissa
__device__ __forceinline__ double int_to_double(int d) {
    return __dsub_rn(__hiloint2double(0x43300000, d ^ 0x80000000), 4503601774854144.0);
}
// oracle/torus.c tor_decompose for ONE level of logB <= 31 bits: only the high word of x matters
__device__ __forceinline__ int digit1(u64 x, u32 logB) {
    const u32 hi = (u32)(x >> 32);
    const u32 res = (((hi >> (31 - logB)) + 1u) >> 1) & ((1u << logB) - 1u);
    const u32 carry = ((res - 1u) & res) >> (logB - 1);
    return (int)res - (int)(carry << logB);
}
// the same digit from the high word of x, as the double the transform consumes: with half = 2^(31-logB),
// (int)(hi - half) >> (32 - logB) is the balanced digit minus one (the wrap of the 32-bit word is the reduction mod B,
// the arithmetic shift the sign extension; tor_decompose maps B/2 to +B/2, hence the shift by one); the +1 is folded into
// the exact int -> double constant.  Three integer instructions and one DADD instead of ten and one.
__device__ __forceinline__ double digit1_double(u32 hi, u32 half, u32 sh) {
    const int dm1 = (int)(hi - half) >> sh;
    return __dsub_rn(__hiloint2double(0x43300000, dm1 ^ 0x80000000), 4503601774854143.0);   // 2^52 + 2^31 - 1
}
// the same digit when the rounding offset is already IN the word: hi_off = hi - half + 2^31 (the caller folds that constant into
// the subtraction that produces hi, where the carry chain has a free operand).  The +2^31 turns the arithmetic shift into a
// logical one, (hi_off >> sh) = digit - 1 + 2^(31 - sh) exactly, and the magic constant takes the offset back out: one shift and
// one DADD per coefficient.  magic = 2^52 + 2^(31 - sh) - 1.
__device__ __forceinline__ double digit1_double_off(u32 hi_off, u32 sh, double magic) {
    return __dsub_rn(__hiloint2double(0x43300000, (int)(hi_off >> sh)), magic);
}
// round(x 2N / 2^64) in [0, 2N]      crypto/lwe.rs:28-34
__device__ __forceinline__ u32 mod_switch(u64 x) {
    const u64 a = x >> (64 - LOGN - 2);
    return (u32)((a >> 1) + (a & 1));
}
// coefficient p of X^sh * poly, wrapping negation
__device__ __forceinline__ u64 rotated(const u64 *poly, u32 p, u32 sh) {
    const u32 i2 = (p - sh) & (2 * N - 1);
    const u64 v = poly[i2 & (N - 1)];
    return (i2 & N) ? (u64)0 - v : v;
}

// ------------------------------------------------------------------------------------------ butterflies
__device__ __forceinline__ void ct_bfly(double2 &u, double2 &v, double wr, double wi) {
    const double tr = __fma_rn(-wi, v.y, u.x);
    const double nr = __fma_rn(wr, v.x, tr);
    const double ti = __fma_rn(wi, v.x, u.y);
    const double ni = __fma_rn(wr, v.y, ti);
    v.x = __fma_rn(2.0, u.x, -nr);
    v.y = __fma_rn(2.0, u.y, -ni);
    u.x = nr;
    u.y = ni;
}
__device__ __forceinline__ void gs_bfly(double2 &u, double2 &v, double wr, double wi) {
    const double sr = __dadd_rn(u.x, v.x), si = __dadd_rn(u.y, v.y);
    const double dr = __dsub_rn(u.x, v.x), di = __dsub_rn(u.y, v.y);
    const double m0 = __dmul_rn(di, wi), m1 = __dmul_rn(dr, wi);
    u.x = sr;
    u.y = si;
    v.x = __fma_rn(dr, wr, m0);
    v.y = __fma_rn(di, wr, -m1);
}
template <bool INV>
__device__ __forceinline__ void bfly(double2 &u, double2 &v, double wr, double wi) {
    if (INV)
        gs_bfly(u, v, wr, wi);
    else
        ct_bfly(u, v, wr, wi);
}

// Butterflies are issued in groups of G, LEVEL by level (all first fused multiply-adds of the group, then all second
// ones, ...): ptxas otherwise emits each butterfly as its own chain of dependent DFMAs and a warp that is alone on its
// scheduler waits out the FP64 latency three times per butterfly (profiles/r02a: stall_wait 20 % of samples).
// The arithmetic per butterfly is unchanged, so results are bit-identical.
#ifndef FFT64_GROUP
#define FFT64_GROUP 4
#endif
// FFT64_ABLATE (timing experiments only, results are meaningless): bit 0 drops the butterflies, bit 1 the transposes
#ifndef FFT64_ABLATE
#define FFT64_ABLATE 0
#endif
template <bool INV, int G>
__device__ __forceinline__ void bfly_group(double2 (&x)[32], const int (&iu)[G], const int (&iv)[G], const double (&wr)[G], const double (&wi)[G]) {
    if (FFT64_ABLATE & 1) return;
    if (!INV) {
        double tr[G], ti[G];
#pragma unroll
        for (int k = 0; k < G; k++) {
            tr[k] = __fma_rn(-wi[k], x[iv[k]].y, x[iu[k]].x);
            ti[k] = __fma_rn(wi[k], x[iv[k]].x, x[iu[k]].y);
        }
#pragma unroll
        for (int k = 0; k < G; k++) {
            tr[k] = __fma_rn(wr[k], x[iv[k]].x, tr[k]);
            ti[k] = __fma_rn(wr[k], x[iv[k]].y, ti[k]);
        }
#pragma unroll
        for (int k = 0; k < G; k++) {
            x[iv[k]].x = __fma_rn(2.0, x[iu[k]].x, -tr[k]);
            x[iv[k]].y = __fma_rn(2.0, x[iu[k]].y, -ti[k]);
            x[iu[k]].x = tr[k];
            x[iu[k]].y = ti[k];
        }
    } else {
        double dr[G], di[G], m0[G], m1[G];
#pragma unroll
        for (int k = 0; k < G; k++) {
            dr[k] = __dsub_rn(x[iu[k]].x, x[iv[k]].x);
            di[k] = __dsub_rn(x[iu[k]].y, x[iv[k]].y);
            x[iu[k]].x = __dadd_rn(x[iu[k]].x, x[iv[k]].x);
            x[iu[k]].y = __dadd_rn(x[iu[k]].y, x[iv[k]].y);
        }
#pragma unroll
        for (int k = 0; k < G; k++) {
            m0[k] = __dmul_rn(di[k], wi[k]);
            m1[k] = __dmul_rn(dr[k], wi[k]);
        }
#pragma unroll
        for (int k = 0; k < G; k++) {
            x[iv[k]].x = __fma_rn(dr[k], wr[k], m0[k]);
            x[iv[k]].y = __fma_rn(di[k], wr[k], -m1[k]);
        }
    }
}

// Phase marks.  With a.trace set (lux_debug_fft_trace) lane 0 of every consumer warp adds the cycles since its previous mark
// to a.trace[phase]; scripts/quick_torus.py prints the table.  Phases: 0 decompose, 1 forward pass A, 2 transpose, 3 forward
// pass B, 4 swap write + barrier, 5 key MAC (incl. waits for key rows and the half-size inverse stages), 6 barrier, 7 inverse
// pass B, 8 transpose, 9 inverse pass A, 10 update.
// The marks stay in the production kernel on purpose: each one is a (never taken, uniform) branch that ends a basic block,
// and ptxas schedules the phases separately instead of mixing them.  Measured on B200 at B = 4,736: 85.9 ms with the marks,
// 93.0 ms without, 90.1 ms with a bare never-taken branch (profiles/r02b_fft_variants.txt).  -DFFT64_NO_MARKS removes them.
// MK (a template argument of everything that contains marks) = 0 compiles them out of one instantiation: a warp that is
// alone on its scheduler (one ciphertext per CTA, the latency case) is faster with the free schedule.
// MK >> 8 = marks left out of that instantiation (bit k = mark k).  Which boundaries ptxas should see is an empirical matter
// per kernel (profiles/r02d_mark_sweep.txt): the four-ciphertext kernel is 4.1 % faster without marks 1, 7 and 8 (forward pass A
// flows into its transpose, the inverse pass B into the transpose and the inverse pass A: 386 << 8), the three-ciphertext kernel
// 5 % slower with the same mask.
#ifndef FFT64_SKIP_MARKS
#define FFT64_SKIP_MARKS 0   // bit k set: phase mark k is left out (codegen experiments)
#endif
#ifndef FFT64_NO_MARKS
#define FFT64_MARK(k)                                                                  \
    do {                                                                               \
        if ((MK & 1) && !(((FFT64_SKIP_MARKS | (MK >> 8)) >> (k)) & 1) && a.trace && lane == 0) {                                      \
            long long now_;                                                            \
            asm volatile("mov.u64 %0, %%clock64;" : "=l"(now_)::"memory");             \
            atomicAdd((unsigned long long *)a.trace + (k), (unsigned long long)(now_ - mark_)); \
            mark_ = now_;                                                              \
        }                                                                              \
    } while (0)
#else
#define FFT64_MARK(k) asm volatile("" ::: "memory")
#endif

// optional marks between the stages of a pass (experiment: -DFFT64_FINE_MARKS)
#ifdef FFT64_FINE_MARKS
#define FFT64_STAGE_MARK() FFT64_MARK(15)
#else
#define FFT64_STAGE_MARK() \
    do {                   \
    } while (0)
#endif

// pass A: tree levels m = 1..16 on the column layout; node m + blk pairs registers blk*2t + jj and + t, t = 16/m
template <bool INV, int MK = 1>
__device__ __forceinline__ void pass_a(double2 (&x)[32], const Args &a, int lane, long long &mark_) {
    constexpr int G = (MK & 2) ? 8 : FFT64_GROUP;   // MK bit 1: groups of 8 (a warp alone on its scheduler hides the FP64 latency itself)
#pragma unroll
    for (int s0 = 0; s0 < 5; s0++) {
        const int s = INV ? 4 - s0 : s0;
        const int m = 1 << s, t = 16 >> s;
#pragma unroll
        for (int b0 = 0; b0 < 16; b0 += G) {
            int iu[G], iv[G];
            double wr[G], wi[G];
#pragma unroll
            for (int k = 0; k < G; k++) {
                const int b = b0 + k, blk = b / t, jj = b % t, node = m + blk;
                iu[k] = blk * 2 * t + jj;
                iv[k] = iu[k] + t;
                if (node > 1 && (node & 1)) {   // tw[2v+1] = i * tw[2v]
                    wr[k] = -a.twA[node - 1].y;
                    wi[k] = a.twA[node - 1].x;
                } else {
                    wr[k] = a.twA[node].x;
                    wi[k] = a.twA[node].y;
                }
            }
            bfly_group<INV, G>(x, iu, iv, wr, wi);
        }
        FFT64_STAGE_MARK();
    }
}
// pass B: tree levels m = 32 g, g = 1..16, on the row layout; lane r owns nodes m + r g + b, b < g
// butterflies [B0, B1) of the 16 of one stage: for g >= 2 the rows 0..15 (butterflies 0..7) and 16..31 (8..15) are independent
template <bool INV, int GI, int B0 = 0, int B1 = 16, int MK = 1>
__device__ __forceinline__ void pass_b_stage(double2 (&y)[32], const double2 *twB, int lane) {
    constexpr int G0 = (MK & 2) ? 8 : FFT64_GROUP;
    constexpr int G = G0 < (B1 - B0) ? G0 : (B1 - B0);
    constexpr int g = 1 << GI, t = 16 >> GI, NW = g >= 2 ? g / 2 : 1;
    constexpr int E0 = (B0 / t) / 2, E1 = ((B1 - 1) / t) / 2 + 1;
    double2 w[NW];
#pragma unroll
    for (int e = E0; e < E1; e++) w[e] = twB[(g == 1 ? 0 : 16 * g) + e * 32 + lane];
#pragma unroll
    for (int b0 = B0; b0 < B1; b0 += G) {
        int iu[G], iv[G];
        double wr[G], wi[G];
#pragma unroll
        for (int k = 0; k < G; k++) {
            const int b = b0 + k, blk = b / t, jj = b % t;
            iu[k] = blk * 2 * t + jj;
            iv[k] = iu[k] + t;
            if (blk & 1) {
                wr[k] = -w[blk / 2].y;
                wi[k] = w[blk / 2].x;
            } else {
                wr[k] = w[blk / 2].x;
                wi[k] = w[blk / 2].y;
            }
        }
        bfly_group<INV, G>(y, iu, iv, wr, wi);
    }
}
template <bool INV, int MK = 1>
__device__ __forceinline__ void pass_b(double2 (&y)[32], const double2 *twB, int lane, const Args &a, long long &mark_) {
    if (!INV) {
        pass_b_stage<false, 0, 0, 16, MK>(y, twB, lane);
        FFT64_STAGE_MARK();
        pass_b_stage<false, 1, 0, 16, MK>(y, twB, lane);
        FFT64_STAGE_MARK();
        pass_b_stage<false, 2, 0, 16, MK>(y, twB, lane);
        FFT64_STAGE_MARK();
        pass_b_stage<false, 3, 0, 16, MK>(y, twB, lane);
        FFT64_STAGE_MARK();
        pass_b_stage<false, 4, 0, 16, MK>(y, twB, lane);
    } else {
        pass_b_stage<true, 4, 0, 16, MK>(y, twB, lane);
        pass_b_stage<true, 3, 0, 16, MK>(y, twB, lane);
        pass_b_stage<true, 2, 0, 16, MK>(y, twB, lane);
        pass_b_stage<true, 1, 0, 16, MK>(y, twB, lane);
        pass_b_stage<true, 0, 0, 16, MK>(y, twB, lane);
    }
}
// inverse pass B without its last stage, on rows 16 H .. 16 H + 15 only (the key MAC of the other half can still be pending)
template <int H, int MK = 1>
__device__ __forceinline__ void pass_b_inv_half(double2 (&y)[32], const double2 *twB, int lane, const Args &a, long long &mark_) {
    pass_b_stage<true, 4, 8 * H, 8 * H + 8, MK>(y, twB, lane);
    FFT64_STAGE_MARK();
    pass_b_stage<true, 3, 8 * H, 8 * H + 8, MK>(y, twB, lane);
    FFT64_STAGE_MARK();
    pass_b_stage<true, 2, 8 * H, 8 * H + 8, MK>(y, twB, lane);
    FFT64_STAGE_MARK();
    pass_b_stage<true, 1, 8 * H, 8 * H + 8, MK>(y, twB, lane);
}
// transposes through the warp's padded buffer
__device__ __forceinline__ void col_to_row(double2 (&x)[32], double2 *buf, int lane) {
#pragma unroll
    for (int i = 0; i < 32; i++) buf[i * ROWP + lane] = x[i];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) x[i] = buf[lane * ROWP + i];
}
__device__ __forceinline__ void row_to_col(double2 (&y)[32], double2 *buf, int lane) {
#pragma unroll
    for (int i = 0; i < 32; i++) buf[lane * ROWP + i] = y[i];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) y[i] = buf[i * ROWP + lane];
}
// half-size buffer (HB): the real parts cross first, then the imaginary parts, through 32 x 33 doubles (8,448 B per warp
// instead of 16,896): same bytes through the crossbar, twice the instructions; this is what lets FOUR ciphertexts (eight
// consumer warps, two on every scheduler) live in one CTA's shared memory.  Conflict-free both ways (row stride 33 doubles).
constexpr int ROWD = 33;
constexpr int HBUF_BYTES = 32 * ROWD * 8;
__device__ __forceinline__ void col_to_row_h(double2 (&x)[32], double *b, int lane) {
    if (FFT64_ABLATE & 2) return;
#pragma unroll
    for (int i = 0; i < 32; i++) b[i * ROWD + lane] = x[i].x;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) x[i].x = b[lane * ROWD + i];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) b[i * ROWD + lane] = x[i].y;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) x[i].y = b[lane * ROWD + i];
}
__device__ __forceinline__ void row_to_col_h(double2 (&y)[32], double *b, int lane) {
    if (FFT64_ABLATE & 2) return;
#pragma unroll
    for (int i = 0; i < 32; i++) b[lane * ROWD + i] = y[i].x;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) y[i].x = b[i * ROWD + lane];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) b[lane * ROWD + i] = y[i].y;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 32; i++) y[i].y = b[i * ROWD + lane];
}
// forward: column layout in (z[lane + 32 i]), row layout out (Z[32 lane + i], oracle index order)
template <int HB = 0>
__device__ __forceinline__ void fft_forward(double2 (&x)[32], double2 *buf, const double2 *twB, const Args &a, int lane) {
    long long mark_ = 0;
    pass_a<false>(x, a, lane, mark_);
    if (HB)
        col_to_row_h(x, (double *)buf, lane);
    else
        col_to_row(x, buf, lane);
    pass_b<false>(x, twB, lane, a, mark_);
}
// inverse (unnormalised): row layout in, column layout out
template <int HB = 0>
__device__ __forceinline__ void fft_inverse(double2 (&y)[32], double2 *buf, const double2 *twB, const Args &a, int lane) {
    long long mark_ = 0;
    pass_b<true>(y, twB, lane, a, mark_);
    if (HB)
        row_to_col_h(y, (double *)buf, lane);
    else
        row_to_col(y, buf, lane);
    pass_a<true>(y, a, lane, mark_);
}

// ------------------------------------------------------------------------------------------ blind rotation
// ring stage = RS rows of the GGSW: [RS][4 polys][32 lanes] double2 = RS * 2 KiB
// TWS: pass-B twiddle table in shared memory (1) or read from global memory through L1 (0: the 8 KiB go to the key ring)
// HB: half-size per-warp buffers (transposes in two rounds, spectrum swap in two rounds of 16 rows)
// GROUPS: the CTA's teams form GROUPS independent groups, each with its own key ring (NS stages), its own producer warp
// and a staggered start: the groups are never in the same phase of the CMUX, so the integer (decompose), FP64 (butterflies)
// and shared-memory (transposes) phases of one group overlap other phases of the other group instead of queueing on one pipe
// (profiles/r02b_overlap_ubench.txt: the FP64 and shared-memory pipes run concurrently at full rate when both have work).
// KL: no key ring at all: every warp loads its key rows straight from L2 / L1 (ld.global.nc) into registers.  Nothing then
// couples the teams of a CTA (no shared ring, no producer warp, no mbarriers): they drift out of phase and the integer,
// FP64 and shared-memory phases of different teams overlap.  Costs TEAMS x the L2 -> SM key traffic of the ring.
// KL == 5: tensor memory as the key cache, filled by the copy engines: the producer warp streams the key into a small staging
// ring with cp.async.bulk and moves every staged piece on with tcgen05.cp (shared -> tensor memory, 64 lanes x 128 bit with
// the warpx2::02_13 multicast: the 32 lanes of key column r = 0 land in lane quadrants 0 and 2, those of r = 1 in 1 and 3 -
// exactly the quadrants of the consumer warps that need them); completion comes back through tcgen05.commit on mbarriers.
// No thread's registers or load/store unit touch the key until a consumer's tcgen05.ld (mac_tmem).
// KL == 4: the ring WITHOUT a producer warp: lane 0 of consumer warp 0 issues the bulk copies (RingFeed below).  A ninth warp
// puts three warps on one scheduler, and a scheduler's 16 K registers then allow 168 per thread; eight warps may use 255.
template <int TEAMS, int NS, int RS, int TWS = 1, int HB = 0, int GROUPS = 1, int KL = 0>
struct Cfg {
    static constexpr int CONSUMER_WARPS = 2 * TEAMS;
    static constexpr int THREADS = 32 * (CONSUMER_WARPS + (KL == 3 ? 4 : KL == 5 ? 1 : KL ? 0 : GROUPS));   // KL == 3: one producer warp per lane quadrant
    static constexpr int TEAMS_PER_GROUP = TEAMS / GROUPS;
    static constexpr int STAGE_POINTS = RS * 4 * 32;
    static constexpr int STAGES_PER_GGSW = 32 / RS;
    static constexpr int WARP_BUF_POINTS = HB ? HBUF_BYTES / 16 : BUF_POINTS;   // in double2 units
    static constexpr size_t SMEM =
        sizeof(double2) * TWB_ENTRIES * TWS + (size_t)TEAMS * (sizeof(u64) * 2 * N + sizeof(double2) * 2 * WARP_BUF_POINTS) +
        (KL == 3 ? sizeof(u64) * 16 : KL == 5 ? sizeof(double2) * (size_t)NS * STAGE_POINTS + sizeof(u64) * (2 * NS + 4) : (KL && KL != 4) ? 0 : sizeof(double2) * (size_t)GROUPS * NS * STAGE_POINTS + sizeof(u64) * 2 * NS * GROUPS + (KL == 4 ? sizeof(u32) * NS : 0)) + 16;
};

// ------------------------------------------------------------------------------------------ tensor memory as the key cache
// KL == 3: the Fourier key does not pass through shared memory at all.  Tensor memory (256 KB per SM, 128 lanes x 512
// 32-bit columns, a datapath of its own: tcgen05.ld / tcgen05.st) holds, for every lane quadrant, two whole GGSWs of the key
// column that quadrant's consumer warps need (a warp reaches only the lanes 32 (warp % 4) .. +31, and warp parity = output
// polynomial r, so a quadrant serves ONE r: 32 rows x 2 (j) x 4 words = 256 columns per GGSW, double-buffered in 512).
// Column of (buffer b, row i, j, word c) = 256 b + 8 i + 4 j + c.  Four producer warps (one per quadrant) load the rows of
// the NEXT iteration straight from L2 (ld.global.nc, lane-contiguous 512-byte rows) and tcgen05.st them while the consumers
// transform; the consumers tcgen05.ld two rows at a time in the key MAC.  The crossbar loses a third of the MAC's traffic
// (64 of 96 LDS.128 per warp), the ring, its waits and its lock step with it.
__device__ __forceinline__ void tmem_ld16(u32 taddr, u32 (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st16(u32 taddr, const u32 (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
        "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// whole-CTA allocation of all 512 columns by one warp (the kernel wrapper calls these around the body: every CTA must free
// what it took, or the next CTA on the SM spins on its allocation for ever)
__device__ __forceinline__ void tmem_alloc_512(u32 *smem_slot) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tk_smem_addr(smem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free_512(u32 taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}

// shared -> tensor memory: 64 rows of 16 bytes (contiguous: no swizzle, 8-row core matrices 128 bytes apart); rows 0..31 go to
// lanes 0..31 of quadrants 0 and 2, rows 32..63 to quadrants 1 and 3, four columns from taddr (scripts/ubench/tcgen05_cp_probe.cu)
__device__ __forceinline__ void tmem_cp_64x128(u32 taddr, const void *smem_src) {
    const u64 desc = (u64)((tk_smem_addr(smem_src) >> 4) & 0x3FFFu) | ((u64)8 << 32) | ((u64)1 << 46);
    asm volatile("tcgen05.cp.cta_group::1.64x128b.warpx2::02_13 [%0], %1;" ::"r"(taddr), "l"(desc) : "memory");
}
// the mbarrier gets one arrival when every tcgen05 operation this thread issued before has completed
__device__ __forceinline__ void tmem_commit(u64 *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tk_smem_addr(bar)) : "memory");
}

// multiply-accumulate of the swapped spectra with the staged key rows: out = X0 K[0][r] + X1 K[1][r], in the oracle's order;
// the RS rows of a stage advance together, one fused multiply-add of each chain at a time
// HB: the partner's rows sit as [row - 16 * (row0 / 16)][lane] (half-size buffer, two rounds of 16 rows)
// IL (interleave, RS == 4 only): a butterfly stage on the OTHER sixteen rows is issued between the loads of this chunk and its
// multiply-adds, so that the FP64 pipe works while the key rows cross the shared-memory crossbar (in lock step all warps of the
// SM otherwise load at the same time and multiply at the same time).  IL = 1: forward stage ILS of rows 16..31 (butterflies
// 8..15; the chunk's own rows are 0..15); IL = 2: inverse stage ILS of rows 0..15 (the chunk's rows are 16..31).
template <int R, int RS, int HB = 0, int KG = 0, int IL = 0, int ILS = 0, int MK = 1>
__device__ __forceinline__ void mac_rows(double2 (&y)[32], const double2 *oth, const double2 *stage, int row0, int lane,
                                         const u32 *kw = nullptr, const double2 *twB = nullptr) {
    double2 k0[RS], k1[RS], x0[RS], x1[RS];
    double re[RS], im[RS];
    // KG == 2: kw = the 16 words of two rows (RS == 2) of this quadrant's key column, already loaded from tensor memory
    static_assert(KG != 2 || RS == 2, "tensor-memory key rows come two at a time");
#pragma unroll
    for (int ii = 0; ii < RS; ii++) {
        const int i = row0 + ii;   // compile-time after unrolling of the caller
        if (KG == 2) {
            k0[ii] = make_double2(__hiloint2double((int)kw[(ii & 1) * 8 + 1], (int)kw[(ii & 1) * 8 + 0]),
                                  __hiloint2double((int)kw[(ii & 1) * 8 + 3], (int)kw[(ii & 1) * 8 + 2]));
            k1[ii] = make_double2(__hiloint2double((int)kw[(ii & 1) * 8 + 5], (int)kw[(ii & 1) * 8 + 4]),
                                  __hiloint2double((int)kw[(ii & 1) * 8 + 7], (int)kw[(ii & 1) * 8 + 6]));
        } else {
            k0[ii] = KG ? __ldg(stage + (ii * 4 + 0 + R) * 32 + lane) : stage[(ii * 4 + 0 + R) * 32 + lane];
            k1[ii] = KG ? __ldg(stage + (ii * 4 + 2 + R) * 32 + lane) : stage[(ii * 4 + 2 + R) * 32 + lane];
        }
        const double2 o = HB ? oth[(i & 15) * 32 + lane] : oth[lane * ROWP + i];
        x0[ii] = R == 0 ? y[i] : o;
        x1[ii] = R == 0 ? o : y[i];
    }
    if (IL == 1) pass_b_stage<false, ILS, 8, 16, MK>(y, twB, lane);
    if (IL == 2) pass_b_stage<true, ILS, 0, 8, MK>(y, twB, lane);
#pragma unroll
    for (int ii = 0; ii < RS; ii++) {
        re[ii] = __fma_rn(x0[ii].x, k0[ii].x, 0.0);
        im[ii] = __fma_rn(x0[ii].x, k0[ii].y, 0.0);
    }
#pragma unroll
    for (int ii = 0; ii < RS; ii++) {
        re[ii] = __fma_rn(-x0[ii].y, k0[ii].y, re[ii]);
        im[ii] = __fma_rn(x0[ii].y, k0[ii].x, im[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < RS; ii++) {
        re[ii] = __fma_rn(x1[ii].x, k1[ii].x, re[ii]);
        im[ii] = __fma_rn(x1[ii].x, k1[ii].y, im[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < RS; ii++) {
        re[ii] = __fma_rn(-x1[ii].y, k1[ii].y, re[ii]);
        im[ii] = __fma_rn(x1[ii].y, k1[ii].x, im[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < RS; ii++) y[row0 + ii] = make_double2(re[ii], im[ii]);
}

// KL: chunks [C0, C1) of the GGSW at `ggsw` in global memory, no staging
template <int R, int RS, int HB, int C0, int C1>
__device__ __forceinline__ void mac_global(double2 (&y)[32], const double2 *oth, const double2 *ggsw, int lane) {
#pragma unroll
    for (int c = C0; c < C1; c++) mac_rows<R, RS, HB, 1>(y, oth, ggsw + (size_t)c * RS * 4 * 32, c * RS, lane);
}
// KL == 3: chunks [C0, C1) of two rows each from tensor memory (taddr = quadrant lane base + buffer column base)
// The loads are software-pipelined: the two rows of chunk c + 1 are requested before the multiply-adds of chunk c
// (tcgen05.wait::ld waits for every earlier load of the thread, so the wait sits right before the next request).
template <int R, int HB, int C0, int C1>
__device__ __forceinline__ void mac_tmem(double2 (&y)[32], const double2 *oth, u32 taddr, int lane) {
    u32 ka[16], kb[16];
    tmem_ld16(taddr + 16u * C0, ka);
#pragma unroll
    for (int c = C0; c < C1; c += 2) {
        tmem_wait_ld();
        if (c + 1 < C1) tmem_ld16(taddr + 16u * (c + 1), kb);
        mac_rows<R, 2, HB, 2>(y, oth, nullptr, c * 2, lane, ka);
        if (c + 1 < C1) {
            tmem_wait_ld();
            if (c + 2 < C1) tmem_ld16(taddr + 16u * (c + 2), ka);
            mac_rows<R, 2, HB, 2>(y, oth, nullptr, (c + 1) * 2, lane, kb);
        }
    }
}
// two rows of one ring stage (and the partner's two rows) in registers: the unit of the software-pipelined MAC below
struct KeyRows2 {
    double2 k0[2], k1[2], o[2];
};
template <int R, int HB>
__device__ __forceinline__ void load_rows2(KeyRows2 &d, const double2 *stage, const double2 *oth, int ii0, int row, int lane) {
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        d.k0[ii] = stage[((ii0 + ii) * 4 + 0 + R) * 32 + lane];
        d.k1[ii] = stage[((ii0 + ii) * 4 + 2 + R) * 32 + lane];
        d.o[ii] = HB ? oth[((row + ii) & 15) * 32 + lane] : oth[lane * ROWP + row + ii];
    }
}
// out = X0 K[0][r] + X1 K[1][r] for rows row, row + 1: per point the oracle's chain of four fused multiply-adds
template <int R>
__device__ __forceinline__ void mac_rows2(double2 (&y)[32], const KeyRows2 &d, int row) {
    double re[2], im[2];
    double2 x0[2], x1[2];
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        x0[ii] = R == 0 ? y[row + ii] : d.o[ii];
        x1[ii] = R == 0 ? d.o[ii] : y[row + ii];
    }
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        re[ii] = __fma_rn(x0[ii].x, d.k0[ii].x, 0.0);
        im[ii] = __fma_rn(x0[ii].x, d.k0[ii].y, 0.0);
    }
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        re[ii] = __fma_rn(-x0[ii].y, d.k0[ii].y, re[ii]);
        im[ii] = __fma_rn(x0[ii].y, d.k0[ii].x, im[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        re[ii] = __fma_rn(x1[ii].x, d.k1[ii].x, re[ii]);
        im[ii] = __fma_rn(x1[ii].x, d.k1[ii].y, im[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < 2; ii++) {
        re[ii] = __fma_rn(-x1[ii].y, d.k1[ii].y, re[ii]);
        im[ii] = __fma_rn(x1[ii].y, d.k1[ii].x, im[ii]);
    }
#pragma unroll
    for (int ii = 0; ii < 2; ii++) y[row + ii] = make_double2(re[ii], im[ii]);
}

// KL == 4: the key ring fed by its consumers.  Piece f of the key goes into stage f % NS once every consumer warp has
// released piece f - NS, and the warp whose arrival completes that phase of the stage's `empty` barrier is the one that
// requests it: after its own arrival a warp tests the phase (it can only be the phase it arrived in, or complete) and, if
// complete, claims the refill with a compare-and-swap on claim[stage] (= the piece the stage holds or has been promised), so
// exactly one lane issues the bulk copy, at the moment the stage becomes free — no producer warp, no polling.
struct RingFeed {
    const double2 *bsk;
    double2 *ring;
    u32 *claim;
    u32 total;
};
__device__ __forceinline__ bool mbar_test(u64 *bar, u32 parity) {
    u32 ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(ok)
        : "r"(tk_smem_addr(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
template <int NS, int SP>
__device__ __forceinline__ void feed_issue(const RingFeed &fd, u32 f, u64 *full) {
    constexpr u32 bytes = (u32)(sizeof(double2) * SP);
    const u32 s = f % NS;
    mbar_expect_tx(full + s, bytes);
    bulk_g2s(fd.ring + (size_t)s * SP, fd.bsk + (size_t)f * SP, bytes, full + s);
}
// lane 0 of a consumer warp, right after its arrival on empty[t % NS] for piece t
template <int NS, int SP>
__device__ __forceinline__ void feed_after_release(const RingFeed &fd, u32 t, u64 *full, u64 *empty) {
    const u32 s = t % NS, f = t + NS;
    if (f < fd.total && mbar_test(empty + s, (t / NS) & 1) && atomicCAS(fd.claim + s, t, f) == t) feed_issue<NS, SP>(fd, f, full);
}
// chunks [C0, C1) of the 32 / RS key-row chunks of one GGSW.
// -DFFT64_MAC_PIPE=1 (RS == 4): software-pipelined — the loads of the next two rows (key rows and partner rows, the next
// stage's full barrier first where they are its rows) are issued before the multiply-adds of the current two, so the
// shared-memory latency of the MAC (ncu: 10 % of all stall samples sit on the first multiply-add of a chunk) overlaps the
// FP64 work.  Measured: the MAC phase drops from 5.75 k to 4.9 k cycles but the 48 extra live registers slow every other
// phase (94.5 ms vs 85.8 ms at B = 4,736): off by default (profiles/r02b_fft_variants.txt).
#ifndef FFT64_MAC_PIPE
#define FFT64_MAC_PIPE 0
#endif
// IL (RS == 4, four chunks): chunk c carries forward stage 1 + c of rows 16..31 (IL = 1) or inverse stage 4 - c of rows 0..15 (IL = 2)
template <int R, int NS, int RS, int HB = 0, int C0 = 0, int C1 = 32 / RS, int IL = 0, int MK = 1>
__device__ __forceinline__ void mac_all(double2 (&y)[32], const double2 *oth, const double2 *ring, u64 *full, u64 *empty, u32 &t, int lane,
                                        const RingFeed *fd = nullptr, const double2 *twB = nullptr) {
    constexpr int SP = RS * 4 * 32;
    if (FFT64_MAC_PIPE && RS == 4) {
        KeyRows2 A, B;
        u32 s = t % NS;
        mbar_wait(full + s, (t / NS) & 1);
        const double2 *stage = ring + (size_t)s * SP;
        load_rows2<R, HB>(A, stage, oth, 0, C0 * RS, lane);
#pragma unroll
        for (int c = C0; c < C1; c++) {
            load_rows2<R, HB>(B, stage, oth, 2, c * RS + 2, lane);
            mac_rows2<R>(y, A, c * RS);
            const u32 s_cur = s;
            if (c + 1 < C1) {
                s = (t + 1) % NS;
                mbar_wait(full + s, ((t + 1) / NS) & 1);
                stage = ring + (size_t)s * SP;
                load_rows2<R, HB>(A, stage, oth, 0, (c + 1) * RS, lane);
            }
            mac_rows2<R>(y, B, c * RS + 2);
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(empty + s_cur);
                if (fd) feed_after_release<NS, SP>(*fd, t, full, empty);
            }
            t++;
        }
        return;
    }
    if constexpr (IL != 0) {
        static_assert(RS == 4 && C1 - C0 == 4, "the interleave pairs four chunks with four butterfly stages");
        auto chunk = [&](auto cc) {
            constexpr int c = decltype(cc)::value;
            const u32 s = t % NS;
            mbar_wait(full + s, (t / NS) & 1);
            mac_rows<R, RS, HB, 0, IL, (IL == 1 ? 1 + c : 4 - c), MK>(y, oth, ring + (size_t)s * SP, (C0 + c) * RS, lane, nullptr, twB);
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(empty + s);
                if (fd) feed_after_release<NS, SP>(*fd, t, full, empty);
            }
            t++;
        };
        chunk(std::integral_constant<int, 0>{});
        chunk(std::integral_constant<int, 1>{});
        chunk(std::integral_constant<int, 2>{});
        chunk(std::integral_constant<int, 3>{});
        return;
    }
#pragma unroll
    for (int c = C0; c < C1; c++) {
        const u32 s = t % NS;
        mbar_wait(full + s, (t / NS) & 1);
        mac_rows<R, RS, HB>(y, oth, ring + (size_t)s * SP, c * RS, lane);
        __syncwarp();
        if (lane == 0) {
            mbar_arrive(empty + s);
            if (fd) feed_after_release<NS, SP>(*fd, t, full, empty);
        }
        t++;
    }
}

// LAG (experiments, profiles/r02d_overlap_variants.txt): the teams of the upper half of a full CTA (the second warp of every scheduler)
// run BEHIND the lower half instead of in lock step with it, so that a scheduler's two warps are in different phases (one in
// an FP64 phase, the other in a shared-memory or integer phase).  LAG = 1: a start delay of a.stagger_ns cycles only;
// LAG = 2: the offset is also held by two named barriers (13: a trailer starts iteration i after the leaders passed point
// a.stagger_ns / 10 of iteration i; 14: a leader starts iteration i + 1 after the trailers passed point a.stagger_ns % 10 of i).
template <int TEAMS, int NS, int RS, int TWS, int HB = 0, int GROUPS = 1, int KL = 0, int MK = 1, int LAG = 0>
__device__ __forceinline__ void blind_rotate_body(const Args &a, unsigned char *smem_raw, u32 tmem_base = 0) {
    using C = Cfg<TEAMS, NS, RS, TWS, HB, GROUPS, KL>;
    static_assert(TEAMS % GROUPS == 0, "teams must split evenly into groups");
    constexpr int TPG = C::TEAMS_PER_GROUP;
    constexpr int BUF_POINTS = C::WARP_BUF_POINTS;   // shadows the full-size constant
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    double2 *tw_smem = (double2 *)smem_raw;
    const double2 *twB = TWS ? tw_smem : a.twB;
    u64 *acc_base = (u64 *)(tw_smem + TWB_ENTRIES * TWS);
    double2 *buf_base = (double2 *)(acc_base + (size_t)TEAMS * 2 * N);
    double2 *ring_all = buf_base + (size_t)TEAMS * 2 * BUF_POINTS;
    u64 *full_all = (u64 *)(ring_all + (size_t)GROUPS * NS * C::STAGE_POINTS);
    u64 *empty_all = full_all + GROUPS * NS;

    const u32 first_ct = (TEAMS < 4 ? a.ct0 : 0u) + blockIdx.x * TEAMS;
    const u32 left = a.B - first_ct;
    const u32 live_teams = left < (u32)TEAMS ? left : (u32)TEAMS;
    if (TWS)
        for (int i = tid; i < TWB_ENTRIES; i += C::THREADS) tw_smem[i] = a.twB[i];
    // live teams of group g: teams [g TPG, (g + 1) TPG) below live_teams
    auto live_in_group = [&](u32 g) -> u32 {
        const u32 lo = g * TPG;
        return live_teams <= lo ? 0u : (live_teams - lo < (u32)TPG ? live_teams - lo : (u32)TPG);
    };
    // KL == 3: full / empty barrier of (quadrant q, buffer b) at index 2 q + b, in the place of the ring's barriers
    u64 *tfull = (u64 *)ring_all, *tempty = tfull + 8;
    auto quadrant_consumers = [&](u32 q) -> u32 {   // live consumer warps w = q, q + 4, ... of this CTA
        u32 c = 0;
        for (u32 w = q; w < (u32)C::CONSUMER_WARPS; w += 4) c += (w >> 1) < live_teams ? 1u : 0u;
        return c;
    };
    if (KL == 3 && tid == 0) {
        for (u32 q = 0; q < 4; q++)
            for (u32 b = 0; b < 2; b++) {
                mbar_init(tfull + 2 * q + b, 1);
                mbar_init(tempty + 2 * q + b, quadrant_consumers(q) ? quadrant_consumers(q) : 1);
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (KL == 5 && tid == 0) {   // staging ring: full (TMA bytes) / empty (tcgen05.commit); key buffers: tfull (commit) / tempty (consumer warps)
        for (int s = 0; s < NS; s++) {
            mbar_init(full_all + s, 1);
            mbar_init(empty_all + s, 1);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(empty_all + NS + b, 1);
            mbar_init(empty_all + NS + 2 + b, 2 * live_teams);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if ((!KL || KL == 4) && tid == 0) {
        for (int g = 0; g < GROUPS; g++)
            for (int s = 0; s < NS; s++) {
                mbar_init(full_all + g * NS + s, 1);
                mbar_init(empty_all + g * NS + s, 2 * live_in_group(g));
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const u32 total = a.steps * C::STAGES_PER_GGSW;
    if (KL == 4 && tid == 0) {   // the first NS pieces; claim[s] = the piece stage s holds
        const RingFeed f0{a.bsk, ring_all, (u32 *)(empty_all + GROUPS * NS), total};
        for (u32 f = 0; f < (u32)NS && f < total; f++) {
            f0.claim[f] = f;
            feed_issue<NS, C::STAGE_POINTS>(f0, f, full_all);
        }
    }
    __syncthreads();

    if (KL == 3 && warp >= C::CONSUMER_WARPS) {   // ---- tensor-memory producers: warp CONSUMER_WARPS + q feeds quadrant q
        const u32 q = warp - C::CONSUMER_WARPS, pr = q & 1;
        if (quadrant_consumers(q) == 0) return;
        const u32 tq = tmem_base + ((32u * q) << 16);
#pragma unroll 1
        for (u32 it = 0; it < a.steps; it++) {
            const u32 b = it & 1;
            if (it >= 2) mbar_wait(tempty + 2 * q + b, ((it >> 1) - 1) & 1);
            tmem_fence_after();
            const double2 *g = a.bsk + (size_t)it * 32 * 4 * 32;
#pragma unroll 4
            for (int i0 = 0; i0 < 32; i0 += 2) {
                u32 v[16];
#pragma unroll
                for (int ii = 0; ii < 2; ii++)
#pragma unroll
                    for (int j = 0; j < 2; j++) {
                        const double2 k = __ldg(g + ((i0 + ii) * 4 + 2 * j + pr) * 32 + lane);
                        v[ii * 8 + j * 4 + 0] = (u32)__double2loint(k.x);
                        v[ii * 8 + j * 4 + 1] = (u32)__double2hiint(k.x);
                        v[ii * 8 + j * 4 + 2] = (u32)__double2loint(k.y);
                        v[ii * 8 + j * 4 + 3] = (u32)__double2hiint(k.y);
                    }
                tmem_st16(tq + 256u * b + 8u * (u32)i0, v);
                if (a.stagger_ns) __nanosleep(a.stagger_ns);   // spread the refill over the iteration instead of one burst after the MAC
            }
            tmem_wait_st();
            tmem_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tfull + 2 * q + b);
        }
        return;
    }
    if (KL == 5 && warp >= C::CONSUMER_WARPS) {   // ---- one lane: key -> staging ring (TMA) -> tensor memory (tcgen05.cp)
        if (lane == 0) {
            constexpr u32 bytes = (u32)(sizeof(double2) * C::STAGE_POINTS);
            constexpr u32 SPG = C::STAGES_PER_GGSW;
            u64 *tfull5 = empty_all + NS, *tempty5 = empty_all + NS + 2;
            u32 issued = 0;
            const u32 total5 = a.stagger_ns == 7777u ? (total < 2 * SPG ? total : 2 * SPG) : total;
            for (u32 t = 0; t < total5; t++) {
                while (issued < total5 && issued < t + (u32)NS) {   // keep the staging ring requested NS pieces ahead
                    const u32 s2 = issued % NS;
                    if (issued >= (u32)NS) mbar_wait(empty_all + s2, ((issued / NS) - 1) & 1);
                    mbar_expect_tx(full_all + s2, bytes);
                    bulk_g2s(ring_all + (size_t)s2 * C::STAGE_POINTS, a.bsk + (size_t)issued * C::STAGE_POINTS, bytes, full_all + s2);
                    issued++;
                }
                const u32 s = t % NS, itp = t / SPG, c = t % SPG, b = itp & 1;
                if (c == 0 && itp >= 2) mbar_wait(tempty5 + b, ((itp >> 1) - 1) & 1);   // the consumers are done with GGSW itp - 2
                mbar_wait(full_all + s, (t / NS) & 1);
                tmem_fence_after();
                const double2 *st = ring_all + (size_t)s * C::STAGE_POINTS;
#pragma unroll
                for (int ii = 0; ii < RS; ii++)
#pragma unroll
                    for (int j = 0; j < 2; j++)
                        tmem_cp_64x128(tmem_base + 256u * b + 8u * (c * RS + ii) + 4u * j, st + (ii * 4 + 2 * j) * 32);
                tmem_commit(empty_all + s);
                if (c == SPG - 1) tmem_commit(tfull5 + b);
                if (a.stagger_ns) __nanosleep(a.stagger_ns);   // pace the refill over the iteration instead of one burst after the MAC
            }
        }
        return;
    }
    if (!KL && warp >= C::CONSUMER_WARPS) {   // ---- producer warps (one per group): one lane streams the key rows
        const u32 g = warp - C::CONSUMER_WARPS;
        double2 *ring = ring_all + (size_t)g * NS * C::STAGE_POINTS;
        u64 *full = full_all + g * NS, *empty = empty_all + g * NS;
        if (lane == 0 && live_in_group(g)) {
            constexpr u32 bytes = (u32)(sizeof(double2) * C::STAGE_POINTS);
            for (u32 t = 0; t < total; t++) {
                const u32 s = t % NS;
                if (t >= (u32)NS) mbar_wait(empty + s, ((t / NS) - 1) & 1);
                mbar_expect_tx(full + s, bytes);
                bulk_g2s(ring + (size_t)s * C::STAGE_POINTS, a.bsk + (size_t)t * C::STAGE_POINTS, bytes, full + s);
            }
        }
        return;
    }

    const int team = warp >> 1, r = warp & 1;
    const u32 ct = first_ct + team;
    if (ct >= a.B) return;
    const int grp = team / TPG;
    const double2 *ring = ring_all + (size_t)grp * NS * C::STAGE_POINTS;
    u64 *full = full_all + grp * NS, *empty = empty_all + grp * NS;
    u64 *acc = acc_base + (size_t)team * 2 * N + (size_t)r * N;   // this warp's accumulator polynomial
    double2 *buf = buf_base + ((size_t)team * 2 + r) * BUF_POINTS;
    const double2 *buf_oth = buf_base + ((size_t)team * 2 + (r ^ 1)) * BUF_POINTS;
    const u64 *lwe = a.lwe + (size_t)ct * (a.n + 1);
    const u64 *tv = a.testv + (size_t)(a.lut_idx ? a.lut_idx[ct / a.lut_div] : 0) * N;

    // acc0 = X^{ms(-b)} (0, testv)      ivc_based_vpbs.rs:109-124
    {
        const u32 sh0 = mod_switch((u64)0 - lwe[a.n]);
        for (int p = lane; p < N; p += 32) acc[p] = r ? rotated(tv, (u32)p, sh0) : 0;
    }
    __syncwarp();

    u32 t = 0;
    const RingFeed feed{a.bsk, ring_all, (u32 *)(empty_all + GROUPS * NS), total};
    const RingFeed *fdp = KL == 4 ? &feed : nullptr;
    u64 mask = a.steps ? lwe[0] : 0;
    const int barrier_id = 1 + team;
    const u32 tq = tmem_base + ((32u * ((u32)warp & 3u)) << 16);   // KL == 3: this warp's lane quadrant of tensor memory
    if (GROUPS > 1 && grp) __nanosleep(a.stagger_ns * grp);   // start the groups out of phase
    const bool lag_on = LAG && live_teams == (u32)TEAMS && TEAMS >= 2;
    const bool trailer = lag_on && team >= TEAMS / 2, leader = lag_on && team < TEAMS / 2;
    const u32 lag_x = a.stagger_ns / 10, lag_y = a.stagger_ns % 10;
    constexpr u32 LAG_THREADS = 64u * TEAMS;
    if (LAG == 1 && trailer) {
        const long long c0 = clock64();
        while (clock64() - c0 < (long long)a.stagger_ns) {
        }
    }
#define FFT64_LAG_POINT(k)                                                                                                   \
    do {                                                                                                                     \
        if (LAG == 2) {                                                                                                      \
            if (leader && lag_x == (k)) asm volatile("bar.arrive 13, %0;" ::"r"(LAG_THREADS) : "memory");                    \
            if (trailer && lag_y == (k) && it + 1 < a.steps) asm volatile("bar.arrive 14, %0;" ::"r"(LAG_THREADS) : "memory"); \
        }                                                                                                                    \
    } while (0)
    long long mark_ = a.trace ? clock64() : 0;
    // MK bit 2: the older decomposition arithmetic (the three-ciphertext kernel, whose six warps leave the ALU pipe slack, is
    // 3 % faster with it: profiles/r02d_mark_sweep.txt)
    constexpr bool OLD_DEC = (MK & 4) != 0;
    // MK bit 4: the update reads the old accumulator value from TENSOR MEMORY instead of shared memory (a mirror of this lane's 64
    // coefficients: 128 columns per warp, the two warps of a lane quadrant side by side; written back with tcgen05.st): shared
    // memory then only sees the store that the next iteration's rotated gather needs, 64 LDS.64 per lane and iteration fewer
    constexpr bool TMACC = (MK & 16) != 0;
    const u32 tacc = tmem_base + ((32u * ((u32)warp & 3u)) << 16) + ((u32)warp >= 4u ? 128u : 0u);
    const u32 dg_half = 1u << (31 - a.logB), dg_sh = 32 - a.logB;
    const u64 dg_off = (u64)(0x80000000u - dg_half) << 32;                                   // folded into the subtraction below
    const double dg_magic = 4503599627370496.0 + (double)(1u << (31 - dg_sh)) - 1.0;         // 2^52 + 2^(31 - sh) - 1
    const unsigned char *accb = (const unsigned char *)acc;
    // own[2 i + h] = acc[lane + 32 (i + 32 h)] - dg_off: this lane's coefficients stay in registers from the update of one iteration
    // to the decomposition of the next (they take the place of x[], which is dead in between)
    u64 own[64];
#pragma unroll
    for (int k = 0; k < 64; k++) own[k] = acc[lane + 32 * ((k >> 1) + 32 * (k & 1))] - (OLD_DEC ? (u64)0 : dg_off);
    if (TMACC) {
#pragma unroll
        for (int c = 0; c < 8; c++) {
            u32 tv[16];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const u64 v = acc[lane + 32 * (((8 * c + k) >> 1) + 32 * (k & 1))];
                tv[2 * k] = (u32)v;
                tv[2 * k + 1] = (u32)(v >> 32);
            }
            tmem_st16(tacc + 16u * c, tv);
        }
        tmem_wait_st();
    }
#pragma unroll 1
    for (u32 it = 0; it < a.steps; it++) {
        // KL == 2: no ring, but the live consumer warps of the CTA meet once per iteration: they then ask for the same key
        // rows at about the same time (one L2 request, L1 hits for the rest) and keep sharing instruction fetches
        if (KL == 2 || KL == 3 || KL == 5) asm volatile("bar.sync 15, %0;" ::"r"(64u * live_teams) : "memory");
        if (LAG == 2) {
            if (trailer) asm volatile("bar.sync 13, %0;" ::"r"(LAG_THREADS) : "memory");
            if (leader && it) asm volatile("bar.sync 14, %0;" ::"r"(LAG_THREADS) : "memory");
        }
        const u32 sh = mod_switch(mask);
        if (it + 1 < a.steps) mask = lwe[it + 1];
        double2 x[32];
        // digits of X^sh acc_r - acc_r, folded: z[q] = (d[q], d[q + M]), q = lane + 32 i.
        // Source of coefficient lane + 32 row: with sh = 32 a + b, coefficient lp + 32 ((row - ap) mod 128) of the
        // 4096-periodic signed extension, lp = (lane - b) mod 32, ap = a + [lane < b]; rows >= 64 are the negated polynomial.
        {
            const u32 b = sh & 31u;
            const u32 lp8 = (((u32)lane - b) & 31u) << 3;
            if (OLD_DEC) {   // the form of the first half of round 2: 12 ALU instructions per coefficient
                const u32 ap8 = (((sh >> 5) + ((u32)lane < b ? 1u : 0u)) & 127u) << 8;
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    double dd[2];
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const u32 tt = (u32)((i + 32 * h) << 8) - ap8;
                        const u64 v = *(const u64 *)(accb + ((tt & 0x3F00u) | lp8));
                        const u32 c = (tt >> 14) & 1u;
                        const u64 d = (v ^ ((u64)0 - (u64)c)) + c - own[2 * i + h];
                        dd[h] = digit1_double((u32)(d >> 32), dg_half, dg_sh);
                    }
                    x[i] = make_double2(dd[0], dd[1]);
                }
            } else {
            // the row counter lives in the TOP seven bits of a word (row << 25): its wrap mod 128 is the word's own, bit 31 says
            // "negated half", so one arithmetic shift gives the negation mask (0 / -1) and (v ^ m) - m is the conditional negation
            const u32 ap25 = ((sh >> 5) + ((u32)lane < b ? 1u : 0u)) << 25;
#pragma unroll
            for (int i = 0; i < 32; i++) {
                double dd[2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const u32 tt = ((u32)(i + 32 * h) << 25) - ap25;
                    const u64 v = *(const u64 *)(accb + (((tt >> 17) & 0x3F00u) | lp8));
                    const u32 m32 = (u32)((int)tt >> 31);
                    const u64 m = ((u64)m32 << 32) | m32;
                    const u64 d = (v ^ m) - m - own[2 * i + h];   // own[] already carries -dg_off
                    dd[h] = digit1_double_off((u32)(d >> 32), dg_sh, dg_magic);
                }
                x[i] = make_double2(dd[0], dd[1]);
            }
            }
        }
        FFT64_MARK(0);
        FFT64_LAG_POINT(0);
        pass_a<false, MK>(x, a, lane, mark_);
        FFT64_MARK(1);
        FFT64_LAG_POINT(1);
        if (HB)
            col_to_row_h(x, (double *)buf, lane);
        else
            col_to_row(x, buf, lane);
        FFT64_MARK(2);
        // MK bit 3 (ring kernels with half-size buffers, four chunks per half): forward stages 1..4 of rows 16..31 are issued inside
        // the first half of the key MAC and inverse stages 4..1 of rows 0..15 inside the second half (mac_rows, IL)
        constexpr bool ILV = (MK & 8) != 0 && HB == 1 && RS == 4 && (KL == 0 || KL == 4);
        if constexpr (ILV) {
            pass_b_stage<false, 0, 0, 16, MK>(x, twB, lane);
            pass_b_stage<false, 1, 0, 8, MK>(x, twB, lane);
            pass_b_stage<false, 2, 0, 8, MK>(x, twB, lane);
            pass_b_stage<false, 3, 0, 8, MK>(x, twB, lane);
            pass_b_stage<false, 4, 0, 8, MK>(x, twB, lane);
        } else {
            pass_b<false, MK>(x, twB, lane, a, mark_);
        }
        FFT64_MARK(3);
        FFT64_LAG_POINT(3);
        if (KL == 3) {   // this iteration's key rows sit in buffer it & 1 of the quadrant (stored during the previous iteration)
            mbar_wait(tfull + 2 * ((u32)warp & 3u) + (it & 1u), (it >> 1) & 1u);
            tmem_fence_after();
        }
        if (KL == 5 && (a.stagger_ns != 7777u || it < 2)) {   // 7777: timing experiment, the refill stops after two GGSWs (wrong results)
            mbar_wait(empty_all + NS + (it & 1u), (it >> 1) & 1u);
            tmem_fence_after();
        }
        // swap spectra with the partner warp
        if (!HB) {
#pragma unroll
            for (int i = 0; i < 32; i++) buf[lane * ROWP + i] = x[i];
            asm volatile("bar.sync %0, 64;" ::"r"(barrier_id) : "memory");
            FFT64_MARK(4);
            // rows 0..15, then the inverse butterflies that only involve them while the key rows of the second half
            // stream in (the ring holds half a GGSW), then rows 16..31
            constexpr int CH = 16 / RS;
            if (r == 0)
                ((KL == 3 || KL == 5) ? mac_tmem<0, 0, (0) * RS / 2, (CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<0, RS, 0, 0, CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<0, NS, RS, 0, 0, CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            else
                ((KL == 3 || KL == 5) ? mac_tmem<1, 0, (0) * RS / 2, (CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<1, RS, 0, 0, CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<1, NS, RS, 0, 0, CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            pass_b_inv_half<0, MK>(x, twB, lane, a, mark_);
            if (r == 0)
                ((KL == 3 || KL == 5) ? mac_tmem<0, 0, (CH) * RS / 2, (2 * CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<0, RS, 0, CH, 2 * CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<0, NS, RS, 0, CH, 2 * CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            else
                ((KL == 3 || KL == 5) ? mac_tmem<1, 0, (CH) * RS / 2, (2 * CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<1, RS, 0, CH, 2 * CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<1, NS, RS, 0, CH, 2 * CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            FFT64_MARK(5);
            asm volatile("bar.sync %0, 64;" ::"r"(barrier_id) : "memory");
            FFT64_MARK(6);
            FFT64_LAG_POINT(6);
            pass_b_inv_half<1, MK>(x, twB, lane, a, mark_);
        } else {   // two rounds of 16 rows through the half-size buffers, layout [row][lane]
            constexpr int CH = 16 / RS;
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 16; i++) buf[i * 32 + lane] = x[i];
            asm volatile("bar.sync %0, 64;" ::"r"(barrier_id) : "memory");
            FFT64_MARK(4);
            if constexpr (ILV) {
                if (r == 0)
                    mac_all<0, NS, RS, 1, 0, CH, 1, MK>(x, buf_oth, ring, full, empty, t, lane, fdp, twB);
                else
                    mac_all<1, NS, RS, 1, 0, CH, 1, MK>(x, buf_oth, ring, full, empty, t, lane, fdp, twB);
            } else
            if (r == 0)
                ((KL == 3 || KL == 5) ? mac_tmem<0, 1, (0) * RS / 2, (CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<0, RS, 1, 0, CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<0, NS, RS, 1, 0, CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            else
                ((KL == 3 || KL == 5) ? mac_tmem<1, 1, (0) * RS / 2, (CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<1, RS, 1, 0, CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<1, NS, RS, 1, 0, CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            FFT64_MARK(5);
            asm volatile("bar.sync %0, 64;" ::"r"(barrier_id) : "memory");
#pragma unroll
            for (int i = 0; i < 16; i++) buf[i * 32 + lane] = x[16 + i];
            asm volatile("bar.sync %0, 64;" ::"r"(barrier_id) : "memory");
            FFT64_MARK(4);
            if constexpr (ILV) {
                if (r == 0)
                    mac_all<0, NS, RS, 1, CH, 2 * CH, 2, MK>(x, buf_oth, ring, full, empty, t, lane, fdp, twB);
                else
                    mac_all<1, NS, RS, 1, CH, 2 * CH, 2, MK>(x, buf_oth, ring, full, empty, t, lane, fdp, twB);
            } else {
            pass_b_inv_half<0, MK>(x, twB, lane, a, mark_);   // rows 0..15 are final: their inverse butterflies cover the wait for the key rows
            if (r == 0)
                ((KL == 3 || KL == 5) ? mac_tmem<0, 1, (CH) * RS / 2, (2 * CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<0, RS, 1, CH, 2 * CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<0, NS, RS, 1, CH, 2 * CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            else
                ((KL == 3 || KL == 5) ? mac_tmem<1, 1, (CH) * RS / 2, (2 * CH) * RS / 2>(x, buf_oth, tq + 256u * (it & 1u), lane)
                 : (KL && KL != 4) ? mac_global<1, RS, 1, CH, 2 * CH>(x, buf_oth, a.bsk + (size_t)it * 32 * 4 * 32, lane)
                    : mac_all<1, NS, RS, 1, CH, 2 * CH>(x, buf_oth, ring, full, empty, t, lane, fdp));
            }
            FFT64_MARK(5);
            asm volatile("bar.sync %0, 64;" ::"r"(barrier_id) : "memory");
            FFT64_MARK(6);
            FFT64_LAG_POINT(6);
            pass_b_inv_half<1, MK>(x, twB, lane, a, mark_);
        }
        if (KL == 3) {   // the buffer may be refilled (with the rows of iteration it + 2)
            tmem_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + 2 * ((u32)warp & 3u) + (it & 1u));
        }
        if (KL == 5) {
            tmem_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(empty_all + NS + 2 + (it & 1u));
        }
        pass_b_stage<true, 0, 0, 16, MK>(x, twB, lane);
        FFT64_MARK(7);
        FFT64_LAG_POINT(7);
        if (HB)
            row_to_col_h(x, (double *)buf, lane);
        else
            row_to_col(x, buf, lane);
        FFT64_MARK(8);
        FFT64_LAG_POINT(8);
        pass_a<true, MK>(x, a, lane, mark_);
        FFT64_MARK(9);
        FFT64_LAG_POINT(9);
#pragma unroll
        for (int i0 = 0; i0 < 32; i0 += 4) {   // four points = eight independent conversion chains at a time
            double t[8], f[8];
            u32 tv[16];
            if (TMACC) tmem_ld16(tacc + 4u * i0, tv);   // own indices 2 i0 .. 2 i0 + 7 = columns 4 i0 .. 4 i0 + 15
#pragma unroll
            for (int k = 0; k < 8; k++) t[k] = __dadd_rn((k & 1) ? x[i0 + k / 2].y : x[i0 + k / 2].x, MAGIC);
#pragma unroll
            for (int k = 0; k < 8; k++) t[k] = __dsub_rn(t[k], MAGIC);
#pragma unroll
            for (int k = 0; k < 8; k++) f[k] = __dsub_rn((k & 1) ? x[i0 + k / 2].y : x[i0 + k / 2].x, t[k]);
#pragma unroll
            for (int k = 0; k < 8; k++) f[k] = __fma_rn(f[k], TWO51, MAGIC);
            if (TMACC) tmem_wait_ld();
#pragma unroll
            for (int k = 0; k < 8; k++) {
                // tor_from_double: (bits - 0x4338000000000000) << 13; the constant's set bits all leave the word, so the
                // shift alone is the same value
                const u32 q = (u32)lane + 32u * (i0 + k / 2) + ((k & 1) ? M : 0);
                const u64 old = TMACC ? (((u64)tv[2 * k + 1] << 32) | tv[2 * k]) : acc[q];
                const u64 nv = old + ((u64)__double_as_longlong(f[k]) << 13);
                acc[q] = nv;
                if (TMACC) {
                    tv[2 * k] = (u32)nv;
                    tv[2 * k + 1] = (u32)(nv >> 32);
                }
                // the digit's rounding offset rides in the register copy: one add on the high word here, where shared memory binds,
                // instead of one in the ALU-bound decomposition (the empty asm keeps ptxas from sinking it back there)
                u64 ov = nv;
                if (!OLD_DEC) {
                    ov = nv - dg_off;
                    asm volatile("" : "+l"(ov));
                }
                own[2 * (i0 + k / 2) + (k & 1)] = ov;
            }
            if (TMACC) tmem_st16(tacc + 4u * i0, tv);
        }
        if (TMACC) tmem_wait_st();
        __syncwarp();
        FFT64_MARK(10);
    }

    if (a.out_mode == 2) {
        u64 *o = a.out + (size_t)ct * 2 * N + (size_t)r * N;
        for (int p = lane; p < N; p += 32) o[p] = acc[p];
    } else {   // sample extraction of coefficient 0      crypto/glwe.rs:96-106
        u64 *o = a.out + (size_t)ct * (N + 1);
        if (r == 0) {
            for (int p = lane; p < N; p += 32) o[p] = p ? (u64)0 - acc[N - p] : acc[0];
        } else if (lane == 0) {
            o[N] = acc[0] + a.body_add;
        }
    }
}

// ------------------------------------------------------------------------------------------ utility kernels
// One warp per polynomial: forward / inverse transform of folded complex data (parity probe), in the oracle's index order.
__device__ __forceinline__ void fft_probe_body(const double2 *in, double2 *out, int inverse, double2 *buf, const double2 *twB, const Args &a, int lane) {
    double2 x[32];
    if (!inverse) {
#pragma unroll
        for (int i = 0; i < 32; i++) x[i] = in[lane + 32 * i];
        fft_forward(x, buf, twB, a, lane);
#pragma unroll
        for (int i = 0; i < 32; i++) out[32 * lane + i] = x[i];
    } else {
#pragma unroll
        for (int i = 0; i < 32; i++) x[i] = in[32 * lane + i];
        fft_inverse(x, buf, twB, a, lane);
#pragma unroll
        for (int i = 0; i < 32; i++) out[lane + 32 * i] = x[i];
    }
}

// key polynomial (coefficient domain, torus words) -> Fourier row in the ring layout   oracle/torus.c tor_ggsw_to_fourier
// poly index pi = ((ggsw * 2 + j) * ell + lvl) * 2 + r with ell = 1; destination [ggsw][row i][j*2 + r][lane]
__device__ __forceinline__ void key_to_fourier_body(const u64 *src, double2 *dst_ggsw, int poly_in_ggsw, double scale, double2 *buf, const double2 *twB,
                                                    const Args &a, int lane) {
    double2 x[32];
#pragma unroll
    for (int i = 0; i < 32; i++) {
        const int q = lane + 32 * i;
        x[i] = make_double2(__dmul_rn(__ll2double_rn((long long)src[q]), scale), __dmul_rn(__ll2double_rn((long long)src[q + M]), scale));
    }
    fft_forward(x, buf, twB, a, lane);
    const bool neg = (poly_in_ggsw >> 1) == 0;   // mask GLev j = 0 (k = 1)
#pragma unroll
    for (int i = 0; i < 32; i++) {
        double2 v = x[i];
        if (neg) v = make_double2(-v.x, -v.y);
        dst_ggsw[((size_t)i * 4 + poly_in_ggsw) * 32 + lane] = v;
    }
}

}  // namespace fft64
