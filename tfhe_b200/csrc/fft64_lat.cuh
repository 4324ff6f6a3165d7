// fft64_lat.cuh — latency form of the torus blind rotation: ONE ciphertext per CTA, eight warps (sm_100a).
//
// The throughput kernel (fft64.cuh) gives a polynomial to one warp, 32 points per lane: a lone warp then walks a chain of
// ~5,000 dependent-ish instructions per CMUX (5.4 ms per bootstrap).  A batch that does not fill the GPU (B <= #SMs: a single
// bootstrap, the levels of a small radix circuit) wants the opposite trade: here 128 threads share one polynomial, 8 points
// per thread, and the 10 tree levels run as passes of 3 + 3 + 3 + 1 levels with the points exchanged through shared memory
// in between (one ciphertext per SM: the crossbar is idle anyway).  The arithmetic is the throughput kernel's and the
// oracle's, operation for operation (same butterflies, same twiddle doubles, same MAC chain, same conversions), so the
// results are bit-identical; only who computes which butterfly differs.
//
// Point index q = b9..b0 (in-place index of oracle/torus.c).  Level s pairs q and q + (512 >> s) with twiddle node
// 2^s + (q >> (10 - s)).  A thread holds 8 points that differ in 3 "register bits"; t = thread within the polynomial's group.
//   forward  F1  q = 128 reg + t                     levels 0,1,2   (bits 9,8,7)      twiddles uniform over the group
//            F2  q = 128 (t >> 4) + 16 reg + (t & 15)  levels 3,4,5  (bits 6,5,4)
//            F3  q = 16 (t >> 1) + 2 reg + (t & 1)     levels 6,7,8  (bits 3,2,1)
//   swap: both groups publish their spectrum, every thread then holds X0 and X1 at
//            M   q = 32 (t & 31) + 8 (t >> 5) + reg    level 9 on both (bit 0), key MAC, inverse levels 9,8,7 (bits 0,1,2)
//            the key row of point q is row q % 32 = 8 (t >> 5) + reg at lane q / 32 = t & 31: coalesced 512-byte loads,
//            issued at the top of the iteration
//   inverse  I2  q = 64 (t >> 3) + 8 reg + (t & 7)     levels 6,5,4  (bits 3,4,5)
//            I3  q = 512 (t >> 6) + 64 reg + (t & 63)  levels 3,2,1  (bits 6,7,8)
//            I4  q = 512 (reg & 1) + 128 (reg >> 1) + t  level 0 (bit 9), then the update: the same 8 points per thread as F1
#pragma once
#include "fft64.cuh"

namespace fft64 {

constexpr int LAT_THREADS = 256;
constexpr int LAT_PAD_POINTS = M;   // exchange buffer: no padding, XOR-swizzled (lat_pad)
constexpr int LAT_TWM = 128 + 256 + 512;   // levels 7, 8, 9 of the twiddle tree once more, lane-major for layout M
constexpr size_t LAT_SMEM = sizeof(u64) * 2 * N + sizeof(double2) * (size_t)(M + LAT_TWM + 6 * LAT_PAD_POINTS) + 16;

struct LatArgs {
    Args a;
    const double2 *tw_full;   // [M] the whole twiddle tree (node v at index v), device memory
};

// position of point q in an exchange buffer: the low three bits (which 16-byte slot of a 128-byte line) are XORed with bits
// 3..5 and 5..7, which makes every layout below conflict-free for 16-byte accesses (checked exhaustively: each quarter-warp
// of each access covers eight distinct slots)
__device__ __forceinline__ int lat_pad(int q) { return q ^ (((q >> 3) ^ (q >> 5)) & 7); }

// three in-register levels on 8 points: level s0 + j pairs register bit BIT[j]; NODE(j, reg) = twiddle node of the pair
template <bool INV, typename NodeFn>
__device__ __forceinline__ void lat_pass3(double2 (&x)[8], const double2 *tw, const int (&bit)[3], NodeFn node) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
        const int b = bit[j];
#pragma unroll
        for (int reg = 0; reg < 8; reg++) {
            if (reg & (1 << b)) continue;
            const double2 w = tw[node(j, reg)];
            bfly<INV>(x[reg], x[reg | (1 << b)], w.x, w.y);
        }
    }
}

__device__ __forceinline__ void group_sync(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

// V bit 0: the thread's own 16 accumulator coefficients stay in registers (F1 and I4 cover the same points)
// V bit 1: the key rows are requested after the second forward pass instead of at the top of the iteration
// V bit 2: phase marks — lane 0 of every warp adds the cycles since its previous mark to a.trace[phase] (lux_debug_fft_trace)
#define LAT_MARK(k)                                                                              \
    do {                                                                                         \
        if ((V & 4) && a.trace && lane == 0) {                                                   \
            const long long now_ = clock64();                                                    \
            atomicAdd((unsigned long long *)a.trace + (k), (unsigned long long)(now_ - mark_));  \
            mark_ = now_;                                                                        \
        }                                                                                        \
    } while (0)
template <int V>
__device__ __forceinline__ void blind_rotate_lat_body(const LatArgs &la, unsigned char *smem_raw) {
    const Args &a = la.a;
    const int tid = threadIdx.x, r = tid >> 7, t = tid & 127, lane = tid & 31, w4 = t >> 5;
    u64 *acc_all = (u64 *)smem_raw;
    double2 *tw = (double2 *)(acc_all + 2 * N);
    double2 *twM = tw + M;                                    // level L = 7..9: node 2^L + lane * 2^(L-5) + rest at (2^L - 128) + 32 rest + lane
    double2 *bufs = twM + LAT_TWM;                            // [group][3][LAT_PAD_POINTS]
    double2 *b0 = bufs + (size_t)(r * 3 + 0) * LAT_PAD_POINTS, *b1 = bufs + (size_t)(r * 3 + 1) * LAT_PAD_POINTS,
            *bx = bufs + (size_t)(r * 3 + 2) * LAT_PAD_POINTS;
    const double2 *bx0 = bufs + (size_t)2 * LAT_PAD_POINTS, *bx1 = bufs + (size_t)5 * LAT_PAD_POINTS;
    const u32 ct = a.ct0 + blockIdx.x;
    u64 *acc = acc_all + (size_t)r * N;
    const u64 *lwe = a.lwe + (size_t)ct * (a.n + 1);
    const u64 *tv = a.testv + (size_t)(a.lut_idx ? a.lut_idx[ct / a.lut_div] : 0) * N;

    for (int i = tid; i < M; i += LAT_THREADS) tw[i] = la.tw_full[i];
    // layout M reads node 2^L + lane * 2^(L-5) + rest: in tree order the 32 lanes of a load sit 2^(L-5) * 16 bytes apart (an
    // 8-way bank conflict at level 9); the copy below puts them side by side
    for (int i = tid; i < LAT_TWM; i += LAT_THREADS) {
        const int L = i < 128 ? 7 : i < 384 ? 8 : 9, o = i - ((1 << L) - 128), ln = o & 31, rest = o >> 5;
        twM[i] = la.tw_full[(1 << L) + (ln << (L - 5)) + rest];
    }
    {   // acc0 = X^{ms(-b)} (0, testv)
        const u32 sh0 = mod_switch((u64)0 - lwe[a.n]);
        for (int p = t; p < N; p += 128) acc[p] = r ? rotated(tv, (u32)p, sh0) : 0;
    }
    __syncthreads();

    const u32 dg_half = 1u << (31 - a.logB), dg_sh = 32 - a.logB;
    u64 mask = a.steps ? lwe[0] : 0;
    u64 own[16];   // V & 1: acc[128 reg + t] and acc[128 reg + t + M], reg = 0..7
    if (V & 1) {
#pragma unroll
        for (int reg = 0; reg < 8; reg++) {
            own[2 * reg] = acc[128 * reg + t];
            own[2 * reg + 1] = acc[128 * reg + t + M];
        }
    }
    long long mark_ = (V & 4) ? clock64() : 0;
#pragma unroll 1
    for (u32 it = 0; it < a.steps; it++) {
        const u32 sh = mod_switch(mask);
        if (it + 1 < a.steps) mask = lwe[it + 1];
        // key rows of this thread's MAC points (layout M), requested early, used after the forward transform
        double2 k0[8], k1[8];
        auto request_keys = [&]() {
            const double2 *g = a.bsk + (size_t)it * 32 * 4 * 32;
#pragma unroll
            for (int reg = 0; reg < 8; reg++) {
                const int row = 8 * w4 + reg;
                k0[reg] = __ldg(g + (row * 4 + 0 + r) * 32 + lane);
                k1[reg] = __ldg(g + (row * 4 + 2 + r) * 32 + lane);
            }
        };
        if (!(V & 2)) request_keys();
        double2 x[8];
        // F1 points q = 128 reg + t: digits of X^sh acc - acc at coefficients q and q + M
#pragma unroll
        for (int reg = 0; reg < 8; reg++) {
            const u32 q = 128u * reg + (u32)t;
            const u64 d0 = rotated(acc, q, sh) - ((V & 1) ? own[2 * reg] : acc[q]);
            const u64 d1 = rotated(acc, q + M, sh) - ((V & 1) ? own[2 * reg + 1] : acc[q + M]);
            x[reg] = make_double2(digit1_double((u32)(d0 >> 32), dg_half, dg_sh), digit1_double((u32)(d1 >> 32), dg_half, dg_sh));
        }
        LAT_MARK(0);
        {   // levels 0,1,2: register bits 2,1,0 <-> index bits 9,8,7; node = 2^j + (reg >> (3 - j))
            const int bit[3] = {2, 1, 0};
            lat_pass3<false>(x, tw, bit, [](int j, int reg) { return (1 << j) + (reg >> (3 - j)); });
        }
#pragma unroll
        for (int reg = 0; reg < 8; reg++) b0[lat_pad(128 * reg + t)] = x[reg];
        LAT_MARK(1);
        group_sync(1 + r);
        LAT_MARK(2);
        {   // F2: q = 128 hi + 16 reg + lo
            const int hi = t >> 4, lo = t & 15;
#pragma unroll
            for (int reg = 0; reg < 8; reg++) x[reg] = b0[lat_pad(128 * hi + 16 * reg + lo)];
            const int bit[3] = {2, 1, 0};
            lat_pass3<false>(x, tw, bit, [hi](int j, int reg) { return (8 << j) + (hi << j) + (reg >> (3 - j)); });
#pragma unroll
            for (int reg = 0; reg < 8; reg++) b1[lat_pad(128 * hi + 16 * reg + lo)] = x[reg];
        }
        LAT_MARK(3);
        group_sync(1 + r);
        LAT_MARK(4);
        if (V & 2) request_keys();
        {   // F3: q = 16 hi + 2 reg + lo
            const int hi = t >> 1, lo = t & 1;
#pragma unroll
            for (int reg = 0; reg < 8; reg++) x[reg] = b1[lat_pad(16 * hi + 2 * reg + lo)];
            const int bit[3] = {2, 1, 0};
            lat_pass3<false>(x, tw, bit, [hi](int j, int reg) { return (64 << j) + (hi << j) + (reg >> (3 - j)); });
#pragma unroll
            for (int reg = 0; reg < 8; reg++) bx[lat_pad(16 * hi + 2 * reg + lo)] = x[reg];
        }
        LAT_MARK(5);
        __syncthreads();   // both spectra (before their last level) are published
        LAT_MARK(6);
        {   // layout M: q = 32 lane + 8 w4 + reg; level 9 on X0 and X1, MAC with key column r, inverse levels 9, 8, 7
            double2 x0[8], x1[8];
            const int qb = 32 * lane + 8 * w4;
#pragma unroll
            for (int reg = 0; reg < 8; reg++) {
                x0[reg] = bx0[lat_pad(qb + reg)];
                x1[reg] = bx1[lat_pad(qb + reg)];
            }
#pragma unroll
            for (int reg = 0; reg < 8; reg += 2) {
                const double2 wv = twM[384 + 32 * (4 * w4 + (reg >> 1)) + lane];
                ct_bfly(x0[reg], x0[reg + 1], wv.x, wv.y);
                ct_bfly(x1[reg], x1[reg + 1], wv.x, wv.y);
            }
#pragma unroll
            for (int reg = 0; reg < 8; reg++) {   // the oracle's chain: x0 terms first, then x1 terms
                double re = __fma_rn(x0[reg].x, k0[reg].x, 0.0), im = __fma_rn(x0[reg].x, k0[reg].y, 0.0);
                re = __fma_rn(-x0[reg].y, k0[reg].y, re);
                im = __fma_rn(x0[reg].y, k0[reg].x, im);
                re = __fma_rn(x1[reg].x, k1[reg].x, re);
                im = __fma_rn(x1[reg].x, k1[reg].y, im);
                re = __fma_rn(-x1[reg].y, k1[reg].y, re);
                im = __fma_rn(x1[reg].y, k1[reg].x, im);
                x[reg] = make_double2(re, im);
            }
            const int bit[3] = {0, 1, 2};   // levels 9, 8, 7 <-> register bits 0, 1, 2
            // node (512 >> j) + ((qb + reg) >> (1 + j)) of the tree, from the lane-major copy
            lat_pass3<true>(x, twM, bit, [w4, lane](int j, int reg) { return ((512 >> j) - 128) + 32 * (((8 * w4 + reg) >> (1 + j))) + lane; });
#pragma unroll
            for (int reg = 0; reg < 8; reg++) b0[lat_pad(qb + reg)] = x[reg];
        }
        LAT_MARK(7);
        __syncthreads();   // also: every read of the published spectra is done before bx is written again
        LAT_MARK(8);
        {   // I2: q = 64 hi + 8 reg + lo, levels 6, 5, 4 <-> register bits 0, 1, 2
            const int hi = t >> 3, lo = t & 7;
#pragma unroll
            for (int reg = 0; reg < 8; reg++) x[reg] = b0[lat_pad(64 * hi + 8 * reg + lo)];
            const int bit[3] = {0, 1, 2};
            lat_pass3<true>(x, tw, bit, [hi](int j, int reg) { return (64 >> j) + ((8 * hi + reg) >> (1 + j)); });
#pragma unroll
            for (int reg = 0; reg < 8; reg++) b1[lat_pad(64 * hi + 8 * reg + lo)] = x[reg];
        }
        LAT_MARK(9);
        group_sync(1 + r);
        LAT_MARK(10);
        {   // I3: q = 512 hi + 64 reg + lo, levels 3, 2, 1 <-> register bits 0, 1, 2
            const int hi = t >> 6, lo = t & 63;
#pragma unroll
            for (int reg = 0; reg < 8; reg++) x[reg] = b1[lat_pad(512 * hi + 64 * reg + lo)];
            const int bit[3] = {0, 1, 2};
            lat_pass3<true>(x, tw, bit, [hi](int j, int reg) { return (8 >> j) + ((8 * hi + reg) >> (1 + j)); });
#pragma unroll
            for (int reg = 0; reg < 8; reg++) b0[lat_pad(512 * hi + 64 * reg + lo)] = x[reg];
        }
        LAT_MARK(11);
        group_sync(1 + r);
        LAT_MARK(12);
        {   // I4: q = 512 (reg & 1) + 128 (reg >> 1) + t, level 0 <-> register bit 0, then acc += torus(x)
#pragma unroll
            for (int reg = 0; reg < 8; reg++) x[reg] = b0[lat_pad(512 * (reg & 1) + 128 * (reg >> 1) + t)];
            const double2 w1 = tw[1];
#pragma unroll
            for (int reg = 0; reg < 8; reg += 2) gs_bfly(x[reg], x[reg + 1], w1.x, w1.y);
#pragma unroll
            for (int reg = 0; reg < 8; reg++) {
                const u32 q = 512u * (reg & 1) + 128u * (reg >> 1) + (u32)t;
                if (V & 1) {   // point q is F1's register 4 (reg & 1) + (reg >> 1)
                    const int f = 4 * (reg & 1) + (reg >> 1);
                    own[2 * f] += from_double(x[reg].x);
                    own[2 * f + 1] += from_double(x[reg].y);
                    acc[q] = own[2 * f];
                    acc[q + M] = own[2 * f + 1];
                } else {
                    acc[q] += from_double(x[reg].x);
                    acc[q + M] += from_double(x[reg].y);
                }
            }
        }
        LAT_MARK(13);
        group_sync(1 + r);   // the next decomposition gathers rotated coefficients written by other threads of the group
        LAT_MARK(14);
    }

    if (a.out_mode == 2) {
        u64 *o = a.out + (size_t)ct * 2 * N + (size_t)r * N;
        for (int p = t; p < N; p += 128) o[p] = acc[p];
    } else {   // sample extraction of coefficient 0      crypto/glwe.rs:96-106
        u64 *o = a.out + (size_t)ct * (N + 1);
        if (r == 0) {
            for (int p = t; p < N; p += 128) o[p] = p ? (u64)0 - acc[N - p] : acc[0];
        } else if (t == 0) {
            o[N] = acc[0] + a.body_add;
        }
    }
}

}  // namespace fft64
