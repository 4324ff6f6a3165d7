// ks_mma.cuh — batched LWE key switch as an exact int8 tensor-core contraction (sm_100a, device only).
//
// out[b] = (0,..,0,body_b) + sum_i sum_j digit_j(a_{b,i}) * KSK[i][j]          (vtfhe/mod.rs:31-54, lev_ct.rs:63-76)
//
// For a batch this is a dense contraction  D[B][K] x KSK[K][n+1]  with K = kN*ell_ks rows, signed digits
// |d| <= 2^(logB_ks-1) and 64-bit key words.  Splitting every key word into its 8 bytes turns it into an
// s8 x u8 -> s32 GEMM  D[B][K] x KB[K][8(n+1)]: the byte-plane sums are exact in int32 as long as
// K * 2^(logB_ks-1) * 255 < 2^31 (S2: 8192 * 8 * 255 = 1.7e7), and
//     out[b][c] = sum_{byte} S[b][8c+byte] * 2^(8*byte)   (mod q)
// is recombined in the epilogue, so the result is the same element of Z_q the reference computes.
// The north-star rule for tensor cores ("only where ncu shows the dense contraction beats the IMAD pipe") is
// met by this one op: 20.5 ms on the integer pipes (lwe_keyswitch_kernel) vs the figure in profiles/ for this kernel.
//
// Operands in HBM:
//   dig  [Mpad][K]  s8, K contiguous   (ks_digits_kernel, per call; rows >= B are zero)
//   kskT [Npad][K]  u8, K contiguous   (ks_bytes_transpose_kernel, once per key load; kskT[8c+byte][r] = byte of KSK[r][c])
// Both are "K-major", i.e. the row.col form of mma.sync.m16n8k32 with plain (non-transposing) ldmatrix.
// This is the mma.sync form: since round 2 the default is the tcgen05 / tensor-memory kernel in ks_tc5.cuh (3.1x faster at
// B = 4,440); this one stays as its checker (lux_debug_ks_mode 1) and as the kernel of lux_pbs_batch_device_static.
#pragma once
#include "gl64.cuh"

namespace ksmma {

constexpr int BM = 128, BN = 128, BK = 128, STAGES = 3, THREADS = 256;
constexpr int C_STRIDE = 136;  // int32 words per row of the epilogue tile (conflict-free 64-bit fragment stores)
constexpr size_t SMEM_BYTES = (size_t)STAGES * (BM + BN) * BK;  // 96 KiB >= BM * C_STRIDE * 4 (68 KiB)

struct Args {
    const int8_t *dig;    // [Mpad][K]
    const uint8_t *kskT;  // [Npad][K]
    const u64 *in;        // [B][kN+1]   (for the body)
    u64 *out;             // [B][n+1]
    u32 B, n, kN, K;
    u32 torus;            // 0: words are Goldilocks elements; 1: words are 2^64-torus elements (wrapping recombination)
};

// digits of the mask coefficients as signed bytes: dig[b][i*ell + j] = digit_{L-ell+j}(in[b][i])
// torus != 0: tfhe-rs SignedDecomposer digits of the top ell*logB bits instead (oracle/torus.c tor_decompose), least significant level first
__global__ void ks_digits_kernel(const u64 *in, int8_t *dig, u32 B, u32 Mpad, u32 kN, u32 ell, u32 logB, u32 L, u32 torus) {
    const bool centred = (L * logB == 64);
    const size_t total = (size_t)Mpad * kN;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const u32 b = (u32)(idx / kN), i = (u32)(idx % kN);
        const u64 x = b < B ? in[(size_t)b * (kN + 1) + i] : 0;
        int8_t *o = dig + idx * ell;
        if (torus) {
            const u32 rep = ell * logB;
            u64 state = rep >= 64 ? x : (((x >> (63 - rep)) + 1) >> 1) & ((1ull << rep) - 1);
            const u64 mask = (1ull << logB) - 1;
            for (u32 j = 0; j < ell; j++) {
                const u64 res = state & mask;
                state >>= logB;
                const u64 carry = (((res - 1ull) | state) & res) >> (logB - 1);
                state += carry;
                o[j] = (int8_t)((long long)res - (long long)(carry << logB));
            }
            continue;
        }
        for (u32 j = 0; j < ell; j++) {
            const u64 d = gl_digit(x, L - ell + j, logB, centred);  // canonical field element, |d| <= 2^(logB-1)
            o[j] = (d > GL_Q / 2) ? (int8_t)(-(int)(u32)(GL_Q - d)) : (int8_t)(u32)d;
        }
    }
}

// kskT[c8][r] = byte (c8 & 7) of ksk[r][c8 >> 3]; zero for padded columns.  One thread per (c8, r), r fastest.
__global__ void ks_bytes_transpose_kernel(const u64 *ksk, uint8_t *kskT, u32 K, u32 cols, u32 Npad) {
    const size_t total = (size_t)Npad * K;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const u32 c8 = (u32)(idx / K), r = (u32)(idx % K);
        const u32 c = c8 >> 3, byte = c8 & 7;
        kskT[idx] = c < cols ? (uint8_t)(ksk[(size_t)r * cols + c] >> (8 * byte)) : (uint8_t)0;
    }
}

__device__ __forceinline__ u32 smem_u32(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(u32 dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldmatrix_x4(u32 addr, u32 &r0, u32 &r1, u32 &r2, u32 &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_s8u8(int (&c)[4], u32 a0, u32 a1, u32 a2, u32 a3, u32 b0, u32 b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// tile rows are BK = 128 bytes = 8 chunks of 16 bytes; chunk c of row r sits at chunk c ^ (r & 7)
__device__ __forceinline__ u32 tile_off(int row, int chunk) { return (u32)(row * BK + ((chunk ^ (row & 7)) << 4)); }

// signed byte-plane sum (|x| < 2^50) -> canonical field element
__device__ __forceinline__ u64 signed_to_field(long long x) { return x < 0 ? GL_Q + (u64)x : (u64)x; }

__global__ void __launch_bounds__(THREADS, 2) lwe_keyswitch_mma_kernel(const Args a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 1, wn = warp >> 1;  // 2 x 4 warps, warp tile 64 (M) x 32 (N)
    const u32 m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int8_t *gA = a.dig + (size_t)m0 * a.K;
    const uint8_t *gB = a.kskT + (size_t)n0 * a.K;
    const u32 sbase = smem_u32(smem);
    const u32 KT = a.K / BK;

    auto load_stage = [&](u32 kt, int s) {
        const u32 sA = sbase + (u32)s * (BM + BN) * BK, sB = sA + BM * BK;
        const size_t k0 = (size_t)kt * BK;
#pragma unroll
        for (int it = 0; it < (BM * 8) / THREADS; it++) {
            const int idx = tid + it * THREADS, row = idx >> 3, ch = idx & 7;
            cp_async16(sA + tile_off(row, ch), gA + (size_t)row * a.K + k0 + ch * 16);
        }
#pragma unroll
        for (int it = 0; it < (BN * 8) / THREADS; it++) {
            const int idx = tid + it * THREADS, row = idx >> 3, ch = idx & 7;
            cp_async16(sB + tile_off(row, ch), gB + (size_t)row * a.K + k0 + ch * 16);
        }
    };

    int acc[4][4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int r = 0; r < 4; r++) acc[i][j][r] = 0;

#pragma unroll
    for (int s = 0; s < STAGES - 1; s++) {
        if ((u32)s < KT) load_stage(s, s);
        cp_async_commit();
    }

    const int mat = lane >> 3, rr = lane & 7;
    for (u32 kt = 0; kt < KT; kt++) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        // refill the stage consumed in the previous iteration
        if (kt + STAGES - 1 < KT) load_stage(kt + STAGES - 1, (int)((kt + STAGES - 1) % STAGES));
        cp_async_commit();
        const int s = (int)(kt % STAGES);
        const u32 sA = sbase + (u32)s * (BM + BN) * BK, sB = sA + BM * BK;
#pragma unroll
        for (int kk = 0; kk < BK / 32; kk++) {
            u32 af[4][4], bf[4][2];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int row = wm * 64 + i * 16 + rr + (mat & 1) * 8, ch = kk * 2 + (mat >> 1);
                ldmatrix_x4(sA + tile_off(row, ch), af[i][0], af[i][1], af[i][2], af[i][3]);
            }
#pragma unroll
            for (int jp = 0; jp < 2; jp++) {
                const int row = wn * 32 + jp * 16 + rr + (mat >> 1) * 8, ch = kk * 2 + (mat & 1);
                ldmatrix_x4(sB + tile_off(row, ch), bf[2 * jp][0], bf[2 * jp][1], bf[2 * jp + 1][0], bf[2 * jp + 1][1]);
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) mma_s8u8(acc[i][j], af[i][0], af[i][1], af[i][2], af[i][3], bf[j][0], bf[j][1]);
        }
    }
    cp_async_wait<0>();
    __syncthreads();  // every warp is done with the operand stages: reuse them for the int32 tile

    int *C = (int *)smem;
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int row = wm * 64 + i * 16 + g, col = wn * 32 + j * 8 + 2 * t;
            *reinterpret_cast<int2 *>(C + row * C_STRIDE + col) = make_int2(acc[i][j][0], acc[i][j][1]);
            *reinterpret_cast<int2 *>(C + (row + 8) * C_STRIDE + col) = make_int2(acc[i][j][2], acc[i][j][3]);
        }
    __syncthreads();

    // one output word = the 8 byte planes of one key column
    const u32 cols = a.n + 1;
    for (int idx = tid; idx < BM * (BN / 8); idx += THREADS) {
        const int row = idx / (BN / 8), c = idx % (BN / 8);
        const u32 b = m0 + row, col = n0 / 8 + c;
        if (b >= a.B || col >= cols) continue;
        const int4 lo4 = *reinterpret_cast<const int4 *>(C + row * C_STRIDE + c * 8);
        const int4 hi4 = *reinterpret_cast<const int4 *>(C + row * C_STRIDE + c * 8 + 4);
        const long long P = (long long)lo4.x + ((long long)lo4.y << 8) + ((long long)lo4.z << 16) + ((long long)lo4.w << 24);
        const long long Q = (long long)hi4.x + ((long long)hi4.y << 8) + ((long long)hi4.z << 16) + ((long long)hi4.w << 24);
        u64 v;
        if (a.torus) {
            v = (u64)P + ((u64)Q << 32);
            if (col == a.n) v += a.in[(size_t)b * (a.kN + 1) + a.kN];
        } else {
            v = gl_canon(gl_add(gl_mul_lazy(signed_to_field(Q), 1ULL << 32), signed_to_field(P)));
            if (col == a.n) v = gl_canon(gl_add(v, a.in[(size_t)b * (a.kN + 1) + a.kN]));
        }
        a.out[(size_t)b * cols + col] = v;
    }
}

}  // namespace ksmma
