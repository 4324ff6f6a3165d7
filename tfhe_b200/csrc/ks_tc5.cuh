// ks_tc5.cuh — the batched LWE key switch as a Blackwell kernel: tcgen05.mma.kind::i8 with tensor-memory accumulators
// (sm_100a, device only).  Same contraction and the same exact arithmetic as ks_mma.cuh (which stays as the mma.sync checker
// form and the form of the static-scratch entry point):
//
//   out[b] = (0,..,0,body_b) + sum_i sum_j digit_j(a_{b,i}) * KSK[i][j]          (vtfhe/mod.rs:31-54, lev_ct.rs:63-76)
//   = s8 digits D[B][K] x u8 key byte planes KB[K][8(n+1)] -> s32 byte-plane sums, recombined to one word per 8 columns.
//
// Structure (one persistent CTA per SM, 192 threads, warp-specialised):
//   warp 0, one lane : producer.  Both operands sit in HBM already cut into stage tiles in the shared-memory image the tensor
//                      core wants (K-major, 128-byte rows, SWIZZLE_128B: 16-byte chunk c of row r at chunk c ^ (r & 7)), so one
//                      stage = ONE contiguous 16 KiB piece of digits + ONE contiguous 32 KiB piece of key bytes, fetched by the
//                      copy engine with cp.async.bulk onto an mbarrier (no tensor map: the tiling is done once, by the kernels
//                      that write the operands, instead of per load by the TMA unit).
//   warp 1, one lane : issues tcgen05.mma.cta_group::1.kind::i8, M = 128, N = 256, K = 32 per instruction, four per stage
//                      (descriptor start address advanced by 32 bytes inside the 128-byte swizzle atom), accumulating in one of
//                      two 256-column tensor-memory buffers; tcgen05.commit releases the stage / publishes the accumulator.
//   warps 2..5       : epilogue.  Warp w reads lane quadrant w % 4 with tcgen05.ld (lane = ciphertext row, 8 consecutive
//                      columns = the 8 byte planes of one output word), recombines and stores, while warp 1 is already
//                      multiplying the next tile into the other buffer.
// Tile 128 x 256 x 128: at S2 (B = 4,440: 35 x 29 tiles, 64 stages each) the operands cross L2 -> SM 3.2 GB per call; that, not
// the 2.7e11 MAC (0.12 ms at the dense int8 rate), bounds the kernel.
#pragma once
#include "gl64.cuh"

namespace kstc5 {

constexpr int BM = 128, BN = 256, BK = 128, NS = 4, THREADS = 192;
constexpr u32 A_BYTES = BM * BK, B_BYTES = BN * BK, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr size_t SMEM_BYTES = 1024 + (size_t)NS * STAGE_BYTES + 256;   // 1 KiB of slack to align the tiles to the swizzle period
// instruction descriptor (kind::i8): D = s32 (bits 4-5 = 2), A = s8 (bits 7-9 = 1), B = u8 (bits 10-12 = 0), both K-major,
// N >> 3 at bit 17, M >> 4 at bit 24
constexpr u32 IDESC = (2u << 4) | (1u << 7) | (0u << 10) | ((u32)(BN >> 3) << 17) | ((u32)(BM >> 4) << 24);

struct Args {
    const int8_t *digT;    // [Mpad/128][K/128][128][128 swizzled]   (ks_digits_tiled_kernel, per call)
    const uint8_t *kskT;   // [Npad/256][K/128][256][128 swizzled]   (ks_bytes_tiled_kernel, once per key load)
    const u64 *in;         // [B][kN+1]   (for the body)
    u64 *out;              // [B][n+1]
    u32 B, n, kN, K, torus, mtiles, ntiles;
};

// byte offset of (row r of the tile, byte kb of the 128-byte K slice) inside a stage tile
__host__ __device__ __forceinline__ u32 swz(u32 r, u32 kb) { return r * BK + ((((kb >> 4) ^ (r & 7u)) << 4) | (kb & 15u)); }

// digits of the mask coefficients as signed bytes, written as stage tiles: digit (b, k = i*ell + j) goes to tile
// (b / 128, k / 128).  Same digit rules as ksmma::ks_digits_kernel (Goldilocks: gl_digit of levels L-ell..L-1; torus: tfhe-rs
// SignedDecomposer, least significant level first, oracle/torus.c tor_decompose).  Rows >= B are zero.
__global__ void ks_digits_tiled_kernel(const u64 *in, int8_t *digT, u32 B, u32 Mpad, u32 kN, u32 ell, u32 logB, u32 L, u32 torus) {
    const bool centred = (L * logB == 64);
    const size_t total = (size_t)Mpad * kN;
    const u32 KT = kN * ell / BK;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const u32 b = (u32)(idx / kN), i = (u32)(idx % kN);
        const u64 x = b < B ? in[(size_t)b * (kN + 1) + i] : 0;
        int8_t *tile_row = digT + (size_t)(b >> 7) * KT * A_BYTES;
        const u32 r = b & 127u;
        u64 state = 0;
        const u64 mask = (1ull << logB) - 1;
        if (torus) {
            const u32 rep = ell * logB;
            state = rep >= 64 ? x : (((x >> (63 - rep)) + 1) >> 1) & ((1ull << rep) - 1);
        }
        for (u32 j = 0; j < ell; j++) {
            int8_t d;
            if (torus) {
                const u64 res = state & mask;
                state >>= logB;
                const u64 carry = (((res - 1ull) | state) & res) >> (logB - 1);
                state += carry;
                d = (int8_t)((long long)res - (long long)(carry << logB));
            } else {
                const u64 f = gl_digit(x, L - ell + j, logB, centred);
                d = (f > GL_Q / 2) ? (int8_t)(-(int)(u32)(GL_Q - f)) : (int8_t)(u32)f;
            }
            const u32 k = i * ell + j;
            tile_row[(size_t)(k >> 7) * A_BYTES + swz(r, k & 127u)] = d;
        }
    }
}

// key byte planes as stage tiles: byte (c8 & 7) of ksk[k][c8 >> 3] goes to tile (c8 / 256, k / 128); zero for padded columns
__global__ void ks_bytes_tiled_kernel(const u64 *ksk, uint8_t *kskT, u32 K, u32 cols, u32 Npad) {
    const size_t total = (size_t)Npad * K;
    const u32 KT = K / BK;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const u32 c8 = (u32)(idx / K), k = (u32)(idx % K);
        const u32 c = c8 >> 3, byte = c8 & 7;
        const uint8_t v = c < cols ? (uint8_t)(ksk[(size_t)k * cols + c] >> (8 * byte)) : (uint8_t)0;
        kskT[((size_t)(c8 >> 8) * KT + (k >> 7)) * B_BYTES + swz(c8 & 255u, k & 127u)] = v;
    }
}

__device__ __forceinline__ u32 saddr(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(u32 bar, u32 count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_wait(u32 bar, u32 parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "KT5_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra KT5_DONE_%=;\n\t"
        "bra KT5_WAIT_%=;\n\t"
        "KT5_DONE_%=:\n\t"
        "}" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bar_arrive(u32 bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void bar_expect_tx(u32 bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(u32 dst, const void *src, u32 bytes, u32 bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// one arrival on the mbarrier once every tcgen05 operation this thread issued before has completed
__device__ __forceinline__ void tc_commit(u32 bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor: K-major rows of 128 bytes, SWIZZLE_128B (bits 61-63 = 2), 8-row groups 1024 bytes apart
// (stride byte offset, bits 32-45, in 16-byte units), leading byte offset unused for this layout (1), descriptor version 1 (bit 46)
__device__ __forceinline__ u64 smem_desc(u32 addr) {
    return (u64)((addr >> 4) & 0x3FFFu) | ((u64)1 << 16) | ((u64)(1024 >> 4) << 32) | ((u64)1 << 46) | ((u64)2 << 61);
}
// D[tmem] (+)= A[smem] * B[smem]^T : 128 x 256 x 32, s8 x u8 -> s32
__device__ __forceinline__ void mma_i8(u32 d_tmem, u64 adesc, u64 bdesc, u32 accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(IDESC), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld32(u32 taddr, int (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
        "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
          "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ u64 signed_to_field(long long x) { return x < 0 ? GL_Q + (u64)x : (u64)x; }

__global__ void __launch_bounds__(THREADS, 1) lwe_keyswitch_tc5_kernel(const Args a) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ u32 tmem_slot;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 base = (saddr(smem_raw) + 1023u) & ~1023u;   // tiles start on a multiple of the swizzle period
    const u32 bars = base + NS * STAGE_BYTES;              // full[NS], empty[NS], tfull[2], tempty[2]
    auto full = [&](u32 s) { return bars + 8u * s; };
    auto empty = [&](u32 s) { return bars + 8u * (NS + s); };
    auto tfull = [&](u32 b) { return bars + 8u * (2 * NS + b); };
    auto tempty = [&](u32 b) { return bars + 8u * (2 * NS + 2 + b); };
    const u32 KT = a.K / BK, tiles = a.mtiles * a.ntiles;

    if (tid == 0) {
        for (u32 s = 0; s < (u32)NS; s++) {
            bar_init(full(s), 1);
            bar_init(empty(s), 1);
        }
        for (u32 b = 0; b < 2; b++) {
            bar_init(tfull(b), 1);
            bar_init(tempty(b), 4);   // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {   // all 512 columns: two 128 x 256 s32 accumulators (one CTA per SM by shared memory, nobody else asks)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(saddr(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const u32 tmem = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {   // ---- producer: stage ring <- HBM / L2
            u32 it = 0;
            for (u32 tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
                const u32 mt = tile % a.mtiles, nt = tile / a.mtiles;
                const int8_t *ga = a.digT + (size_t)mt * KT * A_BYTES;
                const uint8_t *gb = a.kskT + (size_t)nt * KT * B_BYTES;
                for (u32 kt = 0; kt < KT; kt++, it++) {
                    const u32 s = it % NS;
                    if (it >= (u32)NS) bar_wait(empty(s), ((it / NS) - 1) & 1);
                    bar_expect_tx(full(s), STAGE_BYTES);
                    const u32 sa = base + s * STAGE_BYTES;
                    bulk_load(sa, ga + (size_t)kt * A_BYTES, A_BYTES, full(s));
                    bulk_load(sa + A_BYTES, gb + (size_t)kt * B_BYTES, B_BYTES, full(s));
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {   // ---- tensor-core issue
            u32 it = 0, ti = 0;
            for (u32 tile = blockIdx.x; tile < tiles; tile += gridDim.x, ti++) {
                const u32 acc = ti & 1;
                if (ti >= 2) {   // the epilogue has drained this accumulator
                    bar_wait(tempty(acc), ((ti >> 1) - 1) & 1);
                    tc_fence_after();
                }
                const u32 d_tmem = tmem + acc * (u32)BN;
                for (u32 kt = 0; kt < KT; kt++, it++) {
                    const u32 s = it % NS;
                    bar_wait(full(s), (it / NS) & 1);
                    tc_fence_after();
                    const u32 sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
#pragma unroll
                    for (u32 k = 0; k < BK / 32; k++) mma_i8(d_tmem, smem_desc(sa + 32 * k), smem_desc(sb + 32 * k), (kt | k) != 0 ? 1u : 0u);
                    tc_commit(empty(s));   // the stage is free once these four have read it
                }
                tc_commit(tfull(acc));
            }
        }
    } else {   // ---- epilogue: tensor memory -> words
        const u32 q = (u32)warp & 3u, row = 32 * q + (u32)lane, cols = a.n + 1;
        u32 ti = 0;
        for (u32 tile = blockIdx.x; tile < tiles; tile += gridDim.x, ti++) {
            const u32 mt = tile % a.mtiles, nt = tile / a.mtiles, acc = ti & 1;
            bar_wait(tfull(acc), (ti >> 1) & 1);
            tc_fence_after();
            const u32 taddr = tmem + ((32 * q) << 16) + acc * (u32)BN;
            const u32 b = mt * BM + row;
            const bool live = b < a.B;
            u64 *orow = a.out + (size_t)b * cols;
            const u64 body = live ? a.in[(size_t)b * (a.kN + 1) + a.kN] : 0;
#pragma unroll 1
            for (u32 c4 = 0; c4 < BN / 32; c4++) {
                int v[32];
                tmem_ld32(taddr + 32 * c4, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (u32 h = 0; h < 4; h++) {
                    const u32 col = nt * (BN / 8) + 4 * c4 + h;
                    const long long P = (long long)v[8 * h] + ((long long)v[8 * h + 1] << 8) + ((long long)v[8 * h + 2] << 16) + ((long long)v[8 * h + 3] << 24);
                    const long long Q = (long long)v[8 * h + 4] + ((long long)v[8 * h + 5] << 8) + ((long long)v[8 * h + 6] << 16) + ((long long)v[8 * h + 7] << 24);
                    u64 w;
                    if (a.torus) {
                        w = (u64)P + ((u64)Q << 32);
                        if (col == a.n) w += body;
                    } else {
                        w = gl_canon(gl_add(gl_mul_lazy(signed_to_field(Q), 1ULL << 32), signed_to_field(P)));
                        if (col == a.n) w = gl_canon(gl_add(w, body));
                    }
                    if (live && col < cols) orow[col] = w;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(tempty(acc));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

}  // namespace kstc5
