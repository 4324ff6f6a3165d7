// ubench.cu — integer-pipe instruction throughput on sm_100a, to cost the Goldilocks butterfly.
// Each pattern runs 8 independent chains per thread, fully unrolled, no memory traffic; the result is
// reported as SM sub-partition cycles per warp instruction (SASS instruction counts per chain step are
// listed in PATTERNS below and were read off `cuobjdump -sass`).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench ubench.cu && ./ubench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32;
typedef uint64_t u64;

#define CHAINS 8
#define UNROLL 8

template <int K>
__device__ __forceinline__ void step(u32 &lo, u32 &hi, u32 m, u32 k) {
    if (K == 0) {  // 32-bit add
        asm volatile("add.u32 %0, %0, %1;" : "+r"(lo) : "r"(m));
    } else if (K == 1) {  // 64-bit add: add.cc + addc
        asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(lo), "+r"(hi) : "r"(m), "r"(k));
    } else if (K == 2) {  // 64-bit add + carry materialised + folded (gl_add shape)
        asm volatile(
            "{\n\t.reg .u32 c;\n\t"
            "add.cc.u32 %0, %0, %2;\n\taddc.cc.u32 %1, %1, %3;\n\taddc.u32 c, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, c;\n\tsubc.u32 %1, %1, 0;\n\tadd.u32 %1, %1, c;\n\t}"
            : "+r"(lo), "+r"(hi)
            : "r"(m), "r"(k));
    } else if (K == 3) {  // mad.wide.u32 with 64-bit addend
        u64 x = ((u64)hi << 32) | lo;
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x) : "r"(lo), "r"(m));
        lo = (u32)x;
        hi = (u32)(x >> 32);
    } else if (K == 4) {  // mad.lo.u32
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(m), "r"(k));
    } else if (K == 5) {  // mad.hi.u32
        asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(m), "r"(k));
    } else if (K == 6) {  // lop3
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(lo) : "r"(m), "r"(k));
    } else if (K == 7) {  // funnel shift
        asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(lo) : "r"(hi), "r"(k));
    } else if (K == 8) {  // setp + selp
        asm volatile("{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %0, %1;\n\tselp.u32 %0, %2, %0, p;\n\t}" : "+r"(lo) : "r"(m), "r"(k));
    } else if (K == 9) {  // mad.lo.cc + madc.hi.cc (wide multiply-add with carry out) + addc
        asm volatile(
            "{\n\t.reg .u32 c;\n\t"
            "mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.cc.u32 %1, %2, %3, %1;\n\taddc.u32 c, 0, 0;\n\tadd.u32 %1, %1, c;\n\t}"
            : "+r"(lo), "+r"(hi)
            : "r"(m), "r"(k));
    } else if (K == 10) {  // 1 wide + 2 plain adds (pipe overlap)
        u64 x = ((u64)hi << 32) | lo;
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x) : "r"(lo), "r"(m));
        lo = (u32)x;
        hi = (u32)(x >> 32);
        asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(lo), "+r"(hi) : "r"(m), "r"(k));
    } else if (K == 11) {  // 1 wide + 4 carry adds
        u64 x = ((u64)hi << 32) | lo;
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x) : "r"(lo), "r"(m));
        lo = (u32)x;
        hi = (u32)(x >> 32);
        asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(lo), "+r"(hi) : "r"(m), "r"(k));
        asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(lo), "+r"(hi) : "r"(k), "r"(m));
    } else if (K == 12) {  // 64-bit subtract with borrow fold (gl_sub shape)
        asm volatile(
            "{\n\t.reg .u32 c;\n\t"
            "sub.cc.u32 %0, %0, %2;\n\tsubc.cc.u32 %1, %1, %3;\n\tsubc.u32 c, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, c;\n\tsubc.u32 %1, %1, 0;\n\t}"
            : "+r"(lo), "+r"(hi)
            : "r"(m), "r"(k));
    } else if (K == 13) {  // 3-input add
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(lo) : "r"(m), "r"(k));
    } else if (K == 14) {  // umax
        asm volatile("max.u32 %0, %0, %1;" : "+r"(lo) : "r"(m));
    } else if (K == 15) {  // independent IMAD (lo chain) and IADD3 (hi chain), 1:1
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(m), "r"(k));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(m), "r"(k));
    } else if (K == 16) {  // independent IMAD and 3 x IADD3
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(m), "r"(k));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(m), "r"(k));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(k), "r"(lo));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(m), "r"(m));
    } else if (K == 17) {  // IMAD.WIDE on (lo,hi) is too register hungry here: IMAD.HI (4-cycle fma op) + 2 x IADD3 on the other chain
        asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(m), "r"(k));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(m), "r"(k));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(k), "r"(m));
    } else if (K == 18) {  // independent LOP3 (lo) and IADD3 (hi): both alu pipe
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(lo) : "r"(m), "r"(k));
        asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %0, %1;\n\tadd.u32 %0, t, %2;\n\t}" : "+r"(hi) : "r"(m), "r"(k));
    } else if (K == 19) {  // FFMA-free check: IMAD + IMAD.HI independent (both fma pipe)
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo) : "r"(m), "r"(k));
        asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(hi) : "r"(m), "r"(k));
    }
}

template <int K>
__global__ void __launch_bounds__(256) bench_kernel(u32 *sink, u32 iters, u32 seed, long long *cycles) {
    u32 lo[CHAINS], hi[CHAINS];
    const u32 t = seed + blockIdx.x * 256 + threadIdx.x;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
        lo[i] = t * (2 * i + 3) + i;
        hi[i] = t ^ (0x9E3779B9u * (i + 1));
    }
    const u32 m = t | 1, k = (t >> 3) | 5;
    const long long c0 = clock64();
    for (u32 it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < UNROLL; r++) {
#pragma unroll
            for (int i = 0; i < CHAINS; i++) step<K>(lo[i], hi[i], m, k);
        }
    }
    const long long c1 = clock64();
    u32 acc = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) acc ^= lo[i] ^ hi[i];
    if (acc == 0x12345677u) sink[0] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = c1 - c0;
}

struct Pattern {
    const char *name;
    double sass_per_step;  // warp instructions per chain step (from cuobjdump -sass of this file, nvcc 12.9)
};
static const Pattern PATTERNS[] = {
    {"add.u32 (ptxas pairs them: IADD3 a+m+m)", 0.5},
    {"add.cc+addc (IADD3 + 0.5 IADD3.X)", 1.5},
    {"gl_add shape (2 IADD3, 2 IADD3.X, SEL)", 5},
    {"mad.wide.u32 (IMAD.WIDE.U32)", 1},
    {"mad.lo.u32 (IMAD)", 1},
    {"mad.hi.u32 (IMAD.HI.U32)", 1},
    {"lop3 (LOP3.LUT)", 1},
    {"shf.l.wrap (SHF)", 1},
    {"setp+selp (ISETP + SEL)", 2},
    {"mad.lo.cc+madc.hi.cc+addc+add (WIDE + ~2.6)", 3.6},
    {"wide + add64 (WIDE, 2 IADD3, IADD3.X)", 4},
    {"wide + 2 add64 (WIDE, 3 IADD3, .X, IMAD.X)", 6},
    {"gl_sub shape (2 IADD3, 2 IADD3.X, IMAD.X)", 5},
    {"3-input add (IADD3)", 1},
    {"max.u32 (0.5 VIMNMX3)", 0.5},
    {"indep IMAD + IADD3", 2},
    {"indep IMAD + 3 IADD3", 4},
    {"indep IMAD.HI + 2 IADD3", 3},
    {"indep LOP3 + IADD3 (both alu)", 2},
    {"indep IMAD + IMAD.HI (both fma)", 2},
};

template <int K>
static void run(u32 *sink, long long *dcyc, int sms, int warps_per_smsp) {
    const u32 iters = 2048;
    const int blocks_per_sm = warps_per_smsp * 4 / 8;  // 256 threads = 8 warps
    const int grid = sms * (blocks_per_sm ? blocks_per_sm : 1);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e30f;
    long long cyc = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(a);
        bench_kernel<K><<<grid, 256>>>(sink, iters, 1, dcyc);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (rep && ms < best) {
            best = ms;
            cudaMemcpy(&cyc, dcyc, sizeof(cyc), cudaMemcpyDeviceToHost);
        }
    }
    // block 0's own cycle count: its 8 warps share an SM with (blocks_per_sm - 1) other blocks
    const double steps = (double)iters * UNROLL * CHAINS;               // per thread
    const double warp_steps_per_smsp = steps * warps_per_smsp;         // all warps of one sub-partition
    const double cyc_per_step = (double)cyc / warp_steps_per_smsp;
    printf("%-44s warps/SMSP %2d  %8.3f ms  %7.3f cyc/step/SMSP  (%.1f SASS/step -> %6.3f cyc/inst)  %7.2f G steps/s\n", PATTERNS[K].name,
           warps_per_smsp, best, cyc_per_step, PATTERNS[K].sass_per_step, cyc_per_step / PATTERNS[K].sass_per_step,
           steps * 256.0 * grid / (best * 1e-3) / 1e9);
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    printf("%s, %d SMs, %d MHz\n", prop.name, prop.multiProcessorCount, prop.clockRate / 1000);
    u32 *sink;
    long long *dcyc;
    cudaMalloc(&sink, 64);
    cudaMalloc(&dcyc, 64);
    const int sms = prop.multiProcessorCount;
    for (int w : {2, 8}) {
        run<0>(sink, dcyc, sms, w);
        run<1>(sink, dcyc, sms, w);
        run<2>(sink, dcyc, sms, w);
        run<3>(sink, dcyc, sms, w);
        run<4>(sink, dcyc, sms, w);
        run<5>(sink, dcyc, sms, w);
        run<6>(sink, dcyc, sms, w);
        run<7>(sink, dcyc, sms, w);
        run<8>(sink, dcyc, sms, w);
        run<9>(sink, dcyc, sms, w);
        run<10>(sink, dcyc, sms, w);
        run<11>(sink, dcyc, sms, w);
        run<12>(sink, dcyc, sms, w);
        run<13>(sink, dcyc, sms, w);
        run<14>(sink, dcyc, sms, w);
        run<15>(sink, dcyc, sms, w);
        run<16>(sink, dcyc, sms, w);
        run<17>(sink, dcyc, sms, w);
        run<18>(sink, dcyc, sms, w);
        run<19>(sink, dcyc, sms, w);
        printf("\n");
    }
    cudaError_t st = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(st));
    return st != cudaSuccess;
}
