// ubench_overlap.cu — do the FP64 pipe and the shared-memory pipe of one sm_100a SM run at the same time?
// The torus blind rotation (fft64.cuh) needs both about equally (profiles/r02b: fp64 42 % + lsu 50 % of the cycles), and its
// time per iteration is close to the SUM of the two.  This probe runs, on every SM, W warps of pure DFMA work and W warps of
// pure LDS.128 / STS.128 work, alone and together; if the two pipes overlap the combined time is the maximum, not the sum.
// Also: one warp doing both, interleaved instruction by instruction (what an unrolled transform could do).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_overlap ubench_overlap.cu && ./ubench_overlap
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned saddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sts128(void *p, double2 v) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(saddr(p)), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ double2 lds128(const void *p) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(saddr(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(void *p, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(saddr(p)), "d"(v) : "memory"); }
__device__ __forceinline__ double lds64(const void *p) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(saddr(p)) : "memory");
    return v;
}
// per iteration: an FP64 warp issues 512 DFMA (16 chains x 32), a shared-memory warp 32 stores + 32 loads
// mode bit 0: FP64 warps work; bit 1: shared-memory warps work; bit 2: every warp does both, interleaved
template <int LDW>   // shared-memory access width in bytes (8 or 16)
__global__ void __launch_bounds__(512) overlap_kernel(double *sink, int iters, int mode, int warps_per_kind) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_fp = warp < warps_per_kind;
    double x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = 1.0 + i + lane;
    const double a = 1.0000001, b = 1e-9;
    double2 v[8];
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = make_double2(i, lane);
    unsigned char *mine = smem + (size_t)(warp % 8) * 8192;   // 8 KiB per warp slot
    if (mode & 4) {
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < 8; r++) {
#pragma unroll
                for (int i = 0; i < 64; i++) x[i & 15] = __fma_rn(x[i & 15], a, b);
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    if (LDW == 16) {
                        sts128((double2 *)mine + r * 32 + lane, v[r]);
                        v[r] = lds128((double2 *)mine + ((r + 1 + q) & 7) * 32 + lane);
                    } else {
                        sts64((double *)mine + r * 32 + lane, v[r].x);
                        v[r].x = lds64((double *)mine + ((r + 1 + q) & 7) * 32 + lane);
                    }
                }
            }
        }
    } else if (is_fp) {
        if (mode & 1)
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int r = 0; r < 8; r++) {
#pragma unroll
                    for (int i = 0; i < 64; i++) x[i & 15] = __fma_rn(x[i & 15], a, b);
                }
            }
    } else {
        if (mode & 2)
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int r = 0; r < 8; r++) {
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        if (LDW == 16) {
                            sts128((double2 *)mine + r * 32 + lane, v[r]);
                            v[r] = lds128((double2 *)mine + ((r + 1 + q) & 7) * 32 + lane);
                        } else {
                            sts64((double *)mine + r * 32 + lane, v[r].x);
                            v[r].x = lds64((double *)mine + ((r + 1 + q) & 7) * 32 + lane);
                        }
                    }
                }
            }
    }
    double acc = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) acc += x[i];
#pragma unroll
    for (int i = 0; i < 8; i++) acc += v[i].x + v[i].y;
    if (acc == 0.123) sink[0] = acc;
}

template <int LDW>
static float run(int sms, int mode, int wpk, int iters) {
    double *sink;
    cudaMalloc(&sink, 8);
    cudaFuncSetAttribute(overlap_kernel<LDW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0);
        overlap_kernel<LDW><<<sms, 64 * wpk, 65536>>>(sink, iters, mode, wpk);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    cudaFree(sink);
    return best;
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount, iters = 1024;
    printf("%s, %d SMs; per CTA: W warps of DFMA (512 per iteration) + W warps of shared-memory traffic (32 STS + 32 LDS per iteration)\n", prop.name, sms);
    for (int wpk : {4, 8}) {
        const float fp = run<16>(sms, 1, wpk, iters);
        const float m16 = run<16>(sms, 2, wpk, iters), both16 = run<16>(sms, 3, wpk, iters);
        const float m8 = run<8>(sms, 2, wpk, iters), both8 = run<8>(sms, 3, wpk, iters);
        const float il16 = run<16>(sms, 4, wpk, iters), il8 = run<8>(sms, 4, wpk, iters);
        const double clk = prop.clockRate * 1e3;
        printf("W=%d  DFMA alone %.3f ms (%.1f lane-dfma/SM/clk)\n", wpk, fp, (double)wpk * 32 * 512 * iters / (fp * 1e-3) / clk);
        printf("      LDS/STS.128 alone %.3f ms (%.1f B/SM/clk)   together %.3f ms  -> overlap factor %.2f (1 = max, 0 = sum)\n", m16,
               (double)wpk * 64 * 512 * iters / (m16 * 1e-3) / clk, both16, (fp + m16 - both16) / (fp < m16 ? fp : m16));
        printf("      LDS/STS.64  alone %.3f ms (%.1f B/SM/clk)   together %.3f ms  -> overlap factor %.2f\n", m8,
               (double)wpk * 64 * 256 * iters / (m8 * 1e-3) / clk, both8, (fp + m8 - both8) / (fp < m8 ? fp : m8));
        printf("      interleaved in every warp (2W warps, twice the work): .128 %.3f ms (sum of alone x2 = %.3f, max x2 = %.3f); .64 %.3f ms (sum x2 = %.3f)\n",
               il16, 2 * (fp + m16), 2 * (fp > m16 ? fp : m16), il8, 2 * (fp + m8));
    }
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
