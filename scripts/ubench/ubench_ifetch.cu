// ubench_ifetch.cu — how fast can an sm_100a scheduler issue STRAIGHT-LINE code that does not fit its instruction caches?
// The torus blind rotation is one fully unrolled loop body of ~90 KB (register-resident transforms cannot be rolled);
// ncu shows `no_instruction` stalls wherever that body is dense in cheap integer instructions.  This probe runs a loop
// whose body is BODY independent integer instructions (16 B each) and reports instructions per cycle per scheduler for
// bodies from 2 KB to 256 KB, with 1 / 2 / 4 warps per scheduler executing the same code.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_ifetch ubench_ifetch.cu && ./ubench_ifetch
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int N>
struct Body {
    __device__ __forceinline__ static void run(unsigned (&x)[8], unsigned m) {
        Body<N / 2>::run(x, m);
        Body<N - N / 2>::run(x, m);
    }
};
template <>
struct Body<8> {
    __device__ __forceinline__ static void run(unsigned (&x)[8], unsigned m) {
        // 8 independent chains, alternating the ALU pipe (add / xor) and the FMA pipe (mad)
        asm volatile("add.u32 %0, %0, %1;" : "+r"(x[0]) : "r"(m));
        asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(x[1]) : "r"(m));
        asm volatile("xor.b32 %0, %0, %1;" : "+r"(x[2]) : "r"(m));
        asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(x[3]) : "r"(m));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(x[4]) : "r"(m));
        asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(x[5]) : "r"(m));
        asm volatile("xor.b32 %0, %0, %1;" : "+r"(x[6]) : "r"(m));
        asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(x[7]) : "r"(m));
    }
};

template <int BODY>
__global__ void __launch_bounds__(512) ifetch_kernel(unsigned *sink, int iters, unsigned m) {
    unsigned x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x + i;
#pragma unroll 1
    for (int it = 0; it < iters; it++) Body<BODY>::run(x, m);
    unsigned acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc ^= x[i];
    if (acc == 0x12345) sink[0] = acc;
}

template <int BODY>
static void run(int sms, double clk) {
    unsigned *sink;
    cudaMalloc(&sink, 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    printf("body %6d instr (%4d KB):", BODY, BODY * 16 / 1024);
    for (int wps : {1, 2, 4}) {
        const int iters = (1 << 22) / BODY;
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            ifetch_kernel<BODY><<<sms, 128 * wps>>>(sink, iters, 3u);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep && ms < best) best = ms;
        }
        const double ipc = (double)iters * BODY * wps / (best * 1e-3 * clk);
        printf("   %d warp/sched: %.3f instr/clk/sched", wps, ipc);
    }
    printf("\n");
    cudaFree(sink);
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const double clk = prop.clockRate * 1e3;
    printf("%s, %d SMs: straight-line integer code in a loop, all warps of a CTA run the same code\n", prop.name, prop.multiProcessorCount);
    run<128>(prop.multiProcessorCount, clk);
    run<512>(prop.multiProcessorCount, clk);
    run<1024>(prop.multiProcessorCount, clk);
    run<2048>(prop.multiProcessorCount, clk);
    run<4096>(prop.multiProcessorCount, clk);
    run<6144>(prop.multiProcessorCount, clk);
    run<8192>(prop.multiProcessorCount, clk);
    run<16384>(prop.multiProcessorCount, clk);
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
