// Probe: where does tcgen05.cp.64x128b.warpx2::02_13 put the bytes of a shared-memory block?  (one B200, sm_100a)
//   nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/cp_probe tcgen05_cp_probe.cu && /tmp/cp_probe
// Shared memory holds 1 KiB whose 32-bit word w has the value w (word w = bytes 4w..4w+3).  The block is copied with one
// tcgen05.cp (no-swizzle descriptor, stride byte offset SBO) to tensor-memory columns 0..3 and every warp reads its 32 lanes back
// with tcgen05.ld.32x32b.x4.  Printed: for lanes 0, 1, 8, 31 of each warp the four words it sees.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(uint32_t *out, uint32_t sbo16, uint32_t lbo16) {
    __shared__ __align__(1024) uint32_t blk[1024];   // 4 KiB
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 1024; i += blockDim.x) blk[i] = i;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_addr(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes of blk -> async proxy reads
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = slot;
    // zero the columns first so that untouched lanes are visible
    {
        uint32_t z = 0xdeadbeefu;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %1, %1, %1};" ::"r"(tbase + ((32u * (warp & 3)) << 16)), "r"(z) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (tid == 0) {
        uint64_t desc = 0;
        desc |= (uint64_t)((smem_addr(blk) >> 4) & 0x3FFF);
        desc |= (uint64_t)(lbo16 & 0x3FFF) << 16;
        desc |= (uint64_t)(sbo16 & 0x3FFF) << 32;
        desc |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
        asm volatile("tcgen05.cp.cta_group::1.64x128b.warpx2::02_13 [%0], %1;" ::"r"(tbase), "l"(desc) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&bar)) : "memory");
    }
    // everybody waits for the copy
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(smem_addr(&bar))
                : "memory");
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t v0, v1, v2, v3;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3) : "r"(tbase + ((32u * (warp & 3)) << 16)) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    out[tid * 4 + 0] = v0;
    out[tid * 4 + 1] = v1;
    out[tid * 4 + 2] = v2;
    out[tid * 4 + 3] = v3;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tbase) : "memory");
}

int main() {
    uint32_t *d, h[512];
    cudaMalloc(&d, sizeof(h));
    const uint32_t cfgs[][2] = {{8, 0}, {8, 8}, {16, 0}, {1, 8}};
    for (auto &c : cfgs) {
        cudaMemset(d, 0, sizeof(h));
        probe<<<1, 128>>>(d, c[0], c[1]);
        cudaError_t e = cudaDeviceSynchronize();
        printf("SBO=%u LBO=%u (16-byte units): %s\n", c[0], c[1], cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        for (int w = 0; w < 4; w++)
            for (int l : {0, 1, 7, 8, 9, 31})
                printf("  warp %d lane %2d: words %4u %4u %4u %4u\n", w, l, h[(w * 32 + l) * 4], h[(w * 32 + l) * 4 + 1], h[(w * 32 + l) * 4 + 2], h[(w * 32 + l) * 4 + 3]);
    }
    return 0;
}
