// br_fft.cu — kernels and launchers of the 2^64-torus / f64-FFT arithmetic (fft64.cuh).
#include "engine.h"
#include "fft64.cuh"
#include "fft64_lat.cuh"

#include <cmath>
#include <vector>

using namespace fft64;

template <int TEAMS, int NS, int RS, int TWS, int HB = 0, int GROUPS = 1, int KL = 0, int MK = 1, int LAG = 0>
__global__ void __launch_bounds__(Cfg<TEAMS, NS, RS, TWS, HB, GROUPS, KL>::THREADS, 1) blind_rotate_fft_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (KL == 3 || KL == 5 || (MK & 16)) {   // tensor memory (key cache / accumulator mirror): all 512 columns, taken and returned by warp 0 of EVERY CTA
        __shared__ u32 tmem_slot;
        if (threadIdx.x < 32) tmem_alloc_512(&tmem_slot);
        tmem_fence_before();
        __syncthreads();
        tmem_fence_after();
        const u32 tbase = tmem_slot;
        blind_rotate_body<TEAMS, NS, RS, TWS, HB, GROUPS, KL, MK, LAG>(a, smem, tbase);   // its early exits return to here
        tmem_fence_before();
        __syncthreads();
        if (threadIdx.x < 32) tmem_free_512(tbase);
        return;
    }
    blind_rotate_body<TEAMS, NS, RS, TWS, HB, GROUPS, KL, MK, LAG>(a, smem);
}

template <int V>
__global__ void __launch_bounds__(LAT_THREADS, 1) blind_rotate_fft_lat_kernel(const __grid_constant__ LatArgs la) {
    extern __shared__ __align__(16) unsigned char smem[];
    blind_rotate_lat_body<V>(la, smem);
}

template <int V>
static int launch_lat(lux_engine *e, const LatArgs &la, cudaStream_t s) {
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_fft_lat_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LAT_SMEM));
        configured[e->device & 63] = true;
    }
    blind_rotate_fft_lat_kernel<V><<<la.a.B - la.a.ct0, LAT_THREADS, LAT_SMEM, s>>>(la);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

constexpr int UTIL_WARPS = 4;
constexpr size_t UTIL_SMEM = sizeof(double2) * (TWB_ENTRIES + (size_t)UTIL_WARPS * BUF_POINTS);

__global__ void __launch_bounds__(32 * UTIL_WARPS) fft_probe_kernel(const __grid_constant__ Args a, const double2 *in, double2 *out, u32 n_polys, int inverse) {
    extern __shared__ __align__(16) unsigned char smem[];
    double2 *twB = (double2 *)smem;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < TWB_ENTRIES; i += blockDim.x) twB[i] = a.twB[i];
    __syncthreads();
    double2 *buf = twB + TWB_ENTRIES + (size_t)warp * BUF_POINTS;
    for (u32 p = blockIdx.x * UTIL_WARPS + warp; p < n_polys; p += gridDim.x * UTIL_WARPS)
        fft_probe_body(in + (size_t)p * M, out + (size_t)p * M, inverse, buf, twB, a, lane);
}

// src [n_ggsw][4 polys][N] torus words (ell = 1: poly = j * 2 + r) -> dst [n_ggsw][32][4][32] double2
__global__ void __launch_bounds__(32 * UTIL_WARPS) key_to_fourier_kernel(const __grid_constant__ Args a, const u64 *src, double2 *dst, u32 n_polys, double scale) {
    extern __shared__ __align__(16) unsigned char smem[];
    double2 *twB = (double2 *)smem;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < TWB_ENTRIES; i += blockDim.x) twB[i] = a.twB[i];
    __syncthreads();
    double2 *buf = twB + TWB_ENTRIES + (size_t)warp * BUF_POINTS;
    for (u32 p = blockIdx.x * UTIL_WARPS + warp; p < n_polys; p += gridDim.x * UTIL_WARPS)
        key_to_fourier_body(src + (size_t)p * N, dst + (size_t)(p / 4) * 4 * M, (int)(p % 4), scale, buf, twB, a, lane);
}

// torus digits / mod switch probes (parity with oracle/torus.c tor_decompose, tor_mod_switch)
__global__ void torus_digit_kernel(const u64 *in, long long *out, size_t count, u32 logB) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) out[i] = digit1(in[i], logB);
}

// ------------------------------------------------------------------------------------------ host side
// merged-twist twiddle tree, exactly as oracle/torus.c tor_twiddles derives it (same expression, same libm)
static void host_twiddles(std::vector<double2> &tw) {
    tw.assign(M, make_double2(0.0, 0.0));
    std::vector<u32> c(M, 0);
    c[1] = N / 2;
    for (u32 v = 1; v < (u32)M; v++) {
        const u32 e = c[v] / 2;
        if (v == 1 || (v & 1) == 0) {
            const double ang = M_PI * (double)e / (double)N;
            tw[v].x = cos(ang);
            tw[v].y = sin(ang);
        } else {
            tw[v].x = -tw[v - 1].y;
            tw[v].y = tw[v - 1].x;
        }
        if (2 * v < (u32)M) {
            c[2 * v] = e;
            c[2 * v + 1] = e + N;
        }
    }
}

int lux_fft_init(lux_engine *e) {
    std::vector<double2> tw;
    host_twiddles(tw);
    for (int v = 0; v < 32; v++) e->fft_twA[2 * v] = tw[v].x, e->fft_twA[2 * v + 1] = tw[v].y;
    std::vector<double2> twB(TWB_ENTRIES);
    for (int lane = 0; lane < 32; lane++) twB[lane] = tw[32 + lane];
    for (int g = 2; g <= 16; g <<= 1)
        for (int ee = 0; ee < g / 2; ee++)
            for (int lane = 0; lane < 32; lane++) twB[16 * g + ee * 32 + lane] = tw[32 * g + lane * g + 2 * ee];
    if (cudaMalloc(&e->d_fft_twB, sizeof(double2) * TWB_ENTRIES) != cudaSuccess) return LUX_ERR_OOM;
    if (cudaMemcpy(e->d_fft_twB, twB.data(), sizeof(double2) * TWB_ENTRIES, cudaMemcpyHostToDevice) != cudaSuccess) return LUX_ERR_CUDA;
    // the whole tree for the latency kernel (fft64_lat.cuh): node v at index v, the same doubles
    if (cudaMalloc(&e->d_fft_twFull, sizeof(double2) * M) != cudaSuccess) return LUX_ERR_OOM;
    if (cudaMemcpy(e->d_fft_twFull, tw.data(), sizeof(double2) * M, cudaMemcpyHostToDevice) != cudaSuccess) return LUX_ERR_CUDA;
    return LUX_OK;
}

static void fill_tables(const lux_engine *e, Args &a) {
    a.twB = (const double2 *)e->d_fft_twB;
    for (int v = 0; v < 32; v++) a.twA[v] = make_double2(e->fft_twA[2 * v], e->fft_twA[2 * v + 1]);
}

template <int TEAMS, int NS, int RS, int TWS, int HB = 0, int GROUPS = 1, int KL = 0, int MK = 1, int LAG = 0>
static int launch_cfg(lux_engine *e, const Args &a, cudaStream_t s) {
    using C = Cfg<TEAMS, NS, RS, TWS, HB, GROUPS, KL>;
    static_assert(C::SMEM <= 232448, "configuration exceeds the 227 KiB of shared memory a CTA can opt into");
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_fft_kernel<TEAMS, NS, RS, TWS, HB, GROUPS, KL, MK, LAG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        configured[e->device & 63] = true;
    }
    const unsigned grid = (a.B - a.ct0 + TEAMS - 1) / TEAMS;
    blind_rotate_fft_kernel<TEAMS, NS, RS, TWS, HB, GROUPS, KL, MK, LAG><<<grid, C::THREADS, C::SMEM, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

// Ciphertexts per CTA.  Measured on B200 at the named set (profiles/r02b_fft_phase_cycles.txt), cycles per CMUX iteration of a
// CTA: 1 team 17.0k, 2 teams 15.3k, 3 teams 18.6k, 4 teams (half-size buffers) 21.5k.  A batch runs in
// ceil(B / (teams x SMs)) waves, so the choice minimises waves x cycles; tiny batches spread over the SMs first.
static int pick_teams(const lux_engine *e, size_t B) {
    if (e->force_teams >= 1 && e->force_teams <= 4) return e->force_teams;
    const size_t sms = (size_t)e->sm_count;
    if (B <= sms) return 1;
    if (B <= 2 * sms) return 2;
    const double c3 = (double)((B + 3 * sms - 1) / (3 * sms)) * 18.6, c4 = (double)((B + 4 * sms - 1) / (4 * sms)) * 21.5;
    return c3 < c4 ? 3 : 4;
}

size_t lux_fft_wave(const lux_engine *e, size_t B) { return (size_t)e->sm_count * pick_teams(e, B); }

// one launch over the ciphertexts [a.ct0, a.B)
// template arguments: <teams, ring stages, key rows per stage, twiddles in shared memory, half-size buffers, groups, no ring>
static int launch_br_range(lux_engine *e, Args a, cudaStream_t s) {
    const u32 count = a.B - a.ct0;
    const int variant = e->fft_variant;   // 0 = the default of each team count; the others are the measured alternatives
    // a batch that leaves SMs idle anyway: one ciphertext per CTA, eight warps per ciphertext (fft64_lat.cuh)
    if ((e->fft_lat_mode == 1 || (e->fft_lat_mode < 0 && e->force_teams == 0 && count <= (u32)e->sm_count)) && count > 0) {
        LatArgs la{a, (const double2 *)e->d_fft_twFull};
        switch (variant) {   // measured alternatives (scripts/quick_torus.py, cfg = 10 x variant)
            case 1: return launch_lat<1>(e, la, s);
            case 2: return launch_lat<2>(e, la, s);
            case 3: return launch_lat<3>(e, la, s);
            case 4: return launch_lat<4>(e, la, s);   // phase marks (lux_debug_fft_trace)
            default: return launch_lat<0>(e, la, s);
        }
    }
    const int teams = pick_teams(e, count);
    if (teams >= 4 && a.ct0 != 0) LUX_FAIL(e, LUX_ERR_INVALID, "internal: the four-per-CTA kernel starts at ciphertext 0");
    switch (teams) {
        case 1:   // a warp alone on its scheduler: free instruction schedule (no phase marks)
            if (variant == 1) return launch_cfg<1, 4, 8, 1>(e, a, s);
            return launch_cfg<1, 4, 8, 1, 0, 1, 0, 0>(e, a, s);
        case 2:   // one warp per scheduler as well: 7.73 ms without the marks, 8.35 ms with them at B = 296
            if (variant == 1) return launch_cfg<2, 4, 8, 1>(e, a, s);
            return launch_cfg<2, 4, 8, 1, 0, 1, 0, 0>(e, a, s);
        case 3:
            if (variant == 1) return launch_cfg<3, 8, 4, 1, 1>(e, a, s);      // half-size buffers, ring = one whole GGSW
            if (variant == 13) return launch_cfg<3, 4, 4, 0>(e, a, s);        // every phase mark
            if (variant == 14) return launch_cfg<3, 4, 4, 0, 0, 1, 0, 1 | (384 << 8)>(e, a, s);   // the lighter decomposition
            return launch_cfg<3, 4, 4, 0, 0, 1, 0, 1 | 4 | (384 << 8)>(e, a, s);   // marks 7, 8 left out: 48.2k -> 49.6k PBS/s at B = 4,440
        default:
            if (variant == 1) return launch_cfg<4, 3, 4, 1, 1>(e, a, s);      // twiddles in shared memory, ring 3 x 8 KiB
            if (variant == 2) return launch_cfg<4, 1, 4, 1, 1, 1, 1>(e, a, s);   // no ring: key rows straight from L2 / L1
            if (variant == 3) return launch_cfg<4, 2, 4, 0, 1, 2>(e, a, s);   // two independent groups, ring 2 x 8 KiB each
            if (variant == 4) return launch_cfg<4, 1, 4, 1, 1, 1, 2>(e, a, s);   // no ring, CTA-wide barrier per iteration
            if (variant == 5) {   // key rows through tensor memory (four producer warps); refill paced at 400 ns per two rows unless set
                if (e->fft_stagger_ns == 6000) a.stagger_ns = 400;
                return launch_cfg<4, 1, 4, 1, 1, 1, 3>(e, a, s);
            }
            // LAG experiments (fft64.cuh): the upper two teams run behind the lower two; stagger = delay in cycles (6, 7) or
            // 10 x leader point + trailer point of the two coupling barriers (8)
            if (variant == 6) return launch_cfg<4, 4, 4, 0, 1, 1, 0, 1, 1>(e, a, s);   // ring + start delay
            if (variant == 7) return launch_cfg<4, 1, 4, 1, 1, 1, 1, 1, 1>(e, a, s);   // no ring + start delay
            if (variant == 8) {   // no ring + coupled offset; a point code that names no mark would leave the barriers unmatched
                auto ok = [](u32 k) { return k == 0 || k == 1 || k == 3 || (k >= 6 && k <= 9); };
                if (!ok(a.stagger_ns / 10) || !ok(a.stagger_ns % 10) || a.stagger_ns >= 100) a.stagger_ns = 16;
                return launch_cfg<4, 1, 4, 1, 1, 1, 1, 1, 2>(e, a, s);
            }
            // key through tensor memory, filled by TMA + tcgen05.cp (no thread touches the key before its tcgen05.ld)
            if (variant >= 10 && variant <= 12 && e->fft_stagger_ns == 6000) a.stagger_ns = 0;   // pacing in ns per 8 KiB piece (fft_stagger)
            if (variant == 10) return launch_cfg<4, 2, 4, 1, 1, 1, 5>(e, a, s);   // staging 2 x 8 KiB, twiddles in shared memory
            if (variant == 11) return launch_cfg<4, 4, 4, 0, 1, 1, 5>(e, a, s);   // staging 4 x 8 KiB, twiddles through L1
            if (variant == 12) return launch_cfg<4, 3, 4, 1, 1, 1, 5>(e, a, s);   // staging 3 x 8 KiB, twiddles in shared memory
            if (variant == 9) return launch_cfg<4, 4, 4, 0, 1, 1, 4>(e, a, s);   // ring fed by consumer warp 0: eight warps, up to 255 registers
            if (variant == 13) return launch_cfg<4, 4, 4, 0, 1>(e, a, s);     // every phase mark (the default until r02d)
            if (variant == 20) return launch_cfg<4, 4, 4, 0, 1, 1, 0, 1 | 16 | (386 << 8)>(e, a, s);   // accumulator mirror in tensor memory
            if (variant == 21) return launch_cfg<4, 4, 4, 0, 1, 1, 0, 1 | 16>(e, a, s);
            if (variant == 15) return launch_cfg<4, 4, 4, 0, 1, 1, 0, 1 | 8 | (386 << 8)>(e, a, s);   // butterfly stages inside the key MAC
            if (variant == 16) return launch_cfg<4, 4, 4, 0, 1, 1, 0, 1 | 8>(e, a, s);                // same, every mark
            if (variant == 17) return launch_cfg<4, 4, 4, 0, 1, 1, 4, 1 | 8 | (386 << 8)>(e, a, s);   // same on the consumer-fed ring (230 registers)
            if (variant == 18) return launch_cfg<4, 4, 4, 0, 1, 1, 4, 1 | 8>(e, a, s);
            if (variant == 19) return launch_cfg<4, 4, 4, 0, 1, 1, 4, 1 | (386 << 8)>(e, a, s);       // consumer-fed ring, marks 1, 7, 8 out
            return launch_cfg<4, 4, 4, 0, 1, 1, 0, 1 | (386 << 8)>(e, a, s);  // ring 4 x 8 KiB, twiddles through L1, marks 1, 7, 8 left out
    }
}

// Cycles per CMUX iteration of a CTA by ciphertexts per CTA (profiles/r02b_fft_phase_cycles.txt, r02c_latency.txt, r02d):
// latency kernel 8.2k (one per CTA, up to #SMs ciphertexts), 2 -> 15.3k, 3 -> 18.6k, 4 -> 21.5k.
static double range_cost(const lux_engine *e, size_t count) {
    const size_t sms = (size_t)e->sm_count;
    if (count == 0) return 0.0;
    if (count <= sms) return 8.2;
    if (count <= 2 * sms) return 15.3;
    const double c3 = (double)((count + 3 * sms - 1) / (3 * sms)) * 18.6, c4 = (double)((count + 4 * sms - 1) / (4 * sms)) * 21.5;
    return c3 < c4 ? c3 : c4;
}

// A batch that is not a whole number of waves: the whole waves run four ciphertexts per CTA, the rest runs as a launch of its
// own with the kernel that suits ITS size (a level of a radix circuit is nodes x integers ciphertexts, rarely a multiple of
// 592; 5,000 = 8 waves + 264, and 264 ciphertexts on 132 CTAs of two finish in 15.3k cycles per iteration instead of 21.5k).
// Same words either way: every kernel computes a ciphertext independently of its neighbours.
int lux_fft_launch_br(lux_engine *e, Args a, cudaStream_t s) {
    fill_tables(e, a);
    a.stagger_ns = (u32)e->fft_stagger_ns;
    a.trace = (u64 *)e->d_fft_trace;
    const size_t count = a.B - a.ct0, wave4 = (size_t)4 * e->sm_count;
    if (e->fft_tail_split && e->force_teams == 0 && e->fft_variant == 0 && e->fft_lat_mode < 0 && count > wave4 && count % wave4) {
        const size_t tail = count % wave4, full = count / wave4;
        if ((double)full * 21.5 + range_cost(e, tail) < range_cost(e, count)) {
            Args m = a, t = a;
            m.B = a.ct0 + (u32)(count - tail);
            t.ct0 = m.B;
            int rc = launch_br_range(e, m, s);
            if (rc) return rc;
            return launch_br_range(e, t, s);
        }
    }
    return launch_br_range(e, a, s);
}

int lux_fft_key_convert(lux_engine *e, const u64 *d_src, void *d_dst, size_t n_polys, cudaStream_t s) {
    Args a{};
    fill_tables(e, a);
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(key_to_fourier_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UTIL_SMEM));
        CUDA_TRY(e, cudaFuncSetAttribute(fft_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UTIL_SMEM));
        configured[e->device & 63] = true;
    }
    const double scale = ldexp(1.0, -64 - (LOGN - 1));
    const unsigned grid = (unsigned)std::min<size_t>((n_polys + UTIL_WARPS - 1) / UTIL_WARPS, (size_t)e->sm_count * 2);
    key_to_fourier_kernel<<<grid, 32 * UTIL_WARPS, UTIL_SMEM, s>>>(a, d_src, (double2 *)d_dst, (u32)n_polys, scale);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

int lux_fft_probe(lux_engine *e, const void *d_in, void *d_out, size_t n_polys, int inverse, cudaStream_t s) {
    Args a{};
    fill_tables(e, a);
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(fft_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UTIL_SMEM));
        configured[e->device & 63] = true;
    }
    const unsigned grid = (unsigned)std::min<size_t>((n_polys + UTIL_WARPS - 1) / UTIL_WARPS, (size_t)e->sm_count * 2);
    fft_probe_kernel<<<grid, 32 * UTIL_WARPS, UTIL_SMEM, s>>>(a, (const double2 *)d_in, (double2 *)d_out, (u32)n_polys, inverse);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

int lux_fft_digits(lux_engine *e, const u64 *d_in, long long *d_out, size_t count, u32 logB, cudaStream_t s) {
    torus_digit_kernel<<<(unsigned)std::min<size_t>((count + 255) / 256, 4096), 256, 0, s>>>(d_in, d_out, count, logB);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}
