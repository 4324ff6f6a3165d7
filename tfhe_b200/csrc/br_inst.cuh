// br_inst.cuh — instantiates the templated kernels for ONE ring size (LUX_LOGN) and defines its
// launchers.  Included by br_n9.cu .. br_n12.cu, which exist only so that nvcc compiles the ring
// sizes in parallel.
#include "engine.h"

template <int LOGN, int TEAMS, int ELL = 0, int LOGB = 0, int R = 16>
__global__ void __launch_bounds__((Tile<LOGN>::N / R) * TEAMS, 1) blind_rotate_kernel(const BRArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    blind_rotate_body<LOGN, TEAMS, ELL, LOGB, R>(a, threadIdx.x, blockIdx.x, smem);
}

// gadgets of the named parameter sets get their own instantiation (compile-time digit rule):
//   S2 (N = 2048): one 23-bit digit, limb convention logB = 41        S1 (N = 1024, main.rs:23-30): ell = 4, logB = 5
template <int LOGN> struct NamedGadget { static constexpr int ELL = 0, LOGB = 0; };
template <> struct NamedGadget<11> { static constexpr int ELL = 1, LOGB = 41; };
template <> struct NamedGadget<10> { static constexpr int ELL = 4, LOGB = 5; };

template <int LOGN, int TEAMS>
__global__ void __launch_bounds__(Tile<LOGN>::T *TEAMS, 1) ntt_batch_kernel(const NttArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    ntt_batch_body<LOGN, TEAMS>(a, threadIdx.x, blockIdx.x, gridDim.x, smem);
}

// GGSW (x) GLWE for independent pairs: team `ct` multiplies ggsw[ct] with glwe[ct]
template <int LOGN, int TEAMS>
__global__ void __launch_bounds__(Tile<LOGN>::T *TEAMS, 1)
    external_product_kernel(const BRArgs a, const u64 *glwe_in, size_t ggsw_stride) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int N = Tile<LOGN>::N, T = Tile<LOGN>::T;
    u64 *tw = (u64 *)smem;
    const int tid = threadIdx.x, team = tid / T, tau = tid % T;
    u64 *acc = tw + N + (size_t)team * 3 * N;
    u64 *xb = acc + 2 * N;
    for (int i = tid; i < N; i += T * TEAMS) tw[i] = a.tw_fwd[i];
    __syncthreads();
    const u32 ct = blockIdx.x * TEAMS + team;
    if (ct >= a.B) return;
    for (int p = tau; p < 2 * N; p += T) acc[p] = glwe_in[(size_t)ct * 2 * N + p];
    team_sync(team, T);
    KeyStream none{};
    team_external_product<LOGN, false>(a, a.bsk + (size_t)ct * ggsw_stride, none, false, 0, acc, xb, tw, tau, team, tid & 31);
    for (int p = tau; p < 2 * N; p += T) a.out[(size_t)ct * 2 * N + p] = acc[p];
}


// NTT-domain polynomials in the reference's order -> resident key rows, in place: fold 1/N and the sign
// of the mask GLev (vtfhe/ggsw_ct.rs:108-111, crypto/poly.rs:62-63), permute to key_pos.  One CTA per polynomial.
template <int LOGN>
__global__ void __launch_bounds__(Tile<LOGN>::T) key_format_kernel(u64 *polys, u32 ell, u64 ninv) {
    constexpr int N = Tile<LOGN>::N;
    u64 *poly = polys + (size_t)blockIdx.x * N;
    const u32 glev = (blockIdx.x / (2 * ell)) & 1;
    const int tau = threadIdx.x;
    u64 x[16];
#pragma unroll
    for (int e = 0; e < 16; e++) x[e] = poly[posC<LOGN>(tau, e)];
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 16; e++) poly[key_pos<LOGN>(tau, e)] = ggsw_fold_elem(x[e], ninv, glev);
}

template <int LOGN, int TEAMS, int GE, int GB, int R>
static int launch_br_narrow(lux_engine *e, const BRArgs &a, cudaStream_t s, u32 grid, size_t smem) {
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_kernel<LOGN, TEAMS, GE, GB, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[e->device & 63] = true;
    }
    blind_rotate_kernel<LOGN, TEAMS, GE, GB, R><<<grid, (Tile<LOGN>::N / R) * TEAMS, smem, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

template <int LOGN, int TEAMS>
static int launch_br_t(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    constexpr size_t smem = blind_rotate_smem_bytes<LOGN, TEAMS>();
    static bool configured[64] = {};
    constexpr int GE = NamedGadget<LOGN>::ELL, GB = NamedGadget<LOGN>::LOGB;
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_kernel<LOGN, TEAMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (GE) CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_kernel<LOGN, TEAMS, GE, GB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[e->device & 63] = true;
    }
    const u32 grid = (a.B + TEAMS - 1) / TEAMS;
    // narrow tiles (8 / 4 coefficients per thread, 2x / 4x the warps per ciphertext): named gadgets only
    if (GE && a.ell == (u32)GE && a.logB == (u32)GB && e->br_regs != 16) {
        if constexpr (GE != 0 && (Tile<LOGN>::N / 4) * TEAMS <= 1024) {
            if (e->br_regs == 4) return launch_br_narrow<LOGN, TEAMS, GE, GB, 4>(e, a, s, grid, smem);
        }
        if constexpr (GE != 0 && (Tile<LOGN>::N / 8) * TEAMS <= 1024) {
            return launch_br_narrow<LOGN, TEAMS, GE, GB, 8>(e, a, s, grid, smem);
        }
    }
    if (GE && a.ell == (u32)GE && a.logB == (u32)GB && !e->generic_gadget)
        blind_rotate_kernel<LOGN, TEAMS, GE, GB><<<grid, Tile<LOGN>::T * TEAMS, smem, s>>>(a);
    else
        blind_rotate_kernel<LOGN, TEAMS><<<grid, Tile<LOGN>::T * TEAMS, smem, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

// does this launch take the narrow-tile instantiation of a named gadget?
static bool narrow_named(const lux_engine *e, u32 ell, u32 logB) {
    constexpr int GE = NamedGadget<LUX_LOGN>::ELL, GB = NamedGadget<LUX_LOGN>::LOGB;
    return GE != 0 && ell == (u32)GE && logB == (u32)GB && e->br_regs != 16;
}

// ciphertexts per CTA: as few as keep the batch within one wave of CTAs, up to the SMEM limit
static int pick_teams(const lux_engine *e, size_t B, int max_teams) {
    if (e->force_teams > 0 && e->force_teams <= max_teams) return e->force_teams;
    for (int t = 1; t < max_teams; t++)
        if (B <= (size_t)e->sm_count * t) return t;
    return max_teams;
}

#if LUX_LOGN == 10
// N = 1024, narrow tile: four ciphertexts per CTA (512 threads, 128 registers, no spills) take 47.0 ms per wave of
// 4 * #SMs, six (768 threads, 80 registers, spills) 72.9 ms per wave of 6 * #SMs: 705.5 against 728.7 ms for 8,880
// gate bootstraps.  Pick the cheaper wave plan; other gadgets keep 1 / 3 / 6.
static int pick_teams_10(const lux_engine *e, size_t B, u32 ell, u32 logB) {
    const int t = pick_teams(e, B, 6);
    if (!narrow_named(e, ell, logB) || t <= 3) return t;
    if (e->force_teams > 0) return e->force_teams == 4 ? 4 : t;
    const size_t sm = (size_t)e->sm_count;
    const size_t w4 = (B + 4 * sm - 1) / (4 * sm), w6 = (B + 6 * sm - 1) / (6 * sm);
    return (w4 * 470 <= w6 * 729) ? 4 : 6;
}
#endif

#if LUX_LOGN == 11
// N = 2048, narrow tile: two ciphertexts per CTA (512 threads, 128 registers, no spills) take 23.65 ms per wave of
// 2 * #SMs, three (768 threads, 80 registers, some spills) 36.05 ms per wave of 3 * #SMs (profiles/r01e_lockstep.txt):
// 354.8 against 360.5 ms for 4,440 ciphertexts, but 47.3 against 36.1 ms for 444.  Pick the cheaper wave plan.
static int pick_teams_11(const lux_engine *e, size_t B, u32 ell, u32 logB) {
    if (!narrow_named(e, ell, logB)) return pick_teams(e, B, 3);
    if (e->force_teams >= 1 && e->force_teams <= 3) return e->force_teams;
    const size_t sm = (size_t)e->sm_count;
    if (B <= sm) return 1;
    if (B <= 2 * sm) return 2;
    const size_t w2 = (B + 2 * sm - 1) / (2 * sm), w3 = (B + 3 * sm - 1) / (3 * sm);
    return (w2 * 2365 <= w3 * 3605) ? 2 : 3;
}
#endif

template <int LOGN, int TEAMS>
static int launch_ntt_t(lux_engine *e, const NttArgs &a, cudaStream_t s) {
    constexpr size_t smem = sizeof(u64) * ((size_t)Tile<LOGN>::N + (size_t)TEAMS * Tile<LOGN>::N);
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(ntt_batch_kernel<LOGN, TEAMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[e->device & 63] = true;
    }
    u32 grid = (a.n_polys + TEAMS - 1) / TEAMS;
    const u32 cap = (u32)e->sm_count * 8;
    if (grid > cap) grid = cap;
    if (grid == 0) return LUX_OK;
    ntt_batch_kernel<LOGN, TEAMS><<<grid, Tile<LOGN>::T * TEAMS, smem, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

template <int LOGN>
static int launch_xp_t(lux_engine *e, const BRArgs &a, const u64 *d_glwe, size_t stride, cudaStream_t s) {
    constexpr int TEAMS = 1;
    constexpr size_t smem = blind_rotate_smem_bytes<LOGN, TEAMS>();
    CUDA_TRY(e, cudaFuncSetAttribute(external_product_kernel<LOGN, TEAMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    external_product_kernel<LOGN, TEAMS><<<a.B, Tile<LOGN>::T * TEAMS, smem, s>>>(a, d_glwe, stride);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

// latency mode (tfhe_kernels.cuh "latency mode"): one ciphertext per CTA, G = 2*ell teams
template <int LOGN, int G, int ELL = 0, int LOGB = 0, int R = 16>
__global__ void __launch_bounds__((Tile<LOGN>::N / R) * G, 1) blind_rotate_split_kernel(const BRArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    blind_rotate_split_body<LOGN, G, ELL, LOGB, R>(a, threadIdx.x, blockIdx.x, smem);
}

template <int LOGN, int G>
static int launch_br_split_t(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    constexpr size_t smem = blind_rotate_split_smem_bytes<LOGN, G>();
    static_assert(smem <= 227 * 1024 && Tile<LOGN>::T * G <= 1024, "latency-mode configuration does not fit an SM");
    static bool configured[64] = {};
    constexpr int GE = NamedGadget<LOGN>::ELL, GB = NamedGadget<LOGN>::LOGB;
    constexpr bool NAMED = GE != 0 && G == 2 * GE;
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_split_kernel<LOGN, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (NAMED) CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_split_kernel<LOGN, G, NAMED ? GE : 0, NAMED ? GB : 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[e->device & 63] = true;
    }
    // narrow tile (8 coefficients per thread) for the named gadgets when the CTA still fits 1024 threads
    if constexpr (NAMED && (Tile<LOGN>::N / 8) * G <= 1024) {
        if (a.ell == (u32)GE && a.logB == (u32)GB && !e->generic_gadget && e->cluster_regs != 16) {
            static bool configured8[64] = {};
            if (!configured8[e->device & 63]) {
                CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_split_kernel<LOGN, G, GE, GB, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                configured8[e->device & 63] = true;
            }
            blind_rotate_split_kernel<LOGN, G, GE, GB, 8><<<a.B, (Tile<LOGN>::N / 8) * G, smem, s>>>(a);
            e->launches++;
            CUDA_TRY(e, cudaGetLastError());
            return LUX_OK;
        }
    }
    if (NAMED && a.ell == (u32)GE && a.logB == (u32)GB && !e->generic_gadget)
        blind_rotate_split_kernel<LOGN, G, NAMED ? GE : 0, NAMED ? GB : 0><<<a.B, Tile<LOGN>::T * G, smem, s>>>(a);
    else
        blind_rotate_split_kernel<LOGN, G><<<a.B, Tile<LOGN>::T * G, smem, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

// cluster latency mode (tfhe_kernels.cuh "cluster latency mode"): one ciphertext per 2-CTA cluster, TPC = ell teams per
// CTA, R coefficients per thread (N/R threads per team)
template <int LOGN, int TPC, int ELL = 0, int LOGB = 0, int R = 16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((Tile<LOGN>::N / R) * TPC, 1) blind_rotate_cluster_kernel(const BRArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    blind_rotate_cluster_body<LOGN, TPC, ELL, LOGB, R>(a, threadIdx.x, blockIdx.x, smem);
}

template <int LOGN, int TPC, int R>
static int launch_br_cluster_r(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    constexpr size_t smem = blind_rotate_cluster_smem_bytes<LOGN, TPC>();
    constexpr int threads = (Tile<LOGN>::N / R) * TPC;
    static_assert(smem <= 227 * 1024 && threads <= 1024, "cluster latency-mode configuration does not fit an SM");
    constexpr int GE = NamedGadget<LOGN>::ELL, GB = NamedGadget<LOGN>::LOGB;
    constexpr bool NAMED = GE != 0 && TPC == GE;
    static bool configured[64] = {};
    if (!configured[e->device & 63]) {
        CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_cluster_kernel<LOGN, TPC, 0, 0, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (NAMED) CUDA_TRY(e, cudaFuncSetAttribute(blind_rotate_cluster_kernel<LOGN, TPC, NAMED ? GE : 0, NAMED ? GB : 0, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[e->device & 63] = true;
    }
    if (NAMED && a.ell == (u32)GE && a.logB == (u32)GB && !e->generic_gadget)
        blind_rotate_cluster_kernel<LOGN, TPC, NAMED ? GE : 0, NAMED ? GB : 0, R><<<2 * a.B, threads, smem, s>>>(a);
    else
        blind_rotate_cluster_kernel<LOGN, TPC, 0, 0, R><<<2 * a.B, threads, smem, s>>>(a);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

// coefficients per thread: the narrowest tile whose CTA still fits 1024 threads, unless a test pins it
template <int LOGN, int TPC>
static int launch_br_cluster_t(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    // measured (profiles/r01e_latency_probe.txt): 8 per thread is best with one team per CTA (S2: 12.6 / 7.87 / 7.97 ms for
    // 16 / 8 / 4), 4 per thread with several teams (S1: 11.1 / 8.17 / 7.91 ms)
    constexpr int T16 = Tile<LOGN>::T;
    const int r = e->cluster_regs ? e->cluster_regs : (TPC == 1 ? 8 : 4);
    if constexpr (T16 * 4 * TPC <= 1024) {
        if (r == 4) return launch_br_cluster_r<LOGN, TPC, 4>(e, a, s);
    }
    if constexpr (T16 * 2 * TPC <= 1024) {
        if (r == 8 || r == 4) return launch_br_cluster_r<LOGN, TPC, 8>(e, a, s);
    }
    return launch_br_cluster_r<LOGN, TPC, 16>(e, a, s);
}

// 1 = launched, 0 = not applicable for this ring size / gadget
static int try_launch_cluster(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    int rc = 0;
    bool done = true;
#if LUX_LOGN <= 10
    if (a.ell == 1) rc = launch_br_cluster_t<LUX_LOGN, 1>(e, a, s);
    else if (a.ell == 2) rc = launch_br_cluster_t<LUX_LOGN, 2>(e, a, s);
    else if (a.ell == 3) rc = launch_br_cluster_t<LUX_LOGN, 3>(e, a, s);
    else if (a.ell == 4) rc = launch_br_cluster_t<LUX_LOGN, 4>(e, a, s);
    else done = false;
#elif LUX_LOGN == 11
    if (a.ell == 1) rc = launch_br_cluster_t<11, 1>(e, a, s);
    else if (a.ell == 2) rc = launch_br_cluster_t<11, 2>(e, a, s);
    else done = false;
#else
    if (a.ell == 1) rc = launch_br_cluster_t<12, 1>(e, a, s);
    else done = false;
#endif
    if (rc) return rc;
    return done ? 1 : 0;
}

// 1 = launched in latency mode, 0 = not applicable (caller falls through to the throughput kernel), < 0 = error
static int try_launch_split(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    // automatic: only when the batch leaves SMs idle anyway, and not when a tuning run pinned the team count
    if (e->split_mode == 0 || (e->split_mode < 0 && (a.B > (u32)e->sm_count || e->force_teams > 0))) return 0;
    // two SMs per ciphertext while the batch leaves at least half of them idle (or when forced: mode 2)
    if (e->split_mode == 2 || (e->split_mode < 0 && 2 * a.B <= (u32)e->sm_count)) {
        const int c = try_launch_cluster(e, a, s);
        if (c) return c;
    }
    int rc = 0;
    bool done = true;
#if LUX_LOGN <= 10
    if (a.ell == 1) rc = launch_br_split_t<LUX_LOGN, 2>(e, a, s);
    else if (a.ell == 2) rc = launch_br_split_t<LUX_LOGN, 4>(e, a, s);
    else if (a.ell == 3) rc = launch_br_split_t<LUX_LOGN, 6>(e, a, s);
    else if (a.ell == 4) rc = launch_br_split_t<LUX_LOGN, 8>(e, a, s);
    else done = false;
#elif LUX_LOGN == 11
    if (a.ell == 1) rc = launch_br_split_t<11, 2>(e, a, s);
    else if (a.ell == 2) rc = launch_br_split_t<11, 4>(e, a, s);
    else done = false;
#else
    done = false;
#endif
    if (rc) return rc;
    return done ? 1 : 0;
}

#define LUX_CAT_(a, b) a##b
#define LUX_CAT(a, b) LUX_CAT_(a, b)

// TEAMS variants per ring size: SMEM = 16N + 32N*TEAMS bytes <= 227 KiB
int LUX_CAT(lux_launch_br_, LUX_LOGN)(lux_engine *e, const BRArgs &a, cudaStream_t s) {
    {
        const int sp = try_launch_split(e, a, s);
        if (sp) return sp < 0 ? sp : LUX_OK;
    }
#if LUX_LOGN == 9
    const int t = pick_teams(e, a.B, 12);
    if (t <= 1) return launch_br_t<9, 1>(e, a, s);
    if (t <= 4) return launch_br_t<9, 4>(e, a, s);
    return launch_br_t<9, 12>(e, a, s);
#elif LUX_LOGN == 10
    const int t = pick_teams_10(e, a.B, a.ell, a.logB);
    if (t <= 1) return launch_br_t<10, 1>(e, a, s);
    if (t <= 3) return launch_br_t<10, 3>(e, a, s);
    if (t == 4) return launch_br_t<10, 4>(e, a, s);
    return launch_br_t<10, 6>(e, a, s);
#elif LUX_LOGN == 11
    const int t = pick_teams_11(e, a.B, a.ell, a.logB);
    if (t <= 1) return launch_br_t<11, 1>(e, a, s);
    if (t <= 2) return launch_br_t<11, 2>(e, a, s);
    return launch_br_t<11, 3>(e, a, s);
#else
    return launch_br_t<12, 1>(e, a, s);
#endif
}

size_t LUX_CAT(lux_br_wave_, LUX_LOGN)(const lux_engine *e, size_t B) {
    if (e->split_mode != 0 && (e->split_mode > 0 || (B <= (size_t)e->sm_count && e->force_teams <= 0))) {
        const u32 ell = e->p.ell;
        if ((e->split_mode == 2 || (e->split_mode < 0 && 2 * B <= (size_t)e->sm_count)) &&
            ((LUX_LOGN <= 10 && ell <= 4) || (LUX_LOGN == 11 && ell <= 2) || (LUX_LOGN == 12 && ell == 1)))
            return (size_t)e->sm_count / 2;
        if ((LUX_LOGN <= 10 && ell <= 4) || (LUX_LOGN == 11 && ell <= 2)) return (size_t)e->sm_count;
    }
#if LUX_LOGN == 9
    const int t = pick_teams(e, B, 12);
    return (size_t)e->sm_count * (t <= 1 ? 1 : t <= 4 ? 4 : 12);
#elif LUX_LOGN == 10
    const int t = pick_teams_10(e, B, e->p.ell, e->p.logB);
    return (size_t)e->sm_count * (t <= 1 ? 1 : t <= 3 ? 3 : t == 4 ? 4 : 6);
#elif LUX_LOGN == 11
    const int t = pick_teams_11(e, B, e->p.ell, e->p.logB);
    return (size_t)e->sm_count * (t <= 1 ? 1 : t <= 2 ? 2 : 3);
#else
    return (size_t)e->sm_count;
#endif
}

int LUX_CAT(lux_launch_ntt_, LUX_LOGN)(lux_engine *e, const NttArgs &a, cudaStream_t s) {
#if LUX_LOGN <= 10
    return launch_ntt_t<LUX_LOGN, 4>(e, a, s);
#elif LUX_LOGN == 11
    return launch_ntt_t<11, 2>(e, a, s);
#else
    return launch_ntt_t<12, 1>(e, a, s);
#endif
}

int LUX_CAT(lux_launch_keyfmt_, LUX_LOGN)(lux_engine *e, u64 *d_polys, size_t n_polys, cudaStream_t s) {
    if (n_polys == 0) return LUX_OK;
    key_format_kernel<LUX_LOGN><<<(unsigned)n_polys, Tile<LUX_LOGN>::T, 0, s>>>(d_polys, e->p.ell, e->ninv);
    e->launches++;
    CUDA_TRY(e, cudaGetLastError());
    return LUX_OK;
}

int LUX_CAT(lux_launch_xp_, LUX_LOGN)(lux_engine *e, const BRArgs &a, const u64 *d_glwe, size_t stride, cudaStream_t s) {
    return launch_xp_t<LUX_LOGN>(e, a, d_glwe, stride, s);
}
