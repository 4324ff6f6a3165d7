// engine.h — internal state behind the opaque lux_engine handle (include/lux_tfhe.h).
#pragma once
#include <cuda_runtime.h>

#include <cstdio>

#include <map>
#include <mutex>
#include <vector>

#include "../../include/lux_tfhe.h"
#include "tfhe_kernels.cuh"
#include "fft64.cuh"

#define LUX_SCRATCH_SLOTS 10

struct lux_engine {
    // every extern "C" entry point that takes a non-const engine holds this for its whole duration: calls on one engine
    // from several host threads are serialised by the library ("blocking and thread-safe per engine").
    // Recursive because host-pointer entry points call the device-pointer ones.
    std::recursive_mutex mtx;
    lux_params p{};
    int device = 0;
    int sm_count = 148;
    u32 logN = 0, L = 0, L_ks = 0;
    u64 ninv = 0;
    cudaStream_t stream = nullptr;
    // NTT tables (reference convention), device
    u64 *d_tw_fwd = nullptr, *d_tw_inv = nullptr;
    // evaluation keys, one device blob: [header 8][bsk][glwe ksk][lwe ksk][ksum]
    u64 *d_blob = nullptr;
    size_t blob_words = 0, bsk_words = 0, gksk_words = 0, lksk_words = 0, ksum_words = 0;
    u64 *d_bsk = nullptr, *d_gksk = nullptr, *d_lksk = nullptr, *d_ksum = nullptr;
    bool has_bsk = false, has_gksk = false, has_lksk = false;
    // grow-only device staging for the host-pointer entry points
    void *scratch[LUX_SCRATCH_SLOTS] = {};
    size_t scratch_bytes[LUX_SCRATCH_SLOTS] = {};
    std::map<u32, u64 *> any_tables;  // lux_ntt_any: N -> [ROOTS | INVROOTS]
    // accounting
    uint64_t launches = 0;
    bool timing = false;
    cudaEvent_t ev[4] = {};
    float last_ms[2] = {0.f, 0.f};
    int force_teams = 0;  // debug/tuning override of ciphertexts per CTA
    int ks_splits = 0;    // key-switch input slices: 0 = automatic, > 0 forced (tests)
    // tensor-core key switch (ks_mma.cuh): byte planes of the LWE key switch key, K-major, derived at key load
    uint8_t *d_kskT = nullptr;
    u32 ks_npad = 0;          // rows of d_kskT = 8 * (n + 1) rounded up to the N tile
    int8_t *d_ks_dig = nullptr;  // grow-only [Mpad][K] digit buffer
    size_t ks_dig_bytes = 0;
    int ks_mode = -1;         // -1 = automatic (tcgen05 kernel whenever the parameter set is eligible), 0 = integer-pipe
                              // kernel only, 1 = mma.sync kernel, 2 = tcgen05 kernel (tests)
    uint8_t *d_kskT5 = nullptr;  // the same byte planes cut into 256 x 128 swizzled stage tiles (ks_tc5.cuh)
    u32 ks_npad5 = 0;
    int br_regs = 8;              // throughput blind rotate of the named gadgets: coefficients per thread (8 = narrow tile,
                                  // default; 16 = wide tile; 4: tuning).  Other gadgets always run the wide tile.
    int cluster_regs = 0;         // cluster latency mode: coefficients per thread, 0 = automatic, 16 / 8 / 4 pinned (tests, tuning)
    bool generic_gadget = false;  // tests: run the runtime-gadget kernel even where a compile-time instantiation exists
    int split_mode = -1;  // latency-mode blind rotate: -1 = automatic (2B <= #SMs: cluster kernel, B <= #SMs: split kernel), 0 = never,
                          // 1 = split kernel always, 2 = cluster kernel always (tests)
    // torus / f64-FFT arithmetic (fft64.cuh): twiddle tree nodes 0..31 as (re, im) pairs for the kernel-parameter bank,
    // pass-B table in device memory
    double fft_twA[64] = {};
    void *d_fft_twB = nullptr;
    void *d_fft_twFull = nullptr;   // whole twiddle tree (latency kernel)
    int fft_lat_mode = -1;          // -1 = latency kernel when B <= #SMs, 0 = never, 1 = always (tests)
    int fft_default_teams = 3;
    int fft_tail_split = 1;         // cut a batch into whole waves + a tail launch of its own (br_fft.cu); 0 = one launch (tests, A/B)
    void *d_fft_trace = nullptr;  // 16 phase-cycle counters (FFT64_TRACE_BUILD builds; lux_debug_fft_trace)
    int fft_stagger_ns = 6000;    // start delay between the groups of a CTA (kernels with GROUPS > 1)
    int fft_variant = 0;          // tuning: ring depth / twiddle placement of the 3-ciphertext kernel (br_fft.cu)
    struct HostPipe *pipe = nullptr;     // copy streams / pinned staging of the host-pointer PBS entry point (engine.cu)
    struct RadixState *radix = nullptr;  // euintN layer (radix.cu), created on first use
    char err[512] = "";
};

void lux_host_ntt_tables(u32 N, std::vector<u64> &roots, std::vector<u64> &invroots, u64 &ninv);
// PBS on device buffers where ciphertext ct uses test vector d_idx[ct / lut_div] (engine.cu)
int lux_pbs_device_glwe(lux_engine *e, const u64 *d_in, const u64 *d_testvs, const u32 *d_idx, u32 lut_div, size_t B, u64 *d_glwe, cudaStream_t s);
int lux_pbs_device_grouped(lux_engine *e, const u64 *d_in, const u64 *d_testvs, const u32 *d_idx, u32 lut_div, size_t B,
                           u64 *d_out, cudaStream_t s, u64 body_add = 0);
int lux_ensure_scratch(lux_engine *e, int slot, size_t bytes);
void lux_radix_destroy(lux_engine *e);  // radix.cu
extern "C" int lux_debug_force_teams(lux_engine *e, int teams);
extern "C" int lux_debug_split_mode(lux_engine *e, int mode);
extern "C" int lux_debug_ks_splits(lux_engine *e, int splits);
extern "C" int lux_debug_ks_mode(lux_engine *e, int mode);
extern "C" int lux_debug_generic_gadget(lux_engine *e, int on);
extern "C" int lux_debug_cluster_regs(lux_engine *e, int regs);
extern "C" int lux_debug_br_regs(lux_engine *e, int regs);
extern "C" int lux_debug_fft_variant(lux_engine *e, int variant);
extern "C" int lux_debug_fft_stagger(lux_engine *e, int ns);
extern "C" int lux_debug_fft_lat_mode(lux_engine *e, int mode);
extern "C" int lux_debug_fft_tail_split(lux_engine *e, int on);
extern "C" int lux_debug_fft_trace(lux_engine *e, unsigned long long out[16]);
extern "C" int lux_debug_int_peak(lux_engine *e, double out[3]);
extern "C" int lux_debug_field_probe(lux_engine *e, const uint64_t *a, const uint64_t *b, size_t count, uint64_t *out);

// br_fft.cu
int lux_fft_init(lux_engine *e);
int lux_fft_launch_br(lux_engine *e, fft64::Args a, cudaStream_t s);
size_t lux_fft_wave(const lux_engine *e, size_t B);
int lux_fft_key_convert(lux_engine *e, const u64 *d_src, void *d_dst, size_t n_polys, cudaStream_t s);
int lux_fft_probe(lux_engine *e, const void *d_in, void *d_out, size_t n_polys, int inverse, cudaStream_t s);
int lux_fft_digits(lux_engine *e, const u64 *d_in, long long *d_out, size_t count, u32 logB, cudaStream_t s);
inline bool lux_is_torus(const lux_engine *e) { return e->p.arith == LUX_ARITH_TORUS64_FFT; }

// the message is kept per engine and, for lux_last_error on the failing thread, per thread
void lux_note_error(const lux_engine *e);
#define LUX_FAIL(e, code, ...)                                  \
    do {                                                        \
        snprintf((e)->err, sizeof((e)->err), __VA_ARGS__);      \
        lux_note_error(e);                                      \
        return (code);                                          \
    } while (0)

#define LUX_LOCK(e)                                      \
    std::unique_lock<std::recursive_mutex> lux_lock_;    \
    if (e) lux_lock_ = std::unique_lock<std::recursive_mutex>((e)->mtx)

#define CUDA_TRY(e, call)                                                                              \
    do {                                                                                               \
        cudaError_t _st = (call);                                                                      \
        if (_st != cudaSuccess)                                                                        \
            LUX_FAIL(e, LUX_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_st), __FILE__, __LINE__); \
    } while (0)


// per-ring-size launchers, defined in br_n*.cu
#define LUX_DECL_LAUNCHERS(LN)                                                                      \
    int lux_launch_br_##LN(lux_engine *e, const BRArgs &a, cudaStream_t s);                         \
    int lux_launch_ntt_##LN(lux_engine *e, const NttArgs &a, cudaStream_t s);                       \
    size_t lux_br_wave_##LN(const lux_engine *e, size_t B);                                         \
    int lux_launch_xp_##LN(lux_engine *e, const BRArgs &a, const u64 *d_glwe, size_t stride, cudaStream_t s); \
    int lux_launch_keyfmt_##LN(lux_engine *e, u64 *d_polys, size_t n_polys, cudaStream_t s);
LUX_DECL_LAUNCHERS(9)
LUX_DECL_LAUNCHERS(10)
LUX_DECL_LAUNCHERS(11)
LUX_DECL_LAUNCHERS(12)
