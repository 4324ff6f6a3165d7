// emu.cpp — host-thread emulator for the device bodies in tfhe_b200/csrc/tfhe_kernels.cuh.
// Test infrastructure: compiles the SAME kernel bodies as plain C++ (-DTFHE_EMU), one OS thread
// per CUDA thread, std::barrier for bar.sync, so index arithmetic, layouts, swizzles and barrier
// placement are verified against the oracle on a box without a GPU.  PTX paths are not covered
// here (they fall back to portable C); those are checked on the GPU by tests/test_gpu_*.py.
#define TFHE_EMU 1
#include "../../tfhe_b200/csrc/tfhe_kernels.cuh"

#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

thread_local EmuBlock *g_emu_blk = nullptr;

// run `body(tid)` for one CTA of nthreads threads split into teams of team_size
static void run_cta(int nthreads, int team_size, const std::function<void(int)> &body) {
    const int nteams = (nthreads + team_size - 1) / team_size;
    std::barrier<> cta(nthreads);
    std::vector<std::unique_ptr<std::barrier<>>> teams;
    EmuBlock blk{};
    blk.cta = &cta;
    for (int t = 0; t < nteams; t++) {
        teams.emplace_back(new std::barrier<>(team_size));
        blk.team[t] = teams.back().get();
    }
    std::vector<std::thread> th;
    th.reserve(nthreads);
    for (int tid = 0; tid < nthreads; tid++)
        th.emplace_back([&, tid] {
            g_emu_blk = &blk;
            body(tid);
        });
    for (auto &t : th) t.join();
}

template <int LOGN, int TEAMS>
static void emu_ntt_t(const NttArgs &a, int nblocks) {
    constexpr int N = Tile<LOGN>::N, T = Tile<LOGN>::T;
    std::vector<unsigned char> smem(sizeof(u64) * (N + TEAMS * N));
    for (int bid = 0; bid < nblocks; bid++)
        run_cta(T * TEAMS, T, [&](int tid) { ntt_batch_body<LOGN, TEAMS>(a, tid, bid, nblocks, smem.data()); });
}

template <int LOGN, int TEAMS>
static void emu_br_t(const BRArgs &a) {
    constexpr int T = Tile<LOGN>::T;
    std::vector<unsigned char> smem(blind_rotate_smem_bytes<LOGN, TEAMS>());
    const int nblocks = (a.B + TEAMS - 1) / TEAMS;
    for (int bid = 0; bid < nblocks; bid++)
        run_cta(T * TEAMS, T, [&](int tid) { blind_rotate_body<LOGN, TEAMS>(a, tid, bid, smem.data()); });
}

template <int LOGN, int TEAMS, int R>
static void emu_br_r_t(const BRArgs &a) {
    constexpr int T = Tile<LOGN>::N / R;
    std::vector<unsigned char> smem(blind_rotate_smem_bytes<LOGN, TEAMS>());
    const int nblocks = (a.B + TEAMS - 1) / TEAMS;
    for (int bid = 0; bid < nblocks; bid++)
        run_cta(T * TEAMS, T, [&](int tid) { blind_rotate_body<LOGN, TEAMS, 0, 0, R>(a, tid, bid, smem.data()); });
}

template <int LOGN, int G>
static void emu_br_split_t(const BRArgs &a) {
    constexpr int T = Tile<LOGN>::T;
    std::vector<unsigned char> smem(blind_rotate_split_smem_bytes<LOGN, G>());
    for (u32 bid = 0; bid < a.B; bid++)
        run_cta(T * G, T, [&](int tid) { blind_rotate_split_body<LOGN, G>(a, tid, (int)bid, smem.data()); });
}

template <int LOGN, int R>
static void emu_ntt_r_t(const NttArgs &a) {
    constexpr int N = TileR<LOGN, R>::N, T = TileR<LOGN, R>::T;
    std::vector<unsigned char> smem(sizeof(u64) * 2 * N);
    for (u32 bid = 0; bid < a.n_polys; bid++)
        run_cta(T, T, [&](int tid) { ntt_r_body<LOGN, R>(a, tid, (int)bid, smem.data()); });
}

// reference-order NTT-domain GGSWs [n_ggsw][2][ell][2][N] -> resident key rows (fold + key_pos), as key_format_kernel
template <int LOGN>
static void to_key_t(u64 *ggsw, size_t n_polys, u32 ell, u64 ninv) {
    constexpr int N = Tile<LOGN>::N, T = Tile<LOGN>::T;
    std::vector<u64> tmp(N);
    for (size_t pi = 0; pi < n_polys; pi++) {
        u64 *poly = ggsw + pi * N;
        const u32 glev = (u32)((pi / (2 * ell)) & 1);
        for (int tau = 0; tau < T; tau++)
            for (int e = 0; e < 16; e++) tmp[key_pos<LOGN>(tau, e)] = ggsw_fold_elem(poly[posC<LOGN>(tau, e)], ninv, glev);
        memcpy(poly, tmp.data(), sizeof(u64) * N);
    }
}
bool g_emu_force_slow = false;
extern "C" {
void emu_force_slow(int on) { g_emu_force_slow = on != 0; }


int emu_ntt(int logn, int teams, int nblocks, const u64 *in, u64 *out, const u64 *tw, u64 scale, u32 n_polys, u32 inverse) {
    NttArgs a{in, out, tw, scale, n_polys, inverse};
    if (logn == 9 && teams == 2) emu_ntt_t<9, 2>(a, nblocks);
    else if (logn == 10 && teams == 2) emu_ntt_t<10, 2>(a, nblocks);
    else if (logn == 11 && teams == 1) emu_ntt_t<11, 1>(a, nblocks);
    else if (logn == 11 && teams == 2) emu_ntt_t<11, 2>(a, nblocks);
    else if (logn == 12 && teams == 1) emu_ntt_t<12, 1>(a, nblocks);
    else return -1;
    return 0;
}

// narrow-tile transform (R = 4 or 8 registers per thread), one polynomial per CTA
int emu_ntt_r(int logn, int r, const u64 *in, u64 *out, const u64 *tw, u64 scale, u32 n_polys, u32 inverse) {
    NttArgs a{in, out, tw, scale, n_polys, inverse};
    if (logn == 9 && r == 8) emu_ntt_r_t<9, 8>(a);
    else if (logn == 9 && r == 4) emu_ntt_r_t<9, 4>(a);
    else if (logn == 10 && r == 8) emu_ntt_r_t<10, 8>(a);
    else if (logn == 10 && r == 4) emu_ntt_r_t<10, 4>(a);
    else if (logn == 11 && r == 8) emu_ntt_r_t<11, 8>(a);
    else if (logn == 11 && r == 4) emu_ntt_r_t<11, 4>(a);
    else if (logn == 12 && r == 8) emu_ntt_r_t<12, 8>(a);
    else return -1;
    return 0;
}

int emu_blind_rotate(int logn, int teams, const BRArgs *a) {
    // teams >= 100: narrow tile, teams = 100 * R + TEAMS
    if (logn == 9 && teams == 803) emu_br_r_t<9, 3, 8>(*a);
    else if (logn == 10 && teams == 802) emu_br_r_t<10, 2, 8>(*a);
    else if (logn == 9 && teams == 403) emu_br_r_t<9, 3, 4>(*a);
    else if (logn == 9 && teams == 3) emu_br_t<9, 3>(*a);
    else if (logn == 10 && teams == 2) emu_br_t<10, 2>(*a);
    else if (logn == 11 && teams == 2) emu_br_t<11, 2>(*a);
    else return -1;
    return 0;
}

// latency-mode body: one ciphertext per CTA, G = 2*ell teams
int emu_blind_rotate_split(int logn, int g, const BRArgs *a) {
    if ((int)(2 * a->ell) != g) return -2;
    if (logn == 9 && g == 2) emu_br_split_t<9, 2>(*a);
    else if (logn == 9 && g == 4) emu_br_split_t<9, 4>(*a);
    else if (logn == 10 && g == 8) emu_br_split_t<10, 8>(*a);
    else return -1;
    return 0;
}

int emu_keyswitch(const KSArgs *a, int tb, int tx, int rc) {
    if (!(tb == 4 && tx == 32 && rc == 16)) return -1;
    constexpr int TB = 4, TX = 32, RC = 16;
    std::vector<unsigned char> smem(sizeof(u32) * RC * a->ell * TB);
    const int gx = (a->n + 1 + TX - 1) / TX, gy = (a->B + TB - 1) / TB;
    const int gz = a->splits > 1 ? (int)a->splits : 1;
    for (int bz = 0; bz < gz; bz++)
        for (int by = 0; by < gy; by++)
            for (int bx = 0; bx < gx; bx++)
                run_cta(TX, TX, [&](int tid) { lwe_keyswitch_body<TB, TX, RC>(*a, tid, bx, by, smem.data(), bz); });
    if (gz > 1)
        for (size_t idx = 0; idx < (size_t)a->B * (a->n + 1); idx++) ks_finish_body(*a, idx);
    return 0;
}

void emu_ksk_colsum(const u64 *ksk, u64 *ksum, u32 rows, u32 cols) {
    for (u32 c = 0; c < cols; c++) ksk_colsum_body(ksk, ksum, rows, cols, c);
}

int emu_ggsw_to_key(u64 *ggsw, size_t n_ggsw, u32 ell, u32 N, u64 ninv) {
    const size_t n_polys = n_ggsw * 4 * ell;
    switch (N) {
        case 512: to_key_t<9>(ggsw, n_polys, ell, ninv); return 0;
        case 1024: to_key_t<10>(ggsw, n_polys, ell, ninv); return 0;
        case 2048: to_key_t<11>(ggsw, n_polys, ell, ninv); return 0;
        case 4096: to_key_t<12>(ggsw, n_polys, ell, ninv); return 0;
    }
    return -1;
}

u64 emu_digit(u64 x, u32 li, u32 logB, int centred) { return gl_digit(x, li, logB, centred != 0); }
u32 emu_mod_switch(u64 x, u32 logN) { return gl_mod_switch(x, logN); }
u64 emu_mul(u64 a, u64 b) { return gl_mul(a, b); }
u64 emu_add(u64 a, u64 b) { return gl_add(a, b); }
u64 emu_sub(u64 a, u64 b) { return gl_sub(a, b); }
size_t emu_sizeof_brargs(void) { return sizeof(BRArgs); }
size_t emu_sizeof_ksargs(void) { return sizeof(KSArgs); }
}
