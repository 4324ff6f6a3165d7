// tfhe_kernels.cuh — device bodies of the TFHE hot path for sm_100a (B200).
//
// What is here (reference lines are under /root/reference/research/verifiable-fhe/src):
//   * register-tiled negacyclic NTT over Goldilocks, 16 coefficients per thread, three
//     passes with two shared-memory exchanges          (crypto/poly.rs:9-64, ntt/mod.rs:12-70)
//   * external product / CMUX with the accumulator resident in shared memory
//                                                      (vtfhe/ggsw_ct.rs:98-112, glev_ct.rs:92-110,
//                                                       vtfhe/mod.rs:80-136, glwe_poly.rs:28-50)
//   * whole blind rotation in one launch: mod-switch at entry, n CMUX iterations, optional
//     GLWE key switch, sample extraction at exit       (ivc_based_vpbs.rs:109-124,277-371,
//                                                       crypto/glwe.rs:96-113, lwe.rs:28-34)
//   * batched LWE key switch                           (vtfhe/mod.rs:31-54, lev_ct.rs:63-76)
//
// A "team" of T = N/16 threads owns one ciphertext; TEAMS teams share a CTA, its twiddle table and
// the staged bootstrap-key stream.  Teams synchronise internally with named barriers and with each
// other only through the key stream's mbarriers (at most NSTAGE row pairs apart).
//
// The file also compiles as host C++ (-DTFHE_EMU): tests/emu runs the same bodies on host
// threads so that all index arithmetic is checked on the CPU box.
#pragma once
#include "gl64.cuh"

// Lock step (TK_LOCKSTEP).  The loop body of a blind rotation is ~135 KB of straight-line SASS, four times the 32 KB
// L1.5 instruction cache, so every team streams it from L2 once per CMUX; teams that drift apart each pay for their own
// stream (ncu: stall_no_instruction 0.14..0.40 per issue).  Keeping the teams of a CTA at the same place of the code
// lets them share one stream.  Level 0 = teams only meet at the key ring; 1 = CTA-wide barrier once per CMUX;
// 2 = once per NTT; 3 = every team barrier is CTA-wide.  Measured on B200 in profiles/r01e_lockstep.txt.
#ifndef TK_LOCKSTEP
#define TK_LOCKSTEP 1
#endif
// A team handle packs the team index (low 8 bits) and, for kernels whose teams run in lock step, the number of
// live threads of the CTA (bits 8 and up; 0 = no lock step: batched NTT, external-product kernel, emulator).
GL_HD int tk_team(int team, int live_threads) { return team | (live_threads << 8); }

#if defined(TFHE_EMU)
#include <barrier>
struct EmuBlock {
    std::barrier<> *cta;
    std::barrier<> *team[16];
};
extern thread_local EmuBlock *g_emu_blk;
inline void cta_sync() { g_emu_blk->cta->arrive_and_wait(); }
inline void team_sync(int team, int /*nthreads*/) { g_emu_blk->team[team & 0xff]->arrive_and_wait(); }
template <int LEVEL> inline void lockstep_point(int /*team*/) {}
inline void team_sync_outer(int team, int nthreads) { team_sync(team, nthreads); }
#define TK_DEV inline
#define TK_UNROLL
#else
__device__ __forceinline__ void cta_sync() { __syncthreads(); }
__device__ __forceinline__ void live_sync(int live_threads) {
    asm volatile("bar.sync 15, %0;" ::"r"(live_threads) : "memory");
}
__device__ __forceinline__ void team_sync(int team, int nthreads) {
    if (TK_LOCKSTEP >= 3 && (team >> 8)) {
        live_sync(team >> 8);
        return;
    }
    if (nthreads == 32)
        __syncwarp();
    else
        asm volatile("bar.sync %0, %1;" ::"r"((team & 0xff) + 1), "r"(nthreads) : "memory");
}
// extra CTA-wide meeting point of lock-step level LEVEL (no-op at other levels and for plain team handles)
template <int LEVEL>
__device__ __forceinline__ void lockstep_point(int team) {
    if (TK_LOCKSTEP == LEVEL && (team >> 8)) live_sync(team >> 8);
}
// the team barrier that closes a CMUX: CTA-wide from level 1 on
__device__ __forceinline__ void team_sync_outer(int team, int nthreads) {
    if (TK_LOCKSTEP >= 1 && (team >> 8))
        live_sync(team >> 8);
    else
        team_sync(team, nthreads);
}
#define TK_DEV __device__ __forceinline__
#define TK_UNROLL _Pragma("unroll")
#endif

// ---------------------------------------------------------------------------------- tile geometry
template <int LOGN>
struct Tile {
    static_assert(LOGN >= 9 && LOGN <= 12, "register-tiled NTT supports N = 512..4096");
    static constexpr int N = 1 << LOGN;
    static constexpr int T = N / 16;      // threads per polynomial / ciphertext
    static constexpr int S3 = LOGN - 8;   // stages of the last forward pass
    static constexpr int LB = LOGN - 8;   // bits of the low index in layout B
};

// exchange-buffer swizzle: XOR the bank-pair index with the next 4 address bits; makes the three
// access patterns below conflict-free for 8-byte accesses (see DESIGN.md "shared memory")
GL_HD int tk_swz(int p) { return p ^ ((p >> 4) & 15); }

// position of register e of thread tau in the three layouts
template <int LOGN> GL_HD int posA(int tau, int e) { return tau + Tile<LOGN>::T * e; }
template <int LOGN> GL_HD int posB(int tau, int e) {
    return (tau >> Tile<LOGN>::LB) * (Tile<LOGN>::N / 16) + e * (Tile<LOGN>::N / 256) + (tau & ((1 << Tile<LOGN>::LB) - 1));
}
template <int LOGN> GL_HD int posC(int tau, int e) { return 16 * tau + e; }

// one radix-2 stage over the 16 registers; blocks of 2^SH registers.  Block `blk` of the stage uses the
// reference's twiddle ROOTS[m + i0 + blk] (forward) or INVROOTS[m + i0 + blk] (inverse).  Only ROOTS is
// kept on chip: INVROOTS[m + i] = psi^-bitrev(m+i) = -psi^(N - bitrev(m+i)) = -ROOTS[2m - 1 - i], and the
// sign is absorbed by the inverse butterfly (u, v) -> (u + v, (v - u) * ROOTS[2m-1-i]).
template <int SH, bool INV>
TK_DEV void tile_stage(u64 (&x)[16], const u64 *tw, int m, int i0) {
    constexpr int half = 1 << (SH - 1);
    const u64 *p = INV ? tw + 2 * m - 1 - i0 : tw + m + i0;
    TK_UNROLL
    for (int blk = 0; blk < (16 >> SH); blk++) {
        const u64 w = INV ? p[-blk] : p[blk];
        TK_UNROLL
        for (int j = 0; j < half; j++) {
            const int e0 = (blk << SH) + j;
            if (INV)
                gl_gs_butterfly_neg(x[e0], x[e0 + half], w);
            else
                gl_ct_butterfly(x[e0], x[e0 + half], w);
        }
    }
}

// last forward pass / first inverse pass: stages 8..LOGN-1 on 16 consecutive coefficients
template <int LOGN, bool INV>
TK_DEV void tile_pass_low(u64 (&x)[16], const u64 *tw, int tau) {
    constexpr int S3 = Tile<LOGN>::S3;
    // stage with blocks of 2^SH registers: s' = S3 - SH, m = 256 << s', first block (tau << (4 - SH))
    if constexpr (!INV) {
        if constexpr (S3 >= 4) tile_stage<4, false>(x, tw, 256 << (S3 >= 4 ? S3 - 4 : 0), tau << 0);
        if constexpr (S3 >= 3) tile_stage<3, false>(x, tw, 256 << (S3 >= 3 ? S3 - 3 : 0), tau << 1);
        if constexpr (S3 >= 2) tile_stage<2, false>(x, tw, 256 << (S3 >= 2 ? S3 - 2 : 0), tau << 2);
        if constexpr (S3 >= 1) tile_stage<1, false>(x, tw, 256 << (S3 - 1), tau << 3);
    } else {
        if constexpr (S3 >= 1) tile_stage<1, true>(x, tw, 256 << (S3 - 1), tau << 3);
        if constexpr (S3 >= 2) tile_stage<2, true>(x, tw, 256 << (S3 >= 2 ? S3 - 2 : 0), tau << 2);
        if constexpr (S3 >= 3) tile_stage<3, true>(x, tw, 256 << (S3 >= 3 ? S3 - 3 : 0), tau << 1);
        if constexpr (S3 >= 4) tile_stage<4, true>(x, tw, 256 << (S3 >= 4 ? S3 - 4 : 0), tau << 0);
    }
}
// middle pass: stages 4..7 in layout B (b = high 4 bits of the position)
template <int LOGN, bool INV>
TK_DEV void tile_pass_mid(u64 (&x)[16], const u64 *tw, int tau) {
    const int b = tau >> Tile<LOGN>::LB;
    if (!INV) {
        tile_stage<4, false>(x, tw, 16, b);
        tile_stage<3, false>(x, tw, 32, 2 * b);
        tile_stage<2, false>(x, tw, 64, 4 * b);
        tile_stage<1, false>(x, tw, 128, 8 * b);
    } else {
        tile_stage<1, true>(x, tw, 128, 8 * b);
        tile_stage<2, true>(x, tw, 64, 4 * b);
        tile_stage<3, true>(x, tw, 32, 2 * b);
        tile_stage<4, true>(x, tw, 16, b);
    }
}
// stage with compile-time power-of-two twiddles (gl64.cuh: ROOTS[j] = 2^gl_root_log2(j) for j < 16): block BLK of the
// stage with m = M uses ROOTS[M + BLK] (forward) or, for the inverse against the forward table, ROOTS[2M - 1 - BLK]
template <int SH, bool INV, int M, int BLK = 0>
TK_DEV void tile_stage_pow2(u64 (&x)[16]) {
    constexpr int half = 1 << (SH - 1);
    constexpr int E = gl_root_log2(INV ? 2 * M - 1 - BLK : M + BLK);
    TK_UNROLL
    for (int j = 0; j < half; j++) {
        const int e0 = (BLK << SH) + j;
        if (INV)
            gl_gs_butterfly_neg_pow2<E>(x[e0], x[e0 + half]);
        else
            gl_ct_butterfly_pow2<E>(x[e0], x[e0 + half]);
    }
    if constexpr (BLK + 1 < (16 >> SH)) tile_stage_pow2<SH, INV, M, BLK + 1>(x);
}

// first forward pass / last inverse pass: stages 0..3 in layout A (twiddles uniform over threads)
template <bool INV>
TK_DEV void tile_pass_high(u64 (&x)[16], const u64 *tw) {
    // measured on B200 (profiles/r01e_arith_variants.txt): shifts trade four IMAD.WIDE for ~1 more ALU instruction
    // per butterfly, and the ALU pipe is the one that binds: 390.6 -> 403.9 ms.  Kept as an opt-in variant.
#if defined(GL_POW2_TWIDDLES)
    (void)tw;
    if (!INV) {
        tile_stage_pow2<4, false, 1>(x);
        tile_stage_pow2<3, false, 2>(x);
        tile_stage_pow2<2, false, 4>(x);
        tile_stage_pow2<1, false, 8>(x);
    } else {
        tile_stage_pow2<1, true, 8>(x);
        tile_stage_pow2<2, true, 4>(x);
        tile_stage_pow2<3, true, 2>(x);
        tile_stage_pow2<4, true, 1>(x);
    }
    return;
#endif
    if (!INV) {
        tile_stage<4, false>(x, tw, 1, 0);
        tile_stage<3, false>(x, tw, 2, 0);
        tile_stage<2, false>(x, tw, 4, 0);
        tile_stage<1, false>(x, tw, 8, 0);
    } else {
        tile_stage<1, true>(x, tw, 8, 0);
        tile_stage<2, true>(x, tw, 4, 0);
        tile_stage<3, true>(x, tw, 2, 0);
        tile_stage<4, true>(x, tw, 1, 0);
    }
}

// The exchanges of one transform go through ONE buffer xb[N] per team: the second exchange writes
// exactly the positions the same thread read in the first one, so only the first write needs a
// barrier against the previous user of the buffer (the leading team_sync).
//
// forward NTT of one polynomial held by a team: in  x[e] = a[posA(tau,e)] (natural order)
//                                               out x[e] = A[posC(tau,e)] (reference's bit-reversed order)
template <int LOGN>
TK_DEV void team_ntt_forward(u64 (&x)[16], u64 *xb, const u64 *tw, int tau, int team) {
    constexpr int T = Tile<LOGN>::T;
    lockstep_point<2>(team);
    tile_pass_high<false>(x, tw);
    team_sync(team, T);
    TK_UNROLL
    for (int e = 0; e < 16; e++) xb[tk_swz(posA<LOGN>(tau, e))] = x[e];
    team_sync(team, T);
    TK_UNROLL
    for (int e = 0; e < 16; e++) x[e] = xb[tk_swz(posB<LOGN>(tau, e))];
    tile_pass_mid<LOGN, false>(x, tw, tau);
    TK_UNROLL
    for (int e = 0; e < 16; e++) xb[tk_swz(posB<LOGN>(tau, e))] = x[e];
    team_sync(team, T);
    TK_UNROLL
    for (int e = 0; e < 16; e++) x[e] = xb[tk_swz(posC<LOGN>(tau, e))];
    tile_pass_low<LOGN, false>(x, tw, tau);
}

// inverse NTT WITHOUT the final 1/N: in x[e] = A[posC(tau,e)], out x[e] = N*a[posA(tau,e)] (lazy)
template <int LOGN>
TK_DEV void team_ntt_inverse(u64 (&x)[16], u64 *xb, const u64 *tw, int tau, int team) {
    constexpr int T = Tile<LOGN>::T;
    lockstep_point<2>(team);
    tile_pass_low<LOGN, true>(x, tw, tau);
    team_sync(team, T);
    TK_UNROLL
    for (int e = 0; e < 16; e++) xb[tk_swz(posC<LOGN>(tau, e))] = x[e];
    team_sync(team, T);
    TK_UNROLL
    for (int e = 0; e < 16; e++) x[e] = xb[tk_swz(posB<LOGN>(tau, e))];
    tile_pass_mid<LOGN, true>(x, tw, tau);
    TK_UNROLL
    for (int e = 0; e < 16; e++) xb[tk_swz(posB<LOGN>(tau, e))] = x[e];
    team_sync(team, T);
    TK_UNROLL
    for (int e = 0; e < 16; e++) x[e] = xb[tk_swz(posA<LOGN>(tau, e))];
    tile_pass_high<true>(x, tw);
}

struct NttArgs {
    const u64 *in;
    u64 *out;
    const u64 *tw;   // ROOTS [N] (both directions)
    u64 scale;       // every output multiplied by this (NINV for the reference's inverse; 1 = none)
    u32 n_polys;
    u32 inverse;
};

// ---------------------------------------------------------------- narrow tiles (latency modes): R = 4 or 8 registers
// Same transform with R coefficients per thread and T = N/R threads per polynomial: log2(N)/log2(R) passes of
// log2(R) stages (the last one shorter), one shared-memory exchange between passes.  A thread's share of the
// dependent instruction stream is R/16 of the wide tile's, which is what a single bootstrap waits for.
// Pass P works on the index bits [lo, lo + log2 R) with lo = max(logN - (P+1) log2 R, 0): register e of thread tau
// sits at position  ((tau >> lo) << (lo + log2 R)) | (e << lo) | (tau & (2^lo - 1)).
template <int LOGN, int R>
struct TileR {
    static_assert(R == 4 || R == 8, "narrow tiles have 4 or 8 registers");
    static constexpr int LR = (R == 8) ? 3 : 2;
    static constexpr int N = 1 << LOGN;
    static constexpr int T = N / R;
    static constexpr int NP = (LOGN + LR - 1) / LR;
    GL_HD static constexpr int lo(int pass) { return LOGN - (pass + 1) * LR < 0 ? 0 : LOGN - (pass + 1) * LR; }
};
template <int LOGN, int R, int P>
GL_HD int pos_r(int tau, int e) {
    constexpr int lo = TileR<LOGN, R>::lo(P), LR = TileR<LOGN, R>::LR;
    return ((tau >> lo) << (lo + LR)) | (e << lo) | (tau & ((1 << lo) - 1));
}

// OPT: optimistic butterflies (gl64.cuh) that skip the canonicalisation and raise `flag` when it would have mattered
template <int R, int SH, bool INV, bool OPT>
TK_DEV void tile_stage_r(u64 (&x)[R], const u64 *tw, int m, int i0, bool &flag) {
    constexpr int half = 1 << (SH - 1);
    const u64 *p = INV ? tw + 2 * m - 1 - i0 : tw + m + i0;   // see tile_stage: INVROOTS[m+i] = -ROOTS[2m-1-i]
    TK_UNROLL
    for (int blk = 0; blk < (R >> SH); blk++) {
        const u64 w = INV ? p[-blk] : p[blk];
        TK_UNROLL
        for (int j = 0; j < half; j++) {
            const int e0 = (blk << SH) + j;
            if (INV) {
                if (OPT)
                    gl_gs_butterfly_neg_opt(x[e0], x[e0 + half], w, flag);
                else
                    gl_gs_butterfly_neg(x[e0], x[e0 + half], w);
            } else {
                if (OPT)
                    gl_ct_butterfly_opt(x[e0], x[e0 + half], w, flag);
                else
                    gl_ct_butterfly(x[e0], x[e0 + half], w);
            }
        }
    }
}

// stage S of the transform inside pass P (no-op when S is outside the pass)
template <int LOGN, int R, int P, int S, bool INV, bool OPT>
TK_DEV void pass_stage_r(u64 (&x)[R], const u64 *tw, int base, bool &flag) {
    constexpr int LR = TileR<LOGN, R>::LR, lo = TileR<LOGN, R>::lo(P);
    if constexpr (S >= P * LR && S < (P + 1) * LR && S < LOGN)
        tile_stage_r<R, LOGN - S - lo, INV, OPT>(x, tw, 1 << S, base >> (LOGN - S), flag);
}
template <int LOGN, int R, int P, bool INV, bool OPT>
TK_DEV void pass_r(u64 (&x)[R], const u64 *tw, int tau, bool &flag) {
    constexpr int LR = TileR<LOGN, R>::LR;
    const int base = pos_r<LOGN, R, P>(tau, 0);
    if constexpr (!INV) {
        pass_stage_r<LOGN, R, P, P * LR + 0, false, OPT>(x, tw, base, flag);
        pass_stage_r<LOGN, R, P, P * LR + 1, false, OPT>(x, tw, base, flag);
        if constexpr (LR == 3) pass_stage_r<LOGN, R, P, P * LR + 2, false, OPT>(x, tw, base, flag);
    } else {
        if constexpr (LR == 3) pass_stage_r<LOGN, R, P, P * LR + 2, true, OPT>(x, tw, base, flag);
        pass_stage_r<LOGN, R, P, P * LR + 1, true, OPT>(x, tw, base, flag);
        pass_stage_r<LOGN, R, P, P * LR + 0, true, OPT>(x, tw, base, flag);
    }
}

// The exact pass, out of line and memory to memory: recomputes pass P of a thread from its inputs, which are still
// parked in its own slots of the exchange buffer, and leaves the outputs there.  Runs when an optimistic pass raised
// its flag (about once per 2^32 butterflies), so it must not cost the hot loop registers or instruction-cache space.
#if defined(TFHE_EMU)
extern bool g_emu_force_slow;   // emulator tests: take the exact path every time
#define TK_NOINLINE
#define TK_FORCE_SLOW_FLAG g_emu_force_slow
#else
#define TK_NOINLINE __device__ __noinline__
#if defined(TK_FORCE_SLOW)
#define TK_FORCE_SLOW_FLAG true
#else
#define TK_FORCE_SLOW_FLAG false
#endif
#endif
template <int LOGN, int R, int P, bool INV>
TK_NOINLINE void pass_r_slow(u64 *xb, const u64 *tw, int tau) {
    u64 x[R];
    bool unused = false;
    TK_UNROLL
    for (int e = 0; e < R; e++) x[e] = xb[tk_swz(pos_r<LOGN, R, P>(tau, e))];
    pass_r<LOGN, R, P, INV, false>(x, tw, tau, unused);
    TK_UNROLL
    for (int e = 0; e < R; e++) xb[tk_swz(pos_r<LOGN, R, P>(tau, e))] = x[e];
}

// one pass of a transform: optimistic, with the exact recomputation behind a (practically never taken) branch
template <int LOGN, int R, int P, bool INV, bool OPT = true>
TK_DEV void pass_checked_r(u64 (&x)[R], u64 *xb, const u64 *tw, int tau) {
    if constexpr (!OPT) {   // latency kernels with one or two teams per CTA: the flag chain and its branch cost more
        bool none = false;  // on their critical path than the canonicalisation (S2 cluster mode: 7.7 vs 8.1 ms)
        pass_r<LOGN, R, P, INV, false>(x, tw, tau, none);
        return;
    }
#if defined(TK_NO_OPT)
    bool none = false;   // A/B: exact butterflies in line, no fallback
    pass_r<LOGN, R, P, INV, false>(x, tw, tau, none);
    return;
#endif
    bool flag = TK_FORCE_SLOW_FLAG;
    pass_r<LOGN, R, P, INV, true>(x, tw, tau, flag);
    if (flag) {
        pass_r_slow<LOGN, R, P, INV>(xb, tw, tau);
        TK_UNROLL
        for (int e = 0; e < R; e++) x[e] = xb[tk_swz(pos_r<LOGN, R, P>(tau, e))];
    }
}

// Every pass but the first finds its inputs parked in the thread's own slots of xb (it has just read them there), which
// is what lets pass_checked_r recompute it; the first pass of a transform runs the exact butterflies.
template <int LOGN, int R, int P>
TK_DEV void park_r(const u64 (&x)[R], u64 *xb, int tau) {
    TK_UNROLL
    for (int e = 0; e < R; e++) xb[tk_swz(pos_r<LOGN, R, P>(tau, e))] = x[e];
}
template <int LOGN, int R, int P>
TK_DEV void fetch_r(u64 (&x)[R], const u64 *xb, int tau) {
    TK_UNROLL
    for (int e = 0; e < R; e++) x[e] = xb[tk_swz(pos_r<LOGN, R, P>(tau, e))];
}

// forward: in x[e] = a[pos_r<0>(tau, e)] (natural order), out x[e] = A[pos_r<NP-1>(tau, e)] = A[R tau + e]
template <int LOGN, int R, bool OPT = true, int P = 0>
TK_DEV void team_ntt_forward_r(u64 (&x)[R], u64 *xb, const u64 *tw, int tau, int team) {
    constexpr int T = TileR<LOGN, R>::T;
    if constexpr (P == 0) {
        // the first pass has its inputs in registers only: exact butterflies, nothing to park
        bool none = false;
        pass_r<LOGN, R, 0, false, false>(x, tw, tau, none);
        team_sync(team, T);   // the buffer may still be read by its previous user
    } else {
        pass_checked_r<LOGN, R, P, false, OPT>(x, xb, tw, tau);
    }
    if constexpr (P + 1 < TileR<LOGN, R>::NP) {
        park_r<LOGN, R, P>(x, xb, tau);
        team_sync(team, T);
        fetch_r<LOGN, R, P + 1>(x, xb, tau);
        team_ntt_forward_r<LOGN, R, OPT, P + 1>(x, xb, tw, tau, team);
    }
}
// inverse WITHOUT the final 1/N: in x[e] = A[R tau + e], out x[e] = N * a[pos_r<0>(tau, e)] (lazy)
template <int LOGN, int R, bool OPT = true, int P = TileR<LOGN, R>::NP - 1>
TK_DEV void team_ntt_inverse_r(u64 (&x)[R], u64 *xb, const u64 *tw, int tau, int team) {
    constexpr int T = TileR<LOGN, R>::T;
    if constexpr (P == TileR<LOGN, R>::NP - 1) {
        bool none = false;
        pass_r<LOGN, R, P, true, false>(x, tw, tau, none);
        team_sync(team, T);
    } else {
        pass_checked_r<LOGN, R, P, true, OPT>(x, xb, tw, tau);
    }
    if constexpr (P > 0) {
        park_r<LOGN, R, P>(x, xb, tau);
        team_sync(team, T);
        fetch_r<LOGN, R, P - 1>(x, xb, tau);
        team_ntt_inverse_r<LOGN, R, OPT, P - 1>(x, xb, tw, tau, team);
    }
}

// one polynomial per CTA, narrow tile: parity utility for the emulator and the GPU tests
template <int LOGN, int R>
TK_DEV void ntt_r_body(const NttArgs &a, int tid, int bid, unsigned char *smem_raw) {
    constexpr int N = TileR<LOGN, R>::N, T = TileR<LOGN, R>::T, NP = TileR<LOGN, R>::NP;
    u64 *tw = (u64 *)smem_raw, *xb = tw + N;
    for (int i = tid; i < N; i += T) tw[i] = a.tw[i];
    cta_sync();
    const u64 *src = a.in + (size_t)bid * N;
    u64 *dst = a.out + (size_t)bid * N;
    u64 x[R];
    if (!a.inverse) {
        TK_UNROLL
        for (int e = 0; e < R; e++) x[e] = src[pos_r<LOGN, R, 0>(tid, e)];
        team_ntt_forward_r<LOGN, R>(x, xb, tw, tid, 0);
        TK_UNROLL
        for (int e = 0; e < R; e++) dst[pos_r<LOGN, R, NP - 1>(tid, e)] = (a.scale == 1) ? gl_canon(x[e]) : gl_mul(x[e], a.scale);
    } else {
        TK_UNROLL
        for (int e = 0; e < R; e++) x[e] = src[pos_r<LOGN, R, NP - 1>(tid, e)];
        team_ntt_inverse_r<LOGN, R>(x, xb, tw, tid, 0);
        TK_UNROLL
        for (int e = 0; e < R; e++) dst[pos_r<LOGN, R, 0>(tid, e)] = (a.scale == 1) ? gl_canon(x[e]) : gl_mul(x[e], a.scale);
    }
}

// ---------------------------------------------------------------------------- batched NTT kernel

template <int LOGN, int TEAMS>
TK_DEV void ntt_batch_body(const NttArgs &a, int tid, int bid, int nblocks, unsigned char *smem_raw) {
    constexpr int N = Tile<LOGN>::N, T = Tile<LOGN>::T;
    u64 *tw = (u64 *)smem_raw;
    const int team = tid / T, tau = tid % T;
    u64 *xb = tw + N + team * N;
    for (int i = tid; i < N; i += T * TEAMS) tw[i] = a.tw[i];
    cta_sync();
    // polynomials are dealt round-robin to teams; every team of a CTA runs the same trip count
    for (u32 base = (u32)bid * TEAMS; base < a.n_polys; base += (u32)nblocks * TEAMS) {
        const u32 pi = base + team;
        const bool live = pi < a.n_polys;
        const u64 *src = a.in + (size_t)(live ? pi : 0) * N;
        u64 *dst = a.out + (size_t)(live ? pi : 0) * N;
        u64 x[16];
        if (!a.inverse) {
            TK_UNROLL
            for (int e = 0; e < 16; e++) x[e] = src[posA<LOGN>(tau, e)];
            team_ntt_forward<LOGN>(x, xb, tw, tau, team);
            if (live) {
                TK_UNROLL
                for (int e = 0; e < 16; e++)
                    dst[posC<LOGN>(tau, e)] = (a.scale == 1) ? gl_canon(x[e]) : gl_mul(x[e], a.scale);
            }
        } else {
            TK_UNROLL
            for (int e = 0; e < 16; e++) x[e] = src[posC<LOGN>(tau, e)];
            team_ntt_inverse<LOGN>(x, xb, tw, tau, team);
            if (live) {
                TK_UNROLL
                for (int e = 0; e < 16; e++)
                    dst[posA<LOGN>(tau, e)] = (a.scale == 1) ? gl_canon(x[e]) : gl_mul(x[e], a.scale);
            }
        }
    }
}

// ------------------------------------------------------------------- external product / blind rotate
// Resident key format.  A GGSW is 2*ell "row pairs" (one per input polynomial j and level lvl); a row
// pair is [row to output poly 0 : N][row to output poly 1 : N], NTT domain (reference order), with 1/N
// and the minus sign of the mask GLev folded in.  Inside a row the element multiplying register e of
// thread tau sits at key_pos(tau, e), so that one 16-byte access per thread and register pair is
// contiguous across the lanes of a warp (coalesced from L2, conflict-free from shared memory).
template <int LOGN> GL_HD int key_pos(int tau, int e) { return (e >> 1) * (2 * Tile<LOGN>::T) + 2 * tau + (e & 1); }

struct BRArgs {
    u32 n;         // CMUX iterations available = small LWE dimension
    u32 steps;     // iterations actually run (== n; fewer only for per-iteration parity tests)
    u32 ell, logB; // gadget of the external product
    u32 L;         // ceil(64/logB)
    u32 B;         // ciphertexts in the batch
    u32 out_mode;  // 0 = full sample extract [N+1]; 1 = partial extract [n_out+1]; 2 = raw accumulator [2N]
    u32 n_out;     // partial extract length
    const u64 *bsk;     // [n][2*ell row pairs][2][N]  resident key format
    const u64 *ksk;     // one GGSW in the same format, or nullptr          (ivc_based_vpbs.rs:355-371)
    const u64 *lwe;     // [B][n+1] ciphertexts under the small key
    const u64 *testv;   // [n_luts][N]
    const u32 *lut_idx; // [ceil(B/lut_div)] or nullptr: ciphertext ct uses test vector lut_idx[ct / lut_div]
    u32 lut_div;        // >= 1 (1 = one index per ciphertext; the radix layer uses one per slot of B ciphertexts)
    u64 *out;           // [B][out_len]
    const u64 *tw_fwd;  // ROOTS    [N]
    u64 body_add;       // canonical constant added to the extracted body (boolean gates: the +q/8 after the sign bootstrap)
};

// coefficient p of X^sh * poly (negacyclic), poly canonical in shared memory   (vtfhe/mod.rs:80-107)
template <int LOGN>
TK_DEV u64 rotated_coeff(const u64 *poly, int p, u32 sh) {
    constexpr int N = 1 << LOGN;
    const u32 i2 = ((u32)p - sh) & (2 * N - 1);
    const u64 v = poly[i2 & (N - 1)];
    return (i2 & N) ? gl_neg(v) : v;
}

// ---- the CTA-shared stream of key row pairs -------------------------------------------------------
// All teams of a CTA consume the same sequence of row pairs (BSK_0 .. BSK_{steps-1}, then the GLWE key
// switch key).  One elected thread streams them from L2/HBM into a ring of NSTAGE shared-memory stages
// with bulk asynchronous copies (cp.async.bulk -> UBLKCP) that complete on an mbarrier; every warp
// releases a stage through a second mbarrier once it has multiplied with it, so each key byte is fetched
// once per CTA and reused by TEAMS ciphertexts.
// ring depth: as many stages as shared memory leaves next to TEAMS ciphertexts (16N bytes per stage): one at N = 4096,
// three at N = 2048 with at most two ciphertexts per CTA (measured -0.5 % against two), two otherwise
template <int LOGN, int TEAMS = 0> struct StageCfg {
    static constexpr int NSTAGE = (LOGN >= 12) ? 1 : ((LOGN == 11 && TEAMS >= 1 && TEAMS <= 2) ? 3 : 2);
};

struct KeyStream {
    u64 *stage;        // [NSTAGE][2N] shared
    u64 *full, *empty; // [NSTAGE] mbarriers, shared
    const u64 *bsk, *ksk;
    u32 bsk_pairs;     // row pairs coming from bsk; the rest come from ksk
    u32 total;         // row pairs in the whole launch
    u32 t;             // next row pair this thread consumes
    bool producer;
};

#if !defined(TFHE_EMU)
__device__ __forceinline__ u32 tk_smem_addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tk_smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *bar, u32 parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "TK_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra TK_DONE_%=;\n\t"
        "bra TK_WAIT_%=;\n\t"
        "TK_DONE_%=:\n\t"
        "}" ::"r"(tk_smem_addr(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive(u64 *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tk_smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tk_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, u32 bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tk_smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(tk_smem_addr(bar))
                 : "memory");
}

template <int LOGN, int NS>
__device__ __forceinline__ void stream_issue(const KeyStream &ks, u32 t) {
    constexpr int N = 1 << LOGN;
    if (t >= ks.total) return;
    const u64 *src = (t < ks.bsk_pairs) ? ks.bsk + (size_t)t * 2 * N : ks.ksk + (size_t)(t - ks.bsk_pairs) * 2 * N;
    u64 *dst = ks.stage + (size_t)(t % NS) * 2 * N;
    u64 *bar = ks.full + (t % NS);
    constexpr u32 bytes = 2 * N * sizeof(u64), chunk = bytes / 2;
    mbar_expect_tx(bar, bytes);
    bulk_g2s(dst, src, chunk, bar);
    bulk_g2s((unsigned char *)dst + chunk, (const unsigned char *)src + chunk, chunk, bar);
}
#endif

// One GGSW (x) GLWE on the team's accumulator (k = 1).
//   cmux = true : acc <- acc + GGSW (x) (X^sh * acc - acc)            (vtfhe/mod.rs:119-136)
//   cmux = false: acc <- GGSW (x) acc                                  (ivc_based_vpbs.rs:116-121)
// STAGED: the GGSW's row pairs are the next 2*ell entries of the CTA-shared key stream; otherwise they
// are read from `ggsw` in global memory (independent keys per ciphertext: external_product_kernel, emulator).
// ELL / LOGB != 0 fix the gadget at compile time (the named parameter sets): the digit rule then folds to a few
// shifts instead of 64-bit variable shifts and selects per coefficient.  0 = read it from BRArgs.
// R = coefficients per thread (16 = wide tile; 8 / 4 = narrow tiles with N/R threads per team)
template <int LOGN, bool STAGED, int ELL = 0, int LOGB = 0, int R = 16, int TEAMS = 0>
TK_DEV void team_external_product(const BRArgs &a, const u64 *ggsw, KeyStream &ks, bool cmux, u32 sh, u64 *acc, u64 *xb,
                                  const u64 *tw, int tau, int team, int lane) {
    constexpr int N = Tile<LOGN>::N, T = N / R, T16 = N / 16, R16 = 16 / R;
    const u32 g_ell = ELL ? (u32)ELL : a.ell, g_logB = LOGB ? (u32)LOGB : a.logB;
    const u32 g_L = LOGB ? (u32)((64 + (LOGB ? LOGB : 1) - 1) / (LOGB ? LOGB : 1)) : a.L;
    const bool centred = (g_L * g_logB == 64);
    u64 o0[R], o1[R];
    TK_UNROLL
    for (int e = 0; e < R; e++) o0[e] = o1[e] = 0;

    for (int j = 0; j < 2; j++) {
        const u64 *poly = acc + j * N;
#if !defined(TFHE_EMU)
#pragma unroll 1
#endif
        for (u32 lvl = 0; lvl < g_ell; lvl++) {
            u64 x[R];
            TK_UNROLL
            for (int e = 0; e < R; e++) {
                const int p = e * T + tau;   // first-pass layout (posA / pos_r<0>)
                u64 d = poly[p];
                // both operands canonical, so gl_sub's result (a - b, plus q on borrow) already is
                if (cmux) d = gl_sub(rotated_coeff<LOGN>(poly, p, sh), d);
                x[e] = gl_digit(d, g_L - g_ell + lvl, g_logB, centred);   // glev_ct.rs:97-99: top ell limbs
            }
            if constexpr (R == 16)
                team_ntt_forward<LOGN>(x, xb, tw, tau, team);
            else
                team_ntt_forward_r<LOGN, R>(x, xb, tw, tau, team);
            const u64 *rows;
#if !defined(TFHE_EMU)
            constexpr int NS = StageCfg<LOGN, TEAMS>::NSTAGE;
            if (STAGED) {
                mbar_wait(ks.full + (ks.t % NS), (ks.t / NS) & 1);
                rows = ks.stage + (size_t)(ks.t % NS) * 2 * N;
            } else
#endif
                rows = ggsw + ((size_t)(j * g_ell + lvl) * 2) * N;
            TK_UNROLL
            for (int h = 0; h < R / 2; h++) {
                // x[2h], x[2h+1] sit at in-place positions R*tau + 2h (+1) = elements key_pos(tau / R16, (R/2)(tau % R16)
                // + 2h (+1)) of both rows: one 16-byte access each
                const u64 *r0 = rows + (size_t)((R / 2) * (tau & (R16 - 1)) + h) * 2 * T16 + 2 * (tau / R16);
                const u64 *r1 = r0 + N;
#if !defined(TFHE_EMU)
                const ulonglong2 v0 = *reinterpret_cast<const ulonglong2 *>(r0);
                const ulonglong2 v1 = *reinterpret_cast<const ulonglong2 *>(r1);
                const u64 a0 = v0.x, a1 = v0.y, b0 = v1.x, b1 = v1.y;
#else
                const u64 a0 = r0[0], a1 = r0[1], b0 = r1[0], b1 = r1[1];
#endif
                o0[2 * h] = gl_mac(o0[2 * h], x[2 * h], a0);
                o0[2 * h + 1] = gl_mac(o0[2 * h + 1], x[2 * h + 1], a1);
                o1[2 * h] = gl_mac(o1[2 * h], x[2 * h], b0);
                o1[2 * h + 1] = gl_mac(o1[2 * h + 1], x[2 * h + 1], b1);
            }
#if !defined(TFHE_EMU)
            if (STAGED) {
                // release the stage (one arrival per warp), then the producer refills the stage that was
                // released one step earlier with the row pair NSTAGE steps ahead of it
                __syncwarp();
                if (lane == 0) mbar_arrive(ks.empty + (ks.t % NS));
                if (ks.producer && ks.t + 1 >= (u32)NS) {
                    const u32 r = ks.t + 1 - NS;
                    mbar_wait(ks.empty + (r % NS), (r / NS) & 1);
                    stream_issue<LOGN, NS>(ks, r + NS);
                }
                ks.t++;
            }
#endif
        }
    }
    // both output polynomials go through ONE inverse-NTT body (the instruction footprint of the loop is
    // what the three teams of an SM fetch over and over)
#if !defined(TFHE_EMU)
#pragma unroll 1
#endif
    for (int r = 0; r < 2; r++) {
        if constexpr (R == 16)
            team_ntt_inverse<LOGN>(o0, xb, tw, tau, team);
        else
            team_ntt_inverse_r<LOGN, R>(o0, xb, tw, tau, team);
        TK_UNROLL
        for (int e = 0; e < R; e++) {
            const int p = r * N + e * T + tau;
            acc[p] = cmux ? gl_canon(gl_add(o0[e], acc[p])) : gl_canon(o0[e]);
            o0[e] = o1[e];
        }
    }
    team_sync_outer(team, T);  // next iteration gathers rotated coefficients written by other threads
}

// shared memory: [tw N] then per team [acc 2N][xb N], then the key ring [NSTAGE][2N] and its mbarriers
template <int LOGN, int TEAMS>
constexpr size_t blind_rotate_smem_bytes() {
    return sizeof(u64) * ((size_t)Tile<LOGN>::N + (size_t)TEAMS * 3 * Tile<LOGN>::N + (size_t)StageCfg<LOGN, TEAMS>::NSTAGE * 2 * Tile<LOGN>::N) + 64;
}

template <int LOGN, int TEAMS, int ELL = 0, int LOGB = 0, int R = 16>
TK_DEV void blind_rotate_body(const BRArgs &a, int tid, int bid, unsigned char *smem_raw) {
    constexpr int N = Tile<LOGN>::N, T = N / R, NS = StageCfg<LOGN, TEAMS>::NSTAGE;
    u64 *tw = (u64 *)smem_raw;
    const int team = tid / T, tau = tid % T, lane = tid & 31;
    u64 *acc = tw + N + (size_t)team * 3 * N;
    u64 *xb = acc + 2 * N;
    KeyStream ks{};
    ks.stage = tw + N + (size_t)TEAMS * 3 * N;
    ks.full = ks.stage + (size_t)NS * 2 * N;
    ks.empty = ks.full + NS;
    ks.bsk = a.bsk;
    ks.ksk = a.ksk;
    ks.bsk_pairs = a.steps * 2 * a.ell;
    ks.total = ks.bsk_pairs + (a.ksk ? 2 * a.ell : 0);
    ks.t = 0;
    ks.producer = (tid == 0);
    for (int i = tid; i < N; i += T * TEAMS) tw[i] = a.tw_fwd[i];
#if !defined(TFHE_EMU)
    {
        const u32 left = a.B - (u32)bid * TEAMS;
        const u32 live_teams = left < (u32)TEAMS ? left : (u32)TEAMS;
        if (tid == 0) {
            for (int s = 0; s < NS; s++) {
                mbar_init(ks.full + s, 1);
                mbar_init(ks.empty + s, live_teams * (T / 32));
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
#endif
    cta_sync();
    const u32 ct = (u32)bid * TEAMS + team;
    // team handle for the barriers: with lock step it carries the live thread count of this CTA (ragged last CTA)
#if !defined(TFHE_EMU)
    // (only the compile-time-gadget instantiations: the longer runtime-gadget body measured slower in lock step)
    // and only the wide tile: the narrow-tile body is half the size and measured faster with drifting teams)
#ifndef TK_LOCKSTEP_NARROW
#define TK_LOCKSTEP_NARROW 0
#endif
    const int th = tk_team(team, (ELL && (R == 16 || TK_LOCKSTEP_NARROW)) ? (int)((a.B - (u32)bid * TEAMS < (u32)TEAMS ? a.B - (u32)bid * TEAMS : (u32)TEAMS) * T) : 0);
#else
    const int th = team;
#endif
    if (ct >= a.B) return;  // whole team leaves; only team barriers and the key-stream mbarriers are used from here on
#if !defined(TFHE_EMU)
    if (ks.producer)
        for (u32 t = 0; t < (u32)NS; t++) stream_issue<LOGN, NS>(ks, t);
#endif

    const u64 *lwe = a.lwe + (size_t)ct * (a.n + 1);
    const u64 *tv = a.testv + (size_t)(a.lut_idx ? a.lut_idx[ct / a.lut_div] : 0) * N;

    // acc0 = X^{ms(-b)} * (0, testv)      ivc_based_vpbs.rs:109-124 (first_step negates the body)
    {
        const u32 sh0 = gl_mod_switch(gl_neg(lwe[a.n]), LOGN);
        TK_UNROLL
        for (int e = 0; e < R; e++) {
            const int p = e * T + tau;
            acc[p] = 0;
            acc[N + p] = rotated_coeff<LOGN>(tv, p, sh0);
        }
    }
    team_sync(th, T);

    // steps CMUX iterations, then (in-tree order) the GLWE key switch as one more external product: one
    // call site, so the loop body exists once in the instruction stream
    const size_t ggsw_len = (size_t)4 * a.ell * N;
    const u32 iters = a.steps + (a.ksk ? 1u : 0u);
    u64 mask = a.steps ? lwe[0] : 0;
    for (u32 i = 0; i < iters; i++) {
        const bool cmux = i < a.steps;
        const u32 sh = cmux ? gl_mod_switch(mask, LOGN) : 0;
        if (i + 1 < a.steps) mask = lwe[i + 1];  // fetched one iteration ahead
        team_external_product<LOGN, true, ELL, LOGB, R, TEAMS>(a, cmux ? a.bsk + i * ggsw_len : a.ksk, ks, cmux, sh, acc, xb, tw, tau, th, lane);
    }

    // exit: sample extraction of coefficient 0      crypto/glwe.rs:96-113
    if (a.out_mode == 2) {
        u64 *o = a.out + (size_t)ct * 2 * N;
        for (int p = tau; p < 2 * N; p += T) o[p] = acc[p];
    } else {
        const u32 len = a.out_mode == 1 ? a.n_out : (u32)N;
        u64 *o = a.out + (size_t)ct * (len + 1);
        for (u32 p = tau; p < len; p += T) o[p] = p ? gl_neg(acc[N - p]) : acc[0];
        if (tau == 0) o[len] = gl_canon(gl_add(acc[N], a.body_add));
    }
}

// ---- latency mode: ONE ciphertext per CTA, its external product spread over G = 2*ell teams --------------
// The throughput kernel above gives a ciphertext to one team, which runs the 2*ell forward and 2 inverse NTTs
// of a CMUX one after the other.  When the batch is smaller than the GPU (B <= #SMs) that serial chain is the
// latency of a bootstrap, so here team g = (j, lvl) transforms digit polynomial (j, lvl), multiplies it with
// its row pair of the GGSW (read straight from L2: every team needs a different pair at the same time) and
// leaves two partial output polynomials in shared memory; teams 0 and 1 then add up the G partials of output
// polynomial 0 / 1, run the inverse NTT and update the accumulator.  Per CMUX the critical path is one forward
// and one inverse NTT instead of 2*ell + 2.  Sums are exact in Z_q, so results are bit-identical to the
// throughput kernel's.
// shared memory: [tw N][acc 2N] then per team a 2N region: forward exchange buffer (first N), then the team's
// partials [2][N]; inverse team r reuses half r of its own region as exchange buffer (the other inverse team
// only reads the other half).
template <int LOGN, int G>
constexpr size_t blind_rotate_split_smem_bytes() {
    return sizeof(u64) * ((size_t)3 * Tile<LOGN>::N + (size_t)G * 2 * Tile<LOGN>::N);
}

template <int LOGN, int G, int ELL = 0, int LOGB = 0, int R = 16>
TK_DEV void blind_rotate_split_body(const BRArgs &a, int tid, int bid, unsigned char *smem_raw) {
    constexpr int N = Tile<LOGN>::N, T = N / R, T16 = N / 16, R16 = 16 / R;
    u64 *tw = (u64 *)smem_raw;
    u64 *acc = tw + N;
    u64 *part = acc + 2 * N;  // [G][2][N]
    const int g = tid / T, tau = tid % T;
    const u32 g_ell = ELL ? (u32)ELL : a.ell, g_logB = LOGB ? (u32)LOGB : a.logB;
    const u32 g_L = LOGB ? (u32)((64 + (LOGB ? LOGB : 1) - 1) / (LOGB ? LOGB : 1)) : a.L;
    const int j = g / (int)g_ell;
    const u32 lvl = (u32)g % g_ell;
    u64 *mine = part + (size_t)g * 2 * N;
    const bool centred = (g_L * g_logB == 64);
    for (int i = tid; i < N; i += T * G) tw[i] = a.tw_fwd[i];
    const u32 ct = (u32)bid;
    const u64 *lwe = a.lwe + (size_t)ct * (a.n + 1);
    const u64 *tv = a.testv + (size_t)(a.lut_idx ? a.lut_idx[ct / a.lut_div] : 0) * N;
    {
        const u32 sh0 = gl_mod_switch(gl_neg(lwe[a.n]), LOGN);
        for (int p = tid; p < N; p += T * G) {
            acc[p] = 0;
            acc[N + p] = rotated_coeff<LOGN>(tv, p, sh0);
        }
    }
    cta_sync();
    const size_t ggsw_len = (size_t)4 * g_ell * N;
    const u32 iters = a.steps + (a.ksk ? 1u : 0u);
    u64 mask = a.steps ? lwe[0] : 0;
    for (u32 i = 0; i < iters; i++) {
        const bool cmux = i < a.steps;
        const u32 sh = cmux ? gl_mod_switch(mask, LOGN) : 0;
        if (i + 1 < a.steps) mask = lwe[i + 1];
        {   // phase 1: every team, its own digit polynomial
            const u64 *poly = acc + j * N;
            u64 x[R];
            TK_UNROLL
            for (int e = 0; e < R; e++) {
                const int p = e * T + tau;
                u64 d = poly[p];
                if (cmux) d = gl_sub(rotated_coeff<LOGN>(poly, p, sh), d);
                x[e] = gl_digit(d, g_L - g_ell + lvl, g_logB, centred);
            }
            if constexpr (R == 16)
                team_ntt_forward<LOGN>(x, mine, tw, tau, g);
            else
                team_ntt_forward_r<LOGN, R, (G >= 4)>(x, mine, tw, tau, g);
            const u64 *rows = (cmux ? a.bsk + i * ggsw_len : a.ksk) + (size_t)g * 2 * N;
            team_sync(g, T);  // the exchange buffer is about to be overwritten with the partials
            TK_UNROLL
            for (int h = 0; h < R / 2; h++) {
                const u64 *r0 = rows + (size_t)((R / 2) * (tau & (R16 - 1)) + h) * 2 * T16 + 2 * (tau / R16);
                const u64 *r1 = r0 + N;
#if !defined(TFHE_EMU)
                const ulonglong2 v0 = *reinterpret_cast<const ulonglong2 *>(r0);
                const ulonglong2 v1 = *reinterpret_cast<const ulonglong2 *>(r1);
                const u64 a0 = v0.x, a1 = v0.y, b0 = v1.x, b1 = v1.y;
#else
                const u64 a0 = r0[0], a1 = r0[1], b0 = r1[0], b1 = r1[1];
#endif
                mine[(2 * h) * T + tau] = gl_mul(x[2 * h], a0);
                mine[(2 * h + 1) * T + tau] = gl_mul(x[2 * h + 1], a1);
                mine[N + (2 * h) * T + tau] = gl_mul(x[2 * h], b0);
                mine[N + (2 * h + 1) * T + tau] = gl_mul(x[2 * h + 1], b1);
            }
        }
        cta_sync();
        if (g < 2) {  // phase 2: team r = g owns output polynomial r
            u64 o[R];
            TK_UNROLL
            for (int e = 0; e < R; e++) o[e] = part[(size_t)g * N + e * T + tau];
            for (int gg = 1; gg < G; gg++) {
                const u64 *src = part + ((size_t)gg * 2 + g) * N;
                TK_UNROLL
                for (int e = 0; e < R; e++) o[e] = gl_add(o[e], src[e * T + tau]);
            }
            if constexpr (R == 16)
                team_ntt_inverse<LOGN>(o, mine + (size_t)g * N, tw, tau, g);
            else
                team_ntt_inverse_r<LOGN, R, (G >= 4)>(o, mine + (size_t)g * N, tw, tau, g);
            TK_UNROLL
            for (int e = 0; e < R; e++) {
                const int p = g * N + e * T + tau;
                acc[p] = cmux ? gl_canon(gl_add(o[e], acc[p])) : gl_canon(o[e]);
            }
        }
        cta_sync();
    }
    if (a.out_mode == 2) {
        u64 *o = a.out + (size_t)ct * 2 * N;
        for (int p = tid; p < 2 * N; p += T * G) o[p] = acc[p];
    } else {
        const u32 len = a.out_mode == 1 ? a.n_out : (u32)N;
        u64 *o = a.out + (size_t)ct * (len + 1);
        for (u32 p = tid; p < len; p += T * G) o[p] = p ? gl_neg(acc[N - p]) : acc[0];
        if (tid == 0) o[len] = gl_canon(gl_add(acc[N], a.body_add));
    }
}

// ---- cluster latency mode: ONE ciphertext per CLUSTER of two CTAs (two SMs) ------------------------------------
// The latency-mode kernel above still runs the 2*ell forward and the 2 inverse transforms of a CMUX on one SM, whose
// integer pipes are what bounds it.  Here CTA r of a 2-CTA thread-block cluster owns accumulator polynomial r:
//   phase 1  its ell teams decompose (X^a * acc_r - acc_r), transform the digits and multiply them with the row pairs
//            (r, lvl) of the GGSW: partial products towards BOTH output polynomials, left in its own shared memory;
//   cluster barrier (release/acquire);
//   phase 2  team 0 adds the ell local partials of output r and the ell partials of the peer CTA (read through
//            distributed shared memory, ld.shared::cluster), runs the inverse transform and updates acc_r.
// Partials are double-buffered by iteration parity, so one cluster barrier per CMUX orders everything: a CTA
// overwrites buffer `par` two iterations later, after its peer passed the barrier in between, i.e. after the peer's
// reads.  Exact sums in Z_q: same words as the other two kernels.
// shared memory per CTA: [tw N][acc_r N] then per team [xb N][part: 2 parities x 2 outputs x N]
#if !defined(TFHE_EMU)
template <int LOGN, int TPC>
constexpr size_t blind_rotate_cluster_smem_bytes() {
    return sizeof(u64) * ((size_t)2 * Tile<LOGN>::N + (size_t)TPC * 5 * Tile<LOGN>::N);
}
__device__ __forceinline__ u32 cluster_ctarank() {
    u32 r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ u32 cluster_map(const void *p, u32 rank) {
    u32 out;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(tk_smem_addr(p)), "r"(rank));
    return out;
}
__device__ __forceinline__ u64 cluster_ld(u32 addr) {
    u64 v;
    asm volatile("ld.shared::cluster.u64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
    return v;
}

// R = coefficients per thread: 16 (the wide tile of the throughput kernel) or 8 / 4 (narrow tiles, R/16 of the
// dependent instruction stream per thread and 16/R times the warps per polynomial)
template <int LOGN, int TPC, int ELL = 0, int LOGB = 0, int R = 16>
__device__ __forceinline__ void blind_rotate_cluster_body(const BRArgs &a, int tid, int bid, unsigned char *smem_raw) {
    constexpr int N = Tile<LOGN>::N, T = N / R, T16 = N / 16, R16 = 16 / R;
    const u32 g_ell = ELL ? (u32)ELL : a.ell, g_logB = LOGB ? (u32)LOGB : a.logB;
    const u32 g_L = LOGB ? (u32)((64 + (LOGB ? LOGB : 1) - 1) / (LOGB ? LOGB : 1)) : a.L;
    const bool centred = (g_L * g_logB == 64);
    const u32 r = cluster_ctarank();   // which accumulator polynomial this CTA owns (0 = mask, 1 = body)
    const u32 ct = (u32)bid >> 1;
    u64 *tw = (u64 *)smem_raw;
    u64 *acc = tw + N;                 // acc_r only
    const int lvl = tid / T, tau = tid % T;
    u64 *xb = acc + N + (size_t)lvl * 5 * N;
    u64 *part = xb + N;                // [2 parities][2 outputs][N]
    for (int i = tid; i < N; i += T * TPC) tw[i] = a.tw_fwd[i];
    const u64 *lwe = a.lwe + (size_t)ct * (a.n + 1);
    const u64 *tv = a.testv + (size_t)(a.lut_idx ? a.lut_idx[ct / a.lut_div] : 0) * N;
    {
        const u32 sh0 = gl_mod_switch(gl_neg(lwe[a.n]), LOGN);
        for (int p = tid; p < N; p += T * TPC) acc[p] = r ? rotated_coeff<LOGN>(tv, p, sh0) : 0;
    }
    cta_sync();
    // peer's partial buffers towards MY output polynomial: same offsets in the peer's shared memory
    const u32 peer_part0 = cluster_map(acc + N + N, r ^ 1u);   // team 0's `part` of the peer
    const size_t ggsw_len = (size_t)4 * g_ell * N;
    const u32 iters = a.steps + (a.ksk ? 1u : 0u);
    u64 mask = a.steps ? lwe[0] : 0;
    for (u32 i = 0; i < iters; i++) {
        const bool cmux = i < a.steps;
        const u32 sh = cmux ? gl_mod_switch(mask, LOGN) : 0;
        if (i + 1 < a.steps) mask = lwe[i + 1];
        const u32 par = i & 1u;
        {   // phase 1: team lvl, digit polynomial (r, lvl)
            u64 x[R];
            TK_UNROLL
            for (int e = 0; e < R; e++) {
                const int p = e * T + tau;   // first-pass layout (posA / pos_r<0>)
                u64 d = acc[p];
                if (cmux) d = gl_sub(rotated_coeff<LOGN>(acc, p, sh), d);
                x[e] = gl_digit(d, g_L - g_ell + (u32)lvl, g_logB, centred);
            }
            if constexpr (R == 16)
                team_ntt_forward<LOGN>(x, xb, tw, tau, lvl);
            else
                team_ntt_forward_r<LOGN, R, (TPC >= 2)>(x, xb, tw, tau, lvl);
            const u64 *rows = (cmux ? a.bsk + i * ggsw_len : a.ksk) + ((size_t)r * g_ell + (u32)lvl) * 2 * N;
            u64 *mine = part + (size_t)par * 2 * N;
            TK_UNROLL
            for (int h = 0; h < R / 2; h++) {
                // x[2h], x[2h+1] sit at in-place positions R*tau + 2h (+1) = element (tau / R16, (R/2)(tau % R16) + h)
                // of the resident key format (key_pos)
                const u64 *r0 = rows + (size_t)((R / 2) * (tau & (R16 - 1)) + h) * 2 * T16 + 2 * (tau / R16);
                const ulonglong2 v0 = *reinterpret_cast<const ulonglong2 *>(r0);
                const ulonglong2 v1 = *reinterpret_cast<const ulonglong2 *>(r0 + N);
                mine[(2 * h) * T + tau] = gl_mul(x[2 * h], v0.x);
                mine[(2 * h + 1) * T + tau] = gl_mul(x[2 * h + 1], v0.y);
                mine[N + (2 * h) * T + tau] = gl_mul(x[2 * h], v1.x);
                mine[N + (2 * h + 1) * T + tau] = gl_mul(x[2 * h + 1], v1.y);
            }
        }
        cluster_sync_all();
        if (lvl == 0) {  // phase 2: output polynomial r = my partials towards r + the peer's partials towards r
            u64 o[R];
            const size_t off = ((size_t)par * 2 + r) * N;
            TK_UNROLL
            for (int e = 0; e < R; e++) o[e] = part[off + e * T + tau];
            for (u32 l2 = 1; l2 < g_ell; l2++) {
                const u64 *src = part + (size_t)l2 * 5 * N + off;
                TK_UNROLL
                for (int e = 0; e < R; e++) o[e] = gl_add(o[e], src[e * T + tau]);
            }
            for (u32 l2 = 0; l2 < g_ell; l2++) {
                const u32 base = peer_part0 + (u32)(((size_t)l2 * 5 * N + off) * sizeof(u64));
                u64 v[R];
                TK_UNROLL
                for (int e = 0; e < R; e++) v[e] = cluster_ld(base + (u32)((e * T + tau) * sizeof(u64)));
                TK_UNROLL
                for (int e = 0; e < R; e++) o[e] = gl_add(o[e], v[e]);
            }
            if constexpr (R == 16)
                team_ntt_inverse<LOGN>(o, xb, tw, tau, 0);
            else
                team_ntt_inverse_r<LOGN, R, (TPC >= 2)>(o, xb, tw, tau, 0);
            TK_UNROLL
            for (int e = 0; e < R; e++) {
                const int p = e * T + tau;
                acc[p] = cmux ? gl_canon(gl_add(o[e], acc[p])) : gl_canon(o[e]);
            }
        }
        if (TPC == 1)
            team_sync(0, T);
        else
            cta_sync();
    }
    cluster_sync_all();   // the peer may still be reading my partials of the last iteration
    if (a.out_mode == 2) {
        u64 *o = a.out + (size_t)ct * 2 * N + (size_t)r * N;
        for (int p = tid; p < N; p += T * TPC) o[p] = acc[p];
    } else {
        const u32 len = a.out_mode == 1 ? a.n_out : (u32)N;
        u64 *o = a.out + (size_t)ct * (len + 1);
        if (r == 0) {
            for (u32 p = tid; p < len; p += T * TPC) o[p] = p ? gl_neg(acc[N - p]) : acc[0];
        } else if (tid == 0) {
            o[len] = gl_canon(gl_add(acc[0], a.body_add));
        }
    }
}
#endif

// ---------------------------------------------------------------------------------- LWE key switch
// out[b] = (0,..,0,body_b) + sum_i sum_j digit_j(a_{b,i}) * KSK[i][j]         (vtfhe/mod.rs:31-54)
// Signed digits d in [-B/2, B/2] are offset to d' = d + B/2 >= 0 so that every product is an
// unsigned 32x32 IMAD.WIDE into two 64-bit lanes (low and high KSK halves); the offset is removed
// once at the end with ksum[x] = sum of all KSK rows.  Exact in Z_q, hence bit-identical to the
// reference's field arithmetic.
struct KSArgs {
    u32 n;       // output dimension
    u32 kN;      // input dimension
    u32 ell, logB, L;
    u32 B;
    const u64 *ksk;   // [kN][ell][n+1]
    const u64 *ksum;  // [n+1] column sums of ksk (mod q)
    const u64 *in;    // [B][kN+1]
    u64 *out;         // [B][n+1]
    u32 splits;       // > 1: the kN input coefficients are cut into `splits` ranges (grid z); each CTA writes its exact
    u64 *partial;     //      partial sum to partial[z][B][n+1] and ks_finish_body adds them up (small batches: latency)
};

GL_HD u64 ks_fold(u64 a0, u64 a1) {  // a0 + a1 * 2^32 -> canonical
    u64 lo = a0 + (a1 << 32);
    u64 hi = (a1 >> 32) + (lo < a0 ? 1 : 0);
    return gl_canon(gl_reduce128(lo, hi));
}

template <int TB, int TX, int RC>
TK_DEV void lwe_keyswitch_body(const KSArgs &a, int tid, int bx, int by, unsigned char *smem_raw, int bz = 0) {
    u32 *dig = (u32 *)smem_raw;  // [RC*ell][TB]
    const u32 col = (u32)bx * TX + tid;
    const u32 b0 = (u32)by * TB;
    const bool centred = (a.L * a.logB == 64);
    const u32 halfB = 1u << (a.logB - 1);
    // rows that fit before the 64-bit lanes could overflow: each product < 2^(logB+1+32)
    const u32 max_rows = (a.logB + 33 >= 63) ? 1u : (1u << (63 - a.logB - 33 > 20 ? 20 : 63 - a.logB - 33));
    u64 A0[TB], A1[TB], F[TB];
    TK_UNROLL
    for (int b = 0; b < TB; b++) A0[b] = A1[b] = F[b] = 0;
    u32 rows_pending = 0;
    // input-coefficient range of this CTA: all of them, or slice bz of `splits` (multiples of RC)
    const u32 nsplit = a.splits > 1 ? a.splits : 1;
    const u32 per = ((a.kN + nsplit - 1) / nsplit + RC - 1) / RC * RC;
    const u32 c_begin = (u32)bz * per < a.kN ? (u32)bz * per : a.kN;
    const u32 c_end = c_begin + per < a.kN ? c_begin + per : a.kN;
    for (u32 c0 = c_begin; c0 < c_end; c0 += RC) {
        const u32 rc = (c_end - c0 < RC) ? c_end - c0 : RC;
        cta_sync();
        for (u32 idx = tid; idx < rc * TB; idx += TX) {
            const u32 i = idx / TB, b = idx % TB;
            const u64 x = (b0 + b < a.B) ? a.in[(size_t)(b0 + b) * (a.kN + 1) + c0 + i] : 0;
            for (u32 j = 0; j < a.ell; j++) {
                const u64 d = gl_digit(x, a.L - a.ell + j, a.logB, centred);  // canonical field element
                const u32 ds = (d > GL_Q / 2) ? (u32)(halfB - (u32)(GL_Q - d)) : (u32)(halfB + (u32)d);
                dig[(i * a.ell + j) * TB + b] = ds;
            }
        }
        cta_sync();
        const u32 rows = rc * a.ell;
        for (u32 r0 = 0; r0 < rows; r0 += max_rows) {
            const u32 r1 = (rows - r0 < max_rows) ? rows : r0 + max_rows;
            if (rows_pending + (r1 - r0) > max_rows) {
                TK_UNROLL
                for (int b = 0; b < TB; b++) {
                    F[b] = gl_add(F[b], ks_fold(A0[b], A1[b]));
                    A0[b] = A1[b] = 0;
                }
                rows_pending = 0;
            }
            for (u32 r = r0; r < r1; r++) {
                const u64 kv = (col <= a.n) ? a.ksk[((size_t)c0 * a.ell + r) * (a.n + 1) + col] : 0;
                const u32 klo = (u32)kv, khi = (u32)(kv >> 32);
                const u32 *dr = dig + r * TB;
                TK_UNROLL
                for (int b = 0; b < TB; b++) {
                    A0[b] += (u64)dr[b] * klo;
                    A1[b] += (u64)dr[b] * khi;
                }
            }
            rows_pending += r1 - r0;
        }
    }
    if (nsplit > 1) {
        if (col <= a.n)
            for (int b = 0; b < TB; b++) {
                if (b0 + b >= a.B) break;
                a.partial[((size_t)bz * a.B + b0 + b) * (a.n + 1) + col] = gl_canon(gl_add(F[b], ks_fold(A0[b], A1[b])));
            }
        return;
    }
    if (col <= a.n) {
        const u64 off = gl_mul((u64)halfB, a.ksum[col]);
        for (int b = 0; b < TB; b++) {
            if (b0 + b >= a.B) break;
            u64 v = gl_canon(gl_add(F[b], ks_fold(A0[b], A1[b])));
            v = gl_canon(gl_sub(v, off));
            if (col == a.n) v = gl_canon(gl_add(v, a.in[(size_t)(b0 + b) * (a.kN + 1) + a.kN]));
            a.out[(size_t)(b0 + b) * (a.n + 1) + col] = v;
        }
    }
}

// split key switch, second step: out[b][col] = sum_z partial[z][b][col] - halfB * ksum[col] (+ body)
TK_DEV void ks_finish_body(const KSArgs &a, size_t idx) {
    const u32 cols = a.n + 1;
    if (idx >= (size_t)a.B * cols) return;
    const u32 col = (u32)(idx % cols);
    const size_t b = idx / cols;
    u64 v = 0;
    for (u32 z = 0; z < a.splits; z++) v = gl_canon(gl_add(v, a.partial[(size_t)z * a.B * cols + idx]));
    const u64 off = gl_mul((u64)(1u << (a.logB - 1)), a.ksum[col]);
    v = gl_canon(gl_sub(v, off));
    if (col == a.n) v = gl_canon(gl_add(v, a.in[b * (a.kN + 1) + a.kN]));
    a.out[idx] = v;
}

// column sums of the key-switch key: ksum[x] = sum_r ksk[r][x] mod q        (one thread per column)
TK_DEV void ksk_colsum_body(const u64 *ksk, u64 *ksum, u32 rows, u32 cols, u32 col) {
    if (col >= cols) return;
    u64 s = 0;
    for (u32 r = 0; r < rows; r++) s = gl_canon(gl_add(s, ksk[(size_t)r * cols + col]));
    ksum[col] = s;
}

// fold 1/N and the sign of the mask GLevs into an NTT-domain GGSW array (vtfhe/ggsw_ct.rs:108-111:
// result = GLev_k - sum_{j<k} GLev_j, and crypto/poly.rs:62-63: final * NINV).  k = 1.
TK_DEV u64 ggsw_fold_elem(u64 v, u64 ninv, u32 glev_index /* 0 = mask, 1 = body */) {
    const u64 s = gl_mul(v, ninv);
    return glev_index == 0 ? gl_neg(s) : s;
}

// reference-order NTT-domain polynomial (element of thread tau, register e at posC = 16*tau + e) -> resident
// key row: fold 1/N and sign, move to key_pos(tau, e).  `x` holds the thread's 16 elements.
template <int LOGN>
TK_DEV void ggsw_row_to_key_format(const u64 *src, u64 *dst, int tau, u64 ninv, u32 glev_index) {
    u64 x[16];
    TK_UNROLL
    for (int e = 0; e < 16; e++) x[e] = src[posC<LOGN>(tau, e)];
    TK_UNROLL
    for (int e = 0; e < 16; e++) dst[key_pos<LOGN>(tau, e)] = ggsw_fold_elem(x[e], ninv, glev_index);
}
