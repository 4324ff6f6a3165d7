// gl64.cuh — Goldilocks field (q = 2^64 - 2^32 + 1) arithmetic for sm_100a.
//
// The reference's field is plonky2's GoldilocksField (research/verifiable-fhe/Cargo.toml:7,
// src/ntt/gen_param_file.sage:118).  Representation used on the device:
//   "canonical"  u64 in [0, q)
//   "lazy"       any u64, standing for its residue mod q
// Every routine states which it takes and returns; data at rest in HBM/SMEM is canonical.
//
// The same header compiles as plain host C++ (TFHE_EMU) so that kernels can be run by a
// host-thread emulator in the CPU test-suite; there the PTX paths fall back to portable C.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__) && !defined(TFHE_EMU)
#define GL_DEV __device__ __forceinline__
#define GL_HD __host__ __device__ __forceinline__
#else
#define GL_DEV inline
#define GL_HD inline
#endif

typedef uint64_t u64;
typedef uint32_t u32;

#define GL_Q 0xFFFFFFFF00000001ULL
#define GL_EPS 0xFFFFFFFFULL  // 2^64 mod q

#if defined(__CUDA_ARCH__) && !defined(TFHE_EMU) && !defined(GL_PORTABLE)
#define GL_PTX 1
#else
#define GL_PTX 0
#endif

// ---- 64x64 -> 128 multiply -------------------------------------------------------------
GL_HD void gl_mul_wide(u64 a, u64 b, u64 &lo, u64 &hi) {
#if defined(__CUDA_ARCH__) && !defined(TFHE_EMU)
    lo = a * b;
    hi = __umul64hi(a, b);
#else
    unsigned __int128 p = (unsigned __int128)a * b;
    lo = (u64)p;
    hi = (u64)(p >> 64);
#endif
}

// ---- lazy add / sub ----------------------------------------------------------------------
// a lazy, b canonical -> lazy.  (a + b) wraps at most once; the wrap 2^64 == eps (mod q) is
// added back and cannot wrap again because a + b - 2^64 < q.
GL_HD u64 gl_add(u64 a, u64 b) {
#if GL_PTX
    // c = carry; s += c * eps  ==  s - c + (c << 32)   (mod 2^64; cannot wrap as a whole, see above)
    // (the carry of an add chain must not be fed to subc: ptxas keeps borrows inverted)
    u32 lo, hi;
    asm("{\n\t"
        ".reg .u32 c;\n\t"
        "add.cc.u32 %0, %2, %4;\n\t"
        "addc.cc.u32 %1, %3, %5;\n\t"
        "addc.u32 c, 0, 0;\n\t"
        "sub.cc.u32 %0, %0, c;\n\t"
        "subc.u32 %1, %1, 0;\n\t"
        "add.u32 %1, %1, c;\n\t"
        "}"
        : "=&r"(lo), "=&r"(hi)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return ((u64)hi << 32) | lo;
#else
    u64 s = a + b;
    return s + ((s < a) ? GL_EPS : 0);
#endif
}

// a lazy, b canonical -> lazy.  A borrow means the true value is d - 2^64 == d - eps.
// `subc.u32 m, 0, 0` after the chain yields m = 0 - borrow = 0 or 0xFFFFFFFF = eps, so the
// correction is a plain 64-bit subtract of the zero-extended mask (never borrows again: a borrow
// implies d >= 2^64 - q + 1 > eps).
GL_HD u64 gl_sub(u64 a, u64 b) {
#if GL_PTX
    u32 lo, hi;
    asm("{\n\t"
        ".reg .u32 m;\n\t"
        "sub.cc.u32 %0, %2, %4;\n\t"
        "subc.cc.u32 %1, %3, %5;\n\t"
        "subc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 %0, %0, m;\n\t"
        "subc.u32 %1, %1, 0;\n\t"
        "}"
        : "=&r"(lo), "=&r"(hi)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return ((u64)hi << 32) | lo;
#else
    u64 d = a - b;
    return d - ((a < b) ? GL_EPS : 0);
#endif
}

// lazy -> canonical
GL_HD u64 gl_canon(u64 a) {
#if GL_PTX
    // a >= q  <=>  high word all ones and low word non-zero; then a - q = (low - 1, 0)
    u32 lo = (u32)a, hi = (u32)(a >> 32);
    const bool p = (hi == 0xFFFFFFFFu) && (lo != 0);
    if (p) {
        lo -= 1;
        hi = 0;
    }
    return ((u64)hi << 32) | lo;
#else
    return a >= GL_Q ? a - GL_Q : a;
#endif
}

// canonical -> canonical
GL_HD u64 gl_neg(u64 a) { return a ? GL_Q - a : 0; }

// 128-bit (lo,hi) -> lazy.   x = lo + hl*2^64 + hh*2^96 == lo + hl*eps - hh  (mod q)
GL_HD u64 gl_reduce128(u64 lo, u64 hi) {
#if GL_PTX
    // t = lo + hl*eps (carry -> += eps, cannot carry again: t <= 2^64 - 2^33 then), then t -= hh
    // (borrow -> -= eps, cannot borrow again: t > 2^64 - 2^32 then).  Corrections as in gl_add / gl_sub.
    u32 t0, t1;
    asm("{\n\t"
        ".reg .u32 m;\n\t"
        "mad.lo.cc.u32 %0, %2, 0xFFFFFFFF, %3;\n\t"
        "madc.hi.cc.u32 %1, %2, 0xFFFFFFFF, %4;\n\t"
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 %0, %0, m;\n\t"
        "subc.u32 %1, %1, 0;\n\t"
        "add.u32 %1, %1, m;\n\t"
        "sub.cc.u32 %0, %0, %5;\n\t"
        "subc.cc.u32 %1, %1, 0;\n\t"
        "subc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 %0, %0, m;\n\t"
        "subc.u32 %1, %1, 0;\n\t"
        "}"
        : "=&r"(t0), "=&r"(t1)
        : "r"((u32)hi), "r"((u32)lo), "r"((u32)(lo >> 32)), "r"((u32)(hi >> 32)));
    return ((u64)t1 << 32) | t0;
#else
    u64 hh = hi >> 32, hl = hi & GL_EPS;
    u64 t0 = lo - hh;
    if (lo < hh) t0 -= GL_EPS;
    u64 t1 = hl * GL_EPS;
    return gl_add(t0, t1);
#endif
}

// lazy x lazy -> lazy
GL_HD u64 gl_mul_lazy(u64 a, u64 b) {
    u64 lo, hi;
    gl_mul_wide(a, b, lo, hi);
    return gl_reduce128(lo, hi);
}
// lazy x lazy -> canonical
GL_HD u64 gl_mul(u64 a, u64 b) { return gl_canon(gl_mul_lazy(a, b)); }

// lazy o + lazy x * lazy a -> lazy: the addend rides in the 128-bit product (x*a + o < 2^128 always), so a
// multiply-accumulate costs one reduction and needs no canonical operand
GL_HD u64 gl_mac(u64 o, u64 x, u64 a) {
    u64 lo, hi;
    gl_mul_wide(x, a, lo, hi);
    lo += o;
    hi += (lo < o) ? 1 : 0;
    return gl_reduce128(lo, hi);
}

// ---- butterflies -----------------------------------------------------------------------------
// Cooley-Tukey (forward, crypto/poly.rs:17-22): (u, v) -> (u + v*w, u - v*w).  u,v lazy in, lazy out.
GL_HD void gl_ct_butterfly(u64 &u, u64 &v, u64 w) {
    u64 t = gl_mul(v, w);
    v = gl_sub(u, t);
    u = gl_add(u, t);
}
// Gentleman-Sande (inverse, crypto/poly.rs:45-50): (u, v) -> (u + v, (u - v)*w).  lazy in, lazy out.
GL_HD void gl_gs_butterfly(u64 &u, u64 &v, u64 w) {
    u64 vc = gl_canon(v);
    u64 d = gl_sub(u, vc);
    u = gl_add(u, vc);
    v = gl_mul_lazy(d, w);
}

// Inverse butterfly against the FORWARD table: INVROOTS[m+i] = -ROOTS[2m-1-i] (see tile_stage), so
// (u, v) -> (u + v, (u - v) * w_inv) = (u + v, (v - u) * w) with w = ROOTS[2m-1-i].  lazy in, lazy out.
GL_HD void gl_gs_butterfly_neg(u64 &u, u64 &v, u64 w) {
    u64 uc = gl_canon(u);
    u64 d = gl_sub(v, uc);
    u = gl_add(v, uc);
    v = gl_mul_lazy(d, w);
}

// ---- optimistic butterflies -----------------------------------------------------------------------------------
// gl_canon is 5 of the ~22 ALU instructions of a butterfly, and it changes its operand only when the operand is in
// [q, 2^64), i.e. when its high word is 0xFFFFFFFF (probability 2^-32 on transform data).  The optimistic forms skip
// it and instead record "the high word was all ones" in `flag` (one ISETP with an OR-ed predicate).  When the flag stays
// clear the skipped canonicalisation was the identity, so the result is the same word for word; when it is set the
// caller recomputes the pass with the exact forms (tfhe_kernels.cuh pass_r_slow).
GL_HD void gl_ct_butterfly_opt(u64 &u, u64 &v, u64 w, bool &flag) {
    const u64 t = gl_mul_lazy(v, w);
    flag |= ((u32)(t >> 32) == 0xFFFFFFFFu);
    v = gl_sub(u, t);
    u = gl_add(u, t);
}
GL_HD void gl_gs_butterfly_neg_opt(u64 &u, u64 &v, u64 w, bool &flag) {
    flag |= ((u32)(u >> 32) == 0xFFFFFFFFu);
    const u64 d = gl_sub(v, u);
    u = gl_add(v, u);
    v = gl_mul_lazy(d, w);
}

// ---- power-of-two twiddles -------------------------------------------------------------------------
// 2 has order 192 in Z_q (2^96 = -1), so every 32nd root of unity is a power of 2^6.  The reference's first sixteen
// table entries ROOTS[j] = psi^bitrev(j) (ntt/gen_param_file.sage:86-88) are the 32nd roots omega^bitrev4(j) with
// omega = 7^((q-1)/32) for EVERY ring size, i.e. the twiddles of NTT stages 0..3 are these constants:
//     ROOTS[j] = 2^gl_root_log2(j),  j < 16        (checked against the generated table in lux_engine_create)
// and a butterfly of those stages multiplies by a shift instead of four IMAD.WIDE.
GL_HD constexpr int gl_root_log2(int j) {
    constexpr int t[16] = {0, 48, 120, 168, 156, 12, 84, 132, 78, 126, 6, 54, 42, 90, 162, 18};
    return t[j & 15];
}

// x * 2^S mod q for a compile-time S in [0, 96).  lazy -> lazy.
// With S = 32a + r and y = x << r = y0 + y1 W + y2 W^2 (W = 2^32; W^2 = W - 1, W^3 = -1, W^4 = -W mod q):
//   a = 0:  (y1:y0) + y2 * eps            a = 1:  (y0:0) + y1 * eps - y2            a = 2:  y0 * eps - (y2:y1)
template <int S>
GL_HD u64 gl_mul_pow2(u64 x) {
    static_assert(S >= 0 && S < 96, "exponent must be reduced mod 96 (2^96 = -1)");
    if (S == 0) return x;
#if GL_PTX
    constexpr int r = S & 31, a = S >> 5;
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32);
    u32 y0, y1, y2;
    if (r == 0) {
        y0 = x0;
        y1 = x1;
        y2 = 0;
    } else {
        y0 = x0 << r;
        y1 = __funnelshift_l(x0, x1, r);
        y2 = x1 >> ((32 - r) & 31);
    }
    if (a == 1) return gl_reduce128((u64)y0 << 32, ((u64)y2 << 32) | y1);
    u32 t0, t1;
    if (a == 0) {
        // carry -> += eps (cannot carry again, as in gl_reduce128)
        asm("{\n\t"
            ".reg .u32 m;\n\t"
            "mad.lo.cc.u32 %0, %4, 0xFFFFFFFF, %2;\n\t"
            "madc.hi.cc.u32 %1, %4, 0xFFFFFFFF, %3;\n\t"
            "addc.u32 m, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, m;\n\t"
            "subc.u32 %1, %1, 0;\n\t"
            "add.u32 %1, %1, m;\n\t"
            "}"
            : "=&r"(t0), "=&r"(t1)
            : "r"(y0), "r"(y1), "r"(y2));
    } else {
        // y0 * eps = (y0:0) - y0 < q; then minus (y2:y1), borrow -> -= eps (cannot borrow again, as in gl_sub)
        asm("{\n\t"
            ".reg .u32 m;\n\t"
            "sub.cc.u32 %0, 0, %2;\n\t"
            "subc.u32 %1, %2, 0;\n\t"
            "sub.cc.u32 %0, %0, %3;\n\t"
            "subc.cc.u32 %1, %1, %4;\n\t"
            "subc.u32 m, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, m;\n\t"
            "subc.u32 %1, %1, 0;\n\t"
            "}"
            : "=&r"(t0), "=&r"(t1)
            : "r"(y0), "r"(y1), "r"(y2));
    }
    return ((u64)t1 << 32) | t0;
#else
    return gl_mul_lazy(x, S < 64 ? (1ULL << (S & 63)) : (GL_EPS << ((S - 64) & 31)));
#endif
}

// the butterflies of tile_stage with w = ROOTS[j] = 2^E, E = gl_root_log2(j) in [0, 192); E >= 96 means w = -2^(E-96)
template <int E>
GL_HD void gl_ct_butterfly_pow2(u64 &u, u64 &v) {
    const u64 t = gl_canon(gl_mul_pow2<E % 96>(v));
    const u64 lo = gl_sub(u, t), hi = gl_add(u, t);
    u = (E >= 96) ? lo : hi;
    v = (E >= 96) ? hi : lo;
}
// (u, v) -> (u + v, (v - u) * w)
template <int E>
GL_HD void gl_gs_butterfly_neg_pow2(u64 &u, u64 &v) {
    u64 d;
    if (E >= 96) {  // (v - u) * -2^(E-96) = (u - v) * 2^(E-96)
        const u64 vc = gl_canon(v);
        d = gl_sub(u, vc);
        u = gl_add(u, vc);
    } else {
        const u64 uc = gl_canon(u);
        d = gl_sub(v, uc);
        u = gl_add(v, uc);
    }
    v = gl_mul_pow2<E % 96>(d);
}

// ---- the reference's digit rule (vtfhe/glwe_poly.rs:28-50) ------------------------------------
// digit `li` (0 = least significant) of canonical x for base 2^logB with L = ceil(64/logB) limbs,
// returned as a canonical field element.  digit_i = k_i + MSB(k_{i-1}) - MSB(k_i)*B, where k_i are
// the raw logB-bit fields of x' = (sgn ? -x : x) and sgn = bit 63 of x iff L*logB == 64.
GL_HD u64 gl_digit(u64 x, u32 li, u32 logB, bool centred /* L*logB == 64 */) {
    bool sgn = centred && (x >> 63);
    u64 xc = sgn ? GL_Q - x : x;
    u32 sh = li * logB;
    u64 f = (xc >> sh) & ((1ULL << logB) - 1);
    u64 msb = (f >> (logB - 1)) & 1ULL;
    u64 prev = li ? ((xc >> (sh - 1)) & 1ULL) : 0ULL;
    int64_t d = (int64_t)f - (int64_t)(msb << logB) + (int64_t)prev;
    if (sgn) d = -d;
    return d < 0 ? GL_Q + (u64)d : (u64)d;
}

// lwe.rs:28-34 / vtfhe/mod.rs:80-107: round the canonical u64 to [0, 2N]
GL_HD u32 gl_mod_switch(u64 x, u32 logN) {
    u64 s = x >> (64 - logN - 2);
    return (u32)((s >> 1) + (s & 1));
}
