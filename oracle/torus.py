"""ctypes binding of oracle/torus.c: the CPU restatement of the 2^64-torus / f64-FFT bootstrap.

TEST INFRASTRUCTURE ONLY (same rules as oracle/__init__.py).  The API mirrors oracle.Ctx / KeySet / keygen so
that the parity tests of the two arithmetics read alike.  Parity status: see oracle/torus.h.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import OrcParams, OrcRng, Rng, lib as _base_lib, u64

ORDER_LWEKS_PBS = 1
ORDER_PBS_LWEKS = 2
NO_NOISE = -1

_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_c128p = np.ctypeslib.ndpointer(dtype=np.complex128, flags="C_CONTIGUOUS")
_ready = False


def lib():
    global _ready
    L = _base_lib()
    if _ready:
        return L
    u64_, u32, vp, dbl, i32 = C.c_uint64, C.c_uint32, C.c_void_p, C.c_double, C.c_int
    rng = C.POINTER(OrcRng)
    sig = {
        "tor_decompose": (None, [u64_, u32, u32, _i64p]),
        "tor_closest_representable": (u64_, [u64_, u32, u32]),
        "tor_mod_switch": (u32, [u64_, u32]),
        "tor_rotate": (None, [u32, _u64p, u32, _u64p]),
        "tor_delta": (u64_, [u32]),
        "tor_testv_lut": (None, [u32, u32, u64_, _u64p, _u64p]),
        "tor_twiddles": (None, [u32, _c128p]),
        "tor_fft_forward": (None, [u32, _c128p, _c128p]),
        "tor_fft_inverse": (None, [u32, _c128p, _c128p]),
        "tor_from_double": (u64_, [dbl]),
        "tor_ctx_create": (vp, [C.POINTER(OrcParams)]),
        "tor_ctx_destroy": (None, [vp]),
        "tor_ctx_set_bsk": (i32, [vp, _u64p]),
        "tor_ctx_set_lwe_ksk": (i32, [vp, _u64p]),
        "tor_ctx_bsk_fourier": (C.POINTER(dbl), [vp]),
        "tor_ggsw_to_fourier": (None, [vp, _u64p, _c128p]),
        "tor_external_product": (None, [vp, _c128p, _u64p, _u64p]),
        "tor_cmux_step": (None, [vp, _c128p, u64_, _u64p]),
        "tor_blind_rotate_steps": (None, [vp, _u64p, _u64p, u32, _u64p]),
        "tor_lwe_keyswitch": (None, [vp, _u64p, _u64p]),
        "tor_sample_extract": (None, [u32, _u64p, _u64p]),
        "tor_pbs": (i32, [vp, _u64p, _u64p, _u64p]),
        "tor_pbs_batch": (i32, [vp, C.c_size_t, _u64p, _u64p, vp, _u64p]),
        "tor_ct_len": (C.c_size_t, [vp]),
        "tor_rng_tuniform": (C.c_int64, [rng, u32]),
        "tor_lwe_encrypt": (None, [rng, u32, _u64p, u64_, i32, _u64p]),
        "tor_lwe_decrypt": (u64_, [u32, _u64p, _u64p]),
        "tor_lwe_encrypt_batch": (None, [u64_, u32, _u64p, C.c_size_t, _u64p, i32, _u64p]),
        "tor_lwe_decrypt_batch": (None, [u32, _u64p, C.c_size_t, _u64p, _u64p]),
        "tor_decode": (u64_, [u64_, u64_, u32]),
        "tor_glwe_decrypt": (None, [u32, _u64p, _u64p, _u64p]),
        "tor_ggsw_encrypt": (None, [vp, rng, _u64p, u64_, i32, _u64p]),
        "tor_bsk_gen": (None, [vp, rng, _u64p, _u64p, i32, _u64p]),
        "tor_lwe_ksk_gen": (None, [vp, rng, _u64p, _u64p, i32, _u64p]),
        "tor_centered_diff": (dbl, [u64_, u64_]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _ready = True
    return L


@dataclass
class Params:
    """(ell, logB) and (ell_ks, logB_ks) are tfhe-rs's (level, base_log)."""
    n: int
    N: int
    k: int = 1
    ell: int = 1
    logB: int = 23
    ell_ks: int = 4
    logB_ks: int = 4
    order: int = ORDER_LWEKS_PBS

    def c(self) -> OrcParams:
        return OrcParams(self.n, self.N, self.k, self.ell, self.logB, self.ell_ks, self.logB_ks, self.order)

    @property
    def ggsw_len(self) -> int:
        return (self.k + 1) * self.ell * (self.k + 1) * self.N

    @property
    def glwe_len(self) -> int:
        return (self.k + 1) * self.N

    @property
    def lwe_ksk_len(self) -> int:
        return self.k * self.N * self.ell_ks * (self.n + 1)

    @property
    def ct_len(self) -> int:
        return self.k * self.N + 1 if self.order == ORDER_LWEKS_PBS else self.n + 1


def decompose(x: int, logB: int, ell: int) -> np.ndarray:
    out = np.zeros(ell, np.int64)
    lib().tor_decompose(x, logB, ell, out)
    return out


def closest_representable(x: int, logB: int, ell: int) -> int:
    return lib().tor_closest_representable(x, logB, ell)


def mod_switch(x: int, N: int) -> int:
    return lib().tor_mod_switch(x, N)


def rotate(poly, shift: int) -> np.ndarray:
    poly = u64(poly)
    out = np.zeros_like(poly)
    lib().tor_rotate(len(poly), poly, shift, out)
    return out


def delta(space: int) -> int:
    return lib().tor_delta(space)


def testv_lut(N: int, p: int, delta_: int, lut) -> np.ndarray:
    out = np.zeros(N, np.uint64)
    lib().tor_testv_lut(N, p, delta_, u64(lut), out)
    return out


def twiddles(N: int) -> np.ndarray:
    tw = np.zeros(N // 2, np.complex128)
    lib().tor_twiddles(N, tw)
    return tw


def fft_forward(N: int, z, tw=None) -> np.ndarray:
    z = np.ascontiguousarray(z, np.complex128).copy()
    lib().tor_fft_forward(N, twiddles(N) if tw is None else tw, z)
    return z


def fft_inverse(N: int, z, tw=None) -> np.ndarray:
    z = np.ascontiguousarray(z, np.complex128).copy()
    lib().tor_fft_inverse(N, twiddles(N) if tw is None else tw, z)
    return z


def from_double(y: float) -> int:
    return lib().tor_from_double(y)


def decode(noisy: int, delta_: int, space: int) -> int:
    return lib().tor_decode(noisy, delta_, space)


def centered_diff(a: int, b: int) -> float:
    return lib().tor_centered_diff(a, b)


def lwe_encrypt(rng: Rng, s, m: int, noise_log2: int = NO_NOISE) -> np.ndarray:
    s = u64(s)
    ct = np.zeros(len(s) + 1, np.uint64)
    lib().tor_lwe_encrypt(rng.ref, len(s), s, m, noise_log2, ct)
    return ct


def lwe_decrypt(s, ct) -> int:
    s = u64(s)
    return lib().tor_lwe_decrypt(len(s), s, u64(ct))


def lwe_encrypt_batch(seed: int, s, msgs, noise_log2: int = NO_NOISE) -> np.ndarray:
    s, msgs = u64(s), u64(msgs)
    cts = np.zeros((len(msgs), len(s) + 1), np.uint64)
    lib().tor_lwe_encrypt_batch(seed, len(s), s, len(msgs), msgs, noise_log2, cts.reshape(-1))
    return cts


def lwe_decrypt_batch(s, cts) -> np.ndarray:
    s, cts = u64(s), u64(cts)
    out = np.zeros(cts.shape[0], np.uint64)
    lib().tor_lwe_decrypt_batch(len(s), s, cts.shape[0], cts.reshape(-1), out)
    return out


def glwe_decrypt(N: int, s, ct) -> np.ndarray:
    m = np.zeros(N, np.uint64)
    lib().tor_glwe_decrypt(N, u64(s), u64(ct).reshape(-1), m)
    return m


class Ctx:
    def __init__(self, params: Params):
        self.p = params
        cp = params.c()
        self.h = lib().tor_ctx_create(C.byref(cp))
        if not self.h:
            raise ValueError(f"bad torus oracle parameters {params}")

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().tor_ctx_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def set_bsk(self, bsk):
        assert u64(bsk).size == self.p.n * self.p.ggsw_len
        assert lib().tor_ctx_set_bsk(self.h, u64(bsk).reshape(-1)) == 0

    def set_lwe_ksk(self, ksk):
        assert u64(ksk).size == self.p.lwe_ksk_len
        assert lib().tor_ctx_set_lwe_ksk(self.h, u64(ksk).reshape(-1)) == 0

    def bsk_fourier(self) -> np.ndarray:
        cnt = self.p.n * 4 * self.p.ell * (self.p.N // 2)
        return np.ctypeslib.as_array(lib().tor_ctx_bsk_fourier(self.h), (cnt * 2,)).view(np.complex128).copy()

    def ggsw_to_fourier(self, ggsw) -> np.ndarray:
        out = np.zeros(4 * self.p.ell * (self.p.N // 2), np.complex128)
        lib().tor_ggsw_to_fourier(self.h, u64(ggsw).reshape(-1), out)
        return out

    def external_product(self, ggsw_f, glwe) -> np.ndarray:
        out = np.zeros(self.p.glwe_len, np.uint64)
        lib().tor_external_product(self.h, np.ascontiguousarray(ggsw_f, np.complex128), u64(glwe).reshape(-1), out)
        return out

    def cmux_step(self, ggsw_f, mask_element: int, acc) -> np.ndarray:
        acc = u64(acc).reshape(-1).copy()
        lib().tor_cmux_step(self.h, np.ascontiguousarray(ggsw_f, np.complex128), mask_element, acc)
        return acc

    def blind_rotate(self, lwe_small, testv, steps: int | None = None) -> np.ndarray:
        out = np.zeros(self.p.glwe_len, np.uint64)
        lib().tor_blind_rotate_steps(self.h, u64(lwe_small), u64(testv), self.p.n if steps is None else steps, out)
        return out

    def lwe_keyswitch(self, lwe_big) -> np.ndarray:
        out = np.zeros(self.p.n + 1, np.uint64)
        lib().tor_lwe_keyswitch(self.h, u64(lwe_big), out)
        return out

    def pbs(self, lwe_in, testv) -> np.ndarray:
        out = np.zeros(self.p.ct_len, np.uint64)
        if lib().tor_pbs(self.h, u64(lwe_in), u64(testv), out):
            raise RuntimeError("tor_pbs failed (keys not set?)")
        return out

    def pbs_batch(self, lwe_in, testvs, lut_idx=None):
        lwe_in = u64(lwe_in).reshape(-1, self.p.ct_len)
        testvs = u64(testvs).reshape(-1, self.p.N)
        out = np.zeros_like(lwe_in)
        idx = None
        if lut_idx is not None:
            idx_arr = np.ascontiguousarray(lut_idx, dtype=np.uint32)
            idx = idx_arr.ctypes.data_as(C.c_void_p)
        rc = lib().tor_pbs_batch(self.h, lwe_in.shape[0], lwe_in.reshape(-1), testvs.reshape(-1), idx, out.reshape(-1))
        if rc < 0:
            raise RuntimeError("tor_pbs_batch failed (keys not set?)")
        return out, rc

    # ---- fixtures
    def ggsw_encrypt(self, rng: Rng, s_glwe, m: int, noise_log2: int = NO_NOISE) -> np.ndarray:
        out = np.zeros(self.p.ggsw_len, np.uint64)
        lib().tor_ggsw_encrypt(self.h, rng.ref, u64(s_glwe), m, noise_log2, out)
        return out

    def bsk_gen(self, rng: Rng, s_lwe, s_glwe, noise_log2: int = NO_NOISE) -> np.ndarray:
        out = np.zeros(self.p.n * self.p.ggsw_len, np.uint64)
        lib().tor_bsk_gen(self.h, rng.ref, u64(s_lwe), u64(s_glwe), noise_log2, out)
        return out

    def lwe_ksk_gen(self, rng: Rng, s_big, s_small, noise_log2: int = NO_NOISE) -> np.ndarray:
        out = np.zeros(self.p.lwe_ksk_len, np.uint64)
        lib().tor_lwe_ksk_gen(self.h, rng.ref, u64(s_big), u64(s_small), noise_log2, out)
        return out


@dataclass
class KeySet:
    params: Params
    s_lwe: np.ndarray
    s_glwe: np.ndarray
    bsk: np.ndarray
    lwe_ksk: np.ndarray

    @property
    def s_in(self):
        return self.s_glwe if self.params.order == ORDER_LWEKS_PBS else self.s_lwe


def keygen(ctx: Ctx, seed: int, glwe_noise_log2: int = NO_NOISE, lwe_noise_log2: int = NO_NOISE) -> KeySet:
    from . import glwe_keygen, lwe_keygen
    p = ctx.p
    rng = Rng(seed)
    s_lwe = lwe_keygen(rng, p.n)
    s_glwe = glwe_keygen(rng, p.N, p.k)
    bsk = ctx.bsk_gen(rng, s_lwe, s_glwe, glwe_noise_log2)
    ksk = ctx.lwe_ksk_gen(rng, s_glwe, s_lwe, lwe_noise_log2)
    return KeySet(p, s_lwe, s_glwe, bsk, ksk)


def install_keys(ctx: Ctx, ks: KeySet):
    ctx.set_bsk(ks.bsk)
    ctx.set_lwe_ksk(ks.lwe_ksk)


# PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 (named at sdk/relayer/src/tfhe.ts:15; values from tfhe-rs,
# not in the tree: SURVEY App. B): n = 918, N = 2048, k = 1, PBS base 2^23 x 1 level, KS base 2^4 x 4 levels,
# LWE noise TUniform(45), GLWE noise TUniform(17), message 2 bits + carry 2 bits + 1 padding bit.
S2 = Params(n=918, N=2048, k=1, ell=1, logB=23, ell_ks=4, logB_ks=4, order=ORDER_LWEKS_PBS)
S2_LWE_NOISE = 45
S2_GLWE_NOISE = 17
