"""ctypes binding of the CPU oracle (oracle/oracle.c).

TEST INFRASTRUCTURE ONLY.  The product (tfhe_b200/) never imports this package; only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do,
as the checker.  See oracle/oracle.h for the reference lines each function restates.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from dataclasses import dataclass
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "liboracle.so"
Q = 0xFFFFFFFF00000001

ORDER_PBS_GLWEKS = 0
ORDER_LWEKS_PBS = 1


def build(force: bool = False) -> Path:
    """Compile oracle.c with the committed Makefile (gcc, no CUDA)."""
    src_mtime = max((_HERE / f).stat().st_mtime for f in ("oracle.c", "oracle.h", "torus.c", "torus.h", "Makefile"))
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src_mtime:
        subprocess.run(["make", "-C", str(_HERE), "-B" if force else "-s"], check=True, capture_output=True)
    return _LIB_PATH


class OrcParams(C.Structure):
    _fields_ = [(f, C.c_uint32) for f in ("n", "N", "k", "ell", "logB", "ell_ks", "logB_ks", "order")]


class OrcRng(C.Structure):
    _fields_ = [("s", C.c_uint64 * 4), ("have_spare", C.c_int), ("spare", C.c_double)]


_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    try:
        build()
    except Exception:
        if not _LIB_PATH.exists():
            raise
    L = C.CDLL(str(_LIB_PATH))
    u64, u32, vp, dbl = C.c_uint64, C.c_uint32, C.c_void_p, C.c_double
    sig = {
        "orc_add": (u64, [u64, u64]), "orc_sub": (u64, [u64, u64]), "orc_mul": (u64, [u64, u64]),
        "orc_mul_slow": (u64, [u64, u64]), "orc_neg": (u64, [u64]), "orc_pow": (u64, [u64, u64]),
        "orc_inv": (u64, [u64]),
        "orc_ntt_tables": (C.c_int, [u32, _u64p, _u64p, C.POINTER(u64)]),
        "orc_ntt_forward": (None, [u32, _u64p, _u64p]),
        "orc_ntt_backward": (None, [u32, _u64p, u64, _u64p]),
        "orc_num_limbs": (u32, [u32]),
        "orc_decompose": (None, [u64, u32, _u64p]),
        "orc_decompose_signed": (None, [u64, u32, _i64p]),
        "orc_mod_switch": (u32, [u64, u32]),
        "orc_rotate": (None, [u32, _u64p, u32, _u64p]),
        "orc_delta": (u64, [u32]),
        "orc_testv": (None, [u32, u32, u64, _u64p]),
        "orc_testv_lut": (None, [u32, u32, u64, _u64p, _u64p]),
        "orc_sample_extract": (None, [u32, u32, _u64p, _u64p]),
        "orc_partial_sample_extract": (None, [u32, u32, u32, _u64p, _u64p]),
        "orc_ctx_create": (vp, [C.POINTER(OrcParams)]),
        "orc_ctx_destroy": (None, [vp]),
        "orc_ctx_ninv": (u64, [vp]),
        "orc_ctx_roots": (C.POINTER(u64), [vp]),
        "orc_ctx_invroots": (C.POINTER(u64), [vp]),
        "orc_ctx_bsk_ntt": (C.POINTER(u64), [vp]),
        "orc_ctx_set_bsk": (C.c_int, [vp, _u64p, C.c_int]),
        "orc_ctx_set_glwe_ksk": (C.c_int, [vp, _u64p, C.c_int]),
        "orc_ctx_set_lwe_ksk": (C.c_int, [vp, _u64p]),
        "orc_external_product": (None, [vp, _u64p, _u64p, _u64p]),
        "orc_cmux_step": (None, [vp, _u64p, u64, _u64p]),
        "orc_blind_rotate": (None, [vp, _u64p, _u64p, _u64p]),
        "orc_blind_rotate_steps": (None, [vp, _u64p, _u64p, u32, _u64p]),
        "orc_lwe_keyswitch": (None, [vp, _u64p, _u64p]),
        "orc_pbs": (C.c_int, [vp, _u64p, _u64p, _u64p]),
        "orc_pbs_batch": (C.c_int, [vp, C.c_size_t, _u64p, _u64p, vp, _u64p]),
        "orc_lwe_in_len": (C.c_size_t, [vp]),
        "orc_max_threads": (C.c_int, []),
        "orc_rng_seed": (None, [C.POINTER(OrcRng), u64]),
        "orc_rng_u64": (u64, [C.POINTER(OrcRng)]),
        "orc_rng_field": (u64, [C.POINTER(OrcRng)]),
        "orc_rng_error": (u64, [C.POINTER(OrcRng), dbl]),
        "orc_lwe_keygen": (None, [C.POINTER(OrcRng), u32, _u64p]),
        "orc_glwe_keygen": (None, [C.POINTER(OrcRng), u32, u32, _u64p]),
        "orc_glwe_partial_key": (None, [C.POINTER(OrcRng), u32, u32, u32, _u64p]),
        "orc_lwe_encrypt": (None, [C.POINTER(OrcRng), u32, _u64p, u64, dbl, _u64p]),
        "orc_lwe_decrypt": (u64, [u32, _u64p, _u64p]),
        "orc_decode": (u64, [u64, u64, u32]),
        "orc_lwe_encrypt_batch": (None, [u64, u32, _u64p, C.c_size_t, _u64p, dbl, _u64p]),
        "orc_lwe_decrypt_batch": (None, [u32, _u64p, C.c_size_t, _u64p, _u64p]),
        "orc_glwe_encrypt": (None, [vp, C.POINTER(OrcRng), _u64p, _u64p, dbl, _u64p]),
        "orc_glwe_decrypt": (None, [vp, _u64p, _u64p, _u64p]),
        "orc_ggsw_encrypt": (None, [vp, C.POINTER(OrcRng), _u64p, _u64p, dbl, _u64p]),
        "orc_bsk_gen": (None, [vp, C.POINTER(OrcRng), _u64p, _u64p, dbl, _u64p]),
        "orc_glwe_ksk_gen": (None, [vp, C.POINTER(OrcRng), _u64p, _u64p, dbl, _u64p]),
        "orc_lwe_ksk_gen": (None, [vp, C.POINTER(OrcRng), _u64p, _u64p, dbl, _u64p]),
        "orc_glwe_max_error": (dbl, [vp, _u64p, _u64p, _u64p]),
        "orc_glwe_avg_error": (dbl, [vp, _u64p, _u64p, _u64p]),
        "orc_centered_diff": (dbl, [u64, u64]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def u64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint64)


def ntt_tables(N: int):
    roots = np.zeros(N, np.uint64)
    inv = np.zeros(N, np.uint64)
    ninv = C.c_uint64(0)
    rc = lib().orc_ntt_tables(N, roots, inv, C.byref(ninv))
    if rc:
        raise ValueError(f"orc_ntt_tables({N}) -> {rc}")
    return roots, inv, int(ninv.value)


def ntt_forward(a, roots=None):
    a = u64(a).copy()
    if roots is None:
        roots = ntt_tables(len(a))[0]
    lib().orc_ntt_forward(len(a), roots, a)
    return a


def ntt_backward(a, invroots=None, ninv=None):
    a = u64(a).copy()
    if invroots is None:
        _, invroots, ninv = ntt_tables(len(a))
    lib().orc_ntt_backward(len(a), invroots, ninv, a)
    return a


def decompose(x: int, logB: int) -> np.ndarray:
    out = np.zeros(lib().orc_num_limbs(logB), np.uint64)
    lib().orc_decompose(x, logB, out)
    return out


def decompose_signed(x: int, logB: int) -> np.ndarray:
    out = np.zeros(lib().orc_num_limbs(logB), np.int64)
    lib().orc_decompose_signed(x, logB, out)
    return out


def rotate(poly, shift: int) -> np.ndarray:
    poly = u64(poly)
    out = np.zeros_like(poly)
    lib().orc_rotate(len(poly), poly, shift, out)
    return out


def testv_lut(N: int, p: int, delta: int, lut) -> np.ndarray:
    out = np.zeros(N, np.uint64)
    lib().orc_testv_lut(N, p, delta, u64(lut), out)
    return out


class Rng:
    def __init__(self, seed: int):
        self.st = OrcRng()
        lib().orc_rng_seed(C.byref(self.st), seed)

    @property
    def ref(self):
        return C.byref(self.st)

    def u64(self) -> int:
        return lib().orc_rng_u64(self.ref)

    def field(self, count: int | None = None):
        if count is None:
            return lib().orc_rng_field(self.ref)
        return np.array([lib().orc_rng_field(self.ref) for _ in range(count)], dtype=np.uint64)


@dataclass
class Params:
    n: int
    N: int
    k: int = 1
    ell: int = 1
    logB: int = 8
    ell_ks: int = 0
    logB_ks: int = 0
    order: int = ORDER_PBS_GLWEKS

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
        """words per ciphertext entering/leaving one PBS"""
        return self.k * self.N + 1 if self.order == ORDER_LWEKS_PBS else self.n + 1


class Ctx:
    """Parameters + NTT tables + evaluation keys; thin OO wrapper over orc_ctx."""

    def __init__(self, params: Params):
        self.p = params
        cp = params.c()
        self.h = lib().orc_ctx_create(C.byref(cp))
        if not self.h:
            raise ValueError(f"bad oracle parameters {params}")

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_ctx_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # ---- tables
    @property
    def roots(self):
        return np.ctypeslib.as_array(lib().orc_ctx_roots(self.h), (self.p.N,)).copy()

    @property
    def invroots(self):
        return np.ctypeslib.as_array(lib().orc_ctx_invroots(self.h), (self.p.N,)).copy()

    @property
    def ninv(self):
        return lib().orc_ctx_ninv(self.h)

    # ---- keys
    def set_bsk(self, bsk, ntt_domain=False):
        assert u64(bsk).size == self.p.n * self.p.ggsw_len
        assert lib().orc_ctx_set_bsk(self.h, u64(bsk).reshape(-1), int(ntt_domain)) == 0

    def set_glwe_ksk(self, ksk, ntt_domain=False):
        assert u64(ksk).size == self.p.ggsw_len
        assert lib().orc_ctx_set_glwe_ksk(self.h, u64(ksk).reshape(-1), int(ntt_domain)) == 0

    def set_lwe_ksk(self, ksk):
        assert u64(ksk).size == self.p.lwe_ksk_len
        assert lib().orc_ctx_set_lwe_ksk(self.h, u64(ksk).reshape(-1)) == 0

    def bsk_ntt(self):
        return np.ctypeslib.as_array(lib().orc_ctx_bsk_ntt(self.h), (self.p.n * self.p.ggsw_len,)).copy()

    # ---- hot path
    def external_product(self, ggsw_ntt, glwe):
        out = np.zeros(self.p.glwe_len, np.uint64)
        lib().orc_external_product(self.h, u64(ggsw_ntt).reshape(-1), u64(glwe).reshape(-1), out)
        return out

    def cmux_step(self, ggsw_ntt, mask_element: int, acc):
        acc = u64(acc).reshape(-1).copy()
        lib().orc_cmux_step(self.h, u64(ggsw_ntt).reshape(-1), mask_element, acc)
        return acc

    def blind_rotate(self, lwe_small, testv, steps: int | None = None):
        out = np.zeros(self.p.glwe_len, np.uint64)
        if steps is None:
            lib().orc_blind_rotate(self.h, u64(lwe_small), u64(testv), out)
        else:
            lib().orc_blind_rotate_steps(self.h, u64(lwe_small), u64(testv), steps, out)
        return out

    def lwe_keyswitch(self, lwe_big):
        out = np.zeros(self.p.n + 1, np.uint64)
        lib().orc_lwe_keyswitch(self.h, u64(lwe_big), out)
        return out

    def pbs(self, lwe_in, testv):
        out = np.zeros(self.p.ct_len, np.uint64)
        rc = lib().orc_pbs(self.h, u64(lwe_in), u64(testv), out)
        if rc:
            raise RuntimeError(f"orc_pbs -> {rc} (keys not set?)")
        return out

    def pbs_batch(self, lwe_in, testvs, lut_idx=None):
        lwe_in = u64(lwe_in).reshape(-1, self.p.ct_len)
        testvs = u64(testvs).reshape(-1, self.p.N)
        out = np.zeros_like(lwe_in)
        idx = None
        if lut_idx is not None:
            idx_arr = np.ascontiguousarray(lut_idx, dtype=np.uint32)
            idx = idx_arr.ctypes.data_as(C.c_void_p)
        rc = lib().orc_pbs_batch(self.h, lwe_in.shape[0], lwe_in.reshape(-1), testvs.reshape(-1), idx, out.reshape(-1))
        if rc < 0:
            raise RuntimeError("orc_pbs_batch failed (keys not set?)")
        return out, rc

    # ---- fixtures
    def glwe_encrypt(self, rng: Rng, s, m, sigma=0.0):
        ct = np.zeros(self.p.glwe_len, np.uint64)
        lib().orc_glwe_encrypt(self.h, rng.ref, u64(s).reshape(-1), u64(m), sigma, ct)
        return ct

    def glwe_decrypt(self, s, ct):
        m = np.zeros(self.p.N, np.uint64)
        lib().orc_glwe_decrypt(self.h, u64(s).reshape(-1), u64(ct).reshape(-1), m)
        return m

    def ggsw_encrypt(self, rng: Rng, s, m, sigma=0.0):
        out = np.zeros(self.p.ggsw_len, np.uint64)
        lib().orc_ggsw_encrypt(self.h, rng.ref, u64(s).reshape(-1), u64(m), sigma, out)
        return out

    def bsk_gen(self, rng: Rng, s_lwe, s_glwe, sigma=0.0):
        out = np.zeros(self.p.n * self.p.ggsw_len, np.uint64)
        lib().orc_bsk_gen(self.h, rng.ref, u64(s_lwe), u64(s_glwe).reshape(-1), sigma, out)
        return out

    def glwe_ksk_gen(self, rng: Rng, s_to, s_from, sigma=0.0):
        out = np.zeros(self.p.ggsw_len, np.uint64)
        lib().orc_glwe_ksk_gen(self.h, rng.ref, u64(s_to).reshape(-1), u64(s_from).reshape(-1), sigma, out)
        return out

    def lwe_ksk_gen(self, rng: Rng, s_big, s_small, sigma=0.0):
        out = np.zeros(self.p.lwe_ksk_len, np.uint64)
        lib().orc_lwe_ksk_gen(self.h, rng.ref, u64(s_big).reshape(-1), u64(s_small), sigma, out)
        return out


def lwe_keygen(rng: Rng, n: int):
    s = np.zeros(n, np.uint64)
    lib().orc_lwe_keygen(rng.ref, n, s)
    return s


def glwe_keygen(rng: Rng, N: int, k: int):
    s = np.zeros(k * N, np.uint64)
    lib().orc_glwe_keygen(rng.ref, N, k, s)
    return s


def glwe_partial_key(rng: Rng, N: int, k: int, nz: int):
    s = np.zeros(k * N, np.uint64)
    lib().orc_glwe_partial_key(rng.ref, N, k, nz, s)
    return s


def lwe_encrypt(rng: Rng, s, m: int, sigma=0.0):
    s = u64(s)
    ct = np.zeros(len(s) + 1, np.uint64)
    lib().orc_lwe_encrypt(rng.ref, len(s), s, m, sigma, ct)
    return ct


def lwe_decrypt(s, ct) -> int:
    s = u64(s)
    return lib().orc_lwe_decrypt(len(s), s, u64(ct))


def lwe_encrypt_batch(seed: int, s, msgs, sigma=0.0) -> np.ndarray:
    """[B][n+1] encryptions of the field elements `msgs` (already scaled by delta), OpenMP over ciphertexts"""
    s, msgs = u64(s), u64(msgs)
    cts = np.zeros((len(msgs), len(s) + 1), np.uint64)
    lib().orc_lwe_encrypt_batch(seed, len(s), s, len(msgs), msgs, sigma, cts.reshape(-1))
    return cts


def lwe_decrypt_batch(s, cts) -> np.ndarray:
    """noisy plaintexts (field elements) of [B][n+1] ciphertexts"""
    s, cts = u64(s), u64(cts)
    out = np.zeros(cts.shape[0], np.uint64)
    lib().orc_lwe_decrypt_batch(len(s), s, cts.shape[0], cts.reshape(-1), out)
    return out


def decode(noisy: int, delta: int, space: int) -> int:
    return lib().orc_decode(noisy, delta, space)


def delta(space: int) -> int:
    return lib().orc_delta(space)


# ---- boolean gates (BASELINE config 1).  The reference's gate layer (Go luxfi/tfhe) is out of tree; this is
# the textbook TFHE gate bootstrap over the in-tree encoding (bit m -> m * q/4, main.rs:48-51):
#   gate(a, b) = sign-PBS(coef * (a + b) + konst) + q/8, constant test vector q/8.
GATE_COEFFS = {"nand": (-1, 3), "and": (1, -3), "or": (1, -1), "nor": (-1, 1), "xor": (2, -2), "xnor": (-2, 2)}
GATE_TRUTH = {"nand": lambda x, y: 1 - (x & y), "and": lambda x, y: x & y, "or": lambda x, y: x | y,
              "nor": lambda x, y: 1 - (x | y), "xor": lambda x, y: x ^ y, "xnor": lambda x, y: 1 - (x ^ y)}


def gate_batch(ctx: "Ctx", name: str, a, b) -> np.ndarray:
    """[B][ct_len] x [B][ct_len] -> [B][ct_len], the oracle's PBS on the gate's linear combination"""
    coef, kh = GATE_COEFFS[name]
    h = Q >> 3
    a, b = u64(a), u64(b)
    x = ((a.astype(object) + b.astype(object)) * coef) % Q
    x[:, -1] = (x[:, -1] + kh * h) % Q
    tv = np.full((1, ctx.p.N), h, dtype=np.uint64)
    out, _ = ctx.pbs_batch(x.astype(np.uint64), tv, np.zeros(len(x), np.uint32))
    out = out.astype(object)
    out[:, -1] = (out[:, -1] + h) % Q
    return out.astype(np.uint64)


@dataclass
class KeySet:
    """A full seeded fixture: secret keys (test side only) + evaluation keys in the standard domain."""
    params: Params
    s_lwe: np.ndarray      # [n]      small key
    s_glwe: np.ndarray     # [k*N]    GLWE key
    bsk: np.ndarray        # [n] GGSW, standard domain
    glwe_ksk: np.ndarray | None   # order 0
    lwe_ksk: np.ndarray | None    # order 1
    s_to: np.ndarray | None       # order 0: partial GLWE key whose first n coeffs are s_lwe

    @property
    def s_in(self):
        """key the PBS input/output ciphertexts are under"""
        return self.s_glwe if self.params.order == ORDER_LWEKS_PBS else self.s_lwe


def keygen(ctx: Ctx, seed: int, sigma_glwe=0.0, sigma_lwe=0.0) -> KeySet:
    """Fixture keys following main.rs:40-46 (order 0) or the KS->PBS layout (order 1)."""
    p = ctx.p
    rng = Rng(seed)
    if p.order == ORDER_PBS_GLWEKS:
        s_to = glwe_partial_key(rng, p.N, p.k, p.n)          # main.rs:40
        s_lwe = s_to[: p.n].copy()                            # main.rs:41 flatten_partial_key
        s_glwe = glwe_keygen(rng, p.N, p.k)                   # main.rs:44
        bsk = ctx.bsk_gen(rng, s_lwe, s_glwe, sigma_glwe)     # main.rs:45
        ksk = ctx.glwe_ksk_gen(rng, s_to, s_glwe, sigma_lwe)  # main.rs:46
        return KeySet(p, s_lwe, s_glwe, bsk, ksk, None, s_to)
    s_lwe = lwe_keygen(rng, p.n)
    s_glwe = glwe_keygen(rng, p.N, p.k)
    bsk = ctx.bsk_gen(rng, s_lwe, s_glwe, sigma_glwe)
    ksk = ctx.lwe_ksk_gen(rng, s_glwe, s_lwe, sigma_lwe)
    return KeySet(p, s_lwe, s_glwe, bsk, None, ksk, None)


def install_keys(ctx: Ctx, ks: KeySet):
    ctx.set_bsk(ks.bsk)
    if ks.glwe_ksk is not None:
        ctx.set_glwe_ksk(ks.glwe_ksk)
    if ks.lwe_ksk is not None:
        ctx.set_lwe_ksk(ks.lwe_ksk)
