"""ctypes binding of the C ABI in include/lux_tfhe.h (libluxtfhe_b200.so).

Host-side only: marshals numpy / torch buffers into the ABI.  All arithmetic happens in the CUDA
library; nothing here computes on ciphertexts and nothing falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from pathlib import Path

import numpy as np

Q = 0xFFFFFFFF00000001
ORDER_PBS_GLWEKS = 0
ORDER_LWEKS_PBS = 1
ORDER_PBS_LWEKS = 2
ARITH_GOLDILOCKS_NTT = 0
ARITH_TORUS64_FFT = 1
DOMAIN_STANDARD = 0
DOMAIN_NTT = 1

_HERE = Path(__file__).resolve().parent


def lib_path() -> Path:
    """The in-tree CUDA library.  LUX_TFHE_LIB selects another build of the same sources (tuning variants
    made by scripts/build_variant.sh); it is never a different implementation."""
    override = os.environ.get("LUX_TFHE_LIB")
    return Path(override) if override else _HERE / "libluxtfhe_b200.so"


class LuxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"lux error {code}: {msg}")
        self.code = code


class _CParams(C.Structure):
    _fields_ = [(f, C.c_uint32) for f in ("n", "N", "k", "ell", "logB", "ell_ks", "logB_ks", "order", "arith")]


@dataclass(frozen=True)
class Params:
    """lux_params.  (ell, logB) follow the reference's limb convention: L = ceil(64/logB) limbs of
    logB bits counted from bit 0, the external product keeps the top `ell` (glev_ct.rs:97-99)."""
    n: int
    N: int
    k: int = 1
    ell: int = 1
    logB: int = 8
    ell_ks: int = 0
    logB_ks: int = 0
    order: int = ORDER_PBS_GLWEKS
    arith: int = ARITH_GOLDILOCKS_NTT   # ARITH_TORUS64_FFT: (ell, logB) are tfhe-rs's (level, base_log)

    def c(self) -> _CParams:
        return _CParams(self.n, self.N, self.k, self.ell, self.logB, self.ell_ks, self.logB_ks, self.order, self.arith)

    @property
    def modulus(self) -> int:
        return 1 << 64 if self.arith == ARITH_TORUS64_FFT else Q

    @property
    def ggsw_len(self) -> int:
        return (self.k + 1) * self.ell * (self.k + 1) * self.N

    @property
    def ct_len(self) -> int:
        return self.k * self.N + 1 if self.order == ORDER_LWEKS_PBS else self.n + 1

    @property
    def lwe_ksk_len(self) -> int:
        return self.k * self.N * self.ell_ks * (self.n + 1)


GATES = {"nand": 0, "and": 1, "or": 2, "nor": 3, "xor": 4, "xnor": 5}   # lux_gate (include/lux_tfhe.h)

# S1: the reference's in-tree boolean set, research/verifiable-fhe/src/main.rs:23-30
S1_BOOLEAN = Params(n=728, N=1024, k=1, ell=4, logB=5, ell_ks=4, logB_ks=5, order=ORDER_PBS_GLWEKS)
# S2: shape of PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 (sdk/relayer/src/tfhe.ts:15) over the
# reference's Goldilocks modulus: n=918, N=2048, one 23-bit PBS digit at 2^41 (logB=41 in the limb
# convention => L=2, top limb = bits 41..63), key switch 4 levels of 4 bits (limbs 12..15 of 16).
S2_SHORTINT = Params(n=918, N=2048, k=1, ell=1, logB=41, ell_ks=4, logB_ks=4, order=ORDER_LWEKS_PBS)
# S2 on the arithmetic the named set is defined on (tfhe-rs: 2^64 torus, f64 FFT): n = 918, N = 2048, PBS base 2^23 x 1
# level, key switch base 2^4 x 4 levels, LWE noise TUniform(45), GLWE noise TUniform(17)   
S2_TORUS = Params(n=918, N=2048, k=1, ell=1, logB=23, ell_ks=4, logB_ks=4, order=ORDER_LWEKS_PBS, arith=ARITH_TORUS64_FFT)
TEST_SMALL_TORUS = Params(n=20, N=2048, k=1, ell=1, logB=23, ell_ks=4, logB_ks=4, order=ORDER_LWEKS_PBS, arith=ARITH_TORUS64_FFT)
TEST_SMALL_TORUS_O2 = Params(n=20, N=2048, k=1, ell=1, logB=23, ell_ks=4, logB_ks=4, order=ORDER_PBS_LWEKS, arith=ARITH_TORUS64_FFT)
# small sets for fast parity tests
TEST_SMALL_O0 = Params(n=24, N=512, k=1, ell=4, logB=8, ell_ks=4, logB_ks=8, order=ORDER_PBS_GLWEKS)
TEST_SMALL_O1 = Params(n=24, N=512, k=1, ell=2, logB=16, ell_ks=8, logB_ks=4, order=ORDER_LWEKS_PBS)

_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_lib = None

# every symbol include/lux_tfhe.h declares: name -> (restype, argtypes)
_vp, _sz, _u32, _u64, _int = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_int
ABI = {
    "lux_version": (C.c_char_p, []),
    "lux_engine_create": (_int, [C.POINTER(_CParams), _int, C.POINTER(_vp)]),
    "lux_engine_destroy": (None, [_vp]),
    "lux_last_error": (C.c_char_p, [_vp]),
    "lux_engine_params": (_int, [_vp, C.POINTER(_CParams)]),
    "lux_ct_words": (_sz, [_vp]),
    "lux_launch_count": (_u64, [_vp]),
    "lux_wave_ciphertexts": (_sz, [_vp, _sz]),
    "lux_set_timing": (_int, [_vp, _int]),
    "lux_last_timing": (_int, [_vp, C.POINTER(C.c_float * 2)]),
    "lux_synchronize": (_int, [_vp]),
    "lux_load_bsk": (_int, [_vp, _u64p, _sz, _int]),
    "lux_load_glwe_ksk": (_int, [_vp, _u64p, _sz, _int]),
    "lux_load_lwe_ksk": (_int, [_vp, _u64p, _sz]),
    "lux_key_blob_bytes": (_sz, [_vp]),
    "lux_export_keys_device": (_int, [_vp, _vp, _sz]),
    "lux_import_keys_device": (_int, [_vp, _vp, _sz]),
    "lux_pbs_batch": (_int, [_vp, _vp, _vp, _u32, _vp, _sz, _vp]),
    "lux_pbs_batch_device": (_int, [_vp, _vp, _vp, _u32, _vp, _sz, _vp, _vp]),
    "lux_pbs_batch_device_static": (_int, [_vp, _vp, _vp, _u32, _vp, _sz, _vp, _vp]),
    "lux_gate_batch": (_int, [_vp, _int, _vp, _vp, _sz, _vp]),
    "lux_gate_batch_device": (_int, [_vp, _int, _vp, _vp, _sz, _vp, _vp]),
    "lux_gate_not": (_int, [_vp, _vp, _sz, _vp]),
    "lux_make_testv": (_int, [_u32, _u32, _u64, _u64p, _u64p]),
    "lux_delta": (_u64, [_u32]),
    "lux_make_testv_arith": (_int, [_u32, _u32, _u32, _u64, _u64p, _u64p]),
    "lux_delta_arith": (_u64, [_u32, _u32]),
    "lux_fft": (_int, [_vp, _vp, _sz, _int, _vp]),
    "lux_torus_decompose": (_int, [_vp, _u64p, _sz, _vp]),
    "lux_export_bsk_fourier": (_int, [_vp, _vp, _sz]),
    "lux_ntt_forward": (_int, [_vp, _u64p, _sz, _u64p]),
    "lux_ntt_backward": (_int, [_vp, _u64p, _sz, _u64p]),
    "lux_ntt_any": (_int, [_vp, _u32, _int, _u64p, _sz, _u64p]),
    "lux_decompose": (_int, [_vp, _u64p, _sz, _u32, _u64p]),
    "lux_mod_switch": (_int, [_vp, _u64p, _sz, _u32p]),
    "lux_lwe_keyswitch": (_int, [_vp, _u64p, _sz, _u64p]),
    "lux_blind_rotate": (_int, [_vp, _u64p, _u64p, _u32, _vp, _sz, _u32, _u64p]),
    "lux_external_product": (_int, [_vp, _u64p, _int, _u64p, _sz, _u64p]),
    "lux_serialized_bytes": (_sz, [C.POINTER(_CParams), _u32, _u32, _u64]),
    "lux_serialize": (_int, [C.POINTER(_CParams), _u32, _u32, _u64, _vp, _vp, _sz]),
    "lux_blob_inspect": (_int, [_vp, _sz, _vp]),
    "lux_deserialize": (_int, [_vp, _sz, _vp, _u32, _vp, _sz, _vp]),
    "lux_load_key_blob": (_int, [_vp, _vp, _sz]),
    "lux_engine_create_multi": (_int, [C.POINTER(_CParams), _int, C.POINTER(_vp)]),
    "lux_multi_destroy": (None, [_vp]),
    "lux_multi_gpus": (_int, [_vp]),
    "lux_multi_engine": (_vp, [_vp, _int]),
    "lux_multi_last_error": (C.c_char_p, [_vp]),
    "lux_multi_load_bsk": (_int, [_vp, _u64p, _sz, _int]),
    "lux_multi_load_glwe_ksk": (_int, [_vp, _u64p, _sz, _int]),
    "lux_multi_load_lwe_ksk": (_int, [_vp, _u64p, _sz]),
    "lux_multi_replicate_keys": (_int, [_vp]),
    "lux_multi_pbs_batch": (_int, [_vp, _vp, _vp, _u32, _vp, _sz, _vp]),
}


def lib():
    """Load the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        path = lib_path()
        if not path.exists():
            raise LuxError(-100, f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                                 f"or `make -C tfhe_b200/csrc` (there is no CPU fallback)")
        L = C.CDLL(str(path))
        for name, (res, args) in ABI.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if hasattr(L, "lux_debug_force_teams"):
            L.lux_debug_force_teams.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_split_mode"):
            L.lux_debug_split_mode.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_ks_splits"):
            L.lux_debug_ks_splits.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_br_regs"):
            L.lux_debug_br_regs.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_cluster_regs"):
            L.lux_debug_cluster_regs.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_generic_gadget"):
            L.lux_debug_generic_gadget.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_ks_mode"):
            L.lux_debug_ks_mode.argtypes = [_vp, _int]
        if hasattr(L, "lux_debug_int_peak"):
            L.lux_debug_int_peak.argtypes = [_vp, C.POINTER(C.c_double * 3)]
            L.lux_debug_int_peak.restype = _int
        if hasattr(L, "lux_debug_fp64_peak"):
            L.lux_debug_fp64_peak.argtypes = [_vp, C.POINTER(C.c_double * 1)]
            L.lux_debug_fp64_peak.restype = _int
        if hasattr(L, "lux_debug_field_probe"):
            L.lux_debug_field_probe.argtypes = [_vp, _u64p, _u64p, _sz, _u64p]
            L.lux_debug_field_probe.restype = _int
        _lib = L
    return _lib


def _u64arr(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint64)


def make_testv(N: int, p: int, delta_: int, lut, arith: int = ARITH_GOLDILOCKS_NTT) -> np.ndarray:
    """get_testv (crypto/mod.rs:17-27) generalised to an arbitrary p-entry table."""
    out = np.zeros(N, np.uint64)
    rc = lib().lux_make_testv_arith(arith, N, p, delta_, _u64arr(lut), out)
    if rc:
        raise LuxError(rc, "lux_make_testv: bad arguments")
    return out


def delta(space: int, arith: int = ARITH_GOLDILOCKS_NTT) -> int:
    return lib().lux_delta_arith(arith, space)


def _ptr(x):
    """host numpy array or torch tensor -> raw address"""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return x.data_ptr()  # torch tensor (host pinned or device)


class Engine:
    """One engine per GPU (lux_engine)."""

    def __init__(self, params: Params, device: int = 0):
        self.p = params
        self._h = _vp()
        cp = params.c()
        rc = lib().lux_engine_create(C.byref(cp), device, C.byref(self._h))
        if rc:
            raise LuxError(rc, lib().lux_last_error(None).decode())
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            lib().lux_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int):
        if rc:
            raise LuxError(rc, lib().lux_last_error(self._h).decode())

    # ---- keys ------------------------------------------------------------------------------
    def load_bsk(self, bsk, domain=DOMAIN_STANDARD):
        a = _u64arr(bsk).reshape(-1)
        self._ck(lib().lux_load_bsk(self._h, a, a.size, domain))

    def load_glwe_ksk(self, ksk, domain=DOMAIN_STANDARD):
        a = _u64arr(ksk).reshape(-1)
        self._ck(lib().lux_load_glwe_ksk(self._h, a, a.size, domain))

    def load_lwe_ksk(self, ksk):
        a = _u64arr(ksk).reshape(-1)
        self._ck(lib().lux_load_lwe_ksk(self._h, a, a.size))

    def load_keyset(self, ks):
        """ks: any object with .bsk and .glwe_ksk / .lwe_ksk (standard domain)"""
        self.load_bsk(ks.bsk)
        if getattr(ks, "glwe_ksk", None) is not None:
            self.load_glwe_ksk(ks.glwe_ksk)
        if getattr(ks, "lwe_ksk", None) is not None:
            self.load_lwe_ksk(ks.lwe_ksk)

    def load_key_blob(self, blob: bytes):
        """one serialized evaluation key (serialize(..., BLOB_BSK / BLOB_LWE_KSK / BLOB_GLWE_KSK)): refused unless it was
        made for this engine's parameter set and arithmetic"""
        buf = np.frombuffer(blob, np.uint8)
        self._ck(lib().lux_load_key_blob(self._h, buf.ctypes.data, buf.size))

    @property
    def key_blob_bytes(self) -> int:
        return lib().lux_key_blob_bytes(self._h)

    def export_keys_device(self, dev_tensor):
        self._ck(lib().lux_export_keys_device(self._h, dev_tensor.data_ptr(), dev_tensor.numel() * dev_tensor.element_size()))

    def import_keys_device(self, dev_tensor):
        self._ck(lib().lux_import_keys_device(self._h, dev_tensor.data_ptr(), dev_tensor.numel() * dev_tensor.element_size()))

    # ---- PBS ----------------------------------------------------------------------------------
    def pbs_batch(self, lwe_in, testvs, lut_idx=None, out=None):
        """Host buffers (numpy, or pinned torch tensors viewed as int64/uint64)."""
        if isinstance(lwe_in, np.ndarray):
            lwe_in = _u64arr(lwe_in).reshape(-1, self.p.ct_len)
            B = lwe_in.shape[0]
            if out is None:
                out = np.empty_like(lwe_in)
        else:
            B = lwe_in.numel() // self.p.ct_len
            assert out is not None
        if isinstance(testvs, np.ndarray):
            testvs = _u64arr(testvs).reshape(-1, self.p.N)
            n_luts = testvs.shape[0]
        else:
            n_luts = testvs.numel() // self.p.N
        if lut_idx is not None and isinstance(lut_idx, (list, tuple, np.ndarray)):
            lut_idx = np.ascontiguousarray(lut_idx, dtype=np.uint32)
        self._ck(lib().lux_pbs_batch(self._h, _ptr(lwe_in), _ptr(testvs), n_luts, _ptr(lut_idx), B, _ptr(out)))
        return out

    # ---- boolean gates: one bootstrap each (BASELINE config 1); bits are m * delta(4) ------------
    def gate(self, name: str, a, b):
        """NAND / AND / OR / NOR / XOR / XNOR on [B][ct_len] host arrays of encrypted bits"""
        a = _u64arr(a).reshape(-1, self.p.ct_len)
        b = _u64arr(b).reshape(-1, self.p.ct_len)
        assert a.shape == b.shape
        out = np.empty_like(a)
        self._ck(lib().lux_gate_batch(self._h, GATES[name.lower()], _ptr(a), _ptr(b), a.shape[0], _ptr(out)))
        return out

    def gate_device(self, name: str, d_a, d_b, B: int, d_out, stream=None):
        self._ck(lib().lux_gate_batch_device(self._h, GATES[name.lower()], d_a.data_ptr(), d_b.data_ptr(), B,
                                             d_out.data_ptr(), stream))

    def gate_not(self, a):
        a = _u64arr(a).reshape(-1, self.p.ct_len)
        out = np.empty_like(a)
        self._ck(lib().lux_gate_not(self._h, _ptr(a), a.shape[0], _ptr(out)))
        return out

    def pbs_batch_device(self, d_in, d_testvs, d_lut_idx, B: int, d_out, n_luts: int | None = None, stream=None):
        """torch CUDA tensors (int64 storage of the u64 words); asynchronous on `stream`/engine stream."""
        if n_luts is None:
            n_luts = d_testvs.numel() // self.p.N
        self._ck(lib().lux_pbs_batch_device(self._h, d_in.data_ptr(), d_testvs.data_ptr(), n_luts,
                                            None if d_lut_idx is None else d_lut_idx.data_ptr(), B, d_out.data_ptr(),
                                            stream))

    def pbs_batch_device_static(self, d_in, d_testvs, d_lut_idx, B: int, d_out, n_luts: int | None = None, stream=None):
        """pbs_batch_device without any allocation inside the library (lux_pbs_batch_device_static: scratch in a static
        device array of 16 MiB; torus engines, B <= 1024 at the named set; overlapping calls on one GPU race)."""
        if n_luts is None:
            n_luts = d_testvs.numel() // self.p.N
        self._ck(lib().lux_pbs_batch_device_static(self._h, d_in.data_ptr(), d_testvs.data_ptr(), n_luts,
                                                   None if d_lut_idx is None else d_lut_idx.data_ptr(), B, d_out.data_ptr(),
                                                   stream))

    def synchronize(self):
        self._ck(lib().lux_synchronize(self._h))

    def set_timing(self, on: bool):
        self._ck(lib().lux_set_timing(self._h, int(on)))

    def last_timing(self):
        buf = (C.c_float * 2)()
        self._ck(lib().lux_last_timing(self._h, C.byref(buf)))
        return float(buf[0]), float(buf[1])

    @property
    def launch_count(self) -> int:
        return lib().lux_launch_count(self._h)

    def int_peak(self):
        """measured integer-pipe rates of this GPU: (IMAD.WIDE/s, IADD/s, register-only Goldilocks butterflies/s)"""
        buf = (C.c_double * 3)()
        self._ck(lib().lux_debug_int_peak(self._h, C.byref(buf)))
        return float(buf[0]), float(buf[1]), float(buf[2])

    def fp64_peak(self) -> float:
        """measured FP64-pipe rate of this GPU: DFMA lane-operations per second (register-only)"""
        buf = (C.c_double * 1)()
        self._ck(lib().lux_debug_fp64_peak(self._h, C.byref(buf)))
        return float(buf[0])

    def field_probe(self, a, b):
        """device field arithmetic on raw words: rows = a+b, a-b, a*b, CT butterfly (2), GS butterfly (2)"""
        a, b = _u64arr(a).reshape(-1), _u64arr(b).reshape(-1)
        out = np.empty((7, a.size), np.uint64)
        self._ck(lib().lux_debug_field_probe(self._h, a, b, a.size, out.reshape(-1)))
        return out

    def wave_ciphertexts(self, B: int) -> int:
        """ciphertexts resident on the GPU at once in the blind-rotate launch for a batch of B"""
        return int(lib().lux_wave_ciphertexts(self._h, B))

    def split_mode(self, mode: int):
        """latency-mode blind rotate (one ciphertext per CTA, 2*ell teams): -1 automatic (B <= #SMs), 0 never, 1 always"""
        self._ck(lib().lux_debug_split_mode(self._h, mode))

    def ks_splits(self, n: int):
        """key-switch input slices for small batches: 0 automatic, n > 0 forced (1 = single pass)"""
        self._ck(lib().lux_debug_ks_splits(self._h, n))

    def br_regs(self, regs: int):
        """throughput blind rotate: coefficients per thread (16 = wide tile, 8 = narrow tile; named sets only)"""
        self._ck(lib().lux_debug_br_regs(self._h, regs))

    def cluster_regs(self, regs: int):
        """cluster latency mode: coefficients per thread (16, 8, 4; 0 = automatic: the narrowest that fits a CTA)"""
        self._ck(lib().lux_debug_cluster_regs(self._h, regs))

    def generic_gadget(self, on: bool):
        """tests / A-B: run the blind-rotate instantiation that reads the gadget from its arguments even where a
        compile-time instantiation for the named parameter set exists"""
        self._ck(lib().lux_debug_generic_gadget(self._h, int(bool(on))))

    def ks_mode(self, mode: int):
        """LWE key switch kernel: -1 automatic (int8 tensor-core contraction whenever the parameter set is eligible),
        0 integer-pipe kernel only, 1 mma.sync kernel (ks_mma.cuh), 2 tcgen05 / tensor-memory kernel (ks_tc5.cuh)"""
        self._ck(lib().lux_debug_ks_mode(self._h, mode))

    def fft_variant(self, v: int):
        lib().lux_debug_fft_variant.argtypes = [_vp, _int]
        self._ck(lib().lux_debug_fft_variant(self._h, v))

    def fft_trace(self):
        """phase-cycle counters of the torus blind rotation since the last call (FFT64_TRACE_BUILD builds)"""
        buf = (C.c_ulonglong * 16)()
        lib().lux_debug_fft_trace.argtypes = [_vp, C.POINTER(C.c_ulonglong * 16)]
        self._ck(lib().lux_debug_fft_trace(self._h, C.byref(buf)))
        return [int(v) for v in buf]

    def fft_lat_mode(self, mode: int):
        """-1 = latency kernel (one ciphertext per CTA, eight warps) when the batch leaves SMs idle, 0 = never, 1 = always"""
        lib().lux_debug_fft_lat_mode.argtypes = [_vp, _int]
        self._ck(lib().lux_debug_fft_lat_mode(self._h, mode))

    def fft_tail_split(self, on: bool):
        """torus blind rotation of a batch that is not a whole number of waves: whole waves + a tail launch of its own
        (default) or one launch (A/B, tests)"""
        lib().lux_debug_fft_tail_split.argtypes = [_vp, _int]
        self._ck(lib().lux_debug_fft_tail_split(self._h, int(bool(on))))

    def fft_stagger(self, ns: int):
        lib().lux_debug_fft_stagger.argtypes = [_vp, _int]
        self._ck(lib().lux_debug_fft_stagger(self._h, ns))

    def force_teams(self, t: int):
        lib().lux_debug_force_teams(self._h, t)

    # ---- pieces -------------------------------------------------------------------------------
    def ntt_forward(self, polys):
        a = _u64arr(polys).reshape(-1, self.p.N)
        out = np.empty_like(a)
        self._ck(lib().lux_ntt_forward(self._h, a.reshape(-1), a.shape[0], out.reshape(-1)))
        return out

    def ntt_backward(self, polys):
        a = _u64arr(polys).reshape(-1, self.p.N)
        out = np.empty_like(a)
        self._ck(lib().lux_ntt_backward(self._h, a.reshape(-1), a.shape[0], out.reshape(-1)))
        return out

    def ntt_any(self, polys, N: int, inverse=False):
        a = _u64arr(polys).reshape(-1, N)
        out = np.empty_like(a)
        self._ck(lib().lux_ntt_any(self._h, N, int(inverse), a.reshape(-1), a.shape[0], out.reshape(-1)))
        return out

    def decompose(self, values, logB: int):
        a = _u64arr(values).reshape(-1)
        L = (64 + logB - 1) // logB
        out = np.empty((a.size, L), np.uint64)
        self._ck(lib().lux_decompose(self._h, a, a.size, logB, out.reshape(-1)))
        return out

    def mod_switch(self, values):
        a = _u64arr(values).reshape(-1)
        out = np.empty(a.size, np.uint32)
        self._ck(lib().lux_mod_switch(self._h, a, a.size, out))
        return out

    def lwe_keyswitch(self, cts):
        a = _u64arr(cts).reshape(-1, self.p.k * self.p.N + 1)
        out = np.empty((a.shape[0], self.p.n + 1), np.uint64)
        self._ck(lib().lux_lwe_keyswitch(self._h, a.reshape(-1), a.shape[0], out.reshape(-1)))
        return out

    def blind_rotate(self, lwe_small, testvs, lut_idx=None, steps: int | None = None):
        a = _u64arr(lwe_small).reshape(-1, self.p.n + 1)
        tv = _u64arr(testvs).reshape(-1, self.p.N)
        idx = None if lut_idx is None else np.ascontiguousarray(lut_idx, dtype=np.uint32)
        out = np.empty((a.shape[0], 2 * self.p.N), np.uint64)
        self._ck(lib().lux_blind_rotate(self._h, a.reshape(-1), tv.reshape(-1), tv.shape[0], _ptr(idx), a.shape[0],
                                        self.p.n if steps is None else steps, out.reshape(-1)))
        return out

    # ---- pieces of the torus / f64-FFT arithmetic
    def fft(self, z, inverse=False):
        """[n_polys][N/2] complex128: forward natural -> engine order, or the unnormalised inverse"""
        a = np.ascontiguousarray(z, dtype=np.complex128).reshape(-1, self.p.N // 2)
        out = np.empty_like(a)
        self._ck(lib().lux_fft(self._h, a.ctypes.data, a.shape[0], int(inverse), out.ctypes.data))
        return out

    def torus_decompose(self, values):
        a = _u64arr(values).reshape(-1)
        out = np.empty(a.size, np.int64)
        self._ck(lib().lux_torus_decompose(self._h, a, a.size, out.ctypes.data))
        return out

    def bsk_fourier(self):
        """resident Fourier key rows [n][32 rows][4 polys (j*2+r)][32 lanes] complex128"""
        out = np.empty((self.p.n, 32, 4, 32), np.complex128)
        self._ck(lib().lux_export_bsk_fourier(self._h, out.ctypes.data, out.size * 2))
        return out

    def external_product(self, ggsw, glwe, domain=DOMAIN_STANDARD):
        g = _u64arr(ggsw).reshape(-1, self.p.ggsw_len)
        x = _u64arr(glwe).reshape(-1, 2 * self.p.N)
        assert g.shape[0] == x.shape[0]
        out = np.empty_like(x)
        self._ck(lib().lux_external_product(self._h, g.reshape(-1), domain, x.reshape(-1), x.shape[0], out.reshape(-1)))
        return out


# ---- serialized containers (include/lux_tfhe.h "serialized keys and ciphertexts"); host-only ----------------------------
BLOB_LWE, BLOB_RADIX, BLOB_BSK, BLOB_GLWE_KSK, BLOB_LWE_KSK = 1, 2, 3, 4, 5
ERR_FORMAT, ERR_MISMATCH = -6, -7


class _CBlobInfo(C.Structure):
    _fields_ = [("version", C.c_uint32), ("kind", C.c_uint32), ("params", _CParams), ("aux", C.c_uint32),
                ("count", C.c_uint64), ("payload_words", C.c_uint64)]


def serialize(params: Params, kind: int, words, aux: int = 0, count: int = 1) -> bytes:
    """words of `count` objects of `kind` -> one self-describing container (bytes)"""
    w = _u64arr(words).reshape(-1)
    cp = params.c()
    n = lib().lux_serialized_bytes(C.byref(cp), kind, aux, count)
    if n == 0 or n != 80 + 8 * w.size:
        raise LuxError(-1, f"{w.size} words do not make {count} object(s) of kind {kind} (aux {aux}) at these parameters")
    out = np.empty(n, np.uint8)
    rc = lib().lux_serialize(C.byref(cp), kind, aux, count, w.ctypes.data, out.ctypes.data, n)
    if rc:
        raise LuxError(rc, "lux_serialize failed")
    return out.tobytes()


def deserialize(blob: bytes, expect: Params | None = None, expect_kind: int = 0):
    """-> (words, info dict); raises LuxError(ERR_FORMAT) for a damaged container, LuxError(ERR_MISMATCH) for a valid
    one that belongs to another parameter set / arithmetic / kind"""
    buf = np.frombuffer(blob, np.uint8)
    info = _CBlobInfo()
    rc = lib().lux_blob_inspect(buf.ctypes.data, buf.size, C.byref(info))
    if rc:
        raise LuxError(rc, "not a valid container (bad magic, version, length or header checksum)")
    out = np.empty(info.payload_words, np.uint64)
    cp = expect.c() if expect is not None else None
    rc = lib().lux_deserialize(buf.ctypes.data, buf.size, C.byref(cp) if cp is not None else None, expect_kind,
                               out.ctypes.data, out.size, C.byref(info))
    if rc:
        raise LuxError(rc, "container is for another parameter set / arithmetic / kind" if rc == ERR_MISMATCH
                       else "container payload is damaged (checksum)")
    p = info.params
    return out, {"version": info.version, "kind": info.kind, "aux": info.aux, "count": info.count,
                 "params": Params(p.n, p.N, p.k, p.ell, p.logB, p.ell_ks, p.logB_ks, p.order, p.arith)}


class MultiEngine:
    """Several GPUs of one box behind one handle (lux_engine_create_multi): keys converted on GPU 0 and copied device to
    device, batches sliced over the GPUs inside the library with one host thread per GPU.  No torch, no NCCL: this is the
    path a Node / ctypes host uses.  n_gpus <= 0 takes every visible device."""

    def __init__(self, params: Params, n_gpus: int = 0):
        self.p = params
        h = _vp()
        cp = params.c()
        rc = lib().lux_engine_create_multi(C.byref(cp), n_gpus, C.byref(h))
        if rc:
            raise LuxError(rc, lib().lux_multi_last_error(None).decode())
        self._h = h

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().lux_multi_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise LuxError(rc, lib().lux_multi_last_error(self._h).decode())

    @property
    def n_gpus(self) -> int:
        return lib().lux_multi_gpus(self._h)

    def load_keyset(self, ks):
        """ks: any object with .bsk and .glwe_ksk / .lwe_ksk (standard domain); converts on GPU 0, replicates to all"""
        b = _u64arr(ks.bsk).reshape(-1)
        self._ck(lib().lux_multi_load_bsk(self._h, b, b.size, DOMAIN_STANDARD))
        if getattr(ks, "glwe_ksk", None) is not None:
            g = _u64arr(ks.glwe_ksk).reshape(-1)
            self._ck(lib().lux_multi_load_glwe_ksk(self._h, g, g.size, DOMAIN_STANDARD))
        if getattr(ks, "lwe_ksk", None) is not None:
            k = _u64arr(ks.lwe_ksk).reshape(-1)
            self._ck(lib().lux_multi_load_lwe_ksk(self._h, k, k.size))
        self._ck(lib().lux_multi_replicate_keys(self._h))

    def pbs_batch(self, lwe_in, testvs, lut_idx=None, out=None):
        x = _u64arr(lwe_in).reshape(-1, self.p.ct_len)
        tv = _u64arr(testvs).reshape(-1, self.p.N)
        idx = None if lut_idx is None else np.ascontiguousarray(lut_idx, dtype=np.uint32)
        if out is None:
            out = np.empty_like(x)
        self._ck(lib().lux_multi_pbs_batch(self._h, x.ctypes.data, tv.ctypes.data, tv.shape[0],
                                           None if idx is None else idx.ctypes.data, x.shape[0], out.ctypes.data))
        return out
