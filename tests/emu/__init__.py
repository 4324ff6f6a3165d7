"""ctypes loader for the host-thread kernel emulator (tests/emu/emu.cpp). Test infrastructure."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_SRC = [_HERE / "emu.cpp", _HERE.parent.parent / "tfhe_b200" / "csrc" / "tfhe_kernels.cuh",
        _HERE.parent.parent / "tfhe_b200" / "csrc" / "gl64.cuh"]
_LIB = _HERE / "libemu.so"
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")


class BRArgs(C.Structure):
    _fields_ = [("n", C.c_uint32), ("steps", C.c_uint32), ("ell", C.c_uint32), ("logB", C.c_uint32),
                ("L", C.c_uint32), ("B", C.c_uint32), ("out_mode", C.c_uint32), ("n_out", C.c_uint32),
                ("bsk", C.c_void_p), ("ksk", C.c_void_p), ("lwe", C.c_void_p), ("testv", C.c_void_p),
                ("lut_idx", C.c_void_p), ("lut_div", C.c_uint32), ("out", C.c_void_p), ("tw_fwd", C.c_void_p),
                ("body_add", C.c_uint64)]


class KSArgs(C.Structure):
    _fields_ = [("n", C.c_uint32), ("kN", C.c_uint32), ("ell", C.c_uint32), ("logB", C.c_uint32),
                ("L", C.c_uint32), ("B", C.c_uint32),
                ("ksk", C.c_void_p), ("ksum", C.c_void_p), ("in_", C.c_void_p), ("out", C.c_void_p),
                ("splits", C.c_uint32), ("partial", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        newest = max(p.stat().st_mtime for p in _SRC)
        if not _LIB.exists() or _LIB.stat().st_mtime < newest:
            subprocess.run(["/usr/bin/g++", "-std=c++20", "-O2", "-fPIC", "-shared", "-pthread", "-o", str(_LIB),
                            str(_SRC[0])], check=True, capture_output=True)
        L = C.CDLL(str(_LIB))
        L.emu_ntt.argtypes = [C.c_int, C.c_int, C.c_int, _u64p, _u64p, _u64p, C.c_uint64, C.c_uint32, C.c_uint32]
        L.emu_force_slow.argtypes = [C.c_int]
        L.emu_ntt_r.argtypes = [C.c_int, C.c_int, _u64p, _u64p, _u64p, C.c_uint64, C.c_uint32, C.c_uint32]
        L.emu_blind_rotate.argtypes = [C.c_int, C.c_int, C.POINTER(BRArgs)]
        L.emu_blind_rotate_split.argtypes = [C.c_int, C.c_int, C.POINTER(BRArgs)]
        L.emu_keyswitch.argtypes = [C.POINTER(KSArgs), C.c_int, C.c_int, C.c_int]
        L.emu_ksk_colsum.argtypes = [_u64p, _u64p, C.c_uint32, C.c_uint32]
        L.emu_ggsw_to_key.argtypes = [_u64p, C.c_size_t, C.c_uint32, C.c_uint32, C.c_uint64]
        L.emu_ggsw_to_key.restype = C.c_int
        L.emu_digit.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_int]
        L.emu_digit.restype = C.c_uint64
        L.emu_mod_switch.argtypes = [C.c_uint64, C.c_uint32]
        L.emu_mod_switch.restype = C.c_uint32
        for f in ("emu_mul", "emu_add", "emu_sub"):
            getattr(L, f).argtypes = [C.c_uint64, C.c_uint64]
            getattr(L, f).restype = C.c_uint64
        L.emu_sizeof_brargs.restype = C.c_size_t
        L.emu_sizeof_ksargs.restype = C.c_size_t
        assert L.emu_sizeof_brargs() == C.sizeof(BRArgs)
        assert L.emu_sizeof_ksargs() == C.sizeof(KSArgs)
        _lib = L
    return _lib


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None
