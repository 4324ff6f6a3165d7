"""ctypes binding of include/lux_radix.h — euintN operations over the batched PBS.

Host-side mirror of the reference's operator surface (packages/wasm/src/index.ts:19-22 add/sub/eq/lt,
js/v2-sdk/src/luxd/client.ts:34-39 the /evaluate ops, FHE.select/le/... of WrappingERC20.sol:86-102).
Marshals numpy buffers only; every ciphertext operation runs in the CUDA library.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from .engine import Engine, LuxError, lib as _engine_lib

OPS = {"add": 0, "sub": 1, "mul": 2, "neg": 3, "eq": 4, "ne": 5, "lt": 6, "le": 7, "gt": 8, "ge": 9, "min": 10, "max": 11,
       "and": 12, "or": 13, "xor": 14, "not": 15, "shl": 16, "shr": 17, "rotl": 18, "rotr": 19, "div": 20, "rem": 21}
PREDICATES = {"eq", "ne", "lt", "le", "gt", "ge"}
UNARY = {"neg", "not"}
SCALAR_ONLY = {"div", "rem"}   # plaintext right operand only (fhEVM div/rem overloads); circuit depends on the divisor
PROG_SELECT, PROG_ERC20 = 100, 101
BLOCK_P = 16          # plaintext values per block (2 message + 2 carry bits); padding bit on top

_vp, _sz, _u64, _int, _uint = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int, C.c_uint
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")

# every symbol include/lux_radix.h declares: name -> (restype, argtypes)
ABI = {
    "lux_radix_blocks": (_sz, [_uint]),
    "lux_op_is_predicate": (_int, [_int]),
    "lux_op_is_unary": (_int, [_int]),
    "lux_op_is_scalar_only": (_int, [_int]),
    "lux_eval_scalar_pbs_count": (_u64, [_int, _uint, _u64]),
    "lux_radix_program_scalar": (_sz, [_int, _uint, _vp, _sz, _vp, _sz]),
    "lux_eval_scalar_wide": (_int, [_vp, _int, _uint, _vp, _vp, _sz, _sz, _vp]),
    "lux_eval_pbs_count": (_u64, [_int, _uint]),
    "lux_eval_depth": (_u64, [_int, _uint]),
    "lux_eval": (_int, [_vp, _int, _uint, _vp, _vp, _sz, _vp]),
    "lux_eval_device": (_int, [_vp, _int, _uint, _vp, _vp, _sz, _vp]),
    "lux_eval_scalar": (_int, [_vp, _int, _uint, _vp, _vp, _sz, _vp]),
    "lux_select": (_int, [_vp, _uint, _vp, _vp, _vp, _sz, _vp]),
    "lux_select_device": (_int, [_vp, _uint, _vp, _vp, _vp, _sz, _vp]),
    "lux_cast": (_int, [_vp, _uint, _uint, _vp, _sz, _vp]),
    "lux_trivial_encrypt": (_int, [_vp, _uint, _vp, _sz, _vp]),
    "lux_erc20_transfer": (_int, [_vp, _vp, _vp, _vp, _sz, _vp, _vp]),
    "lux_erc20_transfer_device": (_int, [_vp, _vp, _vp, _vp, _sz, _vp, _vp]),
    "lux_erc20_pbs_count": (_u64, []),
    "lux_erc20_depth": (_u64, []),
    "lux_radix_set_pool_bytes": (_int, [_vp, _sz]),
    "lux_multi_eval": (_int, [_vp, _int, _uint, _vp, _vp, _sz, _vp]),
    "lux_multi_select": (_int, [_vp, _uint, _vp, _vp, _vp, _sz, _vp]),
    "lux_multi_erc20_transfer": (_int, [_vp, _vp, _vp, _vp, _sz, _vp, _vp]),
    "lux_radix_program": (_sz, [_int, _uint, _vp, _sz]),
    "lux_radix_lut": (_int, [_int, _vp]),
    "lux_radix_lut_signed": (_int, [_int]),
    "lux_radix_lut_dual": (_int, [_int]),
}
_bound = False


def lib():
    global _bound
    L = _engine_lib()
    if not _bound:
        for name, (res, args) in ABI.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _bound = True
    return L


def _limbs(v: int, n: int) -> np.ndarray:
    return np.array([(v >> (64 * i)) & (2**64 - 1) for i in range(n)], dtype=np.uint64)


def blocks(bit_width: int) -> int:
    return int(lib().lux_radix_blocks(bit_width))


def pbs_count(op: str, bit_width: int, scalar: int | None = None) -> int:
    if scalar is not None:
        return int(lib().lux_eval_scalar_pbs_count(OPS[op], bit_width, int(scalar) & (2**64 - 1)))
    return int(lib().lux_eval_pbs_count(OPS[op], bit_width))


def depth(op: str, bit_width: int) -> int:
    return int(lib().lux_eval_depth(OPS[op], bit_width))


# ---- circuit introspection (no GPU needed): the program the GPU executes, for the CPU oracle --------
@dataclass
class Lin:
    konst: int
    terms: list  # [(slot, coef)]


@dataclass
class Program:
    in_blocks: list
    out_blocks: list
    levels: list = field(default_factory=list)   # [[(lut, out_slot, Lin)]]
    outs: list = field(default_factory=list)     # [[Lin per block] per output operand]

    @property
    def n_pbs(self):
        return sum(len(lv) for lv in self.levels)


PROG_SHALLOW = 0x1000   # lux_radix.h LUX_PROG_SHALLOW
PROG_NODUAL = 0x2000    # lux_radix.h LUX_PROG_NODUAL


def runs_shallow(bit_width: int, batch: int) -> bool:
    """does lux_eval* / lux_erc20_transfer* take the shallow variant for this batch? (radix.cu pick_variant)"""
    return batch * blocks(bit_width) <= 296


def program(op, bit_width: int, scalar: int | None = None, shallow: bool = False, dual: bool = True) -> Program:
    """the circuit of `op`; div/rem need the plaintext divisor `scalar` (the circuit is divisor-specific);
    shallow = the parallel-prefix-carry variant the library runs for very small batches; dual = False: the variant without
    dual tables that engines whose bootstraps do not start with the key switch run"""
    code = OPS[op] if isinstance(op, str) else int(op)
    if shallow:
        code |= PROG_SHALLOW
    if not dual:
        code |= PROG_NODUAL
    sc = _limbs(int(scalar or 0), 4)
    n = lib().lux_radix_program_scalar(code, bit_width, sc.ctypes.data, 4, None, 0)
    if n == 0:
        raise LuxError(-1, f"no circuit for op {op} at width {bit_width}" + (f", scalar {scalar}" if scalar is not None else ""))
    buf = np.zeros(n, np.int64)
    lib().lux_radix_program_scalar(code, bit_width, sc.ctypes.data, 4, buf.ctypes.data, n)
    s = buf.tolist()
    pos = 0

    def take():
        nonlocal pos
        v = s[pos]
        pos += 1
        return v

    def lin():
        k, nt = take(), take()
        return Lin(k, [(take(), take()) for _ in range(nt)])

    in_blocks = [take() for _ in range(take())]
    out_blocks = [take() for _ in range(take())]
    p = Program(in_blocks, out_blocks)
    for _ in range(take()):
        lv = []
        for _ in range(take()):
            lut, out = take(), take()
            lv.append((lut, out, lin()))
        p.levels.append(lv)
    for nb in out_blocks:
        p.outs.append([lin() for _ in range(nb)])
    assert pos == len(s)
    return p


def luts() -> np.ndarray:
    n = lib().lux_radix_lut(0, None)
    t = np.zeros((n, 16), np.uint64)
    for i in range(n):
        rc = lib().lux_radix_lut(i, t[i].ctypes.data)
        assert rc == 0
    return t


def dual_luts() -> set:
    """ids of the dual tables: their nodes write slots out and out + 1 from one blind rotation (lux_radix_lut_dual)"""
    return {i for i in range(lib().lux_radix_lut(0, None)) if lib().lux_radix_lut_dual(i)}


def signed_luts() -> set:
    """ids of the tables that circuits feed with x in (-16, 16) (negacyclic extension T(x) = -T(x + 16))"""
    return {i for i in range(lib().lux_radix_lut(0, None)) if lib().lux_radix_lut_signed(i)}


# ---- evaluation -------------------------------------------------------------------------------------
def _arr(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


class Radix:
    """euintN evaluator bound to one Engine (one GPU).  Buffers: [batch][blocks][ct_words] u64."""

    def __init__(self, engine: Engine):
        self.e = engine
        self.w = engine.p.ct_len
        lib()

    def _ck(self, rc):
        self.e._ck(rc)

    def _shape(self, a, nb):
        if nb == 0:
            raise LuxError(-1, "unsupported bit width (even, 2..256)")
        a = _arr(a).reshape(-1, nb, self.w)
        return a, a.shape[0]

    def eval(self, op: str, bit_width: int, lhs, rhs=None):
        nb = blocks(bit_width)
        lhs, B = self._shape(lhs, nb)
        if op in UNARY:
            rp = None
        else:
            rhs, Br = self._shape(rhs, nb)
            assert Br == B
            rp = rhs.ctypes.data
        out = np.empty((B, 1 if op in PREDICATES else nb, self.w), np.uint64)
        self._ck(lib().lux_eval(self.e._h, OPS[op], bit_width, lhs.ctypes.data, rp, B, out.ctypes.data))
        return out

    def eval_scalar(self, op: str, bit_width: int, lhs, scalars):
        nb = blocks(bit_width)
        lhs, B = self._shape(lhs, nb)
        out = np.empty((B, 1 if op in PREDICATES else nb, self.w), np.uint64)
        if op in SCALAR_ONLY and bit_width > 64:
            sc = np.concatenate([_limbs(int(s), 4) for s in scalars])
            assert sc.size == 4 * B
            self._ck(lib().lux_eval_scalar_wide(self.e._h, OPS[op], bit_width, lhs.ctypes.data, sc.ctypes.data, 4, B,
                                                out.ctypes.data))
            return out
        sc = _arr([int(s) & (2**64 - 1) for s in scalars])
        assert sc.size == B
        self._ck(lib().lux_eval_scalar(self.e._h, OPS[op], bit_width, lhs.ctypes.data, sc.ctypes.data, B, out.ctypes.data))
        return out

    def select(self, bit_width: int, cond, a, b):
        nb = blocks(bit_width)
        a, B = self._shape(a, nb)
        b, _ = self._shape(b, nb)
        cond = _arr(cond).reshape(B, self.w)
        out = np.empty_like(a)
        self._ck(lib().lux_select(self.e._h, bit_width, cond.ctypes.data, a.ctypes.data, b.ctypes.data, B, out.ctypes.data))
        return out

    def cast(self, from_width: int, to_width: int, x):
        x, B = self._shape(x, blocks(from_width))
        out = np.empty((B, blocks(to_width), self.w), np.uint64)
        self._ck(lib().lux_cast(self.e._h, from_width, to_width, x.ctypes.data, B, out.ctypes.data))
        return out

    def trivial_encrypt(self, bit_width: int, values):
        v = _arr([int(x) & (2**64 - 1) for x in values])
        out = np.empty((v.size, blocks(bit_width), self.w), np.uint64)
        self._ck(lib().lux_trivial_encrypt(self.e._h, bit_width, v.ctypes.data, v.size, out.ctypes.data))
        return out

    def erc20_transfer(self, from_bal, to_bal, amount):
        f, B = self._shape(from_bal, 32)
        t, _ = self._shape(to_bal, 32)
        a, _ = self._shape(amount, 32)
        nf, nt = np.empty_like(f), np.empty_like(t)
        self._ck(lib().lux_erc20_transfer(self.e._h, f.ctypes.data, t.ctypes.data, a.ctypes.data, B, nf.ctypes.data, nt.ctypes.data))
        return nf, nt

    def set_pool_bytes(self, n: int):
        self._ck(lib().lux_radix_set_pool_bytes(self.e._h, n))


class MultiRadix:
    """euintN evaluator over every GPU of a MultiEngine: shares of whole integers, sliced inside the library."""

    def __init__(self, multi):
        self.m = multi
        self.w = multi.p.ct_len
        lib()

    def eval(self, op: str, bit_width: int, lhs, rhs=None):
        nb = blocks(bit_width)
        lhs = _arr(lhs).reshape(-1, nb, self.w)
        B = lhs.shape[0]
        rp = None
        if op not in UNARY:
            rhs = _arr(rhs).reshape(B, nb, self.w)
            rp = rhs.ctypes.data
        out = np.empty((B, 1 if op in PREDICATES else nb, self.w), np.uint64)
        self.m._ck(lib().lux_multi_eval(self.m._h, OPS[op], bit_width, lhs.ctypes.data, rp, B, out.ctypes.data))
        return out

    def select(self, bit_width: int, cond, a, b):
        nb = blocks(bit_width)
        a = _arr(a).reshape(-1, nb, self.w)
        B = a.shape[0]
        b = _arr(b).reshape(B, nb, self.w)
        cond = _arr(cond).reshape(B, self.w)
        out = np.empty_like(a)
        self.m._ck(lib().lux_multi_select(self.m._h, bit_width, cond.ctypes.data, a.ctypes.data, b.ctypes.data, B, out.ctypes.data))
        return out

    def erc20_transfer(self, from_bal, to_bal, amount):
        f = _arr(from_bal).reshape(-1, 32, self.w)
        B = f.shape[0]
        t, a = _arr(to_bal).reshape(B, 32, self.w), _arr(amount).reshape(B, 32, self.w)
        nf, nt = np.empty_like(f), np.empty_like(t)
        self.m._ck(lib().lux_multi_erc20_transfer(self.m._h, f.ctypes.data, t.ctypes.data, a.ctypes.data, B, nf.ctypes.data, nt.ctypes.data))
        return nf, nt
