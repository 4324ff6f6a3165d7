"""CPU interpreter for euintN circuits — TEST INFRASTRUCTURE ONLY (see oracle.h).

The reference's integer layer (Go luxfi/tfhe) is not in the tree, so there is no reference algorithm to
restate for it: what is pinned by the reference is the *semantics* of each op (the fhEVM known-answer
tests, tests/golden/fhevm_kats.json; mocks/foundry/src/MockFHE.sol:130-508).  This module therefore
interprets the circuit the product publishes through lux_radix_program():

  interp_plain   on plaintext integers (tables applied to values) — checks the circuit's semantics
                 against the known answers without any cryptography;
  interp_cipher  on ciphertexts with the oracle's own PBS (orc_pbs_batch) and numpy field arithmetic —
                 the bit-exact expectation for what the GPU executor must produce.
"""
from __future__ import annotations

import numpy as np

Q = 0xFFFFFFFF00000001
BLOCK_P = 16


def split_blocks(value: int, nb: int):
    return [(value >> (2 * i)) & 3 for i in range(nb)]


def join_blocks(blocks) -> int:
    return sum(int(b) << (2 * i) for i, b in enumerate(blocks))


def interp_plain(prog, luts, operands, signed=(), dual=()):
    """operands: list of block lists (ints).  Returns list of output block lists.  Raises if a PBS input
    leaves [0,16) (the padding bit would be hit) — except for the tables listed in `signed` (lux_radix_lut_signed),
    which circuits feed with x in (-16, 16): a blind rotation is negacyclic, so x < 0 reads -T(x + 16).
    Tables listed in `dual` (lux_radix_lut_dual) write two slots from one bootstrap: T(x) to `out`, T(x + 1) to `out + 1`."""
    slots = {}
    s = 0
    for blocks in operands:
        for b in blocks:
            slots[s] = int(b)
            s += 1

    def lin(l):
        return l.konst + sum(c * slots[sl] for sl, c in l.terms)

    for lv in prog.levels:
        new = {}
        for lut, out, l in lv:
            x = lin(l)
            if lut in signed and -BLOCK_P < x < 0:
                new[out] = -int(luts[lut][x + BLOCK_P])
                continue
            if not 0 <= x < BLOCK_P:
                raise ValueError(f"PBS input {x} outside [0,16)")
            new[out] = int(luts[lut][x])
            if lut in dual:
                if x & 1:
                    raise ValueError(f"dual table read at odd input {x}")
                new[out + 1] = int(luts[lut][x + 1])
        slots.update(new)
    return [[lin(l) for l in o] for o in prog.outs]


def _fmul_small(a: np.ndarray, c: int) -> np.ndarray:
    """a * c mod q for a small non-negative integer c, exact (object arithmetic)."""
    return ((a.astype(object) * c) % Q).astype(np.uint64)


def sample_extract(acc, c: int, modulus: int):
    """LWE sample of coefficient c of the accumulator (mask | body), crypto/glwe.rs:96-106 generalised from c = 0:
    a'_p = mask[c - p] for p <= c, -mask[N + c - p] for p > c; b' = body[c]"""
    N = acc.shape[0] // 2
    mask, body = acc[:N].astype(object), acc[N:].astype(object)
    out = np.empty(N + 1, dtype=object)
    for p in range(N):
        out[p] = mask[c - p] if p <= c else (modulus - mask[N + c - p]) % modulus
    out[N] = body[c]
    return out.astype(np.uint64)


def interp_cipher(ctx, prog, luts, operands, testv_lut, delta, modulus=Q, dual=()):
    """operands: list of arrays [B][blocks][w] (uint64).  ctx: oracle Ctx (oracle.Ctx or oracle.torus.Ctx) with keys
    installed; modulus = Q (Goldilocks) or 2**64 (torus).  testv_lut(lut_row) -> test vector polynomial.
    A level that holds a table listed in `dual` is evaluated the way the executor does it: key switch, blind rotation to the
    raw accumulator, then one sample extraction per output (coefficient 0, and N/16 for the second output of a dual table).
    Returns list of arrays [B][out_blocks][w]."""
    B = operands[0].shape[0]
    w = operands[0].shape[2]
    slots = {}
    s = 0
    for op in operands:
        for i in range(op.shape[1]):
            slots[s] = op[:, i, :].astype(object)
            s += 1

    def lin(l):
        acc = np.zeros((B, w), dtype=object)
        for sl, c in l.terms:
            acc = (acc + slots[sl] * c) % modulus
        acc[:, w - 1] = (acc[:, w - 1] + (l.konst % 32) * delta) % modulus
        return acc

    tvs = np.stack([testv_lut(luts[i]) for i in range(len(luts))])
    for lv in prog.levels:
        if not lv:
            continue
        x = np.concatenate([lin(l) for _, _, l in lv]).astype(np.uint64)           # slot-major, like the GPU
        if any(lut in dual for lut, _, _ in lv):
            if getattr(ctx.p, "order", 1) != 1:
                raise ValueError("dual tables: key-switch-first order only (take the program variant with dual=False)")
            N = tvs.shape[1]
            for j, (lut, out, _) in enumerate(lv):
                accs = [ctx.blind_rotate(ctx.lwe_keyswitch(x[j * B + b]), tvs[lut]) for b in range(B)]
                slots[out] = np.stack([sample_extract(a, 0, modulus) for a in accs]).astype(object)
                if lut in dual:
                    slots[out + 1] = np.stack([sample_extract(a, N // BLOCK_P, modulus) for a in accs]).astype(object)
            continue
        idx = np.repeat(np.array([lut for lut, _, _ in lv], dtype=np.uint32), B)
        y, _ = ctx.pbs_batch(x, tvs, idx)
        for j, (_, out, _) in enumerate(lv):
            slots[out] = y[j * B:(j + 1) * B].astype(object)
    return [np.stack([lin(l).astype(np.uint64) for l in o], axis=1) for o in prog.outs]
