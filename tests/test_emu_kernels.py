"""Runs the CUDA kernel bodies (tfhe_b200/csrc/tfhe_kernels.cuh) on host threads against the
oracle.  Covers layouts, swizzles, twiddle indexing, barrier placement, the CMUX data flow, the
folded BSK format and the offset-digit key switch.  No GPU needed; PTX arithmetic is covered by
the -m gpu tests."""
import random

import numpy as np
import pytest

import oracle as O
from tests import emu as E

Q = O.Q


def test_scalar_helpers_match_oracle():
    rnd = random.Random(0)
    L = E.lib()
    vals = [0, 1, Q - 1, 2**63, 2**63 - 1, 2**32 - 1, 2**32, Q // 2] + [rnd.randrange(Q) for _ in range(200)]
    for a in vals:
        for b in vals[:30]:
            assert L.emu_mul(a, b) == a * b % Q
            assert L.emu_add(a, b) % Q == (a + b) % Q       # lazy result
            assert L.emu_sub(a, b) % Q == (a - b) % Q
        for logB in (4, 5, 8, 16, 23, 41):
            nl = (64 + logB - 1) // logB
            want = O.decompose(a, logB)
            for li in range(nl):
                assert L.emu_digit(a, li, logB, int(nl * logB == 64)) == int(want[li])
        for logN in (9, 10, 11):
            assert L.emu_mod_switch(a, logN) == O.lib().orc_mod_switch(a, 1 << logN)


@pytest.mark.parametrize("logn,teams,nblocks,npolys", [(9, 2, 2, 5), (10, 2, 1, 3), (11, 1, 1, 2), (11, 2, 1, 2), (12, 1, 1, 1)])
def test_emu_ntt_matches_oracle_and_golden(golden, logn, teams, nblocks, npolys):
    N = 1 << logn
    roots, inv, ninv = O.ntt_tables(N)
    rng = O.Rng(logn)
    polys = np.stack([rng.field(N) for _ in range(npolys)])
    if f"n{N}_testg" in golden.files:
        polys[0] = golden[f"n{N}_testg"]
    out = np.zeros_like(polys)
    assert E.lib().emu_ntt(logn, teams, nblocks, polys.reshape(-1), out.reshape(-1), roots, 1, npolys, 0) == 0
    for i in range(npolys):
        assert np.array_equal(out[i], O.ntt_forward(polys[i], roots))
    if f"n{N}_testghat" in golden.files:
        assert np.array_equal(out[0], golden[f"n{N}_testghat"])
    back = np.zeros_like(polys)
    # the kernels keep ROOTS only: INVROOTS[m+i] = -ROOTS[2m-1-i] is absorbed by the inverse butterfly
    assert E.lib().emu_ntt(logn, teams, nblocks, out.reshape(-1), back.reshape(-1), roots, ninv, npolys, 1) == 0
    assert np.array_equal(back, polys)


@pytest.fixture(params=[False, True], ids=["optimistic", "exact-path"])
def slow_path(request):
    """narrow-tile passes run optimistic butterflies and fall back to an exact out-of-line pass when a canonicalisation
    they skipped would have mattered; `exact-path` forces that fallback on every pass"""
    E.lib().emu_force_slow(int(request.param))
    yield request.param
    E.lib().emu_force_slow(0)


def test_emu_narrow_tile_handles_non_canonical_intermediates(slow_path):
    """inputs chosen so that intermediate sums land in [q, 2^64): all coefficients q-1 and 2^64-1 style lazy words"""
    logn, r = 9, 8
    N = 1 << logn
    roots, inv, ninv = O.ntt_tables(N)
    polys = np.stack([np.full(N, Q - 1, np.uint64), np.arange(N, dtype=np.uint64) * np.uint64(0xFFFFFFFF) % np.uint64(Q),
                      np.where(np.arange(N) % 2 == 0, Q - 1, 1).astype(np.uint64)])
    out = np.zeros_like(polys)
    assert E.lib().emu_ntt_r(logn, r, polys.reshape(-1), out.reshape(-1), roots, 1, 3, 0) == 0
    for i in range(3):
        assert np.array_equal(out[i], O.ntt_forward(polys[i], roots))
    back = np.zeros_like(polys)
    assert E.lib().emu_ntt_r(logn, r, out.reshape(-1), back.reshape(-1), roots, ninv, 3, 1) == 0
    assert np.array_equal(back, polys)


@pytest.mark.parametrize("logn,r", [(9, 8), (9, 4), (10, 8), (10, 4), (11, 8), (11, 4), (12, 8)])
def test_emu_narrow_tile_ntt_matches_oracle_and_golden(golden, logn, r, slow_path):
    """the R = 4 / 8 register tiles of the latency modes: same in-place positions as the reference's transform"""
    N = 1 << logn
    roots, inv, ninv = O.ntt_tables(N)
    rng = O.Rng(100 + logn + r)
    polys = np.stack([rng.field(N) for _ in range(2)])
    if f"n{N}_testg" in golden.files:
        polys[0] = golden[f"n{N}_testg"]
    out = np.zeros_like(polys)
    assert E.lib().emu_ntt_r(logn, r, polys.reshape(-1), out.reshape(-1), roots, 1, 2, 0) == 0
    for i in range(2):
        assert np.array_equal(out[i], O.ntt_forward(polys[i], roots))
    if f"n{N}_testghat" in golden.files:
        assert np.array_equal(out[0], golden[f"n{N}_testghat"])
    back = np.zeros_like(polys)
    assert E.lib().emu_ntt_r(logn, r, out.reshape(-1), back.reshape(-1), roots, ninv, 2, 1) == 0
    assert np.array_equal(back, polys)


def _device_bsk(ctx, ggsw_std, count):
    """std-domain GGSWs -> the engine's device format, via the same fold the GPU applies."""
    p = ctx.p
    ntt = np.concatenate([O.ntt_forward(pp, ctx.roots) for pp in ggsw_std.reshape(-1, p.N)])
    assert E.lib().emu_ggsw_to_key(ntt, count, p.ell, p.N, ctx.ninv) == 0
    return ntt


# teams < 0: the latency-mode body (one ciphertext per CTA, G = -teams = 2*ell teams)
@pytest.mark.parametrize("logn,teams,ell,logB,B,steps", [(9, 3, 2, 16, 4, 3), (10, 2, 4, 5, 2, 2), (9, 3, 1, 41, 3, 2),
                                                         (9, -4, 2, 16, 3, 3), (9, -2, 1, 41, 3, 2), (10, -8, 4, 5, 2, 2),
                                                         # teams >= 100: narrow tiles, 100 * R + TEAMS
                                                         (9, 803, 2, 16, 4, 3), (10, 802, 4, 5, 3, 2), (9, 403, 1, 41, 4, 2)])
def test_emu_blind_rotate_matches_oracle(logn, teams, ell, logB, B, steps):
    N = 1 << logn
    n = 5
    P = O.Params(n=n, N=N, k=1, ell=ell, logB=logB, ell_ks=ell, logB_ks=logB)
    ctx = O.Ctx(P)
    ks = O.keygen(ctx, 3, 1e-12, 1e-9)
    O.install_keys(ctx, ks)
    rng = O.Rng(77)
    lwe = np.stack([rng.field(n + 1) for _ in range(B)])
    lwe[0, 0] = Q - 1          # mod-switch to exactly 2N
    lwe[0, n] = 0              # -0 = 0
    tvs = np.stack([rng.field(N) for _ in range(2)])
    idx = np.array([i % 2 for i in range(B)], dtype=np.uint32)
    bsk_dev = _device_bsk(ctx, ks.bsk, n)
    ksk_dev = _device_bsk(ctx, ks.glwe_ksk, 1)
    roots, inv, _ = O.ntt_tables(N)

    def run(steps, with_ksk, out_mode, n_out=0):
        out_len = {0: N + 1, 1: n_out + 1, 2: 2 * N}[out_mode]
        out = np.zeros((B, out_len), np.uint64)
        a = E.BRArgs(n=n, steps=steps, ell=ell, logB=logB, L=(64 + logB - 1) // logB, B=B, out_mode=out_mode,
                     n_out=n_out, bsk=E.ptr(bsk_dev), ksk=E.ptr(ksk_dev) if with_ksk else None, lwe=E.ptr(lwe),
                     testv=E.ptr(tvs), lut_idx=E.ptr(idx), lut_div=1, out=E.ptr(out), tw_fwd=E.ptr(roots))
        if teams < 0:
            assert E.lib().emu_blind_rotate_split(logn, -teams, a) == 0
        else:
            assert E.lib().emu_blind_rotate(logn, teams, a) == 0
        return out

    raw = run(steps, False, 2)
    for b in range(B):
        assert np.array_equal(raw[b], ctx.blind_rotate(lwe[b], tvs[idx[b]], steps)), f"ct {b}"
    # full PBS in the in-tree order: all n steps, GLWE key switch, partial extract
    full = run(n, True, 1, n_out=n)
    for b in range(B):
        assert np.array_equal(full[b], ctx.pbs(lwe[b], tvs[idx[b]])), f"ct {b}"
    # full extract without key switch
    ext = run(n, False, 0)
    for b in range(B):
        acc = ctx.blind_rotate(lwe[b], tvs[idx[b]])
        want = np.zeros(N + 1, np.uint64)
        O.lib().orc_sample_extract(N, 1, acc, want)
        assert np.array_equal(ext[b], want)


@pytest.mark.parametrize("logB,ell,n,kN,B", [(4, 4, 37, 64, 5), (3, 5, 40, 64, 9), (8, 8, 33, 32, 4), (16, 2, 20, 32, 3)])
def test_emu_keyswitch_matches_oracle(logB, ell, n, kN, B):
    P = O.Params(n=n, N=kN, k=1, ell=1, logB=16, ell_ks=ell, logB_ks=logB, order=O.ORDER_LWEKS_PBS)
    if kN & (kN - 1):
        # oracle ctx needs a power-of-two N for its tables; emulate odd sizes by zero-padding the key
        pytest.skip("non power-of-two input dimension not representable in the oracle context")
    ctx = O.Ctx(P)
    rng = O.Rng(logB * 100 + ell)
    s_big = O.glwe_keygen(rng, kN, 1)
    s_small = O.lwe_keygen(rng, n)
    ksk = ctx.lwe_ksk_gen(rng, s_big, s_small, 1e-9)
    ctx.set_lwe_ksk(ksk)
    cts = np.stack([rng.field(kN + 1) for _ in range(B)])
    cts[0, :4] = [0, Q - 1, 2**63, 2**63 - 1]
    ksum = np.zeros(n + 1, np.uint64)
    E.lib().emu_ksk_colsum(ksk, ksum, kN * ell, n + 1)
    out = np.zeros((B, n + 1), np.uint64)
    a = E.KSArgs(n=n, kN=kN, ell=ell, logB=logB, L=(64 + logB - 1) // logB, B=B, ksk=E.ptr(ksk), ksum=E.ptr(ksum),
                 in_=E.ptr(cts), out=E.ptr(out))
    assert E.lib().emu_keyswitch(a, 4, 32, 16) == 0
    for b in range(B):
        assert np.array_equal(out[b], ctx.lwe_keyswitch(cts[b])), f"ct {b}"
    # split form (small batches): input coefficients cut into slices, exact partial sums added afterwards;
    # 3 slices of a 64-coefficient input = ranges of 32, 32, 0 (a trailing empty slice must contribute zero)
    for splits in (2, 3):
        out2 = np.zeros((B, n + 1), np.uint64)
        partial = np.zeros((splits, B, n + 1), np.uint64)
        a2 = E.KSArgs(n=n, kN=kN, ell=ell, logB=logB, L=(64 + logB - 1) // logB, B=B, ksk=E.ptr(ksk), ksum=E.ptr(ksum),
                      in_=E.ptr(cts), out=E.ptr(out2), splits=splits, partial=E.ptr(partial))
        assert E.lib().emu_keyswitch(a2, 4, 32, 16) == 0
        assert np.array_equal(out2, out), splits
