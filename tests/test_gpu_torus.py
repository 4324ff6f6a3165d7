"""GPU parity tests of the 2^64-torus / f64-FFT arithmetic (lux_params.arith = LUX_ARITH_TORUS64_FFT): the CUDA path
through the C ABI against oracle/torus.c on the same seeded inputs.  Every comparison is BIT-EXACT: the oracle fixes the
order and rounding of every floating-point operation and the kernels reproduce it (fft64.cuh).  Against tfhe-rs itself
parity is unpinned (oracle/torus.h); the decrypt-exact / noise-variance criterion is tested at the end.
Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

import oracle as O
import tfhe_b200 as T
from oracle import torus as TO

pytestmark = pytest.mark.gpu


def oparams(p: T.Params) -> TO.Params:
    return TO.Params(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ell_ks=p.ell_ks, logB_ks=p.logB_ks, order=p.order)


class Rig:
    def __init__(self, params: T.Params, seed: int, glwe_noise=TO.S2_GLWE_NOISE, lwe_noise=TO.S2_LWE_NOISE):
        self.p = params
        self.ctx = TO.Ctx(oparams(params))
        self.ks = TO.keygen(self.ctx, seed, glwe_noise, lwe_noise)
        TO.install_keys(self.ctx, self.ks)
        self.eng = T.Engine(params)
        self.eng.load_keyset(self.ks)
        self.glwe_noise = glwe_noise


@pytest.fixture(scope="module")
def rig_small():
    return Rig(T.TEST_SMALL_TORUS, 21)


@pytest.fixture(scope="module")
def rig_small_o2():
    return Rig(T.TEST_SMALL_TORUS_O2, 22)


@pytest.fixture(scope="module")
def rig_s2():
    return Rig(T.S2_TORUS, 23)


def u64_rand(rs, shape):
    return rs.randint(0, 2**63, size=shape, dtype=np.int64).astype(np.uint64) * np.uint64(2) + rs.randint(0, 2, size=shape).astype(np.uint64)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


# ------------------------------------------------------------------------------------------- FFT
def test_fft_forward_inverse_bit_exact(rig_small):
    e, N = rig_small.eng, rig_small.p.N
    rs = np.random.RandomState(1)
    z = np.empty((9, N // 2), np.complex128)
    z[:4] = rs.randint(-2**22, 2**22, size=(4, N // 2)) + 1j * rs.randint(-2**22, 2**22, size=(4, N // 2))   # digit polynomials
    z[4:8] = (rs.standard_normal((4, N // 2)) + 1j * rs.standard_normal((4, N // 2))) * 2.0**-12            # key-like
    z[8] = 0
    z[8, 0] = 1.0
    tw = TO.twiddles(N)
    fwd = e.fft(z)
    want = np.stack([TO.fft_forward(N, row, tw) for row in z])
    assert np.array_equal(bits(fwd), bits(want)), "forward transform differs from oracle/torus.c"
    inv = e.fft(fwd, inverse=True)
    want_inv = np.stack([TO.fft_inverse(N, row, tw) for row in want])
    assert np.array_equal(bits(inv), bits(want_inv)), "inverse transform differs from oracle/torus.c"
    # and it IS a transform pair: inverse(forward(z)) = (N/2) z up to rounding
    assert np.abs(inv[:4] / (N // 2) - z[:4]).max() < 1e-6


def test_fft_is_the_negacyclic_product(rig_small):
    """pointwise product in the transform domain = product in Z[X]/(X^N+1) (the property the external product rests on)"""
    e, N = rig_small.eng, rig_small.p.N
    rs = np.random.RandomState(2)
    a = rs.randint(-50, 50, N).astype(np.int64)
    b = np.zeros(N, np.int64)
    b[[0, 3, N - 1]] = [2, -1, 5]
    fold = lambda p: (p[: N // 2] + 1j * p[N // 2:]).astype(np.complex128)
    A, Bf = e.fft(fold(a))[0], e.fft(fold(b))[0]
    c = e.fft(A * Bf, inverse=True)[0] / (N // 2)
    c = np.rint(np.concatenate([c.real, c.imag])).astype(np.int64)
    ref = np.zeros(N, np.int64)
    for j in (0, 3, N - 1):
        rot = np.concatenate([-a[N - j:], a[: N - j]]) if j else a.copy()
        ref += b[j] * rot
    assert np.array_equal(c, ref)


def test_torus_decompose_vs_oracle(rig_small):
    e, p = rig_small.eng, rig_small.p
    rs = np.random.RandomState(3)
    half = 1 << (63 - p.logB)   # rounding boundary of the closest representable
    edge = [0, 1, 2**64 - 1, 2**63, 2**63 - 1, half, half - 1, half + 1, 2**64 - half, 2**64 - half - 1,
            (1 << 63) + half, (1 << 63) - half, (1 << 63) + half - 1, 3 << 62, (1 << (64 - p.logB)) - 1]
    vals = np.concatenate([np.array(edge, np.uint64), u64_rand(rs, 4000)])
    got = e.torus_decompose(vals)
    want = np.array([TO.decompose(int(v), p.logB, 1)[0] for v in vals], np.int64)
    assert np.array_equal(got, want)
    rec = (got.astype(np.uint64) << np.uint64(64 - p.logB))
    assert all(int(r) == TO.closest_representable(int(v), p.logB, 1) for r, v in zip(rec, vals))


def test_mod_switch_vs_oracle(rig_small):
    e, p = rig_small.eng, rig_small.p
    rs = np.random.RandomState(4)
    vals = np.concatenate([np.array([0, 1, 2**64 - 1, 2**63, 2**52 - 1, 2**52, 2**51], np.uint64), u64_rand(rs, 2000)])
    got = e.mod_switch(vals)
    assert [int(g) for g in got] == [TO.mod_switch(int(v), p.N) for v in vals]


# ------------------------------------------------------------------------------------------- keys
def test_bsk_fourier_rows_bit_exact(rig_small):
    """the resident key rows are the oracle's Fourier rows (scaled, sign folded) in the ring layout [row][poly][lane]"""
    e, p = rig_small.eng, rig_small.p
    got = e.bsk_fourier()                                                  # [n][32 i][4][32 lane]
    want = rig_small.ctx.bsk_fourier().reshape(p.n, 4, 32, 32)             # [n][poly][lane r][i]: index 32 r + i
    want = np.ascontiguousarray(want.transpose(0, 3, 1, 2))
    assert np.array_equal(bits(got), bits(want))


def test_lwe_keyswitch_vs_oracle(rig_small):
    e, p = rig_small.eng, rig_small.p
    rs = np.random.RandomState(5)
    for B in (1, 3, 130):
        cts = u64_rand(rs, (B, p.N + 1))
        got = e.lwe_keyswitch(cts)
        want = np.stack([rig_small.ctx.lwe_keyswitch(c) for c in cts])
        assert np.array_equal(got, want), B


@pytest.mark.parametrize("which", ["small", "s2"])
def test_lwe_keyswitch_tcgen05_form(rig_small, rig_s2, which):
    """tcgen05.mma.kind::i8 kernel with tensor-memory accumulators (ks_tc5.cuh): the same words as the mma.sync kernel and
    the oracle, on ragged batches (against the 128-row tile, one and several tiles per CTA) and on edge-value masks"""
    r = rig_small if which == "small" else rig_s2
    e, p = r.eng, r.p
    rs = np.random.RandomState(31)
    for B in ((1, 129, 300, 1500) if which == "small" else (1, 129, 1400, 4440)):   # 1400 at the named set: 319 tiles, three per CTA on some SMs; 4440: 1,015 tiles
        cts = u64_rand(rs, (B, p.k * p.N + 1))
        cts[0, :6] = np.array([0, 1, 2**63, 2**63 - 1, 2**64 - 1, 2**59], np.uint64)
        cts[B - 1, -7:-1] = np.array([2**64 - 1, 2**63 + 1, 2**60 - 1, 2**60, 7, 0], np.uint64)
        try:
            e.ks_mode(2)
            got = e.lwe_keyswitch(cts)
            again = e.lwe_keyswitch(cts)
            e.ks_mode(1)
            ref = e.lwe_keyswitch(cts)
        finally:
            e.ks_mode(-1)
        assert np.array_equal(got, ref), B
        assert np.array_equal(again, ref), B
        for b in sorted({0, B // 2, B - 1}):
            assert np.array_equal(got[b], r.ctx.lwe_keyswitch(cts[b])), (B, b)


# ------------------------------------------------------------------------------------------- blind rotation
@pytest.mark.parametrize("steps", [0, 1, 2, 5, None])
def test_blind_rotate_per_iteration_bit_exact(rig_small, steps):
    e, p = rig_small.eng, rig_small.p
    rs = np.random.RandomState(6)
    B = 7
    lwe = u64_rand(rs, (B, p.n + 1))
    lwe[0, :] = 0                       # all shifts zero
    lwe[1, :] = 2**64 - 1               # mod switch rounds up to 2N
    dl = T.delta(32, p.arith)
    tvs = np.stack([T.make_testv(p.N, 16, dl, [(3 * x + s) % 16 for x in range(16)], p.arith) for s in range(3)])
    idx = np.array([0, 1, 2, 0, 1, 2, 0], np.uint32)
    got = e.blind_rotate(lwe, tvs, idx, steps=steps)
    want = np.stack([rig_small.ctx.blind_rotate(lwe[b], tvs[idx[b]], steps=steps) for b in range(B)])
    assert np.array_equal(got, want), f"accumulators differ after {steps} CMUX steps"


@pytest.mark.parametrize("steps", [0, 1, 3, None])
def test_latency_kernel_equals_throughput_kernel_and_oracle(rig_small, steps):
    """the latency form (fft64_lat.cuh: one ciphertext per CTA, eight warps, taken by default when B <= #SMs) against the
    warp-per-polynomial form and the oracle, accumulators after `steps` CMUX iterations, bit for bit"""
    e, p = rig_small.eng, rig_small.p
    rs = np.random.RandomState(16)
    B = 5
    lwe = u64_rand(rs, (B, p.n + 1))
    lwe[0, :] = 0
    lwe[1, :] = 2**64 - 1
    tvs = u64_rand(rs, (2, p.N))
    idx = np.array([0, 1, 1, 0, 1], np.uint32)
    lat = {}
    try:
        e.fft_lat_mode(1)
        for variant in (0, 1, 2, 3):          # accumulator in registers / late key request (fft64_lat.cuh, V)
            e.fft_variant(variant)
            lat[variant] = e.blind_rotate(lwe, tvs, idx, steps=steps)
        e.fft_variant(0)
        e.fft_lat_mode(0)
        thr = e.blind_rotate(lwe, tvs, idx, steps=steps)
    finally:
        e.fft_variant(0)
        e.fft_lat_mode(-1)
    want = np.stack([rig_small.ctx.blind_rotate(lwe[b], tvs[idx[b]], steps=steps) for b in range(B)])
    for variant, got in lat.items():
        assert np.array_equal(got, want), f"latency kernel (variant {variant}) differs from oracle/torus.c"
    assert np.array_equal(thr, want), "throughput kernel differs from oracle/torus.c"


def test_latency_kernel_whole_pbs_both_orders_and_large_batch(rig_small, rig_small_o2, rig_s2):
    """whole PBS through the latency kernel: both orders at the test set, a batch larger than the SM count forced through it
    (grid > #SMs), and the named set (single ciphertext and a ragged small batch) against the oracle"""
    for r, B in ((rig_small, 1), (rig_small, 301), (rig_small_o2, 9)):
        e, p = r.eng, r.p
        rs = np.random.RandomState(17 + B)
        cts = u64_rand(rs, (B, p.ct_len))
        tvs = u64_rand(rs, (3, p.N))
        idx = rs.randint(0, 3, B).astype(np.uint32)
        try:
            e.fft_lat_mode(1)
            got = e.pbs_batch(cts, tvs, idx)
        finally:
            e.fft_lat_mode(-1)
        pick = np.unique(np.concatenate([[0, B - 1], rs.randint(0, B, 6)]))
        want, _ = r.ctx.pbs_batch(cts[pick], tvs, idx[pick])
        assert np.array_equal(got[pick], want), (p.order, B)
    e, p = rig_s2.eng, rig_s2.p
    rs = np.random.RandomState(18)
    for B in (1, 11):
        cts = u64_rand(rs, (B, p.ct_len))
        tvs = u64_rand(rs, (2, p.N))
        idx = rs.randint(0, 2, B).astype(np.uint32)
        got = e.pbs_batch(cts, tvs, idx)            # default mode: B <= #SMs takes the latency kernel
        e.fft_lat_mode(0)
        try:
            thr = e.pbs_batch(cts, tvs, idx)
        finally:
            e.fft_lat_mode(-1)
        want, _ = rig_s2.ctx.pbs_batch(cts, tvs, idx)
        assert np.array_equal(got, want) and np.array_equal(thr, want), B


@pytest.mark.parametrize("teams", [1, 2, 3, 13, 4, 14, 24, 34, 54, 64, 74, 84, 94, 104, 114, 124, 134, 154, 164, 174, 204, 214])
def test_pbs_small_bit_exact_every_team_count(rig_small, teams):
    """whole PBS (key switch, blind rotation, extraction) against the oracle, every ciphertext of a ragged batch,
    for each ciphertexts-per-CTA configuration of the kernel (ring depth and stage size differ between them)"""
    e, p, ks = rig_small.eng, rig_small.p, rig_small.ks
    teams, variant = teams % 10, teams // 10      # tens digit: ring depth / twiddle placement variant of the kernel
    e.force_teams(teams)
    e.fft_variant(variant)
    try:
        dl = T.delta(32, p.arith)
        lut = [(x * x + 1) % 16 for x in range(16)]
        tv = T.make_testv(p.N, 16, dl, lut, p.arith)
        B = 2 * teams + 1
        msgs = np.arange(B, dtype=np.uint64) % np.uint64(16)
        cts = TO.lwe_encrypt_batch(7, ks.s_in, msgs * np.uint64(dl), rig_small.glwe_noise)
        got = e.pbs_batch(cts, tv)
        want, _ = rig_small.ctx.pbs_batch(cts, tv)
        assert np.array_equal(got, want)
        dec = [TO.decode(int(v), dl, 32) for v in TO.lwe_decrypt_batch(ks.s_in, got)]
        assert dec == [lut[int(m)] for m in msgs]
    finally:
        e.force_teams(0)
        e.fft_variant(0)


def test_pbs_order2_bit_exact(rig_small_o2):
    """classic order: blind rotate first, key switch last; ciphertexts under the small key"""
    r = rig_small_o2
    e, p, ks = r.eng, r.p, r.ks
    dl = T.delta(32, p.arith)
    lut = [15 - x for x in range(16)]
    tv = T.make_testv(p.N, 16, dl, lut, p.arith)
    msgs = np.arange(9, dtype=np.uint64)
    cts = TO.lwe_encrypt_batch(8, ks.s_in, msgs * np.uint64(dl), 30)
    assert cts.shape[1] == p.n + 1
    got = e.pbs_batch(cts, tv)
    want, _ = r.ctx.pbs_batch(cts, tv)
    assert np.array_equal(got, want)
    dec = [TO.decode(int(v), dl, 32) for v in TO.lwe_decrypt_batch(ks.s_in, got)]
    assert dec == [lut[int(m)] for m in msgs]


# ------------------------------------------------------------------------------------------- the named set
def test_pbs_s2_bit_exact_across_the_batch(rig_s2):
    """PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 with its own noise: a ragged multi-wave batch; 24 ciphertexts spread
    over the whole batch (first and last CTA, last wave, ragged tail) meet the oracle bit for bit, all of them decrypt"""
    e, p, ks = rig_s2.eng, rig_s2.p, rig_s2.ks
    B = 3 * 148 * 2 + 3 * 40 + 2            # two full waves of 3 per CTA, a partial wave and a ragged last CTA
    dl = T.delta(32, p.arith)
    luts = [[x for x in range(16)], [(x * x) % 16 for x in range(16)], [x >> 2 for x in range(16)], [x & 3 for x in range(16)]]
    tvs = np.stack([T.make_testv(p.N, 16, dl, l, p.arith) for l in luts])
    rs = np.random.RandomState(9)
    msgs = rs.randint(0, 16, B).astype(np.uint64)
    idx = rs.randint(0, 4, B).astype(np.uint32)
    cts = TO.lwe_encrypt_batch(10, ks.s_in, msgs * np.uint64(dl), TO.S2_GLWE_NOISE)
    got = e.pbs_batch(cts, tvs, idx)
    noisy = TO.lwe_decrypt_batch(ks.s_in, got)
    dec = np.array([TO.decode(int(v), dl, 32) for v in noisy])
    want_dec = np.array([luts[i][int(m)] for i, m in zip(idx, msgs)])
    assert np.array_equal(dec, want_dec), f"{int((dec != want_dec).sum())} wrong decryptions"
    pick = np.unique(np.concatenate([[0, 1, 2, 3, 443, 444, 887, 888, B - 3, B - 2, B - 1], rs.randint(0, B, 13)]))
    want, _ = rig_s2.ctx.pbs_batch(cts[pick], tvs, idx[pick])
    assert np.array_equal(got[pick], want), "ciphertext words differ from oracle/torus.c"
    err = np.array([TO.centered_diff(int(v), (int(w) * dl) % 2**64) for v, w in zip(noisy, want_dec)])
    assert err.std() < 2.0**-12, err.std()     # measured 2^-14.7; a failed transform would be ~2^-2


def test_pbs_s2_bench_batch_random_indices(rig_s2):
    """the bench batch of the default kernel (8 waves of 4 x 148) plus a ragged tail: 16 ciphertexts at random indices over
    the whole batch (every wave boundary, last CTA) bit-exact vs the oracle"""
    e, p = rig_s2.eng, rig_s2.p
    B = 4736 + 7
    rs = np.random.RandomState(4)
    cts = u64_rand(rs, (B, p.ct_len))
    tvs = u64_rand(rs, (4, p.N))
    idx = rs.randint(0, 4, B).astype(np.uint32)
    got = e.pbs_batch(cts, tvs, idx)
    wave = e.wave_ciphertexts(B)
    assert wave == 4 * 148
    pick = np.unique(np.concatenate([[0, wave - 1, wave, 7 * wave, 8 * wave, B - 1], rs.randint(0, B, 10)]))
    want, _ = rig_s2.ctx.pbs_batch(cts[pick], tvs, idx[pick])
    assert np.array_equal(got[pick], want)


def test_pbs_s2_whole_waves_plus_tail_launch(rig_s2):
    """a batch that is not a whole number of waves runs as whole waves (four ciphertexts per CTA) + a tail launch of its own
    (latency kernel up to 148 ciphertexts, two per CTA up to 296, three up to 444): same words as the single launch, and the
    ciphertexts either side of the cut, the lookup-table indices included, bit-exact vs the oracle"""
    e, p = rig_s2.eng, rig_s2.p
    rs = np.random.RandomState(44)
    wave = 4 * 148
    for tail in (1, 100, 250, 420):
        B = wave + tail
        cts = u64_rand(rs, (B, p.ct_len))
        tvs = u64_rand(rs, (4, p.N))
        idx = rs.randint(0, 4, B).astype(np.uint32)
        n0 = e.launch_count
        got = e.pbs_batch(cts, tvs, idx)
        n_split = e.launch_count - n0
        try:
            e.fft_tail_split(False)
            n0 = e.launch_count
            one = e.pbs_batch(cts, tvs, idx)
            n_one = e.launch_count - n0
        finally:
            e.fft_tail_split(True)
        assert n_split == n_one + 1, (tail, n_split, n_one)       # the tail really was a launch of its own
        assert np.array_equal(got, one), tail
        pick = np.unique([0, wave - 1, wave, B - 1, wave + tail // 2])
        want, _ = rig_s2.ctx.pbs_batch(cts[pick], tvs, idx[pick])
        assert np.array_equal(got[pick], want), tail


# ------------------------------------------------------------------------------------------- the decryption gate
@pytest.mark.parametrize("arith", ["torus", "goldilocks"])
def test_decryption_gate_65536_trials(arith):
    """north_star correctness bullet 3 in the driver-run suite: 65,536 chained bootstraps of encrypted 4-bit messages at
    the named set (4 rounds, each round's outputs are the next round's inputs), zero wrong decryptions, measured output
    noise variance within the bound p_fail <= 2^-128 needs; samples of every round bit-exact vs the oracle.
    (scripts/decrypt_trials.py runs the same gate at >= 10^6 trials: profiles/r02_decrypt_trials_*.json)"""
    import importlib.util
    from pathlib import Path

    spec = importlib.util.spec_from_file_location("decrypt_trials", Path(__file__).resolve().parent.parent / "scripts" / "decrypt_trials.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    res = mod.run(arith, trials=65536, chunk=16384, oracle_sample=6, verbose=False)
    assert res["trials"] == 65536 and res["wrong_decryptions"] == 0, res
    assert res["variance_within_bound"], res
    assert res["bit_exact_vs_oracle_checked"] >= 4 * 6


# ------------------------------------------------------------------------------------------- boundary: threads, several GPUs
def test_concurrent_calls_on_one_engine_are_serialised(rig_small):
    """SURVEY 8b "blocking and thread-safe per engine": four host threads hammer ONE engine with different batches (ctypes
    releases the GIL, so the calls really overlap); every result equals the serial one"""
    import threading

    e, p, ks = rig_small.eng, rig_small.p, rig_small.ks
    dl = T.delta(32, p.arith)
    tvs = np.stack([T.make_testv(p.N, 16, dl, [(x * k + 1) % 16 for x in range(16)], p.arith) for k in range(4)])
    jobs = []
    for t in range(4):
        msgs = (np.arange(5 + 3 * t, dtype=np.uint64) * np.uint64(t + 1)) % np.uint64(16)
        cts = TO.lwe_encrypt_batch(100 + t, ks.s_in, msgs * np.uint64(dl), rig_small.glwe_noise)
        idx = (np.arange(len(msgs)) % 4).astype(np.uint32)
        jobs.append((cts, idx, e.pbs_batch(cts, tvs, idx)))
    got = [None] * 4
    errs = []

    def work(t):
        try:
            for _ in range(6):
                got[t] = e.pbs_batch(jobs[t][0], tvs, jobs[t][1])
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)

    th = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errs, errs
    for t in range(4):
        assert np.array_equal(got[t], jobs[t][2]), f"thread {t}: concurrent result differs from the serial one"


def test_multi_gpu_handle_matches_single_engine(rig_small):
    """lux_engine_create_multi: keys converted on GPU 0 and copied to the others inside the library, the batch sliced over
    the GPUs (every visible one; a 1-GPU box still runs the whole path): same words as the single-engine call, which the
    other tests pin to the oracle.  Ragged batch so that the shares differ in size."""
    import torch

    p, ks = rig_small.p, rig_small.ks
    n = torch.cuda.device_count()
    m = T.MultiEngine(p, 0)
    assert m.n_gpus == n
    m.load_keyset(ks)
    dl = T.delta(32, p.arith)
    tvs = np.stack([T.make_testv(p.N, 16, dl, [(x + k) % 16 for x in range(16)], p.arith) for k in range(3)])
    B = 7 * n + 3
    msgs = np.arange(B, dtype=np.uint64) % np.uint64(16)
    idx = (np.arange(B) % 3).astype(np.uint32)
    cts = TO.lwe_encrypt_batch(77, ks.s_in, msgs * np.uint64(dl), rig_small.glwe_noise)
    got = m.pbs_batch(cts, tvs, idx)
    want = rig_small.eng.pbs_batch(cts, tvs, idx)
    assert np.array_equal(got, want)
    ora, _ = rig_small.ctx.pbs_batch(cts, tvs, idx)
    assert np.array_equal(got, ora)
    # euintN ops through the same handle
    import tfhe_b200.radix as R
    from oracle.radix import split_blocks
    mr, sr = R.MultiRadix(m), R.Radix(rig_small.eng)
    vals = [(37 * i + 5) & 0xFF for i in range(2 * n + 1)]
    dig = np.array([split_blocks(v, 4) for v in vals], dtype=np.uint64).reshape(-1)
    a = TO.lwe_encrypt_batch(78, ks.s_in, dig * np.uint64(dl), rig_small.glwe_noise).reshape(len(vals), 4, p.ct_len)
    assert np.array_equal(mr.eval("add", 8, a, a[::-1].copy()), sr.eval("add", 8, a, a[::-1].copy()))
    assert np.array_equal(mr.eval("lt", 8, a, a[::-1].copy()), sr.eval("lt", 8, a, a[::-1].copy()))
