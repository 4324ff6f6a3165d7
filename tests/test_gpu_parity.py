"""GPU parity tests: the CUDA path through the C ABI (include/lux_tfhe.h) against the CPU oracle on
the same seeded inputs, bit-exact (integer arithmetic mod q).  Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

import oracle as O
import tfhe_b200 as T

pytestmark = pytest.mark.gpu
Q = O.Q


def oparams(p: T.Params) -> O.Params:
    return O.Params(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ell_ks=p.ell_ks, logB_ks=p.logB_ks, order=p.order)


class Rig:
    """oracle context + keys + engine with the same keys loaded"""

    def __init__(self, params: T.Params, seed: int, sigma_glwe=0.0, sigma_lwe=0.0):
        self.p = params
        self.ctx = O.Ctx(oparams(params))
        self.ks = O.keygen(self.ctx, seed, sigma_glwe, sigma_lwe)
        O.install_keys(self.ctx, self.ks)
        self.eng = T.Engine(params)
        self.eng.load_keyset(self.ks)


@pytest.fixture(scope="module")
def rig_small0():
    return Rig(T.TEST_SMALL_O0, 11, 1e-12, 1e-9)


@pytest.fixture(scope="module")
def rig_small1():
    return Rig(T.TEST_SMALL_O1, 12, 1e-12, 1e-9)


@pytest.fixture(scope="module")
def rig_s1():
    return Rig(T.S1_BOOLEAN, 2, 4.99027217501041e-8, 0.0000117021618159313)   # main.rs:29-30


@pytest.fixture(scope="module")
def rig_s2():
    # TUniform(17)/TUniform(45) of the named tfhe-rs set expressed as Gaussian stddev / q
    return Rig(T.S2_SHORTINT, 3, 2.0**-47.8, 2.0**-19.8)


def edge_values():
    return [0, 1, 2, Q - 1, Q - 2, 2**63, 2**63 - 1, 2**63 + 1, 2**32 - 1, 2**32, Q // 2, Q // 2 + 1, 2**52, 2**53 - 1]



# ------------------------------------------------------------------------------- field arithmetic
def test_field_ops_vs_python_ints(rig_small0):
    """gl64.cuh's PTX carry chains against Python big-int arithmetic mod q (edge words + random)."""
    e = rig_small0.eng
    edge = edge_values() + [2**64 - 1, 2**64 - 2, Q, Q + 1, 2**64 - 2**32, 2**64 - 2**32 - 1, 2**33, 2**31]
    rs = np.random.RandomState(7)
    rnd = [int(x) for x in rs.randint(0, 2**63, size=4000, dtype=np.int64).astype(np.uint64) * np.uint64(2) + rs.randint(0, 2, size=4000).astype(np.uint64)]
    a = [x for x in edge for _ in edge] + rnd
    b = [y % Q for _ in edge for y in edge] + [x % Q for x in rnd[::-1]]
    a, b = a[: len(b)], b[: len(a)]
    got = e.field_probe(np.array(a, np.uint64), np.array(b, np.uint64))
    n = len(a)
    for i in range(n):
        x, y, v = a[i], b[i], a[(i + 1) % n]
        t = v * y % Q
        want = [(x + y) % Q, (x - y) % Q, x * y % Q, (x + t) % Q, (x - t) % Q, (x + v) % Q, (x - v) * y % Q]
        for r in range(7):
            assert int(got[r, i]) == want[r], (r, hex(x), hex(y), hex(v), hex(int(got[r, i])), hex(want[r]))

# ------------------------------------------------------------------------------------------- NTT
@pytest.mark.parametrize("N", [8, 16, 32, 64, 128, 256, 512, 1024, 2048])
def test_ntt_any_golden(golden, rig_small0, N):
    e = rig_small0.eng
    g, gh = golden[f"n{N}_testg"], golden[f"n{N}_testghat"]
    assert np.array_equal(e.ntt_any(g, N)[0], gh)                    # crypto/poly.rs:205
    assert np.array_equal(e.ntt_any(gh, N, inverse=True)[0], g)      # crypto/poly.rs:206
    rng = O.Rng(N)
    polys = np.stack([rng.field(N) for _ in range(5)])
    polys[0, : min(N, 14)] = edge_values()[: min(N, 14)]
    fw = e.ntt_any(polys, N)
    for i in range(5):
        assert np.array_equal(fw[i], O.ntt_forward(polys[i]))
    assert np.array_equal(e.ntt_any(fw, N, inverse=True), polys)


@pytest.mark.parametrize("N", [512, 1024, 2048, 4096])
def test_tiled_ntt_matches_oracle_and_golden(golden, N):
    eng = T.Engine(T.Params(n=8, N=N, k=1, ell=2, logB=16))
    rng = O.Rng(N + 1)
    npolys = 37   # ragged: not a multiple of the teams per CTA
    polys = np.stack([rng.field(N) for _ in range(npolys)])
    if f"n{N}_testg" in golden.files:
        polys[0] = golden[f"n{N}_testg"]
    polys[1, :14] = edge_values()
    polys[2] = Q - 1
    polys[3] = 0
    fw = eng.ntt_forward(polys)
    roots, inv, ninv = O.ntt_tables(N)
    for i in range(npolys):
        assert np.array_equal(fw[i], O.ntt_forward(polys[i], roots)), i
    if f"n{N}_testghat" in golden.files:
        assert np.array_equal(fw[0], golden[f"n{N}_testghat"])
    assert np.array_equal(eng.ntt_backward(fw), polys)
    assert eng.ntt_forward(np.zeros((0, N), np.uint64)).shape == (0, N)   # empty input


def test_ntt_linearity_large_batch():
    N = 2048
    eng = T.Engine(T.Params(n=8, N=N, k=1, ell=2, logB=16))
    rs = np.random.RandomState(5)
    a = (rs.randint(0, 2**62, size=(3000, N), dtype=np.uint64) % np.uint64(2**61))
    b = (rs.randint(0, 2**62, size=(3000, N), dtype=np.uint64) % np.uint64(2**61))
    fa, fb, fs = eng.ntt_forward(a), eng.ntt_forward(b), eng.ntt_forward(a + b)
    s = (fa.astype(object) + fb.astype(object)) % Q
    assert np.array_equal(fs.astype(object), s)
    assert np.array_equal(eng.ntt_backward(fa), a)


# ------------------------------------------------------------------------- decomposition, mod switch
@pytest.mark.parametrize("logB", [1, 3, 4, 5, 8, 16, 22, 23, 32, 41])
def test_decompose_matches_oracle(rig_small0, logB):
    rng = O.Rng(logB)
    vals = np.array(edge_values() + [rng.field() for _ in range(3000)], dtype=np.uint64)
    got = rig_small0.eng.decompose(vals, logB)
    for i, v in enumerate(vals[:400]):
        assert np.array_equal(got[i], O.decompose(int(v), logB)), (int(v), logB)
    # recomposition property on everything (glwe_poly.rs:262-279)
    L = got.shape[1]
    d = got.astype(object)
    d = np.where(d > Q // 2, d - Q, d)
    rec = sum(d[:, i] * (1 << (i * logB)) for i in range(L)) % Q
    assert np.array_equal(rec, vals.astype(object))


@pytest.mark.parametrize("N", [512, 1024, 2048])
def test_mod_switch_matches_oracle(N):
    eng = T.Engine(T.Params(n=8, N=N, k=1, ell=2, logB=16))
    rng = O.Rng(N)
    vals = np.array(edge_values() + [rng.field() for _ in range(2000)], dtype=np.uint64)
    got = eng.mod_switch(vals)
    want = [O.lib().orc_mod_switch(int(v), N) for v in vals]
    assert got.tolist() == want


# ------------------------------------------------------------------------------------ external product
def test_external_product_matches_oracle(rig_small0):
    r = rig_small0
    rng = O.Rng(8)
    N = r.p.N
    count = 5
    ggsw, glwe = [], []
    for i in range(count):
        m = np.zeros(N, np.uint64)
        m[0] = i & 1
        ggsw.append(r.ctx.ggsw_encrypt(rng, r.ks.s_glwe, m, 1e-12))
        glwe.append(np.concatenate([rng.field(N), rng.field(N)]))
    ggsw, glwe = np.stack(ggsw), np.stack(glwe)
    got = r.eng.external_product(ggsw, glwe)
    for i in range(count):
        gn = np.concatenate([O.ntt_forward(pp, r.ctx.roots) for pp in ggsw[i].reshape(-1, N)])
        assert np.array_equal(got[i], r.ctx.external_product(gn, glwe[i])), i
    # NTT-domain input (compute_bsk's output format) gives the same result
    gn_all = np.stack([np.concatenate([O.ntt_forward(pp, r.ctx.roots) for pp in g.reshape(-1, N)]) for g in ggsw])
    assert np.array_equal(r.eng.external_product(gn_all, glwe, domain=T.DOMAIN_NTT), got)


# ---------------------------------------------------------------------------------------- blind rotation
@pytest.mark.parametrize("teams", [1, 4, 12])
def test_blind_rotate_small_all_team_variants(rig_small0, teams):
    r = rig_small0
    r.eng.force_teams(teams)
    rng = O.Rng(100 + teams)
    B = 29
    lwe = np.stack([rng.field(r.p.n + 1) for _ in range(B)])
    lwe[0, 0] = Q - 1    # rotation by exactly 2N
    lwe[0, -1] = 0       # -0
    lwe[1, :] = 0
    tvs = np.stack([rng.field(r.p.N) for _ in range(3)])
    idx = np.array([i % 3 for i in range(B)], dtype=np.uint32)
    for steps in (0, 1, 2, r.p.n):
        got = r.eng.blind_rotate(lwe, tvs, idx, steps)
        for b in range(B if steps <= 2 else 6):
            assert np.array_equal(got[b], r.ctx.blind_rotate(lwe[b], tvs[idx[b]], steps)), (steps, b)
    r.eng.force_teams(0)


def test_blind_rotate_s1_first_iterations(rig_s1):
    r = rig_s1
    rng = O.Rng(5)
    lwe = np.stack([rng.field(r.p.n + 1) for _ in range(7)])
    tv = O.testv_lut(r.p.N, 2, O.delta(4), [0, 1])
    got = r.eng.blind_rotate(lwe, tv, None, 3)
    for b in range(7):
        assert np.array_equal(got[b], r.ctx.blind_rotate(lwe[b], tv, 3)), b


def test_blind_rotate_s2_first_iterations(rig_s2):
    r = rig_s2
    rng = O.Rng(6)
    lwe = np.stack([rng.field(r.p.n + 1) for _ in range(4)])
    tv = O.testv_lut(r.p.N, 16, O.delta(32), list(range(16)))
    got = r.eng.blind_rotate(lwe, tv, None, 2)
    for b in range(4):
        assert np.array_equal(got[b], r.ctx.blind_rotate(lwe[b], tv, 2)), b


# -------------------------------------------------------------------------------------------- key switch
def test_lwe_keyswitch_small(rig_small1):
    r = rig_small1
    rng = O.Rng(9)
    B = 19   # ragged against the 8-ciphertext tile
    cts = np.stack([rng.field(r.p.N + 1) for _ in range(B)])
    cts[0, :14] = edge_values()
    got = r.eng.lwe_keyswitch(cts)
    for b in range(B):
        assert np.array_equal(got[b], r.ctx.lwe_keyswitch(cts[b])), b


def test_lwe_keyswitch_s2(rig_s2):
    r = rig_s2
    rng = O.Rng(10)
    cts = np.stack([rng.field(r.p.N + 1) for _ in range(9)])
    got = r.eng.lwe_keyswitch(cts)
    for b in range(9):
        assert np.array_equal(got[b], r.ctx.lwe_keyswitch(cts[b])), b


def test_lwe_keyswitch_split_forms_agree(rig_s2, rig_small1):
    """small batches cut the input coefficients into slices (exact partial sums, second kernel): same words"""
    for r, B in ((rig_s2, 3), (rig_small1, 11)):
        rng = O.Rng(15)
        cts = np.stack([rng.field(r.p.N + 1) for _ in range(B)])
        outs = []
        try:
            r.eng.ks_mode(0)          # the integer-pipe kernel (the default is the tensor-core form)
            for n in (1, 2, 5, 0):
                r.eng.ks_splits(n)
                outs.append(r.eng.lwe_keyswitch(cts))
        finally:
            r.eng.ks_splits(0)
            r.eng.ks_mode(-1)
        for o in outs[1:]:
            assert np.array_equal(o, outs[0])
        assert np.array_equal(outs[0][0], r.ctx.lwe_keyswitch(cts[0]))


@pytest.mark.parametrize("B", [1, 129, 300])
def test_lwe_keyswitch_tensor_core_form(rig_s2, rig_small1, B):
    """int8 tensor-core contraction (ks_mma.cuh): same words as the integer-pipe kernel and the oracle, on ragged
    batches (1, 129, 300 against the 128-row tile) and edge-value masks"""
    for r in (rig_s2, rig_small1):
        rng = O.Rng(21 + B)
        cts = np.stack([rng.field(r.p.N + 1) for _ in range(B)])
        cts[0, :14] = edge_values()
        cts[B - 1, -14:] = edge_values()
        try:
            r.eng.ks_mode(1)
            got = r.eng.lwe_keyswitch(cts)
            r.eng.ks_mode(0)
            ref = r.eng.lwe_keyswitch(cts)
        finally:
            r.eng.ks_mode(-1)
        assert np.array_equal(got, ref)
        for b in sorted({0, B // 2, B - 1}):
            assert np.array_equal(got[b], r.ctx.lwe_keyswitch(cts[b])), b
        assert np.array_equal(r.eng.lwe_keyswitch(cts), got)   # automatic dispatch takes the tensor-core form


@pytest.mark.parametrize("B", [1, 129, 300])
def test_lwe_keyswitch_tcgen05_form(rig_s2, rig_small1, B):
    """tcgen05.mma.kind::i8 kernel (ks_tc5.cuh) on the Goldilocks words: equal to the mma.sync kernel, the integer-pipe kernel
    and the oracle"""
    for r in (rig_s2, rig_small1):
        rng = O.Rng(41 + B)
        cts = np.stack([rng.field(r.p.N + 1) for _ in range(B)])
        cts[0, :14] = edge_values()
        cts[B - 1, -14:] = edge_values()
        try:
            r.eng.ks_mode(2)
            got = r.eng.lwe_keyswitch(cts)
            r.eng.ks_mode(1)
            ref = r.eng.lwe_keyswitch(cts)
            r.eng.ks_mode(0)
            ref0 = r.eng.lwe_keyswitch(cts)
        finally:
            r.eng.ks_mode(-1)
        assert np.array_equal(got, ref)
        assert np.array_equal(got, ref0)
        for b in sorted({0, B // 2, B - 1}):
            assert np.array_equal(got[b], r.ctx.lwe_keyswitch(cts[b])), b


# ---------------------------------------------------------------------------------------------- whole PBS
def _pbs_case(r: Rig, B: int, p: int, luts, seed: int, sigma: float, check: int):
    rng = O.Rng(seed)
    delta = O.delta(2 * p)
    tvs = np.stack([O.testv_lut(r.p.N, p, delta, lut) for lut in luts])
    # the product's own test-vector builder agrees with the oracle's
    for i, lut in enumerate(luts):
        assert np.array_equal(T.make_testv(r.p.N, p, T.delta(2 * p), lut), tvs[i])
    msgs = [int(rng.u64() % p) for _ in range(B)]
    idx = np.array([int(rng.u64() % len(luts)) for _ in range(B)], dtype=np.uint32)
    cts = np.stack([O.lwe_encrypt(rng, r.ks.s_in, (m * delta) % Q, sigma) for m in msgs])
    got = r.eng.pbs_batch(cts, tvs, idx)
    for b in range(B):   # decryption correct for every ciphertext
        assert O.decode(O.lwe_decrypt(r.ks.s_in, got[b]), delta, 2 * p) == luts[idx[b]][msgs[b]], b
    want, _ = r.ctx.pbs_batch(cts[:check], tvs, idx[:check])
    assert np.array_equal(got[:check], want)   # bit-exact ciphertexts


def test_pbs_small_order0(rig_small0):
    _pbs_case(rig_small0, 41, 4, [[0, 1, 2, 3], [3, 2, 1, 0], [1, 1, 2, 0]], 21, 1e-9, 41)


def test_pbs_small_order1(rig_small1):
    _pbs_case(rig_small1, 41, 4, [[0, 1, 2, 3], [2, 0, 3, 1]], 22, 1e-9, 41)


def test_pbs_s1_boolean_bit_exact(rig_s1):
    """BASELINE config 1 parameters (main.rs:23-30), identity LUT on one bit + the NAND test vector."""
    nand_like = [1, 1]   # constant-1 block table stands for the sign test vector q/8
    _pbs_case(rig_s1, 12, 2, [[0, 1], nand_like], 23, 0.0000117021618159313, 4)


@pytest.mark.parametrize("which", ["small0", "small1", "s1"])
def test_boolean_gates_bit_exact_and_truth_tables(which, rig_small0, rig_small1, rig_s1):
    """BASELINE config 1: NAND (and the other two-input gates) = one sign bootstrap; the GPU gate (linear part,
    PBS and +q/8 all on the device) against the oracle's, bit for bit, and the decrypted truth tables; at the
    reference's own boolean parameter set with its noise (main.rs:23-30) and on both PBS orders."""
    r = {"small0": rig_small0, "small1": rig_small1, "s1": rig_s1}[which]
    sigma = 0.0000117021618159313 if which == "s1" else 1e-9
    delta = O.delta(4)
    rng = O.Rng(77)
    xs, ys = [0, 0, 1, 1, 1, 0], [0, 1, 0, 1, 1, 0]
    a = np.stack([O.lwe_encrypt(rng, r.ks.s_in, (x * delta) % Q, sigma) for x in xs])
    b = np.stack([O.lwe_encrypt(rng, r.ks.s_in, (y * delta) % Q, sigma) for y in ys])
    dec = lambda cts: [O.decode(O.lwe_decrypt(r.ks.s_in, c), delta, 4) for c in cts]
    for name, truth in O.GATE_TRUTH.items():
        got = r.eng.gate(name, a, b)
        assert dec(got) == [truth(x, y) for x, y in zip(xs, ys)], name
        if which != "s1" or name == "nand":
            assert np.array_equal(got, O.gate_batch(r.ctx, name, a, b)), name
    assert dec(r.eng.gate_not(a)) == [1 - x for x in xs]
    chained = r.eng.gate("nand", r.eng.gate("nand", a, b), r.eng.gate_not(b))   # outputs feed gates again
    assert dec(chained) == [1 - ((1 - (x & y)) & (1 - y)) for x, y in zip(xs, ys)]
    with pytest.raises(T.LuxError):
        r.eng._ck(T.lib().lux_gate_batch(r.eng._h, 17, a.ctypes.data, b.ctypes.data, 1, a.ctypes.data))


@pytest.mark.parametrize("which", ["small0", "small1", "s1", "s2"])
def test_latency_mode_equals_throughput_mode(which, rig_small0, rig_small1, rig_s1, rig_s2):
    """the latency-mode blind rotate (one ciphertext per CTA, 2*ell teams, partial sums in shared memory) and the
    throughput kernel produce the same words, and both equal the oracle's"""
    r = {"small0": rig_small0, "small1": rig_small1, "s1": rig_s1, "s2": rig_s2}[which]
    p = 2 if which == "s1" else 4
    delta = O.delta(2 * p)
    rng = O.Rng(91)
    luts = [[(i + 1) % p for i in range(p)], list(range(p))]
    tvs = np.stack([O.testv_lut(r.p.N, p, delta, lut) for lut in luts])
    B = 7
    msgs = [int(rng.u64() % p) for _ in range(B)]
    idx = np.array([b % 2 for b in range(B)], dtype=np.uint32)
    cts = np.stack([O.lwe_encrypt(rng, r.ks.s_in, (m * delta) % Q, 1e-9) for m in msgs])
    try:
        r.eng.split_mode(1)
        fast = r.eng.pbs_batch(cts, tvs, idx)
        r.eng.cluster_regs(16)       # the one-CTA latency kernel with the wide tile (named sets default to the narrow one)
        assert np.array_equal(r.eng.pbs_batch(cts, tvs, idx), fast)
        r.eng.cluster_regs(0)
        r.eng.split_mode(2)          # one ciphertext per 2-CTA cluster, partials exchanged through distributed shared memory
        pair = r.eng.pbs_batch(cts, tvs, idx)
        for regs in (16, 8, 4):      # wide tile and the narrow tiles (8 / 4 coefficients per thread)
            r.eng.cluster_regs(regs)
            assert np.array_equal(r.eng.pbs_batch(cts, tvs, idx), pair), regs
        r.eng.cluster_regs(0)
        r.eng.split_mode(0)
        bulk = r.eng.pbs_batch(cts, tvs, idx)
        r.eng.generic_gadget(True)   # runtime-gadget instantiations of all three kernels
        r.eng.split_mode(2)
        pair_g = r.eng.pbs_batch(cts, tvs, idx)
        r.eng.split_mode(0)
        bulk_g = r.eng.pbs_batch(cts, tvs, idx)
    finally:
        r.eng.generic_gadget(False)
        r.eng.cluster_regs(0)
        r.eng.split_mode(-1)
    assert np.array_equal(fast, bulk)
    assert np.array_equal(pair, bulk)
    assert np.array_equal(pair_g, bulk)
    assert np.array_equal(bulk_g, bulk)
    want, _ = r.ctx.pbs_batch(cts[:2], tvs, idx[:2])
    assert np.array_equal(fast[:2], want)
    assert [O.decode(O.lwe_decrypt(r.ks.s_in, c), delta, 2 * p) for c in fast] == [luts[idx[b]][msgs[b]] for b in range(B)]


@pytest.mark.parametrize("which", ["s1", "s2"])
def test_tile_widths_agree(which, rig_s1, rig_s2):
    """throughput blind rotate of the named sets: narrow tiles (8 and 4 coefficients per thread), the wide tile
    (16, teams in lock step) and the runtime-gadget instantiation give the same words; ragged against every
    ciphertexts-per-CTA count; first outputs equal the oracle's"""
    r = {"s1": rig_s1, "s2": rig_s2}[which]
    rng = O.Rng(57)
    B = 601      # above 3 * 148, so that S1 takes its four-per-CTA plan on its own; ragged against every team count
    cts = np.stack([rng.field(r.p.ct_len) for _ in range(B)])
    tvs = np.stack([rng.field(r.p.N) for _ in range(3)])
    idx = np.array([b % 3 for b in range(B)], dtype=np.uint32)
    outs = {}
    try:
        r.eng.split_mode(0)
        for regs in (8, 16, 4):
            r.eng.br_regs(regs)
            for teams in (0, 1, 2, 3, 4, 6):
                r.eng.force_teams(teams)
                outs[(regs, teams)] = r.eng.pbs_batch(cts, tvs, idx)
        r.eng.force_teams(0)
        r.eng.generic_gadget(True)
        r.eng.br_regs(16)
        outs["generic"] = r.eng.pbs_batch(cts, tvs, idx)
    finally:
        r.eng.generic_gadget(False)
        r.eng.br_regs(8)
        r.eng.force_teams(0)
        r.eng.split_mode(-1)
    base = outs[(8, 0)]
    for k, v in outs.items():
        assert np.array_equal(v, base), k
    want, _ = r.ctx.pbs_batch(cts[:2], tvs, idx[:2])
    assert np.array_equal(base[:2], want)


def test_pbs_s2_shortint_bit_exact(rig_s2):
    """BASELINE config 2 shape: 4-bit LUTs at N=2048 (identity, x^2 mod 16, carry extract, message extract)."""
    luts = [list(range(16)), [(x * x) % 16 for x in range(16)], [x >> 2 for x in range(16)], [x & 3 for x in range(16)]]
    _pbs_case(rig_s2, 10, 16, luts, 24, 2.0**-19.8, 3)


def test_pbs_device_path_matches_host_path(rig_small1):
    torch = pytest.importorskip("torch")
    r = rig_small1
    rng = O.Rng(31)
    B = 50
    cts = np.stack([rng.field(r.p.ct_len) for _ in range(B)])
    tvs = np.stack([rng.field(r.p.N) for _ in range(2)])
    idx = np.array([b & 1 for b in range(B)], dtype=np.uint32)
    host = r.eng.pbs_batch(cts, tvs, idx)
    d_in = torch.from_numpy(cts.view(np.int64)).cuda()
    d_tv = torch.from_numpy(tvs.view(np.int64)).cuda()
    d_idx = torch.from_numpy(idx.view(np.int32)).cuda()
    d_out = torch.empty_like(d_in)
    torch.cuda.synchronize()
    r.eng.pbs_batch_device(d_in, d_tv, d_idx, B, d_out)
    r.eng.synchronize()
    assert np.array_equal(d_out.cpu().numpy().view(np.uint64), host)


# ------------------------------------------------------------------------------------- boundary behaviour
def test_errors_and_edges(rig_small0):
    r = rig_small0
    e2 = T.Engine(r.p)
    cts = np.zeros((2, r.p.ct_len), np.uint64)
    tv = np.zeros((1, r.p.N), np.uint64)
    with pytest.raises(T.LuxError) as ei:
        e2.pbs_batch(cts, tv)
    assert ei.value.code == -4 and "key" in str(ei.value)          # LUX_ERR_NO_KEY
    with pytest.raises(T.LuxError):
        r.eng.pbs_batch(cts, tv, np.array([0, 5], dtype=np.uint32))    # lut index out of range
    with pytest.raises(T.LuxError):
        e2.load_bsk(np.zeros(7, np.uint64))                             # wrong size
    with pytest.raises(T.LuxError):
        T.Engine(T.Params(n=8, N=768, k=1, ell=2, logB=8))              # unsupported ring
    with pytest.raises(T.LuxError):
        T.Engine(T.Params(n=8, N=1024, k=2, ell=2, logB=8))             # unsupported k
    out = r.eng.pbs_batch(np.zeros((0, r.p.ct_len), np.uint64), tv)     # empty batch
    assert out.shape == (0, r.p.ct_len)
    before = r.eng.launch_count
    r.eng.pbs_batch(cts, tv)
    assert r.eng.launch_count == before + 1     # order 0: one launch per blind rotate (KS fused)


def test_key_blob_roundtrip(rig_small1):
    torch = pytest.importorskip("torch")
    r = rig_small1
    blob = torch.empty(r.eng.key_blob_bytes, dtype=torch.uint8, device="cuda")
    r.eng.export_keys_device(blob)
    e2 = T.Engine(r.p)
    e2.import_keys_device(blob)
    rng = O.Rng(77)
    cts = np.stack([rng.field(r.p.ct_len) for _ in range(5)])
    tv = rng.field(r.p.N)
    assert np.array_equal(e2.pbs_batch(cts, tv), r.eng.pbs_batch(cts, tv))


@pytest.mark.parametrize("which,B", [("s1", 5920 + 5), ("s2", 4440 + 3)])
def test_whole_batch_random_indices_meet_the_oracle(which, B, rig_s1, rig_s2):
    """at the real parameter sets and the bench batch sizes (plus a ragged tail) the oracle sees 16
    ciphertexts at random indices over the WHOLE batch — first and last CTA, later waves, the ragged last CTA, the key ring
    after wrap-around — not only the head; every variant-independent index arithmetic (ct, lut_idx, output offset) is
    under test here."""
    r = {"s1": rig_s1, "s2": rig_s2}[which]
    rs = np.random.RandomState(B)
    cts = (rs.randint(0, 2**63, size=(B, r.p.ct_len), dtype=np.int64).astype(np.uint64) % np.uint64(Q))
    tvs = (rs.randint(0, 2**63, size=(5, r.p.N), dtype=np.int64).astype(np.uint64) % np.uint64(Q))
    idx = rs.randint(0, 5, B).astype(np.uint32)
    got = r.eng.pbs_batch(cts, tvs, idx)
    wave = r.eng.wave_ciphertexts(B)
    last_wave = (B - 1) // wave * wave
    pick = np.unique(np.concatenate([[0, 1, wave - 1, wave, last_wave, B - 2, B - 1], rs.randint(0, B, 9)]))
    want, _ = r.ctx.pbs_batch(cts[pick], tvs, idx[pick])
    assert np.array_equal(got[pick], want), "GPU PBS differs from the oracle away from the head of the batch"
