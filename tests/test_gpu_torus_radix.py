"""euintN layer on the 2^64-torus / f64-FFT arithmetic (LUX_ARITH_TORUS64_FFT) — GPU box: pytest -m gpu.

Same bar as tests/test_gpu_radix.py, on the arithmetic the named shortint set is defined on:
 * ciphertext parity: the GPU executor vs the CPU interpreter (oracle/radix.py) running the published circuit with
   oracle/torus.c's PBS and wrapping 64-bit linear algebra, bit-exact;
 * semantics at the named set WITH its noise (TUniform(17) inputs, TUniform(45)/(17) keys): add / compare / select /
   mul and the ERC20 transfer vectors of encryptedERC20/EncryptedERC20.ts:98,114 decrypt to the plaintext result."""
import numpy as np
import pytest

import oracle.torus as TO
import tfhe_b200 as T
import tfhe_b200.radix as R
from oracle.radix import interp_cipher, join_blocks, split_blocks

pytestmark = pytest.mark.gpu
MOD = 1 << 64


class Rig:
    def __init__(self, params, seed, glwe_noise=TO.NO_NOISE, lwe_noise=TO.NO_NOISE):
        self.p = params
        self.ctx = TO.Ctx(TO.Params(n=params.n, N=params.N, k=params.k, ell=params.ell, logB=params.logB,
                                    ell_ks=params.ell_ks, logB_ks=params.logB_ks, order=params.order))
        self.ks = TO.keygen(self.ctx, seed, glwe_noise, lwe_noise)
        TO.install_keys(self.ctx, self.ks)
        self.eng = T.Engine(params)
        self.eng.load_keyset(self.ks)
        self.rx = R.Radix(self.eng)
        self.noise = glwe_noise
        self.delta = T.delta(32, params.arith)
        self.seed = seed + 1000
        self.luts = R.luts()
        self.dual = R.dual_luts()

    def encrypt(self, values, width):
        nb = R.blocks(width)
        digits = np.array([split_blocks(int(v), nb) for v in values], dtype=np.uint64).reshape(-1)
        self.seed += 1
        cts = TO.lwe_encrypt_batch(self.seed, self.ks.s_in, digits * np.uint64(self.delta), self.noise)
        return cts.reshape(len(values), nb, self.p.ct_len)

    def decrypt(self, cts):
        cts = np.asarray(cts)
        B, nb, w = cts.shape
        noisy = TO.lwe_decrypt_batch(self.ks.s_in, cts.reshape(B * nb, w))
        dec = [TO.decode(int(v), self.delta, 32) for v in noisy]
        return [join_blocks(dec[i * nb:(i + 1) * nb]) for i in range(B)]

    def testv(self, lut_row):
        return TO.testv_lut(self.p.N, 16, self.delta, [int(x) for x in lut_row])


@pytest.fixture(scope="module")
def small():
    return Rig(T.TEST_SMALL_TORUS, 51)


@pytest.fixture(scope="module")
def s2():
    return Rig(T.S2_TORUS, 52, TO.S2_GLWE_NOISE, TO.S2_LWE_NOISE)


@pytest.mark.parametrize("op", ["add", "sub", "mul", "lt", "eq", "max", "xor", "shr", "neg"])
def test_torus_ciphertexts_match_cpu_interpreter(small, op):
    r, width = small, 8
    a_vals, b_vals = [71, 255, 0, 200, 13], [66, 1, 0, 57, 250]
    a, b = r.encrypt(a_vals, width), r.encrypt(b_vals, width)
    prog = R.program(op, width, shallow=R.runs_shallow(width, len(a_vals)))
    if op in R.UNARY:
        got = r.rx.eval(op, width, a)
        want = interp_cipher(r.ctx, prog, r.luts, [a], r.testv, r.delta, MOD, dual=r.dual)[0]
    else:
        got = r.rx.eval(op, width, a, b)
        want = interp_cipher(r.ctx, prog, r.luts, [a, b], r.testv, r.delta, MOD, dual=r.dual)[0]
    assert np.array_equal(got, want)


def test_torus_erc20_matches_cpu_interpreter(small):
    r = small
    f, t, a = r.encrypt([10000, 1000], 64), r.encrypt([0, 5], 64), r.encrypt([1337, 1337], 64)
    nf, nt = r.rx.erc20_transfer(f, t, a)
    want = interp_cipher(r.ctx, R.program(R.PROG_ERC20, 64, shallow=R.runs_shallow(64, 2)), r.luts, [f, t, a], r.testv, r.delta, MOD, dual=r.dual)
    assert np.array_equal(nf, want[0]) and np.array_equal(nt, want[1])
    assert r.decrypt(nf) == [8663, 1000] and r.decrypt(nt) == [1337, 5]       # EncryptedERC20.ts:98,114
    tr = r.rx.trivial_encrypt(16, [4242, 65535])
    assert r.decrypt(tr) == [4242, 65535]
    assert r.decrypt(r.rx.eval_scalar("add", 16, tr, [1, 1])) == [4243, 0]


def test_torus_s2_named_set_with_noise(s2):
    """the named set with its own noise: add is bit-exact against the interpreter; a mixed batch decrypts right"""
    r, width = s2, 8
    av, bv = [71, 200], [66, 100]
    a, b = r.encrypt(av, width), r.encrypt(bv, width)
    got = r.rx.eval("add", width, a, b)
    want = interp_cipher(r.ctx, R.program("add", width, shallow=True), r.luts, [a, b], r.testv, r.delta, MOD, dual=r.dual)[0]
    assert np.array_equal(got, want)
    assert r.decrypt(got) == [137, (200 + 100) & 255]
    assert r.decrypt(r.rx.eval("lt", width, a, b)) == [0, 0]
    assert r.decrypt(r.rx.eval("mul", width, a, b)) == [(71 * 66) & 255, (200 * 100) & 255]
    c = r.rx.eval("ge", width, a, b)
    assert r.decrypt(r.rx.select(width, c, a, b)) == [71, 200]
    rs = np.random.RandomState(3)
    xs, ys = rs.randint(0, 1 << 16, 200).tolist(), rs.randint(0, 1 << 16, 200).tolist()      # deep (ripple) variant
    x, y = r.encrypt(xs, 16), r.encrypt(ys, 16)
    assert r.decrypt(r.rx.eval("add", 16, x, y)) == [(p + q) & 0xFFFF for p, q in zip(xs, ys)]
    assert r.decrypt(r.rx.eval("le", 16, x, y)) == [int(p <= q) for p, q in zip(xs, ys)]


def test_torus_s2_erc20_reference_vectors(s2):
    r = s2
    f, t, a = r.encrypt([10000, 1000, 5], 64), r.encrypt([0, 0, (1 << 64) - 3], 64), r.encrypt([1337, 1337, 5], 64)
    nf, nt = r.rx.erc20_transfer(f, t, a)
    assert r.decrypt(nf) == [8663, 1000, 0]
    assert r.decrypt(nt) == [1337, 0, 2]
