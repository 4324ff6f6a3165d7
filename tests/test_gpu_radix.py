"""GPU parity tests of the euintN layer (include/lux_radix.h) — run on the B200 box: pytest -m gpu.

 * ciphertext parity: the GPU executor against the CPU interpreter (oracle/radix.py) running the SAME
   published circuit with the oracle's PBS, bit-exact, on seeded inputs;
 * semantics: every fhEVM known-answer test (tests/golden/fhevm_kats.json) encrypted, evaluated on the
   GPU and decrypted; the ERC20 vectors of encryptedERC20/EncryptedERC20.ts;
 * the S2 (N=2048) parameter shape: bit-exact add, decrypt-correct compare/select/ERC20."""
import json
from pathlib import Path

import numpy as np
import pytest

import oracle as O
import tfhe_b200 as T
import tfhe_b200.radix as R
from oracle.radix import interp_cipher, join_blocks, split_blocks

pytestmark = pytest.mark.gpu
Q = O.Q
KATS = json.loads((Path(__file__).parent / "golden" / "fhevm_kats.json").read_text())
# a small ring with enough gadget precision for 4-bit blocks + padding bit
SMALL = T.Params(n=32, N=512, k=1, ell=2, logB=16, ell_ks=8, logB_ks=4, order=T.ORDER_LWEKS_PBS)
SMALL0 = T.Params(n=32, N=512, k=1, ell=4, logB=8, ell_ks=4, logB_ks=8, order=T.ORDER_PBS_GLWEKS)


class Rig:
    def __init__(self, params, seed, sg, sl):
        self.p = params
        self.ctx = O.Ctx(O.Params(n=params.n, N=params.N, k=params.k, ell=params.ell, logB=params.logB,
                                  ell_ks=params.ell_ks, logB_ks=params.logB_ks, order=params.order))
        self.ks = O.keygen(self.ctx, seed, sg, sl)
        O.install_keys(self.ctx, self.ks)
        self.eng = T.Engine(params)
        self.eng.load_keyset(self.ks)
        self.rx = R.Radix(self.eng)
        self.sigma = sl
        self.delta = O.delta(32)
        self.rng = O.Rng(seed + 1000)
        self.luts = R.luts()
        self.dual = R.dual_luts()

    def encrypt(self, values, width):
        nb = R.blocks(width)
        out = np.empty((len(values), nb, self.p.ct_len), np.uint64)
        for b, v in enumerate(values):
            for i, d in enumerate(split_blocks(int(v), nb)):
                out[b, i] = O.lwe_encrypt(self.rng, self.ks.s_in, (d * self.delta) % Q, self.sigma)
        return out

    def decrypt_blocks(self, cts):
        cts = np.asarray(cts)
        return [[O.decode(O.lwe_decrypt(self.ks.s_in, c), self.delta, 32) for c in row] for row in cts]

    def decrypt(self, cts):
        return [join_blocks(r) for r in self.decrypt_blocks(cts)]

    def testv(self, lut_row):
        return O.testv_lut(self.p.N, 16, self.delta, [int(x) for x in lut_row])


@pytest.fixture(scope="module")
def small():
    return Rig(SMALL, 41, 1e-13, 1e-11)


@pytest.fixture(scope="module")
def small0():
    return Rig(SMALL0, 42, 1e-13, 1e-11)


@pytest.fixture(scope="module")
def s2():
    return Rig(T.S2_SHORTINT, 3, 2.0**-47.8, 2.0**-19.8)


def width_of(t):
    return int(t.lstrip("eu").replace("int", ""))


# ------------------------------------------------------------------------------ ciphertext parity
@pytest.mark.parametrize("op", ["add", "sub", "mul", "lt", "eq", "ne", "min", "and", "shl", "rotr", "neg", "not"])
def test_ciphertexts_match_cpu_interpreter(small, op):
    r, width = small, 8
    a_vals, b_vals = [71, 255, 0, 200, 13], [66, 1, 0, 57, 250]
    a, b = r.encrypt(a_vals, width), r.encrypt(b_vals, width)
    prog = R.program(op, width)
    if op in R.UNARY:
        got = r.rx.eval(op, width, a)
        want = interp_cipher(r.ctx, prog, r.luts, [a], r.testv, r.delta, dual=r.dual)[0]
    else:
        got = r.rx.eval(op, width, a, b)
        want = interp_cipher(r.ctx, prog, r.luts, [a, b], r.testv, r.delta, dual=r.dual)[0]
    assert np.array_equal(got, want)


def test_ciphertexts_match_order0(small0):
    """the in-tree order (PBS then GLWE key switch, ciphertexts under the small key)"""
    r, width = small0, 8
    a, b = r.encrypt([9, 250, 77], width), r.encrypt([8, 17, 77], width)
    for op in ("add", "le"):
        got = r.rx.eval(op, width, a, b)
        prog = R.program(op, width, dual=False)   # engines of this order run the variant without dual tables (pick_variant)
        want = interp_cipher(r.ctx, prog, r.luts, [a, b], r.testv, r.delta, dual=r.dual)[0]
        assert np.array_equal(got, want), op
    assert r.decrypt(r.rx.eval("add", width, a, b)) == [17, 11, 154]


def test_select_and_erc20_match_cpu_interpreter(small):
    r = small
    f, t, a = r.encrypt([10000, 1000], 64), r.encrypt([0, 5], 64), r.encrypt([1337, 1337], 64)
    nf, nt = r.rx.erc20_transfer(f, t, a)
    assert R.runs_shallow(64, 2)      # two transfers: the library runs the parallel-prefix-carry variant (8 levels)
    want = interp_cipher(r.ctx, R.program(R.PROG_ERC20, 64, shallow=True), r.luts, [f, t, a], r.testv, r.delta, dual=r.dual)
    assert np.array_equal(nf, want[0]) and np.array_equal(nt, want[1])
    assert r.decrypt(nf) == [8663, 1000] and r.decrypt(nt) == [1337, 5]       # EncryptedERC20.ts:98,114 / :136-183
    cond = r.rx.eval("le", 64, a, f)
    sel = r.rx.select(64, cond, a, t)
    assert r.decrypt(sel) == [1337, 5]


@pytest.mark.parametrize("op,width", [("add", 16), ("sub", 64), ("neg", 32), ("add", 128)])
def test_shallow_and_deep_variants(small, op, width):
    """tiny batches run the shallow (parallel-prefix carry) circuit, large ones the ripple circuit: each bit-exact
    against the CPU interpreter of the same circuit, both decrypting to the plaintext result"""
    r = small
    m = (1 << width) - 1
    rnd = np.random.RandomState(width)
    few = [m, 1, int("0011" * (width // 4), 2)][:2 if width > 64 else 3]
    many = [int(rnd.randint(0, 2**31)) * int(rnd.randint(0, 2**31)) & m for _ in range(300 // R.blocks(width) + 2)]
    for vals in (few, many):
        others = [(v * 7 + 1) & m if i else int("1101" * (width // 4), 2) for i, v in enumerate(vals)]
        a, b = r.encrypt(vals, width), r.encrypt(others, width)
        shallow = R.runs_shallow(width, len(vals))
        assert shallow == (vals is few)
        prog = R.program(op, width, shallow=shallow)
        if op == "neg":
            got = r.rx.eval(op, width, a)
            want = interp_cipher(r.ctx, prog, r.luts, [a[:2]], r.testv, r.delta, dual=r.dual)[0]
            plain = [(-v) & m for v in vals]
        else:
            got = r.rx.eval(op, width, a, b)
            want = interp_cipher(r.ctx, prog, r.luts, [a[:2], b[:2]], r.testv, r.delta, dual=r.dual)[0]
            plain = [((v + o) if op == "add" else (v - o)) & m for v, o in zip(vals, others)]
        assert np.array_equal(got[:2], want)
        assert r.decrypt(got) == plain


def test_div_rem_by_plaintext_match_cpu_interpreter(small):
    """div / rem by a plaintext divisor (fhevmOperations8.ts overloads): per-divisor circuits, mixed divisors in
    one call (grouped internally), bit-exact against the CPU interpreter of the same circuits"""
    r, width = small, 8
    vals, divs = [208, 176, 180, 255, 7, 99], [76, 180, 180, 1, 8, 76]
    a = r.encrypt(vals, width)
    for op in ("div", "rem"):
        got = r.rx.eval_scalar(op, width, a, divs)
        for i, d in enumerate(divs):
            want = interp_cipher(r.ctx, R.program(op, width, scalar=d), r.luts, [a[i:i + 1]], r.testv, r.delta, dual=r.dual)[0]
            assert np.array_equal(got[i:i + 1], want), (op, i)
        assert r.decrypt(got) == [v // d if op == "div" else v % d for v, d in zip(vals, divs)]
    with pytest.raises(T.LuxError):
        r.rx.eval_scalar("div", width, a, [1, 2, 0, 4, 5, 6])     # zero divisor: the reference reverts
    with pytest.raises(T.LuxError):
        r.rx.eval("div", width, a, a)                            # encrypted divisors are not in the reference's op set
    # wide plaintext divisor (euint128 overload)
    m = (1 << 128) - 1
    x, d = [340282366920938463463367979625242002479 & m, 12345678901234567890123456789], [340282366920938463463370379842289591609 & m, 98765432109876543]
    got = r.rx.eval_scalar("rem", 128, r.encrypt(x, 128), d)
    assert r.decrypt(got) == [xi % di for xi, di in zip(x, d)]


def test_chunked_execution_is_identical(small):
    r, width = small, 16
    rs = np.random.RandomState(5)
    av, bv = rs.randint(0, 1 << 16, 23).tolist(), rs.randint(0, 1 << 16, 23).tolist()
    a, b = r.encrypt(av, width), r.encrypt(bv, width)
    whole = r.rx.eval("mul", width, a, b)
    per_int = (R.pbs_count("mul", width) + 40) * r.p.ct_len * 8
    r.rx.set_pool_bytes(per_int * 5)                    # forces 5 chunks of <= 5 integers
    chunked = r.rx.eval("mul", width, a, b)
    r.rx.set_pool_bytes(r.p.N * 16 * 3)                 # one integer per chunk, accumulator buffer of three nodes: the levels
    runs = r.rx.eval("mul", width, a[:4], b[:4])        # with dual tables run as several node runs
    r.rx.set_pool_bytes(16 << 30)
    assert np.array_equal(whole, chunked) and np.array_equal(whole[:4], runs)
    assert r.decrypt(whole) == [(x * y) & 0xFFFF for x, y in zip(av, bv)]


def test_device_entry_matches_host_entry(small):
    torch = pytest.importorskip("torch")
    r, width = small, 16
    a, b = r.encrypt([1, 65535, 4242], width), r.encrypt([65535, 65535, 1111], width)
    host = r.rx.eval("sub", width, a, b)
    da, db = torch.from_numpy(a.view(np.int64)).cuda(), torch.from_numpy(b.view(np.int64)).cuda()
    do = torch.empty_like(da)
    torch.cuda.synchronize()
    rc = R.lib().lux_eval_device(r.eng._h, R.OPS["sub"], width, da.data_ptr(), db.data_ptr(), 3, do.data_ptr())
    assert rc == 0
    r.eng.synchronize()
    assert np.array_equal(do.cpu().numpy().view(np.uint64), host)


# ------------------------------------------------------------------------------------- semantics
def _kat_groups(max_width):
    groups = {}
    for k in KATS["ops"]:
        if k["op"] not in R.OPS:
            continue
        width = max(width_of(t) for t in k["types"])
        if k["op"] in ("shl", "shr", "rotl", "rotr"):
            width = width_of(k["types"][0])
        if width > max_width:
            continue
        enc = tuple(t.startswith("e") for t in k["types"])
        groups.setdefault((k["op"], width, enc), []).append(k)
    return groups


@pytest.mark.parametrize("key", sorted(_kat_groups(64)), ids=lambda k: f"{k[0]}-{k[1]}-{''.join('e' if e else 'p' for e in k[2])}")
def test_fhevm_known_answers_encrypted(small, key):
    """encrypt -> GPU op -> decrypt == the reference suite's expected value, for every KAT up to 64 bits.
    Mixed-width overloads go through lux_cast, `(euintN, uintN)` overloads through lux_eval_scalar,
    `(uintN, euintN)` through lux_trivial_encrypt."""
    r = small
    op, width, enc = key
    cases = _kat_groups(64)[key]
    exp = [int(k["expected"]) for k in cases]

    def operand(j):
        vals = [int(k["args"][j]) for k in cases]
        ws = {width_of(k["types"][j]) for k in cases}
        out = np.empty((len(cases), R.blocks(width), r.p.ct_len), np.uint64)
        for wj in ws:   # encrypt at the operand's own width, then cast up
            sel = [i for i, k in enumerate(cases) if width_of(k["types"][j]) == wj]
            ct = r.encrypt([vals[i] for i in sel], wj) if enc[j] else r.rx.trivial_encrypt(wj, [vals[i] for i in sel])
            out[sel] = r.rx.cast(wj, width, ct) if wj != width else ct
        return out, vals

    a, av = operand(0)
    if len(enc) == 1:
        got = r.rx.eval(op, width, a)
    elif enc[1]:
        b, _ = operand(1)
        got = r.rx.eval(op, width, a, b)
    else:
        got = r.rx.eval_scalar(op, width, a, [int(k["args"][1]) for k in cases])
    assert r.decrypt(got) == exp, (key, av)


def test_wide_types_encrypted(small):
    r = small
    rs = np.random.RandomState(11)
    for width in (128, 256):
        vals = [int.from_bytes(rs.bytes(width // 8), "little") for _ in range(4)]
        a, b = r.encrypt(vals[:2], width), r.encrypt(vals[2:], width)
        m = (1 << width) - 1
        assert r.decrypt(r.rx.eval("add", width, a, b)) == [(vals[0] + vals[2]) & m, (vals[1] + vals[3]) & m]
        assert r.decrypt(r.rx.eval("gt", width, a, b)) == [int(vals[0] > vals[2]), int(vals[1] > vals[3])]
        assert r.decrypt(r.rx.eval("xor", width, a, b)) == [vals[0] ^ vals[2], vals[1] ^ vals[3]]


def test_radix_errors(small):
    r = small
    a = r.encrypt([1], 8)
    with pytest.raises(T.LuxError):
        r.rx.eval("add", 7, a, a)                       # odd width
    with pytest.raises(T.LuxError):
        R.Radix(T.Engine(SMALL)).eval("add", 8, a, a)   # keys not loaded
    assert r.rx.eval("add", 8, np.zeros((0, 4, r.p.ct_len), np.uint64), np.zeros((0, 4, r.p.ct_len), np.uint64)).shape[0] == 0
    with pytest.raises(T.LuxError):
        R.Radix(r.eng)._ck(R.lib().lux_eval(r.eng._h, 99, 8, a.ctypes.data, a.ctypes.data, 1, a.ctypes.data))


# --------------------------------------------------------------------------------------- S2 shape
def test_s2_add8_bit_exact_and_ops_decrypt(s2):
    r, width = s2, 8
    av, bv = [71, 200], [66, 100]
    a, b = r.encrypt(av, width), r.encrypt(bv, width)
    got = r.rx.eval("add", width, a, b)
    want = interp_cipher(r.ctx, R.program("add", width), r.luts, [a, b], r.testv, r.delta, dual=r.dual)[0]
    assert np.array_equal(got, want)
    assert r.decrypt(got) == [137, (200 + 100) & 255]             # fhevmOperations1.ts:137-150 + wrap
    assert r.decrypt(r.rx.eval("lt", width, a, b)) == [0, 0]
    assert r.decrypt(r.rx.eval("mul", width, a, b)) == [(71 * 66) & 255, (200 * 100) & 255]
    c = r.rx.eval("ge", width, a, b)
    assert r.decrypt(r.rx.select(width, c, a, b)) == [71, 200]


def test_s2_erc20_reference_vectors(s2):
    r = s2
    f, t, a = r.encrypt([10000, 1000, 5], 64), r.encrypt([0, 0, (1 << 64) - 3], 64), r.encrypt([1337, 1337, 5], 64)
    nf, nt = r.rx.erc20_transfer(f, t, a)
    assert r.decrypt(nf) == [8663, 1000, 0]
    assert r.decrypt(nt) == [1337, 0, 2]
