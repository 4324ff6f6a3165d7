"""CPU tests of the euintN circuits (no GPU): the programs the library publishes through
lux_radix_program() are interpreted on plaintext integers and checked against the reference's fhEVM
known-answer tests (tests/golden/fhevm_kats.json, harvested from
/root/reference/tests/fhevm-suite/e2e/test/fhevmOperations/*.ts by tests/golden/make_fhevm_kats.py),
the ERC20 vectors (encryptedERC20/EncryptedERC20.ts:64-183) and wrap-around semantics mod 2^width."""
import json
import random
from pathlib import Path

import pytest

import tfhe_b200.radix as R
from oracle.radix import interp_plain as _interp_plain, join_blocks, split_blocks

KATS = json.loads((Path(__file__).parent / "golden" / "fhevm_kats.json").read_text())
LUTS = R.luts()
SIGNED = R.signed_luts()   # tables fed with x in (-16, 16): lux_radix_lut_signed
DUAL = R.dual_luts()       # tables whose nodes write two slots from one bootstrap: lux_radix_lut_dual


def interp_plain(prog, luts, operands):
    return _interp_plain(prog, luts, operands, SIGNED, DUAL)
SUPPORTED = set(R.OPS)


def width_of(t: str) -> int:
    return int(t.lstrip("eu").replace("int", ""))


def run_plain(op, width, *operands):
    if op in R.SCALAR_ONLY:      # plaintext divisor: divisor-specific circuit, one encrypted operand
        prog = R.program(op, width, scalar=operands[1])
        operands = operands[:1]
    else:
        prog = R.program(op, width)
    nb = R.blocks(width)
    ins = []
    for k, v in enumerate(operands):
        ins.append(split_blocks(v, prog.in_blocks[k]))
    outs = interp_plain(prog, LUTS, ins)
    assert all(0 <= b < 4 for o in outs for b in o), "outputs must be clean blocks"
    return [join_blocks(o) for o in outs], nb


def plain_semantics(op, a, b, width):
    m = (1 << width) - 1
    s = (b % width) if b is not None else 0
    return {
        "add": lambda: (a + b) & m, "sub": lambda: (a - b) & m, "mul": lambda: (a * b) & m, "neg": lambda: (-a) & m,
        "eq": lambda: int(a == b), "ne": lambda: int(a != b), "lt": lambda: int(a < b), "le": lambda: int(a <= b),
        "gt": lambda: int(a > b), "ge": lambda: int(a >= b), "min": lambda: min(a, b), "max": lambda: max(a, b),
        "and": lambda: a & b, "or": lambda: a | b, "xor": lambda: a ^ b, "not": lambda: (~a) & m,
        "shl": lambda: (a << s) & m, "shr": lambda: a >> s,
        "rotl": lambda: ((a << s) | (a >> (width - s))) & m if s else a,
        "rotr": lambda: ((a >> s) | (a << (width - s))) & m if s else a,
        "div": lambda: a // b, "rem": lambda: a % b,
    }[op]()


def test_abi_symbols_present():
    L = R.lib()
    for name in R.ABI:
        assert hasattr(L, name), name
    assert R.blocks(64) == 32 and R.blocks(7) == 0 and R.blocks(258) == 0


def test_kat_harvest_is_complete():
    assert len(KATS["ops"]) == 2394
    ops = {k["op"] for k in KATS["ops"]}
    assert ops <= SUPPORTED                       # every op of the reference's known-answer suite has a circuit


@pytest.mark.parametrize("op", sorted(SUPPORTED))
def test_fhevm_known_answers(op):
    """every known-answer test of the reference for this op, through the circuit (plaintext interpretation)"""
    cases = [k for k in KATS["ops"] if k["op"] == op]
    assert cases
    for k in cases:
        types = k["types"]
        # the suite's overloads mix widths and plaintext scalars: operands are cast to the wider type
        width = max(width_of(t) for t in types)
        if op in ("shl", "shr", "rotl", "rotr"):
            width = width_of(types[0])
        args = [int(a) for a in k["args"]]
        (got,), _ = run_plain(op, width, *args)
        exp = k["expected"]
        exp = int(exp) if not isinstance(exp, bool) else int(exp)
        assert got == exp, (k, got)
        assert got == plain_semantics(op, args[0], args[1] if len(args) > 1 else None, width), k


@pytest.mark.parametrize("width", [4, 8, 16, 32, 64])
def test_wraparound_and_random(width):
    """overflow / underflow wrap mod 2^width (not pinned by the reference's KATs) + random fuzz"""
    m = (1 << width) - 1
    rnd = random.Random(width)
    edge = [0, 1, 2, 3, m, m - 1, m >> 1, (m >> 1) + 1]
    pairs = [(a, b) for a in edge for b in edge] + [(rnd.randrange(m + 1), rnd.randrange(m + 1)) for _ in range(40)]
    ops = sorted(SUPPORTED - {"mul"}) if width == 64 else sorted(SUPPORTED)
    for op in ops:
        for a, b in pairs[: (24 if width >= 32 else len(pairs))]:
            if op in R.UNARY:
                (got,), _ = run_plain(op, width, a)
                assert got == plain_semantics(op, a, None, width), (op, a)
            elif op in R.SCALAR_ONLY and b == 0:
                with pytest.raises(R.LuxError):      # the reference reverts on a zero plaintext divisor
                    R.program(op, width, scalar=0)
            else:
                (got,), _ = run_plain(op, width, a, b)
                assert got == plain_semantics(op, a, b, width), (op, a, b)


def test_mul64_edges():
    m = (1 << 64) - 1
    for a, b in [(m, m), (m, 2), (0x123456789ABCDEF0, 0x0FEDCBA987654321), (1 << 63, 2), (3, 5)]:
        (got,), _ = run_plain("mul", 64, a, b)
        assert got == (a * b) & m


def test_wide_types():
    rnd = random.Random(9)
    for width in (128, 256):
        m = (1 << width) - 1
        for op in ("add", "sub", "lt", "eq", "ne", "xor", "min", "shl", "rotr"):
            a, b = rnd.randrange(m + 1), rnd.randrange(m + 1)
            (got,), _ = run_plain(op, width, a, b)
            assert got == plain_semantics(op, a, b, width), (op, width)


def test_select_program():
    prog = R.program(R.PROG_SELECT, 16)
    for c in (0, 1):
        out = interp_plain(prog, LUTS, [[c], split_blocks(0xBEEF, 8), split_blocks(0x1234, 8)])
        assert join_blocks(out[0]) == (0xBEEF if c else 0x1234)


def test_erc20_vectors():
    """EncryptedERC20.ts: mint 10000, transfer 1337 => 8663 / 1337; mint 1000, transfer 1337 => unchanged"""
    prog = R.program(R.PROG_ERC20, 64)
    rnd = random.Random(3)
    cases = [(v["mint"], 0, v["transfer"], v["from_after"], v["to_after"]) for v in KATS["erc20"]]
    for _ in range(30):   # balances U[0,2^40), amounts U[0,2^41)
        f, t, a = rnd.randrange(1 << 40), rnd.randrange(1 << 40), rnd.randrange(1 << 41)
        cases.append((f, t, a, f - a if a <= f else f, t + a if a <= f else t))
    cases.append(((1 << 64) - 1, (1 << 64) - 1, 1, (1 << 64) - 2, 0))   # recipient wraps mod 2^64
    for f, t, a, nf, nt in cases:
        out = interp_plain(prog, LUTS, [split_blocks(f, 32), split_blocks(t, 32), split_blocks(a, 32)])
        assert (join_blocks(out[0]), join_blocks(out[1])) == (nf, nt), (f, t, a)


@pytest.mark.parametrize("width", [16, 32, 64, 128, 256])
def test_shallow_variants_match_plain_semantics(width):
    """the parallel-prefix-carry variants (LUX_PROG_SHALLOW, run for tiny batches): same results as the plaintext
    semantics on the fhEVM known answers of add / sub / neg plus carry-chain edge cases, clean outputs, every PBS input
    below the padding bit (interp_plain raises otherwise), and log depth"""
    m = (1 << width) - 1
    rnd = random.Random(width)
    kat = [(k["op"], int(k["args"][0]), int(k["args"][1]) if len(k["args"]) > 1 else None) for k in KATS["ops"]
           if k["op"] in ("add", "sub", "neg") and width_of(k["types"][0]) == width
           and all(width_of(t) == width for t in k["types"])]
    edge = [(m, 1), (m, m), (0, 0), (0, 1), (1, m), (m >> 1, 1), (int("01" * (width // 2), 2), int("10" * (width // 2), 2) + 1),
            (int("0011" * (width // 4), 2), int("1100" * (width // 4), 2)), (int("0011" * (width // 4), 2), int("1101" * (width // 4), 2))]
    edge += [(rnd.randrange(m + 1), rnd.randrange(m + 1)) for _ in range(40)]
    for op in ("add", "sub", "neg"):
        prog, ref = R.program(op, width, shallow=True), R.program(op, width)
        assert len(prog.levels) <= max(4, (width // 2 - 1).bit_length() + 2) and len(prog.levels) < len(ref.levels)
        cases = edge + [(a, b if b is not None else 0) for o, a, b in kat if o == op]
        for a, b in cases:
            ins = [split_blocks(a, width // 2)] + ([] if op == "neg" else [split_blocks(b, width // 2)])
            out = interp_plain(prog, LUTS, ins)
            assert all(0 <= d < 4 for d in out[0])
            assert join_blocks(out[0]) == plain_semantics(op, a, b, width), (op, hex(a), hex(b))


@pytest.mark.parametrize("op", ["lt", "le", "gt", "ge", "min", "max", "eq", "ne"])
def test_ordering_on_block_pairs(op):
    """the comparison reads TWO blocks per bootstrap through the signed table (difference in [-15, 15], padding bit = sign):
    exhaustive at 2 / 4 / 6 bits (one block, one pair, a pair and a lone block), random and adjacent values at 10 / 64 / 256
    bits, for both circuit variants"""
    assert SIGNED and all(LUTS[i][0] == 0 for i in SIGNED)      # a signed table is odd: T(0) = 0
    rnd = random.Random(11)
    for width in (2, 4, 6, 10, 64, 256):
        nb = width // 2
        if width <= 6:
            cases = [(a, b) for a in range(1 << width) for b in range(1 << width)]
        else:
            cases = [(rnd.randrange(1 << width), rnd.randrange(1 << width)) for _ in range(60)]
            for _ in range(60):   # equal, adjacent, and differing in one block only
                a = rnd.randrange(1, (1 << width) - 1)
                k = rnd.randrange(nb)
                cases += [(a, a), (a, a + 1), (a + 1, a), (a, a ^ (rnd.randrange(1, 4) << (2 * k)))]
        for shallow in (False, True):
            prog = R.program(op, width, shallow=shallow)
            for a, b in cases:
                out = interp_plain(prog, LUTS, [split_blocks(a, nb), split_blocks(b, nb)])
                assert join_blocks(out[0]) == plain_semantics(op, a, b, width), (op, width, a, b)


@pytest.mark.parametrize("op", ["add", "sub", "neg", "mul", "le", "shl", "shr", "rotl", "rotr"])
def test_dual_and_plain_variants_agree(op):
    """message and carry of a column sum <= 7 from ONE bootstrap (dual table, input doubled, second output one box further)
    against the variant with a bootstrap each: same plaintext results, fewer bootstraps, every dual read at an even input"""
    assert DUAL
    rnd = random.Random(12)
    for width in (8, 16, 64):
        nb = width // 2
        for shallow in (False, True):
            d, n = R.program(op, width, shallow=shallow), R.program(op, width, shallow=shallow, dual=False)
            assert d.n_pbs <= n.n_pbs and not any(lut in DUAL for lv in n.levels for lut, _, _ in lv)
            if op in ("add", "sub", "neg") and not shallow:
                assert d.n_pbs == nb and n.n_pbs == 2 * nb - 1
            for _ in range(25):
                a, b = rnd.randrange(1 << width), rnd.randrange(1 << width)
                ins = [split_blocks(a, nb)] + ([] if op == "neg" else [split_blocks(b, nb)])
                got = join_blocks(interp_plain(d, LUTS, ins)[0])
                assert got == join_blocks(interp_plain(n, LUTS, ins)[0]) == plain_semantics(op, a, b, width), (op, width, a, b)


def test_shallow_erc20_vectors():
    prog, ref = R.program(R.PROG_ERC20, 64, shallow=True), R.program(R.PROG_ERC20, 64)
    assert len(prog.levels) <= 10 < len(ref.levels)
    rnd = random.Random(4)
    cases = [(v["mint"], 0, v["transfer"], v["from_after"], v["to_after"]) for v in KATS["erc20"]]
    for _ in range(40):
        f, t, a = rnd.randrange(1 << 40), rnd.randrange(1 << 40), rnd.randrange(1 << 41)
        cases.append((f, t, a, f - a if a <= f else f, t + a if a <= f else t))
    cases.append(((1 << 64) - 1, (1 << 64) - 1, 1, (1 << 64) - 2, 0))
    cases.append((5, (1 << 64) - 1, 5, 0, 4))
    for f, t, a, nf, nt in cases:
        out = interp_plain(prog, LUTS, [split_blocks(f, 32), split_blocks(t, 32), split_blocks(a, 32)])
        assert (join_blocks(out[0]), join_blocks(out[1])) == (nf, nt), (f, t, a)


def test_pbs_counts_and_depths():
    """circuit sizes reported to bench.py agree with the programs; sanity bounds on the designs"""
    for op in ("add", "sub", "lt", "eq", "and", "mul"):
        for width in (8, 16, 32, 64):
            p = R.program(op, width)
            assert p.n_pbs == R.pbs_count(op, width)
            assert len(p.levels) == R.depth(op, width)
    nb = 32
    assert R.pbs_count("add", 64) == nb                  # ripple: message AND carry of a block from one bootstrap (dual table)
    assert R.program("add", 64, dual=False).n_pbs == 2 * nb - 1   # engines of the other orders: carry + message per block
    assert R.pbs_count("lt", 64) == 16 + 5 + 2 + 1       # signs of block PAIRS (signed table), three flags per reduction step
    assert R.pbs_count("and", 64) == nb
    assert R.depth("lt", 64) == 4
    assert int(R.lib().lux_erc20_pbs_count()) == R.program(R.PROG_ERC20, 64).n_pbs
    assert R.pbs_count("mul", 64) < 2600


def test_every_pbs_input_stays_below_the_padding_bit():
    """interp_plain raises if any bootstrap input leaves [0,16); drive the circuits with worst-case blocks"""
    m = (1 << 32) - 1
    for op in sorted(SUPPORTED - R.UNARY):
        for a, b in [(m, m), (m, 0), (0, m), (0x55555555, 0xAAAAAAAA)]:
            if op in R.SCALAR_ONLY and b == 0:
                continue
            run_plain(op, 32, a, b)


@pytest.mark.parametrize("width", [8, 16, 32, 64, 128, 256])
def test_div_rem_by_plaintext(width):
    """x / d and x % d for plaintext divisors: powers of two (shift path), 1, divisors beyond the type's range
    clipped by the caller, odd/even general divisors (reciprocal path), worst-case dividends"""
    m = (1 << width) - 1
    rnd = random.Random(100 + width)
    divisors = [1, 2, 3, 4, 5, 6, 7, 10, 255, 256, m, m - 1, (m >> 1) + 1, (m >> 1), 1 << (width - 1), (1 << (width // 2)) + 1]
    divisors += [rnd.randrange(1, m + 1) for _ in range(6)] + [rnd.randrange(1, 1 << (width // 3 + 1)) for _ in range(6)]
    n = 0
    for d in divisors:
        if not 1 <= d <= m:
            continue
        xs = [0, 1, d - 1, d, d + 1, m, m - 1, (m // d) * d, (m // d) * d - 1, rnd.randrange(m + 1), rnd.randrange(m + 1)]
        for x in xs[: (6 if width >= 128 else len(xs))]:
            x &= m
            (q,), _ = run_plain("div", width, x, d)
            (r,), _ = run_plain("rem", width, x, d)
            assert (q, r) == (x // d, x % d), (width, x, d, q, r)
            n += 1
    assert n > 50


def test_div_exhaustive_small():
    """every (x, d) pair of a 6-bit integer"""
    for d in range(1, 64):
        for x in range(64):
            (q,), _ = run_plain("div", 6, x, d)
            (r,), _ = run_plain("rem", 6, x, d)
            assert (q, r) == (x // d, x % d), (x, d)


def test_shallow_rule_and_depths():
    """the rule lux_eval* applies by itself (radix.cu pick_variant: batch x blocks <= 296) as mirrored by
    runs_shallow, and what the shallow variants buy: depth log2(blocks) + 2 instead of `blocks`"""
    assert R.runs_shallow(64, 9) and not R.runs_shallow(64, 10)      # 288 / 320 blocks
    assert R.runs_shallow(8, 74) and not R.runs_shallow(8, 75)        # 296 / 300 blocks
    for width in (16, 32, 64):
        nb = R.blocks(width)
        deep, shallow = R.program("add", width), R.program("add", width, shallow=True)
        assert len(deep.levels) == nb
        assert len(shallow.levels) == nb.bit_length() - 1 + 2
        assert shallow.n_pbs > deep.n_pbs
    assert len(R.program(R.PROG_ERC20, 64, shallow=True).levels) == 8
    assert len(R.program(R.PROG_ERC20, 64).levels) == 4 + 1 + 32   # le, the one select, then the two chains side by side
