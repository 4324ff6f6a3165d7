#!/usr/bin/env python
"""BASELINE configs[2..4] (C3, C4, C5) on N GPUs of one box, through the C ABI alone (lux_engine_create_multi +
lux_multi_eval / lux_multi_select / lux_multi_erc20_transfer: the library slices every batch over the GPUs), on the named set
in its own arithmetic (2^64 torus / f64 FFT) with its noise.  EVERY output is decrypted and compared with the plaintext
operation — for C5 that is the plaintext replay of WrappingERC20.sol:86-102 over all transfers plus the two vectors of
encryptedERC20/EncryptedERC20.ts:98,114.

  python scripts/bench_configs.py [--gpus N] [--c3-batch 4096] [--transfers 10000] [--skip c3,c4,c5] > profiles/...jsonl

Host buffers in and out (pageable numpy): the times include H2D / D2H.  One JSON line per measurement."""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--c3-batch", type=int, default=4096, help="integers per GPU for the C3 ops")
    ap.add_argument("--transfers", type=int, default=10000)
    ap.add_argument("--skip", default="")
    ap.add_argument("--arith", default="torus")
    args = ap.parse_args()
    skip = set(args.skip.split(",")) if args.skip else set()

    import bench
    import tfhe_b200 as T
    import tfhe_b200.radix as R

    side = bench.make_side(args.arith)
    ks = side.fixture(21)
    m = T.MultiEngine(side.params, args.gpus)
    n = m.n_gpus
    m.load_keyset(ks)
    mr = R.MultiRadix(m)
    p = side.params
    w = p.ct_len
    rs = np.random.RandomState(7)
    seed = [1000]

    def encrypt(values, width):
        nb = R.blocks(width)
        dig = np.array([[(int(v) >> (2 * i)) & 3 for i in range(nb)] for v in values], dtype=np.uint64).reshape(-1)
        seed[0] += 1
        return side.encrypt(seed[0], dig).reshape(len(values), nb, w)

    def decrypt(cts):
        cts = np.asarray(cts)
        B, nb, _ = cts.shape
        d = side.decrypt(cts.reshape(B * nb, w))
        return [sum(int(b) << (2 * j) for j, b in enumerate(d[i * nb:(i + 1) * nb])) for i in range(B)]

    def rand(width, count):
        mask = (1 << width) - 1
        return [(int(a) * int(b)) & mask for a, b in zip(rs.randint(0, 2 ** 62, count), rs.randint(1, 2 ** 10, count))]

    def emit(config, op, width, batch, dt, pbs_per_op, wrong, extra=None):
        line = {"config": config, "op": op, "bit_width": width, "n_gpus": n, "batch": batch, "seconds": dt, "ops_per_s": batch / dt,
                "pbs_per_op": pbs_per_op, "pbs_per_s": batch * pbs_per_op / dt, "outputs_decrypted": batch, "wrong": wrong,
                "all_correct": wrong == 0, "arith": args.arith, "path": "C ABI (lux_multi_*), host buffers"}
        if extra:
            line.update(extra)
        print(json.dumps(line), flush=True)
        assert wrong == 0, line

    def timed(fn):
        fn()                                  # warm-up: builds the circuit, sizes the pools and the pinned staging
        t0 = time.perf_counter()
        out = fn()
        return out, time.perf_counter() - t0

    if "c3" not in skip:
        for width in (8, 16):
            B = args.c3_batch * n
            mask = (1 << width) - 1
            av, bv = rand(width, B), rand(width, B)
            a, b = encrypt(av, width), encrypt(bv, width)
            for op, plain in (("add", lambda x, y: (x + y) & mask), ("le", lambda x, y: int(x <= y))):
                out, dt = timed(lambda: mr.eval(op, width, a, b))
                got = decrypt(out)
                wrong = sum(int(g != plain(x, y)) for g, x, y in zip(got, av, bv))
                emit("C3", op, width, B, dt, R.pbs_count(op, width), wrong)
            cv = [int(x) for x in rs.randint(0, 2, B)]
            cond = encrypt(cv, 2).reshape(B, w)
            out, dt = timed(lambda: mr.select(width, cond, a, b))
            got = decrypt(out)
            wrong = sum(int(g != (x if c else y)) for g, c, x, y in zip(got, cv, av, bv))
            emit("C3", "select", width, B, dt, R.program(R.PROG_SELECT, width).n_pbs, wrong)

    if "c4" not in skip:
        for width in (32, 64):
            for B in (256, 1024):
                mask = (1 << width) - 1
                av, bv = rand(width, B), rand(width, B)
                a, b = encrypt(av, width), encrypt(bv, width)
                out, dt = timed(lambda: mr.eval("mul", width, a, b))
                got = decrypt(out)
                wrong = sum(int(g != (x * y) & mask) for g, x, y in zip(got, av, bv))
                emit("C4", "mul", width, B, dt, R.pbs_count("mul", width), wrong)

    if "c5" not in skip:
        nt = args.transfers
        f = [int(x) for x in rs.randint(0, 2 ** 40, nt)]          # balances U[0, 2^40), amounts U[0, 2^41)
        t = [int(x) for x in rs.randint(0, 2 ** 40, nt)]
        a = [int(x) for x in rs.randint(0, 2 ** 41, nt)]
        f[0], t[0], a[0] = 10000, 0, 1337                         # EncryptedERC20.ts:98
        f[1], t[1], a[1] = 1000, 0, 1337                          # EncryptedERC20.ts:114
        t0 = time.perf_counter()
        ef, et, ea = encrypt(f, 64), encrypt(t, 64), encrypt(a, 64)
        t_enc = time.perf_counter() - t0
        mr.erc20_transfer(ef[:4 * n], et[:4 * n], ea[:4 * n])     # warm-up on a small slice
        t0 = time.perf_counter()
        nf, nt_ = mr.erc20_transfer(ef, et, ea)
        dt = time.perf_counter() - t0
        gf, gt = decrypt(nf), decrypt(nt_)
        wrong = 0
        for i in range(nt):
            wf, wt = bench.erc20_replay(f[i], t[i], a[i])
            wrong += int(gf[i] != wf) + int(gt[i] != wt)
        moved = sum(int(x <= y) for x, y in zip(a, f))
        emit("C5", "erc20_transfer", 64, nt, dt, int(R.lib().lux_erc20_pbs_count()), wrong,
             {"balances_decrypted": 2 * nt, "transfers_that_moved_funds": moved, "encrypt_cpu_s": t_enc,
              "reference_vectors": {"EncryptedERC20.ts:98": [gf[0], gt[0]], "EncryptedERC20.ts:114": [gf[1], gt[1]]}})


if __name__ == "__main__":
    main()
