#!/usr/bin/env python
"""PBS/s over every GPU of the box through the C ABI ALONE: one process, no torch.distributed, no NCCL — the path a Node /
ctypes host takes (lux_engine_create_multi + lux_multi_pbs_batch, include/lux_tfhe.h).  VERDICT r01 item 4: this number
should sit within a few percent of the torchrun bench (SCALE).

  python scripts/bench_multi_abi.py [--gpus N] [--batch-per-gpu B] [--steps K] [--c5 T]

Prints one JSON line: host-pointer throughput from pageable numpy buffers and from pinned buffers, parity of a sample of
outputs with the CPU oracle, and (--c5) ERC20 transfers/s with every balance decrypted against the plaintext replay."""
import argparse, json, sys, time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--batch-per-gpu", type=int, default=4736)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--arith", default="torus")
    ap.add_argument("--c5", type=int, default=0, help="ERC20 transfers in total (0 = skip)")
    args = ap.parse_args()
    import torch

    import bench
    import tfhe_b200 as T
    import tfhe_b200.radix as R

    side = bench.make_side(args.arith)
    ks = side.fixture()
    m = T.MultiEngine(side.params, args.gpus)
    n = m.n_gpus
    t0 = time.perf_counter()
    m.load_keyset(ks)
    t_keys = time.perf_counter() - t0
    B = args.batch_per_gpu * n
    cts, tvs, idx, pos, msgs = bench.make_inputs(side, B, 4, nreal=64)
    out = np.empty_like(cts)
    for _ in range(2):
        m.pbs_batch(cts, tvs, idx, out=out)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m.pbs_batch(cts, tvs, idx, out=out)
    pageable = B * args.steps / (time.perf_counter() - t0)
    h_in = torch.from_numpy(cts.view(np.int64)).pin_memory()
    h_out = torch.empty_like(h_in).pin_memory()
    pin, pout = h_in.numpy().view(np.uint64), h_out.numpy().view(np.uint64)
    for _ in range(2):
        m.pbs_batch(pin, tvs, idx, out=pout)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m.pbs_batch(pin, tvs, idx, out=pout)
    pinned = B * args.steps / (time.perf_counter() - t0)
    assert np.array_equal(out, pout)
    rs = np.random.RandomState(1)
    pick = np.unique(np.concatenate([[0, B - 1], rs.randint(0, B, 14), [g * args.batch_per_gpu for g in range(n)],
                                     [(g + 1) * args.batch_per_gpu - 1 for g in range(n)]]))
    want, _ = side.oracle_pbs(cts[pick], tvs, idx[pick])
    assert np.array_equal(out[pick], want), "multi-GPU PBS differs from the oracle"
    dec = side.decrypt(out[pos])
    assert all(d == bench.LUTS[idx[i]][int(mm)] for i, mm, d in zip(pos, msgs, dec))
    line = {"metric": "pbs_per_s_batched", "path": "C ABI only (lux_multi_pbs_batch), one process, host buffers", "n_gpus": n,
            "batch_per_gpu": args.batch_per_gpu, "steps": args.steps, "arith": args.arith,
            "e2e_pageable": pageable, "e2e_pinned": pinned, "key_load_and_replicate_s": t_keys,
            "parity": f"{len(pick)} outputs over all shares bit-exact vs oracle, {len(pos)} decryptions correct"}
    if args.c5:
        nt = args.c5
        rs = np.random.RandomState(77)
        f = [int(x) for x in rs.randint(0, 2 ** 62, nt)]
        t = [int(x) for x in rs.randint(0, 2 ** 62, nt)]
        a = [int(x) if i % 3 else int(x) >> 8 for i, x in enumerate(rs.randint(0, 2 ** 62, nt))]
        f[0], t[0], a[0] = 10000, 0, 1337
        f[1], t[1], a[1] = 1000, 0, 1337
        digits = np.array([[(v >> (2 * i)) & 3 for i in range(32)] for arr in (f, t, a) for v in arr], dtype=np.uint64).reshape(-1)
        enc = side.encrypt(991, digits).reshape(3, nt, 32, side.params.ct_len)
        mr = R.MultiRadix(m)
        mr.erc20_transfer(enc[0][:2 * n], enc[1][:2 * n], enc[2][:2 * n])      # warm-up
        t0 = time.perf_counter()
        nf, nt_ = mr.erc20_transfer(enc[0], enc[1], enc[2])
        dt = time.perf_counter() - t0
        dnf = side.decrypt(nf.reshape(-1, side.params.ct_len))
        dnt = side.decrypt(nt_.reshape(-1, side.params.ct_len))
        join = lambda d, i: sum(int(b) << (2 * j) for j, b in enumerate(d[i * 32:(i + 1) * 32]))
        bad = 0
        for i in range(nt):
            wf, wt = bench.erc20_replay(f[i], t[i], a[i])
            bad += int(join(dnf, i) != wf) + int(join(dnt, i) != wt)
        line["c5"] = {"transfers": nt, "seconds": dt, "transfers_per_s": nt / dt, "wrong_balances": bad,
                      "all_balances_correct": bad == 0, "reference_vectors": [join(dnf, 0), join(dnt, 0), join(dnf, 1), join(dnt, 1)],
                      "pbs_per_transfer": int(R.lib().lux_erc20_pbs_count())}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
