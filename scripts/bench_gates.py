#!/usr/bin/env python
"""BASELINE config 1 (C1): the boolean-gate (NAND) programmable bootstrap at the reference's in-tree boolean
parameter set (research/verifiable-fhe/src/main.rs:23-30: n=728, N=1024, k=1, ell=4, logB=5, with its noise).

Reports, on one B200: single-NAND latency (device-resident, CUDA events on the launching stream), batched
gates/s, the CPU oracle's single-thread latency beside it, and a chained-NAND decryption trial
(>= --trials gates, every output decrypted and compared with the plaintext circuit).

  python scripts/bench_gates.py [--trials 1048576] [--chunk 65536] [--out gpurun_out/bench_gates.json]
"""
from __future__ import annotations

import argparse
import json
import statistics
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=1 << 20)
    ap.add_argument("--chunk", type=int, default=1 << 16)
    ap.add_argument("--batch", type=int, default=0, help="throughput batch (default: 10 waves)")
    ap.add_argument("--out", default="gpurun_out/bench_gates.json")
    args = ap.parse_args()

    import torch

    import oracle as O
    import tfhe_b200 as T

    params = T.S1_BOOLEAN
    sg, sl = 4.99027217501041e-8, 0.0000117021618159313        # main.rs:29-30
    ctx = O.Ctx(O.Params(n=params.n, N=params.N, k=params.k, ell=params.ell, logB=params.logB,
                         ell_ks=params.ell_ks, logB_ks=params.logB_ks, order=params.order))
    ks = O.keygen(ctx, 2, sg, sl)
    O.install_keys(ctx, ks)
    eng = T.Engine(params, device=0)
    eng.load_keyset(ks)
    delta = O.delta(4)
    w = params.ct_len
    rs = np.random.RandomState(1)

    def encrypt(bits, seed):
        return O.lwe_encrypt_batch(seed, ks.s_in, np.asarray(bits, dtype=np.uint64) * np.uint64(delta), sl)

    def decrypt(cts):
        noisy = O.lwe_decrypt_batch(ks.s_in, cts)
        return np.array([((int(v) + delta // 2) // delta) % 4 for v in noisy], dtype=np.int64)

    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream

    # ---- single-gate latency
    a1, b1 = encrypt([1], 10), encrypt([1], 11)
    d_a = torch.from_numpy(a1.view(np.int64)).cuda()
    d_b = torch.from_numpy(b1.view(np.int64)).cuda()
    d_o = torch.empty_like(d_a)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    lats = []
    for i in range(8):
        ev0.record(stream)
        eng.gate_device("nand", d_a, d_b, 1, d_o, stream=sptr)
        ev1.record(stream)
        ev1.synchronize()
        if i >= 2:
            lats.append(ev0.elapsed_time(ev1))
    gpu_lat_ms = statistics.median(lats)
    got = d_o.cpu().numpy().view(np.uint64)
    t0 = time.perf_counter()
    want = O.gate_batch(ctx, "nand", a1, b1)
    cpu_lat_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(got, want), "GPU NAND differs from the oracle"
    assert decrypt(got)[0] == 0

    # ---- batched throughput (device resident)
    wave = eng.wave_ciphertexts(1 << 20)
    B = args.batch or 10 * wave
    bits_a, bits_b = rs.randint(0, 2, B), rs.randint(0, 2, B)
    d_a = torch.from_numpy(encrypt(bits_a, 100).view(np.int64)).cuda()
    d_b = torch.from_numpy(encrypt(bits_b, 200).view(np.int64)).cuda()
    d_o = torch.empty_like(d_a)
    for _ in range(2):
        eng.gate_device("nand", d_a, d_b, B, d_o, stream=sptr)
    torch.cuda.synchronize()
    times = []
    for _ in range(3):
        ev0.record(stream)
        eng.gate_device("nand", d_a, d_b, B, d_o, stream=sptr)
        ev1.record(stream)
        ev1.synchronize()
        times.append(ev0.elapsed_time(ev1))
    gates_per_s = B / (statistics.mean(times) * 1e-3)
    assert np.array_equal(decrypt(d_o.cpu().numpy().view(np.uint64)), 1 - (bits_a & bits_b))

    # ---- CPU oracle throughput on a bounded sample (all host threads)
    threads = O.lib().orc_max_threads()
    t0 = time.perf_counter()
    O.gate_batch(ctx, "nand", d_a[:threads].cpu().numpy().view(np.uint64), d_b[:threads].cpu().numpy().view(np.uint64))
    cpu_rate = threads / (time.perf_counter() - t0)

    # ---- chained NAND trials through the host-pointer ABI: x_{r+1} = NAND(x_r, y_r), y_{r+1} = x_r
    C = args.chunk
    rounds = -(-args.trials // C)
    xb, yb = rs.randint(0, 2, C), rs.randint(0, 2, C)
    x, y = encrypt(xb, 1000), encrypt(yb, 2000)
    wrong, total, checked = 0, 0, 0
    t_gpu = 0.0
    for r in range(rounds):
        t0 = time.perf_counter()
        z = eng.gate("nand", x, y)
        t_gpu += time.perf_counter() - t0
        zb = 1 - (xb & yb)
        dz = decrypt(z)
        wrong += int(np.count_nonzero(dz != zb))
        total += C
        if r % 4 == 0:
            assert np.array_equal(z[:4], O.gate_batch(ctx, "nand", x[:4], y[:4])), f"round {r}: not bit-exact"
            checked += 4
        print(f"round {r + 1}/{rounds}: {C} NAND gates, wrong so far {wrong}", file=sys.stderr, flush=True)
        x, y, xb, yb = z, x, dz & 1, xb
    res = {
        "workload": "C1: boolean NAND = one sign bootstrap, S1 set (n=728, N=1024, k=1, ell=4, logB=5; main.rs:23-30 noise), Goldilocks exact NTT",
        "latency_single_nand_ms": {"gpu": gpu_lat_ms, "cpu_1_thread": cpu_lat_ms},
        "gates_per_s_batched": gates_per_s, "batch": B, "wave_ciphertexts": wave,
        "cpu_gates_per_s": cpu_rate, "cpu_threads": threads,
        "trials": total, "wrong_decryptions": wrong, "bit_exact_vs_oracle_checked": checked + 1,
        "gates_per_s_through_host_abi": total / t_gpu if t_gpu else None,
    }
    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    Path(args.out).write_text(json.dumps(res, indent=1))
    print(json.dumps(res))
    if wrong:
        sys.exit(1)


if __name__ == "__main__":
    main()
