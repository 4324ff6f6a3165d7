#!/usr/bin/env python
"""ops/s of the euintN layer at the S2 shape on one B200 (BASELINE configs C3-C5; SURVEY 8d "also report ops/s
for C3-C5 and PBS/op").  Not the headline bench (bench.py is): these are the circuits' PBS counts turned into
measured operation rates, device-resident operands, wall clock around a synchronised batch.

  python scripts/bench_radix.py [--out gpurun_out/bench_radix.json] [--quick]

Operands are real encryptions (oracle keys, test infrastructure) of random integers; a sample of every result
is decrypted and compared with the plaintext operation.
"""
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
    ap.add_argument("--out", default="gpurun_out/bench_radix.json")
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()

    import torch

    import oracle as O
    import tfhe_b200 as T
    import tfhe_b200.radix as R
    from oracle.radix import join_blocks, split_blocks

    params = T.S2_SHORTINT
    ctx = O.Ctx(O.Params(n=params.n, N=params.N, k=params.k, ell=params.ell, logB=params.logB,
                         ell_ks=params.ell_ks, logB_ks=params.logB_ks, order=params.order))
    ks = O.keygen(ctx, 21, 2.0 ** -47.8, 2.0 ** -19.8)
    O.install_keys(ctx, ks)
    eng = T.Engine(params, device=0)
    eng.load_keyset(ks)
    rx = R.Radix(eng)
    L = R.lib()
    delta = O.delta(32)
    w = params.ct_len
    rs = np.random.RandomState(7)

    def encrypt(values, width, seed):
        nb = R.blocks(width)
        digs = np.array([split_blocks(int(v), nb) for v in values], dtype=np.uint64).reshape(-1)
        cts = O.lwe_encrypt_batch(seed, ks.s_in, digs * np.uint64(delta), 2.0 ** -47.8)
        return cts.reshape(len(values), nb, w)

    def decrypt(cts):
        cts = np.asarray(cts)
        B, nb, _ = cts.shape
        noisy = O.lwe_decrypt_batch(ks.s_in, cts.reshape(B * nb, w))
        d = np.array([((int(v) + delta // 2) // delta) % 32 for v in noisy]).reshape(B, nb)
        return [join_blocks(row) for row in d]

    def dev(a):
        return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()

    results = []

    def run(name, width, batch, fn_plain, launch, n_in, pbs_per_op, depth, out_blocks):
        m = (1 << width) - 1
        vals = [[int(x) & m for x in rs.randint(0, 2 ** 62, batch).astype(object) * rs.randint(1, 2 ** 10, batch).astype(object)]
                for _ in range(n_in)]
        if name.startswith("erc20"):   # balances U[0,2^40), amounts U[0,2^41)
            vals = [[int(x) for x in rs.randint(0, 2 ** 40, batch)], [int(x) for x in rs.randint(0, 2 ** 40, batch)],
                    [int(x) for x in rs.randint(0, 2 ** 41, batch)]]
        d_in = [dev(encrypt(v, width, 1000 * (i + 1))) for i, v in enumerate(vals)]
        d_out = [torch.empty((batch, ob, w), dtype=torch.int64, device="cuda") for ob in out_blocks]
        torch.cuda.synchronize()
        launch(d_in, d_out, batch)           # warm-up (builds the circuit, grows the pools)
        eng.synchronize()
        t0 = time.perf_counter()
        launch(d_in, d_out, batch)
        eng.synchronize()
        dt = time.perf_counter() - t0
        ns = min(batch, 16)
        got = [decrypt(o[:ns].cpu().numpy().view(np.uint64)) for o in d_out]
        want = [fn_plain(*[v[i] for v in vals]) for i in range(ns)]
        ok = all(tuple(g[i] for g in got) == (want[i] if isinstance(want[i], tuple) else (want[i],)) for i in range(ns))
        res = {"op": name, "bit_width": width, "batch": batch, "seconds": dt, "ops_per_s": batch / dt, "pbs_per_op": pbs_per_op,
               "levels": depth, "pbs_per_s": batch * pbs_per_op / dt, "sample_decryptions_correct": bool(ok), "sample": ns}
        print(json.dumps(res), flush=True)
        results.append(res)
        assert ok, (name, got, want)

    def binop(op):
        def launch(d_in, d_out, batch):
            rc = L.lux_eval_device(eng._h, R.OPS[op], launch.width, d_in[0].data_ptr(), d_in[1].data_ptr(), batch, d_out[0].data_ptr())
            eng._ck(rc)
        return launch

    def bench_binop(op, width, batch, plain):
        la = binop(op)
        la.width = width
        ob = 1 if op in R.PREDICATES else R.blocks(width)
        prog = R.program(op, width, shallow=R.runs_shallow(width, batch))   # the variant the library runs for this batch
        run(op, width, batch, plain, la, 2, prog.n_pbs, len(prog.levels), [ob])

    scale = 4 if args.quick else 1
    for width in (8, 16):
        m = (1 << width) - 1
        bench_binop("add", width, 4096 // scale, lambda a, b, m=m: (a + b) & m)
        bench_binop("le", width, 4096 // scale, lambda a, b: int(a <= b))

        def sel_launch(d_in, d_out, batch, width=width):
            rc = L.lux_select_device(eng._h, width, d_in[0].data_ptr(), d_in[1].data_ptr(), d_in[2].data_ptr(), batch, d_out[0].data_ptr())
            eng._ck(rc)
        # select(cond, a, b): cond is an ebool = first block of an encrypted 0/1 value
        B = 4096 // scale
        cond_vals = [int(x) for x in rs.randint(0, 2, B)]
        a_vals = [int(x) & m for x in rs.randint(0, 2 ** 30, B)]
        b_vals = [int(x) & m for x in rs.randint(0, 2 ** 30, B)]
        d_c = dev(encrypt(cond_vals, 2, 77).reshape(B, w))
        d_a, d_b = dev(encrypt(a_vals, width, 78)), dev(encrypt(b_vals, width, 79))
        d_o = torch.empty_like(d_a)
        torch.cuda.synchronize()
        sel_launch([d_c, d_a, d_b], [d_o], B)
        eng.synchronize()
        t0 = time.perf_counter()
        sel_launch([d_c, d_a, d_b], [d_o], B)
        eng.synchronize()
        dt = time.perf_counter() - t0
        got = decrypt(d_o[:16].cpu().numpy().view(np.uint64))
        ok = got == [a if c else b for c, a, b in zip(cond_vals[:16], a_vals[:16], b_vals[:16])]
        prog = R.program(R.PROG_SELECT, width)
        res = {"op": "select", "bit_width": width, "batch": B, "seconds": dt, "ops_per_s": B / dt, "pbs_per_op": prog.n_pbs,
               "levels": len(prog.levels), "pbs_per_s": B * prog.n_pbs / dt, "sample_decryptions_correct": bool(ok), "sample": 16}
        print(json.dumps(res), flush=True)
        results.append(res)
        assert ok

    for width, batch in ((32, 256), (64, 64)):
        m = (1 << width) - 1
        bench_binop("mul", width, max(8, batch // scale), lambda a, b, m=m: (a * b) & m)

    def erc_launch(d_in, d_out, batch):
        rc = L.lux_erc20_transfer_device(eng._h, d_in[0].data_ptr(), d_in[1].data_ptr(), d_in[2].data_ptr(), batch,
                                         d_out[0].data_ptr(), d_out[1].data_ptr())
        eng._ck(rc)

    m64 = (1 << 64) - 1
    run("erc20_transfer", 64, 1024 // scale, lambda f, t, a: ((f - a) & m64, (t + a) & m64) if a <= f else (f, t), erc_launch, 3,
        int(L.lux_erc20_pbs_count()), int(L.lux_erc20_depth()), [32, 32])

    # single-operation latency: a batch of ONE integer (levels of a few ciphertexts take the latency-mode kernels:
    # two SMs per ciphertext, narrow tiles); `seconds` is the latency of the operation
    bench_binop("add", 8, 1, lambda a, b: (a + b) & 0xFF)
    bench_binop("add", 64, 1, lambda a, b: (a + b) & m64)
    bench_binop("lt", 64, 1, lambda a, b: int(a < b))
    erc1 = R.program(R.PROG_ERC20, 64, shallow=True)
    run("erc20_transfer", 64, 1, lambda f, t, a: ((f - a) & m64, (t + a) & m64) if a <= f else (f, t), erc_launch, 3,
        erc1.n_pbs, len(erc1.levels), [32, 32])

    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    Path(args.out).write_text(json.dumps({"workload": "S2 shape (n=918, N=2048), one B200, device-resident operands",
                                          "results": results}, indent=1))


if __name__ == "__main__":
    main()
