#!/usr/bin/env python
"""key-switch / blind-rotate milliseconds of small batches in throughput mode (split 0) and latency mode (split 1)"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import oracle as O
import tfhe_b200 as T

for name, params in (("S2", T.S2_SHORTINT), ("S1", T.S1_BOOLEAN)):
    ctx = O.Ctx(O.Params(n=params.n, N=params.N, k=params.k, ell=params.ell, logB=params.logB,
                         ell_ks=params.ell_ks, logB_ks=params.logB_ks, order=params.order))
    ks = O.keygen(ctx, 3, 1e-12, 1e-9)
    O.install_keys(ctx, ks)
    eng = T.Engine(params, device=0)
    eng.load_keyset(ks)
    rs = np.random.RandomState(1)
    tv = torch.from_numpy((rs.randint(0, 2**62, size=(1, params.N), dtype=np.int64))).cuda()
    eng.set_timing(True)
    for B in (1, 8, 64, 74, 148, 296, 444):
        cts = torch.from_numpy(rs.randint(0, 2**62, size=(B, params.ct_len), dtype=np.int64)).cuda()
        out = torch.empty_like(cts)
        row = []
        for mode in (0, 1, 2):
            eng.split_mode(mode)
            ts = []
            for _ in range(4):
                eng.pbs_batch_device(cts, tv, None, B, out, n_luts=1)
                ts.append(eng.last_timing())
            row.append(min(t[0] for t in ts[1:]))
            row.append(min(t[1] for t in ts[1:]))
        if B <= 74:
            eng.split_mode(2)
            for regs in (16, 8, 4):
                eng.cluster_regs(regs)
                ts = []
                for _ in range(4):
                    eng.pbs_batch_device(cts, tv, None, B, out, n_luts=1)
                    ts.append(eng.last_timing())
                print(f"{name} B={B:4d}  cluster mode, {regs:2d} coefficients per thread: br {min(t[1] for t in ts[1:]):8.3f} ms", flush=True)
            eng.cluster_regs(0)
        print(f"{name} B={B:4d}  throughput-mode ks {row[0]:7.3f} br {row[1]:8.3f} ms | latency-mode ks {row[2]:7.3f} br {row[3]:8.3f} ms"
              f" | cluster-mode ks {row[4]:7.3f} br {row[5]:8.3f} ms", flush=True)
    eng.split_mode(-1)
