"""LWE key switch ms per batch size: integer-pipe kernel (split-K for small batches) vs int8 tensor-core kernel"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import tfhe_b200 as T  # noqa: E402

params = T.S2_SHORTINT
eng = T.Engine(params, device=0)
O, ctx, ks = bench.make_fixture(params)
eng.load_keyset(ks)
rs = np.random.RandomState(1)
tv = torch.from_numpy(rs.randint(0, 2**62, size=(1, params.N), dtype=np.int64)).cuda()
eng.set_timing(True)
eng.split_mode(0)
for B in (1, 4, 16, 32, 64, 128, 148, 255, 256, 444, 1024):
    cts = torch.from_numpy(rs.randint(0, 2**62, size=(B, params.ct_len), dtype=np.int64)).cuda()
    out = torch.empty_like(cts)
    row = []
    for mode in (0, 1):
        eng.ks_mode(mode)
        ts = []
        for _ in range(4):
            eng.pbs_batch_device(cts, tv, None, B, out, n_luts=1)
            ts.append(eng.last_timing()[0])
        row.append(min(ts[1:]))
    print(f"B={B:5d}  integer-pipe {row[0]:7.3f} ms   tensor-core {row[1]:7.3f} ms", flush=True)
