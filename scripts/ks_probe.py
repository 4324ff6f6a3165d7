"""LWE key switch ms per batch size at the named set: the int8 tensor-core kernels (mma.sync form, tcgen05 / tensor-memory
form) and, on the Goldilocks arithmetic, the integer-pipe kernel.  Random key words: the time is data independent; the
outputs of the forms are compared word for word.   python scripts/ks_probe.py [torus|goldilocks]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import tfhe_b200 as T  # noqa: E402

torus = len(sys.argv) < 2 or sys.argv[1] == "torus"
p = T.S2_TORUS if torus else T.S2_SHORTINT
eng = T.Engine(p)
rs = np.random.RandomState(1)
if torus:
    r64 = lambda n: rs.randint(0, 2**63, size=n, dtype=np.int64).astype(np.uint64) * np.uint64(2) + rs.randint(0, 2, size=n).astype(np.uint64)
else:
    r64 = lambda n: rs.randint(0, 2**62, size=n, dtype=np.int64).astype(np.uint64)
eng.load_bsk(r64(p.n * p.ggsw_len))
eng.load_lwe_ksk(r64(p.lwe_ksk_len))
tv = torch.from_numpy(r64(p.N).view(np.int64)).cuda()
eng.set_timing(True)
modes = {0: "integer-pipe", 1: "mma.sync", 2: "tcgen05"}
use = (1, 2) if torus else (0, 1, 2)
for B in (1, 16, 128, 148, 256, 444, 1024, 2048, 4440, 4736, 8140):
    cts = torch.from_numpy(r64(B * p.ct_len).view(np.int64)).cuda().reshape(B, p.ct_len)
    row, outs = [], []
    for mode in use:
        eng.ks_mode(mode)
        out = torch.empty_like(cts)
        ts = []
        for _ in range(5):
            eng.pbs_batch_device(cts, tv, None, B, out, n_luts=1)
            ts.append(eng.last_timing()[0])
        row.append(min(ts[1:]))
        outs.append(out.cpu())
    same = all(torch.equal(outs[0], o) for o in outs[1:])
    print(f"B={B:5d}  " + "   ".join(f"{modes[m]} {t:7.3f} ms" for m, t in zip(use, row)) + f"   outputs equal: {same}", flush=True)
eng.ks_mode(-1)
