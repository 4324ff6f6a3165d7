"""A/B of an engine knob inside ONE process on one GPU (same box, same clocks): alternates the two settings and
prints the blind-rotate time of each step.  usage: python scripts/ab_knob.py generic_gadget [--batch 4440]"""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import tfhe_b200 as T  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("knob")
ap.add_argument("--batch", type=int, default=4440)
ap.add_argument("--rounds", type=int, default=3)
ap.add_argument("--set", default="s2", choices=["s1", "s2"])
ap.add_argument("--teams", type=int, default=0, help="force ciphertexts per CTA")
ap.add_argument("--values", default="0,1", help="the two knob settings to alternate")
args = ap.parse_args()
params = T.S2_SHORTINT if args.set == "s2" else T.S1_BOOLEAN
eng = T.Engine(params, device=0)
O, ctx, ks = bench.make_fixture(params)
eng.load_keyset(ks)
if args.teams:
    eng.force_teams(args.teams)
B = args.batch
rs = np.random.RandomState(5)
cts = rs.randint(0, 2 ** 63, size=(B, params.ct_len), dtype=np.int64).astype(np.uint64) % np.uint64(T.Q)
tvs = rs.randint(0, 2 ** 63, size=(4, params.N), dtype=np.int64).astype(np.uint64) % np.uint64(T.Q)
idx = (np.arange(B) % 4).astype(np.uint32)
d_in = torch.from_numpy(cts.view(np.int64)).cuda()
d_tv = torch.from_numpy(tvs.view(np.int64)).cuda()
d_idx = torch.from_numpy(idx.view(np.int32)).cuda()
d_out = [torch.empty_like(d_in), torch.empty_like(d_in)]
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
eng.set_timing(True)
vals = [int(v) for v in args.values.split(",")]
res = {v: [] for v in vals}
d_out = {v: torch.empty_like(d_in) for v in vals}
for r in range(args.rounds + 1):
    for v in vals:
        getattr(eng, args.knob)(v)
        eng.pbs_batch_device(d_in, d_tv, d_idx, B, d_out[v], n_luts=4, stream=stream.cuda_stream)
        torch.cuda.synchronize()
        if r:
            res[v].append(eng.last_timing()[1])
same = bool(torch.equal(d_out[vals[0]], d_out[vals[1]]))
print(json.dumps({"knob": args.knob, "batch": B, "blind_rotate_ms": res, "outputs_identical": same}))
