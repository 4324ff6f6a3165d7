#!/usr/bin/env python3
"""Extract the reference's NTT golden vectors into a compact fixture.

Source: /root/reference/research/verifiable-fhe/src/ntt/params_{8..2048}.rs
(constants NINV, ROOTS, INVROOTS, TESTG, TESTGHAT; consumed by the reference's
own tests crypto/poly.rs:194-208 and src/ntt/mod.rs:82-136).

Run in the build container (the reference tree does not exist on the GPU box):
    python tests/golden/make_ntt_golden.py
writes tests/golden/ntt_golden.npz with arrays  n{N}_{ninv,roots,invroots,testg,testghat}.
"""
import re
import sys
from pathlib import Path

import numpy as np

REF = Path("/root/reference/research/verifiable-fhe/src/ntt")
OUT = Path(__file__).resolve().parent / "ntt_golden.npz"


def parse(path: Path):
    txt = path.read_text()
    out = {}
    m = re.search(r"pub const NINV: u64 = (\d+);", txt)
    out["ninv"] = np.array([int(m.group(1))], dtype=np.uint64)
    for name in ("ROOTS", "INVROOTS", "TESTG", "TESTGHAT"):
        m = re.search(r"pub const %s: \[u64; (\d+)\] = \[([^\]]*)\];" % name, txt)
        n = int(m.group(1))
        vals = [int(v) for v in m.group(2).split(",")]
        assert len(vals) == n, (path, name, len(vals), n)
        out[name.lower()] = np.array(vals, dtype=np.uint64)
    return out


def main():
    if not REF.is_dir():
        sys.exit("reference tree not present; fixture can only be regenerated in the build container")
    arrays = {}
    for n in (8, 16, 32, 64, 128, 256, 512, 1024, 2048):
        for k, v in parse(REF / f"params_{n}.rs").items():
            arrays[f"n{n}_{k}"] = v
    np.savez_compressed(OUT, **arrays)
    print("wrote", OUT, OUT.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
