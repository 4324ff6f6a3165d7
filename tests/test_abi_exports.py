"""The drop-in boundary is a C-ABI shared library: every function include/*.h declares must be exported by the built
libluxtfhe_b200.so (CPU-only check: symbols are looked up, nothing is called that needs a GPU)."""
import ctypes
import re
from pathlib import Path

import tfhe_b200 as T

ROOT = Path(__file__).resolve().parent.parent
DECL = re.compile(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?\b(lux_\w+)\s*\(", re.M)


def declared(header: Path):
    text = re.sub(r"/\*.*?\*/", "", header.read_text(), flags=re.S)     # comments out
    text = re.sub(r"typedef\s+(?:struct|enum)[^;{]*\{.*?\}\s*\w+\s*;", "", text, flags=re.S)
    return sorted({m.group(1) for m in DECL.finditer(text)})


def test_every_declared_function_is_exported():
    lib = ctypes.CDLL(str(T.lib_path()))
    names = []
    for h in sorted((ROOT / "include").glob("*.h")):
        got = declared(h)
        assert got, f"no declarations parsed from {h.name}"
        names += got
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/*.h but not exported: {missing}"
    assert len(names) > 70
    for must in ("lux_pbs_batch", "lux_engine_create_multi", "lux_multi_pbs_batch", "lux_multi_erc20_transfer", "lux_eval"):
        assert must in names


def test_engine_creation_fails_loudly_without_a_gpu():
    """no CPU fallback: on a box without CUDA devices creating an engine is an error, never a silent slow path"""
    import torch

    if torch.cuda.is_available():
        return
    try:
        T.Engine(T.TEST_SMALL_TORUS)
    except T.LuxError:
        pass
    else:
        raise AssertionError("engine creation succeeded without a GPU")
    try:
        T.MultiEngine(T.TEST_SMALL_TORUS, 2)
    except T.LuxError:
        pass
    else:
        raise AssertionError("multi-GPU engine creation succeeded without a GPU")
