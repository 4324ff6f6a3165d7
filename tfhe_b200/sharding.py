"""Batch sharding over the GPUs of one box: one process per GPU, keys replicated once, no collective on
the data path.  Independent ciphertexts (and, for euintN ops, whole integers / whole
transactions) are dealt to ranks as contiguous slices; the converted evaluation keys of rank `src` are
broadcast as one opaque device blob (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

import numpy as np


def shard_bounds(n_units: int, world: int, align: int = 1):
    """contiguous [start, stop) per rank; sizes differ by less than 2*`align`; every unit covered exactly once.
    `align` > 1 rounds slice boundaries to multiples of it (e.g. the per-GPU wave size) where possible."""
    if world < 1 or n_units < 0 or align < 1:
        raise ValueError("bad sharding arguments")
    groups = -(-n_units // align)
    base, extra = divmod(groups, world)
    bounds, start = [], 0
    for r in range(world):
        g = base + (1 if r < extra else 0)
        stop = min(n_units, start + g * align)
        bounds.append((start, stop))
        start = stop
    return bounds


def my_shard(n_units: int, rank: int, world: int, align: int = 1):
    return shard_bounds(n_units, world, align)[rank]


def replicate_keys(engine, rank: int, world: int, src: int = 0, device=None):
    """rank `src` has loaded (and NTT-converted) the keys; every other rank adopts its device blob."""
    if world == 1:
        return
    import torch
    import torch.distributed as dist

    blob = torch.empty(engine.key_blob_bytes, dtype=torch.uint8, device=device if device is not None else "cuda")
    on_gpu = blob.is_cuda
    if rank == src:
        engine.export_keys_device(blob)       # synchronises the engine's own stream before returning
    # the broadcast is ordered on torch's CURRENT stream only, the engine reads the blob on its own non-blocking stream
    # (and its header through the legacy default stream): without these two synchronisations the import could race the
    # NCCL copy whenever the caller runs under a non-default torch stream and load partial keys without any error
    if on_gpu:
        torch.cuda.current_stream().synchronize()
    dist.broadcast(blob, src)
    if on_gpu:
        torch.cuda.current_stream().synchronize()
    if rank != src:
        engine.import_keys_device(blob)


def sharded_apply(fn, operands, rank: int, world: int, align: int = 1):
    """apply `fn(*operand_slices)` to this rank's contiguous slice of the batch (axis 0 of every operand).
    Returns ((start, stop), result)."""
    n = operands[0].shape[0]
    start, stop = my_shard(n, rank, world, align)
    return (start, stop), fn(*[op[start:stop] for op in operands])


def gather_host(result: np.ndarray, bounds, n_units: int, rank: int, world: int, dst: int = 0):
    """collect per-rank result slices on rank `dst` (host arrays; off the hot path: used by tests/examples)."""
    if world == 1:
        return result
    import torch
    import torch.distributed as dist

    parts = [None] * world if rank == dst else None
    dist.gather_object((bounds, result), parts, dst=dst)
    if rank != dst:
        return None
    out = np.empty((n_units,) + result.shape[1:], result.dtype)
    for (s, e), part in parts:
        out[s:e] = part
    return out
