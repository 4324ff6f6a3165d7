"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: slice bounds, key-blob replication protocol,
sharded evaluation + gather.  The engine is replaced by a host stand-in with the same key-blob methods;
the real engine runs the same code on the GPU box (bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest

from tfhe_b200.sharding import gather_host, my_shard, replicate_keys, shard_bounds, sharded_apply


def test_shard_bounds_cover_exactly_once():
    for n in (0, 1, 7, 444, 4096, 10000):
        for world in (1, 2, 3, 4, 8):
            for align in (1, 3, 444):
                b = shard_bounds(n, world, align)
                assert b[0][0] == 0 and b[-1][1] == n
                assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
                sizes = [e - s for s, e in b]
                assert max(sizes) - min(sizes) <= 2 * align - 1
                assert all(s % align == 0 for s, _ in b if s < n)
    assert my_shard(10, 1, 2) == (5, 10)
    with pytest.raises(ValueError):
        shard_bounds(4, 0)


class FakeEngine:
    """key-blob protocol of tfhe_b200.Engine on host tensors"""

    def __init__(self, nbytes):
        self.key_blob_bytes = nbytes
        self.keys = None

    def load(self, seed):
        self.keys = np.random.RandomState(seed).randint(0, 256, self.key_blob_bytes).astype(np.uint8)

    def export_keys_device(self, t):
        t.copy_(__import__("torch").from_numpy(self.keys))

    def import_keys_device(self, t):
        self.keys = t.numpy().copy()

    def pbs(self, x):   # a deterministic "evaluation" that depends on the keys
        return (x.astype(np.uint64) * 3 + int(self.keys[:8].sum())).astype(np.uint64)


def _worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        eng = FakeEngine(4096)
        if rank == 0:
            eng.load(7)
        replicate_keys(eng, rank, world, src=0, device="cpu")
        ref = FakeEngine(4096)
        ref.load(7)
        assert np.array_equal(eng.keys, ref.keys)
        n = 1001
        cts = np.arange(n * 5, dtype=np.uint64).reshape(n, 5)
        bounds, part = sharded_apply(eng.pbs, [cts], rank, world, align=4)
        assert part.shape[0] == bounds[1] - bounds[0]
        full = gather_host(part, bounds, n, rank, world)
        if rank == 0:
            assert np.array_equal(full, ref.pbs(cts))
        q.put((rank, "ok"))
    except Exception as ex:  # surfaced by the parent
        q.put((rank, repr(ex)))
    finally:
        dist.destroy_process_group()


def test_replicate_and_shard_two_ranks_gloo():
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
