"""lux_pbs_batch_device_static (include/lux_tfhe.h): the PBS entry point that allocates nothing — GPU box: pytest -m gpu.

The entry point has no counterpart in the reference; it was added for callers that
cannot call cudaMalloc.  Its scratch (key switch digits, key-switched ciphertexts) is a static device array inside the
library.  The test calls it through the C ABI on device buffers the CALLER owns, twice in a row with different batches (the
array keeps the first call's contents), and checks the output words against the existing lux_pbs_batch_device on the same
inputs — bit for bit, the bar of every parity test here — and against oracle/torus.c for a sample.  The file is named to be
collected last (pytest -x at the end of a round stops at the first failure)."""
import numpy as np
import pytest

import tfhe_b200 as T
from oracle import torus as TO

pytestmark = pytest.mark.gpu


def u64_rand(rs, shape):
    return rs.randint(0, 2**63, size=shape, dtype=np.int64).astype(np.uint64) * np.uint64(2) + rs.randint(0, 2, size=shape).astype(np.uint64)


class Rig:
    def __init__(self, params, seed):
        self.p = params
        self.ctx = TO.Ctx(TO.Params(n=params.n, N=params.N, k=params.k, ell=params.ell, logB=params.logB,
                                    ell_ks=params.ell_ks, logB_ks=params.logB_ks, order=params.order))
        self.ks = TO.keygen(self.ctx, seed, TO.S2_GLWE_NOISE, TO.S2_LWE_NOISE)
        TO.install_keys(self.ctx, self.ks)
        self.eng = T.Engine(params)
        self.eng.load_keyset(self.ks)


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64 if a.dtype == np.uint64 else np.int32)).cuda()


def run_both(rig, B, n_luts, seed):
    """the same seeded batch through the existing entry point and through the static one; returns (existing, static, inputs)"""
    import torch
    e, p = rig.eng, rig.p
    rs = np.random.RandomState(seed)
    cts, tvs = u64_rand(rs, (B, p.ct_len)), u64_rand(rs, (n_luts, p.N))
    idx = rs.randint(0, n_luts, B).astype(np.uint32)
    d_in, d_tv, d_idx = dev(cts), dev(tvs), dev(idx)
    d_ref = torch.zeros(B * p.ct_len, dtype=torch.int64, device="cuda")
    d_got = torch.full((B * p.ct_len,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    e.pbs_batch_device(d_in, d_tv, d_idx, B, d_ref, n_luts=n_luts)
    e.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    e.pbs_batch_device_static(d_in, d_tv, d_idx, B, d_got, n_luts=n_luts)
    e.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    ref = d_ref.cpu().numpy().view(np.uint64).reshape(B, p.ct_len)
    got = d_got.cpu().numpy().view(np.uint64).reshape(B, p.ct_len)
    return ref, got, (cts, tvs, idx), free_before - free_after


@pytest.mark.parametrize("params,seed", [(T.TEST_SMALL_TORUS, 61), (T.TEST_SMALL_TORUS_O2, 62)], ids=["ks_pbs", "pbs_ks"])
def test_static_scratch_pbs_matches_existing_entry_point_twice_in_a_row(params, seed):
    rig = Rig(params, seed)
    # first call: several key-switch row tiles and a ragged tail; second call: a smaller batch over the first call's leftovers
    grown = []
    for call, (B, n_luts) in enumerate([(300, 3), (7, 2)]):
        ref, got, (cts, tvs, idx), used = run_both(rig, B, n_luts, seed + call)
        assert np.array_equal(got, ref), f"call {call}: static-scratch PBS differs from lux_pbs_batch_device"
        pick = np.unique(np.concatenate([[0, B - 1], np.random.RandomState(call).randint(0, B, 4)]))
        want, _ = rig.ctx.pbs_batch(cts[pick], tvs, idx[pick])
        assert np.array_equal(got[pick], want), f"call {call}: differs from oracle/torus.c"
        grown.append(used)
    assert grown[1] <= 0, f"device memory in use grew by {grown[1]} bytes across a call that must not allocate"


def test_static_scratch_pbs_named_set_and_limit():
    """the named set: two calls in a row (one kernel wave of the latency form, then a multi-tile batch), then the size limit"""
    import torch
    rig = Rig(T.S2_TORUS, 63)
    e, p = rig.eng, rig.p
    for call, (B, n_luts) in enumerate([(5, 2), (200, 4)]):
        ref, got, (cts, tvs, idx), used = run_both(rig, B, n_luts, 70 + call)
        assert np.array_equal(got, ref), f"call {call}: static-scratch PBS differs from lux_pbs_batch_device"
        want, _ = rig.ctx.pbs_batch(cts[:2], tvs, idx[:2])
        assert np.array_equal(got[:2], want)
    assert used <= 0, f"device memory in use grew by {used} bytes across a call that must not allocate"
    # B = 1024 is the largest batch the 16 MiB array takes at this set; 1025 is refused before any launch
    B = 1025
    d_in = torch.zeros(B * p.ct_len, dtype=torch.int64, device="cuda")
    d_out = torch.zeros(B * p.ct_len, dtype=torch.int64, device="cuda")
    d_tv = torch.zeros(p.N, dtype=torch.int64, device="cuda")
    launches = e.launch_count
    with pytest.raises(T.LuxError) as err:
        e.pbs_batch_device_static(d_in, d_tv, None, B, d_out, n_luts=1)
    assert "static array" in str(err.value) and "1024" in str(err.value)
    assert e.launch_count == launches
    e.pbs_batch_device_static(d_in, d_tv, None, 1024, d_out, n_luts=1)      # the limit itself runs
    e.synchronize()


def test_static_scratch_pbs_refuses_the_other_arithmetic():
    import torch
    e = T.Engine(T.TEST_SMALL_O1)
    d = torch.zeros(4 * T.TEST_SMALL_O1.ct_len + T.TEST_SMALL_O1.N, dtype=torch.int64, device="cuda")
    with pytest.raises(T.LuxError) as err:
        e.pbs_batch_device_static(d, d, None, 1, d, n_luts=1)
    assert "TORUS64" in str(err.value)
