#!/usr/bin/env python
"""bench.py — PBS/s of the batched programmable bootstrap (BASELINE.json metric) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
                  [--arith torus|goldilocks] [--scaling weak|strong]

A "step" is one pass of the hot path (LWE key switch -> blind rotate -> sample extract) over one
batch of B ciphertexts per GPU.  Workload = BASELINE configs[1]: 4-bit shortint PBS with lookup
tables at N=2048 on the named set PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 (n=918, N=2048, k=1,
PBS base 2^23 x 1, KS base 2^4 x 4), synthetic random ciphertexts, in the arithmetic that set is
defined on (2^64 torus, f64 negacyclic FFT; --arith goldilocks runs the exact-NTT arithmetic of the
in-tree reference code on the same shape instead).

  value        device-resident throughput: inputs already in HBM, CUDA events on the launching stream
  e2e          the same through the host-pointer C-ABI call lux_pbs_batch (pinned host buffers,
               H2D + kernels + D2H inside the timed region); e2e_pageable: the same from plain numpy
  roofline     dominant kernel (the blind rotation) against the resource that binds it: the FP64
               pipe (torus) / the integer pipe (goldilocks); roofline_hbm: its HBM side
  cpu_baseline the CPU oracle (C restatement, OpenMP over ciphertexts) on a bounded sample of the
               same workload, on this box's host cores
  parity       16 ciphertexts at random indices over the whole batch (last wave and ragged last CTA
               included) bit-exact vs the oracle; parity_ranks: ranks whose outputs for rank 0's
               real ciphertexts equal the oracle's (keys arrive there by NCCL broadcast)
  extra.c5     a short slice of BASELINE configs[4] (confidential ERC20 transfers, euint64) per GPU,
               every output balance decrypted and compared with the plaintext replay

Nothing here reads /root/reference.  oracle/ is used only for key/ciphertext fixtures, the parity
checks outside the timed region, and the cpu_baseline / --impl reference legs.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "pbs_per_s_batched"
UNIT = "PBS/s"
SHAPE = "C2: 4-bit shortint PBS with LUT, PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 shape (n=918, N=2048, k=1, ell=1, KS 4x4 bits)"
LUTS = [list(range(16)), [(x * x) % 16 for x in range(16)], [x >> 2 for x in range(16)], [x & 3 for x in range(16)]]


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------ arithmetics
class TorusSide:
    """fixtures / checker for LUX_ARITH_TORUS64_FFT (oracle/torus.c)"""
    name = "torus"
    dtype = "f64 FFT on u64 torus words"
    workload = SHAPE + ", 2^64 torus / f64 FFT (the named set's own arithmetic), TUniform noise"

    def __init__(self):
        import tfhe_b200 as T
        from oracle import torus as TO
        self.T, self.O = T, TO
        self.params = T.S2_TORUS
        self.mod = 1 << 64

    def fixture(self, seed=3):
        TO, p = self.O, self.params
        self.ctx = TO.Ctx(TO.Params(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ell_ks=p.ell_ks, logB_ks=p.logB_ks, order=p.order))
        self.ks = TO.keygen(self.ctx, seed, TO.S2_GLWE_NOISE, TO.S2_LWE_NOISE)
        TO.install_keys(self.ctx, self.ks)
        return self.ks

    def delta(self, space=32):
        return self.T.delta(space, self.params.arith)

    def testvs(self):
        d = self.delta()
        return np.stack([self.T.make_testv(self.params.N, 16, d, l, self.params.arith) for l in LUTS])

    def random_words(self, rs, shape):
        return rs.randint(0, 2 ** 63, size=shape, dtype=np.int64).astype(np.uint64) * np.uint64(2) + rs.randint(0, 2, size=shape).astype(np.uint64)

    def encrypt(self, seed, msgs):
        return self.O.lwe_encrypt_batch(seed, self.ks.s_in, np.asarray(msgs, np.uint64) * np.uint64(self.delta()), self.O.S2_GLWE_NOISE)

    def decrypt(self, cts):
        d = self.delta()
        return [self.O.decode(int(v), d, 32) for v in self.O.lwe_decrypt_batch(self.ks.s_in, cts)]

    def oracle_pbs(self, cts, tvs, idx):
        return self.ctx.pbs_batch(cts, tvs, idx)

    def threads(self):
        import oracle as O
        return O.lib().orc_max_threads()


class GoldilocksSide:
    """fixtures / checker for LUX_ARITH_GOLDILOCKS_NTT (oracle/oracle.c)"""
    name = "goldilocks"
    dtype = "u64 (Goldilocks field)"
    workload = SHAPE + ", Goldilocks exact NTT (the in-tree reference arithmetic)"
    SG, SL = 2.0 ** -47.8, 2.0 ** -19.8

    def __init__(self):
        import oracle as O
        import tfhe_b200 as T
        self.T, self.O = T, O
        self.params = T.S2_SHORTINT
        self.mod = O.Q

    def fixture(self, seed=3):
        O, p = self.O, self.params
        self.ctx = O.Ctx(O.Params(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ell_ks=p.ell_ks, logB_ks=p.logB_ks, order=p.order))
        self.ks = O.keygen(self.ctx, seed, self.SG, self.SL)
        O.install_keys(self.ctx, self.ks)
        return self.ks

    def delta(self, space=32):
        return self.O.delta(space)

    def testvs(self):
        return np.stack([self.O.testv_lut(self.params.N, 16, self.delta(), l) for l in LUTS])

    def random_words(self, rs, shape):
        return rs.randint(0, 2 ** 63, size=shape, dtype=np.int64).astype(np.uint64) % np.uint64(self.O.Q)

    def encrypt(self, seed, msgs):
        return self.O.lwe_encrypt_batch(seed, self.ks.s_in, np.asarray(msgs, np.uint64) * np.uint64(self.delta()), self.SG)

    def decrypt(self, cts):
        d = self.delta()
        return [self.O.decode(self.O.lwe_decrypt(self.ks.s_in, c), d, 32) for c in cts]

    def oracle_pbs(self, cts, tvs, idx):
        return self.ctx.pbs_batch(cts, tvs, idx)

    def threads(self):
        return self.O.lib().orc_max_threads()


def make_side(arith):
    return TorusSide() if arith == "torus" else GoldilocksSide()


def make_inputs(side, B, seed, nreal=32):
    """B ciphertexts: `nreal` real encryptions of 4-bit messages SPREAD over the batch (head, tail and random positions:
    checked against the oracle and decrypted after the run), the rest uniform random words (throughput is data
    independent).  Returns (cts, tvs, idx, real_positions, msgs)."""
    rs = np.random.RandomState(seed)
    cts = side.random_words(rs, (B, side.params.ct_len))
    nreal = min(nreal, B)
    pos = np.unique(np.concatenate([np.arange(min(4, B)), np.arange(max(0, B - 4), B), rs.randint(0, B, nreal)]))[:max(nreal, 1)]
    if B > 8 and (B - 1) not in pos:
        pos[-1] = B - 1
    pos = np.unique(pos)
    msgs = rs.randint(0, 16, len(pos))
    cts[pos] = side.encrypt(seed + 7, msgs)
    idx = (np.arange(B) % len(LUTS)).astype(np.uint32)
    return cts, side.testvs(), idx, pos, msgs


# ----------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, r[3:7]):
                if v.strip().lower() == "active":
                    reasons.add(nme)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# --------------------------------------------------------------------------- algorithmic work
def work_per_pbs(p, arith):
    """arithmetic operations and key bytes of one PBS.
    goldilocks: butterflies + MACs, one modular multiply each.
    torus: FP64 lane-operations of the operation sequence oracle/torus.c fixes (fft64.cuh): per CMUX (k+1)*ell forward
    transforms of N/2 complex points (6 fused multiply-adds per butterfly), k+1 inverse ones (8 operations per
    butterfly: 4 add/sub, 2 mul, 2 fma), (k+1)^2*ell*(N/2) complex multiply-accumulates (4 fma each), one exact
    int->double per digit and four operations per double->torus conversion."""
    import math

    logN = int(math.log2(p.N))
    bsk_bytes = p.n * (p.k + 1) ** 2 * p.ell * p.N * 8
    ksk_bytes = p.k * p.N * p.ell_ks * (p.n + 1) * 8 if p.order == 1 else (p.k + 1) ** 2 * p.ell * p.N * 8
    if arith == "torus":
        M = p.N // 2
        bf = M // 2 * (logN - 1)
        per_iter = (p.k + 1) * p.ell * bf * 6 + (p.k + 1) * bf * 8 + (p.k + 1) ** 2 * p.ell * M * 4 \
            + (p.k + 1) * p.ell * p.N * 1 + (p.k + 1) * p.N * 4
        smem = (p.k + 1) * 12 * M * 16          # bytes through the shared-memory crossbar per CMUX (DESIGN: 12 crossings of 16 KiB per polynomial)
        return dict(br_ops=p.n * per_iter, smem_bytes=p.n * smem, bsk_bytes=bsk_bytes, ksk_bytes=p.k * p.N * p.ell_ks * (p.n + 1))
    bf = p.N // 2 * logN
    per_iter = ((p.k + 1) * p.ell + (p.k + 1)) * bf + (p.k + 1) ** 2 * p.ell * p.N
    return dict(br_ops=p.n * per_iter, bsk_bytes=bsk_bytes, ksk_bytes=ksk_bytes)


# ------------------------------------------------------------------------------- reference arm
def cpu_pbs_rate(side, cts, tvs, idx, budget_s=12.0):
    """time the oracle (OpenMP over ciphertexts) on a bounded sample; returns (PBS/s, threads, sample size, seconds)"""
    threads = side.threads()
    t0 = time.perf_counter()
    side.oracle_pbs(cts[:threads], tvs, idx[:threads])
    t1 = time.perf_counter() - t0                       # one PBS per thread
    rounds = max(1, min(64, int(budget_s / max(t1, 1e-3))))   # ~budget_s seconds of CPU work on all host threads
    nb = min(len(cts), threads * rounds)
    t0 = time.perf_counter()
    _, used = side.oracle_pbs(cts[:nb], tvs, idx[:nb])
    dt = time.perf_counter() - t0
    return nb / dt, int(used), nb, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    side = make_side(args.arith)
    side.fixture()
    threads = side.threads()
    per_step = threads          # one PBS per host thread per step: a bounded sample of the B-ciphertext batch
    cts, tvs, idx, *_ = make_inputs(side, per_step, 4)
    for _ in range(args.warmup):
        side.oracle_pbs(cts, tvs, idx)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        side.oracle_pbs(cts, tvs, idx)
    dt = time.perf_counter() - t0
    v = per_step * args.steps / dt
    sample = f"{per_step} PBS per step (1 per host thread) of the B={args.batch} batch, oracle C port ({side.name}), OpenMP"
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": side.dtype, "data": "synthetic",
        "config": {"workload": side.workload, "batch_per_step": per_step},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference binary (Go luxfi/tfhe / wasm blob; tfhe-rs for the named set) is not in the tree and no go/node/cargo "
                "toolchain exists; this arm is the C restatement of the algorithm (oracle/), kind=port",
    })


# ------------------------------------------------------------------------- C5 slice (ERC20 transfers)
def erc20_replay(f, t, a):
    """WrappingERC20.sol:86-102: canTransfer = amount <= balance; both updates selected on it (wrapping u64)"""
    m = (1 << 64) - 1
    ok = a <= f
    return (f - a if ok else f), ((t + a) & m if ok else t)


def c5_slice(side, eng, n_transfers, rank, world, dist, torch):
    """n_transfers confidential ERC20 transfers on every rank (rank 0's ciphertexts, NCCL-broadcast): device-resident
    timing of lux_erc20_transfer_device, outputs compared across ranks (bit-identical) and, on rank 0, decrypted
    and compared with the plaintext replay."""
    import tfhe_b200.radix as R

    p = side.params
    rs = np.random.RandomState(77)
    shape = (3, n_transfers, 32, p.ct_len)
    if rank == 0:
        f = [int(x) for x in rs.randint(0, 2 ** 62, n_transfers)]
        t = [int(x) for x in rs.randint(0, 2 ** 62, n_transfers)]
        a = [int(x) if i % 3 else int(x) >> 8 for i, x in enumerate(rs.randint(0, 2 ** 62, n_transfers))]
        f[0], t[0], a[0] = 10000, 0, 1337             # EncryptedERC20.ts:98
        if n_transfers > 1:
            f[1], t[1], a[1] = 1000, 0, 1337          # EncryptedERC20.ts:114 (amount > balance: nothing moves)
        digits = np.array([[(v >> (2 * i)) & 3 for i in range(32)] for arr in (f, t, a) for v in arr], dtype=np.uint64).reshape(-1)
        cts = side.encrypt(991, digits).reshape(shape)
        d_in = torch.from_numpy(cts.view(np.int64)).cuda()
    else:
        d_in = torch.empty(shape, dtype=torch.int64, device="cuda")
    if world > 1:
        dist.broadcast(d_in, src=0)
    torch.cuda.synchronize()
    d_nf, d_nt = torch.empty_like(d_in[0]), torch.empty_like(d_in[0])
    L = R.lib()

    def run():
        rc = L.lux_erc20_transfer_device(eng._h, d_in[0].data_ptr(), d_in[1].data_ptr(), d_in[2].data_ptr(), n_transfers,
                                         d_nf.data_ptr(), d_nt.data_ptr())
        eng._ck(rc)
        eng.synchronize()

    run()                                                # warm-up (builds the circuit, sizes the pool)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    h = torch.stack([d_nf.sum(), d_nt.sum(), (d_nf ^ (d_nf >> 17)).sum(), (d_nt ^ (d_nt >> 13)).sum()])
    hs = [h]
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        hs = [torch.empty_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
    if rank != 0:
        return None
    same = all(bool(torch.equal(x, hs[0])) for x in hs)
    nb = 32
    nf = side.decrypt(d_nf.cpu().numpy().view(np.uint64).reshape(n_transfers * nb, p.ct_len))
    nt = side.decrypt(d_nt.cpu().numpy().view(np.uint64).reshape(n_transfers * nb, p.ct_len))
    join = lambda d, i: sum(int(b) << (2 * j) for j, b in enumerate(d[i * nb:(i + 1) * nb]))
    bad = 0
    for i in range(n_transfers):
        wf, wt = erc20_replay(f[i], t[i], a[i])
        bad += int(join(nf, i) != wf) + int(join(nt, i) != wt)
    pbs = int(L.lux_erc20_pbs_count())
    return {"workload": "C5 slice: confidential ERC20 transfers (euint64 le + select + sub/add), device-resident ciphertexts",
            "transfers_per_gpu": n_transfers, "transfers_per_s": world * n_transfers / float(tt.item()),
            "pbs_per_transfer": pbs, "pbs_per_s": world * n_transfers * pbs / float(tt.item()),
            "all_balances_correct": bad == 0 and same, "wrong_balances": bad, "ranks_bit_identical": same,
            "reference_vectors": [join(nf, 0), join(nt, 0)] + ([join(nf, 1), join(nt, 1)] if n_transfers > 1 else [])}


# ------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    import tfhe_b200 as T

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        log(f"note: WORLD_SIZE={world} overrides --gpus {args.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    side = make_side(args.arith)
    params = side.params
    if args.scaling == "strong":
        # ONE global batch split over the ranks (ragged): rank r gets [lo, hi)
        G = args.global_batch
        lo, hi = G * rank // world, G * (rank + 1) // world
        B = hi - lo
    else:
        B = args.batch
    eng = T.Engine(params, device=local)
    if args.teams:
        eng.force_teams(args.teams)      # tuning knob: ciphertexts per CTA in the blind-rotate launch

    # ---- evaluation keys: rank 0 converts on its GPU, NCCL-broadcasts the device blob
    if rank == 0:
        ks = side.fixture()
        eng.load_keyset(ks)
    from tfhe_b200.sharding import replicate_keys

    replicate_keys(eng, rank, world, src=0, device="cuda")
    # ---- inputs: every rank gets its own B ciphertexts (no data-path collective)
    if rank == 0:
        cts, tvs, idx, pos, msgs = make_inputs(side, B, 4)
    else:
        rs = np.random.RandomState(100 + rank)
        cts = side.random_words(rs, (B, params.ct_len))
        tvs = side.random_words(rs, (4, params.N))
        idx = (np.arange(B) % 4).astype(np.uint32)

    h_in = torch.from_numpy(cts.view(np.int64)).pin_memory()
    h_tv = torch.from_numpy(tvs.view(np.int64)).pin_memory()
    h_idx = torch.from_numpy(idx.view(np.int32)).pin_memory()
    h_out = torch.empty_like(h_in).pin_memory()
    d_in, d_tv, d_idx = h_in.cuda(), h_tv.cuda(), h_idx.cuda()
    d_out = torch.empty_like(d_in)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    # one explicit (non-default) stream for every launch and every event of this process: handle 0 would
    # select the engine's own non-blocking stream, which events recorded on torch's default stream do not see
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream
    assert sptr != 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        eng.pbs_batch_device(d_in, d_tv, d_idx, B, d_out, n_luts=4, stream=sptr)

    # ---- device-resident timing: per-step CUDA events on the launching stream, L2 flushed between steps
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    eng.set_timing(True)
    l0 = eng.launch_count
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    ks_ms, br_ms = [], []
    barrier()
    for a, b in ev:
        flush.zero_()                      # untimed: evicts keys and ciphertexts from L2
        a.record(stream)
        step_device()
        b.record(stream)
        b.synchronize()
        t = eng.last_timing()
        ks_ms.append(t[0])
        br_ms.append(t[1])
    barrier()
    launches = eng.launch_count - l0
    eng.set_timing(False)
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    clocks = sampler.stop() if rank == 0 else None
    tt = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    nb_all = torch.tensor([B], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(nb_all, op=dist.ReduceOp.SUM)
    total_ms = float(tt.item())
    global_B = int(nb_all.item())
    value = global_B * args.steps / (total_ms * 1e-3)

    # ---- end to end through the host-pointer ABI call (H2D/D2H inside): pinned buffers, then plain numpy (pageable)
    def e2e_rate(fn):
        for _ in range(2):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()                                          # blocking: returns after the D2H copy
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        return global_B * args.steps / float(te.item())

    e2e = e2e_rate(lambda: eng.pbs_batch(h_in, h_tv, h_idx, out=h_out))
    p_out = np.empty_like(cts)
    e2e_pageable = e2e_rate(lambda: eng.pbs_batch(cts, tvs, idx, out=p_out))
    h2d = h_in.numel() * 8 + h_tv.numel() * 8 + h_idx.numel() * 4
    d2h = h_out.numel() * 8

    # ---- parity on every rank: rank 0's real ciphertexts, bootstrapped under the broadcast keys
    if rank == 0:
        real = torch.from_numpy(cts[pos].view(np.int64)).cuda()
        real_idx = torch.from_numpy(idx[pos].view(np.int32)).cuda()
        n_real = torch.tensor([len(pos)], dtype=torch.int64, device="cuda")
    else:
        n_real = torch.zeros(1, dtype=torch.int64, device="cuda")
    real_tv = d_tv.clone()
    if world > 1:
        dist.broadcast(n_real, src=0)
        if rank != 0:
            real = torch.empty((int(n_real.item()), params.ct_len), dtype=torch.int64, device="cuda")
            real_idx = torch.empty(int(n_real.item()), dtype=torch.int32, device="cuda")
        dist.broadcast(real, src=0)
        dist.broadcast(real_idx, src=0)
        dist.broadcast(real_tv, src=0)
    torch.cuda.synchronize()
    real_out = torch.empty_like(real)
    eng.pbs_batch_device(real, real_tv, real_idx, real.shape[0], real_out, n_luts=4, stream=sptr)
    torch.cuda.synchronize()
    gathered = [real_out]
    if world > 1:
        gathered = [torch.empty_like(real_out) for _ in range(world)]
        dist.all_gather(gathered, real_out)

    # ---- C5 slice
    c5 = None
    if args.c5 > 0:
        try:
            c5 = c5_slice(side, eng, args.c5, rank, world, dist, torch)
        except Exception as ex:  # noqa: BLE001 - reported on the line, never fatal for the headline
            c5 = {"error": repr(ex)[:200]}

    # ---- single-PBS latency (BASELINE metric's second half): one ciphertext, device-resident, CUDA events
    lat_ms = None
    if rank == 0:
        la, lb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        d_one = torch.empty_like(d_in[:1])
        lats = []
        for i in range(6):
            la.record(stream)
            eng.pbs_batch_device(d_in[:1], d_tv, d_idx[:1], 1, d_one, n_luts=4, stream=sptr)
            lb.record(stream)
            lb.synchronize()
            if i:
                lats.append(la.elapsed_time(lb))
        lat_ms = statistics.median(lats)
        torch.cuda.synchronize()
        assert torch.equal(d_one[0], d_out[0]), "single-ciphertext launch disagrees with the batched launch"

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- parity (outside every timed region): device path == host paths == oracle, over the WHOLE batch
    got = h_out.numpy().view(np.uint64)
    assert np.array_equal(got, d_out.cpu().numpy().view(np.uint64)), "device-resident and host-pointer paths disagree"
    assert np.array_equal(got, p_out), "pinned and pageable host paths disagree"
    rs = np.random.RandomState(12)
    wave = eng.wave_ciphertexts(B)
    last_wave = (B - 1) // wave * wave
    ncheck = 16
    pick = np.unique(np.concatenate([[0, B - 1, last_wave, min(B - 1, last_wave + 1)], rs.randint(0, B, ncheck), rs.randint(last_wave, B, 4)]))[:ncheck + 4]
    want, _ = side.oracle_pbs(cts[pick], tvs, idx[pick])
    assert np.array_equal(got[pick], want), "GPU PBS differs from the oracle"
    dec = side.decrypt(got[pos])
    for i, m, dd in zip(pos, msgs, dec):
        assert dd == LUTS[idx[i]][int(m)], (int(i), int(m), dd)
    want_real, _ = side.oracle_pbs(cts[pos], tvs, idx[pos])
    parity_ranks = sum(int(np.array_equal(g.cpu().numpy().view(np.uint64), want_real)) for g in gathered)
    assert parity_ranks == world, f"only {parity_ranks} of {world} ranks reproduce the oracle under the broadcast keys"

    # ---- rooflines for the dominant kernel (blind rotate)
    w = work_per_pbs(params, side.name)
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    waves = -(-B // wave)
    br_s = statistics.mean(br_ms) * 1e-3
    # algorithmic HBM bytes of one launch: the Fourier/NTT key once + ciphertexts in + accumulators out
    br_bytes = w["bsk_bytes"] + B * ((params.n + 1) * 8 + (params.k * params.N + 1) * 8)
    # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel at this batch
    # (profiles/traffic.json, written from the committed ncu summary by scripts/ncu_summary.py traffic)
    traffic, traffic_src = None, None
    tj = ROOT / "profiles" / "traffic.json"
    if tj.exists():
        t = json.loads(tj.read_text())
        if int(t.get("batch", -1)) == B and t.get("arith", "goldilocks") == side.name:
            traffic, traffic_src = float(t["dram_bytes_per_launch"]), t.get("source")
    ops_s = B * w["br_ops"] / br_s
    if side.name == "torus":
        fp64 = eng.fp64_peak()
        sm_clk = (clocks or {}).get("sm_mhz") or 1965.0
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        smem_peak = 128.0 * sms * sm_clk * 1e6          # B/s: 128 B/clk/SM crossbar
        roofline = {"bound": "fp64", "kernel": "blind_rotate_fft_kernel", "achieved": ops_s / 1e12, "peak": fp64 / 1e12,
                    "unit": "T FP64 lane-op/s", "frac": ops_s / fp64, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": "register-only DFMA micro-benchmark on this GPU in this run (lux_debug_fp64_peak); "
                                   "profiles/r02_arith_ubench.txt holds the same measurement (16.9 T/s)",
                    "ops_per_pbs": w["br_ops"], "ms_per_launch": br_s * 1e3,
                    "smem": {"achieved_tbs": B * w["smem_bytes"] / br_s / 1e12, "peak_tbs": smem_peak / 1e12,
                             "frac": B * w["smem_bytes"] / br_s / smem_peak, "bytes_per_pbs": w["smem_bytes"],
                             "note": "second binding resource: shared-memory crossbar, 128 B/clk/SM at the sampled SM clock"},
                    "target_frac": value / world / 125000.0}
    else:
        imad, iadd, bfly = eng.int_peak()
        # pipe-rate bound: ALU instructions per butterfly x 2 cycles per warp instruction per SMSP (profiles/r01e_blind_rotate_sass_mix.txt)
        alu_per_bfly = 21.7
        pipe_bound = torch.cuda.get_device_properties(local).multi_processor_count * 4 * 32 / (alu_per_bfly * 2) * ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
        roofline = {"bound": "int", "kernel": "blind_rotate_kernel", "achieved": ops_s / 1e12, "peak": bfly / 1e12,
                    "unit": "T modmul/s", "frac": ops_s / bfly, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": "register-only Goldilocks butterfly micro-benchmark on this GPU (lux_debug_int_peak)",
                    "peak_pipe_rate": pipe_bound / 1e12, "frac_pipe_rate": ops_s / pipe_bound, "alu_instr_per_butterfly": alu_per_bfly,
                    "imad_wide_tops": imad / 1e12, "iadd3_tops": iadd / 1e12, "modmul_per_pbs": w["br_ops"],
                    "ms_per_launch": br_s * 1e3, "target_frac": value / world / 125000.0}
    roofline_hbm = {"bound": "hbm", "achieved": br_bytes / br_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": br_bytes / br_s / 1e9 / hbm_peak, "bytes_per_launch": br_bytes,
                    "traffic_ratio": (traffic / br_bytes) if traffic else None,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if pk.exists() else "fallback 6650",
                    "note": "key once + ciphertexts in/out; the key is re-read from L2/HBM once per wave of resident ciphertexts "
                            f"({waves} waves of {wave})"}

    # ---- CPU baseline on a bounded sample: rank 0 at N = 1 only (contract); under torchrun the host threads are pinned to 1 per rank
    skip_cpu = args.no_cpu or world > 1
    cpu_v, cores, nb, dt = (None, 0, 0, 0.0) if skip_cpu else cpu_pbs_rate(side, cts, tvs, idx)
    cpu_lat_ms = None
    if not skip_cpu:
        t0 = time.perf_counter()
        side.oracle_pbs(cts[:1], tvs, idx[:1])
        cpu_lat_ms = (time.perf_counter() - t0) * 1e3
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": side.dtype, "data": "synthetic",
        "config": {"workload": side.workload, "batch_per_gpu": B, "global_batch": global_B, "l2": "flushed between timed steps (256 MiB write)",
                   "parallelism": f"batch-sharded x{world}, keys NCCL-broadcast once", "wave_ciphertexts": wave, "arith": side.name},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "e2e_pageable": {"value": e2e_pageable, "unit": UNIT, "note": "same call from plain (pageable) numpy buffers"},
        "gpu_launches": int(launches),
        "kernel_ms": {"lwe_keyswitch": statistics.mean(ks_ms), "blind_rotate": statistics.mean(br_ms)},
        "roofline": roofline, "roofline_hbm": roofline_hbm,
        "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{nb} PBS of the same batch in {dt:.1f} s, oracle C port ({side.name}), OpenMP over ciphertexts"},
        "latency_single_pbs_ms": {"gpu": lat_ms, "cpu_1_thread": cpu_lat_ms},
        "clocks": clocks,
        "parity": f"{len(pick)} outputs at random indices over the whole batch (last wave, ragged tail) bit-exact vs oracle; "
                  f"{len(pos)} decryptions correct",
        "parity_ranks": parity_ranks,
        "extra": {"c5": c5},
    }
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE line, the JSON result: everything else a library prints to fd 1 (NCCL's version
    banner, for one) is sent to stderr."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4736,
                    help="ciphertexts per GPU per step (default 4,736 = 8 full waves of 148 SMs x 4 resident ciphertexts, torus kernel)")
    ap.add_argument("--arith", default="torus", choices=["torus", "goldilocks"],
                    help="torus: 2^64 torus / f64 FFT (the named set's arithmetic, default); goldilocks: exact NTT (in-tree reference arithmetic)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --batch per GPU; strong: --global-batch split over the GPUs (ragged)")
    ap.add_argument("--global-batch", type=int, default=65120, help="strong scaling: total ciphertexts per step")
    ap.add_argument("--c5", type=int, default=296, help="ERC20 transfers per GPU in the C5 slice (0 = skip)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--teams", type=int, default=0, help="force ciphertexts per CTA (tuning)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (tuning sweeps)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        log("note: timing rules want >= 3 warm-up steps")
    if args.impl == "reference":
        # torchrun pins OMP_NUM_THREADS=1 for every rank; the reference arm runs on rank 0 alone and is meant to use
        # all the host threads it can (must be set before the OpenMP runtime of the oracle library loads)
        if int(os.environ.get("RANK", "0")) == 0 and "TORCHELASTIC_RUN_ID" in os.environ:
            os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
