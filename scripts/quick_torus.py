"""Quick device-resident timing of the torus/f64-FFT PBS at S2 (random key words: throughput is data independent).
usage: python scripts/quick_torus.py [B] [cfg...]     cfg = teams + 10 * variant [":" stagger_ns]"""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
import tfhe_b200 as T

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4440
teams_list = sys.argv[2:] or ["0"]   # teams + 10 * variant [: stagger ns]
p = T.S2_TORUS
eng = T.Engine(p)
rs = np.random.RandomState(1)
r64 = lambda n: rs.randint(0, 2**63, size=n, dtype=np.int64).astype(np.uint64) * np.uint64(2)
eng.load_bsk(r64(p.n * p.ggsw_len))
eng.load_lwe_ksk(r64(p.lwe_ksk_len))
d_in = torch.from_numpy(r64(B * p.ct_len).view(np.int64)).cuda()
d_tv = torch.from_numpy(r64(4 * p.N).view(np.int64)).cuda()
d_idx = torch.from_numpy((np.arange(B) % 4).astype(np.int32)).cuda()
d_out = torch.empty_like(d_in)
s = torch.cuda.Stream(); torch.cuda.set_stream(s)
eng.set_timing(True)
if os.environ.get("LAT"):   # small-batch latency: B in LAT (comma list), cfg from the command line
    for b in [int(x) for x in os.environ["LAT"].split(",")]:
        for spec in teams_list:
            t = int(spec.split(":")[0])
            eng.force_teams(t % 10); eng.fft_variant(t // 10)
            ms = []
            for _ in range(5):
                a, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(s); eng.pbs_batch_device(d_in[:b], d_tv, d_idx[:b], b, d_out[:b], n_luts=4, stream=s.cuda_stream); e2.record(s); e2.synchronize()
                ms.append(a.elapsed_time(e2))
            print(f"latency B={b} cfg={spec}: {min(ms[1:]):.3f} ms", flush=True)
            if os.environ.get("FFT64_TRACE") and t // 10 == 4:   # the latency kernel's phase marks (variant 4)
                eng.fft_trace()
                eng.pbs_batch_device(d_in[:b], d_tv, d_idx[:b], b, d_out[:b], n_luts=4, stream=s.cuda_stream)
                tr = eng.fft_trace()
                names = ["decompose", "F1", "bar", "F2", "bar", "F3", "sync", "M", "sync", "I2", "bar", "I3", "bar", "I4+update", "bar"]
                per = [v / (8 * b * p.n) for v in tr[:15]]
                print("   cycles per warp-iteration: " + ", ".join(f"{n} {v:.0f}" for n, v in zip(names, per)) + f"  | total {sum(per):.0f}", flush=True)
    sys.exit(0)
for spec in teams_list:
    t = int(spec.split(":")[0])
    if ":" in spec:
        eng.fft_stagger(int(spec.split(":")[1]))
    eng.force_teams(t % 10)
    eng.fft_variant(t // 10)
    for _ in range(2):
        eng.pbs_batch_device(d_in, d_tv, d_idx, B, d_out, n_luts=4, stream=s.cuda_stream)
    torch.cuda.synchronize()
    ms = []
    for _ in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s); eng.pbs_batch_device(d_in, d_tv, d_idx, B, d_out, n_luts=4, stream=s.cuda_stream); b.record(s); b.synchronize()
        ms.append(a.elapsed_time(b))
    if os.environ.get("FFT64_TRACE"):
        eng.fft_trace()
        eng.pbs_batch_device(d_in, d_tv, d_idx, B, d_out, n_luts=4, stream=s.cuda_stream)
        tr = eng.fft_trace()
        names = ["decompose", "fwd pass A", "transpose", "fwd pass B", "swap+bar", "key MAC", "bar", "inv pass B", "transpose", "inv pass A", "update"]
        per = [v / (2 * B * p.n) for v in tr[:11]]        # cycles per warp-iteration
        print("   phase cycles per warp-iteration: " + ", ".join(f"{n} {v:.0f}" for n, v in zip(names, per)) + f"  | total {sum(per):.0f}", flush=True)
    ks, br = eng.last_timing()
    print(f"B={B} cfg={spec} wave={eng.wave_ciphertexts(B)}: {min(ms):.2f} ms best, {B/min(ms)*1e3:.0f} PBS/s  (ks {ks:.2f} ms, br {br:.2f} ms)", flush=True)
