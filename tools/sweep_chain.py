"""Timing sweep of the mode-N engines over the batch size (run on the GPU box)."""
import sys
import torch
sys.path.insert(0, ".")
import b200spec

engines = [int(e) for e in (sys.argv[1] if len(sys.argv) > 1 else "3").split(",")]
sizes = [int(b) for b in (sys.argv[2] if len(sys.argv) > 2 else "74,148,296,592,1024,2048").split(",")]
g = torch.Generator(device="cuda").manual_seed(2)
big = torch.randn((max(sizes), 20, 10000), generator=g, device="cuda")
big = (big.cumsum(-1) * 0.3 + torch.randn(big.shape, generator=g, device="cuda")).clamp_(-1024, 1024)
for eng in engines:
    p = b200spec.mel_params(engine=eng)
    for B in sizes:
        x = big[:B]
        for _ in range(3):
            y = b200spec.eeg_to_spectrogram(x, p)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n):
            y = b200spec.eeg_to_spectrogram(x, p)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"engine {eng} B {B:5d}  {ms:8.3f} ms  {B / ms * 1e3:10.0f} rec/s  us/record {1e3 * ms / B:7.3f}  hbm frac {B * 1324288 / (ms * 1e-3) / 1e9 / 6546.9:.3f}", flush=True)
