"""Scratch driver for engine 3 (run on the GPU box): parity vs engine 2 / oracle, then timing."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import b200spec
from oracle import melspec_oracle as mo

def eeg_like(rng, shape):
    x = np.cumsum(rng.standard_normal(shape), axis=-1) * 0.3
    return np.clip(x + rng.standard_normal(shape), -1024, 1024).astype(np.float32)

rng = np.random.default_rng(1)
for B in (1, 3, 40):
    x = eeg_like(rng, (B, 20, 10000))
    xt = torch.from_numpy(x).cuda()
    a = b200spec.eeg_to_spectrogram(xt, engine=3); torch.cuda.synchronize()
    b = b200spec.eeg_to_spectrogram(xt, engine=2); torch.cuda.synchronize()
    print("B", B, "engine3 vs engine2 max abs", float((a - b).abs().max()), flush=True)
    if B <= 3:
        want = mo.eeg_to_spectrogram_batch_oracle(x)
        print("   vs oracle", float(np.abs(a.cpu().numpy() - want).max()), flush=True)
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn((1024, 20, 10000), generator=g, device="cuda")
x = (x.cumsum(-1) * 0.3 + torch.randn((1024, 20, 10000), generator=g, device="cuda")).clamp_(-1024, 1024)
p3, p2 = b200spec.mel_params(engine=3), b200spec.mel_params(engine=2)
for name, p in (("engine3", p3), ("engine2", p2)):
    for _ in range(3):
        y = b200spec.eeg_to_spectrogram(x, p)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        y = b200spec.eeg_to_spectrogram(x, p)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(name, "ms/step", ms, "rec/s", 1024 / ms * 1e3, "frac", 1024 * 1324288 / (ms * 1e-3) / 1e9 / 6546.9, flush=True)
