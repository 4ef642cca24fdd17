"""Timing of the raw-EEG transformer kernel (run on the GPU box)."""
import sys
import torch
sys.path.insert(0, ".")
import b200spec

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
g = torch.Generator(device="cuda").manual_seed(3)
x = torch.randn((B, 10000, 20), generator=g, device="cuda") * 40
for kw in (dict(n_feats=19, downsample=5), dict(n_feats=19, downsample=5, channel_major=True), dict(n_feats=8, apply_chris_magic_ch8=True, downsample=5)):
    t = b200spec.EEGTransformer(**kw)
    xin = x[:, :, :8].contiguous() if kw.get('apply_chris_magic_ch8') else x
    for _ in range(3):
        y = t.transform(xin)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        y = t.transform(xin)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    byt = B * (10000 * xin.shape[2] * 4 + y.shape[1] * y.shape[2] * 4)
    print(kw, f"{ms:.3f} ms  {B / ms * 1e3:.0f} rec/s  {byt / ms / 1e6:.0f} GB/s", flush=True)
