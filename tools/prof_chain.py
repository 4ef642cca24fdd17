"""Profiling driver: a few launches of one mode-N engine on a mid-size batch (for ncu)."""
import sys
import torch
sys.path.insert(0, ".")
import b200spec

engine = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = int(sys.argv[2]) if len(sys.argv) > 2 else 296
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn((B, 20, 10000), generator=g, device="cuda")
x = (x.cumsum(-1) * 0.3 + torch.randn((B, 20, 10000), generator=g, device="cuda")).clamp_(-1024, 1024)
p = b200spec.mel_params(engine=engine)
for _ in range(3):
    y = b200spec.eeg_to_spectrogram(x, p)
torch.cuda.synchronize()
print("ok", float(y.mean()))
