import os, sys
import numpy as np, torch
sys.path.insert(0, ".")
os.environ["EEGSPEC_TIMELINE"] = "/tmp/tl.bin"; os.environ["EEGSPEC_TIMELINE_J0"] = "96"
import b200spec
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn((1024, 20, 10000), generator=g, device="cuda")
x = (x.cumsum(-1) * 0.3 + torch.randn(x.shape, generator=g, device="cuda")).clamp_(-1024, 1024)
for _ in range(2): y = b200spec.eeg_to_spectrogram(x)
torch.cuda.synchronize()
tl = np.fromfile("/tmp/tl.bin", dtype=np.int64).reshape(2, 64, 32)
for c in range(2):
    print("CTA", c, "arrival of warps 0..15 relative to the tile's e.mmadone; then xch")
    for j in range(8, 20):
        b = tl[c, j, 8]
        print(j, " ".join(f"{tl[c, j, 16 + w] - b:5d}" for w in range(16)), "| xch", tl[c, j, 9] - b, "| peer start", tl[1 - c, j, 8] - b)
    d = tl[c, 8:56, 16:32] - tl[c, 8:56, 8:9]
    print("mean", " ".join(f"{v:5.0f}" for v in d.mean(0)), "| xch", (tl[c, 8:56, 9] - tl[c, 8:56, 8]).mean())
