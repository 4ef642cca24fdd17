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
t0 = tl[0, 8, 8]
print("tile | CTA0: start main arrive xch folded flushed | CTA1: same   (ns, globaltimer)")
for j in range(8, 24):
    r = []
    for c in range(2):
        r += [tl[c, j, s] - t0 for s in (8, 12, 13, 9, 10, 11)]
    print(j, " ".join(f"{v:7d}" for v in r[:6]), " | ", " ".join(f"{v:7d}" for v in r[6:]))
