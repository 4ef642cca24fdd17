"""Host<->device copy bandwidth of the box next to the end-to-end call (run on the GPU box)."""
import sys, time
import torch
sys.path.insert(0, ".")
import b200spec

def bw(fn, nbytes, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return nbytes * n / (time.perf_counter() - t0) / 1e9

h_in = torch.empty((1024, 20, 10000), dtype=torch.float32).pin_memory()
h_in.normal_()
h_out = torch.empty((1024, 128, 256, 4), dtype=torch.float32).pin_memory()
d_in = torch.empty_like(h_in, device="cuda")
d_out = torch.empty_like(h_out, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
print("h2d GB/s", bw(lambda: d_in.copy_(h_in, non_blocking=True), h_in.nbytes))
print("d2h GB/s", bw(lambda: h_out.copy_(d_out, non_blocking=True), h_out.nbytes))
def both():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)
print("both GB/s (sum)", bw(both, h_in.nbytes + h_out.nbytes))
ts = []
for _ in range(8):
    t0 = time.perf_counter()
    y = b200spec.eeg_to_spectrogram(h_in)
    ts.append((time.perf_counter() - t0) * 1e3)
print("per call ms (result kept alive across the next call):", " ".join(f"{t:.1f}" for t in ts), "pinned:", y.is_pinned())
del y
ts = []
for _ in range(8):
    t0 = time.perf_counter()
    b200spec.eeg_to_spectrogram(h_in)
    ts.append((time.perf_counter() - t0) * 1e3)
print("per call ms (result dropped):", " ".join(f"{t:.1f}" for t in ts))
t0 = time.perf_counter()
o = torch.empty((1024, 128, 256, 4), dtype=torch.float32, pin_memory=True)
print("fresh/cached pinned alloc ms", (time.perf_counter() - t0) * 1e3)
