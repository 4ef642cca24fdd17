"""Mode N on the parquet layout (time-major records) against channel-major (run on the GPU box)."""
import sys
import torch
sys.path.insert(0, ".")
import b200spec

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
g = torch.Generator(device="cuda").manual_seed(3)
x = torch.randn((B, 20, 10000), generator=g, device="cuda") * 40
xt = x.transpose(1, 2).contiguous()
for name, inp, kw in (("channel-major", x, {}), ("time-major", xt, {"time_major": True})):
    for _ in range(3):
        y = b200spec.eeg_to_spectrogram(inp, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = b200spec.eeg_to_spectrogram(inp, **kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(name, f"{ms:.3f} ms  {B / ms * 1e3:.0f} rec/s", flush=True)
    if name == "channel-major":
        ref = y
print("max |tm - cm|", (y - ref).abs().max().item())
