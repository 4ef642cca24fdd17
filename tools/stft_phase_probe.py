"""Time compute_stft on (1024,20,10000) under the EEGSPEC_STFT_DEBUG phase switches (set by the caller)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200spec

class Cfg:
    diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

x = torch.randn(1024, 20, 10000, device="cuda").cumsum(-1).mul_(0.3)
for _ in range(3):
    y = b200spec.compute_stft(x, Cfg)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    y = b200spec.compute_stft(x, Cfg)
e1.record()
torch.cuda.synchronize()
print("EEGSPEC_STFT_DEBUG=%s: %.3f ms" % (os.environ.get("EEGSPEC_STFT_DEBUG", "0"), e0.elapsed_time(e1) / 10))
