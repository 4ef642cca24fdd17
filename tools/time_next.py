"""Timing of the rows (f)2-(f)4 kernels alone (run on the GPU box)."""
import sys; sys.path.insert(0, ".")
import torch, b200spec
g = torch.Generator(device="cuda").manual_seed(1)
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
which = sys.argv[1:] or ["f2", "f3", "f4"]
if "f3" in which:
    win = torch.rand((1024, 400, 300), device="cuda", generator=g) * 40
    print("f3 ms/1024:", timed(lambda: b200spec.process_spectrogram(win, check=False)))
    del win
if "f2" in which:
    img = torch.rand((1024, 19, 33, 2000), device="cuda", generator=g)
    w = torch.randn((19, 16, 3, 3), device="cuda") * 0.1; b = torch.zeros(16, device="cuda")
    print("f2 ms/1024:", timed(lambda: b200spec.spectrogram_condition(img, w, b, 2000)))
    del img
    class Cfg:
        diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}
    for B in (64, 1024):
        x = torch.randn((B, 19, 2000), device="cuda", generator=g)
        print(f"f2 fused stft_condition ms/{B}:", timed(lambda: b200spec.stft_condition(x, Cfg, w, b, 2000)),
              " two calls:", timed(lambda: b200spec.spectrogram_condition(b200spec.compute_stft(x, Cfg), w, b, 2000)))
if "f4" in which:
    for B in (1024, 4096):
        x = torch.randn((B, 10000, 20), device="cuda", generator=g).cumsum_(1)
        print(f"f4 ms/{B}:", timed(lambda: b200spec.process_eeg(x, check=False), 4))
        del x
