"""Attainable HBM bandwidth by direction on this box: pure write (fill), pure read (sum), copy."""
import torch

def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3

n = 5_406_720_000 // 4           # the mode-R image of 1024 records
a = torch.empty(n, dtype=torch.float32, device="cuda")
b = torch.empty(n, dtype=torch.float32, device="cuda")
t = timed(lambda: a.fill_(1.0)); print(f"fill  {4*n/t/1e12:.2f} TB/s written ({t*1e3:.3f} ms)")
t = timed(lambda: a.zero_());    print(f"zero  {4*n/t/1e12:.2f} TB/s written (cudaMemset path) ({t*1e3:.3f} ms)")
t = timed(lambda: a.sum());      print(f"sum   {4*n/t/1e12:.2f} TB/s read ({t*1e3:.3f} ms)")
t = timed(lambda: b.copy_(a));   print(f"copy  {8*n/t/1e12:.2f} TB/s read+written ({t*1e3:.3f} ms)")
