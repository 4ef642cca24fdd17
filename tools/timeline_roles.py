"""Debug timeline of the role-split mode-N kernel (run on the GPU box).

EEGSPEC_TIMELINE makes the launch dump clock64 stamps of the pipeline hand-offs of CTAs 0 and 1;
this prints, per tile, every stamp (nanoseconds of %globaltimer) relative to the tile's MMA completion as seen by the epilogue.
"""
import os, sys
import numpy as np, torch
sys.path.insert(0, ".")
os.environ["EEGSPEC_TIMELINE"] = "/tmp/tl.bin"
os.environ.setdefault("EEGSPEC_TIMELINE_J0", "96")
import b200spec

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn((B, 20, 10000), generator=g, device="cuda")
x = (x.cumsum(-1) * 0.3 + torch.randn(x.shape, generator=g, device="cuda")).clamp_(-1024, 1024)
for _ in range(2):
    y = b200spec.eeg_to_spectrogram(x)
torch.cuda.synchronize()
tl = np.fromfile("/tmp/tl.bin", dtype=np.int64).reshape(2, 64, 40)
names = {0: "b.top", 1: "b.loaded", 2: "b.diffed", 3: "b.tilesfree", 4: "b.built", 5: "m.ops", 6: "m.tmemfree",
         7: "m.issued", 8: "e.mmadone", 9: "e.xch", 10: "e.folded", 11: "e.flushed", 12: "e.main", 13: "e.arrived",
         14: "w15.main", 15: "w15.arrived"}
names[32] = "m.seenprev"; names[33] = "where"
if os.environ.get("TL_FINE"):
    names = {5: "m.ops0", 6: "m.tmemfree", 28: "m.iss0", 25: "m.ops1", 29: "m.iss1", 16: "b2.comp", 17: "b2.free", 18: "b2.st",
             19: "b2.arr", 26: "m.ops2", 30: "m.iss2", 20: "b3.comp", 21: "b3.free", 22: "b3.st", 23: "b3.arr", 27: "m.ops3",
             31: "m.iss3", 8: "e.mmadone", 32: "m.seenprev", 33: "where"}
for cta in range(2):
    t = tl[cta]
    print(f"CTA {cta}: period (e.mmadone j+1 - j): mean {np.diff(t[8:56, 8]).mean():.0f} ns")
    print("tile " + " ".join(f"{names[s]:>11s}" for s in names))
    for j in range(8, 24):
        base = t[j, 8]
        print(f"{j:4d} " + " ".join(f"{(t[j, s] - base) if t[j, s] > 1000 else t[j, s]:11d}" for s in names))
    d = {names[s]: t[8:56, s] - t[8:56, 8] for s in names}
    print("mean " + " ".join(f"{d[names[s]].mean():11.0f}" for s in names))
