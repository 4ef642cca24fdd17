"""Per-region stall breakdown of an ncu report: regions given as start-end SASS line ranges."""
import csv, io, subprocess, sys

rep = sys.argv[1]
ranges = [tuple(int(v) for v in r.split("-")) for r in sys.argv[2:]]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
iS = hdr.index("# Samples")
iI = hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(x[iS]) for x in data)
for lo, hi in ranges:
    seg = data[lo:hi + 1]
    n = sum(int(x[iS]) for x in seg)
    ins = sum(int(x[iI]) for x in seg)
    st = {h: sum(int(x[i] or 0) for x in seg) for i, h in stall_cols}
    top = sorted(st.items(), key=lambda t: -t[1])[:7]
    print(f"lines {lo}-{hi}: samples {100*n/tot:.1f}%  instr {ins}  " + " ".join(f"{k[6:]}={100*v/max(n,1):.0f}%" for k, v in top))
