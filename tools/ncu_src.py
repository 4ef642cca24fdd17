"""Summarise an ncu report: headline metrics, stall reasons, hottest SASS lines / regions."""
import collections, csv, io, subprocess, sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u, r = rows[0], rows[1], rows[2]
for k in ["gpu__time_duration.sum", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
          "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
          "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg"]:
    if k in h:
        print(k, r[h.index(k)], u[h.index(k)])
st = {k.split("issue_stalled_")[1].replace("_per_issue_active.ratio", ""): float(r[i]) for i, k in enumerate(h)
      if "issue_stalled" in k and k.endswith("per_issue_active.ratio")}
print("stalls/issue:", " ".join(f"{k}={v:.2f}" for k, v in sorted(st.items(), key=lambda t: -t[1])[:9]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
iS, iI, iSrc = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
tot_s = sum(int(x[iS]) for x in data)
tot_i = sum(int(x[iI]) for x in data)
print("total instr", tot_i, "samples", tot_s)
print("--- top lines by samples")
for idx, x in sorted(enumerate(data), key=lambda t: -int(t[1][iS]))[:topn]:
    print(f"{idx:5d} {100*int(x[iS])/tot_s:5.1f}% {x[iI]:>9s}  {x[iSrc].strip()[:80]}")
print("--- regions (contiguous, same exec count)")
regs, cur = [], None
for idx, x in enumerate(data):
    c = int(x[iI])
    if cur and cur[2] == c:
        cur[1] = idx; cur[3] += c; cur[4] += int(x[iS])
    else:
        cur = [idx, idx, c, c, int(x[iS])]; regs.append(cur)
for g in sorted(sorted(regs, key=lambda g: -(g[3] + 50 * g[4]))[:topn]):
    print(f"lines {g[0]:5d}-{g[1]:5d} n={g[1]-g[0]+1:4d} exec={g[2]:8d} instr={100*g[3]/tot_i:5.1f}% samples={100*g[4]/tot_s:5.1f}%  {data[g[0]][iSrc].strip()[:40]}")
