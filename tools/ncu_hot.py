#!/usr/bin/env python
"""Development tool: per-instruction stall samples from an Nsight Compute report's source page.

    python tools/ncu_hot.py report.ncu-rep [launch] [top]

Prints the SASS lines with the most warp-stall samples (with executed-instruction counts and the dominant stall
reasons) and the totals per 100-line window, so a hot loop can be told apart from prologue / spin loops."""
import csv, io, subprocess, sys
rep = sys.argv[1]; launch = sys.argv[2] if len(sys.argv) > 2 else "0"; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", launch, "--launch-count", "1"], capture_output=True, text=True).stdout
lines = out.split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
hdr = rows[0]; col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[1:]):
    if len(r) < len(hdr): continue
    def f(k):
        try: return float(r[col[k]])
        except Exception: return 0.0
    data.append((n, r[col["Source"]].strip(), f("Warp Stall Sampling (All Samples)"), f("Instructions Executed"), {s: f(s) for s in stalls}))
tot_s = sum(d[2] for d in data); tot_i = sum(d[3] for d in data)
print(f"total samples {tot_s:.0f}, warp instructions {tot_i:.0f}, sass lines {len(data)}")
agg = {}
for d in data:
    for s, v in d[4].items(): agg[s] = agg.get(s, 0) + v
print("stall totals:", ", ".join(f"{k[6:]}={v:.0f}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0.01 * tot_s))
print("\n-- top lines by samples")
for d in sorted(data, key=lambda d: -d[2])[:top]:
    dom = ", ".join(f"{k[6:]}={v:.0f}" for k, v in sorted(d[4].items(), key=lambda kv: -kv[1])[:3] if v > 0)
    print(f"{d[0]:5d} {d[2]:7.0f} smp {d[3]:9.0f} ins  {d[1][:58]:58s} {dom}")
print("\n-- per 100-line window")
for a in range(0, len(data), 100):
    w = data[a:a + 100]
    print(f"{a:5d}..{a+len(w)-1:5d}: samples {sum(d[2] for d in w):7.0f} ({100*sum(d[2] for d in w)/max(tot_s,1):5.1f}%)  instr {sum(d[3] for d in w):10.0f} ({100*sum(d[3] for d in w)/max(tot_i,1):5.1f}%)")
