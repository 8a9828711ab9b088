#!/usr/bin/env python
"""Summarise an Nsight Compute report (`--set full`) into the few numbers DESIGN.md argues with.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_<kernel>_full.md
    python tools/ncu_summary.py --launches gpurun_out/launches.csv > profiles/rNN_launches.md
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__bytes_read.sum.per_second", "dram read rate"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("sm__cycles_active.avg", "SM active cycles"),
    ("sm__cycles_elapsed.avg", "SM elapsed cycles"),
    ("sm__inst_executed.avg.per_cycle_active", "IPC (active)"),
    ("sm__inst_executed.avg.per_cycle_elapsed", "IPC (elapsed)"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % (active)"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem wavefronts % of peak (elapsed)"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
]
STALLS = "smsp__average_warps_issue_stalled_"


def full(path: str) -> None:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu --set full summary of `{path.split('/')[-1]}` ({len(data)} launches)\n")
    print("| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(data))) + " |")
    print("|---|---|" + "---|" * len(data))
    for name in ("Kernel Name", "Grid Size", "Block Size"):
        if name in col:
            print(f"| {name} | | " + " | ".join(r[col[name]].replace("|", "/")[:60] for r in data) + " |")
    for key, label in KEYS:
        if key in col:
            i = col[key]
            print(f"| {label} (`{key}`) | {units[i]} | " + " | ".join(r[i][:12] for r in data) + " |")
    print("\nWarp stall reasons (cycles per issued instruction):\n")
    print("| stall | " + " | ".join(f"launch {i}" for i in range(len(data))) + " |")
    print("|---|" + "---|" * len(data))
    for h, i in sorted(col.items()):
        if h.startswith(STALLS) and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
            vals = [float(r[i] or 0) for r in data]
            if max(vals) >= 0.2:
                print(f"| {h[len(STALLS):-len('_per_issue_active.ratio')]} | " + " | ".join(f"{v:.2f}" for v in vals) + " |")


def launches(path: str) -> None:
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        k = r[4].split("(")[0].replace("void ", "").replace("b200::", "") + " grid " + r[8]
        agg[k][0] += 1
        agg[k][1] += float(r[-1])
    tot = sum(v[1] for v in agg.values())
    print(f"# ncu launch list `{path.split('/')[-1]}`: {len(rows)} launches, {tot / 1e3:.1f} us summed (serialised, cold cache)\n")
    print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {v[0]} | {v[1] / 1e3:.1f} | {v[1] / tot:.3f} | {v[1] / v[0] / 1e3:.2f} |")


if __name__ == "__main__":
    if sys.argv[1] == "--launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[1])
