"""Summarise an .ncu-rep (raw page + source page) into markdown: key metrics per kernel, stall
reasons, and the executed instruction mix per sample-step.  Usage:
  python tools/ncu_summary.py <rep> <warp_steps_per_launch_of_rollout> > profiles/xxx.md"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
warp_steps = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H, U = rows[0], rows[1]
idx = {h: i for i, h in enumerate(H)}
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "smsp__cycles_active.avg",
        "sm__cycles_elapsed.max", "smsp__inst_executed.sum"]
print(f"# ncu summary of `{rep.split('/')[-1]}` (`--set full --clock-control none`, cold-cache serialised replays)\n")
for r in rows[2:]:
    print(f"## `{r[idx['Kernel Name']][:110]}`\n")
    print("| metric | value | unit |\n|---|---|---|")
    for w in want:
        if w in idx:
            print(f"| {w} | {r[idx[w]]} | {U[idx[w]]} |")
    d = {}
    for h in H:
        if "issue_stalled" in h and "ratio" in h:
            try:
                d[h] = float(r[idx[h]])
            except ValueError:
                pass
    top = sorted(d.items(), key=lambda kv: -kv[1])[:8]
    print("\nTop stall reasons (warps per issue-active cycle): " +
          ", ".join(f"{k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')} {v:.2f}" for k, v in top) + "\n")

if warp_steps:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:rollout"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    H = rows[1]
    si, ei = H.index("Source"), H.index("Instructions Executed")
    ops, tot = collections.Counter(), 0
    for r in rows[2:]:
        if r and r[0] == "Kernel Name":
            break
        if len(r) <= ei or r[0] == "Address":
            continue
        e = int(r[ei])
        tot += e
        ops[[o for o in r[si].split() if not o.startswith("@")][0].split(".")[0]] += e
    print(f"## rollout kernel: executed warp-instructions per warp-step = {tot / warp_steps:.1f} (total {tot})\n")
    print("| opcode | per warp-step |\n|---|---|")
    for op, c in ops.most_common(24):
        print(f"| {op} | {c / warp_steps:.1f} |")
