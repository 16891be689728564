"""Key metrics, stall reasons and opcode mix from the CSV pages exported on the GPU box by tools/r2_ncu.sh.
usage: python tools/ncu_csv.py <raw.csv> [<source.csv> <units per launch (sample-steps)> [kernel regex]]"""
import collections
import csv
import re
import sys

raw = list(csv.reader(open(sys.argv[1])))
H, U = raw[0], raw[1]
idx = {h: i for i, h in enumerate(H)}
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
        "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
        "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum", "sm__sass_inst_executed_op_local.sum"]
for r in raw[2:]:
    print(f"## `{r[idx['Kernel Name']][:120]}`\n")
    print("| metric | value | unit |\n|---|---|---|")
    for w in want:
        if w in idx:
            print(f"| {w} | {r[idx[w]]} | {U[idx[w]]} |")
    d = {}
    for h in H:
        if "issue_stalled" in h and "ratio" in h and "not_issued" not in h:
            try:
                d[h] = float(r[idx[h]])
            except ValueError:
                pass
    top = sorted(d.items(), key=lambda kv: -kv[1])[:8]
    print("\nTop stall reasons (warps per issue-active cycle): " +
          ", ".join(f"{k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')} {v:.2f}" for k, v in top) + "\n")

if len(sys.argv) > 3:
    units = float(sys.argv[3])
    rx = re.compile(sys.argv[4]) if len(sys.argv) > 4 else None
    rows = list(csv.reader(open(sys.argv[2])))
    # the source page holds one table per kernel: "Kernel Name" marker rows separate them
    cur, tables = None, {}
    hdr = None
    for r in rows:
        if not r:
            continue
        if r[0] == "Kernel Name":
            cur = r[1] if len(r) > 1 else "?"
            tables[cur] = []
            hdr = None
            continue
        if r[0] == "Address" or (hdr is None and "Source" in r):
            hdr = r
            tables.setdefault(cur, [])
            tables[cur].append(("H", r))
            continue
        if cur is not None or hdr is not None:
            tables.setdefault(cur, []).append(("R", r))
    for name, tab in tables.items():
        if rx and name and not rx.search(name):
            continue
        hdr = [r for k, r in tab if k == "H"]
        if not hdr:
            continue
        hdr = hdr[0]
        si, ei = hdr.index("Source"), hdr.index("Instructions Executed")
        ti = hdr.index("Thread Instructions Executed") if "Thread Instructions Executed" in hdr else None
        ops, tot, ttot = collections.Counter(), 0, 0
        for k, r in tab:
            if k != "R" or len(r) <= ei:
                continue
            try:
                e = int(r[ei])
            except ValueError:
                continue
            tot += e
            if ti is not None:
                ttot += int(r[ti])
            toks = [o for o in r[si].split() if not o.startswith("@")]
            if toks:
                ops[toks[0].split(".")[0]] += e
        print(f"## `{str(name)[:100]}`: executed warp-instructions {tot}; thread-instructions per unit = {ttot / units:.1f}; warp-instr x 32 per unit = {tot * 32 / units:.1f}\n")
        print("| opcode | warp-instr | thread-slots per unit |\n|---|---|---|")
        for op, c in ops.most_common(28):
            print(f"| {op} | {c} | {c * 32 / units:.2f} |")
