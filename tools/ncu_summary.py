#!/usr/bin/env python
"""Summarise an .ncu-rep: per-kernel headline metrics + top stall reasons / hottest instructions from the source page.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [n_top]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "launch__grid_size", "launch__block_size"]
idx = [(w, next((i for i, h in enumerate(hdr) if h.endswith(w)), None)) for w in want]
for r in rows[2:]:
    print("-" * 100)
    for w, i in idx:
        if i is not None:
            print(f"  {w:75s} {r[i][:90]} {rows[1][i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = src.split('"Kernel Name",')
for blk in blocks[1:]:
    lines = list(csv.reader(io.StringIO('"Kernel Name",' + blk)))
    name = lines[0][1][:90]
    h = lines[1]
    data = [r for r in lines[2:] if len(r) == len(h)]
    isrc, isamp, iex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    stall = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[isamp]) for r in data)
    agg = {}
    for r in data:
        for i, c in stall:
            agg[c] = agg.get(c, 0) + int(r[i])
    print("=" * 100)
    print(name, "| samples", tot, "| SASS instr", len(data), "| executed", sum(int(r[iex]) for r in data))
    print("  stalls:", ", ".join(f"{k[6:]} {100*v/max(tot,1):.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    ops = {}
    for r in data:
        op = r[isrc].split()[0] if r[isrc].split() else "?"
        if op.startswith("@"):
            op = r[isrc].split()[1]
        op = op.split(".")[0]
        o = ops.setdefault(op, [0, 0])
        o[0] += int(r[iex]); o[1] += int(r[isamp])
    print("  op mix (executed, samples):", ", ".join(f"{k} {v[0]/1e6:.1f}M/{v[1]}" for k, v in sorted(ops.items(), key=lambda kv: -kv[1][0])[:10]))
    for r in sorted(data, key=lambda r: -int(r[isamp]))[:ntop]:
        st = sorted(((int(r[i]), c[6:]) for i, c in stall), reverse=True)[:2]
        print(f"   {r[isamp]:>6} {r[iex]:>10}  {r[isrc][:64]:64s} {st}")
