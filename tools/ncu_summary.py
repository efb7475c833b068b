#!/usr/bin/env python
"""Summarise an .ncu-rep (raw + SASS pages) into the few numbers DESIGN.md / profiles/ quote."""
import collections, csv, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum"]
for kernel in rows[2:]:
    print("kernel:", kernel[hdr.index("Kernel Name")][:100])
    for k in keys:
        if k in hdr:
            print(f"  {k:75s} {kernel[hdr.index(k)]:>18s} {units[hdr.index(k)]}")
    st = {h: float(kernel[i]) for i, h in enumerate(hdr) if h.startswith("smsp__pcsamp_warps_issue_stalled") and kernel[i]}
    tot = sum(st.values()) or 1
    print("  stall samples:", ", ".join(f"{k.split('stalled_')[1]} {100 * v / tot:.0f}%" for k, v in sorted(st.items(), key=lambda x: -x[1])[:8]))
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(src.splitlines()))[2:]
    tot = sum(int(r[2]) for r in srows)
    byop = collections.Counter(); ex = collections.Counter()
    for r in srows:
        toks = r[1].split()
        op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
        byop[op] += int(r[2]); ex[op] += int(r[5])
    print("  SASS instrs", len(srows), "samples", tot)
    for op, c in byop.most_common(12):
        print(f"    {op:8s} {100 * c / tot:5.1f}% samples, {ex[op]:>12d} warp-instrs")
    cum = 0
    out = []
    for i, r in enumerate(srows):
        cum += int(r[2])
        if i % 100 == 99:
            out.append(f"{i + 1}:{100 * cum / tot:.0f}%")
    print("  cumulative samples by SASS index:", " ".join(out))
