#!/usr/bin/env python
"""Key metrics per kernel launch of an .ncu-rep:  python profiles/ncu_summary.py <file.ncu-rep> [--md]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = [("gpu__time_duration.sum", "time"), ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_cyc%"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "uni%"),
        ("smsp__inst_executed.sum", "warp_inst"),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%")]
ki = hdr.index("Kernel Name")
for r in data:
    out = [r[ki].split("(")[0].replace("void ssk::", "")]
    for m, short in want:
        if m in hdr:
            i = hdr.index(m); v = r[i]
            try: v = f"{float(v):.4g}"
            except ValueError: pass
            out.append(f"{short}={v}{units[i] if short in ('time', 'dram_rd', 'dram_wr') else ''}")
    print("  ".join(out))
