"""Condenses an .ncu-rep (ncu --set full) into the few per-kernel numbers DESIGN.md quotes.
Usage: python tools/ncu_summary.py report.ncu-rep > profiles/<name>.csv"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("smsp__warps_active.avg.per_cycle_active", "warps_per_smsp"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "pipe_fma_pct"),
    ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "fmaheavy_cycles_pct"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "pipe_alu_pct"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "pipe_xu_pct"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "pipe_lsu_pct"),
    ("smsp__inst_executed.sum", "warp_instructions"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("smsp__sass_inst_executed_op_local_ld.sum", "local_loads"),
]
STALL = "smsp__average_warps_issue_stalled_"

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
w = csv.writer(sys.stdout)
w.writerow(["kernel"] + [k[1] for k in KEYS] + ["top_stalls_per_issue"])
for r in rows[2:]:
    out = [r[col["Kernel Name"]][:90]]
    for key, _ in KEYS:
        out.append("%s %s" % (r[col[key]], units[col[key]]) if key in col else "")
    st = sorted(((float(r[i] or 0), h[len(STALL):-len("_per_issue_active.ratio")]) for h, i in col.items()
                 if h.startswith(STALL) and h.endswith("_per_issue_active.ratio")), reverse=True)[:5]
    out.append("; ".join("%s %.2f" % (n, v) for v, n in st))
    w.writerow(out)
