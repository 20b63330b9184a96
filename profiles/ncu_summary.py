"""Summarise an .ncu-rep (read here, no GPU): python profiles/ncu_summary.py REP [substr ...]"""
import csv, subprocess, sys
rep = sys.argv[1]
want = sys.argv[2:] or ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "fp64", "sm__throughput.avg.pct",
                        "sm__warps_active.avg.pct", "launch__registers", "launch__grid_size", "issue_active.avg.pct", "warp_issue_stalled",
                        "bank_conflicts", "lts__t_bytes.sum ", "sm__cycles_elapsed.max", "smsp__inst_executed.sum ", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum "]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
for r in rows[2:]:
    print("==", r[h.index("Kernel Name")][:70], r[h.index("Grid Size")] if "Grid Size" in h else "")
    for i, k in enumerate(h):
        if any(w.strip() in k for w in want):
            try:
                if float(r[i].replace(",", "")) == 0 and "stalled" in k: continue
            except ValueError: pass
            print(f"  {k:90s} {u[i]:12s} {r[i]}")
