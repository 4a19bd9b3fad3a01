"""selected metrics of an `ncu --set full` report as one CSV row per launch (profiles/*_ncu_full_selected.csv)
usage: ncu -i x.ncu-rep --page raw --csv | python tools/ncu_select.py > out.csv"""
import csv, sys
KEEP = ("ID", "Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "smsp__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__occupancy_limit_",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.avg.per_cycle_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "launch__shared_mem_per_block")
rows = list(csv.reader(sys.stdin))
H = rows[0]
cols = [i for i, h in enumerate(H) if any(h == k or (k.endswith("_") and h.startswith(k)) for k in KEEP)
        and not h.endswith("_not_issued_per_issue_active.ratio") and ".pct_of_peak_sustained_elapsed" not in h.replace("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "")]
w = csv.writer(sys.stdout)
w.writerow([H[i] for i in cols])
for r in rows[2:]:
    if len(r) >= len(H):
        w.writerow([r[i] for i in cols])
