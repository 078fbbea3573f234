"""Prints headline metrics + warp-stall breakdown of every kernel in an .ncu-rep (reads `ncu --page raw --csv`)."""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[0]
keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "sm__inst_executed.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__cycles_elapsed.avg.per_second", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
for r in rows[2:]:
    print(r[h.index("Kernel Name")][:90])
    for k in keys:
        if k in h:
            print(f"  {k:68s} {r[h.index(k)]} {rows[1][h.index(k)]}")
    st = [(float(r[i]), n.split("issue_stalled_")[1].split("_per")[0]) for i, n in enumerate(h)
          if "smsp__average_warps_issue_stalled" in n and "not_issued" not in n and r[i]]
    print("  stalls (warps per issue):", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:9]))
