import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[0]; units=rows[1]
want=['Kernel Name','launch__grid_size','launch__registers_per_thread','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__bytes_read.sum.per_second','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','lts__t_bytes.sum','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','sm__warps_active.avg.pct_of_peak_sustained_active','sm__throughput.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_fp64.sum','smsp__inst_executed.sum','sm__cycles_elapsed.avg.per_second','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__cycles_active.avg','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
idx=[(w,hdr.index(w)) for w in want if w in hdr]
for r in rows[2:]:
    for w,i in idx: print(f"{w:70s} {r[i][:70]} {units[i]}")
    print('---')
