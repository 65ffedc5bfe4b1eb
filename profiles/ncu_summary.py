import csv, sys, subprocess, io
rep = sys.argv[1]
out = subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
r=list(csv.reader(io.StringIO(out)))
hdr=r[0]; units=r[1]
want=['gpu__time_duration.sum','launch__registers_per_thread','launch__grid_size','launch__block_size','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','sm__warps_active.avg.pct_of_peak_sustained_active','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.sum','sm__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed','smsp__thread_inst_executed_per_inst_executed.ratio','sm__cycles_elapsed.max','dram__bytes_read.sum','dram__bytes_write.sum']
stalls=[h for h in hdr if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
ki=hdr.index('Kernel Name')
for row in r[2:]:
    print('-----', row[ki][:60])
    for w in want:
        if w in hdr:
            i=hdr.index(w); print(f"  {w:75s} {row[i]:>16s} {units[i]}")
    st=sorted(((float(row[hdr.index(h)]),h) for h in stalls), reverse=True)[:7]
    print("  stalls:", ", ".join(f"{h.split('stalled_')[1].split('_per_')[0]}={v:.2f}" for v,h in st))
