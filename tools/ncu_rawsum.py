import csv,sys
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','launch__grid_size','launch__block_size','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed.sum','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','lts__t_sector_hit_rate.pct']
short=['dur','rd','wr','grid','blk','regs','warps%','issue%','dram%','winst','fma%','xu%','l2hit']
for f in sys.argv[1:]:
    rows=list(csv.reader(open(f)))
    hdr,units=rows[0],rows[1]
    idx={h:i for i,h in enumerate(hdr)}
    print(f)
    for r in rows[2:]:
        name=r[idx['Kernel Name']][:34]
        vals=[]
        for w,s in zip(want,short):
            if w in idx:
                v=r[idx[w]]; u=units[idx[w]]
                try: v=f"{float(v.replace(',','')):.4g}"
                except: pass
                vals.append(f"{s}={v}{u if s in('dur','rd','wr') else ''}")
        print("  ",name.ljust(34)," ".join(vals))
