"""Key metrics per kernel from `ncu --page raw --csv`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]; units = rows[1]
want = ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread',
 'launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','sm__throughput.avg.pct_of_peak_sustained_elapsed',
 'l1tex__throughput.avg.pct_of_peak_sustained_active','lts__throughput.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct_of_peak_sustained_active',
 'launch__grid_size','launch__block_size','smsp__inst_executed.sum','lts__t_sector_hit_rate.pct','sm__cycles_elapsed.avg.per_second','l1tex__m_xbar2l1tex_read_bytes.sum']
seen = set()
for r in rows[2:]:
    k = r[hdr.index('Kernel Name')][:40]
    if k in seen: continue
    seen.add(k)
    print(k)
    for w in want:
        if w in hdr:
            i = hdr.index(w); print(f"   {w:75s} {r[i][:24]:>24s} {units[i]}")
