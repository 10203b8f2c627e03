"""Print the metrics that matter from an `ncu --page raw --csv` dump (one column per captured launch)."""
import csv
import sys
r = list(csv.reader(open(sys.argv[1])))
hdr, units = r[0], r[1]
pats = sys.argv[2:] or ['dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
                        'launch__registers_per_thread', 'launch__grid_size', 'smsp__thread_inst_executed_per_inst_executed.ratio',
                        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
                        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'lts__t_bytes.sum', 'smsp__inst_executed.sum',
                        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue_stalled', 'occupancy']
for i, h in enumerate(hdr):
    if any(p in h for p in pats):
        print(f"{h:95s} {units[i]:14s}", [x[i] for x in r[2:]])
