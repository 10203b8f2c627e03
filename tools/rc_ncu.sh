#!/bin/bash
# ncu metrics of the extension RayCast kernel alone, per bounce (tools/staged_raycast.py), for a set of scheduler knobs.
# usage: tools/rc_ncu.sh TAG "ENV=..."  -> gpurun_out/TAG.csv
TAG=$1; shift
M=gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts.sum
env "$@" ncu --metrics $M --clock-control none -k regex:"k_trace_extension" --csv --log-file gpurun_out/$TAG.csv python tools/staged_raycast.py > gpurun_out/$TAG.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/$TAG.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value"); ii=hdr.index("ID")
from collections import OrderedDict
d=OrderedDict()
for r in rows[1:]:
    d.setdefault(r[ii],{})[r[mi]]=r[vi]
print("$TAG", "$@")
for k,v in d.items():
    g=lambda n: float(v.get(n,"nan").replace(",",""))
    print("  us %7.1f lanes %5.2f issue%% %5.1f l1pipe%% %5.1f l1hit %5.1f l2hit %5.1f winst %6.1fM occ %5.1f dramMB %6.1f wav %6.1fM smemwav %6.1fM" % (
        g("gpu__time_duration.sum")/1e3, g("smsp__thread_inst_executed_per_inst_executed.ratio"), g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        g("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"), g("l1tex__t_sector_hit_rate.pct"), g("lts__t_sector_hit_rate.pct"),
        g("smsp__inst_executed.sum")/1e6, g("sm__warps_active.avg.pct_of_peak_sustained_active"), (g("dram__bytes_read.sum")+g("dram__bytes_write.sum"))/1e6,
        g("l1tex__data_pipe_lsu_wavefronts.sum")/1e6, g("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")/1e6))
PY
