#!/bin/bash
# ncu: DRAM bytes / duration of the diffuse Shade kernel and both RayCast kernels inside whole frames (tools/wave_frames.py).
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed
ncu --metrics $M --clock-control none -k regex:"k_shade|k_trace" -s 12 -c 12 --csv --log-file gpurun_out/$1.csv python tools/wave_frames.py 2 > gpurun_out/$1.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/$1.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
d={}
for r in rows[1:]: d.setdefault((r[ii],r[ki]),{})[r[mi]]=float(r[vi].replace(",",""))
for (i,k),v in d.items():
    print("%-46s us %7.1f  rd %6.1f MB wr %6.1f MB  issue %4.1f occ %4.1f lanes %4.1f l1 %4.1f" % (k[:46], v["gpu__time_duration.sum"]/1e3, v["dram__bytes_read.sum"]/1e6, v["dram__bytes_write.sum"]/1e6,
      v["smsp__issue_active.avg.pct_of_peak_sustained_active"], v["sm__warps_active.avg.pct_of_peak_sustained_active"], v["smsp__thread_inst_executed_per_inst_executed.ratio"], v["l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"]))
PY
