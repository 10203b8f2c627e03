#!/usr/bin/env python
"""profiles/raycast_traffic.json from an ncu capture of tools/rc_ncu.sh (the extension RayCast kernel launched alone once per bounce
of one staged frame of the bench scene): DRAM bytes, warp instructions, active lanes, issue / L1 / L2 figures per bounce, and the
hash of the kernel sources the capture belongs to (bench.py reports `roofline.traffic` only while that hash still matches).
usage: tools/traffic_json.py gpurun_out/TAG.csv profiles/COPY_OF_THE_CSV.csv"""
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

src, keep = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(src)) if len(r) > 10]
h = rows[0]
ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per = {}
for r in rows[1:]:
    per.setdefault((int(r[ii]), r[ki]), {})[r[mi]] = float(r[vi].replace(",", ""))
launches = sorted(per.items())
# ptb_time_raycast launches the kernel once to time it and staged TRACE launches it again: keep the first of every pair
launches = launches[0::2] if len(launches) == 2 * bench.DEPTH else launches
assert len(launches) == bench.DEPTH, len(launches)
g = lambda name: [v[name] for _, v in launches]
dram = [a + b for a, b in zip(g("dram__bytes_read.sum"), g("dram__bytes_write.sum"))]
lanes = g("smsp__thread_inst_executed_per_inst_executed.ratio")
if os.path.abspath(src) != os.path.abspath(keep):
    shutil.copyfile(src, keep)
out = {"capture": os.path.relpath(keep, ROOT) + " (tools/rc_ncu.sh: the extension RayCast kernel launched alone once per bounce of one staged frame of the bench scene)",
       "kernel_sources_sha16": bench.kernel_sources_hash(),
       "kernels": [k[1][:60] for k, _ in launches],
       "dram_bytes_per_launch": round(sum(dram) / len(dram)),
       "dram_bytes_per_bounce": [round(x) for x in dram],
       "warp_instructions_per_launch": [round(x) for x in g("smsp__inst_executed.sum")],
       "warp_instructions_source": "ncu smsp__inst_executed.sum per launch, same capture",
       "active_lanes_per_bounce": [round(x, 2) for x in lanes],
       "active_lanes_mean": round(sum(lanes) / len(lanes), 2),
       "issue_active_pct_per_bounce": [round(x, 1) for x in g("smsp__issue_active.avg.pct_of_peak_sustained_active")],
       "l1_data_pipe_pct_per_bounce": [round(x, 1) for x in g("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed")],
       "l1_hit_pct_per_bounce": [round(x, 1) for x in g("l1tex__t_sector_hit_rate.pct")],
       "l2_hit_pct_per_bounce": [round(x, 1) for x in g("lts__t_sector_hit_rate.pct")],
       "us_per_bounce_under_ncu": [round(x / 1e3, 1) for x in g("gpu__time_duration.sum")],
       "note": "the stand-alone launch (no Logic) fetches the 64-byte ray + state block of every ray and uses the ray half: 61 MB of the traffic are state "
               "records that the same kernel fused with Logic reads in the frame (there: 253 MB per launch at bounce 1 = 133 B per ray, "
               "profiles/r02_frame_kernels.csv; caches flushed by ncu before every launch, so 28 MB of tree and leaf records are fetched again)"}
json.dump(out, open(os.path.join(ROOT, "profiles", "raycast_traffic.json"), "w"), indent=1)
print(json.dumps({k: out[k] for k in ("kernel_sources_sha16", "dram_bytes_per_launch", "active_lanes_mean")}))
