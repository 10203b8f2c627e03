#!/bin/bash
# Round-2 profiling pass (GPU box).  Every ncu command follows a plain run of the same program that exited 0.
set -x
python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02_ncu_launches.log 2>&1
python tools/wave_frames.py 2 > gpurun_out/r02_wave_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"k_trace_extension_pl|k_trace_shadow_pl|k_shadeILi1" -s 9 -c 3 -o gpurun_out/r02_top3 -f python tools/wave_frames.py 2 > gpurun_out/r02_ncu_full.log 2>&1
ncu -i gpurun_out/r02_top3.ncu-rep --page details > gpurun_out/r02_top3_details.txt 2>&1
ncu -i gpurun_out/r02_top3.ncu-rep --page raw --csv > gpurun_out/r02_top3_raw.csv 2>&1
python tools/staged_raycast.py > gpurun_out/r02_staged_plain.log 2>&1 || exit 1
tools/rc_ncu.sh r02_ext_alone PTB_WIDE=1 > gpurun_out/r02_ext_alone_summary.txt 2>&1
tools/shade_ncu.sh r02_frame_kernels > gpurun_out/r02_frame_kernels_summary.txt 2>&1
ls -la gpurun_out | tail -20
