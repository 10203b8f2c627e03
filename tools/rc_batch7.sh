#!/bin/bash
# Round-2 batch 7: around the new default (4-wide closest-hit walk with persistent lanes, 8 blocks/SM).
python tools/rc_ab.py \
  PTB_WIDE=1 \
  PTB_RC_REFILL=28 \
  PTB_RC_REFILL=20 \
  PTB_OVERLAP_SHADOW=0 \
  PTB_RC_B0=1 \
  PTB_WIDE=3 \
  PTB_SH_MINB=9 \
  PTB_SH_MINB=12 \
  PTB_FUSE_LOGIC=1 \
  PTB_FUSE_LOGIC=0 \
  PTB_WIDE=1 \
  > gpurun_out/rc_batch7.log 2>&1
cat gpurun_out/rc_batch7.log
tools/rc_ncu.sh nc_wide8 PTB_WIDE=1 | awk "NR==1 || NR%2==0"
