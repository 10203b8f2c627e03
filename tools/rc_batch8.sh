#!/bin/bash
# Round-2 batch 8: Logic fused into the 64-register wide persistent-lane kernel, shadow kernel at 9 blocks/SM.
python tools/rc_ab.py \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9,PTB_RC_STEPS=2 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9,PTB_RC_STEPS=8 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=10 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9,PTB_SH_REFILL=28 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9,PTB_OVERLAP_SHADOW=0 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9,PTB_RC_REFILL=20 \
  PTB_FUSE_LOGIC=1,PTB_SH_MINB=9 \
  PTB_FUSE_LOGIC=2,PTB_SH_MINB=10 \
  > gpurun_out/rc_batch8.log 2>&1
cat gpurun_out/rc_batch8.log
