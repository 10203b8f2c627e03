#!/bin/bash
# Round-2 batch 4: Shade's material kernels as parallel graph branches, with the extension-ray scheduler variants.
python tools/rc_ab.py \
  PTB_SHADE_FORK=0 \
  PTB_SHADE_FORK=1 \
  PTB_SHADE_FORK=1,PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_RC_STEPS=4,PTB_FUSE_LOGIC=0 \
  PTB_SHADE_FORK=1,PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_RC_STEPS=4 \
  PTB_SHADE_FORK=1,PTB_RC_MODE=1,PTB_RC_REFILL=28,PTB_RC_STEPS=4,PTB_FUSE_LOGIC=0 \
  PTB_SHADE_FORK=1,PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_RC_STEPS=4,PTB_FUSE_LOGIC=0,PTB_RC_B0=1 \
  PTB_SHADE_FORK=1,PTB_FUSE_LOGIC=0 \
  PTB_SHADE_FORK=1 \
  > gpurun_out/rc_batch4.log 2>&1
cat gpurun_out/rc_batch4.log
