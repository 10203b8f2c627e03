#!/bin/bash
# Round-2 batch 6: persistent-lane variants of both RayCast kernels (blocks per SM, steps per ballot) and the 4-wide closest-hit walk.
python tools/rc_ab.py \
  PTB_WIDE=1 \
  PTB_SH_MINB=9 \
  PTB_SH_MINB=8 \
  PTB_SH_STEPS=8 \
  PTB_SH_STEPS=2 \
  PTB_SH_MINB=9,PTB_SH_STEPS=8 \
  PTB_RC_STEPS=8 \
  PTB_RC_MINB=9 \
  PTB_RC_MODE=3,PTB_RC_MINB=8 \
  PTB_RC_MODE=3,PTB_RC_MINB=9 \
  PTB_RC_MODE=3,PTB_RC_MINB=10 \
  PTB_RC_MODE=3,PTB_RC_MINB=8,PTB_RC_REFILL=16 \
  PTB_RC_MODE=3,PTB_RC_MINB=8,PTB_RC_STEPS=2 \
  > gpurun_out/rc_batch6.log 2>&1
cat gpurun_out/rc_batch6.log
