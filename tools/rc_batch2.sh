#!/bin/bash
# Round-2 batch 2: two-wide fp32 node test (new default build) against the round-1 library, then scheduler knobs in-frame.
set -x
python tools/lib_ab.py ab/base_r01.so cuda_path_tracer_b200/libptb200.so > gpurun_out/rc_batch2_lib.log 2>&1
python tools/rc_ab.py \
  PTB_RC_MODE=0 \
  PTB_RC_MODE=0,PTB_FUSE_LOGIC=0 \
  PTB_RC_MODE=0,PTB_OVERLAP_SHADOW=0 \
  PTB_RC_MODE=1,PTB_RC_REFILL=32 \
  PTB_RC_MODE=1,PTB_RC_REFILL=24 \
  PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_FUSE_LOGIC=0 \
  PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_OVERLAP_SHADOW=0 \
  PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_RC_MINB=9 \
  PTB_RC_MODE=1,PTB_RC_REFILL=24,PTB_RC_MINB=8 \
  PTB_RC_MODE=1,PTB_RC_REFILL=20 \
  PTB_RC_MODE=1,PTB_RC_REFILL=28 \
  > gpurun_out/rc_batch2.log 2>&1
cat gpurun_out/rc_batch2_lib.log; cat gpurun_out/rc_batch2.log
