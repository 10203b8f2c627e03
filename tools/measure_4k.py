"""BASELINE config 5: 3840x2160, 64 spp, tile-partitioned render of the 262k-triangle textured scene across the GPUs of
one box with one NCCL gather of the resolved tiles; rank 0 then renders the same frame alone and compares bit for bit.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/measure_4k.py"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import multigpu, scenegen

# SCENE=stress: BASELINE config 4 instead (10,000,000-triangle height field from arrays, 1920x1080, 4 spp)
STRESS = os.environ.get("SCENE", "sponza") == "stress"
SPP, RES, STRIPE = int(os.environ.get("SPP", "4" if STRESS else "64")), ((1920, 1080) if STRESS else (3840, 2160)), 8
rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if STRESS:
    pos, nor, idx, mat = scenegen.heightfield_arrays(10_000_000)
    scene = ptb.Scene.from_arrays(pos, nor, idx, mat, scenegen.stress_materials(), scenegen.stress_camera(RES))
else:
    scene = ptb.Scene(scenegen.sponza(f"/tmp/m4k_{rank}", res=RES)[0])
W, H = RES
stream = torch.cuda.Stream()
with torch.cuda.stream(stream):
    r = ptb.Renderer(scene, device=local, rank=rank, world=world, stripe_rows=STRIPE)
    r.set_stream(stream.cuda_stream)

    class _One:                      # world == 1: same interface without a process group
        send = None
    if world > 1:
        fg = multigpu.FrameGather(W, H, STRIPE, dist, device="cuda")
        tile = fg.tile
    else:
        tile = torch.empty((W * H, 3), dtype=torch.float32, device="cuda")

    def render():
        r.render(1, SPP)
        r.resolve_device(tile.data_ptr(), SPP)
        return fg.gather() if world > 1 else tile

    render()                          # warm-up (graph upload, NCCL channels)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    frame = render()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        out = {"config": ("4: 1920x1080, %d spp, 10,000,000-triangle stress scene" if STRESS else "5: 3840x2160, %d spp, 262,279-triangle textured scene") % SPP, "n_gpus": world, "ms": round(float(ms[0]), 2),
               "mpaths_per_s": round(W * H * SPP / float(ms[0]) / 1e3, 1), "partition": f"interleaved {STRIPE}-row stripes, one gather"}
        if world > 1:
            one = ptb.Renderer(scene, device=local)
            one.set_stream(stream.cuda_stream)
            full = torch.empty((W * H, 3), dtype=torch.float32, device="cuda")
            one.render(1, 2); torch.cuda.synchronize()
            t0 = time.perf_counter()
            one.render(1, SPP)
            one.resolve_device(full.data_ptr(), SPP)
            torch.cuda.synchronize()
            t1 = time.perf_counter() - t0
            out["one_gpu_ms"] = round(t1 * 1e3, 2)
            out["one_gpu_mpaths_per_s"] = round(W * H * SPP / t1 / 1e6, 1)
            out["speedup"] = round(t1 * 1e3 / float(ms[0]), 2)
            out["bit_identical_to_one_gpu"] = bool(torch.equal(frame.view(torch.int32), full.view(torch.int32)))
        print("MEASURE4K " + json.dumps(out), flush=True)
    r.free()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
