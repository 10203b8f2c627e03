#!/usr/bin/env python
"""bench.py -- Mpaths/s of the wavefront render path on BASELINE.json configs[1]:
synthetic 143,173-triangle Fireplace-like room (PBR / diffuse / mirror / dielectric mix of README.md:34),
1920x1080, 8 bounces, MIS + NEE, 1 sample per pixel per step.

    python bench.py --gpus 1 --steps 20 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
    python bench.py --impl reference ...     # the reference's own wavefront build (oracle/_ref/refdump), same scene

A step = one full frame per GPU: RayGen -> 8 x (RayCast, Logic, Shade) -> resolve.  With N > 1 the ranks take
different sample indices of the same frame (rank r renders iterations r+1, r+1+N, ...) and one NCCL reduce per step
adds the resolved sums on rank 0, so a step is N samples per pixel and per-GPU work is fixed ("scaling": "weak").
The other partition north_star names -- ONE frame split into interleaved 8-row stripes, gathered to rank 0,
bit-identical to the 1-GPU frame -- is timed in the same run and reported under config.tile_partition (strong scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOAD = "fireplace-like 143173 tris (PBR 24421 / diffuse 107572 / mirror 10 / dielectric 11170), 1920x1080, 8 bounces, MIS+NEE, 1 spp/step"
RES = (1920, 1080)
REFDUMP = os.path.join(ROOT, "oracle", "_ref", "refdump")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 6 for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def make_scene(rank):
    from cuda_path_tracer_b200 import scenegen
    return scenegen.fireplace(f"/tmp/ptb_bench_fire_r{rank}", res=RES)[0]


def reference_arm(args, rank):
    """The UNMODIFIED reference wavefront renderer (built from /root/reference into oracle/_ref by oracle/Makefile),
    driven through its own PathtraceInit / Pathtrace entry points on the same scene files.  The reference is a
    single-GPU GPU program with no CPU path; it is timed on ONE GPU whatever --gpus says."""
    if rank != 0:
        return
    if not os.path.exists(REFDUMP):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/refdump not built (needs /root/reference at build time)"}))
        return
    scene = make_scene(0)
    p = subprocess.run([REFDUMP, scene, "/tmp/ptb_bench_ref", "--no-dump", "--bench", str(args.steps), str(args.warmup)],
                       capture_output=True, text=True, timeout=3000)
    line = [l for l in p.stdout.splitlines() if l.startswith("REFBENCH")]
    if p.returncode != 0 or not line:
        print(json.dumps({"impl": "reference", "unavailable": "refdump failed: " + (p.stderr or p.stdout)[-200:].replace("\n", " ")}))
        return
    rb = json.loads(line[0][len("REFBENCH "):])
    v = rb["mpaths_per_s"]
    # the reference's other front-end on the same files (secondary figure; the headline metric is the wavefront path)
    pm = subprocess.run([REFDUMP, scene, "/tmp/ptb_bench_ref", "--no-dump", "--bench", str(args.steps), str(args.warmup), "--mega"],
                        capture_output=True, text=True, timeout=3000)
    lm = [l for l in pm.stdout.splitlines() if l.startswith("REFBENCH")]
    mega = json.loads(lm[0][len("REFBENCH "):]) if lm else None
    print(json.dumps({"impl": "reference", "metric": "Mpaths/s", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": rb["ms_per_frame"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": WORKLOAD, "reference": "s64zxu/CUDA-Path-Tracer wavefront build recompiled -arch=sm_100 (oracle/Makefile), Pathtrace(NULL,0,iter) loop, wall clock around the loop",
                                 "mrays_per_s": rb["mrays_per_s"], "gpus_used": 1,
                                 "megakernel_path": None if mega is None else {"ms_per_frame": mega["ms_per_frame"], "mpaths_per_s": mega["mpaths_per_s"], "mrays_per_s": mega["mrays_per_s"]}},
                      "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": 0, "kind": "reference",
                                       "sample": "the reference has no CPU implementation of this path (SURVEY.md fact 2): this arm is its own CUDA build on one B200; the CPU port is timed in the main arm"},
                      "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def roofline_and_cpu(r, scene, n_gpus, rank, stats_rays_per_path):
    """Rank-0 post-pass on ONE staged frame: (1) k_trace_extension timed alone per bounce with CUDA events on the
    renderer's stream (ptb_time_raycast), algorithmic bytes from SURVEY.md 8(d): 36 B read per ray + 28 B written
    on a hit / 8 B on a miss; (2) the CPU port of the same traversal on a bounded ray sample."""
    import numpy as np

    from cuda_path_tracer_b200 import capi
    depth = 8
    r.stage(capi.STAGE_RAYGEN, 1)
    tot_ms, tot_bytes, tot_rays, per_bounce, sample = 0.0, 0.0, 0, [], None
    for b in range(depth):
        ms, rays = r.time_raycast(5)
        if b <= 3 and rank == 0 and n_gpus == 1:
            q = r.queue("extension")
            ori, dirs = r.path_field("ray_ori")[q], r.path_field("ray_dir_dist")[q]
            sel = np.arange(q.size)                 # whole queues of bounces 0-3: ~7.8 M rays, ~30 core-seconds
            sample = (ori[sel], dirs[sel]) if sample is None else (np.concatenate([sample[0], ori[sel]]), np.concatenate([sample[1], dirs[sel]]))
        r.stage(capi.STAGE_TRACE)
        q = r.queue("extension")
        hits = int((r.path_field("hit_geom_id")[q] >= 0).sum())
        nbytes = 36.0 * rays + 28.0 * hits + 8.0 * (rays - hits)
        per_bounce.append({"rays": rays, "hits": hits, "ms": round(ms, 4)})
        tot_ms += ms; tot_bytes += nbytes; tot_rays += rays
        r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    else:
        peak, peak_src = 7700.0, "fallback nominal HBM3e figure (MEASURED_PEAKS.json absent)"
    traffic = issue = None
    tpath = os.path.join(ROOT, "profiles", "raycast_traffic.json")
    if os.path.exists(tpath):
        prof = json.load(open(tpath))
        traffic = prof.get("dram_bytes_per_launch")
        wi = prof.get("warp_instructions_per_launch")
        if wi and len(wi) == depth:
            # What actually bounds the kernel: issue slots.  Warp-instructions per launch from the committed ncu
            # capture of the same frame (deterministic: same scene, same rays) over the launch times measured live above.
            sms, ghz = 148, 1.965
            try:
                import torch
                sms = torch.cuda.get_device_properties(0).multi_processor_count
            except Exception:  # noqa: BLE001
                pass
            ach = sum(wi) / (tot_ms * 1e-3) / 1e9
            pk = sms * 4 * ghz
            issue = {"bound": "issue slots", "unit": "G warp-instructions/s (peak = SMs x 4 schedulers x 1.965 GHz)", "achieved": round(ach, 1), "peak": round(pk, 1),
                     "frac": round(ach / pk, 3), "warp_instructions_per_step": int(sum(wi)), "source": prof.get("warp_instructions_source")}
    achieved = tot_bytes / (tot_ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": "k_trace_extension", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 5),
            "traffic": traffic, "peak_source": peak_src, "launches_per_step": depth, "avg_launch_ms": round(tot_ms / depth, 4),
            "algorithmic_bytes_per_launch": round(tot_bytes / depth), "mrays_per_s": round(tot_rays / (tot_ms * 1e-3) / 1e6, 1), "per_bounce": per_bounce,
            "note": "BVH (19 MB) is L2-resident at this size: the kernel is bound by issue slots under divergence (see `issue`), the HBM figure counts ray-state streaming only",
            "issue": issue}
    cpu = None
    if sample is not None:
        import _oracle as orc
        pos, im = scene.array("pos"), scene.array("indices_matid")
        bvh = {k: r.lbvh(k) for k in ("aabb_min", "aabb_max", "primitive_indices", "child_nodes", "escape_indices")}
        cores = os.cpu_count() or 1
        t0 = time.perf_counter()
        orc.trace_closest(bvh, pos, im, sample[0], sample[1], threads=cores)
        dt = time.perf_counter() - t0
        mrays = sample[0].shape[0] / dt / 1e6
        cpu = {"value": round(mrays / stats_rays_per_path, 4), "unit": "Mpaths/s", "cores": cores, "kind": "port",
               "sample": f"{sample[0].shape[0]} extension rays (the whole queues of bounces 0-3 of one frame) through oracle/cpu/traverse.c on {cores} threads: {mrays:.3f} Mrays/s in {dt:.1f} s, "
                         f"divided by the {stats_rays_per_path:.2f} rays a path costs (traversal only, no shading: an upper bound for a CPU path tracer)"}
    return roof, cpu


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ptb200")
    ap.add_argument("--no-extras", action="store_true", help="skip the roofline / cpu_baseline post-pass")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, rank)
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch

    import cuda_path_tracer_b200 as ptb
    from cuda_path_tracer_b200 import capi, multigpu
    if capi.lib.ptb_device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; this library has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    scene = ptb.Scene(make_scene(rank))
    stripe = 8
    W, H = RES
    P = W * H
    stream = torch.cuda.Stream()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    host = torch.empty((P, 3), dtype=torch.float32, pin_memory=True) if rank == 0 else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def timed(step, first):
        """W warm-up + K timed steps of `step(k)`; returns (device ms over the K steps, max over ranks)."""
        for k in range(args.warmup):
            step(first + k)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        barrier()
        for k in range(args.steps):
            flush.zero_()                         # L2 flush between timed iterations (outside the event pair)
            ev[k][0].record(stream)
            step(first + args.warmup + k)
            ev[k][1].record(stream)
        barrier()
        return allmax(sum(a.elapsed_time(b) for a, b in ev))

    with torch.cuda.stream(stream):
        # ---- N > 1, secondary figure: TILE partition of ONE frame per step (strong scaling, bit-identical to 1 GPU)
        tile_info = None
        if world > 1:
            rt = ptb.Renderer(scene, device=local_rank, rank=rank, world=world, stripe_rows=stripe)
            rt.set_stream(stream.cuda_stream)
            fg = multigpu.FrameGather(W, H, stripe, dist, device="cuda")

            def tile_step(k):
                rt.render(k + 1)
                rt.resolve_device(fg.tile.data_ptr(), k + 1)
                return fg.gather()
            tms = timed(tile_step, 0) / args.steps
            tile_info = {"what": f"one 1080p frame per step split over {world} GPUs (interleaved {stripe}-row stripes, NCCL gather to rank 0), bit-identical to 1 GPU",
                         "ms_per_frame": round(tms, 4), "mpaths_per_s": round(P / tms / 1e3, 3), "scaling": "strong"}
            rt.free()
            del fg
        # ---- headline: every rank renders full frames; N > 1 = sample-index partition, one reduce per step
        r = ptb.Renderer(scene, device=local_rank)
        r.set_stream(stream.cuda_stream)
        sr = multigpu.SampleReduce(W, H, dist, device="cuda") if world > 1 else None
        tile = sr.sum if world > 1 else torch.empty((P, 3), dtype=torch.float32, device="cuda")

        def step(k):
            if world == 1:
                r.render(k + 1)
                r.resolve_device(tile.data_ptr(), k + 1)
                return tile
            if k == 0:
                r.clear()
            r.render(sr.iteration(k))
            r.resolve_device(tile.data_ptr(), 1, capi.IMG_SUM_RGB32F)
            return sr.reduce((k + 1) * world)

        clocks = ClockSampler(local_rank)
        for k in range(args.warmup):                 # extra warm-up so that stats_reset sits before the timed loop only
            step(k)
        barrier()
        r.stats_reset()
        if rank == 0:
            clocks.start()
        total_ms = timed(step, args.warmup)
        clk = clocks.stop() if rank == 0 else None
        st = r.stats()
        frames_counted = args.steps + args.warmup
        t = torch.tensor([float(st.kernel_launches), float(st.extension_rays + st.shadow_rays)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        launches = int(float(t[0]) * args.steps / frames_counted)          # stats cover warm-up + timed steps of timed()
        rays = float(t[1]) * args.steps / frames_counted
        # ---- end to end through the public API with HOST buffers: render + resolve (+ reduce) + D2H of the frame, every step
        # N = 1: double-buffered -- frame k is resolved on the device, its D2H copy runs on a copy stream while frame k + 1
        # renders, and a buffer is reused only after its copy has landed (the caller-side pipeline of ptb_render +
        # ptb_resolve(dst_on_device) + cudaMemcpyAsync); every step's frame still reaches the pinned host buffer inside the
        # timed region.  N > 1: the reduced frame is copied by rank 0 after each step.
        copy_stream = torch.cuda.Stream()
        dev2 = [tile, torch.empty_like(tile)] if world == 1 else None
        host2 = [host, torch.empty((P, 3), dtype=torch.float32, pin_memory=True)] if (world == 1 and rank == 0) else None
        copied = [None, None]
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            kk = 2 * args.warmup + args.steps + k
            if world == 1:
                b = k & 1
                if copied[b] is not None:
                    stream.wait_event(copied[b])              # the render stream may overwrite dev2[b] only after its copy
                r.render(kk + 1)
                r.resolve_device(dev2[b].data_ptr(), kk + 1)
                done = torch.cuda.Event()
                done.record(stream)
                copy_stream.wait_event(done)
                with torch.cuda.stream(copy_stream):
                    host2[b].copy_(dev2[b], non_blocking=True)
                    copied[b] = torch.cuda.Event()
                    copied[b].record(copy_stream)
            else:
                frame = step(kk)
                if rank == 0:
                    host.copy_(frame, non_blocking=True)
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        barrier()
        e2e_s = allmax(time.perf_counter() - t0)
        roof = cpu = mega_info = None
        if not args.no_extras:
            roof, cpu = roofline_and_cpu(r, scene, world, rank, rays / (args.steps * P * world))
            if world == 1:
                # secondary figure: the same frame through the OTHER render path (ptb_config.megakernel; the reference's
                # pathtrace_megakernel::Pathtrace, src/pathtrace.cu:645-706), timed the same way
                rm = ptb.Renderer(scene, device=local_rank, megakernel=True)
                rm.set_stream(stream.cuda_stream)

                def mega_step(k):
                    rm.render(k + 1)
                    rm.resolve_device(tile.data_ptr(), k + 1)
                    return tile
                mms = timed(mega_step, 0) / args.steps
                sm = rm.stats()
                mega_info = {"what": "ptb_config.megakernel = 1: one persistent kernel per frame with the reference megakernel's estimator (bit-exact vs its d_image, tests/test_gpu_parity.py)",
                             "ms_per_frame": round(mms, 4), "mpaths_per_s": round(P / mms / 1e3, 3),
                             "rays_per_path": round((sm.extension_rays + sm.shadow_rays) / (sm.frames * P), 3)}
                rm.free()
    paths_per_step = P * world
    if rank == 0:
        ms = total_ms / args.steps
        part = "single GPU" if world == 1 else (f"sample-index partition: each of the {world} GPUs renders a full 1080p frame per step (iterations rank+1, rank+1+N, ...), "
                                                f"scene+LBVH replicated, ONE NCCL reduce of the resolved sums to rank 0 per step; a step is {world} spp")
        out = {"metric": "Mpaths/s", "value": round(paths_per_step / ms / 1e3, 3), "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": WORKLOAD, "partition": part, "paths_per_step": paths_per_step,
                          "l2": "256 MB flush write between timed steps", "timing": "CUDA events on the launching stream per step, max over ranks",
                          "mrays_per_s": round(rays / (total_ms * 1e-3) / 1e6, 1), "rays_per_path": round(rays / (args.steps * paths_per_step), 3),
                          "lbvh_build_ms": round(st.lbvh_build_ms, 4), "lbvh_max_depth": st.lbvh_max_depth, "parity": "radiance bit-exact vs reference (tests/test_gpu_parity.py)"},
               "gpu_launches": launches, "clocks": clk,
               "e2e": {"value": round(args.steps * paths_per_step / e2e_s / 1e6, 3), "unit": "Mpaths/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": P * 12,
                       "note": "ptb_render + ptb_resolve (+ reduce) + D2H of the frame into a pinned HOST buffer every step, wall clock incl. the final sync (N = 1: double-buffered, the copy of frame k overlaps the render of frame k + 1); the scene is static, the only per-step input is the iteration index"},
               "roofline": roof}
        if cpu is not None:
            out["cpu_baseline"] = cpu
        if tile_info is not None:
            out["config"]["tile_partition"] = tile_info
        if mega_info is not None:
            out["config"]["megakernel_path"] = mega_info
        print(json.dumps(out))
    r.free()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
