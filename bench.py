#!/usr/bin/env python
"""bench.py -- Mpaths/s of the wavefront render path on BASELINE.json configs[1]:
synthetic 143,173-triangle Fireplace-like room (PBR / diffuse / mirror / dielectric mix of README.md:34),
1920x1080, 8 bounces, MIS + NEE, 1 sample per pixel per GPU per step.

    python bench.py --gpus 1 --steps 20 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
    python bench.py --impl reference ...     # the reference's own wavefront build (oracle/_ref/refdump), same scene files

A step = one full frame per GPU: RayGen -> 8 x (RayCast, Logic, Shade) -> resolve.  With N > 1 three partitions of the work are
timed in the same run, all through the library's peer-store gather (every rank's resolve kernel writes its pixels into rank 0's
frame buffer over NVLink; multigpu.PeerFrame, include/ptb200.h "multi-GPU"), each checked against a 1-GPU render of the same
iterations on rank 0 (scalar keys config.*_bit_identical / *_max_rel_err):
  * sample-index partition (the headline `value`, "scaling": "weak"): every rank renders the full frame, rank r the iterations
    r+1, r+1+N, ...; a step is N samples per pixel;
  * tile partition (config.tile_*, strong scaling): ONE frame per step cut into N interleaved 8-row stripe sets;
  * hybrid (config.hybrid_*): G tile groups x N/G sample groups.
With N > 1 the run also renders BASELINE configs 5 (262,279-triangle textured scene + HDR sky, 3840x2160, 64 spp) and 4
(10,000,000 triangles, 1920x1080, 4 spp) as whole jobs over the N GPUs in the partitions their sample counts allow
(config.config5_* / config.config4_*: job ms, Mpaths/s, speed-up over ONE renderer doing the same job on rank 0, bit-identity /
max relative error against that renderer's frame).
Every secondary figure is a flat scalar key (the driver's record keeps scalars only).
"""
import argparse
import hashlib
import importlib.util
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
DEPTH = 8
REFDUMP = os.path.join(ROOT, "oracle", "_ref", "refdump")
# sources whose change invalidates the committed ncu figures of profiles/raycast_traffic.json
KERNEL_SOURCES = ["raycast.cu", "raycast.cuh", "wide.cuh", "logic.cuh", "common.cuh", "render.cuh"]


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


def load_scenegen():
    """The scene generator loaded BY FILE PATH: pure numpy, no import of the package, so the reference arm never maps libptb200.so."""
    spec = importlib.util.spec_from_file_location("ptb_scenegen_standalone", os.path.join(ROOT, "cuda_path_tracer_b200", "scenegen.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_scene(rank):
    return load_scenegen().fireplace(f"/tmp/ptb_bench_fire_r{rank}", res=RES)[0]


def sha256_file(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    return h.hexdigest()


def reference_arm(args, rank):
    """The UNMODIFIED reference wavefront renderer (built from /root/reference into oracle/_ref by oracle/Makefile),
    driven through its own PathtraceInit / Pathtrace entry points on the same scene files.  The reference is a
    single-GPU GPU program with no CPU path; it is timed on ONE GPU whatever --gpus says.  Nothing of this repo's
    library is imported or mapped in this arm."""
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
    info = {}
    bi = os.path.join(os.path.dirname(REFDUMP), "refdump.buildinfo")
    if os.path.exists(bi):
        info = json.load(open(bi))
    sha = sha256_file(REFDUMP)
    cfg = {"workload": WORKLOAD, "reference": "s64zxu/CUDA-Path-Tracer wavefront build recompiled -arch=sm_100 (oracle/Makefile), Pathtrace(NULL,0,iter) loop, wall clock around the loop",
           "mrays_per_s": rb["mrays_per_s"], "gpus_used": 1, "timing": "wall clock around the frame loop (the reference synchronises after every launch)",
           "refdump_sha256": sha, "refdump_sha256_at_build": info.get("sha256"), "refdump_matches_build": info.get("sha256") == sha if info else None,
           "refdump_nvcc": info.get("nvcc"), "refdump_built": info.get("built")}
    if mega is not None:
        cfg.update({"mega_ms_per_frame": mega["ms_per_frame"], "mega_mpaths_per_s": mega["mpaths_per_s"], "mega_mrays_per_s": mega["mrays_per_s"]})
    print(json.dumps({"impl": "reference", "metric": "Mpaths/s", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": rb["ms_per_frame"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": cfg,
                      "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": 0, "kind": "reference",
                                       "sample": "the reference has no CPU implementation of this path (SURVEY.md fact 2): this arm is its own CUDA build on one B200; the CPU port is timed in the main arm"},
                      "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def kernel_sources_hash():
    h = hashlib.sha256()
    for name in KERNEL_SOURCES:
        with open(os.path.join(ROOT, "cuda_path_tracer_b200", "csrc", "cuda", name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def roofline_and_cpu(r, scene, n_gpus, rank, stats_rays_per_path, sm_count, sm_mhz):
    """Rank-0 post-pass on ONE staged frame: (1) either RayCast kernel timed alone per bounce with CUDA events on the
    renderer's stream (ptb_time_kernel): warm (back to back, the tree stays in L2 -- how it runs inside a frame) and cold
    (256 MB flush before every launch); algorithmic bytes from SURVEY.md 8(d): extension ray 36 B read + 28 B written
    on a hit / 8 B on a miss, shadow ray 32 B read + 32 B (20 B read + 12 B atomics) if unoccluded; (2) the CPU port
    of the same traversal on a bounded ray sample, on all host threads and on one."""
    import numpy as np

    from cuda_path_tracer_b200 import capi
    r.stage(capi.STAGE_RAYGEN, 1)
    ext = {"ms": 0.0, "cold_ms": 0.0, "bytes": 0.0, "rays": 0}
    sh = {"ms": 0.0, "cold_ms": 0.0, "bytes": 0.0, "rays": 0, "launches": 0}
    per_bounce, sample = [], None
    for b in range(DEPTH):
        ms, rays = r.time_kernel("ext", 5)
        cold_ms, _ = r.time_kernel("ext", 3, cold=True)
        sms, srays = r.time_kernel("shadow", 5)
        scold, _ = r.time_kernel("shadow", 3, cold=True)
        if b <= 3 and rank == 0 and n_gpus == 1:
            q = r.queue("extension")
            ori, dirs = r.path_field("ray_ori")[q], r.path_field("ray_dir_dist")[q]
            sample = (ori, dirs) if sample is None else (np.concatenate([sample[0], ori]), np.concatenate([sample[1], dirs]))
        direct_before = r.accumulators()[0] if srays else None
        r.stage(capi.STAGE_TRACE)
        q = r.queue("extension")
        hits = int((r.path_field("hit_geom_id")[q] >= 0).sum())
        ext["ms"] += ms; ext["cold_ms"] += cold_ms; ext["rays"] += rays
        ext["bytes"] += 36.0 * rays + 28.0 * hits + 8.0 * (rays - hits)
        row = {"rays": rays, "hits": hits, "ms": round(ms, 4), "cold_ms": round(cold_ms, 4), "shadow_rays": srays, "shadow_ms": round(sms, 4)}
        if srays:
            lit = int(((r.accumulators()[0] != direct_before).any(axis=1)).sum())       # pixels a visible shadow ray added to
            sh["ms"] += sms; sh["cold_ms"] += scold; sh["rays"] += srays; sh["launches"] += 1
            sh["bytes"] += 32.0 * srays + 32.0 * lit
            row["shadow_unoccluded"] = lit
        per_bounce.append(row)
        r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    else:
        peak, peak_src = 7700.0, "fallback nominal HBM3e figure (MEASURED_PEAKS.json absent)"
    # ncu figures of the committed capture, only while the kernel sources are the captured ones
    traffic = traffic_src = issue_frac = lanes = traffic_note = None
    tpath = os.path.join(ROOT, "profiles", "raycast_traffic.json")
    if os.path.exists(tpath):
        prof = json.load(open(tpath))
        if prof.get("kernel_sources_sha16") == kernel_sources_hash():
            traffic = prof.get("dram_bytes_per_launch")
            traffic_src = "committed ncu capture " + str(prof.get("capture")) + " (same kernel sources)"
            wi = prof.get("warp_instructions_per_launch")
            if wi and len(wi) == DEPTH and sm_count and sm_mhz:
                issue_frac = round(sum(wi) / (ext["ms"] * 1e-3) / (sm_count * 4 * sm_mhz * 1e6), 3)
            lanes = prof.get("active_lanes_mean")
            traffic_note = prof.get("note")
        else:
            traffic_src = "committed ncu capture is of other kernel sources: not reported"
    achieved = ext["bytes"] / (ext["ms"] * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": "k_trace_extension", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 5),
            "traffic": traffic, "traffic_source": traffic_src, "traffic_note": traffic_note, "peak_source": peak_src, "launches_per_step": DEPTH, "avg_launch_ms": round(ext["ms"] / DEPTH, 4),
            "avg_launch_ms_cold": round(ext["cold_ms"] / DEPTH, 4), "timing": "warm: 5 back-to-back launches per bounce (tree in L2, as inside a frame); cold: 256 MB L2 flush before each of 3 launches",
            "algorithmic_bytes_per_launch": round(ext["bytes"] / DEPTH), "mrays_per_s": round(ext["rays"] / (ext["ms"] * 1e-3) / 1e6, 1),
            "frac_cold": round(ext["bytes"] / (ext["cold_ms"] * 1e-3) / 1e9 / peak, 5),
            "issue_frac": issue_frac, "active_lanes": lanes, "per_bounce": per_bounce,
            "note": "BVH (19 MB) is L2-resident at this size: the kernel is bound by issue slots under divergence (issue_frac), the HBM figure counts ray-state streaming only"}
    shadow = None
    if sh["launches"]:
        a = sh["bytes"] / (sh["ms"] * 1e-3) / 1e9
        shadow = {"bound": "hbm", "kernel": "k_trace_shadow", "achieved": round(a, 2), "peak": peak, "unit": "GB/s", "frac": round(a / peak, 5),
                  "launches_per_step": sh["launches"], "avg_launch_ms": round(sh["ms"] / sh["launches"], 4), "avg_launch_ms_cold": round(sh["cold_ms"] / sh["launches"], 4),
                  "algorithmic_bytes_per_launch": round(sh["bytes"] / sh["launches"]), "mrays_per_s": round(sh["rays"] / (sh["ms"] * 1e-3) / 1e6, 1)}
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
        one = sample[0].shape[0] // max(cores, 8)                  # the 1-thread figure on a proportionally smaller slice (same bound on time)
        t0 = time.perf_counter()
        orc.trace_closest(bvh, pos, im, sample[0][:one], sample[1][:one], threads=1)
        dt1 = time.perf_counter() - t0
        mrays1 = one / dt1 / 1e6
        cpu = {"value": round(mrays / stats_rays_per_path, 4), "unit": "Mpaths/s", "cores": cores, "kind": "port",
               "value_1thread": round(mrays1 / stats_rays_per_path, 5), "mrays_per_s": round(mrays, 3), "mrays_per_s_1thread": round(mrays1, 4),
               "sample": f"{sample[0].shape[0]} extension rays (the whole queues of bounces 0-3 of one frame) through oracle/cpu/traverse.c on {cores} threads: {mrays:.3f} Mrays/s in {dt:.1f} s "
                         f"(1 thread on the first {one} of them: {mrays1:.3f} Mrays/s in {dt1:.1f} s), divided by the {stats_rays_per_path:.2f} rays a path costs "
                         f"(traversal only, no shading: an upper bound for a CPU path tracer)"}
    return roof, shadow, cpu


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ptb200")
    ap.add_argument("--no-extras", action="store_true", help="skip the roofline / cpu_baseline / parity post-passes")
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
    sm_count = torch.cuda.get_device_properties(local_rank).multi_processor_count

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

    extra = {}                                    # flat scalar keys merged into config
    out_frame = torch.empty((P, 3), dtype=torch.float32, device="cuda")

    class Partition:
        """One way of cutting the work over the ranks, through the peer-store gather (or, without peer access, NCCL)."""

        def __init__(self, tile_groups, scene_=None, res=None, out=None):
            self.G = tile_groups
            self.S = world // tile_groups
            self.W, self.H = res or (W, H)
            self.P = self.W * self.H
            self.out = out_frame if out is None else out
            self.peer = None
            try:
                self.peer = multigpu.PeerFrame(self.W, self.H, dist, local_rank, tile_groups=tile_groups, stripe_rows=stripe)
                kw = self.peer.renderer_kwargs()
            except multigpu.PeerUnavailable:
                kw = dict(rank=rank % self.G, world=self.G, stripe_rows=stripe)
                self.nccl_tile = multigpu.FrameGather(self.W, self.H, stripe, dist, device="cuda") if self.S == 1 else None
                self.nccl_sum = multigpu.SampleReduce(self.W, self.H, dist, device="cuda") if self.G == 1 else None
                if self.nccl_tile is None and self.nccl_sum is None:
                    raise
            self.r = ptb.Renderer(scene if scene_ is None else scene_, device=local_rank, **kw)
            self.r.set_stream(stream.cuda_stream)
            self.combines = 0

        def step(self, k):
            """Step k (0-based): this rank's iteration of its share, gathered into out_frame on rank 0 as the running mean."""
            it = k * self.S + rank // self.G + 1
            if k == 0 and it != 1:
                self.r.clear()
            self.r.render(it)
            if self.peer is not None:
                self.peer.submit(self.r, k)
                if rank == 0:
                    self.peer.combine_device(self.out.data_ptr(), k + 1, capi.IMG_MEAN_RGB32F, stream.cuda_stream, k)
                    self.combines += 1
                return self.out
            if self.nccl_tile is not None:
                self.r.resolve_device(self.nccl_tile.tile.data_ptr(), k + 1)
                return self.nccl_tile.gather()
            self.r.resolve_device(self.nccl_sum.sum.data_ptr(), 1, capi.IMG_SUM_RGB32F)
            return self.nccl_sum.reduce((k + 1) * world)

        def check(self, frame, steps_done, ref_renderer):
            """Rank 0: the gathered mean frame against ONE renderer's mean over the same iterations 1 .. steps_done * S.
            Returns (bit-identical, max relative error, ms per frame of that one renderer alone on this GPU)."""
            n = steps_done * self.S
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            ref_renderer.render(1, n)
            e1.record(stream)
            ref = torch.empty((self.P, 3), dtype=torch.float32, device="cuda")
            ref_renderer.resolve_device(ref.data_ptr(), n)
            torch.cuda.synchronize()
            a, b = frame.cpu().numpy(), ref.cpu().numpy()
            same = bool(np.array_equal(a.view(np.uint32), b.view(np.uint32)))
            rel = float((np.abs(a - b) / np.maximum(np.abs(b), 1e-3)).max())
            return same, rel, e0.elapsed_time(e1) / n

        def free(self):
            self.r.free()
            if self.peer is not None:
                self.peer.free()

    with torch.cuda.stream(stream):
        # ---- N > 1, secondary figures: tile partition (strong scaling) and hybrid tile x sample partition
        if world > 1:
            ref_r = None
            if rank == 0 and not args.no_extras:
                ref_r = ptb.Renderer(scene, device=local_rank)
                ref_r.set_stream(stream.cuda_stream)
            modes = [("tile", world)]
            if world >= 4:
                modes.append(("hybrid", 2))
            for label, G in modes:
                part = Partition(G)
                ms = timed(part.step, 0) / args.steps
                done = args.warmup + args.steps
                extra[f"{label}_ms_per_step"] = round(ms, 4)
                extra[f"{label}_mpaths_per_s"] = round(part.S * P / ms / 1e3, 3)
                extra[f"{label}_groups"] = f"{G} tile groups x {part.S} sample groups"
                extra[f"{label}_gather"] = "peer stores into rank 0 over NVLink (CUDA IPC), 4-byte NCCL all-reduce as barrier" if part.peer is not None else "NCCL gather (no peer access)"
                if ref_r is not None:
                    frame = part.step(done)            # one more step outside the timed region; its frame is the one compared
                    torch.cuda.synchronize()
                    same, rel, one_ms = part.check(frame, done + 1, ref_r)
                    extra[f"{label}_bit_identical"] = same
                    extra[f"{label}_max_rel_err"] = rel
                    extra["one_gpu_ms_per_frame"] = round(one_ms, 4)          # the same library, one renderer alone on rank 0's GPU (frames back to back)
                    extra[f"{label}_speedup_vs_1gpu"] = round(one_ms * part.S / ms, 3)
                elif not args.no_extras:
                    part.step(done)
                part.free()
        # ---- headline: every rank renders full frames; N > 1 = sample-index partition
        part = Partition(1) if world > 1 else None
        r = part.r if part else ptb.Renderer(scene, device=local_rank)
        if part is None:
            r.set_stream(stream.cuda_stream)

        def step(k):
            if world == 1:
                r.render(k + 1)
                r.resolve_device(out_frame.data_ptr(), k + 1)
                return out_frame
            return part.step(k)

        clocks = ClockSampler(local_rank)
        for k in range(args.warmup):                 # extra warm-up so that stats_reset sits before the timed loop only
            step(k)
        barrier()
        r.stats_reset()
        if part:
            part.combines = 0
        if rank == 0:
            clocks.start()
        total_ms = timed(step, args.warmup)
        clk = clocks.stop() if rank == 0 else None
        st = r.stats()
        frames_counted = args.steps + args.warmup
        combines = part.combines if part else 0
        t = torch.tensor([float(st.kernel_launches + combines), float(st.extension_rays + st.shadow_rays)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        launches = int(float(t[0]) * args.steps / frames_counted)          # stats cover warm-up + timed steps of timed()
        rays = float(t[1]) * args.steps / frames_counted
        steps_done = 2 * args.warmup + args.steps
        if world > 1 and rank == 0 and not args.no_extras:
            frame = step(steps_done)
            torch.cuda.synchronize()
            same, rel, _ = part.check(frame, steps_done + 1, ref_r)
            extra["sample_bit_identical"] = same
            extra["sample_max_rel_err"] = rel
            steps_done += 1
        elif world > 1 and not args.no_extras:
            step(steps_done)
            steps_done += 1
        # ---- end to end through the public API with HOST buffers: render + resolve / gather + D2H of the frame, every step.
        # Double-buffered: frame k is resolved (N > 1: gathered and folded) on the device, its D2H copy runs on a copy stream
        # while frame k + 1 renders, and a device buffer is reused only after its copy has landed; every step's frame still
        # reaches the pinned host buffer inside the timed region.  No host synchronisation inside the loop.
        copy_stream = torch.cuda.Stream()
        dev2 = [out_frame, torch.empty_like(out_frame)]
        host2 = [torch.empty((P, 3), dtype=torch.float32, pin_memory=True) for _ in range(2)] if rank == 0 else None
        copied = [None, None]
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            kk = steps_done + k
            b = k & 1
            if rank == 0 and copied[b] is not None:
                stream.wait_event(copied[b])              # the render stream may overwrite dev2[b] only after its copy
            if world == 1:
                r.render(kk + 1)
                r.resolve_device(dev2[b].data_ptr(), kk + 1)
            else:
                part.r.render(kk * part.S + rank // part.G + 1)
                if part.peer is not None:
                    part.peer.submit(part.r, kk)
                    if rank == 0:
                        part.peer.combine_device(dev2[b].data_ptr(), kk + 1, capi.IMG_MEAN_RGB32F, stream.cuda_stream, kk)
                else:
                    part.r.resolve_device(part.nccl_sum.sum.data_ptr(), 1, capi.IMG_SUM_RGB32F)
                    f = part.nccl_sum.reduce((kk + 1) * world)
                    if rank == 0:
                        dev2[b].copy_(f)
            if rank == 0:
                done = torch.cuda.Event()
                done.record(stream)
                copy_stream.wait_event(done)
                with torch.cuda.stream(copy_stream):
                    host2[b].copy_(dev2[b], non_blocking=True)
                    copied[b] = torch.cuda.Event()
                    copied[b].record(copy_stream)
        torch.cuda.synchronize()
        barrier()
        e2e_s = allmax(time.perf_counter() - t0)
        # ---- N > 1: BASELINE configs 5 and 4 (the multi-GPU cases of north_star) as whole jobs, same gather, same checks
        if world > 1 and not args.no_extras:
            try:
                from cuda_path_tracer_b200 import scenegen as sg
                for tag, spp, res in (("config5", 64, (3840, 2160)), ("config4", 4, (1920, 1080))):
                    if tag == "config5":
                        sdir = "/tmp/ptb_bench_sponza"
                        if rank == 0:
                            sg.sponza(sdir, res=res)
                        barrier()
                        sc = ptb.Scene(os.path.join(sdir, "scene.json"))
                        extra[tag + "_workload"] = "262279-triangle textured scene + HDR sky, 3840x2160, 8 bounces, 64 spp (one job)"
                    else:
                        pos, nor, idx, mat = sg.heightfield_arrays(10_000_000)
                        sc = ptb.Scene.from_arrays(pos, nor, idx, mat, sg.stress_materials(), sg.stress_camera(res))
                        del pos, nor, idx, mat
                        extra[tag + "_workload"] = "10000000-triangle height field, 1920x1080, 8 bounces, 4 spp (one job)"
                    Px = res[0] * res[1]
                    outx = torch.empty((Px, 3), dtype=torch.float32, device="cuda")
                    refx = None
                    if rank == 0:
                        refx = ptb.Renderer(sc, device=local_rank)
                        refx.set_stream(stream.cuda_stream)
                        refx.render(1, 2)                                  # warm-up of the one-GPU comparison render
                        sx = refx.stats()
                        extra[tag + "_lbvh_build_ms"] = round(sx.lbvh_build_ms, 3)
                        extra[tag + "_lbvh_max_depth"] = sx.lbvh_max_depth
                    for label, G in (("tile", world), ("hybrid", 2), ("sample", 1)):
                        S = world // G
                        if G > world or world % G or spp % S or (label == "hybrid" and (world < 4 or G == world)):
                            continue
                        px = Partition(G, sc, res, outx)
                        steps = spp // S
                        for timed_run in (False, True):                    # the whole job twice; the second one is timed
                            barrier()
                            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                            e0.record(stream)
                            for k in range(steps):
                                frame = px.step(k)
                            e1.record(stream)
                            barrier()
                            job_ms = allmax(e0.elapsed_time(e1))
                        key = f"{tag}_{label}"
                        extra[key + "_groups"] = f"{G} tile groups x {S} sample groups"
                        extra[key + "_job_ms"] = round(job_ms, 3)
                        extra[key + "_mpaths_per_s"] = round(Px * spp / job_ms / 1e3, 1)
                        if refx is not None:
                            same, rel, one_ms = px.check(frame, steps, refx)
                            extra[key + "_bit_identical"] = same
                            extra[key + "_max_rel_err"] = rel
                            extra[tag + "_one_gpu_job_ms"] = round(one_ms * spp, 3)
                            extra[key + "_speedup_vs_1gpu"] = round(one_ms * spp / job_ms, 3)
                        px.free()
                    if refx is not None:
                        refx.free()
                    sc.free()
                    del outx
            except Exception as e:                                         # the headline line must survive a failure here
                extra["other_configs_error"] = repr(e)[:300]
        roof = shadow = cpu = None
        if not args.no_extras and world == 1:
            roof, shadow, cpu = roofline_and_cpu(r, scene, world, rank, rays / (args.steps * P * world), sm_count, (clk or {}).get("sm_max_mhz"))
            # secondary figure: the same frame through the OTHER render path (ptb_config.megakernel; the reference's
            # pathtrace_megakernel::Pathtrace, src/pathtrace.cu:645-706), timed the same way
            rm = ptb.Renderer(scene, device=local_rank, megakernel=True)
            rm.set_stream(stream.cuda_stream)

            def mega_step(k):
                rm.render(k + 1)
                rm.resolve_device(out_frame.data_ptr(), k + 1)
                return out_frame
            mms = timed(mega_step, 0) / args.steps
            sm = rm.stats()
            extra.update({"mega_ms_per_frame": round(mms, 4), "mega_mpaths_per_s": round(P / mms / 1e3, 3),
                          "mega_rays_per_path": round((sm.extension_rays + sm.shadow_rays) / (sm.frames * P), 3)})
            rm.free()
    paths_per_step = P * world
    if rank == 0:
        ms = total_ms / args.steps
        part_s = "single GPU" if world == 1 else (f"sample-index partition: each of the {world} GPUs renders a full 1080p frame per step (iterations rank+1, rank+1+N, ...), scene+LBVH replicated; "
                                                  f"every rank's resolve kernel stores its sums into rank 0's buffer over NVLink, rank 0 folds them in rank order; a step is {world} spp")
        cfg = {"workload": WORKLOAD, "partition": part_s, "paths_per_step": paths_per_step,
               "l2": "256 MB flush write between timed steps", "timing": "CUDA events on the launching stream per step, max over ranks",
               "mrays_per_s": round(rays / (total_ms * 1e-3) / 1e6, 1), "rays_per_path": round(rays / (args.steps * paths_per_step), 3),
               "lbvh_build_ms": round(st.lbvh_build_ms, 4), "lbvh_max_depth": st.lbvh_max_depth,
               "parity": "tested by tests/test_gpu_parity.py against the reference build (not checked in this run)"}
        if world > 1:
            cfg["gather"] = "peer stores into rank 0 over NVLink (CUDA IPC), 4-byte NCCL all-reduce as barrier" if part.peer is not None else "NCCL reduce (no peer access)"
        cfg.update(extra)
        if roof is not None:
            cfg.update({"raycast_issue_frac": roof["issue_frac"], "raycast_active_lanes": roof["active_lanes"], "raycast_avg_launch_ms": roof["avg_launch_ms"],
                        "raycast_avg_launch_ms_cold": roof["avg_launch_ms_cold"], "raycast_mrays_per_s": roof["mrays_per_s"]})
        if shadow is not None:
            cfg.update({"shadow_avg_launch_ms": shadow["avg_launch_ms"], "shadow_avg_launch_ms_cold": shadow["avg_launch_ms_cold"], "shadow_mrays_per_s": shadow["mrays_per_s"],
                        "shadow_hbm_frac": shadow["frac"]})
        out = {"metric": "Mpaths/s", "value": round(paths_per_step / ms / 1e3, 3), "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": cfg, "gpu_launches": launches, "clocks": clk,
               "e2e": {"value": round(args.steps * paths_per_step / e2e_s / 1e6, 3), "unit": "Mpaths/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": P * 12,
                       "note": "ptb_render + resolve / gather + D2H of the frame into a pinned HOST buffer every step, wall clock incl. the final sync; double-buffered (the copy of frame k overlaps the render of frame k + 1, no host synchronisation inside the loop); the scene is static, the only per-step input is the iteration index"},
               "roofline": roof}
        if shadow is not None:
            out["roofline_shadow"] = shadow
        if cpu is not None:
            out["cpu_baseline"] = cpu
        print(json.dumps(out))
    if part:
        part.free()
    else:
        r.free()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
