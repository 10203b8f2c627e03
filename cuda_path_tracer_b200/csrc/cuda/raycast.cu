// raycast.cu -- RayCast stage kernels (TraceExtensionRay / TraceShadowRay, reference
// src/kernels/ray_cast.cu:204-234) as persistent kernels: a fixed grid sized from the SM count,
// warps pull 32-ray batches from the queue through a device-side cursor, and the queue length is
// read on the device (no host round trip per bounce).
#include "raycast.cuh"
#include "wide.cuh"
#include "logic.cuh"
#include "render.cuh"

namespace ptb {

#ifndef PTB_SH_MINBLOCKS
#define PTB_SH_MINBLOCKS 12   /* any-hit keeps no hit record: 40 registers without spills, 48 resident warps per SM */
#endif
#ifndef PTB_RC_MINBLOCKS
#define PTB_RC_MINBLOCKS 10
#endif

// one 32-ray batch per warp fetch
__device__ __forceinline__ int warp_fetch(int* cursor) {
    int base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32);
    return __shfl_sync(0xffffffffu, base, 0);
}

// FUSE_LOGIC: run the Logic stage (logic.cuh) on each path the moment its extension ray is done, while the hit
// record is in registers -- the frame graph's form.  Without it this is the reference's TraceExtensionRay alone
// (staged API, ptb_time_raycast).
// WIDE: the 4-wide quantised walk (wide.cuh) instead of trace_closest(); both give the same bits.
template <bool FUSE_LOGIC, bool WIDE>
__global__ void __launch_bounds__(PTB_RC_THREADS, PTB_RC_MINBLOCKS) k_trace_extension(FrameDev f, int bounce, int* cursor) {
    __shared__ int stack[PTB_EXT_STACK * PTB_RC_THREADS];
    int* ctr = f.ctr + bounce * CTR_STRIDE;
    const int count = ctr[CTR_EXT];
    int* my_stack = stack + threadIdx.x;
    for (int base = warp_fetch(cursor); base < count; base = warp_fetch(cursor)) {
        int i = base + (threadIdx.x & 31);
        int queue = -1, path = -1;
        if (i < count) {
            path = __ldg(&f.q_ext[i]);
            float4 o4 = f.pool.ray_ori[path];
            float4 d4 = f.pool.ray_dir_dist[path];
            Hit h = WIDE ? trace_closest_exact(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), my_stack, PTB_RC_THREADS)
                         : trace_closest(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), my_stack, PTB_RC_THREADS);
            // write-back, src/kernels/ray_cast.cu:111-121
            if (h.prim == -1) {
                f.pool.ray_dir_dist[path].w = -1.0f;
                f.pool.hit_geom_id[path] = -1;
            } else {
                f.pool.ray_dir_dist[path].w = h.t;
                f.pool.material_id[path] = h.mat;
                f.pool.hit_geom_id[path] = h.prim;
                f.pool.hit_uv[path] = make_float4(h.u, h.v, 0.0f, 0.0f);
            }
            if (FUSE_LOGIC) queue = logic_path<true>(f, path, h.prim, h.mat, make_float4(d4.x, d4.y, d4.z, h.t));
        }
        if (FUSE_LOGIC) logic_dispatch(f, ctr, queue, path);
    }
}

// (A persistent-lane scheduler -- a warp owns a chunk of the queue and refills lanes whose ray ended -- was measured
// 1.3-1.4x slower than one 32-ray batch per warp fetch and is no longer part of the build)

template <int STACK, bool SHARED, bool WIDE>
__global__ void __launch_bounds__(PTB_RC_THREADS, (SHARED && STACK <= 32) ? PTB_SH_MINBLOCKS : 1) k_trace_shadow(FrameDev f, int bounce) {
    __shared__ int sstack[SHARED ? STACK * PTB_RC_THREADS : 1];
    int lstack[SHARED ? 1 : STACK];
    int* my_stack = SHARED ? (sstack + threadIdx.x) : lstack;
    const int stride = SHARED ? PTB_RC_THREADS : 1;
    int* ctr = f.ctr + bounce * CTR_STRIDE;
    const int count = ctr[CTR_SHADOW];
    for (int base = warp_fetch(&ctr[CTR_WORK_SHADOW]); base < count; base = warp_fetch(&ctr[CTR_WORK_SHADOW])) {
        int i = base + (threadIdx.x & 31);
        if (i >= count) continue;
        float4 o4 = __ldg(&f.shadow.ray_ori_tmax[i]);
        float4 d4 = __ldg(&f.shadow.ray_dir[i]);
        int r = WIDE ? trace_any_exact<STACK>(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), o4.w, my_stack, stride)
                     : trace_any<STACK>(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), o4.w, my_stack, stride);
        if (r == 0) {   // visible: AtomicAddVec3(&d_direct_image[pixel], radiance), :193-199
            int slot = __ldg(&f.shadow.slot[i]);
            float4 rad = __ldg(&f.shadow.radiance[i]);
            float* dst = f.direct + 3 * (size_t)slot;
            atomicAdd(dst + 0, rad.x);
            atomicAdd(dst + 1, rad.y);
            atomicAdd(dst + 2, rad.z);
        } else if (r < 0) {
            atomicExch(f.error_flag, 1);
        }
    }
}

static inline int rc_grid(const LaunchCfg& c) { return c.num_sms * (c.rc_blocks > 0 ? c.rc_blocks : 8); }

static void run_extension(const FrameDev& f, int bounce, int* cursor, const LaunchCfg& c, cudaStream_t s, bool fuse_logic = false) {
    const bool wide = (c.wide & 2) != 0;
    if (fuse_logic && wide) k_trace_extension<true, true><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce, cursor);
    else if (fuse_logic) k_trace_extension<true, false><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce, cursor);
    else if (wide) k_trace_extension<false, true><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce, cursor);
    else k_trace_extension<false, false><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce, cursor);
}

int launch_trace_extension_only(const FrameDev& f, int bounce, int* cursor, const LaunchCfg& c, cudaStream_t s) {
    run_extension(f, bounce, cursor, c, s);
    return 1;
}

int launch_trace_extension(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s, bool fuse_logic) {
    run_extension(f, bounce, f.ctr + bounce * CTR_STRIDE + CTR_WORK_EXT, c, s, fuse_logic);
    return 1;
}

int launch_trace_shadow(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s) {
    if (bounce == 0) return 0;   // no shadow rays exist before the first Shade (counter zeroed by RayGen, src/kernels/ray_gen.cu:121-124)
    // the DFS stack never holds more than one entry per tree level: size it from the measured LBVH depth
    const int g = rc_grid(c);
    const bool wide = (c.wide & 1) != 0;
    if (c.shadow_deep == 3) k_trace_shadow<256, false, false><<<g, PTB_RC_THREADS, 0, s>>>(f, bounce);
    else if (c.shadow_deep == 0 && wide) k_trace_shadow<32, true, true><<<c.num_sms * (c.sh_blocks > 0 ? c.sh_blocks : PTB_SH_MINBLOCKS), PTB_RC_THREADS, 0, s>>>(f, bounce);
    else if (c.shadow_deep == 0) k_trace_shadow<32, true, false><<<c.num_sms * PTB_SH_MINBLOCKS, PTB_RC_THREADS, 0, s>>>(f, bounce);
    else if (c.shadow_deep == 1 && wide) k_trace_shadow<48, true, true><<<g, PTB_RC_THREADS, 0, s>>>(f, bounce);
    else if (c.shadow_deep == 1) k_trace_shadow<48, true, false><<<g, PTB_RC_THREADS, 0, s>>>(f, bounce);
    else if (wide) k_trace_shadow<64, true, true><<<g, PTB_RC_THREADS, 0, s>>>(f, bounce);
    else k_trace_shadow<64, true, false><<<g, PTB_RC_THREADS, 0, s>>>(f, bounce);
    return 1;
}

int launch_trace(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s, bool fuse_logic) {
    return launch_trace_extension(f, bounce, c, s, fuse_logic) + launch_trace_shadow(f, bounce, c, s);
}

// ---- batched entry points (ptb_trace_closest / ptb_trace_occluded) ---------------------------
__global__ void __launch_bounds__(PTB_RC_THREADS) k_batch_closest(BvhDev bvh, const float4* __restrict__ ori, const float4* __restrict__ dir,
                                                                  int n, int* __restrict__ hit, float* __restrict__ t,
                                                                  float* __restrict__ u, float* __restrict__ v, int* work) {
    __shared__ int stack[PTB_EXT_STACK * PTB_RC_THREADS];
    int* my_stack = stack + threadIdx.x;
    for (int base = warp_fetch(work); base < n; base = warp_fetch(work)) {
        int i = base + (threadIdx.x & 31);
        if (i >= n) continue;
        float4 o4 = __ldg(&ori[i]), d4 = __ldg(&dir[i]);
        Hit h = trace_closest(bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), my_stack, PTB_RC_THREADS);
        hit[i] = h.prim;
        t[i] = h.prim == -1 ? -1.0f : h.t;
        u[i] = h.u;
        v[i] = h.v;
    }
}

template <int STACK, bool SHARED>
__global__ void __launch_bounds__(PTB_RC_THREADS) k_batch_occluded(BvhDev bvh, const float4* __restrict__ ori_tmax, const float4* __restrict__ dir,
                                                                   int n, int* __restrict__ occluded, int* work, int* error_flag) {
    __shared__ int sstack[SHARED ? STACK * PTB_RC_THREADS : 1];
    int lstack[SHARED ? 1 : STACK];
    int* my_stack = SHARED ? (sstack + threadIdx.x) : lstack;
    const int stride = SHARED ? PTB_RC_THREADS : 1;
    for (int base = warp_fetch(work); base < n; base = warp_fetch(work)) {
        int i = base + (threadIdx.x & 31);
        if (i >= n) continue;
        float4 o4 = __ldg(&ori_tmax[i]), d4 = __ldg(&dir[i]);
        int r = trace_any<STACK>(bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), o4.w, my_stack, stride);
        if (r < 0) atomicExch(error_flag, 1);
        occluded[i] = r == 1 ? 1 : 0;
    }
}

int launch_trace_batch_closest(const BvhDev& bvh, const float4* ori, const float4* dir, int n, int* hit, float* t, float* u,
                               float* v, int* work, const LaunchCfg& c, cudaStream_t s) {
    cudaMemsetAsync(work, 0, sizeof(int), s);
    k_batch_closest<<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(bvh, ori, dir, n, hit, t, u, v, work);
    return 1;
}

int launch_trace_batch_occluded(const BvhDev& bvh, const float4* ori_tmax, const float4* dir, int n, int* occluded, int* work,
                                int* error_flag, const LaunchCfg& c, cudaStream_t s) {
    cudaMemsetAsync(work, 0, sizeof(int), s);
    if (c.shadow_deep == 3) k_batch_occluded<256, false><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(bvh, ori_tmax, dir, n, occluded, work, error_flag);
    else k_batch_occluded<PTB_SHADOW_STACK, true><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(bvh, ori_tmax, dir, n, occluded, work, error_flag);
    return 1;
}

}  // namespace ptb

#ifdef PTB_WIDE_DEBUG   // development counters of the wide walk (not part of the ABI): rays, bad reciprocal, re-traced, stack overflow
extern "C" int ptb_debug_wide(unsigned long long* out) { return (int)cudaMemcpyFromSymbol(out, ptb::g_wide_dbg, sizeof(unsigned long long) * 8); }
#endif
