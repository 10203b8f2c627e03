// raycast.cu -- RayCast stage kernels (TraceExtensionRay / TraceShadowRay, reference
// src/kernels/ray_cast.cu:204-234) as persistent kernels: a fixed grid sized from the SM count,
// warps pull 32-ray batches from the queue through a device-side cursor, and the queue length is
// read on the device (no host round trip per bounce).
#include "raycast.cuh"
#include "render.cuh"

namespace ptb {

// one 32-ray batch per warp fetch
__device__ __forceinline__ int warp_fetch(int* cursor) {
    int base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32);
    return __shfl_sync(0xffffffffu, base, 0);
}

__global__ void __launch_bounds__(PTB_RC_THREADS) k_trace_extension(FrameDev f, int bounce, int* cursor) {
    __shared__ int stack[PTB_EXT_STACK * PTB_RC_THREADS];
    const int count = f.ctr[bounce * CTR_STRIDE + CTR_EXT];
    int* my_stack = stack + threadIdx.x;
    for (int base = warp_fetch(cursor); base < count; base = warp_fetch(cursor)) {
        int i = base + (threadIdx.x & 31);
        if (i >= count) continue;
        int path = __ldg(&f.q_ext[i]);
        float4 o4 = f.pool.ray_ori[path];
        float4 d4 = f.pool.ray_dir_dist[path];
        Hit h = trace_closest(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), my_stack, PTB_RC_THREADS);
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
    }
}

template <int STACK, bool SHARED>
__global__ void __launch_bounds__(PTB_RC_THREADS) k_trace_shadow(FrameDev f, int bounce) {
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
        int r = trace_any<STACK>(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), o4.w, my_stack, stride);
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

static inline int rc_grid(const LaunchCfg& c) { return c.num_sms * 8; }

int launch_trace_extension_only(const FrameDev& f, int bounce, int* cursor, const LaunchCfg& c, cudaStream_t s) {
    k_trace_extension<<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce, cursor);
    return 1;
}

int launch_trace(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s) {
    k_trace_extension<<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce, f.ctr + bounce * CTR_STRIDE + CTR_WORK_EXT);
    if (bounce == 0) return 1;   // no shadow rays exist before the first Shade (counter zeroed by RayGen, src/kernels/ray_gen.cu:121-124)
    if (c.shadow_deep) k_trace_shadow<256, false><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce);
    else k_trace_shadow<PTB_SHADOW_STACK, true><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(f, bounce);
    return 2;
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
    if (c.shadow_deep) k_batch_occluded<256, false><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(bvh, ori_tmax, dir, n, occluded, work, error_flag);
    else k_batch_occluded<PTB_SHADOW_STACK, true><<<rc_grid(c), PTB_RC_THREADS, 0, s>>>(bvh, ori_tmax, dir, n, occluded, work, error_flag);
    return 1;
}

}  // namespace ptb
