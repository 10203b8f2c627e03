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
            // write-back, src/kernels/ray_cast.cu:111-121: one whole 32-byte hit record (t = -1 on a miss)
            if (h.prim == -1) h.t = -1.0f;
            f.pool.store_hit(path, h.t, h.u, h.v, h.prim, h.mat);
            if (FUSE_LOGIC) queue = logic_path<true>(f, path, h.prim, h.mat, h.u, h.v, make_float4(d4.x, d4.y, d4.z, h.t));
        }
        if (FUSE_LOGIC) logic_dispatch(f, ctr, queue, path);
    }
}

// ---- persistent lanes -----------------------------------------------------------------------------------------------
// A warp keeps walking while lanes finish; once `refill` lanes have nothing to do it retires their rays (write-back + Logic)
// and hands them the next queue entries.  refill == 32 degenerates to one batch per fetch.  The loop has ONE back edge and
// every decision that shapes it is warp-uniform (ballots at the two converged points), so the lanes of a warp reconverge
// after every traversal step whatever their rays do.
struct StackShared {                         // the whole column in shared memory
    static constexpr int kShared = PTB_EXT_STACK + PTB_WIDE_PUSHES;
    // the column's address in the shared window, held in ONE register the compiler cannot re-derive: it would otherwise rebuild it
    // from %tid and the CTA's window base (two S2R, three more instructions) at every push
    unsigned base;
    __device__ __forceinline__ explicit StackShared(int* c) : base((unsigned)__cvta_generic_to_shared(c)) {
        asm volatile("mov.u32 %0, %0;" : "+r"(base));
    }
    __device__ __forceinline__ int get(int i) const {
        int v;
        asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(base + (unsigned)i * (PTB_RC_THREADS * 4u)) : "memory");
        return v;
    }
    __device__ __forceinline__ void set(int i, int v) {
        asm volatile("st.shared.s32 [%0], %1;" :: "r"(base + (unsigned)i * (PTB_RC_THREADS * 4u)), "r"(v) : "memory");
    }
};
template <int SH>
struct StackHybrid {                         // the SH entries nearest the root in shared memory, deeper ones in local memory
    static constexpr int kShared = SH;
    int* col;
    int loc[PTB_EXT_STACK - SH];
    __device__ __forceinline__ explicit StackHybrid(int* c) : col(c) {}
    __device__ __forceinline__ int get(int i) const { return i < SH ? col[i * PTB_RC_THREADS] : loc[i - SH]; }
    __device__ __forceinline__ void set(int i, int v) { if (i < SH) col[i * PTB_RC_THREADS] = v; else loc[i - SH] = v; }
};

template <bool FUSE_LOGIC, class WALK, class STACK, int MINB, int STEPS = 1>
__global__ void __launch_bounds__(PTB_RC_THREADS, MINB) k_trace_extension_pl(FrameDev f, int bounce, int* cursor, int refill) {
    __shared__ int stack[STACK::kShared * PTB_RC_THREADS];
    int* ctr = f.ctr + bounce * CTR_STRIDE;
    const int count = ctr[CTR_EXT];
    STACK st(stack + threadIdx.x);
    const unsigned lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    int path = -1;
    Vec3 dir;
    WALK w;
    w.reset();
    bool exhausted = false;
    for (;;) {
        // ---- converged: hand out queue entries
        const unsigned idle_mask = __ballot_sync(0xffffffffu, path < 0);
        const int n_idle = __popc(idle_mask);
        if (exhausted && n_idle == 32) break;
        if (!exhausted && n_idle >= refill) {
            int base = 0;
            if (lane == 0) base = atomicAdd(cursor, n_idle);
            base = __shfl_sync(0xffffffffu, base, 0);
            exhausted = base + n_idle >= count;
            const int i = base + __popc(idle_mask & lt);
            if (path < 0 && i < count) {
                path = __ldg(&f.q_ext[i]);
                float4 o4 = f.pool.ray_ori[path];
                float4 d4 = f.pool.ray_dir_dist[path];
                dir = Vec3(d4.x, d4.y, d4.z);
                w.init(f.bvh, Vec3(o4.x, o4.y, o4.z), dir);
            }
        }
        // ---- walk until enough lanes are out of work
        const int want = exhausted ? 32 : refill;
        int n_done;
        do {
#pragma unroll
            for (int k = 0; k < STEPS; k++)
                if (w.busy()) w.step(f.bvh, st);
            n_done = __popc(__ballot_sync(0xffffffffu, !w.busy()));
        } while (n_done < want);
        // ---- converged: retire finished rays (write-back, src/kernels/ray_cast.cu:111-121, then Logic)
        const bool fin = path >= 0 && !w.busy();
        int queue = -1;
        if (fin) {
            Hit h = w.finish(f.bvh, st);
            if (h.prim == -1) h.t = -1.0f;
            f.pool.store_hit(path, h.t, h.u, h.v, h.prim, h.mat);
            if (FUSE_LOGIC) queue = logic_path<true>(f, path, h.prim, h.mat, h.u, h.v, make_float4(dir.x, dir.y, dir.z, h.t));
        }
        if (FUSE_LOGIC) logic_dispatch(f, ctr, queue, path);
        if (fin) path = -1;
    }
}

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

// Visibility rays with persistent lanes (the scheduler of k_trace_extension_pl) on the 4-wide quantised tree; trees up to 32
// levels deep (the stack column is PTB_EXT_STACK entries).  A ray the wide walk cannot decide is re-walked on the fp32 records.
template <int MINB, int STEPS>
__global__ void __launch_bounds__(PTB_RC_THREADS, MINB) k_trace_shadow_pl(FrameDev f, int bounce, int refill) {
    __shared__ int stack[StackShared::kShared * PTB_RC_THREADS];
    int* ctr = f.ctr + bounce * CTR_STRIDE;
    const int count = ctr[CTR_SHADOW];
    int* cursor = &ctr[CTR_WORK_SHADOW];
    StackShared st(stack + threadIdx.x);
    const unsigned lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    int ray = -1;
    WalkAnyWide w;
    w.reset();
    bool exhausted = false;
    for (;;) {
        const unsigned idle_mask = __ballot_sync(0xffffffffu, ray < 0);
        const int n_idle = __popc(idle_mask);
        if (exhausted && n_idle == 32) break;
        if (!exhausted && n_idle >= refill) {
            int base = 0;
            if (lane == 0) base = atomicAdd(cursor, n_idle);
            base = __shfl_sync(0xffffffffu, base, 0);
            exhausted = base + n_idle >= count;
            const int i = base + __popc(idle_mask & lt);
            if (ray < 0 && i < count) {
                ray = i;
                float4 o4 = __ldg(&f.shadow.ray_ori_tmax[i]);
                float4 d4 = __ldg(&f.shadow.ray_dir[i]);
                w.init(f.bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), o4.w);
            }
        }
        const int want = exhausted ? 32 : refill;
        int n_done;
        do {
#pragma unroll
            for (int k = 0; k < STEPS; k++)
                if (w.busy()) w.step(f.bvh, st);
            n_done = __popc(__ballot_sync(0xffffffffu, !w.busy()));
        } while (n_done < want);
        if (ray >= 0 && !w.busy()) w.confirm(f.bvh);        // a refused triangle hit: the lane is busy again and keeps its ray
        if (ray >= 0 && !w.busy()) {
            int res = w.result;
            if (res < 0) res = trace_any<PTB_EXT_STACK>(f.bvh, w.r.o, w.r.d, w.tmax, stack + threadIdx.x, PTB_RC_THREADS);
            if (res == 0) {   // visible: AtomicAddVec3(&d_direct_image[pixel], radiance), src/kernels/ray_cast.cu:193-199
                int slot = __ldg(&f.shadow.slot[ray]);
                float4 rad = __ldg(&f.shadow.radiance[ray]);
                float* dst = f.direct + 3 * (size_t)slot;
                atomicAdd(dst + 0, rad.x);
                atomicAdd(dst + 1, rad.y);
                atomicAdd(dst + 2, rad.z);
            } else if (res < 0) {
                atomicExch(f.error_flag, 1);
            }
            ray = -1;
        }
    }
}

static inline int rc_grid(const LaunchCfg& c) { return c.num_sms * (c.rc_blocks > 0 ? c.rc_blocks : 8); }

template <class WALK, class STACK, int MINB, int STEPS = 1>
static void run_extension_pl(const FrameDev& f, int bounce, int* cursor, const LaunchCfg& c, cudaStream_t s, bool fuse_logic) {
    const int refill = c.rc_refill < 1 ? 32 : c.rc_refill > 32 ? 32 : c.rc_refill;
    const int grid = c.num_sms * MINB;
    if (fuse_logic) k_trace_extension_pl<true, WALK, STACK, MINB, STEPS><<<grid, PTB_RC_THREADS, 0, s>>>(f, bounce, cursor, refill);
    else k_trace_extension_pl<false, WALK, STACK, MINB, STEPS><<<grid, PTB_RC_THREADS, 0, s>>>(f, bounce, cursor, refill);
}

static void run_extension(const FrameDev& f, int bounce, int* cursor, const LaunchCfg& c_in, cudaStream_t s, bool fuse_logic = false) {
    LaunchCfg c = c_in;
    if (bounce == 0 && !c.rc_b0) c.rc_mode = 0;      // primary rays are coherent and of equal length: one batch per fetch is the best schedule
    // rc_mode: 0 = one batch per fetch on the fp32 records (below); persistent lanes on 1 = the fp32 records, 2 = the 32-byte
    // quantised binary records, 3 = the 4-wide quantised records.  rc_stack: 1 = hybrid stack (8 entries in shared memory,
    // the rest local).  rc_minb: resident blocks per SM the kernel is built and launched for (10: 48 registers, 9: 56, 8: 64).
    // rc_steps: traversal steps between two ballots.
#define PTB_PL(W, S, B, T) return run_extension_pl<W, S, B, T>(f, bounce, cursor, c, s, fuse_logic)
#define PTB_PL_T(W, S, B) do { if (c.rc_steps >= 8) PTB_PL(W, S, B, 8); if (c.rc_steps >= 4) PTB_PL(W, S, B, 4); if (c.rc_steps >= 2) PTB_PL(W, S, B, 2); PTB_PL(W, S, B, 1); } while (0)
#define PTB_PL_B(W, S) do { if (c.rc_minb == 8) PTB_PL_T(W, S, 8); if (c.rc_minb == 9) PTB_PL_T(W, S, 9); PTB_PL_T(W, S, 10); } while (0)
    if (c.rc_mode == 1) { if (c.rc_stack) PTB_PL_B(WalkBinary, StackHybrid<8>); else PTB_PL_B(WalkBinary, StackShared); }
    if (c.rc_mode == 2) { if (c.rc_stack) PTB_PL_B(WalkBq, StackHybrid<8>); else PTB_PL_B(WalkBq, StackShared); }
    if (c.rc_mode == 3) PTB_PL_B(WalkWide, StackShared);      // (hybrid stacks of 8 / 16 shared entries: 10.52-10.61 ms against 10.19)
#undef PTB_PL_B
#undef PTB_PL_T
#undef PTB_PL
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
    if (c.shadow_deep == 0 && wide && c.sh_mode == 1) {
        const int refill = c.sh_refill < 1 ? 32 : c.sh_refill > 32 ? 32 : c.sh_refill;
#define PTB_SPL(B, T) do { k_trace_shadow_pl<B, T><<<c.num_sms * B, PTB_RC_THREADS, 0, s>>>(f, bounce, refill); return 1; } while (0)
#define PTB_SPL_T(B) do { if (c.sh_steps >= 8) PTB_SPL(B, 8); if (c.sh_steps >= 4) PTB_SPL(B, 4); if (c.sh_steps >= 2) PTB_SPL(B, 2); PTB_SPL(B, 1); } while (0)
        if (c.sh_minb >= 12) PTB_SPL_T(12);
        if (c.sh_minb == 8) PTB_SPL_T(8);
        if (c.sh_minb == 9) PTB_SPL_T(9);
        PTB_SPL_T(10);
#undef PTB_SPL_T
#undef PTB_SPL
    }
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
