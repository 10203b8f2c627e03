// noalloc.cu -- ptb_trace_closest_noalloc: the batched closest-hit RayCast for callers that may not allocate.
//
// No counterpart in the reference.  ptb_trace_closest (renderer.cu) stages the
// caller's rays and the results in a scratch buffer it grows with cudaMalloc, and times the kernel with freshly created
// events.  Code under CUDA graph capture, tight serving loops and harnesses that forbid allocation cannot call that.  This
// variant keeps ALL of its scratch -- the staged rays, the staged results and the work cursor of the persistent kernel -- in
// one file-scope `static __device__` array of a fixed size, and calls no cudaMalloc, cudaFree, cudaEventCreate or any other
// allocating routine.
//
// Rules that follow from the static array:
//   * it is PTB_NOALLOC_MAX_RAYS * 48 bytes + one 256-byte header (12 MB + 256 B), and every GPU that loads the library holds
//     it until the process exits, used or not;
//   * in host code the array's NAME is a host-side stand-in, not a device address -- a kernel handed `g_noalloc_scratch` from
//     the host would fault or, where the GPU can reach host memory, silently use the wrong memory.  The device address comes
//     from cudaGetSymbolAddress, once per device, with that device current (scratch_for_device below), in THIS file;
//   * every call on a device shares the array and it keeps the last call's contents: the work cursor is cleared at the start
//     of every call (the only part that must start at zero; rays and results are overwritten for [0, n) before they are read);
//     two calls that overlap on the same GPU (different streams / host threads) race on it -- callers serialise them.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "../cuda/raycast.cuh"
#include "errors.hpp"
#include "renderer_internal.hpp"

using namespace ptb;

namespace {

constexpr size_t kHeaderBytes = 256;                                   // [0, 4): work cursor; rest of the header unused
constexpr size_t kRayBytes = 16 + 16 + 4 * 4;                          // ori, dir, hit, t, u, v
constexpr size_t kScratchBytes = kHeaderBytes + (size_t)PTB_NOALLOC_MAX_RAYS * kRayBytes;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mutex;
unsigned char* g_device_ptr[kMaxDevices] = {};                         // device address of g_noalloc_scratch, per device

// k_batch_closest (raycast.cu) with its cursor and all six arrays carved out of one scratch block
__global__ void __launch_bounds__(PTB_RC_THREADS) k_batch_closest_static(BvhDev bvh, unsigned char* scratch, int n) {
    __shared__ int stack[PTB_EXT_STACK * PTB_RC_THREADS];
    int* work = reinterpret_cast<int*>(scratch);
    const float4* ori = reinterpret_cast<const float4*>(scratch + kHeaderBytes);
    const float4* dir = ori + PTB_NOALLOC_MAX_RAYS;
    int* hit = reinterpret_cast<int*>(const_cast<float4*>(dir + PTB_NOALLOC_MAX_RAYS));
    float* t = reinterpret_cast<float*>(hit + PTB_NOALLOC_MAX_RAYS);
    float* u = t + PTB_NOALLOC_MAX_RAYS;
    float* v = u + PTB_NOALLOC_MAX_RAYS;
    int* my_stack = stack + threadIdx.x;
    for (;;) {
        int base = 0;
        if ((threadIdx.x & 31) == 0) base = atomicAdd(work, 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const int i = base + (threadIdx.x & 31);
        if (i >= n) continue;
        const float4 o4 = ori[i], d4 = dir[i];
        const Hit h = trace_closest(bvh, Vec3(o4.x, o4.y, o4.z), Vec3(d4.x, d4.y, d4.z), my_stack, PTB_RC_THREADS);
        hit[i] = h.prim;
        t[i] = h.prim == -1 ? -1.0f : h.t;
        u[i] = h.u;
        v[i] = h.v;
    }
}

// Device address of the static array on `device`, which must be the current device.  Once per device.
int scratch_for_device(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) return fail(PTB_ERR_INVALID, "ptb_trace_closest_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_device_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(PTB_ERR_CUDA, std::string("ptb_trace_closest_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_device_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_device_ptr[device];
    return PTB_OK;
}

}  // namespace

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(PTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

extern "C" int ptb_trace_closest_noalloc(ptb_renderer* r, const float* ori4, const float* dir4, int32_t n, int32_t* hit_geom_id, float* t,
                                         float* u, float* v) {
    if (!r || !ori4 || !dir4 || !hit_geom_id || !t || !u || !v) return fail(PTB_ERR_INVALID, "ptb_trace_closest_noalloc: null argument");
    if (n < 0) return fail(PTB_ERR_INVALID, "ptb_trace_closest_noalloc: negative count");
    if (n > PTB_NOALLOC_MAX_RAYS)
        return fail(PTB_ERR_INVALID, "ptb_trace_closest_noalloc: " + std::to_string(n) + " rays exceed PTB_NOALLOC_MAX_RAYS (" +
                                         std::to_string(PTB_NOALLOC_MAX_RAYS) + "): split the batch or use ptb_trace_closest");
    if (n == 0) return PTB_OK;
    const int device = ptb_internal_device(r);
    CU(cudaSetDevice(device));
    unsigned char* scratch = nullptr;
    int rc = scratch_for_device(device, &scratch);
    if (rc) return rc;
    cudaStream_t s = ptb_internal_stream(r);
    const FrameDev& f = ptb_internal_frame(r);
    const size_t N = (size_t)n, M = (size_t)PTB_NOALLOC_MAX_RAYS;
    unsigned char* rays = scratch + kHeaderBytes;
    CU(cudaMemsetAsync(scratch, 0, sizeof(int), s));                                   // the cursor must start at zero
    CU(cudaMemcpyAsync(rays, ori4, 16 * N, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(rays + 16 * M, dir4, 16 * N, cudaMemcpyHostToDevice, s));
    const int grid = ptb_internal_num_sms(r) * 8;
    k_batch_closest_static<<<grid, PTB_RC_THREADS, 0, s>>>(f.bvh, scratch, n);
    CU(cudaGetLastError());
    ptb_internal_count_launches(r, 1);
    CU(cudaMemcpyAsync(hit_geom_id, rays + 32 * M, 4 * N, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(t, rays + 36 * M, 4 * N, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(u, rays + 40 * M, 4 * N, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(v, rays + 44 * M, 4 * N, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return PTB_OK;
}
