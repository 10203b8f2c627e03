// multigpu.cu -- the multi-GPU half of the C ABI (include/ptb200.h, "multi-GPU"): one frame over several B200s with NO
// collective library on the data path.
//
// The reference renders on device 0 of one process (src/main.cpp:168) and keeps the frame in two accumulators it resolves
// with SendImageToPBOKernel (src/pathtrace_wavefront.cu:47-116).  Here every GPU holds a scene + LBVH replica and renders
// its share -- an interleaved stripe set of the frame (tile group) and / or a subset of the sample indices (sample group)
// -- and the RESOLVE KERNEL ITSELF is the gather: each GPU's k_resolve_scatter writes the sums of the pixels it owns
// straight into the root GPU's frame buffer through NVLink peer stores (slot -> global pixel through TileMap), so the
// exchange is the tail of a kernel that runs anyway, not a separate collective with its own staging copies.  The root then
// folds the sample groups in a fixed order (deterministic) and applies the output format.
//   * one process, several devices: ptb_multi_* (cudaDeviceEnablePeerAccess; without P2P the stripes go through
//     cudaMemcpyPeerAsync and are scattered by the root);
//   * one process per GPU (torchrun): ptb_gather_export / ptb_gather_open hand the root's buffer to the other ranks as a
//     CUDA IPC mapping and ptb_resolve_into writes into it; the ranks only need a barrier afterwards (multigpu.py).
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../cuda/render.cuh"
#include "errors.hpp"
#include "renderer_internal.hpp"

using namespace ptb;

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(PTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

namespace ptb {

// direct + indirect of every owned pixel -> its GLOBAL position in a full-frame float3 buffer (possibly on another GPU)
__global__ void __launch_bounds__(256) k_resolve_scatter(const float* __restrict__ direct, const float* __restrict__ indirect, int n, TileMap tile,
                                                         float* __restrict__ frame) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t g = (size_t)tile.global_pixel(i);
    frame[3 * g + 0] = direct[3 * (size_t)i + 0] + indirect[3 * (size_t)i + 0];
    frame[3 * g + 1] = direct[3 * (size_t)i + 1] + indirect[3 * (size_t)i + 1];
    frame[3 * g + 2] = direct[3 * (size_t)i + 2] + indirect[3 * (size_t)i + 2];
}

// a rank's compact tile (owned pixels in slot order) -> full frame: the root-side half of the no-P2P fallback
__global__ void __launch_bounds__(256) k_scatter_tile(const float* __restrict__ tile_rgb, int n, TileMap tile, float* __restrict__ frame) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t g = (size_t)tile.global_pixel(i);
    frame[3 * g + 0] = tile_rgb[3 * (size_t)i + 0];
    frame[3 * g + 1] = tile_rgb[3 * (size_t)i + 1];
    frame[3 * g + 2] = tile_rgb[3 * (size_t)i + 2];
}

// fold the sample groups (fixed order: a deterministic sum) and write the requested format; arithmetic of k_resolve (stages.cu)
__global__ void __launch_bounds__(256) k_combine(const float* __restrict__ slices, int num_slices, size_t pixels, void* dst, int format, int samples) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float s = slices[3 * i + k];
        for (int j = 1; j < num_slices; j++) s = __fadd_rn(s, slices[(size_t)j * 3 * pixels + 3 * i + k]);
        c[k] = s;
    }
    if (format == PTB_IMG_SUM_RGB32F) {
        float* o = (float*)dst + 3 * i;
        o[0] = c[0]; o[1] = c[1]; o[2] = c[2];
        return;
    }
    const float div = (float)samples;
    c[0] = c[0] / div; c[1] = c[1] / div; c[2] = c[2] / div;
    if (format == PTB_IMG_MEAN_RGB32F) {
        float* o = (float*)dst + 3 * i;
        o[0] = c[0]; o[1] = c[1]; o[2] = c[2];
        return;
    }
    uchar4 px;
    px.x = (unsigned char)min(max((int)(powf(c[0], 1.0f / 2.2f) * 255.0f), 0), 255);
    px.y = (unsigned char)min(max((int)(powf(c[1], 1.0f / 2.2f) * 255.0f), 0), 255);
    px.z = (unsigned char)min(max((int)(powf(c[2], 1.0f / 2.2f) * 255.0f), 0), 255);
    px.w = 255;
    ((uchar4*)dst)[i] = px;
}

}  // namespace ptb

// ------------------------------------------------------------------------------------------------ gather buffer
struct ptb_gather {
    int device = 0;              // device the buffer lives on (the root)
    int width = 0, height = 0, slices = 1;
    float* frame = nullptr;      // [slices][height * width][3]
    bool owner = true;           // false: an IPC mapping of another process's buffer
    void* out_scratch = nullptr; // root: output staging for host destinations
    size_t out_bytes = 0;
};

static int enable_peer(int from, int to) {
    if (from == to) return PTB_OK;
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, from, to));
    if (!can) return fail(PTB_ERR_UNSUPPORTED, "no peer access between the two devices");
    CU(cudaSetDevice(from));
    cudaError_t e = cudaDeviceEnablePeerAccess(to, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return PTB_OK; }
    if (e != cudaSuccess) return fail(PTB_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
    return PTB_OK;
}

extern "C" {

int ptb_gather_create(int32_t device, int32_t width, int32_t height, int32_t slices, ptb_gather** out) {
    if (!out || width <= 0 || height <= 0 || slices < 1 || slices > 64) return fail(PTB_ERR_INVALID, "ptb_gather_create: bad argument");
    *out = nullptr;
    if (device < 0 || device >= ptb_device_count()) return fail(PTB_ERR_INVALID, "ptb_gather_create: device ordinal out of range");
    CU(cudaSetDevice(device));
    ptb_gather* g = new ptb_gather();
    g->device = device; g->width = width; g->height = height; g->slices = slices;
    const size_t bytes = (size_t)slices * width * height * 3 * sizeof(float);
    cudaError_t e = cudaMalloc((void**)&g->frame, bytes);        // plain cudaMalloc: exportable through CUDA IPC
    if (e != cudaSuccess) { delete g; return fail(PTB_ERR_CUDA, std::string("ptb_gather_create: ") + cudaGetErrorString(e)); }
    cudaMemset(g->frame, 0, bytes);
    *out = g;
    return PTB_OK;
}

int ptb_gather_export(ptb_gather* g, void* handle, size_t bytes) {
    if (!g || !handle) return fail(PTB_ERR_INVALID, "ptb_gather_export: null argument");
    if (!g->owner) return fail(PTB_ERR_STATE, "ptb_gather_export: only the creating process can export the buffer");
    if (bytes < sizeof(cudaIpcMemHandle_t)) return fail(PTB_ERR_INVALID, "ptb_gather_export: handle buffer smaller than PTB_GATHER_HANDLE_BYTES");
    CU(cudaSetDevice(g->device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, g->frame));
    memcpy(handle, &h, sizeof(h));
    return PTB_OK;
}

int ptb_gather_open(int32_t device, const void* handle, int32_t width, int32_t height, int32_t slices, ptb_gather** out) {
    if (!handle || !out || width <= 0 || height <= 0 || slices < 1) return fail(PTB_ERR_INVALID, "ptb_gather_open: bad argument");
    *out = nullptr;
    CU(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    ptb_gather* g = new ptb_gather();
    g->device = device; g->width = width; g->height = height; g->slices = slices;
    g->frame = (float*)p;
    g->owner = false;
    *out = g;
    return PTB_OK;
}

int ptb_gather_free(ptb_gather* g) {
    if (!g) return PTB_OK;
    cudaSetDevice(g->device);
    if (g->frame) { if (g->owner) cudaFree(g->frame); else cudaIpcCloseMemHandle(g->frame); }
    if (g->out_scratch) cudaFree(g->out_scratch);
    delete g;
    return PTB_OK;
}

int ptb_resolve_into(ptb_renderer* r, ptb_gather* g, int32_t slice) {
    if (!r || !g) return fail(PTB_ERR_INVALID, "ptb_resolve_into: null argument");
    if (slice < 0 || slice >= g->slices) return fail(PTB_ERR_INVALID, "ptb_resolve_into: slice out of range");
    const FrameDev& f = ptb_internal_frame(r);
    if (f.tile.width != g->width || f.tile.height != g->height) return fail(PTB_ERR_INVALID, "ptb_resolve_into: frame size differs from the gather buffer's");
    const int dev = ptb_internal_device(r);
    if (g->owner && dev != g->device) { int rc = enable_peer(dev, g->device); if (rc) return rc; }
    CU(cudaSetDevice(dev));
    float* dst = g->frame + (size_t)slice * 3 * g->width * g->height;
    k_resolve_scatter<<<(f.num_paths + 255) / 256, 256, 0, ptb_internal_stream(r)>>>(f.direct, f.indirect, f.num_paths, f.tile, dst);
    ptb_internal_count_launches(r, 1);
    CU(cudaGetLastError());
    return PTB_OK;
}

int ptb_gather_combine(ptb_gather* g, void* dst, int32_t format, int32_t samples, int32_t dst_on_device, void* cuda_stream) {
    if (!g || !dst) return fail(PTB_ERR_INVALID, "ptb_gather_combine: null argument");
    if (!g->owner) return fail(PTB_ERR_STATE, "ptb_gather_combine: only the process that created the buffer combines it");
    if (format < PTB_IMG_SUM_RGB32F || format > PTB_IMG_RGBA8) return fail(PTB_ERR_INVALID, "ptb_gather_combine: format must be PTB_IMG_SUM_RGB32F, _MEAN_RGB32F or _RGBA8");
    if (samples < 1) return fail(PTB_ERR_INVALID, "ptb_gather_combine: samples must be >= 1");
    CU(cudaSetDevice(g->device));
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const size_t P = (size_t)g->width * g->height;
    const size_t bytes = P * (format == PTB_IMG_RGBA8 ? 4 : 12);
    void* out = dst;
    if (!dst_on_device) {
        if (g->out_bytes < bytes) {
            if (g->out_scratch) cudaFree(g->out_scratch);
            g->out_scratch = nullptr; g->out_bytes = 0;
            CU(cudaMalloc(&g->out_scratch, bytes));
            g->out_bytes = bytes;
        }
        out = g->out_scratch;
    }
    k_combine<<<(unsigned)((P + 255) / 256), 256, 0, s>>>(g->frame, g->slices, P, out, format, samples);
    CU(cudaGetLastError());
    if (!dst_on_device) {
        CU(cudaMemcpyAsync(dst, out, bytes, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
    }
    return PTB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ one process, several devices
struct ptb_multi {
    int n = 0, tile_groups = 1, sample_groups = 1;
    std::vector<int> devices;
    std::vector<ptb_renderer*> r;
    std::vector<cudaEvent_t> done;         // per device: its resolve has been enqueued / finished
    std::vector<bool> peer;                // device k can store into the root's memory
    std::vector<float*> tile_local;        // no-P2P fallback: compact resolved tile on device k ...
    std::vector<float*> tile_root;         // ... and its landing buffer on the root
    ptb_gather* gather = nullptr;
    cudaStream_t root_stream = nullptr;
    int steps_rendered = 0;                // samples per pixel accumulated = steps_rendered * sample_groups
};

extern "C" {

void ptb_multi_config_default(ptb_multi_config* c) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->n_devices = 1;
    c->stripe_rows = 8;
    c->use_graph = 1;
}

int ptb_multi_free(ptb_multi* m) {
    if (!m) return PTB_OK;
    for (size_t k = 0; k < m->r.size(); k++) {
        if (m->r[k]) ptb_free(m->r[k]);
        if (k < m->done.size() && m->done[k]) { cudaSetDevice(m->devices[k]); cudaEventDestroy(m->done[k]); }
        if (k < m->tile_local.size() && m->tile_local[k]) { cudaSetDevice(m->devices[k]); cudaFree(m->tile_local[k]); }
        if (k < m->tile_root.size() && m->tile_root[k]) { cudaSetDevice(m->devices[0]); cudaFree(m->tile_root[k]); }
    }
    if (m->root_stream) { cudaSetDevice(m->devices[0]); cudaStreamDestroy(m->root_stream); }
    ptb_gather_free(m->gather);
    delete m;
    return PTB_OK;
}

int ptb_multi_init(const ptb_scene* scene, const ptb_multi_config* cfg, ptb_multi** out) {
    if (!scene || !cfg || !out) return fail(PTB_ERR_INVALID, "ptb_multi_init: null argument");
    *out = nullptr;
    const int n = cfg->n_devices;
    if (n < 1 || n > PTB_MAX_DEVICES) return fail(PTB_ERR_INVALID, "ptb_multi_init: n_devices must be in [1, 16]");
    int G = cfg->tile_groups > 0 ? cfg->tile_groups : n;
    if (n % G != 0) return fail(PTB_ERR_INVALID, "ptb_multi_init: tile_groups must divide n_devices");
    const int S = n / G;
    const int ndev = ptb_device_count();
    for (int k = 0; k < n; k++) {
        if (cfg->devices[k] < 0 || cfg->devices[k] >= ndev) return fail(PTB_ERR_INVALID, "ptb_multi_init: device ordinal out of range");
    }
    ptb_multi* m = new ptb_multi();
    m->n = n; m->tile_groups = G; m->sample_groups = S;
    m->devices.assign(cfg->devices, cfg->devices + n);
    m->r.assign(n, nullptr); m->done.assign(n, nullptr); m->peer.assign(n, true);
    m->tile_local.assign(n, nullptr); m->tile_root.assign(n, nullptr);
#define TRYM(x) do { int rc_ = (x); if (rc_) { ptb_multi_free(m); return rc_; } } while (0)
#define TRYMCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { ptb_multi_free(m); return fail(PTB_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } } while (0)
    ptb_scene_info info;
    TRYM(ptb_scene_get_info(scene, &info));
    for (int k = 0; k < n; k++) {
        ptb_config c;
        ptb_config_default(&c);
        c.device = m->devices[k];
        c.trace_depth = cfg->trace_depth;
        c.rank = k % G; c.world = G; c.stripe_rows = cfg->stripe_rows > 0 ? cfg->stripe_rows : 8;
        c.use_graph = cfg->use_graph; c.megakernel = cfg->megakernel; c.env_nee = cfg->env_nee;
        TRYM(ptb_init(scene, &c, &m->r[k]));
        TRYMCU(cudaSetDevice(m->devices[k]));
        TRYMCU(cudaEventCreateWithFlags(&m->done[k], cudaEventDisableTiming));
    }
    TRYM(ptb_gather_create(m->devices[0], info.width, info.height, S, &m->gather));
    TRYMCU(cudaSetDevice(m->devices[0]));
    TRYMCU(cudaStreamCreateWithFlags(&m->root_stream, cudaStreamNonBlocking));
    const bool force_copy = std::getenv("PTB_MULTI_NO_P2P") != nullptr;     // exercise the fallback on a P2P box
    for (int k = 1; k < n; k++) {
        if (force_copy || enable_peer(m->devices[k], m->devices[0]) != PTB_OK) {
            m->peer[k] = false;
            ptb_tile_info ti;
            TRYM(ptb_get_tile_info(m->r[k], &ti));
            TRYMCU(cudaSetDevice(m->devices[k]));
            TRYMCU(cudaMalloc((void**)&m->tile_local[k], (size_t)ti.owned_pixels * 3 * sizeof(float)));
            TRYMCU(cudaSetDevice(m->devices[0]));
            TRYMCU(cudaMalloc((void**)&m->tile_root[k], (size_t)ti.owned_pixels * 3 * sizeof(float)));
        }
    }
#undef TRYM
#undef TRYMCU
    *out = m;
    return PTB_OK;
}

int ptb_multi_render(ptb_multi* m, int32_t first_step, int32_t count) {
    if (!m) return fail(PTB_ERR_INVALID, "ptb_multi_render: null handle");
    if (first_step < 1 || count < 0) return fail(PTB_ERR_INVALID, "ptb_multi_render: steps are 1-based");
    const int G = m->tile_groups, S = m->sample_groups;
    for (int k = 0; k < m->n; k++) {
        const int sg = k / G;
        for (int step = first_step; step < first_step + count; step++) {
            const int it = (step - 1) * S + sg + 1;          // sample group sg takes iterations sg+1, sg+1+S, ...
            if (step == 1 && it != 1) { int rc = ptb_clear_accumulators(m->r[k]); if (rc) return rc; }   // iteration 1 clears by itself
            int rc = ptb_render(m->r[k], it);
            if (rc) return rc;
        }
    }
    if (first_step == 1) m->steps_rendered = 0;
    m->steps_rendered = first_step + count - 1;
    return PTB_OK;
}

int ptb_multi_resolve(ptb_multi* m, void* dst, int32_t format, int32_t dst_on_device) {
    if (!m || !dst) return fail(PTB_ERR_INVALID, "ptb_multi_resolve: null argument");
    if (m->steps_rendered < 1) return fail(PTB_ERR_STATE, "ptb_multi_resolve: nothing rendered yet");
    const int G = m->tile_groups;
    const size_t P = (size_t)m->gather->width * m->gather->height;
    CU(cudaSetDevice(m->devices[0]));
    for (int k = 0; k < m->n; k++) {
        const int sg = k / G;
        if (m->peer[k]) {
            int rc = ptb_resolve_into(m->r[k], m->gather, sg);       // peer stores into the root's buffer
            if (rc) return rc;
            CU(cudaSetDevice(m->devices[k]));
            CU(cudaEventRecord(m->done[k], ptb_internal_stream(m->r[k])));
        } else {
            ptb_tile_info ti;
            ptb_get_tile_info(m->r[k], &ti);
            const size_t bytes = (size_t)ti.owned_pixels * 3 * sizeof(float);
            int rc = ptb_resolve(m->r[k], m->tile_local[k], PTB_IMG_SUM_RGB32F, 1, 1);
            if (rc) return rc;
            CU(cudaSetDevice(m->devices[k]));
            cudaStream_t s = ptb_internal_stream(m->r[k]);
            CU(cudaMemcpyPeerAsync(m->tile_root[k], m->devices[0], m->tile_local[k], m->devices[k], bytes, s));
            CU(cudaEventRecord(m->done[k], s));
        }
    }
    CU(cudaSetDevice(m->devices[0]));
    for (int k = 0; k < m->n; k++) {
        CU(cudaStreamWaitEvent(m->root_stream, m->done[k], 0));
        if (!m->peer[k]) {
            const FrameDev& f = ptb_internal_frame(m->r[k]);
            k_scatter_tile<<<(f.num_paths + 255) / 256, 256, 0, m->root_stream>>>(m->tile_root[k], f.num_paths, f.tile,
                                                                                 m->gather->frame + (size_t)(k / G) * 3 * P);
        }
    }
    return ptb_gather_combine(m->gather, dst, format, m->steps_rendered * m->sample_groups, dst_on_device, m->root_stream);
}

int ptb_multi_sync(ptb_multi* m) {
    if (!m) return fail(PTB_ERR_INVALID, "ptb_multi_sync: null handle");
    for (int k = 0; k < m->n; k++) { int rc = ptb_sync(m->r[k]); if (rc) return rc; }
    CU(cudaSetDevice(m->devices[0]));
    CU(cudaStreamSynchronize(m->root_stream));
    return PTB_OK;
}

int ptb_multi_get_renderer(ptb_multi* m, int32_t index, ptb_renderer** out) {
    if (!m || !out || index < 0 || index >= m->n) return fail(PTB_ERR_INVALID, "ptb_multi_get_renderer: bad argument");
    *out = m->r[index];
    return PTB_OK;
}

int ptb_multi_get_info(const ptb_multi* m, ptb_multi_info* out) {
    if (!m || !out) return fail(PTB_ERR_INVALID, "ptb_multi_get_info: null argument");
    memset(out, 0, sizeof(*out));
    out->n_devices = m->n; out->tile_groups = m->tile_groups; out->sample_groups = m->sample_groups;
    out->width = m->gather->width; out->height = m->gather->height;
    out->samples = m->steps_rendered * m->sample_groups;
    for (int k = 0; k < m->n; k++) out->peer_stores += m->peer[k] ? 1 : 0;
    return PTB_OK;
}

}  // extern "C"
