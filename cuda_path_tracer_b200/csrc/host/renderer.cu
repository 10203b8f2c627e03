// renderer.cu -- the ptb_renderer half of the C ABI (include/ptb200.h): device upload, LBVH build,
// per-frame CUDA graph, staged (parity) execution, read-backs.
// Replaces pathtrace_wavefront::{PathtraceInit, Pathtrace, PathtraceFree} (reference
// src/pathtrace_wavefront.cu:144-413) and WavefrontPathTracerState (src/wavefront_internal.cu).
#include <cmath>
#include <utility>
#include <cstdlib>
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <vector>

#include "../cuda/lbvh.cuh"
#include "../cuda/render.cuh"
#include "errors.hpp"
#include "scene.hpp"

using namespace ptb;

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(PTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

struct ptb_renderer {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t side = nullptr;          // forked branch of the frame (shadow rays beside extension rays)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaStream_t shade_side[3] = {nullptr, nullptr, nullptr};   // forked branches of Shade: PBR, mirror, dielectric beside diffuse
    cudaEvent_t ev_sfork = nullptr, ev_sjoin[3] = {nullptr, nullptr, nullptr};
    ptb_config cfg{};

    // scene replica on the device
    std::vector<void*> allocs;               // everything cudaMalloc'ed here (freed in ptb_free)
    std::vector<cudaArray_t> arrays;
    std::vector<cudaTextureObject_t> textures;
    LbvhBuffers lbvh;
    FrameDev frame{};
    LaunchCfg launch{};
    float last_batch_ms = 0.0f;
    bool gbuffer = false;                 // ptb_config.gbuffer: first-bounce G-buffer kept every frame
    GBufferDev gb{};
    bool denoiser = false;                // ptb_config.denoiser: SVGF over a 1-spp frame, every frame
    SvgfDev svgf{};
    std::vector<std::pair<void*, size_t>> svgf_state;   // buffers zeroed when iteration 1 restarts the history
    size_t device_bytes = 0;

    // frame graph
    cudaGraphExec_t graph_exec = nullptr;
    int device_iteration = -1;               // value *frame.iteration will hold when the stream drains

    // staged execution state
    int stage_bounce = 0;                    // bounce the next TRACE / LOGIC / SHADE works on
    int logic_bounce = 0;                    // bounce of the last LOGIC (material queue counters)
    bool stage_traced = false;               // staged frame: TRACE of stage_bounce has run, its SHADE has not
    bool hit_valid = false;                  // the hit records belong to the current frame (false between RAYGEN and the first trace)

    // scratch for batched tracing / timing
    void* scratch = nullptr;
    size_t scratch_bytes = 0;
    void* flush_buf = nullptr;               // ptb_time_kernel(cold): written between launches to flush L2
    int* cursors = nullptr;                  // kMaxTimed work cursors + 1 for batches
    std::vector<int32_t> owned_pixels_cache;

    // stats
    uint64_t frames = 0, kernel_launches = 0;
    float lbvh_build_ms = 0.0f;
    int lbvh_max_depth = 0;
    int launches_per_frame = 0;
};

static constexpr int kMaxTimed = 4096;

#include "renderer_internal.hpp"
const ptb::FrameDev& ptb_internal_frame(const ptb_renderer* r) { return r->frame; }
int ptb_internal_device(const ptb_renderer* r) { return r->device; }
cudaStream_t ptb_internal_stream(const ptb_renderer* r) { return r->stream; }
void ptb_internal_count_launches(ptb_renderer* r, int n) { r->kernel_launches += n; }
int ptb_internal_num_sms(const ptb_renderer* r) { return r->launch.num_sms; }

template <typename T>
static int dev_alloc(ptb_renderer* r, T*& p, size_t count) {
    size_t bytes = (count ? count : 1) * sizeof(T);
    void* q = nullptr;
    CU(cudaMalloc(&q, bytes));
    r->allocs.push_back(q);
    r->device_bytes += bytes;
    p = (T*)q;
    return PTB_OK;
}

template <typename T>
static int dev_upload(ptb_renderer* r, T*& p, const void* src, size_t count) {
    int rc = dev_alloc(r, p, count);
    if (rc) return rc;
    if (count) CU(cudaMemcpyAsync((void*)p, src, count * sizeof(T), cudaMemcpyHostToDevice, r->stream));
    return PTB_OK;
}

static int ensure_scratch(ptb_renderer* r, size_t bytes) {
    if (bytes <= r->scratch_bytes) return PTB_OK;
    if (r->scratch) cudaFree(r->scratch);
    r->scratch = nullptr;
    r->scratch_bytes = 0;
    CU(cudaMalloc(&r->scratch, bytes));
    r->scratch_bytes = bytes;
    return PTB_OK;
}

// Scene::createTexture (src/scene.h:94-145): wrap in X always, Y per texture, linear, normalised
// coordinates, uchar4 -> normalised float.
static int upload_texture(ptb_renderer* r, const HostTexture& t) {
    cudaChannelFormatDesc desc;
    size_t elem;
    const void* src;
    if (!t.is_float) { desc = cudaCreateChannelDesc<uchar4>(); elem = 4; src = t.rgba8.data(); }
    else if (t.channels == 4) { desc = cudaCreateChannelDesc<float4>(); elem = 16; src = t.f32.data(); }
    else { desc = cudaCreateChannelDesc<float>(); elem = 4; src = t.f32.data(); }
    cudaArray_t arr = nullptr;
    CU(cudaMallocArray(&arr, &desc, t.width, t.height));
    r->arrays.push_back(arr);
    CU(cudaMemcpy2DToArray(arr, 0, 0, src, t.width * elem, t.width * elem, t.height, cudaMemcpyHostToDevice));
    cudaResourceDesc res;
    memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray;
    res.res.array.array = arr;
    cudaTextureDesc td;
    memset(&td, 0, sizeof(td));
    td.readMode = t.is_float ? cudaReadModeElementType : cudaReadModeNormalizedFloat;
    td.addressMode[0] = cudaAddressModeWrap;
    td.addressMode[1] = t.clamp_y ? cudaAddressModeClamp : cudaAddressModeWrap;
    td.filterMode = cudaFilterModeLinear;
    td.normalizedCoords = 1;
    cudaTextureObject_t obj = 0;
    CU(cudaCreateTextureObject(&obj, &res, &td, nullptr));
    r->textures.push_back(obj);
    r->device_bytes += (size_t)t.width * t.height * elem;
    return PTB_OK;
}

static int owned_rows(const ptb_config& c, int height) {
    if (c.world <= 1) return height;
    int rows = 0;
    for (int y = 0; y < height; y++)
        if ((y / c.stripe_rows) % c.world == c.rank) rows++;
    return rows;
}

// glm::lookAt / glm::perspective / mat4 * mat4 with GLM 0.9.6.3's operation order (gtc/matrix_transform.inl:212-231,
// 387-412; detail/type_mat4x4.inl:686-704), as the reference evaluates them on the host every frame
// (src/kernels/ray_gen.cu:92-103): view-projection of the scene camera for the G-buffer's motion vectors.
static Mat4 camera_view_proj(const ptb_camera& c) {
    using namespace ptb;
    V3 eye(c.position[0], c.position[1], c.position[2]), center(c.look_at[0], c.look_at[1], c.look_at[2]), up(c.up[0], c.up[1], c.up[2]);
    V3 f = normalize(center - eye);
    V3 s = normalize(cross(f, up));
    V3 u = cross(s, f);
    float view[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
    view[0][0] = s.x; view[1][0] = s.y; view[2][0] = s.z;
    view[0][1] = u.x; view[1][1] = u.y; view[2][1] = u.z;
    view[0][2] = -f.x; view[1][2] = -f.y; view[2][2] = -f.z;
    view[3][0] = -dot(s, eye); view[3][1] = -dot(u, eye); view[3][2] = dot(f, eye);
    const float fovy = c.fov[1] * 0.01745329251994329576923690768489f;      // glm::radians
    const float aspect = float(c.resolution[0]) / float(c.resolution[1]);
    const float z_near = 0.1f, z_far = 1000.0f;
    const float tan_half = std::tan(fovy / 2.0f);
    float proj[4][4] = {};
    proj[0][0] = 1.0f / (aspect * tan_half);
    proj[1][1] = 1.0f / tan_half;
    proj[2][2] = -(z_far + z_near) / (z_far - z_near);
    proj[2][3] = -1.0f;
    proj[3][2] = -(2.0f * z_far * z_near) / (z_far - z_near);
    Mat4 vp;
    for (int j = 0; j < 4; j++)
        for (int i = 0; i < 4; i++)
            vp.m[j][i] = ((proj[0][i] * view[j][0] + proj[1][i] * view[j][1]) + proj[2][i] * view[j][2]) + proj[3][i] * view[j][3];
    return vp;
}

static int enqueue_frame(ptb_renderer* r, cudaStream_t s) {
    const FrameDev& f = r->frame;
    int n = 0;
    if (r->cfg.megakernel) {   // pathtrace_megakernel::Pathtrace (src/pathtrace.cu:645-706): the frame is one kernel
        n += launch_megakernel(f, r->launch, s);
        n += launch_frame_end(f, r->launch, s);
        return n;
    }
    // with the denoiser every frame is a fresh 1-spp estimate (src/pathtrace_wavefront.cu:231-236)
    if (r->denoiser) n += launch_clear_accumulators(f, s);
    n += launch_raygen(f, r->launch, s);
    for (int b = 0; b < f.trace_depth; b++) {
        // whole frames run Logic inside the extension RayCast kernel; the staged API keeps the stages apart
        // Logic in the tail of the RayCast kernel: pays in the batch kernel and in the 64-register persistent-lane kernels (10.4 vs
        // 11.0 ms per frame), costs spills in the 48-register ones (fuse_logic = 2 keeps it to the batch kernel)
        const bool batch_kernel = r->launch.rc_mode == 0 || (b == 0 && !r->launch.rc_b0);
        const bool fuse = r->launch.fuse_logic == 1 || (r->launch.fuse_logic == 2 && batch_kernel);
        if (b > 0 && r->launch.overlap_shadow && r->side) {
            // TraceShadow(b) reads what Shade(b-1) queued and only adds to `direct`; TraceExtension(b) + Logic(b) touch
            // the path pool and `indirect`: independent until Shade(b), so they run as two branches of the graph and
            // each fills the other's tail.
            cudaEventRecord(r->ev_fork, s);
            cudaStreamWaitEvent(r->side, r->ev_fork, 0);
            n += launch_trace_shadow(f, b, r->launch, r->side);
            n += launch_trace_extension(f, b, r->launch, s, fuse);
            if (!fuse) n += launch_logic(f, b, r->launch, s);
            cudaEventRecord(r->ev_join, r->side);
            cudaStreamWaitEvent(s, r->ev_join, 0);
        } else {
            n += launch_trace(f, b, r->launch, s, fuse);
            if (!fuse) n += launch_logic(f, b, r->launch, s);
        }
        if (b == 0 && r->gbuffer) n += launch_gbuffer(f, r->gb, s);   // the reference's Logic writes it on the first bounce
        // Rays produced by the last Shade are never traced (loop bound src/pathtrace_wavefront.cu:281,
        // counters overwritten by the next RayGen, src/kernels/ray_gen.cu:121-124): skip it.
        if (b + 1 < f.trace_depth) {
            if (r->launch.shade_fork && r->shade_side[0]) {
                static const int side_type[3] = {MAT_PBR, MAT_REFLECT, MAT_REFRACT};
                cudaEventRecord(r->ev_sfork, s);
                for (int k = 0; k < 3; k++) {
                    cudaStreamWaitEvent(r->shade_side[k], r->ev_sfork, 0);
                    n += launch_shade_one(f, b, side_type[k], r->launch, r->shade_side[k]);
                    cudaEventRecord(r->ev_sjoin[k], r->shade_side[k]);
                }
                n += launch_shade_one(f, b, MAT_DIFFUSE, r->launch, s);
                for (int k = 0; k < 3; k++) cudaStreamWaitEvent(s, r->ev_sjoin[k], 0);
            } else {
                n += launch_shade(f, b, r->launch, s);
            }
        }
    }
    if (f.env.nee) n += launch_fold_env(f, s);
    if (r->denoiser) n += launch_svgf(f, r->svgf, r->gb, r->launch.num_sms, s);   // reads *f.iteration: before frame_end advances it
    n += launch_frame_end(f, r->launch, s);
    return n;
}

static int check_device_error(ptb_renderer* r) {
    int flag = 0;
    CU(cudaMemcpy(&flag, r->frame.error_flag, sizeof(int), cudaMemcpyDeviceToHost));
    if (flag) {
        cudaMemset(r->frame.error_flag, 0, sizeof(int));
        return fail(PTB_ERR_UNSUPPORTED, "shadow-ray traversal stack overflow (LBVH deeper than the stack)");
    }
    return PTB_OK;
}

extern "C" {

int ptb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

void ptb_config_default(ptb_config* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->device = 0;
    cfg->trace_depth = 0;
    cfg->rank = 0;
    cfg->world = 1;
    cfg->stripe_rows = 8;
    cfg->use_graph = 1;
}

int ptb_free(ptb_renderer* r) {
    if (!r) return PTB_OK;
    cudaSetDevice(r->device);
    // a caller-owned stream (ptb_set_stream) may already be destroyed: only the device-wide wait is safe then
    if (r->stream && r->own_stream) cudaStreamSynchronize(r->stream);
    else cudaDeviceSynchronize();
    if (r->graph_exec) cudaGraphExecDestroy(r->graph_exec);
    for (cudaTextureObject_t t : r->textures) cudaDestroyTextureObject(t);
    for (cudaArray_t a : r->arrays) cudaFreeArray(a);
    for (void* p : r->allocs) cudaFree(p);
    if (r->scratch) cudaFree(r->scratch);
    if (r->flush_buf) cudaFree(r->flush_buf);
    lbvh_free(r->lbvh);
    if (r->own_stream && r->stream) cudaStreamDestroy(r->stream);
    if (r->side) cudaStreamDestroy(r->side);
    if (r->ev_fork) cudaEventDestroy(r->ev_fork);
    if (r->ev_join) cudaEventDestroy(r->ev_join);
    for (int k = 0; k < 3; k++) { if (r->shade_side[k]) cudaStreamDestroy(r->shade_side[k]); if (r->ev_sjoin[k]) cudaEventDestroy(r->ev_sjoin[k]); }
    if (r->ev_sfork) cudaEventDestroy(r->ev_sfork);
    delete r;
    return PTB_OK;
}

int ptb_init(const ptb_scene* scene, const ptb_config* cfg_in, ptb_renderer** out) {
    if (!scene || !out) return fail(PTB_ERR_INVALID, "ptb_init: null argument");
    *out = nullptr;
    ptb_config cfg;
    if (cfg_in) cfg = *cfg_in; else ptb_config_default(&cfg);
    if (cfg.world < 1) cfg.world = 1;
    if (cfg.stripe_rows < 1) cfg.stripe_rows = 8;
    if (cfg.rank < 0 || cfg.rank >= cfg.world) return fail(PTB_ERR_INVALID, "ptb_init: rank outside [0, world)");
    if (cfg.megakernel && (cfg.gbuffer || cfg.denoiser || cfg.env_nee))
        return fail(PTB_ERR_UNSUPPORTED, "ptb_init: the megakernel path has no G-buffer, denoiser or environment NEE (src/pathtrace.cu has none)");
    {
        const std::string why = validate_scene(*scene);
        if (!why.empty()) return fail(PTB_ERR_INVALID, "ptb_init: " + why);
    }
    const int N = (int)scene->indices_matid.size();
    const int V = (int)scene->pos.size();
    if (N < 2) return fail(PTB_ERR_UNSUPPORTED, "ptb_init: the LBVH needs at least 2 triangles (the reference's hierarchy is undefined below that)");
    const int W = scene->camera.resolution[0], H = scene->camera.resolution[1];
    if (W <= 0 || H <= 0) return fail(PTB_ERR_INVALID, "ptb_init: camera resolution not set");
    int ndev = ptb_device_count();
    if (ndev == 0) return fail(PTB_ERR_CUDA, "ptb_init: no CUDA device (this library has no CPU fallback)");
    if (cfg.device < 0 || cfg.device >= ndev) return fail(PTB_ERR_INVALID, "ptb_init: device ordinal out of range");
    CU(cudaSetDevice(cfg.device));

    ptb_renderer* r = new ptb_renderer();
    r->device = cfg.device;
    r->cfg = cfg;
#define TRY(x) do { int rc_ = (x); if (rc_) { ptb_free(r); return rc_; } } while (0)
#define TRYCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { ptb_free(r); return fail(PTB_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } } while (0)
    TRYCU(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
    r->own_stream = true;
    cudaDeviceProp prop;
    TRYCU(cudaGetDeviceProperties(&prop, cfg.device));
    r->launch.num_sms = prop.multiProcessorCount;
    {   // tuning knobs of the RayCast scheduler (defaults are the measured best; env overrides are for A/B runs only)
        auto knob = [](const char* name, int def) { const char* e = std::getenv(name); return e ? std::atoi(e) : def; };
        r->launch.rc_blocks = knob("PTB_RC_BLOCKS", 10);
        r->launch.fuse_logic = knob("PTB_FUSE_LOGIC", 1);   // 0 never, 1 every bounce, 2 only where the one-batch-per-fetch kernel runs (bounce 0)
        r->launch.overlap_shadow = knob("PTB_OVERLAP_SHADOW", 1);
        r->launch.wide = knob("PTB_WIDE", 1);
        r->launch.sh_blocks = knob("PTB_SH_BLOCKS", 12);
        r->launch.sh_mode = knob("PTB_SH_MODE", 1);
        r->launch.sh_refill = knob("PTB_SH_REFILL", 24);
        r->launch.sh_steps = knob("PTB_SH_STEPS", 4);
        r->launch.sh_minb = knob("PTB_SH_MINB", 9);
        r->launch.rc_mode = knob("PTB_RC_MODE", 3);
        r->launch.rc_refill = knob("PTB_RC_REFILL", 24);
        r->launch.rc_stack = knob("PTB_RC_STACK", 0);
        r->launch.rc_minb = knob("PTB_RC_MINB", 8);
        r->launch.rc_steps = knob("PTB_RC_STEPS", 4);
        r->launch.rc_b0 = knob("PTB_RC_B0", 0);
        r->launch.mega_regen = knob("PTB_MEGA_REGEN", 1);
        r->launch.mega_refill = knob("PTB_MEGA_REFILL", 1);
        r->launch.mega_blocks = knob("PTB_MEGA_BLOCKS", 10);
        r->launch.shade_fork = knob("PTB_SHADE_FORK", 0);   // measured: no gain (12.26 vs 12.28 ms), the frame is a sum of issue-bound kernels
        if (r->launch.shade_fork) {
            for (int k = 0; k < 3; k++) {
                TRYCU(cudaStreamCreateWithFlags(&r->shade_side[k], cudaStreamNonBlocking));
                TRYCU(cudaEventCreateWithFlags(&r->ev_sjoin[k], cudaEventDisableTiming));
            }
            TRYCU(cudaEventCreateWithFlags(&r->ev_sfork, cudaEventDisableTiming));
        }
        if (r->launch.overlap_shadow) {
            TRYCU(cudaStreamCreateWithFlags(&r->side, cudaStreamNonBlocking));
            TRYCU(cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming));
            TRYCU(cudaEventCreateWithFlags(&r->ev_join, cudaEventDisableTiming));
        }
    }

    FrameDev& f = r->frame;
    // ---- mesh, materials, lights (initSceneGeometry, src/wavefront_internal.cu:118-167)
    float4 *pos, *nor, *tangent, *nor_geom; float2* uv; int4* im;
    TRY(dev_upload(r, pos, scene->pos.data(), V));
    TRY(dev_upload(r, nor, scene->nor.data(), V));
    TRY(dev_upload(r, tangent, scene->tangent.data(), V));
    TRY(dev_upload(r, uv, scene->uv.data(), V));
    TRY(dev_upload(r, im, scene->indices_matid.data(), N));
    TRY(dev_upload(r, nor_geom, scene->nor_geom.data(), N));
    f.mesh.pos = pos; f.mesh.nor = nor; f.mesh.tangent = tangent; f.mesh.uv = uv;
    f.mesh.indices_matid = im; f.mesh.nor_geom = nor_geom;
    f.mesh.num_vertices = V; f.mesh.num_triangles = N;
    ptb_material* mats;
    TRY(dev_upload(r, mats, scene->materials.data(), scene->materials.size()));
    f.materials = mats;
    int* ltri; float* lcdf;
    TRY(dev_upload(r, ltri, scene->light_tri_idx.data(), scene->light_tri_idx.size()));
    TRY(dev_upload(r, lcdf, scene->light_cdf.data(), scene->light_cdf.size()));
    f.lights.tri_idx = ltri; f.lights.cdf = lcdf;
    f.lights.num_lights = (int)scene->light_tri_idx.size();
    f.lights.total_area = scene->light_total_area;
    // ---- textures
    for (const HostTexture& t : scene->textures) TRY(upload_texture(r, t));
    cudaTextureObject_t* dtex;
    TRY(dev_upload(r, dtex, r->textures.data(), r->textures.size()));
    f.tex.handles = dtex;
    f.tex.env_tex_id = scene->env.env_tex_id;
    f.tex.pdf_map_id = scene->env.pdf_map_id;
    // the alias table the reference builds but never uploads (src/wavefront_internal.cu:206-214); see envmap.cuh
    f.env = EnvDev{};
    const bool have_env = scene->env.env_tex_id >= 0 && scene->env.pdf_map_id >= 0 && !scene->env.probs.empty();
    if (have_env) {
        float* probs; int* aliases;
        TRY(dev_upload(r, probs, scene->env.probs.data(), scene->env.probs.size()));
        TRY(dev_upload(r, aliases, scene->env.aliases.data(), scene->env.aliases.size()));
        f.env.probs = probs; f.env.aliases = aliases;
        f.env.width = scene->env.width; f.env.height = scene->env.height;
        f.env.nee = cfg.env_nee ? 1 : 0;
    }
    // with env_nee a shaded path may queue two shadow rays per bounce, and the environment's land in direct[P..2P)
    const size_t nee_factor = f.env.nee ? 2 : 1;

    // ---- LBVH (initBVH, src/wavefront_internal.cu:180-193 -> BuildLBVH, src/bvh.cu:456-516)
    {
        int e = lbvh_alloc(r->lbvh, N);
        if (e) { ptb_free(r); return fail(PTB_ERR_CUDA, std::string("lbvh_alloc: ") + cudaGetErrorString((cudaError_t)e)); }
        r->device_bytes += r->lbvh.bytes;
        cudaEvent_t e0, e1;
        TRYCU(cudaEventCreate(&e0));
        TRYCU(cudaEventCreate(&e1));
        int launches = 0;
        TRYCU(cudaEventRecord(e0, r->stream));
        e = lbvh_build(r->lbvh, pos, im, r->stream, &launches);
        TRYCU(cudaEventRecord(e1, r->stream));
        TRYCU(cudaStreamSynchronize(r->stream));
        if (e) { ptb_free(r); return fail(PTB_ERR_CUDA, std::string("lbvh_build: ") + cudaGetErrorString((cudaError_t)e)); }
        TRYCU(cudaEventElapsedTime(&r->lbvh_build_ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        r->kernel_launches += launches;
        TRYCU(cudaMemcpy(&r->lbvh_max_depth, r->lbvh.max_depth, sizeof(int), cudaMemcpyDeviceToHost));
        float4 root[2];
        TRYCU(cudaMemcpy(&root[0], r->lbvh.ref.aabb_min, sizeof(float4), cudaMemcpyDeviceToHost));
        TRYCU(cudaMemcpy(&root[1], r->lbvh.ref.aabb_max, sizeof(float4), cudaMemcpyDeviceToHost));
        f.bvh.node = r->lbvh.node;
        f.bvh.tri = r->lbvh.tri;
        f.bvh.wnode = r->lbvh.wnode;
        f.bvh.bnode = r->lbvh.bnode;
        // the quantisation step k_pack_wide used: (root_max - root_min) / PTB_WIDE_GRID, IEEE division on both sides
        f.bvh.wcell[0] = (root[1].x - root[0].x) / PTB_WIDE_GRID;
        f.bvh.wcell[1] = (root[1].y - root[0].y) / PTB_WIDE_GRID;
        f.bvh.wcell[2] = (root[1].z - root[0].z) / PTB_WIDE_GRID;
        f.bvh.root_min = root[0];
        f.bvh.root_max = root[1];
        f.bvh.num_triangles = N;
        if (r->lbvh_max_depth > 256) { ptb_free(r); return fail(PTB_ERR_UNSUPPORTED, "ptb_init: LBVH deeper than 256 levels"); }
        // the megakernel's visibility walk has 32- and 64-entry stacks only (megakernel.cu): refuse here, not at the first sync
        if (cfg.megakernel && r->lbvh_max_depth > 64) { ptb_free(r); return fail(PTB_ERR_UNSUPPORTED, "ptb_init: the megakernel path supports LBVHs up to 64 levels deep (this one: " + std::to_string(r->lbvh_max_depth) + "); use the wavefront path"); }
        // any-hit stack: at most one entry per internal node on the root-to-leaf path, i.e. <= lbvh_max_depth
        const int dmax = r->lbvh_max_depth;
        r->launch.shadow_deep = dmax <= 32 ? 0 : dmax <= 48 ? 1 : dmax <= 64 ? 2 : 3;
        // The reference's closest-hit walk gives up once 32 entries are stacked (src/kernels/ray_cast.cu:106), which can only
        // happen in a tree at least 32 levels deep.  The quantised walks visit in another order and would not give up at
        // the same ray, so extension rays of such trees always take the reference-order walk on the fp32 records.
        if (dmax >= PTB_EXT_STACK_DEPTH) { r->launch.wide &= ~2; if (r->launch.rc_mode >= 2) r->launch.rc_mode = 1; }
    }

    // ---- path pool, queues, accumulators (initWavefrontQueues / initImageSystem,
    //      src/wavefront_internal.cu:70-86,221-254) sized by the pixels this rank owns
    f.tile.width = W; f.tile.height = H;
    f.tile.rank = cfg.rank; f.tile.world = cfg.world; f.tile.stripe_rows = cfg.stripe_rows;
    f.tile.owned_pixels = owned_rows(cfg, H) * W;
    const size_t P = (size_t)f.tile.owned_pixels;
    f.num_paths = (int)P;
    f.trace_depth = cfg.trace_depth > 0 ? cfg.trace_depth : scene->trace_depth;
    if (f.trace_depth < 1) f.trace_depth = 1;
    TRY(dev_alloc(r, f.pool.rs, 4 * P));
    TRY(dev_alloc(r, f.pool.hit, 2 * P));
    f.pool.ray_ori.base = f.pool.rs;
    f.pool.ray_dir_dist.base = f.pool.rs + 1;
    f.pool.throughput_pdf.base = f.pool.rs + 2;
    f.pool.remaining_bounces.base = reinterpret_cast<int*>(f.pool.rs + 3) + 0;
    f.pool.rng_state.base = reinterpret_cast<uint32_t*>(f.pool.rs + 3) + 1;
    TRY(dev_alloc(r, f.shadow.ray_ori_tmax, nee_factor * P));
    TRY(dev_alloc(r, f.shadow.ray_dir, nee_factor * P));
    TRY(dev_alloc(r, f.shadow.radiance, nee_factor * P));
    TRY(dev_alloc(r, f.shadow.slot, nee_factor * P));
    TRY(dev_alloc(r, f.q_ext, P));
    for (int q = 0; q < 4; q++) TRY(dev_alloc(r, f.q_mat[q], P));
    TRY(dev_alloc(r, f.ctr, (size_t)(f.trace_depth + 1) * CTR_STRIDE));
    TRY(dev_alloc(r, f.ray_totals, 2));
    TRY(dev_alloc(r, f.iteration, 1));
    TRY(dev_alloc(r, f.error_flag, 1));
    TRY(dev_alloc(r, f.direct, nee_factor * 3 * P));
    TRY(dev_alloc(r, f.indirect, 3 * P));
    TRY(dev_alloc(r, r->cursors, kMaxTimed + 1));
    TRYCU(cudaMemsetAsync(f.ctr, 0, sizeof(int) * (f.trace_depth + 1) * CTR_STRIDE, r->stream));
    TRYCU(cudaMemsetAsync(f.ray_totals, 0, 2 * sizeof(unsigned long long), r->stream));
    TRYCU(cudaMemsetAsync(f.iteration, 0, sizeof(int), r->stream));
    TRYCU(cudaMemsetAsync(f.error_flag, 0, sizeof(int), r->stream));
    TRYCU(cudaMemsetAsync(f.direct, 0, nee_factor * 3 * P * sizeof(float), r->stream));
    TRYCU(cudaMemsetAsync(f.indirect, 0, 3 * P * sizeof(float), r->stream));
    // InitPathPoolKernel (src/pathtrace_wavefront.cu:183-190): every slot starts dead
    TRYCU(cudaMemsetAsync(f.pool.rs, 0, 4 * P * sizeof(float4), r->stream));
    TRYCU(cudaMemset2DAsync(f.pool.rs + 2, 64, 0xff, 32, P, r->stream));             // state halves: remaining_bounces = -1
    TRYCU(cudaMemsetAsync(f.pool.hit, 0xff, 2 * P * sizeof(float4), r->stream));     // hit_geom_id = material_id = -1

    // ---- first-bounce G-buffer (initSVGF buffers, src/wavefront_internal.cu:286-300), on request
    f.view_proj = nullptr;
    if (cfg.denoiser && cfg.world > 1) { ptb_free(r); return fail(PTB_ERR_UNSUPPORTED, "ptb_init: the denoiser filters across rows; it needs the whole frame on one GPU (world == 1)"); }
    if (cfg.gbuffer || cfg.denoiser) {
        r->gbuffer = true;
        TRY(dev_alloc(r, r->gb.depth, P));
        TRY(dev_alloc(r, r->gb.normal_matid, P));
        TRY(dev_alloc(r, r->gb.albedo, P));
        TRY(dev_alloc(r, r->gb.motion, P));
        TRYCU(cudaMemsetAsync(r->gb.depth, 0, P * sizeof(float), r->stream));
        TRYCU(cudaMemsetAsync(r->gb.normal_matid, 0, P * sizeof(float4), r->stream));
        TRYCU(cudaMemsetAsync(r->gb.albedo, 0, P * sizeof(float4), r->stream));
        TRYCU(cudaMemsetAsync(r->gb.motion, 0, P * sizeof(float2), r->stream));
        Mat4 vp[2];
        vp[0] = camera_view_proj(scene->camera);
        memset(&vp[1], 0, sizeof(Mat4));                       // prev_view_proj_mat starts as glm::mat4() = identity
        for (int i = 0; i < 4; i++) vp[1].m[i][i] = 1.0f;
        TRY(dev_alloc(r, f.view_proj, 2));
        TRYCU(cudaMemcpyAsync(f.view_proj, vp, sizeof(vp), cudaMemcpyHostToDevice, r->stream));
        TRYCU(cudaStreamSynchronize(r->stream));               // vp is a stack object
    }

    // ---- SVGF state (SVGFDenoiser::Init, src/kernels/svgf.cu:549-580) + the denoised image
    if (cfg.denoiser) {
        r->denoiser = true;
        auto zeroed = [&](auto*& p, size_t count) -> int {
            int rc = dev_alloc(r, p, count);
            if (rc) return rc;
            r->svgf_state.push_back({(void*)p, count * sizeof(*p)});
            return PTB_OK;
        };
        for (int i = 0; i < 2; i++) { TRY(zeroed(r->svgf.direct[i], P)); TRY(zeroed(r->svgf.indirect[i], P)); TRY(zeroed(r->svgf.moments[i], P)); }
        TRY(zeroed(r->svgf.history, P));
        TRY(zeroed(r->svgf.prev_normal_matid, P));
        TRY(zeroed(r->svgf.prev_depth, P));
        TRY(zeroed(r->svgf.final_image, 3 * P));
        TRY(zeroed(r->svgf.short_list, P));
        TRY(zeroed(r->svgf.short_count, 1));
        for (auto& b : r->svgf_state) TRYCU(cudaMemsetAsync(b.first, 0, b.second, r->stream));
    }

    // ---- camera
    const ptb_camera& c = scene->camera;
    f.cam.res_x = W; f.cam.res_y = H;
    for (int i = 0; i < 3; i++) { f.cam.pos[i] = c.position[i]; f.cam.view[i] = c.view[i]; f.cam.up[i] = c.up[i]; f.cam.right[i] = c.right[i]; }
    f.cam.pixel_length[0] = c.pixel_length[0];
    f.cam.pixel_length[1] = c.pixel_length[1];

    // ---- one CUDA graph per frame (the reference re-launches ~10 kernels per bounce with a
    //      cudaDeviceSynchronize after each, src/common/cuda_utilities.cu:4)
    if (cfg.use_graph) {
        cudaGraph_t graph = nullptr;
        TRYCU(cudaStreamSynchronize(r->stream));
        TRYCU(cudaStreamBeginCapture(r->stream, cudaStreamCaptureModeThreadLocal));
        r->launches_per_frame = enqueue_frame(r, r->stream);
        cudaError_t ce = cudaStreamEndCapture(r->stream, &graph);
        if (ce != cudaSuccess) { ptb_free(r); return fail(PTB_ERR_CUDA, std::string("graph capture: ") + cudaGetErrorString(ce)); }
        ce = cudaGraphInstantiate(&r->graph_exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) { r->graph_exec = nullptr; ptb_free(r); return fail(PTB_ERR_CUDA, std::string("graph instantiate: ") + cudaGetErrorString(ce)); }
    }
    TRYCU(cudaStreamSynchronize(r->stream));
    r->device_iteration = 0;
#undef TRY
#undef TRYCU
    *out = r;
    return PTB_OK;
}

}  // extern "C"

// ===================================================================== rendering ================
extern "C" {

int ptb_set_stream(ptb_renderer* r, void* cuda_stream) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_set_stream: null renderer");
    CU(cudaSetDevice(r->device));
    CU(cudaStreamSynchronize(r->stream));
    if (r->own_stream) cudaStreamDestroy(r->stream);
    r->stream = (cudaStream_t)cuda_stream;
    r->own_stream = false;
    return PTB_OK;
}

int ptb_sync(ptb_renderer* r) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_sync: null renderer");
    CU(cudaSetDevice(r->device));
    CU(cudaStreamSynchronize(r->stream));
    return check_device_error(r);
}

int ptb_render_many(ptb_renderer* r, int32_t first, int32_t count) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_render: null renderer");
    if (first < 1 || count < 0) return fail(PTB_ERR_INVALID, "ptb_render: iteration is 1-based");
    CU(cudaSetDevice(r->device));
    for (int32_t it = first; it < first + count; it++) {
        // iteration 1 restarts the accumulation (src/pathtrace_wavefront.cu:232-242)
        if (it == 1) {
            r->kernel_launches += launch_clear_accumulators(r->frame, r->stream);
            // WavefrontPathTracerState::clear() re-initialises the denoiser and the final image (src/wavefront_internal.cu:88-98)
            for (auto& b : r->svgf_state) CU(cudaMemsetAsync(b.first, 0, b.second, r->stream));
        }
        if (r->device_iteration != it) r->kernel_launches += launch_set_iteration(r->frame, it, r->stream);
        if (r->graph_exec) {
            CU(cudaGraphLaunch(r->graph_exec, r->stream));
            r->kernel_launches += r->launches_per_frame;
        } else {
            r->kernel_launches += enqueue_frame(r, r->stream);
        }
        r->device_iteration = it + 1;
        r->frames++;
        r->stage_traced = false;
        r->hit_valid = true;
    }
    CU(cudaGetLastError());
    return PTB_OK;
}

int ptb_render(ptb_renderer* r, int32_t iteration) { return ptb_render_many(r, iteration, 1); }

int ptb_clear_accumulators(ptb_renderer* r) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_clear_accumulators: null renderer");
    CU(cudaSetDevice(r->device));
    r->kernel_launches += launch_clear_accumulators(r->frame, r->stream);
    CU(cudaGetLastError());
    return PTB_OK;
}

int ptb_resolve(ptb_renderer* r, void* dst, int32_t format, int32_t iteration, int32_t dst_on_device) {
    if (!r || !dst) return fail(PTB_ERR_INVALID, "ptb_resolve: null argument");
    if (format < PTB_IMG_SUM_RGB32F || format > PTB_IMG_DENOISED_RGBA8 || format == PTB_IMG_STEPS_I32) return fail(PTB_ERR_INVALID, "ptb_resolve: unknown format");
    const bool den = format >= PTB_IMG_DENOISED_RGB32F;
    if (den && !r->denoiser) return fail(PTB_ERR_STATE, "ptb_resolve: denoised output needs ptb_config.denoiser = 1");
    const bool view = !den && format >= PTB_IMG_VIEW_NORMALS;
    if (view && !r->gbuffer) return fail(PTB_ERR_STATE, "ptb_resolve: G-buffer views need ptb_config.gbuffer = 1");
    if (!view && format != PTB_IMG_SUM_RGB32F && iteration < 1) return fail(PTB_ERR_INVALID, "ptb_resolve: iteration must be >= 1");
    CU(cudaSetDevice(r->device));
    const size_t P = (size_t)r->frame.num_paths;
    const size_t bytes = P * ((format == PTB_IMG_RGBA8 || format == PTB_IMG_DENOISED_RGBA8 || view) ? 4 : 12);
    auto run = [&](void* out) {
        if (den) return launch_resolve_final(r->svgf.final_image, (int)P, out, format == PTB_IMG_DENOISED_RGBA8 ? PTB_IMG_RGBA8 : PTB_IMG_SUM_RGB32F, r->stream);
        if (view) return launch_view(r->gb, (int)P, out, format - PTB_IMG_VIEW_NORMALS + 1, r->stream);
        return launch_resolve(r->frame, out, format, iteration, r->stream);
    };
    if (dst_on_device) {
        r->kernel_launches += run(dst);
        CU(cudaGetLastError());
        return PTB_OK;
    }
    int rc = ensure_scratch(r, bytes);
    if (rc) return rc;
    r->kernel_launches += run(r->scratch);
    CU(cudaMemcpyAsync(dst, r->scratch, bytes, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return check_device_error(r);
}

int ptb_get_accumulators(ptb_renderer* r, float* direct, float* indirect) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_get_accumulators: null renderer");
    CU(cudaSetDevice(r->device));
    CU(cudaStreamSynchronize(r->stream));
    const size_t bytes = 3 * (size_t)r->frame.num_paths * sizeof(float);
    if (direct) CU(cudaMemcpy(direct, r->frame.direct, bytes, cudaMemcpyDeviceToHost));
    if (indirect) CU(cudaMemcpy(indirect, r->frame.indirect, bytes, cudaMemcpyDeviceToHost));
    return check_device_error(r);
}

int ptb_get_gbuffer(ptb_renderer* r, int32_t which, void* dst, size_t bytes) {
    if (!r || !dst) return fail(PTB_ERR_INVALID, "ptb_get_gbuffer: null argument");
    if (!r->gbuffer) return fail(PTB_ERR_STATE, "ptb_get_gbuffer: renderer was created without ptb_config.gbuffer");
    const size_t P = (size_t)r->frame.num_paths;
    const void* src = nullptr; size_t elem = 0;
    switch (which) {
        case PTB_GB_DEPTH: src = r->gb.depth; elem = 4; break;
        case PTB_GB_NORMAL_MATID: src = r->gb.normal_matid; elem = 16; break;
        case PTB_GB_ALBEDO: src = r->gb.albedo; elem = 16; break;
        case PTB_GB_MOTION: src = r->gb.motion; elem = 8; break;
        default: return fail(PTB_ERR_INVALID, "ptb_get_gbuffer: unknown buffer");
    }
    if (bytes < P * elem) return fail(PTB_ERR_INVALID, "ptb_get_gbuffer: destination too small");
    CU(cudaSetDevice(r->device));
    CU(cudaStreamSynchronize(r->stream));
    CU(cudaMemcpy(dst, src, P * elem, cudaMemcpyDeviceToHost));
    return check_device_error(r);
}

int ptb_visualize_lbvh(ptb_renderer* r, void* dst, int32_t format, int32_t dst_on_device) {
    if (!r || !dst) return fail(PTB_ERR_INVALID, "ptb_visualize_lbvh: null argument");
    if (format != PTB_IMG_SUM_RGB32F && format != PTB_IMG_STEPS_I32) return fail(PTB_ERR_INVALID, "ptb_visualize_lbvh: format must be PTB_IMG_SUM_RGB32F or PTB_IMG_STEPS_I32");
    CU(cudaSetDevice(r->device));
    const size_t P = (size_t)r->frame.num_paths;
    const size_t bytes = P * (format == PTB_IMG_SUM_RGB32F ? 12 : 4);
    void* out = dst;
    if (!dst_on_device) {
        int rc = ensure_scratch(r, bytes);
        if (rc) return rc;
        out = r->scratch;
    }
    r->kernel_launches += launch_bvh_heatmap(r->frame, format == PTB_IMG_SUM_RGB32F ? (float*)out : nullptr,
                                             format == PTB_IMG_STEPS_I32 ? (int*)out : nullptr, r->stream);
    CU(cudaGetLastError());
    if (dst_on_device) return PTB_OK;
    CU(cudaMemcpyAsync(dst, r->scratch, bytes, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return check_device_error(r);
}

int ptb_sample_environment(ptb_renderer* r, const float* rnd, int32_t count, float* dir, float* pdf, int32_t* pixel) {
    if (!r || !rnd || !dir || !pdf || !pixel) return fail(PTB_ERR_INVALID, "ptb_sample_environment: null argument");
    if (count <= 0) return fail(PTB_ERR_INVALID, "ptb_sample_environment: count must be positive");
    if (!r->frame.env.probs) return fail(PTB_ERR_STATE, "ptb_sample_environment: the scene has no environment map");
    CU(cudaSetDevice(r->device));
    const size_t n = (size_t)count;
    int rc = ensure_scratch(r, n * (2 + 3 + 1 + 1) * 4);
    if (rc) return rc;
    float* d_rnd = (float*)r->scratch;
    float* d_dir = d_rnd + 2 * n;
    float* d_pdf = d_dir + 3 * n;
    int* d_pix = (int*)(d_pdf + n);
    CU(cudaMemcpyAsync(d_rnd, rnd, n * 8, cudaMemcpyHostToDevice, r->stream));
    r->kernel_launches += launch_sample_environment(r->frame, d_rnd, count, d_dir, d_pdf, d_pix, r->stream);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(dir, d_dir, n * 12, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaMemcpyAsync(pdf, d_pdf, n * 4, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaMemcpyAsync(pixel, d_pix, n * 4, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return PTB_OK;
}

int ptb_get_tile_info(const ptb_renderer* r, ptb_tile_info* out) {
    if (!r || !out) return fail(PTB_ERR_INVALID, "ptb_get_tile_info: null argument");
    out->width = r->frame.tile.width;
    out->height = r->frame.tile.height;
    out->owned_pixels = r->frame.tile.owned_pixels;
    out->rank = r->frame.tile.rank;
    out->world = r->frame.tile.world;
    out->stripe_rows = r->frame.tile.stripe_rows;
    return PTB_OK;
}

int ptb_get_owned_pixels(const ptb_renderer* r, int32_t* dst, int32_t capacity) {
    if (!r || !dst) return fail(PTB_ERR_INVALID, "ptb_get_owned_pixels: null argument");
    int n = r->frame.tile.owned_pixels;
    if (capacity < n) n = capacity;
    for (int i = 0; i < n; i++) dst[i] = r->frame.tile.global_pixel(i);
    return r->frame.tile.owned_pixels;
}

// ===================================================================== LBVH view ================
int64_t ptb_get_lbvh(ptb_renderer* r, int which, void* dst, size_t bytes) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_get_lbvh: null renderer");
    const size_t N = (size_t)r->lbvh.num_triangles;
    const void* src = nullptr;
    size_t size = 0;
    switch (which) {
        case PTB_LBVH_AABB_MIN: src = r->lbvh.ref.aabb_min; size = 2 * N * sizeof(float4); break;
        case PTB_LBVH_AABB_MAX: src = r->lbvh.ref.aabb_max; size = 2 * N * sizeof(float4); break;
        case PTB_LBVH_CENTROID: src = r->lbvh.ref.centroid; size = 2 * N * sizeof(float4); break;
        case PTB_LBVH_PRIMITIVE_INDICES: src = r->lbvh.ref.primitive_indices; size = 2 * N * sizeof(int); break;
        case PTB_LBVH_MORTON_CODES: src = r->lbvh.ref.morton_codes; size = N * sizeof(unsigned long long); break;
        case PTB_LBVH_CHILD_NODES: src = r->lbvh.ref.child_nodes; size = (N - 1) * sizeof(int2); break;
        case PTB_LBVH_PARENT: src = r->lbvh.ref.parent; size = 2 * N * sizeof(int); break;
        case PTB_LBVH_ESCAPE_INDICES: src = r->lbvh.ref.escape_indices; size = 2 * N * sizeof(int); break;
        case PTB_LBVH_WORLD_AABB: src = r->lbvh.ref.world_aabb; size = 2 * sizeof(float4); break;
        default: return fail(PTB_ERR_INVALID, "ptb_get_lbvh: unknown array id");
    }
    if (dst && bytes) {
        CU(cudaSetDevice(r->device));
        CU(cudaStreamSynchronize(r->stream));
        CU(cudaMemcpy(dst, src, bytes < size ? bytes : size, cudaMemcpyDeviceToHost));
    }
    return (int64_t)size;
}

int ptb_build_lbvh(int32_t device, const float* pos4, int32_t V, const int32_t* indices_matid4, int32_t N,
                   uint64_t* morton_codes, int32_t* primitive_indices, int32_t* child_nodes, int32_t* parent,
                   int32_t* escape_indices, float* aabb_min4, float* aabb_max4, float* world_aabb8, float* build_ms) {
    if (!pos4 || !indices_matid4) return fail(PTB_ERR_INVALID, "ptb_build_lbvh: null input");
    if (N < 2) return fail(PTB_ERR_UNSUPPORTED, "ptb_build_lbvh: needs at least 2 triangles");
    if (V < 1) return fail(PTB_ERR_INVALID, "ptb_build_lbvh: no vertices");
    for (int64_t i = 0; i < (int64_t)N; i++)
        for (int k = 0; k < 3; k++)
            if (indices_matid4[4 * i + k] < 0 || indices_matid4[4 * i + k] >= V) return fail(PTB_ERR_INVALID, "ptb_build_lbvh: vertex index out of range");
    if (ptb_device_count() == 0) return fail(PTB_ERR_CUDA, "ptb_build_lbvh: no CUDA device (no CPU fallback)");
    CU(cudaSetDevice(device));
    float4* dpos = nullptr; int4* dim = nullptr;
    LbvhBuffers b;
    cudaStream_t s = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = PTB_OK;
    auto cleanup = [&]() {
        if (dpos) cudaFree(dpos);
        if (dim) cudaFree(dim);
        lbvh_free(b);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        if (s) cudaStreamDestroy(s);
    };
#define G(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { rc = fail(PTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); cleanup(); return rc; } } while (0)
    G(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    G(cudaEventCreate(&e0));
    G(cudaEventCreate(&e1));
    G(cudaMalloc((void**)&dpos, (size_t)V * sizeof(float4)));
    G(cudaMalloc((void**)&dim, (size_t)N * sizeof(int4)));
    G((cudaError_t)lbvh_alloc(b, N));
    G(cudaMemcpyAsync(dpos, pos4, (size_t)V * sizeof(float4), cudaMemcpyHostToDevice, s));
    G(cudaMemcpyAsync(dim, indices_matid4, (size_t)N * sizeof(int4), cudaMemcpyHostToDevice, s));
    G(cudaEventRecord(e0, s));
    G((cudaError_t)lbvh_build(b, dpos, dim, s, nullptr));
    G(cudaEventRecord(e1, s));
    G(cudaStreamSynchronize(s));
    if (build_ms) G(cudaEventElapsedTime(build_ms, e0, e1));
    const size_t n = (size_t)N;
    if (morton_codes) G(cudaMemcpy(morton_codes, b.ref.morton_codes, n * 8, cudaMemcpyDeviceToHost));
    if (primitive_indices) G(cudaMemcpy(primitive_indices, b.ref.primitive_indices, 2 * n * 4, cudaMemcpyDeviceToHost));
    if (child_nodes) G(cudaMemcpy(child_nodes, b.ref.child_nodes, (n - 1) * 8, cudaMemcpyDeviceToHost));
    if (parent) G(cudaMemcpy(parent, b.ref.parent, 2 * n * 4, cudaMemcpyDeviceToHost));
    if (escape_indices) G(cudaMemcpy(escape_indices, b.ref.escape_indices, 2 * n * 4, cudaMemcpyDeviceToHost));
    if (aabb_min4) G(cudaMemcpy(aabb_min4, b.ref.aabb_min, 2 * n * 16, cudaMemcpyDeviceToHost));
    if (aabb_max4) G(cudaMemcpy(aabb_max4, b.ref.aabb_max, 2 * n * 16, cudaMemcpyDeviceToHost));
    if (world_aabb8) G(cudaMemcpy(world_aabb8, b.ref.world_aabb, 32, cudaMemcpyDeviceToHost));
#undef G
    cleanup();
    return PTB_OK;
}

// ===================================================================== batched RayCast ==========
namespace {
struct BatchTimer {            // device time of one batched RayCast kernel -> ptb_stats.last_batch_ms
    ptb_renderer* r; cudaEvent_t e0 = nullptr, e1 = nullptr;
    explicit BatchTimer(ptb_renderer* r_) : r(r_) {
        if (cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess) cudaEventRecord(e0, r->stream);
    }
    void stop() { if (e1) cudaEventRecord(e1, r->stream); }
    ~BatchTimer() {
        if (e0 && e1 && cudaEventSynchronize(e1) == cudaSuccess) cudaEventElapsedTime(&r->last_batch_ms, e0, e1);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    }
};
}  // namespace

int ptb_trace_closest(ptb_renderer* r, const float* ori4, const float* dir4, int32_t n, int32_t* hit, float* t, float* u, float* v) {
    if (!r || !ori4 || !dir4 || !hit || !t || !u || !v) return fail(PTB_ERR_INVALID, "ptb_trace_closest: null argument");
    if (n < 0) return fail(PTB_ERR_INVALID, "ptb_trace_closest: negative count");
    if (n == 0) return PTB_OK;
    CU(cudaSetDevice(r->device));
    const size_t N = (size_t)n;
    int rc = ensure_scratch(r, N * (16 + 16 + 4 * 4));
    if (rc) return rc;
    char* base = (char*)r->scratch;
    float4* dori = (float4*)base; float4* ddir = (float4*)(base + 16 * N);
    int* dhit = (int*)(base + 32 * N); float* dt = (float*)(base + 36 * N); float* du = (float*)(base + 40 * N); float* dv = (float*)(base + 44 * N);
    CU(cudaMemcpyAsync(dori, ori4, 16 * N, cudaMemcpyHostToDevice, r->stream));
    CU(cudaMemcpyAsync(ddir, dir4, 16 * N, cudaMemcpyHostToDevice, r->stream));
    BatchTimer timer(r);
    r->kernel_launches += launch_trace_batch_closest(r->frame.bvh, dori, ddir, n, dhit, dt, du, dv, r->cursors + kMaxTimed, r->launch, r->stream);
    timer.stop();
    CU(cudaMemcpyAsync(hit, dhit, 4 * N, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaMemcpyAsync(t, dt, 4 * N, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaMemcpyAsync(u, du, 4 * N, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaMemcpyAsync(v, dv, 4 * N, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return PTB_OK;
}

int ptb_trace_occluded(ptb_renderer* r, const float* ori_tmax4, const float* dir4, int32_t n, int32_t* occluded) {
    if (!r || !ori_tmax4 || !dir4 || !occluded) return fail(PTB_ERR_INVALID, "ptb_trace_occluded: null argument");
    if (n < 0) return fail(PTB_ERR_INVALID, "ptb_trace_occluded: negative count");
    if (n == 0) return PTB_OK;
    CU(cudaSetDevice(r->device));
    const size_t N = (size_t)n;
    int rc = ensure_scratch(r, N * (16 + 16 + 4));
    if (rc) return rc;
    char* base = (char*)r->scratch;
    float4* dori = (float4*)base; float4* ddir = (float4*)(base + 16 * N); int* docc = (int*)(base + 32 * N);
    CU(cudaMemcpyAsync(dori, ori_tmax4, 16 * N, cudaMemcpyHostToDevice, r->stream));
    CU(cudaMemcpyAsync(ddir, dir4, 16 * N, cudaMemcpyHostToDevice, r->stream));
    BatchTimer timer(r);
    r->kernel_launches += launch_trace_batch_occluded(r->frame.bvh, dori, ddir, n, docc, r->cursors + kMaxTimed, r->frame.error_flag, r->launch, r->stream);
    timer.stop();
    CU(cudaMemcpyAsync(occluded, docc, 4 * N, cudaMemcpyDeviceToHost, r->stream));
    CU(cudaStreamSynchronize(r->stream));
    return check_device_error(r);
}

}  // extern "C"

// ===================================================================== staged frame =============
extern "C" {

int ptb_stage_run(ptb_renderer* r, int32_t stage, int32_t iteration) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_stage_run: null renderer");
    if (r->cfg.megakernel) return fail(PTB_ERR_STATE, "ptb_stage_run: the megakernel path has no stages");
    CU(cudaSetDevice(r->device));
    const FrameDev& f = r->frame;
    switch (stage) {
        case PTB_STAGE_RAYGEN:
            if (iteration < 1) return fail(PTB_ERR_INVALID, "ptb_stage_run: iteration is 1-based");
            if (iteration == 1) r->kernel_launches += launch_clear_accumulators(f, r->stream);
            r->kernel_launches += launch_set_iteration(f, iteration, r->stream);
            r->kernel_launches += launch_raygen(f, r->launch, r->stream);
            r->device_iteration = iteration;   // frame_end is not run in staged mode
            r->stage_bounce = 0;
            r->logic_bounce = 0;
            r->stage_traced = false;
            r->hit_valid = false;
            break;
        case PTB_STAGE_TRACE:
            if (r->stage_bounce >= f.trace_depth) return fail(PTB_ERR_STATE, "ptb_stage_run: all bounces of this frame are done");
            r->kernel_launches += launch_trace(f, r->stage_bounce, r->launch, r->stream);
            // the environment's half of `direct` is folded in ONCE per frame, after the last trace, exactly where the frame graph does
            // it (enqueue_frame): folding after every TRACE would round (d + e1) + e2 instead of d + (e1 + e2)
            if (f.env.nee && r->stage_bounce == f.trace_depth - 1) r->kernel_launches += launch_fold_env(f, r->stream);
            r->stage_traced = true;
            r->hit_valid = true;
            break;
        case PTB_STAGE_LOGIC:
            if (r->stage_bounce >= f.trace_depth) return fail(PTB_ERR_STATE, "ptb_stage_run: all bounces of this frame are done");
            r->kernel_launches += launch_logic(f, r->stage_bounce, r->launch, r->stream);
            r->logic_bounce = r->stage_bounce;
            break;
        case PTB_STAGE_SHADE:
            if (r->stage_bounce >= f.trace_depth) return fail(PTB_ERR_STATE, "ptb_stage_run: all bounces of this frame are done");
            r->kernel_launches += launch_shade(f, r->stage_bounce, r->launch, r->stream);
            r->stage_bounce++;
            r->stage_traced = false;
            break;
        default:
            return fail(PTB_ERR_INVALID, "ptb_stage_run: unknown stage");
    }
    CU(cudaGetLastError());
    return PTB_OK;
}

static int queue_counter(ptb_renderer* r, int which, int* count) {
    int row, col;
    switch (which) {
        case PTB_Q_EXTENSION: row = r->stage_bounce; col = CTR_EXT; break;
        case PTB_Q_SHADOW: row = r->stage_bounce; col = CTR_SHADOW; break;
        case PTB_Q_PBR: case PTB_Q_DIFFUSE: case PTB_Q_REFLECTION: case PTB_Q_REFRACTION:
            row = r->logic_bounce; col = CTR_PBR + (which - PTB_Q_PBR); break;
        default: return fail(PTB_ERR_INVALID, "unknown queue id");
    }
    CU(cudaStreamSynchronize(r->stream));
    CU(cudaMemcpy(count, r->frame.ctr + row * CTR_STRIDE + col, sizeof(int), cudaMemcpyDeviceToHost));
    return PTB_OK;
}

int ptb_get_queue_count(ptb_renderer* r, int32_t which, int32_t* count) {
    if (!r || !count) return fail(PTB_ERR_INVALID, "ptb_get_queue_count: null argument");
    CU(cudaSetDevice(r->device));
    return queue_counter(r, which, count);
}

int ptb_get_queue(ptb_renderer* r, int32_t which, int32_t* dst, int32_t capacity) {
    if (!r || !dst) return fail(PTB_ERR_INVALID, "ptb_get_queue: null argument");
    CU(cudaSetDevice(r->device));
    int n = 0;
    int rc = queue_counter(r, which, &n);
    if (rc) return rc;
    int m = n < capacity ? n : capacity;
    const int* src;
    if (which == PTB_Q_EXTENSION) src = r->frame.q_ext;
    else if (which == PTB_Q_SHADOW) src = r->frame.shadow.slot;
    else src = r->frame.q_mat[which - PTB_Q_PBR];
    if (m > 0) CU(cudaMemcpy(dst, src, (size_t)m * sizeof(int), cudaMemcpyDeviceToHost));
    if (r->frame.tile.world > 1)
        for (int i = 0; i < m; i++) dst[i] = r->frame.tile.global_pixel(dst[i]);
    return n;
}

int ptb_get_shadow_queue(ptb_renderer* r, float* ori_tmax4, float* dir4, float* radiance4, int32_t* pixel_idx, int32_t capacity) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_get_shadow_queue: null renderer");
    CU(cudaSetDevice(r->device));
    int n = 0;
    int rc = queue_counter(r, PTB_Q_SHADOW, &n);
    if (rc) return rc;
    size_t m = (size_t)(n < capacity ? n : capacity);
    if (m > 0) {
        if (ori_tmax4) CU(cudaMemcpy(ori_tmax4, r->frame.shadow.ray_ori_tmax, m * 16, cudaMemcpyDeviceToHost));
        if (dir4) CU(cudaMemcpy(dir4, r->frame.shadow.ray_dir, m * 16, cudaMemcpyDeviceToHost));
        if (radiance4) CU(cudaMemcpy(radiance4, r->frame.shadow.radiance, m * 16, cudaMemcpyDeviceToHost));
        if (pixel_idx) {
            CU(cudaMemcpy(pixel_idx, r->frame.shadow.slot, m * 4, cudaMemcpyDeviceToHost));
            if (r->frame.tile.world > 1)
                for (size_t i = 0; i < m; i++) pixel_idx[i] = r->frame.tile.global_pixel(pixel_idx[i]);
        }
    }
    return n;
}

int ptb_get_path_field(ptb_renderer* r, int32_t field, void* dst, size_t bytes) {
    if (!r || !dst) return fail(PTB_ERR_INVALID, "ptb_get_path_field: null argument");
    CU(cudaSetDevice(r->device));
    // The reference's PathState fields as VIEWS of the records of a slot (common.cuh: a 64-byte ray + state block, a 32-byte hit
    // record): src / pitch select a word range of a record; ray_dir_dist.w and hit_normal are assembled below.
    const char* src;
    size_t elem, pitch = 64;
    const PathPool& pool = r->frame.pool;
    switch (field) {
        case PTB_PF_RAY_ORI: src = (const char*)pool.rs; elem = 16; break;
        case PTB_PF_RAY_DIR_DIST: src = (const char*)(pool.rs + 1); elem = 16; break;
        case PTB_PF_HIT_UV: src = (const char*)pool.hit; elem = 16; pitch = 32; break;     // (t, u, v, prim) -> (u, v, 0, 0) below
        case PTB_PF_THROUGHPUT_PDF: src = (const char*)(pool.rs + 2); elem = 16; break;
        case PTB_PF_HIT_GEOM_ID: src = (const char*)pool.hit + 12; elem = 4; pitch = 32; break;
        case PTB_PF_MATERIAL_ID: src = (const char*)pool.hit + 16; elem = 4; pitch = 32; break;
        case PTB_PF_REMAINING_BOUNCES: src = (const char*)(pool.rs + 3); elem = 4; break;
        case PTB_PF_RNG_STATE: src = (const char*)(pool.rs + 3) + 4; elem = 4; break;
        default: return fail(PTB_ERR_INVALID, "ptb_get_path_field: unknown field");
    }
    const TileMap& tm = r->frame.tile;
    const size_t full = (size_t)tm.width * tm.height * elem;
    if (bytes < full) return fail(PTB_ERR_INVALID, "ptb_get_path_field: buffer smaller than width*height elements");
    CU(cudaStreamSynchronize(r->stream));
    const size_t n = (size_t)tm.owned_pixels;
    std::vector<char> tmp(n * elem);
    CU(cudaMemcpy2D(tmp.data(), elem, src, pitch, elem, n, cudaMemcpyDeviceToHost));
    if (field == PTB_PF_HIT_UV) {
        float* p = reinterpret_cast<float*>(tmp.data());
        for (size_t i = 0; i < n; i++) { p[4 * i] = p[4 * i + 1]; p[4 * i + 1] = p[4 * i + 2]; p[4 * i + 2] = 0.0f; p[4 * i + 3] = 0.0f; }
    } else if ((field == PTB_PF_HIT_GEOM_ID || field == PTB_PF_MATERIAL_ID) && !r->hit_valid) {
        // RayGen resets both to -1 in the reference (src/kernels/ray_gen.cu:68-69); here the hit record is undefined until the
        // frame's first trace
        int32_t* p = reinterpret_cast<int32_t*>(tmp.data());
        for (size_t i = 0; i < n; i++) p[i] = -1;
    } else if (field == PTB_PF_RAY_DIR_DIST && r->stage_traced) {
        // between TRACE and SHADE of a staged frame the reference holds the hit distance (or -1) in ray_dir_dist.w of every path
        // of the bounce's extension queue (src/kernels/ray_cast.cu:111-121)
        int count = 0;
        CU(cudaMemcpy(&count, r->frame.ctr + r->stage_bounce * CTR_STRIDE + CTR_EXT, sizeof(int), cudaMemcpyDeviceToHost));
        std::vector<int32_t> q((size_t)(count > 0 ? count : 0));
        std::vector<float> t(n);
        if (count > 0) CU(cudaMemcpy(q.data(), r->frame.q_ext, (size_t)count * sizeof(int32_t), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy2D(t.data(), 4, pool.hit, 32, 4, n, cudaMemcpyDeviceToHost));
        float* p = reinterpret_cast<float*>(tmp.data());
        for (int32_t slot : q) p[4 * (size_t)slot + 3] = t[slot];
    }
    if (tm.world <= 1) {
        memcpy(dst, tmp.data(), n * elem);
        return PTB_OK;
    }
    for (int i = 0; i < tm.owned_pixels; i++)
        memcpy((char*)dst + (size_t)tm.global_pixel(i) * elem, tmp.data() + (size_t)i * elem, elem);
    return PTB_OK;
}

// ===================================================================== stats ====================
int ptb_get_stats(ptb_renderer* r, ptb_stats* out) {
    if (!r || !out) return fail(PTB_ERR_INVALID, "ptb_get_stats: null argument");
    CU(cudaSetDevice(r->device));
    CU(cudaStreamSynchronize(r->stream));
    unsigned long long rays[2] = {0, 0};
    CU(cudaMemcpy(rays, r->frame.ray_totals, sizeof(rays), cudaMemcpyDeviceToHost));
    memset(out, 0, sizeof(*out));
    out->extension_rays = rays[0];
    out->shadow_rays = rays[1];
    out->frames = r->frames;
    out->kernel_launches = r->kernel_launches;
    out->lbvh_build_ms = r->lbvh_build_ms;
    out->lbvh_max_depth = r->lbvh_max_depth;
    out->num_triangles = r->lbvh.num_triangles;
    out->last_batch_ms = r->last_batch_ms;
    return check_device_error(r);
}

int ptb_stats_reset(ptb_renderer* r) {
    if (!r) return fail(PTB_ERR_INVALID, "ptb_stats_reset: null renderer");
    CU(cudaSetDevice(r->device));
    CU(cudaMemsetAsync(r->frame.ray_totals, 0, 2 * sizeof(unsigned long long), r->stream));
    r->frames = 0;
    r->kernel_launches = 0;
    return PTB_OK;
}

int ptb_time_kernel(ptb_renderer* r, int32_t which, int32_t count, int32_t cold, float* avg_ms, int32_t* rays) {
    if (!r || !avg_ms) return fail(PTB_ERR_INVALID, "ptb_time_kernel: null argument");
    if (which != PTB_TIME_EXTENSION && which != PTB_TIME_SHADOW) return fail(PTB_ERR_INVALID, "ptb_time_kernel: unknown kernel");
    if (count < 1 || count > kMaxTimed) return fail(PTB_ERR_INVALID, "ptb_time_kernel: count must be in [1, 4096]");
    if (r->cfg.megakernel) return fail(PTB_ERR_STATE, "ptb_time_kernel: the megakernel path has no RayCast stage");
    if (r->stage_bounce >= r->frame.trace_depth) return fail(PTB_ERR_STATE, "ptb_time_kernel: no live queue");
    CU(cudaSetDevice(r->device));
    const FrameDev& f = r->frame;
    const int b = r->stage_bounce;
    int n = 0;
    CU(cudaStreamSynchronize(r->stream));
    CU(cudaMemcpy(&n, f.ctr + b * CTR_STRIDE + (which == PTB_TIME_SHADOW ? CTR_SHADOW : CTR_EXT), sizeof(int), cudaMemcpyDeviceToHost));
    if (rays) *rays = n;
    *avg_ms = 0.0f;
    if (which == PTB_TIME_SHADOW && (b == 0 || n == 0)) return PTB_OK;      // no shadow rays exist before the first Shade
    const size_t kFlushBytes = (size_t)256 << 20;                             // > the 126 MB L2
    if (cold && !r->flush_buf) CU(cudaMalloc(&r->flush_buf, kFlushBytes));
    // the shadow kernel adds into the direct accumulator: keep a copy and put it back afterwards
    const size_t direct_bytes = (f.env.nee ? 2 : 1) * 3 * (size_t)f.num_paths * sizeof(float);
    if (which == PTB_TIME_SHADOW) {
        int rc = ensure_scratch(r, direct_bytes);
        if (rc) return rc;
        CU(cudaMemcpyAsync(r->scratch, f.direct, direct_bytes, cudaMemcpyDeviceToDevice, r->stream));
    }
    std::vector<cudaEvent_t> ev(2 * (size_t)count);
    for (auto& e : ev) CU(cudaEventCreate(&e));
    CU(cudaMemsetAsync(r->cursors, 0, sizeof(int) * count, r->stream));
    if (!cold) CU(cudaEventRecord(ev[0], r->stream));
    for (int i = 0; i < count; i++) {
        if (which == PTB_TIME_SHADOW) CU(cudaMemsetAsync(f.ctr + b * CTR_STRIDE + CTR_WORK_SHADOW, 0, sizeof(int), r->stream));
        if (cold) {
            CU(cudaMemsetAsync(r->flush_buf, i & 0xff, kFlushBytes, r->stream));
            CU(cudaEventRecord(ev[2 * i], r->stream));
        }
        if (which == PTB_TIME_SHADOW) r->kernel_launches += launch_trace_shadow(f, b, r->launch, r->stream);
        else r->kernel_launches += launch_trace_extension_only(f, b, r->cursors + i, r->launch, r->stream);
        if (cold) CU(cudaEventRecord(ev[2 * i + 1], r->stream));
    }
    if (!cold) CU(cudaEventRecord(ev[1], r->stream));
    if (which == PTB_TIME_SHADOW) {
        CU(cudaMemcpyAsync(f.direct, r->scratch, direct_bytes, cudaMemcpyDeviceToDevice, r->stream));
        CU(cudaMemsetAsync(f.ctr + b * CTR_STRIDE + CTR_WORK_SHADOW, 0, sizeof(int), r->stream));
    }
    CU(cudaStreamSynchronize(r->stream));
    float total = 0.0f;
    if (cold) {
        for (int i = 0; i < count; i++) { float ms = 0.0f; CU(cudaEventElapsedTime(&ms, ev[2 * i], ev[2 * i + 1])); total += ms; }
    } else {
        CU(cudaEventElapsedTime(&total, ev[0], ev[1]));
    }
    for (auto& e : ev) cudaEventDestroy(e);
    *avg_ms = total / (float)count;
    CU(cudaGetLastError());
    return PTB_OK;
}

int ptb_time_raycast(ptb_renderer* r, int32_t count, float* avg_ms, int32_t* rays) {
    return ptb_time_kernel(r, PTB_TIME_EXTENSION, count, 0, avg_ms, rays);
}

}  // extern "C"
