// render.cuh -- per-frame device state of the wavefront core and the stage launchers.
// Replaces WavefrontPathTracerState's queue / counter members (src/wavefront_internal.h:10-92)
// and the stage wrappers of src/kernels/{ray_gen,ray_cast,logic,shading}.h.
#pragma once
#include "common.cuh"
#include "envmap.cuh"

namespace ptb {

// Counter rows: row b holds the queue lengths of bounce b.  Nothing is reset inside a frame:
// Shade(b) appends to row b+1, so the host never reads a counter back
// (the reference does 6 blocking D2H reads + 6 memsets per bounce, src/pathtrace_wavefront.cu:286-343).
enum { CTR_WORK_EXT = 6, CTR_WORK_SHADOW = 7 };   // work-fetch cursors of the persistent RayCast kernels

struct Mat4 { float m[4][4]; };          // glm::mat4: m[column][row]

struct GBufferDev {                      // first-bounce G-buffer, one entry per owned pixel (src/wavefront_internal.cu:286-300)
    float* depth;                        // hit distance, -1000 for the sky
    float4* normal_matid;                // geometric normal, (float)material id (-1 sky)
    float4* albedo;                      // base colour x texture^2.2, w = 1 (sky: 1,1,1,0)
    float2* motion;                      // screen-space motion in pixels: uv(current view-proj) - uv(previous)
};

struct SvgfDev {                         // SVGFDenoiser's owned buffers (src/kernels/svgf.h:55-72), 120 B per pixel + the output
    float4* direct[2];                   // demodulated direct light (rgb) + variance (w): [0] work, [1] history
    float4* indirect[2];
    float4* moments[2];                  // (lum_dir, lum_ind, lum_dir^2, lum_ind^2)
    int* history;                        // frames of valid history per pixel
    float4* prev_normal_matid;           // last frame's G-buffer for the consistency test
    float* prev_depth;
    float* final_image;                  // float3: (direct + indirect) * albedo, no gamma (d_final_image)
    int* short_list;                     // pixels with fewer than 4 frames of history (they get the spatial variance estimate)
    int* short_count;
};

struct FrameDev {
    PathPool pool;
    ShadowQueueDev shadow;
    int* q_ext;
    int* q_mat[4];            // PBR, DIFFUSE, REFLECT, REFRACT
    int* ctr;                 // [(trace_depth + 1) * CTR_STRIDE]
    unsigned long long* ray_totals;   // [2] extension / shadow rays since the last stats reset
    int* iteration;           // 1-based sample index of the frame being rendered; advanced by frame_end
    int* error_flag;          // set by a kernel on traversal-stack overflow
    Mat4* view_proj;          // [2] current / previous view-projection (G-buffer motion vectors); null when the G-buffer is off
    float* direct;            // 3 floats per owned pixel (d_direct_image, src/wavefront_internal.h:26)
    float* indirect;          // (d_indirect_image, :27)
    MeshDev mesh;
    LightDev lights;
    BvhDev bvh;
    const ptb_material* materials;
    TexturesDev tex;
    CameraDev cam;
    TileMap tile;
    int trace_depth;
    int num_paths;            // == tile.owned_pixels
    EnvDev env;               // alias table of the HDR environment + ptb_config.env_nee
};

struct LaunchCfg {
    int num_sms;
    int shadow_deep;          // shadow-stack class from the LBVH depth: 0: <= 32, 1: <= 48, 2: <= 64 (shared memory), 3: <= 256 (local)
    int rc_blocks;            // resident blocks per SM the persistent grids are sized for
    int fuse_logic;           // frame graph: Logic runs in the tail of the extension RayCast kernel
    int overlap_shadow;       // frame graph: TraceShadow(b) runs beside TraceExtension(b) on a forked branch
    int wide;                 // 4-wide quantised traversal (wide.cuh): bit 0 shadow rays, bit 1 extension rays, bit 2 megakernel
    int sh_blocks;            // resident blocks per SM of the wide shadow kernel
    int sh_mode, sh_refill, sh_steps, sh_minb;   // shadow rays: 1 = persistent lanes (k_trace_shadow_pl), refill threshold, steps per ballot, blocks per SM
    int shade_fork;           // frame graph: the four material kernels of a bounce run as parallel branches
    int rc_mode;              // extension-ray scheduler / node format (raycast.cu run_extension)
    int rc_stack, rc_minb, rc_steps, rc_b0;    // see raycast.cu run_extension
    int rc_refill;            // persistent lanes: idle lanes per warp that trigger a refill (32 = one batch per fetch)
    int mega_regen;           // megakernel: a lane whose path ended takes the next pixel at once (path regeneration)
    int mega_refill;          // megakernel: regenerate once this many lanes of a warp are idle
    int mega_blocks;          // megakernel: resident blocks per SM the persistent grid is sized for
};

// Each returns the number of kernels it enqueued (for ptb_stats.kernel_launches).
int launch_clear_accumulators(const FrameDev& f, cudaStream_t s);
int launch_raygen(const FrameDev& f, const LaunchCfg& c, cudaStream_t s);   // iteration is read from *f.iteration
int launch_set_iteration(const FrameDev& f, int iteration, cudaStream_t s);
// extension + shadow rays; fuse_logic: the extension kernel also runs the Logic stage on each finished path
int launch_trace(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s, bool fuse_logic = false);
int launch_trace_extension(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s, bool fuse_logic);
int launch_trace_shadow(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s);   // 0 launches at bounce 0
// Extension rays only, with a caller-provided (zeroed) work cursor: used to time the kernel alone.
int launch_trace_extension_only(const FrameDev& f, int bounce, int* cursor, const LaunchCfg& c, cudaStream_t s);
int launch_logic(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s);
int launch_shade(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s);
int launch_shade_one(const FrameDev& f, int bounce, int type, const LaunchCfg& c, cudaStream_t s);   // type: MAT_*
int launch_fold_env(const FrameDev& f, cudaStream_t s);
int launch_sample_environment(const FrameDev& f, const float* rnd, int n, float* dir, float* pdf, int* pixel, cudaStream_t s);
// ptb_config.megakernel (megakernel.cu): counters reset + the whole frame in one kernel
int launch_megakernel(const FrameDev& f, const LaunchCfg& c, cudaStream_t s);
int launch_frame_end(const FrameDev& f, const LaunchCfg& c, cudaStream_t s);                   // ray statistics
int launch_resolve(const FrameDev& f, void* dst, int format, int iteration, cudaStream_t s);
// headless views (views.cu)
int launch_gbuffer(const FrameDev& f, const GBufferDev& g, cudaStream_t s);
int launch_view(const GBufferDev& g, int n, void* dst_rgba8, int mode, cudaStream_t s);   // SendImageToPBOKernel modes 1-4
int launch_bvh_heatmap(const FrameDev& f, float* out_rgb, int* out_steps, cudaStream_t s);
// SVGF (svgf.cu): one frame of the filter over the 1-spp accumulators, 8 kernels
int launch_svgf(const FrameDev& f, const SvgfDev& s, const GBufferDev& g, int num_sms, cudaStream_t st);
int launch_resolve_final(const float* final_image, int n, void* dst, int format, cudaStream_t s);

// Batched RayCast on caller rays (ptb_trace_closest / ptb_trace_occluded): device arrays.
int launch_trace_batch_closest(const BvhDev& bvh, const float4* ori, const float4* dir, int n, int* hit, float* t, float* u,
                               float* v, int* work, const LaunchCfg& c, cudaStream_t s);
int launch_trace_batch_occluded(const BvhDev& bvh, const float4* ori_tmax, const float4* dir, int n, int* occluded, int* work,
                                int* error_flag, const LaunchCfg& c, cudaStream_t s);

}  // namespace ptb
