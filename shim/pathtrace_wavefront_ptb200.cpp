// shim/pathtrace_wavefront_ptb200.cpp -- source-level drop-in for the reference's src/pathtrace_wavefront.cu.
//
// A maintainer of s64zxu/CUDA-Path-Tracer compiles THIS file instead of src/pathtrace_wavefront.cu, src/wavefront_internal.cu,
// src/bvh.cu and src/kernels/*.cu, against the reference's UNCHANGED headers (src/pathtrace_wavefront.h:6-12, src/scene.h,
// src/common/utilities.h), and links -lptb200.  main.cpp's callers (src/main.cpp:33-60, headless loop :211-224) do not change.
// Built and tested here by oracle/Makefile (target shim) + shim/shim_driver.cpp + tests/test_gpu_parity.py::test_shim_*.
//
// Two ways to hand the scene over, chosen by the environment variable PTB_SHIM_SCENE:
//   "file"  (default)  ptb_scene_load() parses the same scene.json again.  `class Scene` does not keep its file name
//                      (src/scene.h:73-92), so the host application exports the argv[1] it built the Scene from as
//                      `const char* g_scene_file`.
//   "live"             everything comes from the live Scene object: vertices / indices / materialIds / materials / camera
//                      through ptb_scene_from_arrays, every texture read back from the CUDA array behind
//                      Scene::texture_handles (Scene::createTexture, src/scene.h:94-145), the alias table from Scene::env_map.
#include "pathtrace_wavefront.h"      // reference header, unchanged
#include "scene.h"                    // reference header, unchanged
#include "ptb200.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

extern const char* g_scene_file;      // defined by the host application (see above)

namespace pathtrace_wavefront {

static ptb_scene* g_scene = nullptr;          // the reference keeps the same process-global state,
static ptb_renderer* g_renderer = nullptr;    // src/pathtrace_wavefront.cu:118-120
static GuiDataContainer* g_gui = nullptr;

ptb_renderer* ShimRenderer() { return g_renderer; }   // not part of the reference interface: lets a test read the accumulators

static void die(const char* what) {           // the reference's behaviour on failure: print and exit (src/common/cuda_utilities.cu:3-19)
    std::fprintf(stderr, "%s: %s\n", what, ptb_last_error());
    std::exit(EXIT_FAILURE);
}

void InitDataContainer(GuiDataContainer* guiData) { g_gui = guiData; }     // src/pathtrace_wavefront.cu:122-125

static void scene_from_live(Scene* sc) {
    const size_t V = sc->vertices.size(), N = sc->indices.size() / 3;
    std::vector<float> pos(3 * V), nor(3 * V), uv(2 * V), tan(3 * V);
    for (size_t i = 0; i < V; i++) {
        const Vertex& v = sc->vertices[i];
        pos[3 * i] = v.pos.x; pos[3 * i + 1] = v.pos.y; pos[3 * i + 2] = v.pos.z;
        nor[3 * i] = v.nor.x; nor[3 * i + 1] = v.nor.y; nor[3 * i + 2] = v.nor.z;       // normalised at upload, as
        tan[3 * i] = v.tangent.x; tan[3 * i + 1] = v.tangent.y; tan[3 * i + 2] = v.tangent.z;   // src/wavefront_internal.cu:136-144
        uv[2 * i] = v.uv.x; uv[2 * i + 1] = v.uv.y;
    }
    const Camera& c = sc->state.camera;
    ptb_camera cam;
    cam.resolution[0] = c.resolution.x; cam.resolution[1] = c.resolution.y;
    for (int k = 0; k < 3; k++) {
        cam.position[k] = c.position[k]; cam.look_at[k] = c.lookAt[k]; cam.view[k] = c.view[k];
        cam.up[k] = c.up[k]; cam.right[k] = c.right[k];
    }
    cam.fov[0] = c.fov.x; cam.fov[1] = c.fov.y;
    cam.pixel_length[0] = c.pixelLength.x; cam.pixel_length[1] = c.pixelLength.y;
    static_assert(sizeof(Material) == sizeof(ptb_material), "ptb_material must mirror the reference Material (src/scene_structs.h:52-65)");
    if (ptb_scene_from_arrays(pos.data(), nor.data(), uv.data(), tan.data(), (int32_t)V, sc->indices.data(), sc->materialIds.data(), (int32_t)N,
                              reinterpret_cast<const ptb_material*>(sc->materials.data()), (int32_t)sc->materials.size(), &cam,
                              sc->state.traceDepth, &g_scene) != PTB_OK) die("ptb_scene_from_arrays");
    // textures: what the Scene uploaded, read back from its own CUDA arrays, in id order
    for (cudaTextureObject_t h : sc->texture_handles) {
        cudaResourceDesc rd; cudaTextureDesc td; cudaChannelFormatDesc fd; cudaExtent ext; unsigned int flags = 0;
        if (cudaGetTextureObjectResourceDesc(&rd, h) != cudaSuccess || cudaGetTextureObjectTextureDesc(&td, h) != cudaSuccess ||
            cudaArrayGetInfo(&fd, &ext, &flags, rd.res.array.array) != cudaSuccess) { std::fprintf(stderr, "shim: cannot query a texture of the Scene\n"); std::exit(EXIT_FAILURE); }
        int format; size_t elem;
        if (fd.f == cudaChannelFormatKindUnsigned && fd.x == 8 && fd.w == 8) { format = PTB_TEX_RGBA8; elem = 4; }
        else if (fd.f == cudaChannelFormatKindFloat && fd.y == 0) { format = PTB_TEX_R32F; elem = 4; }
        else if (fd.f == cudaChannelFormatKindFloat && fd.w == 32) { format = PTB_TEX_RGBA32F; elem = 16; }
        else { std::fprintf(stderr, "shim: texture format the reference never creates\n"); std::exit(EXIT_FAILURE); }
        std::vector<unsigned char> px(ext.width * ext.height * elem);
        if (cudaMemcpy2DFromArray(px.data(), ext.width * elem, rd.res.array.array, 0, 0, ext.width * elem, ext.height, cudaMemcpyDeviceToHost) != cudaSuccess) {
            std::fprintf(stderr, "shim: texture read-back failed\n"); std::exit(EXIT_FAILURE);
        }
        if (ptb_scene_add_texture(g_scene, px.data(), (int32_t)ext.width, (int32_t)ext.height, format, td.addressMode[1] == cudaAddressModeClamp) < 0) die("ptb_scene_add_texture");
    }
    const EnvMap& e = sc->env_map;
    if (e.width > 0 && e.env_tex_id >= 0)
        if (ptb_scene_set_environment_table(g_scene, e.env_tex_id, e.pdf_map_id, e.probs.data(), e.aliases.data(), e.width, e.height) != PTB_OK) die("ptb_scene_set_environment_table");
}

void PathtraceInit(Scene* scene) {                                         // src/pathtrace_wavefront.cu:144-167
    const char* mode = std::getenv("PTB_SHIM_SCENE");
    if (mode && std::strcmp(mode, "live") == 0) scene_from_live(scene);
    else if (ptb_scene_load(g_scene_file, &g_scene) != PTB_OK) die("ptb_scene_load");
    ptb_config cfg;
    ptb_config_default(&cfg);                                              // device 0 (src/main.cpp:168), full frame, CUDA graph
    cfg.denoiser = (g_gui && g_gui->DenoiserOn) ? 1 : 0;                   // fixed at init here; the reference reads it every frame (:231)
    if (ptb_init(g_scene, &cfg, &g_renderer) != PTB_OK) die("ptb_init");
}

void PathtraceFree() {                                                     // src/pathtrace_wavefront.cu:169-181
    ptb_free(g_renderer); g_renderer = nullptr;
    ptb_scene_free(g_scene); g_scene = nullptr;
}

void InitBVH(Scene*) {}   // declared by the reference header, defined nowhere in the reference (src/pathtrace_wavefront.h:11)

void Pathtrace(uchar4* pbo, int /*frame*/, int iteration) {                // src/pathtrace_wavefront.cu:205-413
    if (!g_renderer) { std::printf("Pathtrace: not initialised\n"); return; }   // :207-210
    ptb_stats before, after;
    const bool stats = g_gui != nullptr;
    std::chrono::steady_clock::time_point t0;
    if (stats) { ptb_get_stats(g_renderer, &before); t0 = std::chrono::steady_clock::now(); }
    if (ptb_render(g_renderer, iteration) != PTB_OK) die("ptb_render");
    if (pbo) {                                                             // :385-393, SendImageToPBOKernel into the caller's device buffer
        const int fmt = (g_gui && g_gui->DenoiserOn) ? PTB_IMG_DENOISED_RGBA8 : PTB_IMG_RGBA8;
        if (ptb_resolve(g_renderer, pbo, fmt, iteration, /*dst_on_device=*/1) != PTB_OK) die("ptb_resolve");
    }
    if (ptb_sync(g_renderer) != PTB_OK) die("ptb_sync");                   // the reference ends a frame synchronised (checkCUDAError, :395)
    if (stats) {                                                           // :371-379: rays of this frame over its wall time
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        ptb_get_stats(g_renderer, &after);
        const double rays = double(after.extension_rays - before.extension_rays) + double(after.shadow_rays - before.shadow_rays);
        g_gui->MraysPerSec = ms > 0.0 ? float(rays / (ms * 1e3)) : 0.0f;
    }
}

}  // namespace pathtrace_wavefront
