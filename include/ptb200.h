/* ptb200.h -- C ABI of the B200-native wavefront path-tracing core.
 *
 * Drop-in boundary for the render path of s64zxu/CUDA-Path-Tracer.  Every entry point cites
 * the reference interface it replaces (file:line into the reference tree).  Conventions:
 *   - plain pointers and sizes only, no C++/torch types; opaque handles, no process-global
 *     renderer state (the reference keeps one static renderer per process,
 *     src/pathtrace_wavefront.cu:118-120);
 *   - return 0 on success, a negative ptb_status on failure, NEVER exit()
 *     (the reference exits: src/common/cuda_utilities.cu:10-18, src/scene.cpp:54,87,126);
 *   - ptb_last_error() gives the message of the last failure on the calling thread;
 *   - there is NO CPU fallback: every compute entry point needs a CUDA device and fails
 *     with PTB_ERR_CUDA without one.
 */
#ifndef PTB200_H
#define PTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTB_ABI_VERSION 1

typedef enum ptb_status {
    PTB_OK = 0,
    PTB_ERR_INVALID = -1,      /* bad argument / handle */
    PTB_ERR_IO = -2,           /* scene / asset file could not be read */
    PTB_ERR_PARSE = -3,        /* malformed scene.json / OBJ / MTL / image */
    PTB_ERR_CUDA = -4,         /* CUDA runtime failure (message has the CUDA error string) */
    PTB_ERR_UNSUPPORTED = -5,  /* input the reference itself leaves undefined (e.g. < 2 triangles) */
    PTB_ERR_STATE = -6         /* call order violated (e.g. render before init) */
} ptb_status;

/* Material POD, 48 bytes, identical layout to the reference `Material`
 * (src/scene_structs.h:52-65).  type: 0 MicrofacetPBR, 1 DIFFUSE, 2 SPECULAR_REFLECTION,
 * 3 SPECULAR_REFRACTION (src/scene_structs.h:8-14). */
typedef struct ptb_material {
    float basecolor[3];
    float metallic;
    float roughness;
    float emittance;
    float ior;
    int32_t type;
    int32_t diffuse_tex_id;
    int32_t normal_tex_id;
    int32_t metallic_roughness_tex_id;
    int32_t _pad;
} ptb_material;

/* Camera, same members as the reference `Camera` (src/scene_structs.h:16-26) after
 * Scene::loadCamera (src/scene.cpp:541-577). */
typedef struct ptb_camera {
    int32_t resolution[2];
    float position[3];
    float look_at[3];
    float view[3];
    float up[3];
    float right[3];
    float fov[2];
    float pixel_length[2];
} ptb_camera;

typedef struct ptb_scene ptb_scene;        /* replaces `class Scene` (src/scene.h:51-92) */
typedef struct ptb_renderer ptb_renderer;  /* replaces WavefrontPathTracerState + the statics of
                                              src/pathtrace_wavefront.cu:118-120 */

/* ------------------------------------------------------------------ library */
int ptb_abi_version(void);
const char* ptb_last_error(void);
/* Number of CUDA devices visible; 0 if none.  Never fails. */
int ptb_device_count(void);

/* -------------------------------------------------------------------- scene */
/* Scene::Scene(string filename) (src/scene.cpp:45-56): same scene.json schema, OBJ/MTL
 * semantics, material-id order, vertex de-duplication order, light CDF and env alias table
 * (src/scene.cpp:83-631).  Pure host work; textures are uploaded by ptb_init. */
int ptb_scene_load(const char* json_path, ptb_scene** out);
/* ptb_scene_load with the vertex de-duplication (the unordered_map<Vertex,int> walk of src/scene.cpp:330-539, numbering at
 * :525-532) and the repack into upload layout with normalised normals / tangents (src/wavefront_internal.cu:118-167) done on GPU
 * `device`: every corner inserts itself into a device hash table that keeps the minimum corner index per key class, an exclusive
 * scan of the first-occurrence flags in file order numbers the vertices.  The resulting scene is bit-identical to
 * ptb_scene_load's.  PTB_ERR_CUDA when the device cannot be used (no fall-back to the host path); at most 2^30 triangle
 * corners.  Setting PTB_LOADER_GPU=<ordinal> in the environment makes ptb_scene_load itself take this route.
 * PTB_LOADER_VERBOSE=1 prints the stage's time on stderr for either route. */
int ptb_scene_load_device(const char* json_path, int32_t device, ptb_scene** out);

/* Direct SoA injection for synthetic meshes (what WavefrontPathTracerState::initSceneGeometry
 * builds from a Scene, src/wavefront_internal.cu:118-167).
 *   pos[3*V], nor[3*V] (unit length), uv[2*V] (may be NULL => 0), tangent[3*V] (may be NULL => 0),
 *   indices[3*N], material_ids[N], materials[M].
 * Geometric normals and the light CDF are derived exactly as src/scene.cpp:518-519,579-631. */
int ptb_scene_from_arrays(const float* pos, const float* nor, const float* uv, const float* tangent,
                          int32_t num_vertices, const int32_t* indices, const int32_t* material_ids,
                          int32_t num_triangles, const ptb_material* materials, int32_t num_materials,
                          const ptb_camera* camera, int32_t trace_depth, ptb_scene** out);
/* Attach an equirect RGB32F environment image (what Scene::buildEnvMapAliasTable does after
 * stbi_loadf, src/scene.cpp:183-300).  rgb has 3*width*height floats, row-major. */
int ptb_scene_set_environment(ptb_scene* s, const float* rgb, int32_t width, int32_t height);
/* Textures and the environment of a scene built with ptb_scene_from_arrays, from images the caller already holds -- what
 * a live reference `Scene` keeps as CUDA arrays behind Scene::texture_handles (Scene::createTexture, src/scene.h:94-145):
 * ptb_scene_add_texture appends one image and returns its id (>= 0; ids count up in call order, as Scene::loadTexture /
 * buildEnvMapAliasTable assign them, src/scene.cpp:58-81,283-297) or a negative status.  PTB_TEX_RGBA8: uchar4 read as
 * normalised float; PTB_TEX_R32F / PTB_TEX_RGBA32F: float / float4 read as they are.  X wraps, Y wraps or clamps (clamp_y).
 * ptb_scene_set_environment_table then names the two textures that are the environment and its pdf map and hands over
 * the Vose alias table Scene::buildEnvMapAliasTable built (EnvMap::probs / aliases, src/scene.h:37-48). */
enum { PTB_TEX_RGBA8 = 0, PTB_TEX_R32F = 1, PTB_TEX_RGBA32F = 2 };
int ptb_scene_add_texture(ptb_scene* s, const void* pixels, int32_t width, int32_t height, int32_t format, int32_t clamp_y);
int ptb_scene_set_environment_table(ptb_scene* s, int32_t env_tex_id, int32_t pdf_map_id, const float* probs,
                                    const int32_t* aliases, int32_t width, int32_t height);
/* The decoded image behind texture `id` as it will be uploaded (what Scene::loadTexture hands to createTexture, src/scene.cpp:148-172:
 * uchar4 with grey replicated and alpha 255 where the file has none): copies min(bytes, size), returns the full size in bytes or
 * a negative status; width / height / format (PTB_TEX_*) may be NULL.  Texture files are decoded by this library's own PNG / BMP /
 * TGA / PNM readers (the reference uses stb_image); a file that exists but cannot be decoded makes ptb_scene_load fail with
 * PTB_ERR_UNSUPPORTED, a missing file leaves the material untextured as in the reference (src/scene.cpp:141-146). */
int64_t ptb_scene_get_texture(const ptb_scene* s, int32_t id, void* dst, size_t bytes, int32_t* width, int32_t* height, int32_t* format);
/* Recompute camera.view/right/up/fov/pixel_length from eye/lookat/up/fovy/resolution
 * (src/scene.cpp:560-572) and store it in the scene. */
int ptb_scene_set_camera(ptb_scene* s, const int32_t resolution[2], float fovy_deg,
                         const float eye[3], const float look_at[3], const float up[3], int32_t trace_depth);
/* Binary scene cache: the loaded scene (geometry in upload layout, materials, light CDF, decoded textures, the
 * environment alias table, camera) written verbatim, so a later start skips OBJ / MTL / image parsing,
 * de-duplication and table building (Scene::Scene does all of that on every run, src/scene.cpp:45-631).
 * A scene read back is bit-identical to the one saved.  Host byte order; PTB_ERR_PARSE for a file that is not a
 * cache of this version, is truncated or fails the index range checks. */
int ptb_scene_save_cache(const ptb_scene* s, const char* path);
int ptb_scene_load_cache(const char* path, ptb_scene** out);
void ptb_scene_free(ptb_scene* s);

/* Image::savePNG / Image::saveHDR (src/image.cpp:22-49) with the transform of main.cpp's saveImage()
 * (src/main.cpp:239-262) as options.  rgb: float3[height*width], row-major.  Every channel is multiplied by
 * `scale` (the reference divides by the sample count), then PTB_SAVE_GAMMA applies pow(max(0, c), 1/2.2) and
 * PTB_SAVE_FLIP_X mirrors the columns (the reference writes pixel x to width-1-x).  The extension picks the
 * format: .png (8-bit RGB, clamp to [0,1] x 255, truncated) or .hdr (Radiance RGBE).  Pure host code. */
enum { PTB_SAVE_GAMMA = 1, PTB_SAVE_FLIP_X = 2 };
int ptb_save_image(const char* path, const float* rgb, int32_t width, int32_t height, float scale, int32_t flags);

typedef struct ptb_scene_info {
    int32_t num_vertices, num_triangles, num_materials, num_lights, num_textures;
    int32_t width, height, trace_depth, iterations;
    int32_t env_width, env_height, env_tex_id, pdf_map_id;
    float light_total_area;
} ptb_scene_info;
int ptb_scene_get_info(const ptb_scene* s, ptb_scene_info* out);
int ptb_scene_get_camera(const ptb_scene* s, ptb_camera* out);

/* Host arrays of the loaded scene, for parity against Scene's public vectors
 * (src/scene.h:77-89) in the device-upload layout of src/wavefront_internal.cu:118-167. */
typedef enum ptb_scene_array {
    PTB_SA_POS = 0,            /* float4[V]  (xyz, 1)            */
    PTB_SA_NOR = 1,            /* float4[V]  (normalize(n), 0)   */
    PTB_SA_TANGENT = 2,        /* float4[V]  (normalize(t), 0)   */
    PTB_SA_UV = 3,             /* float2[V]                      */
    PTB_SA_INDICES_MATID = 4,  /* int4[N]    (v0,v1,v2,mat)      */
    PTB_SA_NOR_GEOM = 5,       /* float4[N]                      */
    PTB_SA_MATERIALS = 6,      /* ptb_material[M]                */
    PTB_SA_LIGHT_TRI_IDX = 7,  /* int[L]                         */
    PTB_SA_LIGHT_CDF = 8,      /* float[L]                       */
    PTB_SA_ENV_PROBS = 9,      /* float[env_w*env_h]             */
    PTB_SA_ENV_ALIASES = 10,   /* int[env_w*env_h]               */
    PTB_SA_ENV_PDF = 11        /* float[env_w*env_h]             */
} ptb_scene_array;
/* Copies min(bytes, size) bytes; returns the full size in bytes (>= 0) or a negative status. */
int64_t ptb_scene_get_array(const ptb_scene* s, int which, void* dst, size_t bytes);

/* ----------------------------------------------------------------- renderer */
typedef struct ptb_config {
    int32_t device;          /* CUDA device ordinal (the reference hard-codes 0, src/main.cpp:168) */
    int32_t trace_depth;     /* <= 0: take the scene's DEPTH (src/scene.cpp:549) */
    /* Multi-GPU pixel partition (new; the reference is single-GPU).  Rank r of `world` owns the
     * image rows y with (y / stripe_rows) % world == r.  Pixel indices, and therefore RNG seeds
     * (src/kernels/ray_gen.cu:41), stay GLOBAL, so the union of all ranks' tiles is bit-identical
     * to a 1-GPU render.  world <= 1 renders the full frame. */
    int32_t rank, world, stripe_rows;
    int32_t use_graph;       /* 1: replay the frame as a CUDA graph (default); 0: plain launches */
    int32_t gbuffer;         /* 1: keep the first-bounce G-buffer the reference's Logic writes every frame (depth, normal +
                                material id, albedo, motion; src/kernels/logic.cu:66-73,119-138); 0 (default): skip it */
    int32_t denoiser;        /* 1: SVGF (src/kernels/svgf.cu) as in the reference's DenoiserOn mode: every frame is a fresh 1-spp
                                estimate (accumulators cleared per frame, src/pathtrace_wavefront.cu:231-236), filtered into the
                                final image; implies gbuffer; iteration 1 restarts the history; world must be 1 */
    int32_t env_nee;         /* 1: next-event estimation of the HDR environment through the alias table (sampleEnvironmentMap,
                                src/interactions.cu:447-488): the half of the environment's MIS pair the reference's Logic already
                                assumes (src/kernels/logic.cu:88-100) but nothing in the reference runs.  CHANGES THE IMAGE: surfaces
                                lit by the environment gain the missing energy, the indirect image and the light sampling stay bit
                                for bit what they are with 0 (default = the reference's result).  No effect without an environment. */
    int32_t megakernel;      /* 1: render with the reference's OTHER path, pathtrace_megakernel::Pathtrace (src/pathtrace.cu:645-706):
                                one kernel carries each path through all bounces.  It is a different estimator, not only a
                                different schedule (own seeding, mirrored camera x axis, unweighted environment hits, NEE that
                                advances the path's random stream; src/pathtrace.cu:234-452), so its image is the reference
                                MEGAKERNEL's image, not the wavefront's.  The sample sum lands in the direct accumulator
                                (its d_image), the indirect one stays zero; ptb_resolve divides by the iteration, which equals the
                                reference's per-pixel sample count for iterations 1..n.  Not combinable with gbuffer, denoiser,
                                env_nee or the staged API (PTB_ERR_UNSUPPORTED / PTB_ERR_STATE). */
    int32_t reserved[4];
} ptb_config;
void ptb_config_default(ptb_config* cfg);

/* pathtrace_wavefront::PathtraceInit(Scene*) (src/pathtrace_wavefront.cu:144-167): uploads mesh,
 * materials and textures, builds the LBVH on the device (src/bvh.cu:456-516) and allocates the
 * path pool / queues (src/wavefront_internal.cu:29-49).  The scene may be freed afterwards. */
int ptb_init(const ptb_scene* scene, const ptb_config* cfg, ptb_renderer** out);
/* pathtrace_wavefront::PathtraceFree() (src/pathtrace_wavefront.cu:169-181). */
int ptb_free(ptb_renderer* r);

/* pathtrace_wavefront::Pathtrace(pbo, frame, iteration) (src/pathtrace_wavefront.cu:205-413):
 * one sample per pixel through trace_depth bounces, accumulated into the direct / indirect
 * buffers.  `iteration` is 1-based; iteration == 1 clears the accumulators (:232-242).
 * Asynchronous: returns after enqueueing on the renderer's stream. */
int ptb_render(ptb_renderer* r, int32_t iteration);
/* Same as calling ptb_render for iteration = first .. first+count-1. */
int ptb_render_many(ptb_renderer* r, int32_t first_iteration, int32_t count);
/* Zero the direct / indirect accumulators without rendering: what iteration == 1 does implicitly
 * (src/pathtrace_wavefront.cu:232-242), for callers that start at another iteration -- a rank of a
 * sample-index partition renders iterations rank+1, rank+1+world, ... and never sees iteration 1. */
int ptb_clear_accumulators(ptb_renderer* r);
/* Wait for everything enqueued on the renderer's stream; reports deferred CUDA errors. */
int ptb_sync(ptb_renderer* r);
/* Use a caller-owned CUDA stream (cudaStream_t passed as void*) instead of the internal one. */
int ptb_set_stream(ptb_renderer* r, void* cuda_stream);

/* SendImageToPBOKernel (src/pathtrace_wavefront.cu:47-116) into a headless buffer.
 * PTB_IMG_SUM_RGB32F : direct+indirect, float3 per pixel (not divided)
 * PTB_IMG_MEAN_RGB32F: (direct+indirect)/iteration               (:74-75)
 * PTB_IMG_RGBA8      : gamma 1/2.2, x255, clamp, alpha 255       (:80,112-115)
 * dst holds owned_pixels (see ptb_get_tile_info) elements when the renderer is a tile of a
 * multi-GPU frame, else width*height.  dst_on_device != 0: dst is a device pointer (written on
 * the renderer's stream, asynchronous); else a host pointer (synchronous).
 * PTB_IMG_VIEW_*     : the other display modes of that kernel (:82-109) as RGBA8, from the G-buffer
 *                      (needs ptb_config.gbuffer): normals*0.5+0.5, 1-depth/500, albedo, motion*100+0.5 */
enum { PTB_IMG_SUM_RGB32F = 0, PTB_IMG_MEAN_RGB32F = 1, PTB_IMG_RGBA8 = 2,
       PTB_IMG_VIEW_NORMALS = 3, PTB_IMG_VIEW_DEPTH = 4, PTB_IMG_VIEW_ALBEDO = 5, PTB_IMG_VIEW_MOTION = 6,
       PTB_IMG_STEPS_I32 = 7 /* ptb_visualize_lbvh only: int32 node visits per pixel */,
       PTB_IMG_DENOISED_RGB32F = 8 /* SVGF output d_final_image, float3, no gamma (needs ptb_config.denoiser) */,
       PTB_IMG_DENOISED_RGBA8 = 9 /* the same through display mode 0: gamma 1/2.2, x255, clamp */ };
int ptb_resolve(ptb_renderer* r, void* dst, int32_t format, int32_t iteration, int32_t dst_on_device);
/* First-bounce G-buffer of the last rendered frame, one element per owned pixel (needs ptb_config.gbuffer):
 * WavefrontPathTracerState::d_depth / d_normal_matid / d_albedo / d_motion_vectors (src/wavefront_internal.cu:286-300). */
enum { PTB_GB_DEPTH = 0 /* float */, PTB_GB_NORMAL_MATID = 1 /* float4 */, PTB_GB_ALBEDO = 2 /* float4 */, PTB_GB_MOTION = 3 /* float2 */ };
int ptb_get_gbuffer(ptb_renderer* r, int32_t which, void* dst, size_t bytes);
/* VisualizeLBVH (src/bvh.cu:518-640): per pixel, the number of LBVH nodes an un-jittered camera ray visits, as the
 * reference's blue-to-red colour (PTB_IMG_SUM_RGB32F, float3, = its d_final_image) or as the raw count
 * (PTB_IMG_STEPS_I32).  Replaces the GL "Show BVH Heatmap" view (src/preview.cpp:159). */
int ptb_visualize_lbvh(ptb_renderer* r, void* dst, int32_t format, int32_t dst_on_device);

/* sampleEnvironmentMap (src/interactions.cu:447-488) for `count` pairs of uniform numbers rnd[2*count] (host): the
 * sampled directions dir[3*count], their solid-angle pdfs pdf[count] and the chosen texels pixel[count] (host).  What
 * ptb_config.env_nee draws from; PTB_ERR_STATE when the scene has no environment map. */
int ptb_sample_environment(ptb_renderer* r, const float* rnd, int32_t count, float* dir, float* pdf, int32_t* pixel);
/* Raw accumulators (WavefrontPathTracerState::d_direct_image / d_indirect_image,
 * src/wavefront_internal.h:26-27) copied to host, 3 floats per owned pixel; either may be NULL. */
int ptb_get_accumulators(ptb_renderer* r, float* direct, float* indirect);

typedef struct ptb_tile_info {
    int32_t width, height;     /* full frame */
    int32_t owned_pixels;      /* pixels this renderer traces */
    int32_t rank, world, stripe_rows;
} ptb_tile_info;
int ptb_get_tile_info(const ptb_renderer* r, ptb_tile_info* out);
/* Global pixel index of every owned pixel, in the order ptb_resolve / ptb_get_accumulators use. */
int ptb_get_owned_pixels(const ptb_renderer* r, int32_t* dst, int32_t capacity);

/* ---------------------------------------------------------------- multi-GPU */
/* New (the reference renders on device 0 of one process, src/main.cpp:168).  Every GPU keeps a scene + LBVH replica and
 * renders a share of the frame; the resolve kernel of each GPU (the counterpart of SendImageToPBOKernel,
 * src/pathtrace_wavefront.cu:47-116) writes the sums of the pixels it owns directly into ONE full-frame buffer on the root
 * GPU through NVLink peer stores, so the gather is the tail of that kernel and no collective library sits on the data
 * path.  Two partitions compose: `tile_groups` G interleaved stripe sets of the frame (ptb_config.rank / world, bit-identical
 * to one GPU) x `n_devices / G` sample groups (group s renders iterations s+1, s+1+S, ...; their sums are folded on the
 * root in group order, so the result is deterministic and differs from one GPU only in float summation order). */
#define PTB_MAX_DEVICES 16
typedef struct ptb_multi ptb_multi;
typedef struct ptb_multi_config {
    int32_t n_devices;                    /* 1 .. PTB_MAX_DEVICES */
    int32_t devices[PTB_MAX_DEVICES];     /* CUDA ordinals; devices[0] is the root: it owns the gathered frame.  An ordinal may
                                             appear more than once (several renderers then share that GPU) */
    int32_t tile_groups;                  /* G; 0 = n_devices (pure tile partition); must divide n_devices */
    int32_t stripe_rows;                  /* rows per stripe of the tile partition (default 8) */
    int32_t trace_depth, use_graph, megakernel, env_nee;   /* as in ptb_config */
    int32_t reserved[6];
} ptb_multi_config;
typedef struct ptb_multi_info {
    int32_t n_devices, tile_groups, sample_groups;
    int32_t width, height;
    int32_t samples;                      /* samples per pixel accumulated so far */
    int32_t peer_stores;                  /* devices whose resolve kernel stores into the root's memory (the others copy) */
    int32_t reserved;
} ptb_multi_info;
void ptb_multi_config_default(ptb_multi_config* cfg);
/* PathtraceInit on every listed device (scene upload + LBVH build per GPU) + the root's frame buffer + peer access. */
int ptb_multi_init(const ptb_scene* scene, const ptb_multi_config* cfg, ptb_multi** out);
int ptb_multi_free(ptb_multi* m);
/* Steps first .. first+count-1 (1-based; step 1 restarts the accumulation).  One step adds `sample_groups` samples per pixel:
 * device k (tile group k % G, sample group k / G) renders iteration (step-1)*S + k/G + 1 of its stripes.  Asynchronous. */
int ptb_multi_render(ptb_multi* m, int32_t first_step, int32_t count);
/* Gather + fold + format into `dst` (full frame: width*height float3, or uchar4 for PTB_IMG_RGBA8), on the root device
 * (dst_on_device != 0, asynchronous on the root's stream; ptb_multi_sync waits) or on the host (synchronous). */
int ptb_multi_resolve(ptb_multi* m, void* dst, int32_t format, int32_t dst_on_device);
int ptb_multi_sync(ptb_multi* m);
/* The per-device renderer (statistics, parity views); owned by the ptb_multi. */
int ptb_multi_get_renderer(ptb_multi* m, int32_t index, ptb_renderer** out);
int ptb_multi_get_info(const ptb_multi* m, ptb_multi_info* out);

/* The same gather for ONE PROCESS PER GPU (torchrun): the root rank creates the full-frame buffer and exports it as a CUDA
 * IPC handle; every other rank opens it and its resolve kernel stores into the root's memory over NVLink.  The ranks
 * only need a barrier between ptb_resolve_into and the root's ptb_gather_combine (cuda_path_tracer_b200/multigpu.py uses
 * a 4-byte NCCL all-reduce on the renderers' stream).  slices = sample groups (1 for a pure tile partition). */
#define PTB_GATHER_HANDLE_BYTES 64
typedef struct ptb_gather ptb_gather;
int ptb_gather_create(int32_t device, int32_t width, int32_t height, int32_t slices, ptb_gather** out);
int ptb_gather_export(ptb_gather* g, void* handle, size_t bytes);
int ptb_gather_open(int32_t device, const void* handle, int32_t width, int32_t height, int32_t slices, ptb_gather** out);
int ptb_gather_free(ptb_gather* g);
/* direct + indirect of every pixel the renderer owns -> slice `slice` of the buffer, at the pixels' global positions. */
int ptb_resolve_into(ptb_renderer* r, ptb_gather* g, int32_t slice);
/* Root only: fold the slices in order and write `format` (PTB_IMG_SUM_RGB32F / _MEAN_RGB32F / _RGBA8; `samples` = samples
 * per pixel the sums hold) on cuda_stream (cudaStream_t as void*); host destinations are synchronous. */
int ptb_gather_combine(ptb_gather* g, void* dst, int32_t format, int32_t samples, int32_t dst_on_device, void* cuda_stream);

/* ------------------------------------------------------- LBVH (parity view) */
/* The LBVH in the reference's own layout (`LBVHData`, src/scene_structs.h:151-169): node index
 * space [0,N-2] internal, N-1 gap, [N,2N-1] leaves; leaf children stored bit-inverted. */
typedef enum ptb_lbvh_array {
    PTB_LBVH_AABB_MIN = 0,           /* float4[2N] */
    PTB_LBVH_AABB_MAX = 1,           /* float4[2N] */
    PTB_LBVH_CENTROID = 2,           /* float4[2N] */
    PTB_LBVH_PRIMITIVE_INDICES = 3,  /* int[2N]    */
    PTB_LBVH_MORTON_CODES = 4,       /* uint64[N]  */
    PTB_LBVH_CHILD_NODES = 5,        /* int2[N-1]  */
    PTB_LBVH_PARENT = 6,             /* int[2N]    */
    PTB_LBVH_ESCAPE_INDICES = 7,     /* int[2N]    */
    PTB_LBVH_WORLD_AABB = 8          /* float4[2] (min, max) */
} ptb_lbvh_array;
int64_t ptb_get_lbvh(ptb_renderer* r, int which, void* dst, size_t bytes);

/* BuildLBVH(LBVHData&, const MeshData&) (src/bvh.h:7, src/bvh.cu:456-516) as a standalone call
 * on HOST arrays: pos float4[V], indices_matid int4[N]; host<->device copies inside.
 * Outputs (any may be NULL) use the layouts listed above.  build_ms (may be NULL) receives the
 * device time of the build itself (CUDA events, excluding the copies). */
int ptb_build_lbvh(int32_t device, const float* pos4, int32_t num_vertices, const int32_t* indices_matid4,
                   int32_t num_triangles, uint64_t* morton_codes, int32_t* primitive_indices,
                   int32_t* child_nodes, int32_t* parent, int32_t* escape_indices,
                   float* aabb_min4, float* aabb_max4, float* world_aabb8, float* build_ms);

/* -------------------------------------------------------- RayCast (batched) */
/* TraceExtensionRay (src/kernels/ray_cast.cu:204-218) over a batch of HOST rays against the
 * renderer's LBVH: ori4/dir4 are float4[n] (xyz used).  Outputs: hit_geom_id[n] (-1 miss),
 * t[n] (-1 miss), u[n], v[n].  Host<->device copies inside (this is the e2e RayCast call). */
int ptb_trace_closest(ptb_renderer* r, const float* ori4, const float* dir4, int32_t n,
                      int32_t* hit_geom_id, float* t, float* u, float* v);
/* TraceShadowRay's visibility test (src/kernels/ray_cast.cu:125-201): ori_tmax4 float4[n]
 * (xyz origin, w = tmax), dir4 float4[n]; occluded[n] receives 0/1. */
int ptb_trace_occluded(ptb_renderer* r, const float* ori_tmax4, const float* dir4, int32_t n,
                       int32_t* occluded);

/* ptb_trace_closest for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same arguments and
 * results as ptb_trace_closest (host arrays in, host arrays out, synchronous), but the staged rays, the staged results and the
 * kernel's work cursor live in ONE file-scope static __device__ array (csrc/host/noalloc.cu) instead of a cudaMalloc'ed scratch
 * buffer, and the call makes no cudaMalloc / cudaFree / event creation or any other allocating call.
 *   - n > PTB_NOALLOC_MAX_RAYS is rejected with PTB_ERR_INVALID (split the batch, or use ptb_trace_closest);
 *   - the array is PTB_NOALLOC_MAX_RAYS * 48 B + 256 B (12 MB) and every GPU that loads the library holds it until exit;
 *   - all calls on a GPU share it and it keeps the previous call's contents (the cursor is cleared per call): calls that
 *     overlap on the same GPU race -- serialise them;
 *   - ptb_stats.last_batch_ms is not updated (timing would need events). */
#define PTB_NOALLOC_MAX_RAYS 262144
int ptb_trace_closest_noalloc(ptb_renderer* r, const float* ori4, const float* dir4, int32_t n,
                              int32_t* hit_geom_id, float* t, float* u, float* v);

/* ---------------------------------------------- staged frame (parity/debug) */
/* The reference's stage wrappers driven one at a time, in the order of
 * src/pathtrace_wavefront.cu:278-348, so that tests can read queues between stages. */
typedef enum ptb_stage {
    PTB_STAGE_RAYGEN = 0,   /* ray_gen::GeneratePathTracingRays   (src/kernels/ray_gen.cu:84-131) */
    PTB_STAGE_TRACE = 1,    /* TraceExtensionRay + TraceShadowRay (src/kernels/ray_cast.cu:204-234) */
    PTB_STAGE_LOGIC = 2,    /* logic::RunPathLogic                (src/kernels/logic.cu:202-227) */
    PTB_STAGE_SHADE = 3     /* shading::Sample* x4                (src/kernels/shading.cu:274-304) */
} ptb_stage;
int ptb_stage_run(ptb_renderer* r, int32_t stage, int32_t iteration);

typedef enum ptb_queue {
    PTB_Q_EXTENSION = 0, PTB_Q_SHADOW = 1, PTB_Q_PBR = 2, PTB_Q_DIFFUSE = 3,
    PTB_Q_REFLECTION = 4, PTB_Q_REFRACTION = 5
} ptb_queue;
/* Current length of a queue (the reference's d_*_counter, src/wavefront_internal.h:48-57). */
int ptb_get_queue_count(ptb_renderer* r, int32_t which, int32_t* count);
/* Queue contents: path indices (global pixel index == path slot, src/kernels/ray_gen.cu:30-31);
 * for PTB_Q_SHADOW the pixel index of each shadow ray.  Order is unspecified (as in the
 * reference, where it follows atomicAdd arrival, src/common/cuda_utilities.h:216-231). */
int ptb_get_queue(ptb_renderer* r, int32_t which, int32_t* dst, int32_t capacity);
/* Shadow-queue payload (ShadowQueue, src/scene_structs.h:121-131): float4[count] each. */
int ptb_get_shadow_queue(ptb_renderer* r, float* ori_tmax4, float* dir4, float* radiance4,
                         int32_t* pixel_idx, int32_t capacity);

typedef enum ptb_path_field {
    PTB_PF_RAY_ORI = 0,            /* float4[P] */
    PTB_PF_RAY_DIR_DIST = 1,       /* float4[P] (dir, t / -1 / FLT_MAX) */
    PTB_PF_HIT_UV = 2,             /* float4[P] (u, v, 0, 0): the reference's `hit_normal` */
    PTB_PF_THROUGHPUT_PDF = 3,     /* float4[P] */
    PTB_PF_HIT_GEOM_ID = 4,        /* int[P] */
    PTB_PF_MATERIAL_ID = 5,        /* int[P] */
    PTB_PF_REMAINING_BOUNCES = 6,  /* int[P] */
    PTB_PF_RNG_STATE = 7           /* uint32[P] */
} ptb_path_field;
/* PathState (src/scene_structs.h:97-117) scattered back to slot == GLOBAL pixel order over the
 * full frame; pixels this renderer does not own are left untouched in dst. */
int ptb_get_path_field(ptb_renderer* r, int32_t field, void* dst, size_t bytes);

/* -------------------------------------------------------------------- stats */
typedef struct ptb_stats {
    uint64_t extension_rays;   /* rays launched since the last ptb_stats_reset, same definition as */
    uint64_t shadow_rays;      /* the reference's MraysPerSec (src/pathtrace_wavefront.cu:305,375-379) */
    uint64_t frames;
    uint64_t kernel_launches;  /* kernels of this library enqueued (graph nodes count individually) */
    float lbvh_build_ms;       /* "LBVH Build Time" (src/bvh.cu:467-498) */
    int32_t lbvh_max_depth;    /* "LBVH Max Tree Depth" (src/bvh.cu:416-454) */
    int32_t num_triangles;
    float last_batch_ms;       /* device time of the kernel of the most recent ptb_trace_closest / ptb_trace_occluded (CUDA events) */
} ptb_stats;
int ptb_get_stats(ptb_renderer* r, ptb_stats* out);
int ptb_stats_reset(ptb_renderer* r);

/* Time `count` launches of the extension-ray RayCast kernel alone on the current extension
 * queue (after PTB_STAGE_RAYGEN or a shade stage), with CUDA events on the renderer's stream.
 * Results are discarded-safe: the kernel is idempotent on the same queue.  Used by bench.py
 * for the roofline figure. */
int ptb_time_raycast(ptb_renderer* r, int32_t count, float* avg_ms, int32_t* rays);
/* The same for either RayCast kernel of the current bounce: PTB_TIME_EXTENSION (TraceExtensionRayKernel's counterpart alone,
 * without the fused Logic) or PTB_TIME_SHADOW (TraceShadowRayKernel's; the direct accumulator it adds into is saved and
 * restored, so the frame is unchanged; 0 rays / 0 ms before the first Shade).  cold != 0: a 256 MB write flushes the L2
 * before every launch and each launch is timed by its own event pair; cold == 0: back-to-back launches, tree warm in L2. */
enum { PTB_TIME_EXTENSION = 0, PTB_TIME_SHADOW = 1 };
int ptb_time_kernel(ptb_renderer* r, int32_t which, int32_t count, int32_t cold, float* avg_ms, int32_t* rays);

#ifdef __cplusplus
}
#endif
#endif /* PTB200_H */
