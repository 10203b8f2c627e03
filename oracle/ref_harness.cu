// oracle/ref_harness.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Drives the UNMODIFIED reference renderer (sources compiled where they lie under
// /root/reference, see oracle/Makefile) and dumps what the parity tests need:
//   * the LBVH arrays right after PathtraceInit   (reference: src/bvh.cu:456-516)
//   * per-bounce queues / counters / PathState     (reference loop: src/pathtrace_wavefront.cu:278-348)
//   * the direct / indirect accumulators           (radiance = (direct+indirect)/iter, :74-75)
// and times the reference's own wavefront / megakernel front-ends for the bench's
// reference arm.
//
// The reference keeps its state in TU-local statics (`pState`, `hst_scene`,
// pathtrace_wavefront.cu:118-120), so the only way to reach them without editing the
// reference is textual inclusion of that TU plus an accessor in the same namespace.
#include "pathtrace_wavefront.cu"   // resolved through -I$(REF)/src ; not a copy

#include "pathtrace.h"
#include <chrono>
#include <string>
#include <vector>
#include <sys/stat.h>

bool g_enableVisualization = false;   // extern in the reference (pathtrace_wavefront.cu:41); defined by the host app

namespace pathtrace_wavefront {
    WavefrontPathTracerState* OracleState() { return pState; }
}
namespace pathtrace_megakernel {   // ref_mega_access.cu
    glm::vec3* OracleImage();
    int* OracleSampleCount();
}

static std::string g_out;
static void dump_dev(const std::string& name, const void* dev, size_t bytes);

// sampleEnvironmentMap (src/interactions.cu:447-488) is defined in the reference but declared in no header and called by
// nothing; the build is -rdc, so a declaration here links against the reference's own definition.
__device__ float3 sampleEnvironmentMap(const EnvMapAliasTable env, cudaTextureObject_t* d_texture_objects, float2 rnd, float& pdf);
__global__ void SampleEnvKernel(EnvMapAliasTable env, cudaTextureObject_t* tex, const float2* rnd, int n, float3* dir, float* pdf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float p = 0.0f;
    float3 d = sampleEnvironmentMap(env, tex, rnd[i], p);
    dir[i] = d;
    pdf[i] = p;
}
// --sample-env RND.bin: the reference's own sampler over the caller's random pairs -> env_dir.bin / env_pdf.bin
static int sample_env(Scene* scene, const std::string& rnd_file) {
    if (scene->env_map.width <= 0) { fprintf(stderr, "refdump: scene has no environment map\n"); return 5; }
    FILE* f = fopen(rnd_file.c_str(), "rb");
    if (!f) { fprintf(stderr, "refdump: cannot read %s\n", rnd_file.c_str()); return 5; }
    fseek(f, 0, SEEK_END); long bytes = ftell(f); fseek(f, 0, SEEK_SET);
    const int n = (int)(bytes / sizeof(float2));
    std::vector<float2> rnd(n);
    if (fread(rnd.data(), sizeof(float2), n, f) != (size_t)n) { fclose(f); return 5; }
    fclose(f);
    const size_t N = (size_t)scene->env_map.width * scene->env_map.height;
    EnvMapAliasTable env;
    env.width = scene->env_map.width; env.height = scene->env_map.height;
    env.pdf_map_id = scene->env_map.pdf_map_id; env.env_tex_id = scene->env_map.env_tex_id;
    float* d_probs; int* d_aliases; cudaTextureObject_t* d_tex; float2* d_rnd; float3* d_dir; float* d_pdf;
    cudaMalloc(&d_probs, N * sizeof(float)); cudaMalloc(&d_aliases, N * sizeof(int));
    cudaMemcpy(d_probs, scene->env_map.probs.data(), N * sizeof(float), cudaMemcpyHostToDevice);     // the upload the reference's
    cudaMemcpy(d_aliases, scene->env_map.aliases.data(), N * sizeof(int), cudaMemcpyHostToDevice);   // initEnvAliasTable leaves out
    env.probs = d_probs; env.aliases = d_aliases;
    cudaMalloc(&d_tex, scene->texture_handles.size() * sizeof(cudaTextureObject_t));
    cudaMemcpy(d_tex, scene->texture_handles.data(), scene->texture_handles.size() * sizeof(cudaTextureObject_t), cudaMemcpyHostToDevice);
    cudaMalloc(&d_rnd, n * sizeof(float2)); cudaMalloc(&d_dir, n * sizeof(float3)); cudaMalloc(&d_pdf, n * sizeof(float));
    cudaMemcpy(d_rnd, rnd.data(), n * sizeof(float2), cudaMemcpyHostToDevice);
    SampleEnvKernel<<<(n + 255) / 256, 256>>>(env, d_tex, d_rnd, n, d_dir, d_pdf);
    cudaError_t e = cudaDeviceSynchronize();
    dump_dev("env_dir", d_dir, n * sizeof(float3));
    dump_dev("env_pdf", d_pdf, n * sizeof(float));
    printf("REFENV {\"samples\": %d, \"cuda\": \"%s\"}\n", n, cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 4;
}

// position-weighted word sum of a device array (the --load-only checksum): what a 10 M-triangle run compares instead of
// writing gigabytes of dumps
static unsigned long long sum_dev(const void* dev, size_t bytes) {
    std::vector<unsigned int> h(bytes / 4 ? bytes / 4 : 1);
    if (bytes) cudaMemcpy(h.data(), dev, bytes, cudaMemcpyDeviceToHost);
    unsigned long long s = 0;
    for (size_t i = 0; i < bytes / 4; i++) s += ((unsigned long long)h[i] + 1ull) * ((unsigned long long)i * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull);
    return s;
}

static void dump_raw(const std::string& name, const void* host, size_t bytes) {
    std::string p = g_out + "/" + name + ".bin";
    FILE* f = fopen(p.c_str(), "wb");
    if (!f) { fprintf(stderr, "refdump: cannot write %s\n", p.c_str()); exit(2); }
    if (bytes) fwrite(host, 1, bytes, f);
    fclose(f);
}

static void dump_dev(const std::string& name, const void* dev, size_t bytes) {
    std::vector<unsigned char> h(bytes ? bytes : 1);
    if (bytes) {
        cudaError_t e = cudaMemcpy(h.data(), dev, bytes, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { fprintf(stderr, "refdump: D2H %s: %s\n", name.c_str(), cudaGetErrorString(e)); exit(3); }
    }
    dump_raw(name, h.data(), bytes);
}

static int read_counter(const int* d) {
    int v = 0;
    cudaMemcpy(&v, d, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}

static void dump_scene_and_bvh(Scene* scene, WavefrontPathTracerState* st) {
    const int N = st->d_mesh_data.num_triangles;
    const int V = st->d_mesh_data.num_vertices;
    dump_dev("mesh_pos", st->d_mesh_data.pos, sizeof(float4) * V);
    dump_dev("mesh_nor", st->d_mesh_data.nor, sizeof(float4) * V);
    dump_dev("mesh_tangent", st->d_mesh_data.tangent, sizeof(float4) * V);
    dump_dev("mesh_uv", st->d_mesh_data.uv, sizeof(float2) * V);
    dump_dev("mesh_indices_matid", st->d_mesh_data.indices_matid, sizeof(int4) * N);
    dump_dev("mesh_nor_geom", st->d_mesh_data.nor_geom, sizeof(float4) * N);
    dump_raw("materials", scene->materials.data(), sizeof(Material) * scene->materials.size());
    dump_raw("light_tri_idx", scene->lightInfo.tri_idx.data(), sizeof(int) * scene->lightInfo.tri_idx.size());
    dump_raw("light_cdf", scene->lightInfo.cdf.data(), sizeof(float) * scene->lightInfo.cdf.size());
    if (scene->env_map.width > 0) {
        dump_raw("env_probs", scene->env_map.probs.data(), sizeof(float) * scene->env_map.probs.size());
        dump_raw("env_aliases", scene->env_map.aliases.data(), sizeof(int) * scene->env_map.aliases.size());
    }

    const LBVHData& b = st->d_bvh_data;
    dump_dev("bvh_aabb_min", b.aabb_min, sizeof(float4) * 2 * N);
    dump_dev("bvh_aabb_max", b.aabb_max, sizeof(float4) * 2 * N);
    dump_dev("bvh_centroid", b.centroid, sizeof(float4) * 2 * N);
    dump_dev("bvh_primitive_indices", b.primitive_indices, sizeof(int) * 2 * N);
    dump_dev("bvh_morton_codes", b.morton_codes, sizeof(unsigned long long) * N);
    dump_dev("bvh_child_nodes", b.child_nodes, sizeof(int2) * (N > 0 ? N - 1 : 0));
    dump_dev("bvh_parent", b.parent, sizeof(int) * 2 * N);
    dump_dev("bvh_escape_indices", b.escape_indices, sizeof(int) * 2 * N);
    float world[8] = { b.world_aabb_min.x, b.world_aabb_min.y, b.world_aabb_min.z, b.world_aabb_min.w,
                       b.world_aabb_max.x, b.world_aabb_max.y, b.world_aabb_max.z, b.world_aabb_max.w };
    dump_raw("bvh_world_aabb", world, sizeof(world));

    const Camera& c = scene->state.camera;
    float cam[17] = { c.position.x, c.position.y, c.position.z, c.view.x, c.view.y, c.view.z,
                      c.up.x, c.up.y, c.up.z, c.right.x, c.right.y, c.right.z,
                      c.fov.x, c.fov.y, c.pixelLength.x, c.pixelLength.y, 0.0f };
    dump_raw("camera", cam, sizeof(cam));

    std::string p = g_out + "/meta.json";
    FILE* f = fopen(p.c_str(), "w");
    fprintf(f, "{\"num_triangles\": %d, \"num_vertices\": %d, \"width\": %d, \"height\": %d, "
               "\"trace_depth\": %d, \"num_materials\": %d, \"num_lights\": %d, \"light_total_area\": %.9g, "
               "\"num_textures\": %d, \"env_tex_id\": %d, \"pdf_map_id\": %d, \"env_width\": %d, \"env_height\": %d}\n",
            N, V, c.resolution.x, c.resolution.y, scene->state.traceDepth, (int)scene->materials.size(),
            scene->lightInfo.num_lights, scene->lightInfo.total_area, (int)scene->texture_handles.size(),
            scene->env_map.env_tex_id, scene->env_map.pdf_map_id, scene->env_map.width, scene->env_map.height);
    fclose(f);
}

static void dump_path_state(const std::string& tag, WavefrontPathTracerState* st, int P) {
    const PathState& ps = st->d_path_state;
    dump_dev(tag + "_ray_ori", ps.ray_ori, sizeof(float4) * P);
    dump_dev(tag + "_ray_dir_dist", ps.ray_dir_dist, sizeof(float4) * P);
    dump_dev(tag + "_hit_normal", ps.hit_normal, sizeof(float4) * P);
    dump_dev(tag + "_throughput_pdf", ps.throughput_pdf, sizeof(float4) * P);
    dump_dev(tag + "_hit_geom_id", ps.hit_geom_id, sizeof(int) * P);
    dump_dev(tag + "_material_id", ps.material_id, sizeof(int) * P);
    dump_dev(tag + "_pixel_idx", ps.pixel_idx, sizeof(int) * P);
    dump_dev(tag + "_remaining_bounces", ps.remaining_bounces, sizeof(int) * P);
    dump_dev(tag + "_rng_state", ps.rng_state, sizeof(unsigned int) * P);
}

// One frame, driving the reference's own stage wrappers in the order of
// pathtrace_wavefront.cu:232-348 with dumps in between.
static void staged_frame(Scene* scene, WavefrontPathTracerState* st, int iter, bool dump) {
    using namespace pathtrace_wavefront;
    const int trace_depth = scene->state.traceDepth;
    const Camera& cam = scene->state.camera;
    const int P = cam.resolution.x * cam.resolution.y;
    if (iter == 1) {
        cudaMemset(st->d_direct_image, 0, P * sizeof(glm::vec3));
        cudaMemset(st->d_indirect_image, 0, P * sizeof(glm::vec3));
        st->clear();
        InitPathPoolKernel<<<(P + 127) / 128, 128>>>(st->d_path_state, P);
    }
    ray_gen::GeneratePathTracingRays(cam, trace_depth, iter, P, st);
    if (dump) dump_path_state("raygen", st, P);

    std::vector<int> counters;   // per step: n_ext, n_sh, n_pbr, n_diffuse, n_refl, n_refr
    for (int step = 0; step < trace_depth; step++) {
        char tag[32]; snprintf(tag, sizeof(tag), "s%d", step);
        std::string T(tag);
        int n_ext = read_counter(st->d_extension_ray_counter);
        int n_sh = read_counter(st->d_shadow_queue_counter);
        if (dump) {
            dump_dev(T + "_ext_queue", st->d_extension_ray_queue, sizeof(int) * n_ext);
            dump_dev(T + "_shadow_ori_tmax", st->d_shadow_queue.ray_ori_tmax, sizeof(float4) * n_sh);
            dump_dev(T + "_shadow_dir", st->d_shadow_queue.ray_dir, sizeof(float4) * n_sh);
            dump_dev(T + "_shadow_radiance", st->d_shadow_queue.radiance, sizeof(float4) * n_sh);
            dump_dev(T + "_shadow_pixel", st->d_shadow_queue.pixel_idx, sizeof(int) * n_sh);
        }
        if (n_ext > 0) TraceExtensionRay(n_ext, st);
        if (n_sh > 0) TraceShadowRay(n_sh, st);
        if (dump) {
            dump_path_state(T + "_traced", st, P);
            dump_dev(T + "_direct", st->d_direct_image, sizeof(glm::vec3) * P);
        }
        st->resetCounters();
        logic::RunPathLogic(P, trace_depth, st, make_float2(cam.resolution.x, cam.resolution.y));
        int n_pbr = read_counter(st->d_pbr_counter);
        int n_dif = read_counter(st->d_diffuse_counter);
        int n_rfl = read_counter(st->d_reflection_counter);
        int n_rfr = read_counter(st->d_refraction_counter);
        if (dump) {
            dump_dev(T + "_pbr_queue", st->d_pbr_queue, sizeof(int) * n_pbr);
            dump_dev(T + "_diffuse_queue", st->d_diffuse_queue, sizeof(int) * n_dif);
            dump_dev(T + "_reflection_queue", st->d_reflection_queue, sizeof(int) * n_rfl);
            dump_dev(T + "_refraction_queue", st->d_refraction_queue, sizeof(int) * n_rfr);
            dump_dev(T + "_indirect", st->d_indirect_image, sizeof(glm::vec3) * P);
            dump_dev(T + "_logic_remaining_bounces", st->d_path_state.remaining_bounces, sizeof(int) * P);
        }
        // The reference sorts the PBR and diffuse queues by material id (:312-334); the
        // order only changes which atomicAdd slot a path lands in, never a value.
        if (n_pbr > 0) {
            GenerateMaterialSortKeysKernel<<<(n_pbr + 127) / 128, 128>>>(st->d_pbr_queue, n_pbr, st->d_path_state, st->d_mat_sort_keys);
            thrust::device_ptr<int> k(st->d_mat_sort_keys), v(st->d_pbr_queue);
            thrust::sort_by_key(k, k + n_pbr, v);
            shading::SamplePBR(n_pbr, trace_depth, st);
        }
        if (n_dif > 0) {
            GenerateMaterialSortKeysKernel<<<(n_dif + 127) / 128, 128>>>(st->d_diffuse_queue, n_dif, st->d_path_state, st->d_mat_sort_keys);
            thrust::device_ptr<int> k(st->d_mat_sort_keys), v(st->d_diffuse_queue);
            thrust::sort_by_key(k, k + n_dif, v);
            shading::SampleDiffuse(n_dif, trace_depth, st);
        }
        if (n_rfl > 0) shading::SampleSpecularReflection(n_rfl, trace_depth, st);
        if (n_rfr > 0) shading::SampleSpecularRefraction(n_rfr, trace_depth, st);
        if (dump) dump_path_state(T + "_shaded", st, P);
        int c6[6] = { n_ext, n_sh, n_pbr, n_dif, n_rfl, n_rfr };
        counters.insert(counters.end(), c6, c6 + 6);
    }
    st->prev_view_proj_mat = st->view_proj_mat;
    if (dump) dump_raw("counters", counters.data(), sizeof(int) * counters.size());
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// --stbi FILE...: the reference's own image decoder (stb_image 2.06, external/include/stb_image.h, compiled by src/stb_image.cpp)
// exactly as Scene::loadTexture calls it (src/scene.cpp:139): size, channel count and a position-weighted sum of the pixel bytes
extern "C" unsigned char* stbi_load(char const* filename, int* x, int* y, int* comp, int req_comp);
extern "C" const char* stbi_failure_reason(void);
static int stbi_report(int argc, char** argv) {
    for (int i = 2; i < argc; i++) {
        int w = 0, h = 0, c = 0;
        unsigned char* px = stbi_load(argv[i], &w, &h, &c, 0);
        if (!px) { printf("REFSTBI {\"file\": \"%s\", \"ok\": false, \"reason\": \"%s\"}\n", argv[i], stbi_failure_reason() ? stbi_failure_reason() : ""); continue; }
        unsigned long long s = 0;
        const size_t n = (size_t)w * h * c;
        for (size_t k = 0; k < n; k++) s += ((unsigned long long)px[k] + 1ull) * ((unsigned long long)k * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull);
        printf("REFSTBI {\"file\": \"%s\", \"ok\": true, \"w\": %d, \"h\": %d, \"channels\": %d, \"sum\": \"%016llx\"}\n", argv[i], w, h, c, s);
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 3 && std::string(argv[1]) == "--stbi") return stbi_report(argc, argv);
    if (argc < 3) {
        fprintf(stderr,
            "usage: refdump SCENE.json OUTDIR [--iters N] [--stages] [--selfcheck] [--bench K W] [--mega] [--no-dump] [--views] [--denoise] [--denoise-bench K]\n");
        return 1;
    }
    std::string scene_file = argv[1];
    g_out = argv[2];
    int iters = 1, bench_k = 0, bench_w = 0;
    bool stages = false, selfcheck = false, mega = false, no_dump = false, views = false, denoise = false;
    int bench_d = 0;
    bool load_only = false, sums = false;
    std::string env_rnd;
    for (int i = 3; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--iters" && i + 1 < argc) iters = atoi(argv[++i]);
        else if (a == "--stages") stages = true;
        else if (a == "--selfcheck") selfcheck = true;
        else if (a == "--mega") mega = true;
        else if (a == "--no-dump") no_dump = true;
        else if (a == "--views") views = true;
        else if (a == "--denoise") denoise = true;
        else if (a == "--sample-env" && i + 1 < argc) env_rnd = argv[++i];   // the reference's sampleEnvironmentMap over the pairs in this file
        else if (a == "--sums") sums = true;                                 // print checksums of the LBVH / accumulators instead of dumping them
        else if (a == "--load-only") load_only = true;   // time the reference's scene loader (host only) and stop
        else if (a == "--denoise-bench" && i + 1 < argc) { denoise = true; bench_d = atoi(argv[++i]); }
        else if (a == "--bench" && i + 2 < argc) { bench_k = atoi(argv[++i]); bench_w = atoi(argv[++i]); }
    }
    mkdir(g_out.c_str(), 0777);
    cudaSetDevice(0);
    cudaFree(0);

    double t_load0 = now_s();
    Scene* scene = new Scene(scene_file);
    double t_load = now_s() - t_load0;
    if (load_only) {
        // position-weighted word sum (vectorisable on the checking side) over what the de-duplication decides:
        // vertex positions and uvs in id order, then the index triples, then the material ids
        unsigned long long h = 0, n = 0;
        auto mix = [&](const void* p, size_t words) {
            const unsigned int* w = (const unsigned int*)p;
            for (size_t i = 0; i < words; i++, n++) h += ((unsigned long long)w[i] + 1ull) * (n * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull);
        };
        for (const Vertex& v : scene->vertices) { mix(&v.pos, 3); mix(&v.uv, 2); }
        mix(scene->indices.data(), scene->indices.size());
        mix(scene->materialIds.data(), scene->materialIds.size());
        printf("REFLOAD {\"load_s\": %.4f, \"triangles\": %zu, \"vertices\": %zu, \"checksum\": \"%016llx\"}\n", t_load, scene->indices.size() / 3,
               scene->vertices.size(), h);
        return 0;
    }
    const Camera& cam = scene->state.camera;
    const int P = cam.resolution.x * cam.resolution.y;
    if (!env_rnd.empty()) return sample_env(scene, env_rnd);

    GuiDataContainer gui;
    gui.DenoiserOn = false;   // ctor default is true, which would switch the frame to SVGF 1-spp mode
    gui.ShowBVH = false;      // left uninitialised by the reference ctor (utilities.h:32-38)

    if (mega) {
        pathtrace_megakernel::InitDataContainer(&gui);
        pathtrace_megakernel::PathtraceInit(scene);
        if (bench_k == 0) {
            // --mega without --bench: the float image of the reference's megakernel after `iters` samples (src/pathtrace.cu:446-451)
            for (int i = 1; i <= iters; i++) pathtrace_megakernel::Pathtrace(NULL, 0, i);
            cudaDeviceSynchronize();
            dump_dev("mega_image", pathtrace_megakernel::OracleImage(), sizeof(glm::vec3) * P);
            dump_dev("mega_samples", pathtrace_megakernel::OracleSampleCount(), sizeof(int) * P);
            cudaError_t e = cudaGetLastError();
            printf("REFDUMP {\"mode\": \"mega\", \"iters\": %d, \"pixels\": %d, \"mrays_per_s\": %.3f, \"cuda\": \"%s\"}\n", iters, P, gui.MraysPerSec, cudaGetErrorString(e));
            pathtrace_megakernel::PathtraceFree();
            return e == cudaSuccess ? 0 : 4;
        }
        double mr = 0;
        for (int i = 1; i <= bench_w; i++) pathtrace_megakernel::Pathtrace(NULL, 0, i);
        cudaDeviceSynchronize();
        double t0 = now_s();
        for (int i = 1; i <= bench_k; i++) { pathtrace_megakernel::Pathtrace(NULL, 0, bench_w + i); mr += gui.MraysPerSec; }
        cudaDeviceSynchronize();
        double dt = now_s() - t0;
        printf("REFBENCH {\"mode\": \"mega\", \"frames\": %d, \"ms_per_frame\": %.6f, \"mpaths_per_s\": %.6f, \"mrays_per_s\": %.6f}\n",
               bench_k, dt * 1e3 / (bench_k > 0 ? bench_k : 1), bench_k * (double)P / dt / 1e6, mr / (bench_k > 0 ? bench_k : 1));
        pathtrace_megakernel::PathtraceFree();
        return 0;
    }

    pathtrace_wavefront::InitDataContainer(&gui);
    double t_init0 = now_s();
    pathtrace_wavefront::PathtraceInit(scene);
    cudaDeviceSynchronize();
    double t_init = now_s() - t_init0;
    WavefrontPathTracerState* st = pathtrace_wavefront::OracleState();
    if (!no_dump) dump_scene_and_bvh(scene, st);
    if (sums) {
        const size_t N = st->d_mesh_data.num_triangles, V = st->d_mesh_data.num_vertices;
        const LBVHData& b = st->d_bvh_data;
        printf("REFSUMS {\"num_triangles\": %zu, \"num_vertices\": %zu, \"load_s\": %.3f, \"init_s\": %.3f, "
               "\"pos\": \"%016llx\", \"indices_matid\": \"%016llx\", \"aabb_min\": \"%016llx\", \"aabb_max\": \"%016llx\", \"centroid\": \"%016llx\", "
               "\"primitive_indices\": \"%016llx\", \"morton_codes\": \"%016llx\", \"child_nodes\": \"%016llx\", \"parent\": \"%016llx\", \"escape_indices\": \"%016llx\"}\n",
               N, V, t_load, t_init, sum_dev(st->d_mesh_data.pos, V * 16), sum_dev(st->d_mesh_data.indices_matid, N * 16),
               sum_dev(b.aabb_min, 2 * N * 16), sum_dev(b.aabb_max, 2 * N * 16), sum_dev(b.centroid, 2 * N * 16), sum_dev(b.primitive_indices, 2 * N * 4),
               sum_dev(b.morton_codes, N * 8), sum_dev(b.child_nodes, (N - 1) * 8), sum_dev(b.parent, 2 * N * 4), sum_dev(b.escape_indices, 2 * N * 4));
        fflush(stdout);
    }

    if (bench_k > 0) {
        double mr = 0;
        for (int i = 1; i <= bench_w; i++) pathtrace_wavefront::Pathtrace(NULL, 0, i);
        cudaDeviceSynchronize();
        double t0 = now_s();
        for (int i = 1; i <= bench_k; i++) { pathtrace_wavefront::Pathtrace(NULL, 0, bench_w + i); mr += gui.MraysPerSec; }
        cudaDeviceSynchronize();
        double dt = now_s() - t0;
        printf("REFBENCH {\"mode\": \"wave\", \"frames\": %d, \"ms_per_frame\": %.6f, \"mpaths_per_s\": %.6f, \"mrays_per_s\": %.6f, \"init_s\": %.4f}\n",
               bench_k, dt * 1e3 / bench_k, bench_k * (double)P / dt / 1e6, mr / bench_k, t_init);
        pathtrace_wavefront::PathtraceFree();
        return 0;
    }

    if (denoise) {
        // the reference's DenoiserOn mode through its untouched front-end: 1-spp frames + SVGF into d_final_image
        gui.DenoiserOn = true;
        for (int i = 1; i <= iters; i++) pathtrace_wavefront::Pathtrace(NULL, 0, i);
        cudaDeviceSynchronize();
        dump_dev("denoised", st->d_final_image, sizeof(glm::vec3) * P);
        dump_dev("direct", st->d_direct_image, sizeof(glm::vec3) * P);
        dump_dev("indirect", st->d_indirect_image, sizeof(glm::vec3) * P);
        uchar4* pbo = nullptr;
        cudaMalloc(&pbo, sizeof(uchar4) * P);
        const dim3 bs(8, 8), grid((cam.resolution.x + 7) / 8, (cam.resolution.y + 7) / 8);
        pathtrace_wavefront::SendImageToPBOKernel<<<grid, bs>>>(pbo, cam.resolution, iters, 0, true, st->d_final_image,
            st->d_direct_image, st->d_indirect_image, st->d_normal_matid, st->d_depth, st->d_albedo, st->d_motion_vectors);
        dump_dev("denoised_view0", pbo, sizeof(uchar4) * P);
        cudaFree(pbo);
        if (bench_d > 0) {   // filter cost alone: the reference's frame with and without the denoiser
            double t[2];
            for (int on = 0; on < 2; on++) {
                gui.DenoiserOn = on != 0;
                for (int i = 1; i <= 3; i++) pathtrace_wavefront::Pathtrace(NULL, 0, iters + i);
                cudaDeviceSynchronize();
                double t0 = now_s();
                for (int i = 1; i <= bench_d; i++) pathtrace_wavefront::Pathtrace(NULL, 0, iters + 3 + i);
                cudaDeviceSynchronize();
                t[on] = (now_s() - t0) * 1e3 / bench_d;
            }
            printf("REFDENOISE {\"frames\": %d, \"ms_per_frame_plain\": %.4f, \"ms_per_frame_denoised\": %.4f, \"svgf_ms\": %.4f}\n", bench_d, t[0], t[1], t[1] - t[0]);
        }
        cudaError_t e = cudaGetLastError();
        printf("REFDUMP {\"iters\": %d, \"pixels\": %d, \"cuda\": \"%s\"}\n", iters, P, cudaGetErrorString(e));
        pathtrace_wavefront::PathtraceFree();
        return e == cudaSuccess ? 0 : 4;
    }

    if (selfcheck) {
        // harness-driven frames must reproduce the untouched Pathtrace() bit for bit
        for (int i = 1; i <= iters; i++) pathtrace_wavefront::Pathtrace(NULL, 0, i);
        cudaDeviceSynchronize();
        std::vector<float> d0(3 * (size_t)P), i0(3 * (size_t)P), d1(3 * (size_t)P), i1(3 * (size_t)P);
        cudaMemcpy(d0.data(), st->d_direct_image, sizeof(float) * 3 * P, cudaMemcpyDeviceToHost);
        cudaMemcpy(i0.data(), st->d_indirect_image, sizeof(float) * 3 * P, cudaMemcpyDeviceToHost);
        for (int i = 1; i <= iters; i++) staged_frame(scene, st, i, false);
        cudaDeviceSynchronize();
        cudaMemcpy(d1.data(), st->d_direct_image, sizeof(float) * 3 * P, cudaMemcpyDeviceToHost);
        cudaMemcpy(i1.data(), st->d_indirect_image, sizeof(float) * 3 * P, cudaMemcpyDeviceToHost);
        size_t bad = 0;
        for (size_t k = 0; k < 3 * (size_t)P; k++)
            if (memcmp(&d0[k], &d1[k], 4) || memcmp(&i0[k], &i1[k], 4)) bad++;
        printf("REFSELFCHECK {\"iters\": %d, \"mismatching_floats\": %zu}\n", iters, bad);
    }

    for (int i = 1; i <= iters; i++) staged_frame(scene, st, i, stages && i == iters);
    cudaDeviceSynchronize();
    dump_dev("direct", st->d_direct_image, sizeof(glm::vec3) * P);
    dump_dev("indirect", st->d_indirect_image, sizeof(glm::vec3) * P);
    if (views) {
        // first-bounce G-buffer of the last frame (written by the reference's Logic, src/kernels/logic.cu:66-73,119-138)
        dump_dev("gb_depth", st->d_depth, sizeof(float) * P);
        dump_dev("gb_normal_matid", st->d_normal_matid, sizeof(float4) * P);
        dump_dev("gb_albedo", st->d_albedo, sizeof(float4) * P);
        dump_dev("gb_motion", st->d_motion_vectors, sizeof(float2) * P);
        dump_raw("view_proj", &st->view_proj_mat, sizeof(glm::mat4));
        // the display modes of the reference's own SendImageToPBOKernel into a plain device buffer
        uchar4* pbo = nullptr;
        cudaMalloc(&pbo, sizeof(uchar4) * P);
        const dim3 bs(8, 8), grid((cam.resolution.x + 7) / 8, (cam.resolution.y + 7) / 8);
        for (int mode = 0; mode <= 4; mode++) {
            pathtrace_wavefront::SendImageToPBOKernel<<<grid, bs>>>(pbo, cam.resolution, iters, mode, false, st->d_final_image,
                st->d_direct_image, st->d_indirect_image, st->d_normal_matid, st->d_depth, st->d_albedo, st->d_motion_vectors);
            char name[16]; snprintf(name, sizeof(name), "view%d", mode);
            dump_dev(name, pbo, sizeof(uchar4) * P);
        }
        cudaFree(pbo);
        // BVH heat map: the reference's VisualizeLBVH into its final image buffer (src/pathtrace_wavefront.cu:244-256)
        VisualizeLBVH(st->d_final_image, cam.resolution.x, cam.resolution.y, cam, st->d_bvh_data, st->d_mesh_data.num_triangles);
        cudaDeviceSynchronize();
        dump_dev("bvh_heatmap", st->d_final_image, sizeof(glm::vec3) * P);
    }
    cudaError_t e = cudaGetLastError();
    printf("REFDUMP {\"iters\": %d, \"pixels\": %d, \"cuda\": \"%s\"}\n", iters, P, cudaGetErrorString(e));
    pathtrace_wavefront::PathtraceFree();
    return e == cudaSuccess ? 0 : 4;
}
