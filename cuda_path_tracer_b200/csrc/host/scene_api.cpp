// scene_api.cpp -- the ptb_scene_* half of the C ABI (include/ptb200.h).  Pure host code.
// Replaces `class Scene` (reference src/scene.h:51-92, src/scene.cpp:45-631).
#include <vector>
#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

#include "errors.hpp"
#include "scene.hpp"

using namespace ptb;

namespace ptb {
thread_local std::string g_last_error;
int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
}  // namespace ptb

namespace ptb {
std::string validate_scene(const ptb_scene& s) {
    const int ntx = (int)s.textures.size();
    const size_t T = s.indices_matid.size();
    for (size_t i = 0; i < s.materials.size(); i++) {
        const ptb_material& m = s.materials[i];
        if (m.type < 0 || m.type > 3) return "material " + std::to_string(i) + ": type outside 0..3";
        const int ids[3] = {m.diffuse_tex_id, m.normal_tex_id, m.metallic_roughness_tex_id};
        for (int id : ids) {
            if (id < -1 || id >= ntx) return "material " + std::to_string(i) + ": texture id " + std::to_string(id) + " with " + std::to_string(ntx) + " textures";
            if (id >= 0 && s.textures[id].is_float) return "material " + std::to_string(i) + ": texture " + std::to_string(id) + " is a float (environment) texture";
        }
    }
    if (s.light_tri_idx.size() != s.light_cdf.size()) return "light triangle list and CDF differ in length";
    for (int32_t t : s.light_tri_idx)
        if (t < 0 || (size_t)t >= T) return "light triangle index out of range";
    if (s.env.width > 0 || s.env.env_tex_id >= 0) {
        const size_t n = (size_t)s.env.width * (size_t)s.env.height;
        if (s.env.width <= 0 || s.env.height <= 0 || s.env.probs.size() != n || s.env.aliases.size() != n) return "environment alias table size differs from width * height";
        if (s.env.env_tex_id < 0 || s.env.env_tex_id >= ntx || s.env.pdf_map_id < 0 || s.env.pdf_map_id >= ntx) return "environment texture id out of range";
        for (int32_t a : s.env.aliases)
            if (a < 0 || (size_t)a >= n) return "environment alias out of range";
    }
    return std::string();
}
}  // namespace ptb

extern "C" {

int ptb_abi_version(void) { return PTB_ABI_VERSION; }
const char* ptb_last_error(void) { return g_last_error.c_str(); }

static int scene_load(const char* fn, const char* json_path, int gpu_device, ptb_scene** out) {
    const std::string who(fn);
    if (!json_path || !out) return fail(PTB_ERR_INVALID, who + ": null argument");
    *out = nullptr;
    std::string path(json_path);
    // Only .json scenes are accepted (src/scene.cpp:48-55)
    if (path.size() < 5 || path.substr(path.size() - 5) != ".json") return fail(PTB_ERR_INVALID, who + ": not a .json scene: " + path);
    ptb_scene* s = new ptb_scene();
    try {
        load_scene_json(path, *s, gpu_device);
    } catch (const std::exception& e) {
        delete s;
        std::string msg = e.what();
        int code = msg.compare(0, 5, "cuda:") == 0                ? PTB_ERR_CUDA
                   : msg.find("unsupported") != std::string::npos ? PTB_ERR_UNSUPPORTED
                   : msg.find("open") != std::string::npos        ? PTB_ERR_IO
                                                                  : PTB_ERR_PARSE;
        return fail(code, who + ": " + msg);
    }
    *out = s;
    return PTB_OK;
}

int ptb_scene_load(const char* json_path, ptb_scene** out) {
    // PTB_LOADER_GPU=<ordinal> routes callers that cannot change their call (the reference-side shim) through the GPU loader
    int gpu_device = -1;
    if (const char* e = getenv("PTB_LOADER_GPU")) {
        char* end = nullptr;
        long v = strtol(e, &end, 10);
        if (end == e || *end || v < 0 || v > 1023) return fail(PTB_ERR_INVALID, std::string("ptb_scene_load: PTB_LOADER_GPU is not a device ordinal: ") + e);
        gpu_device = (int)v;
    }
    return scene_load("ptb_scene_load", json_path, gpu_device, out);
}

int ptb_scene_load_device(const char* json_path, int32_t device, ptb_scene** out) {
    if (device < 0) return fail(PTB_ERR_INVALID, "ptb_scene_load_device: negative device ordinal");
    return scene_load("ptb_scene_load_device", json_path, device, out);
}

int ptb_scene_from_arrays(const float* pos, const float* nor, const float* uv, const float* tangent, int32_t num_vertices,
                          const int32_t* indices, const int32_t* material_ids, int32_t num_triangles,
                          const ptb_material* materials, int32_t num_materials, const ptb_camera* camera,
                          int32_t trace_depth, ptb_scene** out) {
    if (!pos || !nor || !indices || !material_ids || !materials || !camera || !out)
        return fail(PTB_ERR_INVALID, "ptb_scene_from_arrays: null argument");
    if (num_vertices <= 0 || num_triangles <= 0 || num_materials <= 0) return fail(PTB_ERR_INVALID, "ptb_scene_from_arrays: empty scene");
    *out = nullptr;
    for (int64_t i = 0; i < 3 * (int64_t)num_triangles; i++)
        if (indices[i] < 0 || indices[i] >= num_vertices) return fail(PTB_ERR_INVALID, "ptb_scene_from_arrays: vertex index out of range");
    for (int32_t i = 0; i < num_triangles; i++)
        if (material_ids[i] < 0 || material_ids[i] >= num_materials) return fail(PTB_ERR_INVALID, "ptb_scene_from_arrays: material id out of range");
    for (int32_t i = 0; i < num_materials; i++)
        if (materials[i].type < 0 || materials[i].type > 3) return fail(PTB_ERR_INVALID, "ptb_scene_from_arrays: material type outside 0..3");
    ptb_scene* s = new ptb_scene();
    s->pos.resize(num_vertices); s->nor.resize(num_vertices); s->tangent.resize(num_vertices); s->uv.resize(num_vertices);
    for (int32_t i = 0; i < num_vertices; i++) {
        s->pos[i] = f4{pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], 1.0f};
        // upload-time normalisation, src/wavefront_internal.cu:136-144
        V3 n = normalize(V3(nor[3 * i], nor[3 * i + 1], nor[3 * i + 2]));
        s->nor[i] = f4{n.x, n.y, n.z, 0.0f};
        V3 t = tangent ? normalize(V3(tangent[3 * i], tangent[3 * i + 1], tangent[3 * i + 2])) : normalize(V3(0, 0, 0));
        s->tangent[i] = f4{t.x, t.y, t.z, 0.0f};
        s->uv[i] = uv ? f2{uv[2 * i], uv[2 * i + 1]} : f2{0.0f, 0.0f};
    }
    s->indices_matid.resize(num_triangles); s->nor_geom.resize(num_triangles);
    for (int32_t i = 0; i < num_triangles; i++) {
        int a = indices[3 * i], b = indices[3 * i + 1], c = indices[3 * i + 2];
        s->indices_matid[i] = i4{a, b, c, material_ids[i]};
        V3 p0(pos[3 * a], pos[3 * a + 1], pos[3 * a + 2]), p1(pos[3 * b], pos[3 * b + 1], pos[3 * b + 2]), p2(pos[3 * c], pos[3 * c + 1], pos[3 * c + 2]);
        V3 ng = normalize(cross(p1 - p0, p2 - p0));   // src/scene.cpp:518-519
        s->nor_geom[i] = f4{ng.x, ng.y, ng.z, 0.0f};
    }
    s->materials.assign(materials, materials + num_materials);
    s->camera = *camera;
    s->trace_depth = trace_depth > 0 ? trace_depth : 8;
    build_light_cdf(*s);
    *out = s;
    return PTB_OK;
}

int ptb_scene_set_environment(ptb_scene* s, const float* rgb, int32_t width, int32_t height) {
    if (!s || !rgb || width <= 0 || height <= 0) return fail(PTB_ERR_INVALID, "ptb_scene_set_environment: bad argument");
    if (s->env.width > 0) return fail(PTB_ERR_STATE, "ptb_scene_set_environment: environment already set");
    build_env_alias_table(*s, rgb, width, height);
    return PTB_OK;
}

int ptb_scene_add_texture(ptb_scene* s, const void* pixels, int32_t width, int32_t height, int32_t format, int32_t clamp_y) {
    if (!s || !pixels || width <= 0 || height <= 0) return fail(PTB_ERR_INVALID, "ptb_scene_add_texture: bad argument");
    if (format < PTB_TEX_RGBA8 || format > PTB_TEX_RGBA32F) return fail(PTB_ERR_INVALID, "ptb_scene_add_texture: unknown format");
    HostTexture t;
    t.width = width; t.height = height; t.clamp_y = clamp_y != 0;
    const size_t n = (size_t)width * height;
    if (format == PTB_TEX_RGBA8) {
        t.channels = 4; t.is_float = false;
        t.rgba8.assign((const uint8_t*)pixels, (const uint8_t*)pixels + 4 * n);
    } else {
        t.channels = format == PTB_TEX_R32F ? 1 : 4; t.is_float = true;
        t.f32.assign((const float*)pixels, (const float*)pixels + t.channels * n);
    }
    s->textures.push_back(std::move(t));
    return (int)s->textures.size() - 1;
}

int64_t ptb_scene_get_texture(const ptb_scene* s, int32_t id, void* dst, size_t bytes, int32_t* width, int32_t* height, int32_t* format) {
    if (!s) return fail(PTB_ERR_INVALID, "ptb_scene_get_texture: null scene");
    if (id < 0 || id >= (int32_t)s->textures.size()) return fail(PTB_ERR_INVALID, "ptb_scene_get_texture: texture id out of range");
    const HostTexture& t = s->textures[id];
    if (width) *width = t.width;
    if (height) *height = t.height;
    if (format) *format = !t.is_float ? PTB_TEX_RGBA8 : t.channels == 1 ? PTB_TEX_R32F : PTB_TEX_RGBA32F;
    const void* src = t.is_float ? (const void*)t.f32.data() : (const void*)t.rgba8.data();
    const size_t size = t.is_float ? t.f32.size() * sizeof(float) : t.rgba8.size();
    if (dst && bytes) memcpy(dst, src, bytes < size ? bytes : size);
    return (int64_t)size;
}

int ptb_scene_set_environment_table(ptb_scene* s, int32_t env_tex_id, int32_t pdf_map_id, const float* probs, const int32_t* aliases,
                                    int32_t width, int32_t height) {
    if (!s || !probs || !aliases || width <= 0 || height <= 0) return fail(PTB_ERR_INVALID, "ptb_scene_set_environment_table: bad argument");
    if (s->env.width > 0) return fail(PTB_ERR_STATE, "ptb_scene_set_environment_table: environment already set");
    const int nt = (int)s->textures.size();
    if (env_tex_id < 0 || env_tex_id >= nt || pdf_map_id < 0 || pdf_map_id >= nt)
        return fail(PTB_ERR_INVALID, "ptb_scene_set_environment_table: texture id out of range (add the textures first)");
    const HostTexture& et = s->textures[env_tex_id];
    const HostTexture& pt = s->textures[pdf_map_id];
    if (!et.is_float || et.channels != 4 || !pt.is_float || pt.channels != 1 || et.width != width || et.height != height || pt.width != width || pt.height != height)
        return fail(PTB_ERR_INVALID, "ptb_scene_set_environment_table: the environment must be an RGBA32F and its pdf map an R32F texture of the table's size");
    const size_t n = (size_t)width * height;
    for (size_t i = 0; i < n; i++)
        if (aliases[i] < 0 || (size_t)aliases[i] >= n) return fail(PTB_ERR_INVALID, "ptb_scene_set_environment_table: alias out of range");
    s->env.width = width; s->env.height = height;
    s->env.env_tex_id = env_tex_id; s->env.pdf_map_id = pdf_map_id;
    s->env.probs.assign(probs, probs + n);
    s->env.aliases.assign(aliases, aliases + n);
    s->env.pdf_map = pt.f32;
    return PTB_OK;
}

int ptb_scene_set_camera(ptb_scene* s, const int32_t resolution[2], float fovy_deg, const float eye[3], const float look_at[3],
                         const float up[3], int32_t trace_depth) {
    if (!s || !resolution || !eye || !look_at || !up) return fail(PTB_ERR_INVALID, "ptb_scene_set_camera: null argument");
    if (resolution[0] <= 0 || resolution[1] <= 0) return fail(PTB_ERR_INVALID, "ptb_scene_set_camera: bad resolution");
    s->camera.resolution[0] = resolution[0]; s->camera.resolution[1] = resolution[1];
    for (int i = 0; i < 3; i++) { s->camera.position[i] = eye[i]; s->camera.look_at[i] = look_at[i]; s->camera.up[i] = up[i]; }
    derive_camera(s->camera, fovy_deg);
    if (trace_depth > 0) s->trace_depth = trace_depth;
    return PTB_OK;
}

void ptb_scene_free(ptb_scene* s) { delete s; }

int ptb_scene_get_info(const ptb_scene* s, ptb_scene_info* out) {
    if (!s || !out) return fail(PTB_ERR_INVALID, "ptb_scene_get_info: null argument");
    memset(out, 0, sizeof(*out));
    out->num_vertices = (int32_t)s->pos.size();
    out->num_triangles = (int32_t)s->indices_matid.size();
    out->num_materials = (int32_t)s->materials.size();
    out->num_lights = (int32_t)s->light_tri_idx.size();
    out->num_textures = (int32_t)s->textures.size();
    out->width = s->camera.resolution[0];
    out->height = s->camera.resolution[1];
    out->trace_depth = s->trace_depth;
    out->iterations = s->iterations;
    out->env_width = s->env.width;
    out->env_height = s->env.height;
    out->env_tex_id = s->env.env_tex_id;
    out->pdf_map_id = s->env.pdf_map_id;
    out->light_total_area = s->light_total_area;
    return PTB_OK;
}

int ptb_save_image(const char* path, const float* rgb, int32_t width, int32_t height, float scale, int32_t flags) {
    if (!path || !rgb) return fail(PTB_ERR_INVALID, "ptb_save_image: null argument");
    if (width <= 0 || height <= 0) return fail(PTB_ERR_INVALID, "ptb_save_image: empty image");
    std::string p(path);
    const bool png = p.size() > 4 && p.compare(p.size() - 4, 4, ".png") == 0;
    const bool hdr = p.size() > 4 && p.compare(p.size() - 4, 4, ".hdr") == 0;
    if (!png && !hdr) return fail(PTB_ERR_INVALID, "ptb_save_image: path must end in .png or .hdr");
    const size_t n = (size_t)width * height;
    std::vector<float> img(3 * n);
    const float inv_gamma = 1.0f / 2.2f;
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            const float* s = rgb + 3 * ((size_t)y * width + x);
            int dx = (flags & PTB_SAVE_FLIP_X) ? width - 1 - x : x;        // img.setPixel(width - 1 - x, y, ...), src/main.cpp:259
            float* d = &img[3 * ((size_t)y * width + dx)];
            for (int c = 0; c < 3; c++) {
                float v = s[c] * scale;                                      // renderState->image[index] / samples, :254
                if (flags & PTB_SAVE_GAMMA) v = std::pow(std::max(0.0f, v), inv_gamma);   // :256-258
                d[c] = v;
            }
        }
    std::string err;
    bool ok;
    if (hdr) {
        ok = ptb::save_hdr_rgb32f(p, img.data(), width, height, err);
    } else {
        std::vector<uint8_t> bytes(3 * n);                                   // Image::savePNG, src/image.cpp:22-43
        for (size_t i = 0; i < 3 * n; i++) bytes[i] = (uint8_t)(std::min(std::max(img[i], 0.0f), 1.0f) * 255.0f);
        ok = ptb::save_png_rgb8(p, bytes.data(), width, height, err);
    }
    return ok ? PTB_OK : fail(PTB_ERR_IO, "ptb_save_image: " + err);
}

int ptb_scene_get_camera(const ptb_scene* s, ptb_camera* out) {
    if (!s || !out) return fail(PTB_ERR_INVALID, "ptb_scene_get_camera: null argument");
    *out = s->camera;
    return PTB_OK;
}

int64_t ptb_scene_get_array(const ptb_scene* s, int which, void* dst, size_t bytes) {
    if (!s) return fail(PTB_ERR_INVALID, "ptb_scene_get_array: null scene");
    const void* src = nullptr;
    size_t size = 0;
#define VIEW(v) do { src = (v).data(); size = (v).size() * sizeof((v)[0]); } while (0)
    switch (which) {
        case PTB_SA_POS: VIEW(s->pos); break;
        case PTB_SA_NOR: VIEW(s->nor); break;
        case PTB_SA_TANGENT: VIEW(s->tangent); break;
        case PTB_SA_UV: VIEW(s->uv); break;
        case PTB_SA_INDICES_MATID: VIEW(s->indices_matid); break;
        case PTB_SA_NOR_GEOM: VIEW(s->nor_geom); break;
        case PTB_SA_MATERIALS: VIEW(s->materials); break;
        case PTB_SA_LIGHT_TRI_IDX: VIEW(s->light_tri_idx); break;
        case PTB_SA_LIGHT_CDF: VIEW(s->light_cdf); break;
        case PTB_SA_ENV_PROBS: VIEW(s->env.probs); break;
        case PTB_SA_ENV_ALIASES: VIEW(s->env.aliases); break;
        case PTB_SA_ENV_PDF: VIEW(s->env.pdf_map); break;
        default: return fail(PTB_ERR_INVALID, "ptb_scene_get_array: unknown array id");
    }
#undef VIEW
    if (dst && bytes) memcpy(dst, src, bytes < size ? bytes : size);
    return (int64_t)size;
}

}  // extern "C"
