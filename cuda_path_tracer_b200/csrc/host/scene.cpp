// scene.cpp -- scene.json / OBJ / MTL ingestion with the semantics of the reference `Scene`
// (src/scene.cpp:45-631): same schema, same material-id assignment (sorted JSON keys, then one
// appended material block per object), same triangle order, same first-seen vertex
// de-duplication, same light CDF, same Vose alias table.  Triangle order and vertex positions
// feed the LBVH tie-break (src/bvh.cu:95), so they are part of the bit-exact contract.
// Build with -ffp-contract=off.
#include "scene.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "dedup_gpu.hpp"
#include "json_mini.hpp"
#include "obj_loader.hpp"
#include "parallel.hpp"

namespace ptb {

static const float kPI = 3.1415926535897932384626422832795028841971f;   // src/common/utilities.h:12

// ---- vertex de-duplication (src/scene.cpp:18-41, src/scene.h:26-35) -------------------------
// Equality is float == on (pos, nor, uv): -0 == +0 and NaN never matches, exactly like
// Vertex::operator==.  The hash only has to be consistent with that equality.
struct VKey { float v[8]; };
struct VKeyHash {
    size_t operator()(const VKey& k) const {
        uint64_t h = 1469598103934665603ull;
        for (int i = 0; i < 8; i++) {
            float f = k.v[i] == 0.0f ? 0.0f : k.v[i];   // fold -0 onto +0
            uint32_t b;
            memcpy(&b, &f, 4);
            h = (h ^ b) * 1099511628211ull;
        }
        return (size_t)h;
    }
};
struct VKeyEq {
    bool operator()(const VKey& a, const VKey& b) const {
        for (int i = 0; i < 8; i++) if (!(a.v[i] == b.v[i])) return false;
        return true;
    }
};

void derive_camera(ptb_camera& cam, float fovy) {   // src/scene.cpp:560-572
    float yscaled = tanf((fovy * 0.5f) * (kPI / 180.0f));
    float xscaled = (yscaled * cam.resolution[0]) / cam.resolution[1];
    float fovx = (atanf(xscaled) * 180.0f / kPI) * 2.0f;
    cam.fov[0] = fovx; cam.fov[1] = fovy;
    V3 pos(cam.position[0], cam.position[1], cam.position[2]);
    V3 look(cam.look_at[0], cam.look_at[1], cam.look_at[2]);
    V3 up(cam.up[0], cam.up[1], cam.up[2]);
    V3 view = normalize(look - pos);
    V3 right = normalize(cross(view, up));
    up = normalize(cross(right, view));
    cam.view[0] = view.x; cam.view[1] = view.y; cam.view[2] = view.z;
    cam.right[0] = right.x; cam.right[1] = right.y; cam.right[2] = right.z;
    cam.up[0] = up.x; cam.up[1] = up.y; cam.up[2] = up.z;
    cam.pixel_length[0] = 2 * xscaled / (float)cam.resolution[0];
    cam.pixel_length[1] = 2 * yscaled / (float)cam.resolution[1];
}

void build_light_cdf(ptb_scene& s) {   // src/scene.cpp:579-631
    s.light_total_area = 0.0f;
    s.light_tri_idx.clear();
    s.light_cdf.clear();
    const int n = (int)s.indices_matid.size();
    for (int i = 0; i < n; i++) {
        const i4& t = s.indices_matid[i];
        if (t.w < 0 || t.w >= (int)s.materials.size()) continue;
        if (s.materials[t.w].emittance > 0.0f) {
            V3 p0(s.pos[t.x].x, s.pos[t.x].y, s.pos[t.x].z);
            V3 p1(s.pos[t.y].x, s.pos[t.y].y, s.pos[t.y].z);
            V3 p2(s.pos[t.z].x, s.pos[t.z].y, s.pos[t.z].z);
            float area = 0.5f * length(cross(p1 - p0, p2 - p0));
            s.light_total_area += area;
            s.light_tri_idx.push_back(i);
            s.light_cdf.push_back(s.light_total_area);
        }
    }
    if (s.light_total_area > 0.0f) {
        for (float& v : s.light_cdf) v /= s.light_total_area;
        if (!s.light_cdf.empty()) s.light_cdf.back() = 1.0f;
    }
}

static int add_texture(ptb_scene& s, HostTexture&& t) {
    s.textures.push_back(std::move(t));
    return (int)s.textures.size() - 1;
}

void build_env_alias_table(ptb_scene& s, const float* data, int width, int height) {   // src/scene.cpp:194-299
    EnvMapHost& env = s.env;
    env.width = width; env.height = height;
    const int N = width * height;
    HostTexture envtex;
    envtex.width = width; envtex.height = height; envtex.channels = 4; envtex.is_float = true; envtex.clamp_y = true;
    envtex.f32.resize((size_t)N * 4);
    for (int i = 0; i < N; i++) {
        envtex.f32[(size_t)i * 4 + 0] = data[i * 3]; envtex.f32[(size_t)i * 4 + 1] = data[i * 3 + 1];
        envtex.f32[(size_t)i * 4 + 2] = data[i * 3 + 2]; envtex.f32[(size_t)i * 4 + 3] = 1.0f;
    }
    env.probs.assign(N, 0.0f);
    env.aliases.assign(N, 0);
    env.pdf_map.assign(N, 0.0f);
    std::vector<float> energy(N);
    float total = 0.0f;
    for (int v = 0; v < height; ++v) {
        float theta = (v + 0.5f) / (float)height * kPI;
        float sin_theta = sinf(theta);
        for (int u = 0; u < width; ++u) {
            int idx = v * width + u;
            float lum = 0.2126f * data[idx * 3] + 0.7152f * data[idx * 3 + 1] + 0.0722f * data[idx * 3 + 2];
            energy[idx] = std::max(lum, 0.0f) * sin_theta;
            total += energy[idx];
        }
    }
    env.total_sum = total;
    float avg = total / N;
    for (int i = 0; i < N; ++i) energy[i] /= avg;
    // Vose with LIFO work lists (src/scene.cpp:236-267)
    std::vector<int> small, large;
    small.reserve(N); large.reserve(N);
    for (int i = 0; i < N; ++i) (energy[i] < 1.0f ? small : large).push_back(i);
    while (!small.empty() && !large.empty()) {
        int sidx = small.back(); small.pop_back();
        int l = large.back(); large.pop_back();
        env.probs[sidx] = energy[sidx];
        env.aliases[sidx] = l;
        energy[l] = (energy[l] + energy[sidx]) - 1.0f;
        (energy[l] < 1.0f ? small : large).push_back(l);
    }
    while (!large.empty()) { int l = large.back(); large.pop_back(); env.probs[l] = 1.0f; env.aliases[l] = l; }
    while (!small.empty()) { int q = small.back(); small.pop_back(); env.probs[q] = 1.0f; env.aliases[q] = q; }
    float pdf_factor = (total > 0.0f) ? (N / (total * 2.0f * kPI * kPI)) : 0.0f;
    for (int idx = 0; idx < N; ++idx) {
        float lum = 0.2126f * data[idx * 3] + 0.7152f * data[idx * 3 + 1] + 0.0722f * data[idx * 3 + 2];
        env.pdf_map[idx] = std::max(lum, 1e-6f) * pdf_factor;
    }
    env.env_tex_id = add_texture(s, std::move(envtex));
    HostTexture pdftex;
    pdftex.width = width; pdftex.height = height; pdftex.channels = 1; pdftex.is_float = true; pdftex.clamp_y = true;
    pdftex.f32 = env.pdf_map;
    env.pdf_map_id = add_texture(s, std::move(pdftex));
}

static int load_texture(ptb_scene& s, const std::string& path) {   // src/scene.cpp:130-181
    if (path.empty()) return -1;
    auto it = s.texture_path_to_idx.find(path);
    if (it != s.texture_path_to_idx.end()) return it->second;
    int w, h, ch;
    std::vector<uint8_t> px;
    std::string err;
    if (!load_image_u8(path, w, h, ch, px, err)) {
        // a file that is there but that this loader cannot decode must not become a silently untextured render
        if (err.compare(0, 11, "unsupported") == 0) throw std::runtime_error("texture " + err);
        fprintf(stderr, "ptb200: could not load texture %s (%s)\n", path.c_str(), err.c_str());   // missing file: as the reference, id -1
        return -1;
    }
    HostTexture t;
    t.width = w; t.height = h; t.channels = 4; t.is_float = false; t.clamp_y = false;
    t.rgba8.resize((size_t)w * h * 4);
    for (size_t i = 0; i < (size_t)w * h; i++) {
        uint8_t* o = &t.rgba8[i * 4];
        if (ch == 1) { o[0] = o[1] = o[2] = px[i]; o[3] = 255; }
        else if (ch == 3) { o[0] = px[i * 3]; o[1] = px[i * 3 + 1]; o[2] = px[i * 3 + 2]; o[3] = 255; }
        else if (ch == 4) { memcpy(o, &px[i * 4], 4); }
        else { o[0] = 255; o[1] = 0; o[2] = 255; o[3] = 255; }
    }
    int id = add_texture(s, std::move(t));
    s.texture_path_to_idx[path] = id;
    return id;
}

static V3 json_vec3(const JsonValue& a) {
    return V3((float)a.at((size_t)0).number(), (float)a.at((size_t)1).number(), (float)a.at((size_t)2).number());
}

static void load_materials(const JsonValue& mats, ptb_scene& s, std::unordered_map<std::string, uint32_t>& name_to_id) {
    // src/scene.cpp:302-328; std::map iteration == nlohmann's sorted-key order
    for (const auto& kv : mats.obj) {
        const JsonValue& p = kv.second;
        ptb_material m;
        memset(&m, 0, sizeof(m));
        if (p.contains("basecolor")) {
            V3 c = json_vec3(p.at("basecolor"));
            m.basecolor[0] = c.x; m.basecolor[1] = c.y; m.basecolor[2] = c.z;
        }
        m.metallic = p.value_f("metallic", 0.0f);
        m.roughness = p.value_f("roughness", 0.5f);
        m.emittance = p.value_f("emittance", 0.0f);
        m.ior = p.value_f("ior", 1.0f);
        std::string type = p.value_s("Type", "MicrofacetPBR");
        if (type == "DIFFUSE") m.type = 1;
        else if (type == "SPECULAR_REFLECTION") m.type = 2;
        else if (type == "SPECULAR_REFRACTION") m.type = 3;
        else m.type = 0;
        m.diffuse_tex_id = m.normal_tex_id = m.metallic_roughness_tex_id = -1;
        name_to_id[kv.first] = (uint32_t)s.materials.size();
        s.materials.push_back(m);
    }
}

static ptb_material material_from_mtl(ptb_scene& s, const MtlMaterial& t, const std::string& base_dir) {
    // src/scene.cpp:385-465.  Material{} value-initialises, then the in-class initialisers set
    // the three texture ids to -1 (src/scene_structs.h:61-63).
    ptb_material m;
    memset(&m, 0, sizeof(m));
    m.diffuse_tex_id = m.normal_tex_id = m.metallic_roughness_tex_id = -1;
    m.basecolor[0] = t.diffuse[0]; m.basecolor[1] = t.diffuse[1]; m.basecolor[2] = t.diffuse[2];
    V3 emission(t.emission[0], t.emission[1], t.emission[2]);
    if (length(emission) > 0.001f) {
        m.emittance = length(emission);
        m.basecolor[0] = emission.x; m.basecolor[1] = emission.y; m.basecolor[2] = emission.z;
    }
    bool has_tex = false;
    if (!t.diffuse_texname.empty()) { m.diffuse_tex_id = load_texture(s, base_dir + t.diffuse_texname); if (m.diffuse_tex_id >= 0) has_tex = true; }
    if (!t.bump_texname.empty()) { m.normal_tex_id = load_texture(s, base_dir + t.bump_texname); if (m.normal_tex_id >= 0) has_tex = true; }
    if (!t.roughness_texname.empty()) { m.metallic_roughness_tex_id = load_texture(s, base_dir + t.roughness_texname); if (m.metallic_roughness_tex_id >= 0) has_tex = true; }
    if (t.shininess >= 0) m.roughness = 1.0f - std::min(1.0f, t.shininess / 1000.0f);
    else m.roughness = 0.5f;
    float spec_avg = (t.specular[0] + t.specular[1] + t.specular[2]) / 3.0f;
    m.metallic = (spec_avg > 0.1f) ? 1.0f : 0.0f;
    m.ior = t.ior;
    if (m.emittance > 0.0f) {
        m.type = 1;
    } else if (t.illum == 1) {
        m.type = 1;
    } else if (t.illum == 6 || t.illum == 7 || (t.ior > 1.01f && t.dissolve < 1.0f)) {
        m.type = 3;
        m.basecolor[0] = t.transmittance[0]; m.basecolor[1] = t.transmittance[1]; m.basecolor[2] = t.transmittance[2];
    } else if (has_tex) {
        m.type = 0;
    } else {
        if ((m.metallic > 0.9f && m.roughness < 0.02f) || t.illum == 3) {
            m.type = 2;
            m.basecolor[0] = t.specular[0]; m.basecolor[1] = t.specular[1]; m.basecolor[2] = t.specular[2];
            m.roughness = 0.0f;
            m.metallic = 1.0f;
        } else if (m.metallic < 0.1f && m.roughness > 0.8f) {
            m.type = 1;
        } else {
            m.type = 0;
        }
    }
    return m;
}

// ---- order-preserving parallel de-duplication ---------------------------------------------------------------------
// The reference numbers vertices by first occurrence while it walks the corners in file order through one
// unordered_map (src/scene.cpp:330-539) -- 30 M dependent hash probes at 10 M triangles.  The numbering only depends
// on, per corner, the earliest corner with an equal key.  Corners are bucketed by hash into shards (stable, so every
// shard still sees its corners in file order), each shard finds first occurrences in its own small table, and an
// exclusive scan over the "is a first occurrence" flags in corner order gives exactly the sequential ids.
namespace {

struct CornerStage {
    std::vector<VKey> key;          // 3 per triangle
    std::vector<uint64_t> hash;
    std::vector<V3> tangent;        // per triangle
    std::vector<int32_t> material;  // per triangle
};

void dedup_and_emit(CornerStage& st, ptb_scene& s, int threads) {
    const size_t C = st.key.size();
    const size_t T = C / 3;
    if (C == 0) return;
    if (C >= 0x7fffffffull) throw std::runtime_error("scene too large: more than 2^31 triangle corners");
    size_t S = 1;
    while (S < (size_t)threads * 2 && S < 256) S <<= 1;
    auto shard_of = [S](uint64_t h) { return (size_t)(h >> 40) & (S - 1); };

    // stable bucketing of corner indices by shard
    size_t R = std::max<size_t>(1, std::min<size_t>((size_t)threads, (C + 65535) / 65536));
    std::vector<size_t> count(R * S, 0);
    parallel_tasks(R, threads, [&](size_t r) {
        size_t* cnt = &count[r * S];
        for (size_t c = C * r / R; c < C * (r + 1) / R; c++) cnt[shard_of(st.hash[c])]++;
    });
    std::vector<size_t> shard_begin(S + 1, 0);
    {
        size_t run = 0;
        for (size_t sh = 0; sh < S; sh++) {
            shard_begin[sh] = run;
            for (size_t r = 0; r < R; r++) { size_t n = count[r * S + sh]; count[r * S + sh] = run; run += n; }
        }
        shard_begin[S] = run;
    }
    std::vector<uint32_t> order(C);
    parallel_tasks(R, threads, [&](size_t r) {
        size_t* pos = &count[r * S];
        for (size_t c = C * r / R; c < C * (r + 1) / R; c++) order[pos[shard_of(st.hash[c])]++] = (uint32_t)c;
    });

    // first occurrence per corner
    std::vector<uint32_t> first(C);
    VKeyEq eq;
    parallel_tasks(S, threads, [&](size_t sh) {
        const size_t n = shard_begin[sh + 1] - shard_begin[sh];
        if (n == 0) return;
        size_t cap = 16;
        while (cap < n * 2) cap <<= 1;
        std::vector<uint32_t> table(cap, 0xffffffffu);
        for (size_t k = shard_begin[sh]; k < shard_begin[sh + 1]; k++) {
            const uint32_t c = order[k];
            size_t slot = (size_t)st.hash[c] & (cap - 1);
            for (;;) {
                uint32_t o = table[slot];
                if (o == 0xffffffffu) { table[slot] = c; first[c] = c; break; }
                if (st.hash[o] == st.hash[c] && eq(st.key[o], st.key[c])) { first[c] = o; break; }
                slot = (slot + 1) & (cap - 1);
            }
        }
    });

    // ids = exclusive scan of the first-occurrence flags in corner order
    std::vector<uint32_t> id(C);
    std::vector<size_t> part(R + 1, 0);
    parallel_tasks(R, threads, [&](size_t r) {
        size_t n = 0;
        for (size_t c = C * r / R; c < C * (r + 1) / R; c++) n += (first[c] == c);
        part[r + 1] = n;
    });
    for (size_t r = 0; r < R; r++) part[r + 1] += part[r];
    const size_t base = s.pos.size();
    const size_t V = part[R];
    s.pos.resize(base + V);
    s.raw_nor.resize(base + V);
    s.raw_tangent.resize(base + V);
    s.uv.resize(base + V);
    parallel_tasks(R, threads, [&](size_t r) {
        size_t next = base + part[r];
        for (size_t c = C * r / R; c < C * (r + 1) / R; c++) {
            if (first[c] != c) continue;
            const float* k = st.key[c].v;
            id[c] = (uint32_t)next;
            s.pos[next] = f4{k[0], k[1], k[2], 1.0f};
            s.raw_nor[next] = V3(k[3], k[4], k[5]);
            s.raw_tangent[next] = st.tangent[c / 3];   // the tangent of the triangle that introduced the vertex
            s.uv[next] = f2{k[6], k[7]};
            next++;
        }
    });
    const size_t tri_base = s.indices_matid.size();
    s.indices_matid.resize(tri_base + T);
    parallel_ranges(T, threads, [&](size_t b, size_t e) {
        for (size_t t = b; t < e; t++)
            s.indices_matid[tri_base + t] = i4{(int32_t)id[first[3 * t]], (int32_t)id[first[3 * t + 1]], (int32_t)id[first[3 * t + 2]], st.material[t]};
    });
}

}  // namespace

// gpu_device >= 0: vertex de-duplication, numbering, normalisation and the repack into upload layout run on that GPU
// (csrc/cuda/dedup.cu); the arrays are the same bits either way (tests/test_gpu_loader.py).
static void load_objects(const JsonValue& objects, ptb_scene& s, const std::unordered_map<std::string, uint32_t>& name_to_id, int gpu_device) {
    // src/scene.cpp:330-539
    const int threads = loader_threads();
    const bool verbose = getenv("PTB_LOADER_VERBOSE") != nullptr;
    CornerStage st;
    VKeyHash hasher;
    for (const JsonValue& p : objects.arr) {
        bool forced = false;
        int forced_id = -1;
        std::string forced_name = "none";
        if (p.contains("MATERIAL")) forced_name = p.at("MATERIAL").string();
        if (forced_name != "none" && !forced_name.empty()) {
            auto it = name_to_id.find(forced_name);
            if (it != name_to_id.end()) { forced_id = (int)it->second; forced = true; }
            else fprintf(stderr, "ptb200: explicit material '%s' not found, falling back to MTL/default\n", forced_name.c_str());
        }
        V3 translation = json_vec3(p.at("TRANS"));
        V3 rotation = json_vec3(p.at("ROTAT"));
        V3 scl = json_vec3(p.at("SCALE"));
        // utilityCore::buildTransformationMatrix (src/common/utilities.cpp:85-93)
        M4 tm = translate(M4::identity(), translation);
        M4 rm = rotate(M4::identity(), rotation.x * (float)kPI / 180, V3(1, 0, 0));
        rm = rm * rotate(M4::identity(), rotation.y * (float)kPI / 180, V3(0, 1, 0));
        rm = rm * rotate(M4::identity(), rotation.z * (float)kPI / 180, V3(0, 0, 1));
        M4 sm = scale(M4::identity(), scl);
        M4 transform = tm * rm * sm;
        M4 inv_t = inverse_transpose(transform);

        if (!p.contains("FILE")) continue;
        std::string filename = p.at("FILE").string();
        std::string base_dir = filename.substr(0, filename.find_last_of("/\\") + 1);
        ObjData obj;
        std::string err;
        if (!load_obj(filename, base_dir, obj, err)) {
            fprintf(stderr, "ptb200: OBJ skipped (%s)\n", err.c_str());   // reference: `if (!ret) continue;` (:376)
            continue;
        }
        int mat_start = (int)s.materials.size();
        bool mtl_loaded = !obj.materials.empty();
        if (!forced && mtl_loaded) {
            for (const MtlMaterial& t : obj.materials) s.materials.push_back(material_from_mtl(s, t, base_dir));
        } else {
            ptb_material d;   // default grey PBR, appended even under an override (:467-475)
            memset(&d, 0, sizeof(d));
            d.diffuse_tex_id = d.normal_tex_id = d.metallic_roughness_tex_id = -1;
            d.basecolor[0] = d.basecolor[1] = d.basecolor[2] = 0.7f;
            d.roughness = 0.5f; d.metallic = 0.0f; d.type = 0;
            s.materials.push_back(d);
        }
        // per-triangle work has no order dependence: transform, tangent, geometric normal, corner keys
        const size_t ntri = obj.tri_material.size();
        const size_t t0 = st.tangent.size();
        st.key.resize((t0 + ntri) * 3);
        if (gpu_device < 0) st.hash.resize((t0 + ntri) * 3);   // the GPU path hashes on the device
        st.tangent.resize(t0 + ntri);
        st.material.resize(t0 + ntri);
        const size_t g0 = s.nor_geom.size();
        s.nor_geom.resize(g0 + ntri);
        parallel_ranges(ntri, threads, [&](size_t fb, size_t fe) {
            for (size_t f = fb; f < fe; f++) {
                V3 vpos[3], vnor[3];
                V2 vuv[3];
                for (int i = 0; i < 3; i++) {
                    const ObjIndex& ix = obj.tri_indices[f * 3 + i];
                    V4 rp(obj.v[3 * ix.v + 0], obj.v[3 * ix.v + 1], obj.v[3 * ix.v + 2], 1.0f);
                    V4 tp = transform * rp;
                    vpos[i] = V3(tp.x, tp.y, tp.z);
                    if (ix.vt >= 0) { vuv[i].x = obj.vt[2 * ix.vt + 0]; vuv[i].y = 1.0f - obj.vt[2 * ix.vt + 1]; }
                    if (ix.vn >= 0) {
                        V4 rn(obj.vn[3 * ix.vn + 0], obj.vn[3 * ix.vn + 1], obj.vn[3 * ix.vn + 2], 0.0f);
                        V4 tn = inv_t * rn;
                        vnor[i] = normalize(V3(tn.x, tn.y, tn.z));
                    }
                }
                V3 e1 = vpos[1] - vpos[0], e2 = vpos[2] - vpos[0];
                float du1x = vuv[1].x - vuv[0].x, du1y = vuv[1].y - vuv[0].y;
                float du2x = vuv[2].x - vuv[0].x, du2y = vuv[2].y - vuv[0].y;
                float det = du1x * du2y - du2x * du1y;
                V3 tangent(0, 0, 0);
                if (std::abs(det) > 1e-6f) {
                    float inv_det = 1.0f / det;
                    tangent = inv_det * (du2y * e1 - du1y * e2);
                }
                V3 ng = normalize(cross(e1, e2));
                s.nor_geom[g0 + f] = f4{ng.x, ng.y, ng.z, 0.0f};
                st.tangent[t0 + f] = tangent;
                st.material[t0 + f] = forced ? forced_id : (mat_start + std::max(0, obj.tri_material[f]));
                for (int i = 0; i < 3; i++) {
                    VKey k = {{vpos[i].x, vpos[i].y, vpos[i].z, vnor[i].x, vnor[i].y, vnor[i].z, vuv[i].x, vuv[i].y}};
                    st.key[(t0 + f) * 3 + i] = k;
                    if (gpu_device < 0) st.hash[(t0 + f) * 3 + i] = (uint64_t)hasher(k);
                }
            }
        });
    }
    const auto t_dedup = std::chrono::steady_clock::now();
    auto elapsed_ms = [&t_dedup]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_dedup).count(); };
    if (gpu_device >= 0) {
        static_assert(sizeof(VKey) == 32 && sizeof(V3) == 12, "corner keys and tangents are uploaded as they lie");
        GpuDedupTimes tm;
        if (!st.material.empty())
            gpu_dedup_pack(gpu_device, st.key[0].v, &st.tangent[0].x, st.material.data(), st.material.size(), s, &tm);
        if (verbose)
            fprintf(stderr, "ptb200: loader de-duplication + pack on GPU %d: %zu corners -> %zu vertices, %.1f ms (upload %.1f, device %.1f, download %.1f)\n",
                    gpu_device, tm.corners, tm.vertices, elapsed_ms(), tm.upload_ms, tm.device_ms, tm.download_ms);
        return;
    }
    dedup_and_emit(st, s, threads);
    // upload-time normalisation (src/wavefront_internal.cu:136-144): zero vectors become NaN, as there
    s.nor.resize(s.pos.size());
    s.tangent.resize(s.pos.size());
    parallel_ranges(s.pos.size(), threads, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; i++) {
            V3 n = normalize(s.raw_nor[i]);
            V3 t = normalize(s.raw_tangent[i]);
            s.nor[i] = f4{n.x, n.y, n.z, 0.0f};
            s.tangent[i] = f4{t.x, t.y, t.z, 0.0f};
        }
    });
    if (verbose)
        fprintf(stderr, "ptb200: loader de-duplication + pack on %d host threads: %zu corners -> %zu vertices, %.1f ms\n", threads, st.key.size(),
                s.pos.size(), elapsed_ms());
}

void load_scene_json(const std::string& path, ptb_scene& s, int gpu_device) {
    size_t dot = path.find_last_of('.');
    if (dot == std::string::npos || path.substr(dot) != ".json")
        throw std::runtime_error("Couldn't read from " + path + " (unknown extension)");   // src/scene.cpp:48-55
    std::ifstream f(path.c_str(), std::ios::binary);
    if (!f.is_open()) throw std::runtime_error("Could not open file " + path);            // :85-88
    std::stringstream ss;
    ss << f.rdbuf();
    std::string text = ss.str();
    if (text.size() >= 3 && (unsigned char)text[0] == 0xEF && (unsigned char)text[1] == 0xBB && (unsigned char)text[2] == 0xBF)
        text = text.substr(3);
    JsonValue data = JsonParser(text).parse();

    std::unordered_map<std::string, uint32_t> name_to_id;
    if (data.contains("Materials")) load_materials(data.at("Materials"), s, name_to_id);
    if (data.contains("Objects")) load_objects(data.at("Objects"), s, name_to_id, gpu_device);
    if (data.contains("Camera")) {   // src/scene.cpp:541-577
        const JsonValue& c = data.at("Camera");
        s.camera.resolution[0] = (int)c.at("RES").at((size_t)0).number();
        s.camera.resolution[1] = (int)c.at("RES").at((size_t)1).number();
        float fovy = (float)c.at("FOVY").number();
        s.iterations = (int)c.at("ITERATIONS").number();
        s.trace_depth = (int)c.at("DEPTH").number();
        s.image_name = c.at("FILE").string();
        V3 eye = json_vec3(c.at("EYE")), look = json_vec3(c.at("LOOKAT")), up = json_vec3(c.at("UP"));
        s.camera.position[0] = eye.x; s.camera.position[1] = eye.y; s.camera.position[2] = eye.z;
        s.camera.look_at[0] = look.x; s.camera.look_at[1] = look.y; s.camera.look_at[2] = look.z;
        s.camera.up[0] = up.x; s.camera.up[1] = up.y; s.camera.up[2] = up.z;
        derive_camera(s.camera, fovy);
    }
    build_light_cdf(s);
    if (data.contains("Environment Texture")) {   // :112-120
        const JsonValue& e = data.at("Environment Texture");
        if (e.contains("FILE")) {
            int w, h;
            std::vector<float> rgb;
            std::string err;
            if (load_image_hdr(e.at("FILE").string(), w, h, rgb, err)) build_env_alias_table(s, rgb.data(), w, h);
            else fprintf(stderr, "ptb200: failed to load HDR image (%s); environment disabled\n", err.c_str());   // :189-192
        }
    }
}

}  // namespace ptb
