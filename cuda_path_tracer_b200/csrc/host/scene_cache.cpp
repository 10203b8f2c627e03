// scene_cache.cpp -- binary scene cache.
// The reference re-parses OBJ / MTL / images and re-runs the de-duplication, the light CDF and the Vose alias table
// at every start (src/scene.cpp:83-631): 5 s per 2 M triangles.  Everything those steps produce is the ptb_scene
// container, already in upload layout, so the cache is that container written out verbatim: loading it is one read
// per array, and the arrays are bit-identical to a fresh parse by construction (tests/test_host.py).
// Layout: "PTBSCN01", u32 version, u32 byte-order mark, then every field in declaration order; a vector is a u64
// count followed by its elements; the file ends with "PTBSCEND".  Host byte order; not a portable interchange format.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "errors.hpp"
#include "scene.hpp"

using namespace ptb;

namespace {

const char kMagic[8] = {'P', 'T', 'B', 'S', 'C', 'N', '0', '1'};
const char kEnd[8] = {'P', 'T', 'B', 'S', 'C', 'E', 'N', 'D'};
const uint32_t kVersion = 1, kOrder = 0x01020304u;

struct Writer {
    FILE* f;
    void raw(const void* p, size_t n) {
        if (n && fwrite(p, 1, n, f) != n) throw std::runtime_error("write failed");
    }
    template <class T> void pod(const T& v) { static_assert(std::is_trivially_copyable<T>::value, "pod"); raw(&v, sizeof(T)); }
    template <class T> void vec(const std::vector<T>& v) {
        static_assert(std::is_trivially_copyable<T>::value, "pod vector");
        pod<uint64_t>(v.size());
        raw(v.data(), v.size() * sizeof(T));
    }
    void str(const std::string& s) { pod<uint64_t>(s.size()); raw(s.data(), s.size()); }
};

struct Reader {
    FILE* f;
    uint64_t remaining;
    void raw(void* p, size_t n) {
        if (n > remaining) throw std::runtime_error("truncated or corrupt cache");
        if (n && fread(p, 1, n, f) != n) throw std::runtime_error("truncated or corrupt cache");
        remaining -= n;
    }
    template <class T> void pod(T& v) { raw(&v, sizeof(T)); }
    template <class T> void vec(std::vector<T>& v) {
        uint64_t n = 0;
        pod(n);
        if (n > remaining / sizeof(T)) throw std::runtime_error("truncated or corrupt cache");   // before allocating
        v.resize((size_t)n);
        raw(v.data(), (size_t)n * sizeof(T));
    }
    void str(std::string& s) {
        uint64_t n = 0;
        pod(n);
        if (n > remaining) throw std::runtime_error("truncated or corrupt cache");
        s.resize((size_t)n);
        raw(&s[0], (size_t)n);
    }
};

void write_scene(Writer& w, const ptb_scene& s) {
    w.raw(kMagic, 8); w.pod(kVersion); w.pod(kOrder);
    w.vec(s.pos); w.vec(s.nor); w.vec(s.tangent); w.vec(s.uv); w.vec(s.indices_matid); w.vec(s.nor_geom);
    w.vec(s.raw_nor); w.vec(s.raw_tangent);
    w.vec(s.materials); w.vec(s.light_tri_idx); w.vec(s.light_cdf); w.pod(s.light_total_area);
    w.pod<uint64_t>(s.textures.size());
    for (const HostTexture& t : s.textures) {
        w.pod<int32_t>(t.width); w.pod<int32_t>(t.height); w.pod<int32_t>(t.channels);
        w.pod<uint8_t>(t.is_float); w.pod<uint8_t>(t.clamp_y);
        w.vec(t.rgba8); w.vec(t.f32);
    }
    w.pod<uint64_t>(s.texture_path_to_idx.size());
    for (const auto& kv : s.texture_path_to_idx) { w.str(kv.first); w.pod<int32_t>(kv.second); }
    w.pod<int32_t>(s.env.width); w.pod<int32_t>(s.env.height); w.pod<int32_t>(s.env.env_tex_id); w.pod<int32_t>(s.env.pdf_map_id);
    w.pod(s.env.total_sum); w.vec(s.env.probs); w.vec(s.env.aliases); w.vec(s.env.pdf_map);
    w.pod(s.camera); w.pod(s.trace_depth); w.pod(s.iterations); w.str(s.image_name);
    w.raw(kEnd, 8);
}

void read_scene(Reader& r, ptb_scene& s) {
    char magic[8];
    uint32_t version = 0, order = 0;
    r.raw(magic, 8);
    if (memcmp(magic, kMagic, 8) != 0) throw std::runtime_error("not a ptb200 scene cache");
    r.pod(version); r.pod(order);
    if (version != kVersion) throw std::runtime_error("scene cache version " + std::to_string(version) + ", this library reads " + std::to_string(kVersion));
    if (order != kOrder) throw std::runtime_error("scene cache written with another byte order");
    r.vec(s.pos); r.vec(s.nor); r.vec(s.tangent); r.vec(s.uv); r.vec(s.indices_matid); r.vec(s.nor_geom);
    r.vec(s.raw_nor); r.vec(s.raw_tangent);
    r.vec(s.materials); r.vec(s.light_tri_idx); r.vec(s.light_cdf); r.pod(s.light_total_area);
    uint64_t ntex = 0;
    r.pod(ntex);
    if (ntex > r.remaining / 11) throw std::runtime_error("truncated or corrupt cache");
    s.textures.resize((size_t)ntex);
    for (HostTexture& t : s.textures) {
        int32_t w = 0, h = 0, c = 0;
        uint8_t is_float = 0, clamp_y = 0;
        r.pod(w); r.pod(h); r.pod(c); r.pod(is_float); r.pod(clamp_y);
        t.width = w; t.height = h; t.channels = c; t.is_float = is_float != 0; t.clamp_y = clamp_y != 0;
        r.vec(t.rgba8); r.vec(t.f32);
        size_t px = (size_t)std::max(w, 0) * (size_t)std::max(h, 0);
        if (t.is_float ? t.f32.size() != px * (size_t)c : t.rgba8.size() != px * 4) throw std::runtime_error("corrupt cache: texture size");
    }
    uint64_t npaths = 0;
    r.pod(npaths);
    for (uint64_t i = 0; i < npaths; i++) {
        std::string k;
        int32_t v = 0;
        r.str(k); r.pod(v);
        s.texture_path_to_idx[k] = v;
    }
    int32_t ew = 0, eh = 0, et = 0, ep = 0;
    r.pod(ew); r.pod(eh); r.pod(et); r.pod(ep);
    s.env.width = ew; s.env.height = eh; s.env.env_tex_id = et; s.env.pdf_map_id = ep;
    r.pod(s.env.total_sum); r.vec(s.env.probs); r.vec(s.env.aliases); r.vec(s.env.pdf_map);
    r.pod(s.camera); r.pod(s.trace_depth); r.pod(s.iterations); r.str(s.image_name);
    char end[8];
    r.raw(end, 8);
    if (memcmp(end, kEnd, 8) != 0 || r.remaining != 0) throw std::runtime_error("truncated or corrupt cache");
    // the invariants ptb_init relies on
    const size_t V = s.pos.size(), T = s.indices_matid.size();
    if (s.nor.size() != V || s.tangent.size() != V || s.uv.size() != V || s.nor_geom.size() != T) throw std::runtime_error("corrupt cache: array sizes");
    for (const i4& t : s.indices_matid)
        if (t.x < 0 || t.y < 0 || t.z < 0 || (size_t)t.x >= V || (size_t)t.y >= V || (size_t)t.z >= V || t.w < 0 || (size_t)t.w >= s.materials.size())
            throw std::runtime_error("corrupt cache: index out of range");
    const int ntx = (int)s.textures.size();
    if (s.env.env_tex_id >= ntx || s.env.pdf_map_id >= ntx) throw std::runtime_error("corrupt cache: texture id out of range");
    const std::string why = validate_scene(s);
    if (!why.empty()) throw std::runtime_error("corrupt cache: " + why);
}

}  // namespace

extern "C" {

int ptb_scene_save_cache(const ptb_scene* s, const char* path) {
    if (!s || !path) return fail(PTB_ERR_INVALID, "ptb_scene_save_cache: null argument");
    std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(PTB_ERR_IO, std::string("ptb_scene_save_cache: cannot open ") + tmp);
    bool ok = true;
    std::string msg;
    try {
        Writer w{f};
        write_scene(w, *s);
    } catch (const std::exception& e) { ok = false; msg = e.what(); }
    if (fclose(f) != 0 && ok) { ok = false; msg = "close failed"; }
    if (ok && rename(tmp.c_str(), path) != 0) { ok = false; msg = "rename failed"; }   // readers never see a partial file
    if (!ok) { remove(tmp.c_str()); return fail(PTB_ERR_IO, "ptb_scene_save_cache: " + msg); }
    return PTB_OK;
}

int ptb_scene_load_cache(const char* path, ptb_scene** out) {
    if (!path || !out) return fail(PTB_ERR_INVALID, "ptb_scene_load_cache: null argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(PTB_ERR_IO, std::string("ptb_scene_load_cache: cannot open ") + path);
    fseek(f, 0, SEEK_END);
    long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    ptb_scene* s = new ptb_scene();
    try {
        Reader r{f, (uint64_t)std::max(size, 0L)};
        read_scene(r, *s);
    } catch (const std::exception& e) {
        fclose(f);
        delete s;
        return fail(PTB_ERR_PARSE, std::string("ptb_scene_load_cache: ") + e.what());
    }
    fclose(f);
    *out = s;
    return PTB_OK;
}

}  // extern "C"
