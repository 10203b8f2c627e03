// scene.hpp -- host-side scene container behind the opaque `ptb_scene` handle.
// Mirrors the public vectors of the reference `class Scene` (src/scene.h:73-92) but already in
// the SoA upload layout WavefrontPathTracerState::initSceneGeometry produces
// (src/wavefront_internal.cu:118-167), so ptb_init is a straight copy to the device.
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/ptb200.h"
#include "vecmath.hpp"

namespace ptb {

struct f4 { float x, y, z, w; };
struct f2 { float x, y; };
struct i4 { int32_t x, y, z, w; };

// A decoded image waiting for upload as a CUDA texture (src/scene.h:94-145).
struct HostTexture {
    int width = 0, height = 0;
    int channels = 4;                  // 4: RGBA8 normalized-float read; 1/4 with is_float: R32F / RGBA32F
    bool is_float = false;
    bool clamp_y = false;              // address mode Y (X is always wrap, src/scene.h:133-134)
    std::vector<uint8_t> rgba8;        // width*height*4 when !is_float
    std::vector<float> f32;            // width*height*channels when is_float
};

struct EnvMapHost {
    int width = 0, height = 0;
    int env_tex_id = -1, pdf_map_id = -1;
    float total_sum = 0.0f;
    std::vector<float> probs;
    std::vector<int32_t> aliases;
    std::vector<float> pdf_map;
};

}  // namespace ptb

struct ptb_scene {
    // geometry, upload layout
    std::vector<ptb::f4> pos, nor, tangent;
    std::vector<ptb::f2> uv;
    std::vector<ptb::i4> indices_matid;
    std::vector<ptb::f4> nor_geom;
    // the pre-normalisation attributes Scene keeps per Vertex (src/scene.h:26-35); needed for
    // de-duplication across objects and by from_arrays
    std::vector<ptb::V3> raw_nor, raw_tangent;

    std::vector<ptb_material> materials;
    std::vector<int32_t> light_tri_idx;
    std::vector<float> light_cdf;
    float light_total_area = 0.0f;

    std::vector<ptb::HostTexture> textures;
    std::unordered_map<std::string, int> texture_path_to_idx;
    ptb::EnvMapHost env;

    ptb_camera camera{};
    int32_t trace_depth = 8;
    int32_t iterations = 1;
    std::string image_name;
};

namespace ptb {

// Scene::loadFromJSON and friends (src/scene.cpp:83-631).  Throws std::runtime_error.  gpu_device >= 0 runs the vertex
// de-duplication and the repack into upload layout on that GPU (dedup_gpu.hpp); -1 keeps them on the host threads.
void load_scene_json(const std::string& path, ptb_scene& out, int gpu_device = -1);
// Scene::buildLightCDF (src/scene.cpp:579-631)
void build_light_cdf(ptb_scene& s);
// Scene::buildEnvMapAliasTable after the image is decoded (src/scene.cpp:194-299)
void build_env_alias_table(ptb_scene& s, const float* rgb, int width, int height);
// Scene::loadCamera's derived quantities (src/scene.cpp:560-572)
void derive_camera(ptb_camera& cam, float fovy_deg);

// What every kernel relies on and no kernel checks: material types and texture ids, light triangle indices, the environment's
// table sizes and alias range.  Returns an empty string when the scene is consistent, else what is wrong.  Called by
// ptb_scene_load_cache and ptb_init (a scene built with ptb_scene_from_arrays may still gain textures before ptb_init).
std::string validate_scene(const ptb_scene& s);

// image_io.cpp: decoders for the formats the procedural generators emit and stb_image 2.06
// also reads (PNM P5/P6, BMP 24/32, TGA 24/32 uncompressed, Radiance .hdr flat + RLE).
bool load_image_u8(const std::string& path, int& w, int& h, int& channels, std::vector<uint8_t>& px, std::string& err);
bool load_image_hdr(const std::string& path, int& w, int& h, std::vector<float>& rgb, std::string& err);
bool save_png_rgb8(const std::string& path, const uint8_t* rgb, int w, int h, std::string& err);
bool save_hdr_rgb32f(const std::string& path, const float* rgb, int w, int h, std::string& err);

}  // namespace ptb
