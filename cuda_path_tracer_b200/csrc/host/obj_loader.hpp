// obj_loader.hpp -- the OBJ/MTL subset the scene loader needs, restating the behaviour of
// tinyobjloader 2.0.0 (vendored by the reference, external/include/tiny_obj_loader.h) as it is
// used by Scene::loadObjects (src/scene.cpp:366-376): LoadObj with triangulate = true.
//   * numbers: tinyobj's own tryParseDouble (tiny_obj_loader.h:913-1050), NOT strtod -- the two
//     differ in the last bit for some decimal strings, and vertex positions are part of the
//     bit-exact LBVH contract.
//   * faces: file order is preserved across `g`/`o`/`usemtl`; triangles pass through, quads are
//     split along the shorter diagonal (tiny_obj_loader.h:1555-1663); polygons with more than
//     four vertices are rejected (tinyobj ear-clips them; the generators never emit them).
//   * per-face material id = index of the last `usemtl` name in the MTL, -1 if none/unknown.
#pragma once
#include <string>
#include <vector>

namespace ptb {

struct ObjIndex { int v = -1, vt = -1, vn = -1; };

struct MtlMaterial {   // the tinyobj::material_t members Scene::loadObjects reads (src/scene.cpp:385-465)
    std::string name;
    float diffuse[3] = {0, 0, 0};
    float specular[3] = {0, 0, 0};
    float transmittance[3] = {0, 0, 0};
    float emission[3] = {0, 0, 0};
    float shininess = 1.0f;
    float ior = 1.0f;
    float dissolve = 1.0f;
    int illum = 0;
    std::string diffuse_texname, bump_texname, roughness_texname;
};

struct ObjData {
    std::vector<float> v, vn, vt;           // 3, 3, 2 floats per record
    std::vector<ObjIndex> tri_indices;      // 3 per triangle, file order
    std::vector<int> tri_material;          // per triangle, tinyobj material id (-1 = none)
    std::vector<MtlMaterial> materials;     // in MTL file order
    std::string warn;
};

// Returns false and fills `err` on failure (the reference then skips the object, src/scene.cpp:376).
bool load_obj(const std::string& filename, const std::string& base_dir, ObjData& out, std::string& err);

// tinyobj's decimal parser, exposed for tests.
bool tinyobj_parse_double(const char* s, const char* s_end, double* result);

}  // namespace ptb
