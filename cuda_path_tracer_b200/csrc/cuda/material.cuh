// material.cuh -- the 48-byte material record and the MIS weight shared by the Logic and Shading stages.
#pragma once
#include "common.cuh"

namespace ptb {

struct Mat {          // `Material`, src/scene_structs.h:52-65
    Vec3 basecolor;
    float metallic, roughness, emittance, ior;
    int type, diffuse_tex, normal_tex, mr_tex;
};

static __device__ Mat load_material(const ptb_material* __restrict__ mats, int id) {
    const float4* p = reinterpret_cast<const float4*>(mats + id);
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    Mat m;
    m.basecolor = Vec3(a.x, a.y, a.z);
    m.metallic = a.w;
    m.roughness = b.x; m.emittance = b.y; m.ior = b.z; m.type = __float_as_int(b.w);
    m.diffuse_tex = __float_as_int(c.x); m.normal_tex = __float_as_int(c.y); m.mr_tex = __float_as_int(c.z);
    return m;
}

// PowerHeuristic, src/common/cuda_utilities.h:193-197
static __device__ float power_heuristic(float f, float g) {
    float f2 = f * f;
    float g2 = g * g;
    return f2 / (f2 + g2 + 1e-5f);
}

}  // namespace ptb
