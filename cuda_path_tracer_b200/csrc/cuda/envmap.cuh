// envmap.cuh -- importance sampling of the HDR environment (sampleEnvironmentMap, src/interactions.cu:447-488).
// The reference builds the Vose alias table and the solid-angle pdf map on the host (src/scene.cpp:183-300) and its
// Logic stage already weights environment hits of BSDF-sampled rays by pdf_bsdf^2 / (pdf_bsdf^2 + pdf_env^2 + eps)
// (src/kernels/logic.cu:88-100) -- the BSDF half of a two-strategy MIS estimator -- but nothing calls the sampler
// and the wavefront state never uploads the table (src/wavefront_internal.cu:206-214), so the other half is missing
// and environment-lit surfaces come out too dark.  ptb_config.env_nee adds that half; it changes the image, hence
// the flag (off = the reference's result, bit for bit).
#pragma once
#include "common.cuh"

namespace ptb {

struct EnvDev {                 // EnvMapAliasTable, src/scene_structs.h:67-74
    const float* probs;
    const int* aliases;
    int width, height;
    int nee;                    // ptb_config.env_nee and an environment is present
};

// Two uniform numbers -> (direction, solid-angle pdf, texel).  O(1): pick a column of the alias table, keep it or
// take its alias.  The direction is the CENTRE of the chosen texel (the reference does not jitter inside it), with
// the same (u, v) <-> direction convention as Logic's lookup: phi = atan2(z, x), theta = acos(y).
__device__ __forceinline__ Vec3 sample_environment(const EnvDev& env, const TexturesDev& tex, float r0, float r1, float& pdf, int& pixel,
                                                   float& u_coord, float& v_coord) {
    (void)r1;   // the reference takes a float2 and uses only .x: the fractional part of r0 * N decides keep / alias
    const int N = env.width * env.height;
    float u = __fmul_rn(r0, (float)N);   // rounded on its own: `u - idx` below must not become fma(r0, N, -idx)
    int idx = min((int)u, N - 1);
    float xi = u - idx;
    float prob = __ldg(&env.probs[idx]);
    pixel = (xi < prob) ? idx : __ldg(&env.aliases[idx]);
    int py = pixel / env.width;
    int px = pixel % env.width;
    u_coord = (px + 0.5f) / env.width;
    v_coord = (py + 0.5f) / env.height;
    pdf = tex2D<float>(__ldg(&tex.handles[tex.pdf_map_id]), u_coord, v_coord);
    float phi = u_coord * 2.0f * PTB_PI;
    float theta = v_coord * PTB_PI;
    Vec3 dir;
    dir.x = sinf(theta) * cosf(phi);
    dir.y = cosf(theta);
    dir.z = sinf(theta) * sinf(phi);
    return dir;
}

}  // namespace ptb
