// stages.cu -- RayGen, Logic, Shading and resolve kernels of the wavefront core.
// Reference: src/kernels/ray_gen.cu:15-79, src/kernels/logic.cu:31-192, src/kernels/shading.cu:23-259,
// src/common/cuda_utilities.h:233-326, src/pathtrace_wavefront.cu:47-116.
//
// Differences from the reference that do not change a single sample:
//   * queue lengths stay on the device (counter rows, render.cuh); every kernel is launched with a
//     fixed grid and strides over `count` read from device memory, so a whole frame is one CUDA graph;
//   * Logic walks the extension queue of the bounce instead of all P slots (the live slots of a
//     bounce are exactly the paths that were traced, src/common/cuda_utilities.h:253-275);
//   * the four material kernels are ONE launch: each material queue is cut into 32-path chunks so
//     a warp always shades a single material type (type-uniform control flow, no divergence
//     between models);
//   * queue appends are warp-aggregated at a converged point (one atomic per warp per queue);
//   * each path owns its pixel, so the emission / environment adds never contend (they stay
//     atomicAdd = fire-and-forget RED, and stay un-fused with the product that feeds them).
#include "bsdf.cuh"
#include "logic.cuh"
#include "render.cuh"

namespace ptb {

static constexpr int kThreads = 128;

__device__ __forceinline__ Vec3 cam_v3(const float* p) { return Vec3(p[0], p[1], p[2]); }

// ---- frame begin / RayGen ----------------------------------------------------------------------
__global__ void k_clear(float* a, float* b, size_t n) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t stride = gridDim.x * (size_t)blockDim.x;
    for (; i < n; i += stride) { a[i] = 0.0f; b[i] = 0.0f; }
}

__global__ void __launch_bounds__(kThreads) k_raygen(FrameDev f) {
    const int iter = *f.iteration;
    const int qpos = blockIdx.x * blockDim.x + threadIdx.x;
    // counter rows: everything zero except the first extension queue, which holds every path
    int nctr = (f.trace_depth + 1) * CTR_STRIDE;
    if (qpos < nctr) f.ctr[qpos] = (qpos == CTR_EXT) ? f.num_paths : 0;
    if (qpos >= f.num_paths) return;
    const int slot = tiled_slot(qpos, f.tile.width, f.num_paths / f.tile.width);
    int pixel_idx = f.tile.global_pixel(slot);
    int x = pixel_idx % f.cam.res_x;
    int y = pixel_idx / f.cam.res_x;
    // wang_hash((pixel_idx * 19990303) + iter * 719393) in wrapping int arithmetic, :41
    uint32_t seed = wang_hash((uint32_t)pixel_idx * 19990303u + (uint32_t)iter * 719393u);
    if (seed == 0) seed = 1;
    seed = wang_hash(seed);
    float jx = (float)seed / 4294967296.0f - 0.5f;
    seed = wang_hash(seed);
    float jy = (float)seed / 4294967296.0f - 0.5f;
    Vec3 view = cam_v3(f.cam.view), right = cam_v3(f.cam.right), up = cam_v3(f.cam.up);
    Vec3 dir = normalize3(view
        + right * f.cam.pixel_length[0] * ((float)x + jx - (float)f.cam.res_x * 0.5f)
        - up * f.cam.pixel_length[1] * ((float)y + jy - (float)f.cam.res_y * 0.5f));
    stg256(f.pool.ray_record(slot), make_float4(f.cam.pos[0], f.cam.pos[1], f.cam.pos[2], 0.0f), make_float4(dir.x, dir.y, dir.z, FLT_MAX));
    // the reference also resets hit_geom_id / material_id to -1 here (src/kernels/ray_gen.cu:68-69); the hit record is only defined
    // after a trace, and ptb_get_path_field reports -1 until then
    f.pool.store_state(slot, make_float4(1.0f, 1.0f, 1.0f, 0.0f), f.trace_depth, seed);
    f.q_ext[qpos] = slot;
}

__device__ __forceinline__ Vec3 pow3(Vec3 c, float e) { return Vec3(powf(c.x, e), powf(c.y, e), powf(c.z, e)); }

// ---- Logic (PathLogicKernel, src/kernels/logic.cu:31-192) as its own kernel: the staged API --------------
__global__ void __launch_bounds__(kThreads) k_logic(FrameDev f, int bounce) {
    int* ctr = f.ctr + bounce * CTR_STRIDE;
    const int count = ctr[CTR_EXT];
    const int stride = gridDim.x * blockDim.x;
    const int rounds = (count + stride - 1) / stride;
    for (int r = 0; r < rounds; r++) {
        int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        int queue = -1;
        int idx = -1;
        if (i < count) {
            idx = __ldg(&f.q_ext[i]);
            const HitRecord h = f.pool.load_hit(idx);
            queue = logic_path<false>(f, idx, h.prim, h.mat, h.u, h.v, make_float4(0.0f, 0.0f, 0.0f, h.t));
        }
        logic_dispatch(f, ctr, queue, idx);
    }
}

// ---- Shading (src/kernels/shading.cu:23-259 + UpdatePathState, cuda_utilities.h:233-277) --------
struct ShadeOut {
    bool push_ext;
    bool push_shadow;
    float4 sh_ori_tmax, sh_dir, sh_rad;
    bool push_env;                         // ENV only: the environment's next-event ray
    float4 env_dir, env_rad;               // (its origin is the light ray's: P + Ng * eps)
    float4 env_ori_tmax;
};

template <int TYPE, bool ENV>
__device__ __forceinline__ void shade_path(const FrameDev& f, int idx, ShadeOut& out) {
    float4 oo, dd;                                        // three whole records: ray, hit, state
    ldg256_rw(f.pool.ray_record(idx), oo, dd);
    const HitRecord hr = f.pool.load_hit(idx);
    float4 tp; int remaining; uint32_t seed;
    f.pool.load_state(idx, tp, remaining, seed);
    dd.w = hr.t;                                          // the reference's ray_dir_dist.w after the trace
    float4 huv = make_float4(hr.u, hr.v, 0.0f, 0.0f);
    int prim = hr.prim;
    int mat_id = hr.mat;
    Mat m = load_material(f.materials, mat_id);
    Vec3 N; float uvx, uvy;
    surface_properties(f, prim, huv.x, huv.y, m, N, uvx, uvy);
    if (TYPE == MAT_PBR || TYPE == MAT_DIFFUSE) {
        if (m.diffuse_tex >= 0) {
            float4 tc = tex2D<float4>(__ldg(&f.tex.handles[m.diffuse_tex]), uvx, uvy);
            m.basecolor = m.basecolor * pow3(Vec3(tc.x, tc.y, tc.z), 2.2f);
        }
    }
    if (TYPE == MAT_PBR) {
        if (m.mr_tex >= 0) {
            float4 rm = tex2D<float4>(__ldg(&f.tex.handles[m.mr_tex]), uvx, uvy);
            m.roughness *= rm.y;
            m.metallic *= rm.z;
        }
    }
    Vec3 dirv(dd.x, dd.y, dd.z);
    Vec3 P = Vec3(oo.x, oo.y, oo.z) + dirv * dd.w;
    Vec3 wo = -dirv;
    Vec3 Ng = make_v3(__ldg(&f.mesh.nor_geom[prim]));
    if (TYPE != MAT_REFRACT) { if (dot3(Ng, wo) < 0.0f) Ng = -Ng; }
    Vec3 throughput(tp.x, tp.y, tp.z);

    // ---- next-event estimation (ComputeNextEventEstimation, shading.cu:23-84).  The reference
    // passes the seed BY VALUE, so light and BSDF sampling share random numbers; so do we.
    if ((TYPE == MAT_PBR || TYPE == MAT_DIFFUSE) && f.lights.num_lights != 0) {
        uint32_t nee_seed = seed;
        Vec3 lp, ln; float pdf_area; int light_tri;
        sample_light(f.mesh, f.lights, nee_seed, lp, ln, pdf_area, light_tri);
        Vec3 wi = normalize3(lp - P);
        float dist = length3(lp - P);
        float dist_sq = glm_max(0.000001f, dist * dist);
        float cos_surf = fmaxf(dot3(N, wi), 0.0f);
        float cos_light = fmaxf(dot3(ln, -wi), 0.0f);
        if (!(dot3(Ng, wi) <= 0.0f)) {
            if (cos_surf > 0.0f && cos_light > 0.0f && pdf_area > 0.0f) {
                int lmat_id = __ldg(&f.mesh.indices_matid[light_tri]).w;
                Mat lm = load_material(f.materials, lmat_id);
                Vec3 Le = lm.basecolor * lm.emittance;
                Vec3 fr = eval_bsdf(wo, wi, N, m);
                float pdf = pdf_bsdf(wo, wi, N, m);
                if (length3(fr) > 0.0f) {
                    float pdf_sa = pdf_area * dist_sq / cos_light;
                    float weight = power_heuristic(pdf_sa, pdf);
                    Vec3 L = throughput * Le * fr * (cos_surf * cos_light) / dist_sq * weight / pdf_area;
                    if (length3(L) > 0.0f) {
                        out.push_shadow = true;
                        out.sh_ori_tmax = make_float4(P.x + Ng.x * PTB_EPSILON, P.y + Ng.y * PTB_EPSILON, P.z + Ng.z * PTB_EPSILON, dist);
                        out.sh_dir = make_float4(wi.x, wi.y, wi.z, 0.0f);
                        out.sh_rad = make_float4(L.x, L.y, L.z, 0.0f);
                    }
                }
            }
        }
    }

    // ---- the environment's next-event estimate (ptb_config.env_nee; see envmap.cuh).  Its random numbers come from
    // a stream hashed off the path's state, which is not advanced: the BSDF sample below, hence the whole path and
    // the indirect image, are the same with the flag on or off, and the light's estimate above is untouched.
    if (ENV && (TYPE == MAT_PBR || TYPE == MAT_DIFFUSE)) {
        uint32_t env_seed = wang_hash(seed ^ 0x85ebca6bu);
        if (env_seed == 0) env_seed = 1;
        float r0 = rand_float(env_seed), r1 = rand_float(env_seed);
        float pdf_env, eu, ev; int texel;
        Vec3 wi = sample_environment(f.env, f.tex, r0, r1, pdf_env, texel, eu, ev);
        float cos_surf = fmaxf(dot3(N, wi), 0.0f);
        if (dot3(Ng, wi) > 0.0f && cos_surf > 0.0f && pdf_env > 0.0f) {
            Vec3 fr = eval_bsdf(wo, wi, N, m);
            float pdf = pdf_bsdf(wo, wi, N, m);
            if (length3(fr) > 0.0f && !(pdf > PTB_PDF_DIRAC_DELTA)) {
                float4 e4 = tex2D<float4>(__ldg(&f.tex.handles[f.tex.env_tex_id]), eu, ev);
                // the complement of Logic's weight for the BSDF-sampled hit of the same direction (logic.cuh, src/kernels/logic.cu:99)
                float weight = (pdf_env * pdf_env + PTB_EPSILON) / (pdf * pdf + pdf_env * pdf_env + PTB_EPSILON);
                Vec3 L = throughput * Vec3(e4.x, e4.y, e4.z) * fr * (cos_surf * weight / pdf_env);
                if (length3(L) > 0.0f) {
                    out.push_env = true;
                    out.env_ori_tmax = make_float4(P.x + Ng.x * PTB_EPSILON, P.y + Ng.y * PTB_EPSILON, P.z + Ng.z * PTB_EPSILON, FLT_MAX);
                    out.env_dir = make_float4(wi.x, wi.y, wi.z, 0.0f);
                    out.env_rad = make_float4(L.x, L.y, L.z, 0.0f);
                }
            }
        }
    }

    Vec3 next_dir(0.0f);
    float next_pdf = 0.0f;
    Vec3 att;
    if (TYPE == MAT_PBR) att = sample_pbr(wo, next_dir, next_pdf, N, m, seed);
    else if (TYPE == MAT_DIFFUSE) att = sample_diffuse(wo, next_dir, next_pdf, N, m, seed);
    else if (TYPE == MAT_REFLECT) att = sample_mirror(wo, next_dir, next_pdf, N, m);
    else att = sample_dielectric(wo, next_dir, next_pdf, N, m, seed);

    // ---- UpdatePathState
    if (next_pdf > 0.0f && length3(att) > 0.0f) {
        throughput = throughput * att;
        Vec3 bias_n = Ng;
        bool alive = true;
        if (TYPE == MAT_REFRACT) {
            bool exiting = dot3(next_dir, Ng) > 0.0f;
            bias_n = exiting ? Ng : -Ng;
        } else if (dot3(next_dir, Ng) <= 0.0f) {
            alive = false;
        }
        if (alive) {
            tp = make_float4(throughput.x, throughput.y, throughput.z, next_pdf);
            stg256(f.pool.ray_record(idx),
                   make_float4(P.x + bias_n.x * PTB_EPSILON, P.y + bias_n.y * PTB_EPSILON, P.z + bias_n.z * PTB_EPSILON, 0.0f),
                   make_float4(next_dir.x, next_dir.y, next_dir.z, FLT_MAX));
            out.push_ext = true;
        }
    }
    // one whole state record: a path that ends here keeps its throughput, as in the reference (remaining_bounces = -1, seed stored)
    f.pool.store_state(idx, tp, out.push_ext ? remaining - 1 : -1, seed);
}

// One kernel per material queue: a queue's launch carries only its own model's registers and local frame (the
// four models in one kernel cost 95 registers and an 864-byte frame per thread, most of it dead for any given warp).
// 8 resident blocks per SM (64 registers, 24 bytes of spills) instead of the 4 that the natural 121 registers allow: the
// kernel waits on scattered path-state and mesh loads (ncu: issue slots 21 % busy at 25 % occupancy), more warps hide
// more of it: 143k scene 12.53 -> 12.40 ms, textured scene 7.98 -> 7.56 ms per frame, same bits (6: 12.80, 10: 12.59).
#ifndef PTB_SHADE_MINBLOCKS
#define PTB_SHADE_MINBLOCKS 8
#endif
template <int TYPE, bool ENV>
__global__ void __launch_bounds__(kThreads, PTB_SHADE_MINBLOCKS) k_shade(const __grid_constant__ FrameDev f, int bounce) {
    const int* ctr = f.ctr + bounce * CTR_STRIDE;
    int* next = f.ctr + (bounce + 1) * CTR_STRIDE;
    const int n = ctr[CTR_PBR + TYPE];
    const int stride = gridDim.x * blockDim.x;
    const int rounds = (n + stride - 1) / stride;
    for (int r = 0; r < rounds; r++) {
        int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        ShadeOut out;
        out.push_ext = false;
        out.push_shadow = false;
        out.push_env = false;
        int idx = -1;
        if (i < n) {
            idx = __ldg(&f.q_mat[TYPE][i]);
            shade_path<TYPE, ENV>(f, idx, out);
        }
        int sp = warp_append(out.push_shadow, &next[CTR_SHADOW]);
        if (sp >= 0) {
            f.shadow.ray_ori_tmax[sp] = out.sh_ori_tmax;
            f.shadow.ray_dir[sp] = out.sh_dir;
            f.shadow.radiance[sp] = out.sh_rad;
            f.shadow.slot[sp] = idx;
        }
        if (ENV) {   // second shadow-queue entry; its radiance goes to the accumulator's second half (see renderer.cu)
            int se = warp_append(out.push_env, &next[CTR_SHADOW]);
            if (se >= 0) {
                f.shadow.ray_ori_tmax[se] = out.env_ori_tmax;
                f.shadow.ray_dir[se] = out.env_dir;
                f.shadow.radiance[se] = out.env_rad;
                f.shadow.slot[se] = idx + f.num_paths;
            }
        }
        int ep = warp_append(out.push_ext, &next[CTR_EXT]);
        if (ep >= 0) f.q_ext[ep] = idx;
    }
}

// ---- frame end: ray statistics (MraysPerSec's numerator, src/pathtrace_wavefront.cu:305) --------
__global__ void k_frame_end(FrameDev f) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        unsigned long long e = 0, s = 0;
        for (int b = 0; b < f.trace_depth; b++) {
            e += (unsigned long long)f.ctr[b * CTR_STRIDE + CTR_EXT];
            s += (unsigned long long)f.ctr[b * CTR_STRIDE + CTR_SHADOW];
        }
        f.ray_totals[0] += e;
        f.ray_totals[1] += s;
        *f.iteration += 1;
        if (f.view_proj) f.view_proj[1] = f.view_proj[0];   // prev_view_proj_mat = view_proj_mat, src/pathtrace_wavefront.cu:368
    }
}

// ---- resolve (SendImageToPBOKernel mode 0, src/pathtrace_wavefront.cu:66-80,112-115) -----------
__global__ void __launch_bounds__(256) k_resolve(const float* __restrict__ direct, const float* __restrict__ indirect, int n,
                                                 void* dst, int format, int iter) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Vec3 c;
    if (indirect) {
        c = Vec3(direct[3 * (size_t)i] + indirect[3 * (size_t)i], direct[3 * (size_t)i + 1] + indirect[3 * (size_t)i + 1],
                 direct[3 * (size_t)i + 2] + indirect[3 * (size_t)i + 2]);
    } else {   // denoised final image: shown as it is (SendImageToPBOKernel with denoiserOn, src/pathtrace_wavefront.cu:69-71)
        c = Vec3(direct[3 * (size_t)i], direct[3 * (size_t)i + 1], direct[3 * (size_t)i + 2]);
    }
    if (format == PTB_IMG_SUM_RGB32F) {
        float* o = (float*)dst + 3 * (size_t)i;
        o[0] = c.x; o[1] = c.y; o[2] = c.z;
        return;
    }
    if (indirect) c = c / (float)iter;
    if (format == PTB_IMG_MEAN_RGB32F) {
        float* o = (float*)dst + 3 * (size_t)i;
        o[0] = c.x; o[1] = c.y; o[2] = c.z;
        return;
    }
    c = pow3(c, 1.0f / 2.2f);
    uchar4 px;
    px.x = (unsigned char)min(max((int)(c.x * 255.0f), 0), 255);
    px.y = (unsigned char)min(max((int)(c.y * 255.0f), 0), 255);
    px.z = (unsigned char)min(max((int)(c.z * 255.0f), 0), 255);
    px.w = 255;
    ((uchar4*)dst)[i] = px;
}

// ---- launchers ---------------------------------------------------------------------------------
static inline int stage_grid(const LaunchCfg& c) { return c.num_sms * 8; }

int launch_clear_accumulators(const FrameDev& f, cudaStream_t s) {
    size_t n = 3 * (size_t)f.num_paths;
    k_clear<<<1184, 256, 0, s>>>(f.direct, f.indirect, n);
    return 1;
}
__global__ void k_set_iteration(int* p, int v) { *p = v; }
int launch_set_iteration(const FrameDev& f, int iteration, cudaStream_t s) {
    k_set_iteration<<<1, 1, 0, s>>>(f.iteration, iteration);
    return 1;
}
int launch_raygen(const FrameDev& f, const LaunchCfg&, cudaStream_t s) {
    int work = max(f.num_paths, (f.trace_depth + 1) * CTR_STRIDE);
    k_raygen<<<(work + kThreads - 1) / kThreads, kThreads, 0, s>>>(f);
    return 1;
}
int launch_logic(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s) {
    k_logic<<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
    return 1;
}
int launch_shade(const FrameDev& f, int bounce, const LaunchCfg& c, cudaStream_t s) {
    if (f.env.nee) {
        k_shade<MAT_PBR, true><<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
        k_shade<MAT_DIFFUSE, true><<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
    } else {
        k_shade<MAT_PBR, false><<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
        k_shade<MAT_DIFFUSE, false><<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
    }
    k_shade<MAT_REFLECT, false><<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
    k_shade<MAT_REFRACT, false><<<stage_grid(c), kThreads, 0, s>>>(f, bounce);
    return 4;
}
// One material queue's kernel on its own stream: the four queues hold disjoint paths and only share the (atomic) counters of
// the next bounce, so the frame graph runs them as parallel branches (renderer.cu, LaunchCfg::shade_fork).
int launch_shade_one(const FrameDev& f, int bounce, int type, const LaunchCfg& c, cudaStream_t s) {
    const int g = stage_grid(c);
    switch (type) {
        case MAT_PBR: if (f.env.nee) k_shade<MAT_PBR, true><<<g, kThreads, 0, s>>>(f, bounce); else k_shade<MAT_PBR, false><<<g, kThreads, 0, s>>>(f, bounce); break;
        case MAT_DIFFUSE: if (f.env.nee) k_shade<MAT_DIFFUSE, true><<<g, kThreads, 0, s>>>(f, bounce); else k_shade<MAT_DIFFUSE, false><<<g, kThreads, 0, s>>>(f, bounce); break;
        case MAT_REFLECT: k_shade<MAT_REFLECT, false><<<g, kThreads, 0, s>>>(f, bounce); break;
        default: k_shade<MAT_REFRACT, false><<<g, kThreads, 0, s>>>(f, bounce); break;
    }
    return 1;
}
// ptb_config.env_nee: the environment's shadow rays accumulate into direct[P .. 2P) so that every accumulator cell
// still receives at most one add per bounce (a deterministic sum, bit-identical between the tile partitions); the
// frame ends by folding that half into the direct image.
__global__ void __launch_bounds__(256) k_fold_env(float* direct, int n3) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    float e = direct[n3 + i];
    if (e != 0.0f) {
        direct[i] = direct[i] + e;
        direct[n3 + i] = 0.0f;
    }
}
int launch_fold_env(const FrameDev& f, cudaStream_t s) {
    int n3 = 3 * f.num_paths;
    k_fold_env<<<(n3 + 255) / 256, 256, 0, s>>>(f.direct, n3);
    return 1;
}

// sampleEnvironmentMap for a batch of random pairs (ptb_sample_environment: the parity hook for the sampler)
__global__ void __launch_bounds__(256) k_sample_environment(EnvDev env, TexturesDev tex, const float* rnd, int n, float* dir, float* pdf, int* pixel) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float p, eu, ev; int texel;
    Vec3 d = sample_environment(env, tex, rnd[2 * i], rnd[2 * i + 1], p, texel, eu, ev);
    dir[3 * i] = d.x; dir[3 * i + 1] = d.y; dir[3 * i + 2] = d.z;
    pdf[i] = p;
    pixel[i] = texel;
}
int launch_sample_environment(const FrameDev& f, const float* rnd, int n, float* dir, float* pdf, int* pixel, cudaStream_t s) {
    k_sample_environment<<<(n + 255) / 256, 256, 0, s>>>(f.env, f.tex, rnd, n, dir, pdf, pixel);
    return 1;
}

int launch_frame_end(const FrameDev& f, const LaunchCfg&, cudaStream_t s) {
    k_frame_end<<<1, 32, 0, s>>>(f);
    return 1;
}
int launch_resolve(const FrameDev& f, void* dst, int format, int iteration, cudaStream_t s) {
    k_resolve<<<(f.num_paths + 255) / 256, 256, 0, s>>>(f.direct, f.indirect, f.num_paths, dst, format, iteration);
    return 1;
}
int launch_resolve_final(const float* final_image, int n, void* dst, int format, cudaStream_t s) {
    k_resolve<<<(n + 255) / 256, 256, 0, s>>>(final_image, nullptr, n, dst, format, 1);
    return 1;
}

}  // namespace ptb
