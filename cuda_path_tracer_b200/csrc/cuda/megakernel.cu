// megakernel.cu -- the reference's OTHER render path (pathtrace_megakernel::Pathtrace, src/pathtrace.cu:645-706;
// kernel MegakernelPathTrace, :234-452) behind ptb_config.megakernel: one kernel carries a path from the camera
// through every bounce -- closest hit, miss / emission, next-event estimation with its visibility ray, BSDF
// sample -- in registers, and adds the finished sample to the image once.  SURVEY.md section 8f rank 3.
//
// It is a different ESTIMATOR from the wavefront path, not just a different schedule, and this file keeps the
// megakernel's own semantics so that its image is the reference megakernel's image:
//   seed = wang_hash(iter * 19990303 + pixel) and xorshift jitter (:258-263; the wavefront hashes twice more);
//   the camera's x axis is mirrored (view - right * ..., :265; the wavefront adds);
//   environment hits are unweighted and the direction's y is not clamped before acos (:292-304);
//   the geometric normal is flipped to the wo side for EVERY material, textures are applied to every type (:318-334);
//   SampleLight ADVANCES the path's random stream (:366-368; the wavefront's NEE works on a copy);
//   L = T * Le * f * G * w / pdfA with tmax = dist - 2 eps, no geometric-normal cull (:376-401);
//   a sampled direction below the surface is offset to the other side instead of ending the path (:432-435);
//   closest hit stacks up to 64 entries and drops the far child beyond that instead of giving up (:52,106-113);
//   radiance is summed in a register in bounce order and added to the image once, unless it is NaN (:444-451).
//
// What is B200-native: warps pull 8x4-pixel tiles from a device cursor (a persistent grid sized from the SM count, no
// tail of idle blocks), the traversal reads the 64-byte node / triangle records with 256-bit loads and keeps its
// stack in shared memory (the reference: int2 + 4 scattered float4 per node, 256 bytes of local-memory stack per
// thread), and the visibility ray is an ordered stack walk instead of the stackless escape-index walk (the set of
// boxes that pass `tmax > t_box` does not depend on the order, raycast.cuh).
#include "bsdf.cuh"
#include "logic.cuh"
#include "raycast.cuh"
#include "wide.cuh"
#include "render.cuh"

namespace ptb {

#define PTB_MEGA_THREADS 128
#ifndef PTB_MEGA_MINBLOCKS
#define PTB_MEGA_MINBLOCKS 10
#endif

__device__ __forceinline__ Vec3 cam3(const float* p) { return Vec3(p[0], p[1], p[2]); }
__device__ __forceinline__ Vec3 powv(Vec3 c, float e) { return Vec3(powf(c.x, e), powf(c.y, e), powf(c.z, e)); }

// FindNearestIntersection, src/pathtrace.cu:42-120: the wavefront's walk (same order, same tests) with a 64-entry
// stack that drops the far child when full.  STACK is chosen from the LBVH depth (>= depth => never full).
template <int STACK>
__device__ __forceinline__ Hit mega_closest(const BvhDev& bvh, Vec3 o, Vec3 d, int* stack, int stride) {
    Vec3 inv(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    Hit h;
    h.prim = -1; h.mat = -1; h.t = FLT_MAX; h.u = 0.0f; h.v = 0.0f;
    int sp = 0;
    int node = (slab_test(bvh.root_min, bvh.root_max, o, inv) == -1.0f) ? PTB_NODE_NONE : 0;
    while (node != PTB_NODE_NONE || sp > 0) {
        if (node == PTB_NODE_NONE) node = stack[(--sp) * stride];
        if (node < 0) {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float u, v;
            float t = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), o, d, u, v);
            if (t > PTB_EPSILON && t < h.t) {
                h.t = t; h.u = u; h.v = v;
                h.prim = __float_as_int(a.w);
                h.mat = __float_as_int(b.w);
            }
            node = PTB_NODE_NONE;
        } else {
            const NodePair nt = node_test(bvh.node + 4 * (size_t)node, o, inv);
            const float t_l = nt.t_l, t_r = nt.t_r;
            const int left = nt.left, right = nt.right;
            bool hit_l = (t_l != -1.0f && t_l < h.t);
            bool hit_r = (t_r != -1.0f && t_r < h.t);
            if (hit_l && hit_r) {
                bool left_first = t_l < t_r;
                if (sp < STACK) stack[(sp++) * stride] = left_first ? right : left;   // :106-113
                node = left_first ? left : right;
            } else if (hit_l) node = left;
            else if (hit_r) node = right;
            else node = PTB_NODE_NONE;
        }
    }
    return h;
}

__global__ void k_mega_begin(FrameDev f) {
    int nctr = (f.trace_depth + 1) * CTR_STRIDE;
    for (int i = threadIdx.x; i < nctr; i += blockDim.x) f.ctr[i] = 0;
}

// One path in flight on a lane.  Everything the reference's thread keeps across its depth loop (:275-281).
struct MegaPath {
    Vec3 ray_ori, ray_dir, throughput, accumulated;
    float last_pdf;
    uint32_t seed;
    int slot, depth;
};

// Camera ray of queue position `qpos`, :254-281.
__device__ __forceinline__ void mega_start(const FrameDev& f, int iter, int rows, int qpos, MegaPath& p) {
    p.slot = tiled_slot(qpos, f.tile.width, rows);
    const int pixel_index = f.tile.global_pixel(p.slot);
    const int x = pixel_index % f.cam.res_x;
    const int y = pixel_index / f.cam.res_x;
    uint32_t seed = wang_hash((uint32_t)iter * 19990303u + (uint32_t)pixel_index);
    if (seed == 0) seed = 1;
    float jitterX = rand_float(seed) - 0.5f;
    float jitterY = rand_float(seed) - 0.5f;
    Vec3 view = cam3(f.cam.view), right = cam3(f.cam.right), up = cam3(f.cam.up);
    p.ray_dir = normalize3(view
        - right * f.cam.pixel_length[0] * ((float)x + jitterX - (float)f.cam.res_x * 0.5f)
        - up * f.cam.pixel_length[1] * ((float)y + jitterY - (float)f.cam.res_y * 0.5f));
    p.ray_ori = cam3(f.cam.pos);
    p.throughput = Vec3(1.0f);
    p.accumulated = Vec3(0.0f);
    p.last_pdf = 0.0f;
    p.seed = seed;
    p.depth = 0;
}

// One iteration of the reference's depth loop (:283-440) for the path on this lane.  Returns false when the path ended
// (miss, emitter, dead sample, or the bounce budget).
template <int STACK, bool WIDE>
__device__ __forceinline__ bool mega_bounce(const FrameDev& f, MegaPath& p, int* my_stack, int& n_ext, int& n_shadow) {
    n_ext++;
    Hit hit = mega_closest<STACK>(f.bvh, p.ray_ori, p.ray_dir, my_stack, PTB_MEGA_THREADS);

    // miss: the environment, unweighted, :290-309
    if (hit.prim == -1) {
        if (f.tex.env_tex_id >= 0) {
            Vec3 unit_dir = normalize3(p.ray_dir);
            float theta = acosf(unit_dir.y);
            float phi = atan2f(unit_dir.z, unit_dir.x);
            if (phi < 0) phi += 2 * PTB_PI;
            float u = phi * PTB_INV_TWO_PI;
            float v = theta * PTB_INV_PI;
            float4 env = tex2D<float4>(__ldg(&f.tex.handles[f.tex.env_tex_id]), u, v);
            p.accumulated = p.accumulated + p.throughput * Vec3(env.x, env.y, env.z);
        }
        return false;
    }

    // material and geometry, :311-334
    Mat material = load_material(f.materials, hit.mat);
    Vec3 N_shading; float uvx, uvy;
    surface_properties(f, hit.prim, hit.u, hit.v, material, N_shading, uvx, uvy);
    Vec3 Ng = make_v3(__ldg(&f.mesh.nor_geom[hit.prim]));
    Vec3 intersect_point = p.ray_ori + p.ray_dir * hit.t;
    Vec3 wo = -p.ray_dir;
    if (dot3(Ng, wo) < 0.0f) { Ng = -Ng; }
    if (material.diffuse_tex >= 0) {
        float4 tc = tex2D<float4>(__ldg(&f.tex.handles[material.diffuse_tex]), uvx, uvy);
        material.basecolor = material.basecolor * powv(Vec3(tc.x, tc.y, tc.z), 2.2f);
    }
    if (material.mr_tex >= 0) {
        float4 rm = tex2D<float4>(__ldg(&f.tex.handles[material.mr_tex]), uvx, uvy);
        material.roughness *= rm.y;
        material.metallic *= rm.z;
    }

    // emission with the area-light MIS weight, :336-358
    if (material.emittance > 0.0f) {
        float mis_weight = 1.0f;
        if (p.depth > 0 && f.lights.num_lights > 0) {
            bool prev_specular = (p.last_pdf > (PTB_PDF_DIRAC_DELTA * 0.9f));
            if (!prev_specular) {
                float cos_light = fmaxf(dot3(N_shading, wo), 0.0f);
                if (cos_light > PTB_EPSILON) {
                    float pdf_light_area = 1.0f / f.lights.total_area;
                    float pdf_light_sa = pdf_light_area * (hit.t * hit.t) / cos_light;
                    mis_weight = power_heuristic(p.last_pdf, pdf_light_sa);
                } else {
                    mis_weight = 0.0f;
                }
            }
        }
        p.accumulated = p.accumulated + p.throughput * material.basecolor * material.emittance * mis_weight;
        return false;
    }

    // next-event estimation of the mesh lights, :360-411
    if (f.lights.num_lights > 0 && material.type != MAT_REFLECT && material.type != MAT_REFRACT) {
        Vec3 light_pos, light_N; float pdf_light_area; int light_idx;
        sample_light(f.mesh, f.lights, p.seed, light_pos, light_N, pdf_light_area, light_idx);
        Vec3 wi_L = normalize3(light_pos - intersect_point);
        float dist_L = length3(intersect_point - light_pos);          // glm::distance(p0, p1) = length(p1 - p0)
        float cos_surf = fmaxf(dot3(N_shading, wi_L), 0.0f);
        float cos_light = fmaxf(dot3(light_N, -wi_L), 0.0f);
        if (cos_surf > 0.0f && cos_light > 0.0f && pdf_light_area > 0.0f) {
            int light_mat_id = __ldg(&f.mesh.indices_matid[light_idx]).w;
            Mat light_mat = load_material(f.materials, light_mat_id);
            Vec3 Le = light_mat.basecolor * light_mat.emittance;
            Vec3 fr = eval_bsdf(wo, wi_L, N_shading, material);
            float pdf_b = pdf_bsdf(wo, wi_L, N_shading, material);
            if (length3(fr) > 0.0f) {
                float pdf_light_sa = pdf_light_area * (dist_L * dist_L) / cos_light;
                float weight = power_heuristic(pdf_light_sa, pdf_b);
                float G = (cos_surf * cos_light) / (dist_L * dist_L);
                Vec3 L_potential = p.throughput * Le * fr * G * weight / pdf_light_area;
                if (length3(L_potential) > 0.0f) {
                    Vec3 shadow_ori = intersect_point + Ng * PTB_EPSILON;
                    float t_max = dist_L - 2.0f * PTB_EPSILON;
                    n_shadow++;
                    // IsOccluded, :122-169: a leaf occludes iff eps < t < t_max - eps, boxes pass iff t_max > t_box
                    int occ = WIDE ? trace_any_exact<STACK>(f.bvh, shadow_ori, wi_L, t_max, my_stack, PTB_MEGA_THREADS)
                                   : trace_any<STACK>(f.bvh, shadow_ori, wi_L, t_max, my_stack, PTB_MEGA_THREADS);
                    if (occ < 0) atomicExch(f.error_flag, 1);
                    if (occ == 0) p.accumulated = p.accumulated + L_potential;
                }
            }
        }
    }

    // BSDF sample, :413-440
    Vec3 next_dir(0.0f);
    float next_pdf = 0.0f;
    Vec3 attenuation(0.0f);
    if (material.type == MAT_PBR) attenuation = sample_pbr(wo, next_dir, next_pdf, N_shading, material, p.seed);
    else if (material.type == MAT_DIFFUSE) attenuation = sample_diffuse(wo, next_dir, next_pdf, N_shading, material, p.seed);
    else if (material.type == MAT_REFLECT) attenuation = sample_mirror(wo, next_dir, next_pdf, N_shading, material);
    else if (material.type == MAT_REFRACT) attenuation = sample_dielectric(wo, next_dir, next_pdf, N_shading, material, p.seed);
    if (next_pdf <= 0.0f || length3(attenuation) <= 0.0f) return false;
    p.throughput = p.throughput * attenuation;
    p.last_pdf = next_pdf;
    bool is_reflect = dot3(next_dir, Ng) > 0.0f;
    Vec3 bias_n = is_reflect ? Ng : -Ng;
    p.ray_ori = intersect_point + bias_n * PTB_EPSILON;
    p.ray_dir = next_dir;
    p.depth++;
    return p.depth < f.trace_depth;
}

// One add per sample; a NaN sample is dropped (its count is not), :444-451.
__device__ __forceinline__ void mega_finish(const FrameDev& f, const MegaPath& p) {
    if (p.accumulated.x == p.accumulated.x) {
        float* pix = f.direct + 3 * (size_t)p.slot;
        atomicAdd(pix + 0, p.accumulated.x);
        atomicAdd(pix + 1, p.accumulated.y);
        atomicAdd(pix + 2, p.accumulated.z);
    }
}

// REGEN = false: a warp takes an 8x4-pixel tile and carries its 32 paths to the end before it takes the next (lanes
// whose path ended wait for the longest one, as in the reference).  REGEN = true: a lane whose path ended takes the
// next pixel of the frame at once (path regeneration), so a warp stays full until the frame runs out; which lane
// renders a pixel does not enter its arithmetic, so both give the same image.
template <int STACK, bool REGEN, bool WIDE>
__global__ void __launch_bounds__(PTB_MEGA_THREADS, PTB_MEGA_MINBLOCKS) k_megakernel(const __grid_constant__ FrameDev f, int refill_min) {
    __shared__ int stack[STACK * PTB_MEGA_THREADS];
    int* my_stack = stack + threadIdx.x;
    int* cursor = f.ctr + CTR_WORK_EXT;
    const int iter = *f.iteration;
    const int rows = f.num_paths / f.tile.width;
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int n_ext = 0, n_shadow = 0;
    MegaPath p;
    if (!REGEN) {
        for (;;) {
            int base = 0;
            if (lane == 0) base = atomicAdd(cursor, 32);
            base = __shfl_sync(FULL, base, 0);
            if (base >= f.num_paths) break;
            const int qpos = base + lane;
            if (qpos < f.num_paths) {
                mega_start(f, iter, rows, qpos, p);
                while (mega_bounce<STACK, WIDE>(f, p, my_stack, n_ext, n_shadow)) {}
                mega_finish(f, p);
            }
        }
    } else {
        bool active = false, exhausted = false;
        for (;;) {
            unsigned idle = __ballot_sync(FULL, !active);
            if (!exhausted && (idle == FULL || __popc(idle) >= refill_min)) {
                int base = 0;
                if (lane == 0) base = atomicAdd(cursor, __popc(idle));
                base = __shfl_sync(FULL, base, 0);
                const int qpos = base + __popc(idle & ((1u << lane) - 1u));
                if (!active && qpos < f.num_paths) {
                    mega_start(f, iter, rows, qpos, p);
                    active = true;
                }
                if (base + __popc(idle) >= f.num_paths) exhausted = true;
                idle = __ballot_sync(FULL, !active);
            }
            if (idle == FULL) break;
            if (active) {
                if (!mega_bounce<STACK, WIDE>(f, p, my_stack, n_ext, n_shadow)) {
                    mega_finish(f, p);
                    active = false;
                }
            }
        }
    }
    // ray statistics (d_total_rays, :443-445), split into extension and visibility rays
    n_ext = __reduce_add_sync(FULL, n_ext);
    n_shadow = __reduce_add_sync(FULL, n_shadow);
    if (lane == 0) {
        atomicAdd(&f.ctr[CTR_EXT], n_ext);
        atomicAdd(&f.ctr[CTR_SHADOW], n_shadow);
    }
}

// Returns the number of kernels enqueued.  The stack is sized from the LBVH depth: <= 32 levels => 32 entries, else 64
// (deeper trees drop the far child of the closest-hit walk as the reference does; the visibility walk reports an
// overflow through error_flag).
int launch_megakernel(const FrameDev& f, const LaunchCfg& c, cudaStream_t s) {
    k_mega_begin<<<1, 128, 0, s>>>(f);
    const int grid = c.num_sms * (c.mega_blocks > 0 ? c.mega_blocks : PTB_MEGA_MINBLOCKS);
    const bool deep = c.shadow_deep != 0;
    const int refill = c.mega_regen ? c.mega_refill : 0;
    const bool wide = (c.wide & 4) != 0;
#define PTB_MEGA_LAUNCH(S, R, W) k_megakernel<S, R, W><<<grid, PTB_MEGA_THREADS, 0, s>>>(f, refill)
    if (c.mega_regen) {
        if (!deep) { if (wide) PTB_MEGA_LAUNCH(32, true, true); else PTB_MEGA_LAUNCH(32, true, false); }
        else { if (wide) PTB_MEGA_LAUNCH(64, true, true); else PTB_MEGA_LAUNCH(64, true, false); }
    } else {
        if (!deep) { if (wide) PTB_MEGA_LAUNCH(32, false, true); else PTB_MEGA_LAUNCH(32, false, false); }
        else { if (wide) PTB_MEGA_LAUNCH(64, false, true); else PTB_MEGA_LAUNCH(64, false, false); }
    }
#undef PTB_MEGA_LAUNCH
    return 2;
}

}  // namespace ptb
