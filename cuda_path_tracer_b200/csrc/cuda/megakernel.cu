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
#include "render.cuh"

namespace ptb {

#define PTB_MEGA_THREADS 128
#ifndef PTB_MEGA_MINBLOCKS
#define PTB_MEGA_MINBLOCKS 4
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
            const float4* np = bvh.node + 4 * (size_t)node;
            F8 nl = ldg256(np), nr = ldg256(np + 2);
            float4 n0 = nl.a, n1 = nl.b, n2 = nr.a, n3 = nr.b;
            float t_l = slab_test(n0, n1, o, inv);
            float t_r = slab_test(n2, n3, o, inv);
            int left = __float_as_int(n0.w), right = __float_as_int(n1.w);
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

template <int STACK>
__global__ void __launch_bounds__(PTB_MEGA_THREADS, PTB_MEGA_MINBLOCKS) k_megakernel(const __grid_constant__ FrameDev f) {
    __shared__ int stack[STACK * PTB_MEGA_THREADS];
    int* my_stack = stack + threadIdx.x;
    int* cursor = f.ctr + CTR_WORK_EXT;
    const int iter = *f.iteration;
    const int rows = f.num_paths / f.tile.width;
    const Vec3 cam_pos = cam3(f.cam.pos);
    for (;;) {
        int base = 0;
        if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= f.num_paths) break;
        const int qpos = base + (threadIdx.x & 31);
        int n_ext = 0, n_shadow = 0;
        if (qpos < f.num_paths) {
            const int slot = tiled_slot(qpos, f.tile.width, rows);
            const int pixel_index = f.tile.global_pixel(slot);
            const int x = pixel_index % f.cam.res_x;
            const int y = pixel_index / f.cam.res_x;
            // 1. RNG and camera ray, :258-273
            uint32_t seed = wang_hash((uint32_t)iter * 19990303u + (uint32_t)pixel_index);
            if (seed == 0) seed = 1;
            float jitterX = rand_float(seed) - 0.5f;
            float jitterY = rand_float(seed) - 0.5f;
            Vec3 view = cam3(f.cam.view), right = cam3(f.cam.right), up = cam3(f.cam.up);
            Vec3 ray_dir = normalize3(view
                - right * f.cam.pixel_length[0] * ((float)x + jitterX - (float)f.cam.res_x * 0.5f)
                - up * f.cam.pixel_length[1] * ((float)y + jitterY - (float)f.cam.res_y * 0.5f));
            Vec3 ray_ori = cam_pos;
            Vec3 throughput(1.0f);
            Vec3 accumulated(0.0f);
            float last_pdf = 0.0f;

            for (int depth = 0; depth < f.trace_depth; ++depth) {
                n_ext++;
                Hit hit = mega_closest<STACK>(f.bvh, ray_ori, ray_dir, my_stack, PTB_MEGA_THREADS);

                // miss: the environment, unweighted, :290-309
                if (hit.prim == -1) {
                    if (f.tex.env_tex_id >= 0) {
                        Vec3 unit_dir = normalize3(ray_dir);
                        float theta = acosf(unit_dir.y);
                        float phi = atan2f(unit_dir.z, unit_dir.x);
                        if (phi < 0) phi += 2 * PTB_PI;
                        float u = phi * PTB_INV_TWO_PI;
                        float v = theta * PTB_INV_PI;
                        float4 env = tex2D<float4>(__ldg(&f.tex.handles[f.tex.env_tex_id]), u, v);
                        accumulated = accumulated + throughput * Vec3(env.x, env.y, env.z);
                    }
                    break;
                }

                // material and geometry, :311-334
                Mat material = load_material(f.materials, hit.mat);
                Vec3 N_shading; float uvx, uvy;
                surface_properties(f, hit.prim, hit.u, hit.v, material, N_shading, uvx, uvy);
                Vec3 Ng = make_v3(__ldg(&f.mesh.nor_geom[hit.prim]));
                Vec3 intersect_point = ray_ori + ray_dir * hit.t;
                Vec3 wo = -ray_dir;
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
                    if (depth > 0 && f.lights.num_lights > 0) {
                        bool prev_specular = (last_pdf > (PTB_PDF_DIRAC_DELTA * 0.9f));
                        if (!prev_specular) {
                            float cos_light = fmaxf(dot3(N_shading, wo), 0.0f);
                            if (cos_light > PTB_EPSILON) {
                                float pdf_light_area = 1.0f / f.lights.total_area;
                                float pdf_light_sa = pdf_light_area * (hit.t * hit.t) / cos_light;
                                mis_weight = power_heuristic(last_pdf, pdf_light_sa);
                            } else {
                                mis_weight = 0.0f;
                            }
                        }
                    }
                    accumulated = accumulated + throughput * material.basecolor * material.emittance * mis_weight;
                    break;
                }

                // next-event estimation of the mesh lights, :360-411
                if (f.lights.num_lights > 0 && material.type != MAT_REFLECT && material.type != MAT_REFRACT) {
                    Vec3 light_pos, light_N; float pdf_light_area; int light_idx;
                    sample_light(f.mesh, f.lights, seed, light_pos, light_N, pdf_light_area, light_idx);
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
                            Vec3 L_potential = throughput * Le * fr * G * weight / pdf_light_area;
                            if (length3(L_potential) > 0.0f) {
                                Vec3 shadow_ori = intersect_point + Ng * PTB_EPSILON;
                                float t_max = dist_L - 2.0f * PTB_EPSILON;
                                n_shadow++;
                                // IsOccluded, :122-169: a leaf occludes iff eps < t < t_max - eps, boxes pass iff t_max > t_box
                                int occ = trace_any<STACK>(f.bvh, shadow_ori, wi_L, t_max, my_stack, PTB_MEGA_THREADS);
                                if (occ < 0) atomicExch(f.error_flag, 1);
                                if (occ == 0) accumulated = accumulated + L_potential;
                            }
                        }
                    }
                }

                // BSDF sample, :413-440
                Vec3 next_dir(0.0f);
                float next_pdf = 0.0f;
                Vec3 attenuation(0.0f);
                if (material.type == MAT_PBR) attenuation = sample_pbr(wo, next_dir, next_pdf, N_shading, material, seed);
                else if (material.type == MAT_DIFFUSE) attenuation = sample_diffuse(wo, next_dir, next_pdf, N_shading, material, seed);
                else if (material.type == MAT_REFLECT) attenuation = sample_mirror(wo, next_dir, next_pdf, N_shading, material);
                else if (material.type == MAT_REFRACT) attenuation = sample_dielectric(wo, next_dir, next_pdf, N_shading, material, seed);
                if (next_pdf <= 0.0f || length3(attenuation) <= 0.0f) break;
                throughput = throughput * attenuation;
                last_pdf = next_pdf;
                bool is_reflect = dot3(next_dir, Ng) > 0.0f;
                Vec3 bias_n = is_reflect ? Ng : -Ng;
                ray_ori = intersect_point + bias_n * PTB_EPSILON;
                ray_dir = next_dir;
            }

            // one add per sample; a NaN sample is dropped (its count is not), :444-451
            if (accumulated.x == accumulated.x) {
                float* pix = f.direct + 3 * (size_t)slot;
                atomicAdd(pix + 0, accumulated.x);
                atomicAdd(pix + 1, accumulated.y);
                atomicAdd(pix + 2, accumulated.z);
            }
        }
        // ray statistics (d_total_rays, :443-445), split into extension and visibility rays
        n_ext = __reduce_add_sync(0xffffffffu, n_ext);
        n_shadow = __reduce_add_sync(0xffffffffu, n_shadow);
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&f.ctr[CTR_EXT], n_ext);
            atomicAdd(&f.ctr[CTR_SHADOW], n_shadow);
        }
    }
}

// Returns the number of kernels enqueued.  stack_class: 0 => LBVH depth <= 32, else 64 entries (deeper trees drop the
// far child of the closest-hit walk as the reference does; the visibility walk reports an overflow through error_flag).
int launch_megakernel(const FrameDev& f, const LaunchCfg& c, cudaStream_t s) {
    k_mega_begin<<<1, 128, 0, s>>>(f);
    const int grid = c.num_sms * PTB_MEGA_MINBLOCKS;
    if (c.shadow_deep == 0) k_megakernel<32><<<grid, PTB_MEGA_THREADS, 0, s>>>(f);
    else k_megakernel<64><<<grid, PTB_MEGA_THREADS, 0, s>>>(f);
    return 2;
}

}  // namespace ptb
