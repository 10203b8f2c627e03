// logic.cuh -- the per-path body of the Logic stage (PathLogicKernel, src/kernels/logic.cu:31-192) and the two
// helpers it shares with Shading.  It is a device function because it runs in two places: as its own kernel
// (k_logic: the staged API, one stage per call, for stage-by-stage parity) and fused into the tail of the
// extension RayCast kernel (the frame graph), where the hit record is still in registers.
#pragma once
#include "material.cuh"
#include "render.cuh"

namespace ptb {

// ---- surface interpolation (GetSurfaceProperties, src/common/cuda_utilities.h:279-326) ---------
__device__ __forceinline__ void surface_properties(const FrameDev& f, int prim, float u, float v, const Mat& mat, Vec3& N, float& uvx, float& uvy) {
    int4 im = __ldg(&f.mesh.indices_matid[prim]);
    float w = 1.0f - u - v;
    float2 uv0 = __ldg(&f.mesh.uv[im.x]), uv1 = __ldg(&f.mesh.uv[im.y]), uv2 = __ldg(&f.mesh.uv[im.z]);
    uvx = w * uv0.x + u * uv1.x + v * uv2.x;
    uvy = w * uv0.y + u * uv1.y + v * uv2.y;
    Vec3 n0 = make_v3(__ldg(&f.mesh.nor[im.x])), n1 = make_v3(__ldg(&f.mesh.nor[im.y])), n2 = make_v3(__ldg(&f.mesh.nor[im.z]));
    Vec3 Ns = normalize3(w * n0 + u * n1 + v * n2);
    if (mat.normal_tex < 0) { N = Ns; return; }
    Vec3 t0 = make_v3(__ldg(&f.mesh.tangent[im.x])), t1 = make_v3(__ldg(&f.mesh.tangent[im.y])), t2 = make_v3(__ldg(&f.mesh.tangent[im.z]));
    Vec3 Ti = w * t0 + u * t1 + v * t2;
    Vec3 B = normalize3(cross3(Ns, Ti));
    Vec3 T = cross3(B, Ns);
    float4 s = tex2D<float4>(__ldg(&f.tex.handles[mat.normal_tex]), uvx, uvy);
    Vec3 m(s.x * 2.0f - 1.0f, s.y * 2.0f - 1.0f, s.z * 2.0f - 1.0f);
    // glm::mat3(T, B, N) * m = T*m.x + B*m.y + N*m.z (type_mat3x3.inl:487-493)
    N = normalize3(Vec3(T.x * m.x + B.x * m.y + Ns.x * m.z, T.y * m.x + B.y * m.y + Ns.y * m.z, T.z * m.x + B.z * m.y + Ns.z * m.z));
}

// warp-aggregated append at a converged point: returns the slot for lanes with `want`
__device__ __forceinline__ int warp_append(bool want, int* counter) {
    unsigned int mask = __ballot_sync(0xffffffffu, want);
    if (mask == 0) return -1;
    int lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return want ? base + __popc(mask & ((1u << lane) - 1u)) : -1;
}

// One live path after its extension ray: miss -> environment (+ pdf-map MIS) and kill; emissive hit -> emission
// (+ area-light MIS) and kill; otherwise return the material queue (0..3) the path goes to, -1 if it ended.
// hit / mat_id / (hit_u, hit_v): the path's hit record.  HAVE_RAY: the caller also holds (dir.xyz, t) in registers (fused
// RayCast tail); otherwise dd_in.w is t and the direction is read when a branch needs it.
template <bool HAVE_RAY>
__device__ __forceinline__ int logic_path(const FrameDev& f, int idx, int hit, int mat_id, float hit_u, float hit_v, float4 dd_in) {
    float4 tp; int remaining; uint32_t rng_unused;
    f.pool.load_state(idx, tp, remaining, rng_unused);
    Vec3 throughput(tp.x, tp.y, tp.z);
    float last_pdf = tp.w;
    bool first = (remaining == f.trace_depth);
    float* pix = f.indirect + 3 * (size_t)idx;
    if (hit == -1) {
        if (f.tex.env_tex_id >= 0) {
            float4 dd = HAVE_RAY ? dd_in : f.pool.ray_dir_dist[idx];
            // normalize with the dot product's contraction PINNED to the reference's fma(z,z, fma(x,x, y*y)): left to nvcc it depends on
            // the kernel this function is inlined into (the 4-wide persistent-lane kernel got fma(z,z, fma(y,y, x*x)), which moves the
            // environment lookup by an ulp and the filtered texel with it: 168 pixels of the textured scene's indirect image)
            const Vec3 dv(dd.x, dd.y, dd.z);
            Vec3 dir = dv * (1.0f / sqrtf(dot3_pinned(dv, dv)));
            float phi = atan2f(dir.z, dir.x);
            if (phi < 0) phi += PTB_TWO_PI;
            float theta = acosf(clampf(dir.y, -1.0f, 1.0f));
            float eu = phi * PTB_INV_TWO_PI;
            float ev = theta * PTB_INV_PI;
            float4 e4 = tex2D<float4>(__ldg(&f.tex.handles[f.tex.env_tex_id]), eu, ev);
            Vec3 env(e4.x, e4.y, e4.z);
            float mis = 1.0f;
            if (!first) {
                float pdf_bsdf = last_pdf;
                float pdf_env = tex2D<float>(__ldg(&f.tex.handles[f.tex.pdf_map_id]), eu, ev);
                if (pdf_bsdf > 1e10f) mis = 1.0f;
                else mis = (pdf_bsdf * pdf_bsdf) / (pdf_bsdf * pdf_bsdf + pdf_env * pdf_env + PTB_EPSILON);
            }
            Vec3 add = throughput * env * mis;
            atomicAdd(pix + 0, add.x); atomicAdd(pix + 1, add.y); atomicAdd(pix + 2, add.z);   // a separate add: `pix += a*b` would fuse
        }
        f.pool.remaining_bounces[idx] = -1;
        return -1;
    }
    Mat m = load_material(f.materials, mat_id);
    if (m.emittance > 0.0f) {
        Vec3 N; float uvx, uvy;
        surface_properties(f, hit, hit_u, hit_v, m, N, uvx, uvy);
        float4 dd = dd_in;
        if (!HAVE_RAY) { float4 rd = f.pool.ray_dir_dist[idx]; dd = make_float4(rd.x, rd.y, rd.z, dd_in.w); }
        Vec3 wo = -Vec3(dd.x, dd.y, dd.z);
        float mis = 1.0f;
        if (!first && f.lights.num_lights > 0) {
            bool prev_specular = (last_pdf > (PTB_PDF_DIRAC_DELTA * 0.9f));
            if (!prev_specular) {
                float dist = dd.w;
                float cos_light = fmaxf(dot3(N, wo), 0.0f);
                if (cos_light > PTB_EPSILON) {
                    float pdf_area = 1.0f / f.lights.total_area;
                    float pdf_sa = pdf_area * (dist * dist) / cos_light;
                    mis = power_heuristic(last_pdf, pdf_sa);
                } else mis = 0.0f;
            }
        }
        Vec3 add = throughput * m.basecolor * m.emittance * mis;
        atomicAdd(pix + 0, add.x); atomicAdd(pix + 1, add.y); atomicAdd(pix + 2, add.z);   // a separate add: `pix += a*b` would fuse
        f.pool.remaining_bounces[idx] = -1;
        return -1;
    }
    return m.type;
}

// Append the path to its material queue; must be reached by all 32 lanes (queue = -1 for lanes without a path).
__device__ __forceinline__ void logic_dispatch(const FrameDev& f, int* ctr, int queue, int idx) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
        int pos = warp_append(queue == q, &ctr[CTR_PBR + q]);
        if (pos >= 0) f.q_mat[q][pos] = idx;
    }
}

}  // namespace ptb
