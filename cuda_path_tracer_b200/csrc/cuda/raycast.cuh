// raycast.cuh -- BVH traversal primitives of the RayCast stage (device functions).
//
// Same tree and the same visiting ORDER as the reference's closest-hit walk
// (src/kernels/ray_cast.cu:26-107): near child first, ties go right, strict `t < t_min`, and the
// walk stops when 32 entries are stacked.  With identical slab / triangle arithmetic this makes
// the winning triangle, t and (u, v) bit-identical, including for equal-t ties on shared edges.
// What changes is the memory side: one 64-byte node record per internal node (2 x LDG.256,
// instead of int2 + 4 scattered float4), vertices inline in the leaf (LDG.256 + LDG.128,
// instead of primitive_indices -> indices_matid -> 3 x pos), the stack in shared memory instead of
// local memory, and no function calls (the reference's two tests are __noinline__).
#pragma once
#include "common.cuh"

namespace ptb {

#ifdef PTB_WIDE_DEBUG
static __device__ unsigned long long g_wide_dbg[8];
#define PTB_WDBG(i) atomicAdd(&g_wide_dbg[i], 1ull)
#else
#define PTB_WDBG(i)
#endif

#define PTB_NODE_NONE 0x7fffffff
#define PTB_EXT_STACK 32          /* src/kernels/ray_cast.cu:35,106 */
#define PTB_SHADOW_STACK 64
#ifndef PTB_RC_THREADS
#define PTB_RC_THREADS 128
#endif

// BoudingboxIntersetionTest, src/intersections.cu:34-61 (FADD, FMUL, min/max chain; no FMA).
__device__ __forceinline__ float slab_test(float4 mn, float4 mx, Vec3 o, Vec3 inv) {
    float t1x = __fmul_rn(__fsub_rn(mn.x, o.x), inv.x);
    float t2x = __fmul_rn(__fsub_rn(mx.x, o.x), inv.x);
    float t_near = fminf(t1x, t2x);
    float t_far = fmaxf(t1x, t2x);
    float t1y = __fmul_rn(__fsub_rn(mn.y, o.y), inv.y);
    float t2y = __fmul_rn(__fsub_rn(mx.y, o.y), inv.y);
    t_near = fmaxf(t_near, fminf(t1y, t2y));
    t_far = fminf(t_far, fmaxf(t1y, t2y));
    float t1z = __fmul_rn(__fsub_rn(mn.z, o.z), inv.z);
    float t2z = __fmul_rn(__fsub_rn(mx.z, o.z), inv.z);
    t_near = fmaxf(t_near, fminf(t1z, t2z));
    t_far = fminf(t_far, fmaxf(t1z, t2z));
    if (t_near <= t_far && t_far > 0.0f) return fmaxf(0.0f, t_near);
    return -1.0f;
}

// TriangleIntersectionTest, src/intersections.cu:10-32, in the contraction the reference's sm_100
// SASS shows: cross = fma(a,b,-(c*d)), dot = fma(z, fma(x, y*y')), then a
// separate multiply by the IEEE reciprocal of det.
__device__ __forceinline__ float triangle_test(Vec3 v0, Vec3 v1, Vec3 v2, Vec3 o, Vec3 d, float& out_u, float& out_v) {
    Vec3 e1(__fsub_rn(v1.x, v0.x), __fsub_rn(v1.y, v0.y), __fsub_rn(v1.z, v0.z));
    Vec3 e2(__fsub_rn(v2.x, v0.x), __fsub_rn(v2.y, v0.y), __fsub_rn(v2.z, v0.z));
    Vec3 pvec = cross3_pinned(d, e2);
    float det = dot3_pinned(e1, pvec);
    if (fabsf(det) < PTB_EPSILON) return -1.0f;
    float inv_det = __frcp_rn(det);
    Vec3 tvec(__fsub_rn(o.x, v0.x), __fsub_rn(o.y, v0.y), __fsub_rn(o.z, v0.z));
    out_u = __fmul_rn(dot3_pinned(tvec, pvec), inv_det);
    if (out_u < 0.0f || out_u > 1.0f) return -1.0f;
    Vec3 qvec = cross3_pinned(tvec, e1);
    out_v = __fmul_rn(dot3_pinned(d, qvec), inv_det);
    if (out_v < 0.0f || __fadd_rn(out_u, out_v) > 1.0f) return -1.0f;
    float t = __fmul_rn(dot3_pinned(e2, qvec), inv_det);
    return (t > 1e-7f) ? t : -1.0f;
}

// One 32-byte sector per lane per instruction (LDG.E.256, sm_100+): half the L1 wavefronts of two 16-byte loads.
struct F8 { float4 a, b; };
__device__ __forceinline__ F8 ldg256(const float4* p) {
    F8 r;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
        : "l"(p));
    return r;
}

struct Hit {
    int prim;      // triangle index, -1 = miss
    int mat;       // material id of the hit triangle
    float t, u, v;
};

// Closest hit.  `stack` points at this thread's column of a [PTB_EXT_STACK][blockDim] shared array.
__device__ __forceinline__ Hit trace_closest(const BvhDev& bvh, Vec3 o, Vec3 d, int* stack, int stride) {
    Vec3 inv(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));   // 1.0f / ray.direction, :29
    Hit h;
    h.prim = -1; h.mat = -1; h.t = FLT_MAX; h.u = 0.0f; h.v = 0.0f;
    int sp = 0;
    int node = (slab_test(bvh.root_min, bvh.root_max, o, inv) == -1.0f) ? PTB_NODE_NONE : 0;
    while (node != PTB_NODE_NONE || sp > 0) {
        if (node == PTB_NODE_NONE) node = stack[(--sp) * stride];
#ifdef PTB_WIDE_DEBUG   // live lanes of this warp-iteration, in bins of 4 (tools/lane_hist.py)
        { unsigned m_ = __activemask(); if ((threadIdx.x & 31) == __ffs(m_) - 1) atomicAdd(&g_wide_dbg[(__popc(m_) - 1) >> 2], 1ull); }
#endif
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
                stack[(sp++) * stride] = left_first ? right : left;
                node = left_first ? left : right;
            } else if (hit_l) node = left;
            else if (hit_r) node = right;
            else node = PTB_NODE_NONE;
        }
        if (sp >= PTB_EXT_STACK) break;   // :106 -- the reference gives up here; so do we
    }
    return h;
}

// Any hit (TraceShadowRayKernel's visibility test, src/kernels/ray_cast.cu:136-191).  The
// reference walks stacklessly over escape indices; the SET of nodes whose box passes
// `t_box != -1 && tmax > t_box` does not depend on the order, so a stack walk finds an occluder
// iff the reference does.  Returns 1 occluded, 0 visible, -1 stack overflow.
template <int STACK>
__device__ __forceinline__ int trace_any(const BvhDev& bvh, Vec3 o, Vec3 d, float tmax, int* stack, int stride) {
    Vec3 inv(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    const float t_limit = __fsub_rn(tmax, PTB_EPSILON);
    int sp = 0;
    float t_root = slab_test(bvh.root_min, bvh.root_max, o, inv);
    int node = (t_root != -1.0f && tmax > t_root) ? 0 : PTB_NODE_NONE;
    while (node != PTB_NODE_NONE || sp > 0) {
        if (node == PTB_NODE_NONE) node = stack[(--sp) * stride];
        if (node < 0) {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float u, v;
            float t = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), o, d, u, v);
            if (t > PTB_EPSILON && t < t_limit) return 1;
            node = PTB_NODE_NONE;
        } else {
            const float4* np = bvh.node + 4 * (size_t)node;
            F8 nl = ldg256(np), nr = ldg256(np + 2);
            float4 n0 = nl.a, n1 = nl.b, n2 = nr.a, n3 = nr.b;
            float t_l = slab_test(n0, n1, o, inv);
            float t_r = slab_test(n2, n3, o, inv);
            int left = __float_as_int(n0.w), right = __float_as_int(n1.w);
            bool hit_l = (t_l != -1.0f && tmax > t_l);
            bool hit_r = (t_r != -1.0f && tmax > t_r);
            if (hit_l && hit_r) {
                if (sp >= STACK) return -1;
                bool left_first = t_l <= t_r;
                stack[(sp++) * stride] = left_first ? right : left;
                node = left_first ? left : right;
            } else if (hit_l) node = left;
            else if (hit_r) node = right;
            else node = PTB_NODE_NONE;
        }
    }
    return 0;
}

}  // namespace ptb
