// raycast.cuh -- BVH traversal primitives of the RayCast stage (device functions).
//
// Same tree and the same visiting ORDER as the reference's closest-hit walk
// (src/kernels/ray_cast.cu:26-107): near child first, ties go right, strict `t < t_min`, and the
// walk stops when 32 entries are stacked.  With identical slab / triangle arithmetic this makes
// the winning triangle, t and (u, v) bit-identical, including for equal-t ties on shared edges.
// What changes is the memory side: one 64-byte node record per internal node (2 x LDG.256,
// instead of int2 + 4 scattered float4; tested with two-wide fp32 instructions, node_test()), vertices inline in the leaf (LDG.256 + LDG.128,
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

// ---- one internal node = both children's slab tests, on sm_100's two-wide fp32 instructions -----------------------------
// RayCast is bound by issue slots, and 24 of the ~67 instructions of a node visit are the `(plane -
// origin) * inv` of twelve planes.  sub.rn.f32x2 / mul.rn.f32x2 (FADD2 / FMUL2) do two of them per issue slot with the same
// per-element IEEE rounding, so the result is bit for bit slab_test()'s.  The node record is laid out for it:
//   node[4*i+0] = (lmin.x, lmin.y, lmax.x, lmax.y)     node[4*i+1] = (rmin.x, rmin.y, rmax.x, rmax.y)
//   node[4*i+2] = (lmin.z, lmax.z, rmin.z, rmax.z)     node[4*i+3] = (bits(left link), bits(right link), 0, 0)
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
struct U64x4 { u64 a, b, c, d; };
__device__ __forceinline__ U64x4 ldg256_b64(const void* p) {
    U64x4 r;
    asm("ld.global.nc.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(r.a), "=l"(r.b), "=l"(r.c), "=l"(r.d) : "l"(p));
    return r;
}
__device__ __forceinline__ float slab_result(float t1x, float t2x, float t1y, float t2y, float t1z, float t2z) {
    float t_near = fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z));
    float t_far = fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z));
    if (t_near <= t_far && t_far > 0.0f) return fmaxf(0.0f, t_near);
    return -1.0f;
}
struct NodePair { float t_l, t_r; int left, right; };    // slab_test() of either child (-1: missed) and the two links
__device__ __forceinline__ NodePair node_test(const float4* np, Vec3 o, Vec3 inv) {
    const u64 oxy = pack2(o.x, o.y), ixy = pack2(inv.x, inv.y), ozz = pack2(o.z, o.z), izz = pack2(inv.z, inv.z);
    const U64x4 q = ldg256_b64(np), z = ldg256_b64(np + 2);
    const u64 A = mul2(sub2(q.a, oxy), ixy), B = mul2(sub2(q.b, oxy), ixy);     // left:  (t1x, t1y), (t2x, t2y)
    const u64 C = mul2(sub2(q.c, oxy), ixy), D = mul2(sub2(q.d, oxy), ixy);     // right
    const u64 E = mul2(sub2(z.a, ozz), izz), F = mul2(sub2(z.b, ozz), izz);     // (t1z, t2z) left, right
    float t1x, t1y, t2x, t2y, t1z, t2z, fl, fr;
    NodePair r;
    unpack2(A, t1x, t1y); unpack2(B, t2x, t2y); unpack2(E, t1z, t2z);
    r.t_l = slab_result(t1x, t2x, t1y, t2y, t1z, t2z);
    unpack2(C, t1x, t1y); unpack2(D, t2x, t2y); unpack2(F, t1z, t2z);
    r.t_r = slab_result(t1x, t2x, t1y, t2y, t1z, t2z);
    unpack2(z.c, fl, fr);
    r.left = __float_as_int(fl); r.right = __float_as_int(fr);
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
            const NodePair nt = node_test(bvh.node + 4 * (size_t)node, o, inv);
            const float t_l = nt.t_l, t_r = nt.t_r;
            const int left = nt.left, right = nt.right;
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
            const NodePair nt = node_test(bvh.node + 4 * (size_t)node, o, inv);
            const float t_l = nt.t_l, t_r = nt.t_r;
            const int left = nt.left, right = nt.right;
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
