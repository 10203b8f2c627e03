// wide.cuh -- traversal of the 4-wide, 15-bit-quantised collapse of the LBVH (BvhDev::wnode, built by k_pack_wide).
//
// Why: RayCast on B200 is bound by the L1 data pipe and the issue slots, not by HBM: every node
// visit of the binary walk pulls a 64-byte record through L1 to test two boxes.  Here a 64-byte record holds FOUR
// child boxes, so a ray makes about half the dependent round trips and pulls half the bytes per box tested.
//
// Why the results are still the reference's.  The reference tests a triangle iff every box on the leaf's ancestor chain
// passes BoudingboxIntersetionTest (src/intersections.cu:34-61).  That test is monotone under box inclusion in fp32 (each
// (p - o) * inv is monotone in p, min / max are monotone), so the chain passes iff the LEAF's own box passes -- and the
// leaf's box is the min / max of the triangle's vertices (ComputeAABB, src/bvh.cu:13-35), which the leaf record holds.
// So: inner boxes here only have to be conservative (they are: rounded outwards plus a margin larger than the
// reconstruction error), and a triangle hit is accepted only if the reference's own test on the leaf's fp32 box passes
// with the reference's cull (`t_box < t_best`, resp. `tmax > t_box`).  What remains order-dependent in the reference's
// walk is which of two hits within a few ulps of each other wins (exact ties; a leaf whose fp32 box distance rounds
// above its own hit distance can be culled by a hit that is farther by less than that).  trace_closest_wide() keeps
// every candidate within PTB_WIDE_SLACK of the best alive, notices when two accepted hits are that close, and then
// reports `ambiguous`; the caller re-traces such a ray (about one in 10^5) with the reference-order binary walk.
// Rays with a zero direction component (inv = inf; the fma form would produce NaN where the reference's form does
// not) and stack overflows take the same exit.
#pragma once
#include "raycast.cuh"

namespace ptb {

#define PTB_WIDE_SLACK 1.00001f
#define PTB_WIDE_STACK 32
#define PTB_WIDE_PUSHES 3          /* spare stack entries of the persistent-lane walks: a node visit pushes at most three children */
#define PTB_WIDE_SEL_LO 0x7104u   /* PRMT selectors, see wide_plane() */
#define PTB_WIDE_SEL_HI 0x7324u

struct WideRay {
    Vec3 o, d, inv;          // inv = 1 / d (IEEE), as the reference computes it
    float iqx, iqy, iqz;     // inv * cell: t per quantisation step
    float bx, by, bz;        // t(q) = fma(32768 + q, iq, b)
    unsigned near_x, near_y, near_z;   // PRMT selector of an axis' near plane: the `lo` half of a child's word if inv >= 0, else `hi`
    bool ok;                 // false: direction has a zero / non-finite reciprocal -> use the binary walk
};

__device__ __forceinline__ WideRay wide_setup(const BvhDev& bvh, Vec3 o, Vec3 d) {
    WideRay r;
    r.o = o; r.d = d;
    r.inv = Vec3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    r.iqx = r.inv.x * bvh.wcell[0]; r.iqy = r.inv.y * bvh.wcell[1]; r.iqz = r.inv.z * bvh.wcell[2];
    r.bx = (bvh.root_min.x - o.x) * r.inv.x - 32768.0f * r.iqx;
    r.by = (bvh.root_min.y - o.y) * r.inv.y - 32768.0f * r.iqy;
    r.bz = (bvh.root_min.z - o.z) * r.inv.z - 32768.0f * r.iqz;
    r.near_x = r.inv.x < 0.0f ? PTB_WIDE_SEL_HI : PTB_WIDE_SEL_LO;
    r.near_y = r.inv.y < 0.0f ? PTB_WIDE_SEL_HI : PTB_WIDE_SEL_LO;
    r.near_z = r.inv.z < 0.0f ? PTB_WIDE_SEL_HI : PTB_WIDE_SEL_LO;
    const float big = 3.0e38f;
    r.ok = fabsf(r.inv.x) < big && fabsf(r.inv.y) < big && fabsf(r.inv.z) < big &&
           fabsf(r.bx) < big && fabsf(r.by) < big && fabsf(r.bz) < big;     // also false for NaN
    return r;
}

// A child's word holds lo | hi << 16 (15-bit steps).  One PRMT turns either half into the float 32768 + q:
// bytes (0, q.lo, q.hi, 0x47); the selector picks the half, so the ray's octant costs no extra instruction.
// (prmt.b32 itself: __byte_perm() masks the selector with 0x7777 first, one more ALU instruction per plane when the selector
// comes out of a register.)
__device__ __forceinline__ float wide_plane(unsigned word, unsigned sel) {
    unsigned r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(word), "r"(0x47000000u), "r"(sel));
    return __uint_as_float(r);
}

// Entry distances of the four children (max(0, t_near)), or -1 where the ray misses the box or enters beyond t_limit.
template <bool MISS_IS_MAX = false>
__device__ __forceinline__ void wide_test(const WideRay& r, const uint4& qx, const uint4& qy, const uint4& qz, float t_limit, float t[4]) {
    const unsigned far_x = r.near_x ^ 0x0220u, far_y = r.near_y ^ 0x0220u, far_z = r.near_z ^ 0x0220u;
    const unsigned wx[4] = {qx.x, qx.y, qx.z, qx.w}, wy[4] = {qy.x, qy.y, qy.z, qy.w}, wz[4] = {qz.x, qz.y, qz.z, qz.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float tn = fmaxf(fmaxf(fmaf(wide_plane(wx[k], r.near_x), r.iqx, r.bx), fmaf(wide_plane(wy[k], r.near_y), r.iqy, r.by)),
                         fmaf(wide_plane(wz[k], r.near_z), r.iqz, r.bz));
        float tf = fminf(fminf(fmaf(wide_plane(wx[k], far_x), r.iqx, r.bx), fmaf(wide_plane(wy[k], far_y), r.iqy, r.by)),
                         fmaf(wide_plane(wz[k], far_z), r.iqz, r.bz));
        tn = fmaxf(tn, 0.0f);
        tf = fminf(tf, t_limit);
        t[k] = (tn <= tf) ? tn : (MISS_IS_MAX ? FLT_MAX : -1.0f);     // MISS_IS_MAX: misses sort last in the closest-hit walks
    }
}

// The reference's own test on the leaf's fp32 box (= min / max of the vertices, src/bvh.cu:24-25).
__device__ __forceinline__ float leaf_box_test(float4 a, float4 b, float4 c, Vec3 o, Vec3 inv) {
    float4 mn = make_float4(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z), 0.0f);
    float4 mx = make_float4(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z), 0.0f);
    return slab_test(mn, mx, o, inv);
}

#define PTB_WIDE_CSWAP(i, j)                                         \
    {                                                                \
        bool sw_ = key[j] < key[i];                                  \
        float tk_ = sw_ ? key[j] : key[i];                           \
        int tl_ = sw_ ? lnk[j] : lnk[i];                             \
        key[j] = sw_ ? key[i] : key[j];                              \
        lnk[j] = sw_ ? lnk[i] : lnk[j];                              \
        key[i] = tk_;                                                \
        lnk[i] = tl_;                                                \
    }

// Any hit.  Returns 1 occluded, 0 visible, -1: not decidable here (stack overflow), take the binary walk.
template <int STACK>
__device__ __forceinline__ int trace_any_wide(const BvhDev& bvh, const WideRay& r, float tmax, int* stack, int stride) {
    const float t_limit = __fsub_rn(tmax, PTB_EPSILON);
    int sp = 0;
    int node = 0;
    int result = 0;
    // one loop, one exit: the walk ends by emptying (node, sp), so the lanes of a warp reconverge at the loop head
    while (node != PTB_NODE_NONE || sp > 0) {
        if (node == PTB_NODE_NONE) node = stack[(--sp) * stride];
        if (node >= 0) {
            const uint4* np = bvh.wnode + 4 * (size_t)node;
            F8 a = ldg256(reinterpret_cast<const float4*>(np)), b = ldg256(reinterpret_cast<const float4*>(np + 2));
            const uint4 qx = make_uint4(__float_as_uint(a.a.x), __float_as_uint(a.a.y), __float_as_uint(a.a.z), __float_as_uint(a.a.w));
            const uint4 qy = make_uint4(__float_as_uint(a.b.x), __float_as_uint(a.b.y), __float_as_uint(a.b.z), __float_as_uint(a.b.w));
            const uint4 qz = make_uint4(__float_as_uint(b.a.x), __float_as_uint(b.a.y), __float_as_uint(b.a.z), __float_as_uint(b.a.w));
            const int lnk[4] = {__float_as_int(b.b.x), __float_as_int(b.b.y), __float_as_int(b.b.z), __float_as_int(b.b.w)};
            float t[4];
            wide_test(r, qx, qy, qz, tmax, t);
            node = PTB_NODE_NONE;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (t[k] >= 0.0f) {
                    if (node == PTB_NODE_NONE) node = lnk[k];
                    else if (sp < STACK) stack[(sp++) * stride] = lnk[k];
                    else result = -1;
                }
            }
        } else {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float u, v;
            float t = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), r.o, r.d, u, v);
            node = PTB_NODE_NONE;
            if (t > PTB_EPSILON && t < t_limit) {
                float tb = leaf_box_test(a, b, c, r.o, r.inv);          // the reference reaches this leaf iff its own box passes
                if (tb != -1.0f && tmax > tb) result = 1;
            }
        }
        if (result != 0) { node = PTB_NODE_NONE; sp = 0; }
    }
    return result;
}

// Closest hit, nearest child first.  `ambiguous` is set when the answer could depend on the reference's visiting order
// (two accepted hits within PTB_WIDE_SLACK of each other) or the stack overflowed: the caller re-traces with trace_closest().
template <int STACK>
__device__ __forceinline__ Hit trace_closest_wide(const BvhDev& bvh, const WideRay& r, int* stack, int stride, bool& ambiguous) {
    Hit h;
    h.prim = -1; h.mat = -1; h.t = FLT_MAX; h.u = 0.0f; h.v = 0.0f;
    float limit = FLT_MAX;          // h.t * slack
    float second = FLT_MAX;         // nearest accepted candidate other than h
    ambiguous = false;
    int sp = 0;
    int node = 0;
    bool overflow = false;
    while (node != PTB_NODE_NONE || sp > 0) {
        if (node == PTB_NODE_NONE) node = stack[(--sp) * stride];
        if (node >= 0) {
            const uint4* np = bvh.wnode + 4 * (size_t)node;
            F8 a = ldg256(reinterpret_cast<const float4*>(np)), b = ldg256(reinterpret_cast<const float4*>(np + 2));
            const uint4 qx = make_uint4(__float_as_uint(a.a.x), __float_as_uint(a.a.y), __float_as_uint(a.a.z), __float_as_uint(a.a.w));
            const uint4 qy = make_uint4(__float_as_uint(a.b.x), __float_as_uint(a.b.y), __float_as_uint(a.b.z), __float_as_uint(a.b.w));
            const uint4 qz = make_uint4(__float_as_uint(b.a.x), __float_as_uint(b.a.y), __float_as_uint(b.a.z), __float_as_uint(b.a.w));
            int lnk[4] = {__float_as_int(b.b.x), __float_as_int(b.b.y), __float_as_int(b.b.z), __float_as_int(b.b.w)};
            float key[4];
            PTB_WDBG(4);
            wide_test(r, qx, qy, qz, limit, key);
#pragma unroll
            for (int k = 0; k < 4; k++) if (key[k] < 0.0f) key[k] = FLT_MAX;     // misses sort last
            PTB_WIDE_CSWAP(0, 1) PTB_WIDE_CSWAP(2, 3) PTB_WIDE_CSWAP(0, 2) PTB_WIDE_CSWAP(1, 3) PTB_WIDE_CSWAP(1, 2)
            // push the farther hits, farthest first
#pragma unroll
            for (int k = 3; k >= 1; k--) {
                if (key[k] != FLT_MAX) {
                    if (sp < STACK) stack[(sp++) * stride] = lnk[k];
                    else overflow = true;
                }
            }
            node = (key[0] != FLT_MAX) ? lnk[0] : PTB_NODE_NONE;
        } else {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float u, v;
            float t = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), r.o, r.d, u, v);
            PTB_WDBG(5);
            if (t > PTB_EPSILON && t < limit) {
                float tb = leaf_box_test(a, b, c, r.o, r.inv);
                if (tb != -1.0f && tb < limit) {
                    if (t < h.t && tb < h.t) {                       // the reference's acceptance, src/kernels/ray_cast.cu:68,93-94
                        second = fminf(second, h.t);
                        h.t = t; h.u = u; h.v = v;
                        h.prim = __float_as_int(a.w);
                        h.mat = __float_as_int(b.w);
                        limit = h.t * PTB_WIDE_SLACK;
                    } else {
                        second = fminf(second, t);
                    }
                }
            }
            node = PTB_NODE_NONE;
        }
        if (overflow) { node = PTB_NODE_NONE; sp = 0; }
    }
    ambiguous = overflow || ((h.prim != -1) && (second <= limit));
    return h;
}

// ---- resumable walks: one traversal step per call, so that a scheduler can interleave rays of different ages in one warp
// (raycast.cu, persistent lanes).  STACK is a policy with get(i) / set(i, v) (raycast.cu).
// WalkBinary is trace_closest() cut at the loop body: the reference's order, the reference's arithmetic, its 32-entry abort.
struct WalkBinary {
    Vec3 o, d, inv;
    Hit h;
    int node, sp;
    // Boxes entered beyond `ub` are skipped.  FLT_MAX: the reference's walk.  A finite bound is for re-walks that already know the
    // answer lies below it: the reference's decisions among the hits below the bound do not depend on what it met beyond it (its best
    // so far is then either such a far hit or none, and every test a near hit has to pass -- t < best, box distance < best -- holds
    // in both cases), and in trees under 32 levels its stack never aborts.
    float ub;
    __device__ __forceinline__ void reset() { node = PTB_NODE_NONE; sp = 0; }
    __device__ __forceinline__ bool busy() const { return node != PTB_NODE_NONE || sp > 0; }
    __device__ __forceinline__ void init(const BvhDev& bvh, Vec3 o_, Vec3 d_) {
        o = o_; d = d_;
        inv = Vec3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
        h.prim = -1; h.mat = -1; h.t = FLT_MAX; h.u = 0.0f; h.v = 0.0f;
        sp = 0;
        ub = FLT_MAX;
        node = (slab_test(bvh.root_min, bvh.root_max, o, inv) == -1.0f) ? PTB_NODE_NONE : 0;
    }
    template <class STACK>
    __device__ __forceinline__ void step(const BvhDev& bvh, STACK& st) {
        if (node == PTB_NODE_NONE) node = st.get(--sp);
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
            bool hit_l = (t_l != -1.0f && t_l < h.t && t_l <= ub);
            bool hit_r = (t_r != -1.0f && t_r < h.t && t_r <= ub);
            if (hit_l && hit_r) {
                bool left_first = t_l < t_r;
                st.set(sp++, left_first ? right : left);
                node = left_first ? left : right;
            } else if (hit_l) node = left;
            else if (hit_r) node = right;
            else node = PTB_NODE_NONE;
        }
        if (sp >= PTB_EXT_STACK) { node = PTB_NODE_NONE; sp = 0; }   // src/kernels/ray_cast.cu:106
    }
    template <class STACK>
    __device__ __forceinline__ Hit finish(const BvhDev&, STACK&) { return h; }
};

// WalkBq: the same binary tree through the 32-byte quantised records (BvhDev::bnode), near child first.  Exactness as in
// trace_closest_wide(): boxes are conservative, a hit is accepted through the reference's test on the leaf's fp32 box, and a
// result that could depend on the reference's visiting order is re-traced in that order (finish()).
struct WalkBq {
    Vec3 o, d;
    float iqx, iqy, iqz, bx, by, bz;          // t(q) = fma(32768 + q, iq, b)  (WideRay)
    unsigned near_x, near_y, near_z;
    float t, u, v, second;                    // best hit so far (t = FLT_MAX: none) and the nearest other accepted candidate
    int leaf;                                 // sorted leaf position of the best hit
    int node, sp;
    bool bad;
    __device__ __forceinline__ void reset() { node = PTB_NODE_NONE; sp = 0; }
    __device__ __forceinline__ bool busy() const { return node != PTB_NODE_NONE || sp > 0; }
    __device__ __forceinline__ void init(const BvhDev& bvh, Vec3 o_, Vec3 d_) {
        WideRay r = wide_setup(bvh, o_, d_);
        o = o_; d = d_;
        iqx = r.iqx; iqy = r.iqy; iqz = r.iqz; bx = r.bx; by = r.by; bz = r.bz;
        near_x = r.near_x; near_y = r.near_y; near_z = r.near_z;
        t = FLT_MAX; u = 0.0f; v = 0.0f; second = FLT_MAX; leaf = -1;
        sp = 0;
        bad = !r.ok;
        node = bad ? PTB_NODE_NONE : 0;
    }
    template <class STACK>
    __device__ __forceinline__ void step(const BvhDev& bvh, STACK& st) {
        if (node == PTB_NODE_NONE) node = st.get(--sp);
        const float limit = t * PTB_WIDE_SLACK;           // +inf while there is no hit
        if (node >= 0) {
            F8 q = ldg256(reinterpret_cast<const float4*>(bvh.bnode + 2 * (size_t)node));
            const unsigned xl = __float_as_uint(q.a.x), xr = __float_as_uint(q.a.y), yl = __float_as_uint(q.a.z), yr = __float_as_uint(q.a.w);
            const unsigned zl = __float_as_uint(q.b.x), zr = __float_as_uint(q.b.y);
            const int left = __float_as_int(q.b.z), right = __float_as_int(q.b.w);
            const unsigned far_x = near_x ^ 0x0220u, far_y = near_y ^ 0x0220u, far_z = near_z ^ 0x0220u;
            float tn_l = fmaxf(fmaxf(fmaf(wide_plane(xl, near_x), iqx, bx), fmaf(wide_plane(yl, near_y), iqy, by)),
                               fmaxf(fmaf(wide_plane(zl, near_z), iqz, bz), 0.0f));
            float tf_l = fminf(fminf(fmaf(wide_plane(xl, far_x), iqx, bx), fmaf(wide_plane(yl, far_y), iqy, by)),
                               fminf(fmaf(wide_plane(zl, far_z), iqz, bz), limit));
            float tn_r = fmaxf(fmaxf(fmaf(wide_plane(xr, near_x), iqx, bx), fmaf(wide_plane(yr, near_y), iqy, by)),
                               fmaxf(fmaf(wide_plane(zr, near_z), iqz, bz), 0.0f));
            float tf_r = fminf(fminf(fmaf(wide_plane(xr, far_x), iqx, bx), fmaf(wide_plane(yr, far_y), iqy, by)),
                               fminf(fmaf(wide_plane(zr, far_z), iqz, bz), limit));
            const bool hit_l = tn_l <= tf_l, hit_r = tn_r <= tf_r;
            if (hit_l && hit_r) {
                const bool left_first = tn_l <= tn_r;
                if (sp < PTB_WIDE_STACK) st.set(sp++, left_first ? right : left);
                else bad = true;
                node = left_first ? left : right;
            } else if (hit_l) node = left;
            else if (hit_r) node = right;
            else node = PTB_NODE_NONE;
        } else {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float tu, tv;
            float tt = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), o, d, tu, tv);
            if (tt > PTB_EPSILON && tt < limit) {
                const Vec3 inv(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));      // 1.0f / ray.direction, as the reference has it
                float tb = leaf_box_test(a, b, c, o, inv);
                if (tb != -1.0f && tb < limit) {
                    if (tt < t && tb < t) {                          // the reference's acceptance, src/kernels/ray_cast.cu:68,93-94
                        second = fminf(second, t);
                        t = tt; u = tu; v = tv;
                        leaf = ~node;
                    } else {
                        second = fminf(second, tt);
                    }
                }
            }
            node = PTB_NODE_NONE;
        }
        if (bad) { node = PTB_NODE_NONE; sp = 0; }
    }
    template <class STACK>
    __device__ __forceinline__ Hit finish(const BvhDev& bvh, STACK& st) {
        const bool ambiguous = bad || ((leaf >= 0) && (second <= t * PTB_WIDE_SLACK));
        if (ambiguous) {     // about one ray in 10^5: the reference-order walk decides
            WalkBinary w;
            w.init(bvh, o, d);
            while (w.busy()) w.step(bvh, st);
            return w.h;
        }
        Hit h;
        h.prim = -1; h.mat = -1; h.t = t; h.u = u; h.v = v;
        if (leaf >= 0) {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)leaf;
            h.prim = __float_as_int(__ldg(tp).w);
            h.mat = __float_as_int(__ldg(tp + 1).w);
        }
        return h;
    }
};

// WalkWide: trace_closest_wide() cut at the loop body (closest hit on the 4-wide quantised tree, nearest child first); a result
// that could depend on the reference's visiting order is re-traced in that order by finish(), as trace_closest_exact() does.
struct WalkWide {
    WideRay r;
    float t, u, v, second;                    // best hit (t = FLT_MAX: none), nearest other accepted candidate
    int leaf;
    int node, sp;                             // second < 0: the walk cannot vouch for its result (bad reciprocal, stack overflow)
    __device__ __forceinline__ void reset() { node = PTB_NODE_NONE; sp = 0; }
    __device__ __forceinline__ bool busy() const { return node != PTB_NODE_NONE || sp > 0; }
    __device__ __forceinline__ void init(const BvhDev& bvh, Vec3 o, Vec3 d) {
        r = wide_setup(bvh, o, d);
        t = FLT_MAX; u = 0.0f; v = 0.0f; second = FLT_MAX; leaf = -1;
        sp = 0;
        if (!r.ok) second = -1.0f;
        node = r.ok ? 0 : PTB_NODE_NONE;
    }
    // EXACT: the leaf-box test inside the walk (a triangle the reference cannot see is never a candidate); the scheduler's loop runs
    // EXACT = false and finish() verifies the winner.
    template <class STACK, bool EXACT = false>
    __device__ __forceinline__ void step(const BvhDev& bvh, STACK& st) {
        if (node == PTB_NODE_NONE) node = st.get(--sp);
        const float limit = t * PTB_WIDE_SLACK;
        if (node >= 0) {
            const uint4* np = bvh.wnode + 4 * (size_t)node;
            F8 a = ldg256(reinterpret_cast<const float4*>(np)), b = ldg256(reinterpret_cast<const float4*>(np + 2));
            const uint4 qx = make_uint4(__float_as_uint(a.a.x), __float_as_uint(a.a.y), __float_as_uint(a.a.z), __float_as_uint(a.a.w));
            const uint4 qy = make_uint4(__float_as_uint(a.b.x), __float_as_uint(a.b.y), __float_as_uint(a.b.z), __float_as_uint(a.b.w));
            const uint4 qz = make_uint4(__float_as_uint(b.a.x), __float_as_uint(b.a.y), __float_as_uint(b.a.z), __float_as_uint(b.a.w));
            int lnk[4] = {__float_as_int(b.b.x), __float_as_int(b.b.y), __float_as_int(b.b.z), __float_as_int(b.b.w)};
            float key[4];
            wide_test<true>(r, qx, qy, qz, limit, key);
            // nearest child first; the order of the other three decides almost nothing (3 % of the popped nodes could be
            // culled by their entry distance, profiles/r02_raycast_notes.md), so the network stops once key[0] is the minimum
            PTB_WIDE_CSWAP(0, 1) PTB_WIDE_CSWAP(2, 3) PTB_WIDE_CSWAP(0, 2)
#ifdef PTB_WIDE_FULLSORT
            PTB_WIDE_CSWAP(1, 3) PTB_WIDE_CSWAP(1, 2)
#endif
            // the column has PTB_WIDE_PUSHES spare entries: one overflow check per visit instead of one per push
#pragma unroll
            for (int k = 3; k >= 1; k--) {
                if (key[k] != FLT_MAX) { st.set(sp, lnk[k]); sp++; }
            }
            if (sp > PTB_WIDE_STACK) { sp = PTB_WIDE_STACK; second = -1.0f; }   // entries dropped: finish() re-walks
            node = (key[0] != FLT_MAX) ? lnk[0] : PTB_NODE_NONE;
        }
        if (node < 0) {                                             // a popped leaf, or the nearest child just chosen: same step
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float tu, tv;
            float tt = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), r.o, r.d, tu, tv);
            // The reference's test on the leaf's own box (what decides whether it ever sees this triangle) waits until finish(): one
            // lane in 32 needs it here, and it is as long as half a box visit.  Until then the nearest triangle hit is provisional.
            if (tt > PTB_EPSILON && tt < limit) {
                bool seen = true, best = tt < t;
                if (EXACT) {
                    const float tb = leaf_box_test(a, b, c, r.o, r.inv);
                    seen = tb != -1.0f && tb < limit;
                    best = best && tb < t;                           // the reference's acceptance, src/kernels/ray_cast.cu:68,93-94
                }
                if (seen) {
                    if (best) {
                        second = fminf(second, t);
                        t = tt; u = tu; v = tv;
                        leaf = ~node;
                    } else {
                        second = fminf(second, tt);
                    }
                }
            }
            node = PTB_NODE_NONE;
        }
    }
    template <class STACK>
    __device__ __forceinline__ Hit finish(const BvhDev& bvh, STACK& st) {
        // The winner stands if the reference reaches its leaf: its own box passes the reference's test with an entry distance below
        // every best-so-far the reference can hold when it gets there (all of them beyond t * slack, else `second` says so).
        bool verified = leaf < 0;
        int prim = -1, mat = -1;
        if (second >= 0.0f && leaf >= 0) {
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)leaf;
            F8 ab = ldg256(tp);
            const float4 c = __ldg(tp + 2);
            const float tb = leaf_box_test(ab.a, ab.b, c, r.o, r.inv);
            prim = __float_as_int(ab.a.w);
            mat = __float_as_int(ab.b.w);
            verified = tb != -1.0f && tb < t * PTB_WIDE_SLACK;
        }
        if (second >= 0.0f && !verified) {
            // A winner that fails was never seen by the reference, and what it culled on the way is unknown (one ray in 200 on the
            // bench scene: grazing hits on flat, axis-aligned triangles): this walk again, with the test inside it.
            t = FLT_MAX; u = 0.0f; v = 0.0f; second = FLT_MAX; leaf = -1;
            sp = 0;
            node = 0;
            while (busy()) step<STACK, true>(bvh, st);
            verified = true;
            if (leaf >= 0) {
                const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)leaf;
                prim = __float_as_int(__ldg(tp).w);
                mat = __float_as_int(__ldg(tp + 1).w);
            }
        }
        // overflow / unusable reciprocal: the whole reference-order walk; another candidate within the slack of a verified winner: the
        // reference's order decides between them, and nothing beyond them matters (WalkBinary::ub)
        const bool near_tie = second >= 0.0f && leaf >= 0 && second <= t * PTB_WIDE_SLACK;
        if (second < 0.0f || near_tie) {
            WalkBinary w;
            w.init(bvh, r.o, r.d);
            if (near_tie) w.ub = t * (PTB_WIDE_SLACK * PTB_WIDE_SLACK);
            while (w.busy()) w.step(bvh, st);
            return w.h;
        }
        Hit h;
        h.prim = prim; h.mat = mat; h.t = t; h.u = u; h.v = v;
        return h;
    }
};

struct WalkAnyWide {
    WideRay r;
    float tmax;
    int node, sp, result;                     // result 2: a triangle was hit, its leaf `cand` still owes the reference's box test (confirm())
    int cand;
    __device__ __forceinline__ void reset() { node = PTB_NODE_NONE; sp = 0; }
    __device__ __forceinline__ bool busy() const { return result == 0 && (node != PTB_NODE_NONE || sp > 0); }
    __device__ __forceinline__ void init(const BvhDev& bvh, Vec3 o, Vec3 d, float tmax_) {
        r = wide_setup(bvh, o, d);
        tmax = tmax_;
        sp = 0;
        result = r.ok ? 0 : -1;
        node = r.ok ? 0 : PTB_NODE_NONE;
    }
    template <class STACK>
    __device__ __forceinline__ void step(const BvhDev& bvh, STACK& st) {
        if (node == PTB_NODE_NONE) node = st.get(--sp);
        if (node >= 0) {
            const uint4* np = bvh.wnode + 4 * (size_t)node;
            F8 a = ldg256(reinterpret_cast<const float4*>(np)), b = ldg256(reinterpret_cast<const float4*>(np + 2));
            const uint4 qx = make_uint4(__float_as_uint(a.a.x), __float_as_uint(a.a.y), __float_as_uint(a.a.z), __float_as_uint(a.a.w));
            const uint4 qy = make_uint4(__float_as_uint(a.b.x), __float_as_uint(a.b.y), __float_as_uint(a.b.z), __float_as_uint(a.b.w));
            const uint4 qz = make_uint4(__float_as_uint(b.a.x), __float_as_uint(b.a.y), __float_as_uint(b.a.z), __float_as_uint(b.a.w));
            const int lnk[4] = {__float_as_int(b.b.x), __float_as_int(b.b.y), __float_as_int(b.b.z), __float_as_int(b.b.w)};
            float t[4];
            wide_test(r, qx, qy, qz, tmax, t);
            node = PTB_NODE_NONE;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (t[k] >= 0.0f) {
                    if (node == PTB_NODE_NONE) node = lnk[k];
                    else { st.set(sp, lnk[k]); sp++; }
                }
            }
            if (sp > PTB_WIDE_STACK) result = -1;                       // PTB_WIDE_PUSHES spare entries took the excess
        }
        if (node < 0 && result == 0) {                              // a popped leaf, or the first child just chosen: same step
            const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)(~node);
            F8 ab = ldg256(tp);
            float4 a = ab.a, b = ab.b, c = __ldg(tp + 2);
            float u, v;
            float t = triangle_test(Vec3(a.x, a.y, a.z), Vec3(b.x, b.y, b.z), Vec3(c.x, c.y, c.z), r.o, r.d, u, v);
            if (t > PTB_EPSILON && t < __fsub_rn(tmax, PTB_EPSILON)) { result = 2; cand = ~node; }
            node = PTB_NODE_NONE;
        }
    }
    // A lane that stopped at a triangle hit (result 2), in the converged part of the scheduler's loop: the reference reaches that
    // triangle's leaf iff the leaf's own box passes its test.  Passed: occluded (1).  Failed: the reference never saw the triangle;
    // the lane's stack is untouched, so it simply goes on with its walk (result 0, busy() again unless nothing is left).
    __device__ __forceinline__ void confirm(const BvhDev& bvh) {
        if (result != 2) return;
        const float4* tp = bvh.tri + PTB_TRI_STRIDE * (size_t)cand;
        F8 ab = ldg256(tp);
        const float4 c = __ldg(tp + 2);
        const float tb = leaf_box_test(ab.a, ab.b, c, r.o, r.inv);
        result = (tb != -1.0f && tmax > tb) ? 1 : 0;
    }
};

// The two walks behind one call: wide first, the reference-order binary walk where the wide one cannot vouch for the result.
__device__ __forceinline__ Hit trace_closest_exact(const BvhDev& bvh, Vec3 o, Vec3 d, int* stack, int stride) {
    WideRay r = wide_setup(bvh, o, d);
    bool ambiguous = true;
    Hit h;
    PTB_WDBG(0);
    if (!r.ok) PTB_WDBG(1);
    if (r.ok) h = trace_closest_wide<PTB_WIDE_STACK>(bvh, r, stack, stride, ambiguous);
    if (ambiguous) { PTB_WDBG(2); h = trace_closest(bvh, o, d, stack, stride); }
    return h;
}
template <int STACK>
__device__ __forceinline__ int trace_any_exact(const BvhDev& bvh, Vec3 o, Vec3 d, float tmax, int* stack, int stride) {
    WideRay r = wide_setup(bvh, o, d);
    int res = -1;
    if (r.ok) res = trace_any_wide<STACK>(bvh, r, tmax, stack, stride);
    if (res < 0) res = trace_any<STACK>(bvh, o, d, tmax, stack, stride);
    return res;
}

}  // namespace ptb
