// common.cuh -- device-side types shared by every kernel of the wavefront core.
//
// Arithmetic contract.  Radiance parity with the reference (per-pixel rel. err <= 1e-4 at 1 spp)
// needs the same DECISIONS per path, so the float expressions here keep the reference's shape:
// `Vec3` mirrors glm::vec3 component-wise operators, and the three primitives whose contraction
// was read off the reference's sm_100 SASS are pinned with explicit intrinsics:
//   dot(a,b)   = fma(a.z,b.z, fma(a.x,b.x, a.y*b.y))      GLM (x*x' + y*y') + z*z' after nvcc -fmad
//   cross(a,b) = fma(a.y,b.z, -(b.y*a.z)), ...            first product fused, second rounded
//   normalize  = v * (1 / sqrt(dot(v,v)))                 IEEE sqrt and reciprocal
// (reference: external/include/glm/detail/func_geometric.inl:65-72,134-142,154-159; SASS of
// src/intersections.cu:10-32 and src/kernels/ray_gen.cu:60-63.)
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdint>

#include "../../../include/ptb200.h"

namespace ptb {

#define PTB_PI 3.1415926535897932384626422832795028841971f      /* src/common/utilities.h:12 */
#define PTB_TWO_PI 6.2831853071795864769252867665590057683943f  /* :13 */
#define PTB_EPSILON 0.001f                                       /* :15 */
#define PTB_INV_PI 0.31830988618f                                /* :16 */
#define PTB_INV_TWO_PI 0.15915494309189533577f                   /* :17 */
#define PTB_PDF_DIRAC_DELTA 1e10f                                /* :18 */

enum { MAT_PBR = 0, MAT_DIFFUSE = 1, MAT_REFLECT = 2, MAT_REFRACT = 3 };

// Like glm::tvec3 (detail/type_vec3.hpp) this type has user-provided default/copy constructors and
// assignment.  That is deliberate: it makes the type non-trivially-copyable, so a `Vec3` passed "by
// value" to a non-inlined function travels by hidden reference, as glm::vec3 does in the reference's
// evalBSDF / pdfBSDF -- and the loads-vs-registers difference changes nvcc's FMA choices.
struct Vec3 {
    float x, y, z;
    __host__ __device__ Vec3() : x(0.0f), y(0.0f), z(0.0f) {}
    __host__ __device__ Vec3(const Vec3& v) : x(v.x), y(v.y), z(v.z) {}
    __host__ __device__ Vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    __host__ __device__ explicit Vec3(float a) : x(a), y(a), z(a) {}
    __host__ __device__ Vec3& operator=(const Vec3& v) { x = v.x; y = v.y; z = v.z; return *this; }
};
// Operators take their operands by const reference, exactly like GLM's tvec3 operators
// (detail/type_vec3.inl): argument passing changes the order in which nvcc sees the operations and
// with it which product of `a*b + c*d` gets fused (checked on PTX, tools/ptx_alpha.py).
typedef const Vec3& V3R;
__device__ __forceinline__ Vec3 operator+(V3R a, V3R b) { return Vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ Vec3 operator-(V3R a, V3R b) { return Vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ Vec3 operator-(V3R a) { return Vec3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ Vec3 operator*(V3R a, V3R b) { return Vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ Vec3 operator*(V3R a, const float& s) { return Vec3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ Vec3 operator*(const float& s, V3R a) { return Vec3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ Vec3 operator/(V3R a, const float& s) { return Vec3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ Vec3 operator+(V3R a, const float& s) { return Vec3(a.x + s, a.y + s, a.z + s); }
__device__ __forceinline__ Vec3 make_v3(float4 f) { return Vec3(f.x, f.y, f.z); }

// Two flavours of the GLM geometric functions:
//  * dot3 / cross3 / length3 / normalize3 are written exactly like GLM 0.9.6.3 writes them
//    (detail/func_geometric.inl:65-72,134-142,154-159) and left to nvcc's FMA contraction.  Which
//    product ends up rounded depends on the surrounding expression (e.g. dot(N, -wo) contracts
//    differently from dot(N, wo) -- read off the reference's SASS), so the shading code, which must
//    land on the reference's contraction, uses these and mirrors the reference's expression shapes.
//  * *_pinned spell out the contraction of the reference's intersection routines (SURVEY.md
//    Appendix B.3) with explicit intrinsics; the RayCast stage uses them.
__device__ __forceinline__ float dot3(V3R a, V3R b) {
    Vec3 tmp(a.x * b.x, a.y * b.y, a.z * b.z);
    return tmp.x + tmp.y + tmp.z;
}
__device__ __forceinline__ Vec3 cross3(V3R x, V3R y) {
    return Vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
__device__ __forceinline__ float length3(V3R v) { return sqrtf(dot3(v, v)); }
__device__ __forceinline__ Vec3 normalize3(V3R v) { return v * (1.0f / sqrtf(dot3(v, v))); }
__device__ __forceinline__ float dot3_pinned(Vec3 a, Vec3 b) {
    return __fmaf_rn(a.z, b.z, __fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y)));
}
__device__ __forceinline__ Vec3 cross3_pinned(Vec3 a, Vec3 b) {
    return Vec3(__fmaf_rn(a.y, b.z, -__fmul_rn(b.y, a.z)),
                __fmaf_rn(a.z, b.x, -__fmul_rn(b.z, a.x)),
                __fmaf_rn(a.x, b.y, -__fmul_rn(b.x, a.y)));
}
// glm::min / glm::max / glm::clamp (detail/func_common.inl:409-456): plain comparisons, so a NaN
// argument propagates the way it does in the reference (fminf/fmaxf would drop it).
__device__ __forceinline__ float glm_min(float x, float y) { return x < y ? x : y; }
__device__ __forceinline__ float glm_max(float x, float y) { return x > y ? x : y; }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return glm_min(glm_max(x, lo), hi); }
// glm::mix (detail/func_common.inl:74-105): x + a * (y - x)
__device__ __forceinline__ float mixf(float x, float y, float a) { return x + a * (y - x); }
__device__ __forceinline__ Vec3 mix3(V3R x, V3R y, const float& a) { return x + a * (y - x); }
__device__ __forceinline__ Vec3 operator/(V3R a, V3R b) { return Vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
// glm::reflect (detail/func_geometric.inl:176-180): I - N * dot(N, I) * 2
__device__ __forceinline__ Vec3 reflect3(V3R I, V3R N) { return I - N * dot3(N, I) * 2.0f; }

// ---- RNG (src/rng.h:6-27) ---------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t wang_hash(uint32_t seed) {
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}
__device__ __forceinline__ float rand_float(uint32_t& state) {
    state ^= state << 13;
    state ^= state >> 17;
    state ^= state << 5;
    return state * 2.3283064365386963e-10f;
}

// ---- scene on the device ------------------------------------------------------------------------
struct MeshDev {                 // MeshData, src/scene_structs.h:134-149
    const float4* pos;
    const float4* nor;
    const float4* tangent;
    const float2* uv;
    const int4* indices_matid;
    const float4* nor_geom;
    int num_vertices, num_triangles;
};

struct LightDev {                // LightData, src/scene_structs.h:76-81
    const int* tri_idx;
    const float* cdf;
    int num_lights;
    float total_area;
};

// LBVH in the reference's layout (LBVHData, src/scene_structs.h:151-169) -- what parity reads.
struct LbvhRef {
    float4* aabb_min;
    float4* aabb_max;
    float4* centroid;
    int* primitive_indices;
    unsigned long long* morton_codes;
    int2* child_nodes;
    int* parent;
    int* escape_indices;
    float4* world_aabb;          // [2] min, max (device copy; the reference keeps these on the host)
};

// Traversal layout derived from it (DESIGN.md "compressed node layout"): one 64-byte record per
// internal node holding BOTH children's boxes and links, so a node visit is two 32-byte loads
// (sm_100 LDG.256, one full sector per lane per instruction) from one 64-byte record instead of
// the reference's int2 + 4 scattered float4 (src/kernels/ray_cast.cu:78-88).  Leaves carry their
// three vertices inline in a 64-byte record as well (two sectors, 32-byte aligned), which removes
// the primitive_indices -> indices_matid -> pos chain (:55-63).  RayCast is bound by the L1 data
// pipe (wavefronts), not by bytes, so records are laid out per sector, not packed.
//   node[4*i+0] = (lmin.x, lmin.y, lmax.x, lmax.y)     node[4*i+1] = (rmin.x, rmin.y, rmax.x, rmax.y)
//   node[4*i+2] = (lmin.z, lmax.z, rmin.z, rmax.z)     node[4*i+3] = (bits(left link), bits(right link), 0, 0)
//   link >= 0 : internal node index, link < 0 : ~(sorted leaf position); pairs as node_test() (raycast.cuh) consumes them
//   tri[4*k+0]  = (v0.xyz, bits(triangle index)), tri[4*k+1] = (v1.xyz, bits(material id)),
//   tri[4*k+2]  = (v2.xyz, 0),                    tri[4*k+3] = 0
#define PTB_TRI_STRIDE 4
struct BvhDev {
    const float4* node;
    const float4* tri;
    float4 root_min, root_max;   // AABB of node 0 (the root test of src/kernels/ray_cast.cu:40-47)
    int num_triangles;
    // 4-wide collapse of the same tree with 15-bit boxes (wide.cuh); wcell = (root_max - root_min) / PTB_WIDE_GRID
    const uint4* wnode;
    float wcell[3];
    // the binary tree itself with both children's boxes on the same 15-bit grid: ONE 32-byte sector per node visit
    //   bnode[2*i+0] = (x_left, x_right, y_left, y_right)   each word lo | hi << 16 (15-bit steps, rounded outwards)
    //   bnode[2*i+1] = (z_left, z_right, left link, right link)
    const uint4* bnode;
};
#define PTB_EXT_STACK_DEPTH 32       /* entries the reference's closest-hit walk stacks before it gives up (src/kernels/ray_cast.cu:35,106) */
#define PTB_WIDE_GRID 32760.0f      /* quantisation steps across the scene box; +-2 steps of margin stay below 2^15 */
#define PTB_WIDE_EMPTY 0x7fffffff   /* link of an unused child slot */

struct TexturesDev {
    const cudaTextureObject_t* handles;
    int env_tex_id, pdf_map_id;  // EnvMapAliasTable ids, src/scene_structs.h:67-74
};

// ---- path pool and queues -----------------------------------------------------------------------
// PathState (src/scene_structs.h:97-117).  Slot == index into this renderer's owned pixels; the
// GLOBAL pixel index (what seeds the RNG and what the reference calls pixel_idx) is derived from
// the slot by TileMap, so it is not stored.
// Storage is three 32-byte records per slot, one per WRITER, so that every store of a stage is a whole 32-byte sector (a
// partial-sector store costs a fill read plus a whole-sector write-back; the round-1 layout had three of them per ray in
// RayCast and two per path in Shade -- 2.8x the algorithmic DRAM bytes).  The ray and the state record of a slot are the two
// halves of ONE 64-byte block: HBM moves 64 bytes per access (a scattered 32-byte record costs 64 bytes of DRAM traffic either
// way: measured, profiles/r02_raycast_notes.md section 3), RayGen and Shade write both halves, RayCast + Logic and Shade read both.
//   ray  = rs[4*slot]   = ray_ori (xyz, 0)      rs[4*slot+1] = (dir.xyz, FLT_MAX)        written by RayGen / Shade, read by RayCast
//   state= rs[4*slot+2] = throughput_pdf        rs[4*slot+3] = (bits(remaining_bounces), bits(rng_state), 0, 0)
//                                                                                        written by RayGen / Shade (Logic: remaining = -1)
//   hit  [2*slot] = (t, u, v, bits(prim))       hit  [2*slot+1] = (bits(mat), 0, 0, 0)   written by RayCast (t = -1, prim = -1: miss)
// The reference's fields map onto them as: ray_dir_dist.w after a trace = hit.t; hit_normal = (hit.u, hit.v, 0, 0); hit_geom_id =
// hit.prim; material_id = hit.mat (ptb_get_path_field assembles these views).  Members with operator[] keep the reference's names.
struct PoolF4 {                  // one float4 of every 64-byte ray + state block
    float4* base;
    __device__ __forceinline__ float4& operator[](int i) const { return base[4 * (size_t)i]; }
};
template <typename T>
struct PoolI {                   // one 32-bit word of every 64-byte ray + state block
    T* base;
    __device__ __forceinline__ T& operator[](int i) const { return base[16 * (size_t)i]; }
};
// One 32-byte store / load of a whole record (sm_100 STG.256 / LDG.256): a full sector, no partial write.
__device__ __forceinline__ void stg256(float4* p, float4 a, float4 b) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
}
__device__ __forceinline__ void ldg256_rw(const float4* p, float4& a, float4& b) {     // data other kernels of the frame rewrite: no .nc
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p) : "memory");
}
struct HitRecord { float t, u, v; int prim, mat; };
struct PathPool {
    PoolF4 ray_ori;
    PoolF4 ray_dir_dist;         // .w is FLT_MAX (written with the ray); the hit distance lives in the hit record
    PoolF4 throughput_pdf;
    PoolI<int> remaining_bounces;
    PoolI<uint32_t> rng_state;
    float4* rs;                  // [4 * P]: ray (2 float4) + state (2 float4) per slot
    float4* hit;                 // [2 * P]
    __device__ __forceinline__ float4* ray_record(int slot) const { return rs + 4 * (size_t)slot; }
    __device__ __forceinline__ float4* state_record(int slot) const { return rs + 4 * (size_t)slot + 2; }
    __device__ __forceinline__ void store_hit(int slot, float t, float u, float v, int prim, int mat) const {
        stg256(hit + 2 * (size_t)slot, make_float4(t, u, v, __int_as_float(prim)), make_float4(__int_as_float(mat), 0.0f, 0.0f, 0.0f));
    }
    __device__ __forceinline__ HitRecord load_hit(int slot) const {
        float4 a, b;
        ldg256_rw(hit + 2 * (size_t)slot, a, b);
        HitRecord h;
        h.t = a.x; h.u = a.y; h.v = a.z; h.prim = __float_as_int(a.w); h.mat = __float_as_int(b.x);
        return h;
    }
    __device__ __forceinline__ void store_state(int slot, float4 tp, int remaining, uint32_t rng) const {
        stg256(state_record(slot), tp, make_float4(__int_as_float(remaining), __uint_as_float(rng), 0.0f, 0.0f));
    }
    __device__ __forceinline__ void load_state(int slot, float4& tp, int& remaining, uint32_t& rng) const {
        float4 b;
        ldg256_rw(state_record(slot), tp, b);
        remaining = __float_as_int(b.x); rng = __float_as_uint(b.y);
    }
};

struct ShadowQueueDev {          // ShadowQueue, src/scene_structs.h:121-131
    float4* ray_ori_tmax;
    float4* ray_dir;
    float4* radiance;
    int* slot;                   // path slot (== local pixel) the radiance is added to
};

// Per-bounce counters, device resident for the whole frame (the reference reads six of them
// back to the host every bounce, src/pathtrace_wavefront.cu:286-343).
enum { CTR_EXT = 0, CTR_SHADOW = 1, CTR_PBR = 2, CTR_DIFFUSE = 3, CTR_REFLECT = 4, CTR_REFRACT = 5, CTR_STRIDE = 8 };

struct TileMap {                 // rank r owns rows y with (y / stripe_rows) % world == r
    int width, height;
    int rank, world, stripe_rows;
    int owned_pixels;
    __host__ __device__ int global_pixel(int slot) const {
        if (world <= 1) return slot;
        int local_row = slot / width;
        int x = slot - local_row * width;
        int stripe_local = local_row / stripe_rows;
        int row_in = local_row - stripe_local * stripe_rows;
        int y = (stripe_local * world + rank) * stripe_rows + row_in;
        return y * width + x;
    }
};

struct CameraDev {               // Camera, src/scene_structs.h:16-26 (members the kernels read)
    int res_x, res_y;
    float pos[3], view[3], up[3], right[3];
    float pixel_length[2];
};

// Queue position -> path slot.  The reference fills its first queue in pixel order (queue[i] = i,
// src/kernels/ray_gen.cu:76), so a warp of RayCast gets a 32x1 run of pixels.  Queue ORDER is unobservable
// (every per-path result is independent of its position), so the first queue is laid out in 8x4 pixel tiles
// instead: a warp's rays span a quarter of the width, stay on nearby surfaces for the following bounces, and
// touch fewer distinct BVH records per load -- which is what bounds RayCast (L1 wavefronts).
// Pixels outside the largest (8k x 4m) region keep their linear order.
__device__ __forceinline__ int tiled_slot(int i, int width, int rows) {
    const int w8 = width & ~7, h4 = rows & ~3;
    const int tiled = w8 * h4;
    if (i >= tiled) {
        int j = i - tiled;                                   // right strip (rows < h4), then the bottom rows
        int strip = (width - w8) * h4;
        if (j < strip) { int sw = width - w8; return (j / sw) * width + w8 + (j % sw); }
        return h4 * width + (j - strip);
    }
    int tile = i >> 5, lane = i & 31;
    int tiles_per_row = w8 >> 3;
    int ty = tile / tiles_per_row, tx = tile - ty * tiles_per_row;
    return (ty * 4 + (lane >> 3)) * width + tx * 8 + (lane & 7);
}

}  // namespace ptb
