// lbvh.cu -- parallel LBVH build for sm_100a, bit-exact against the reference BuildLBVH
// (src/bvh.cu:456-516) but restructured:
//   * per-primitive bounds and the world-bounds reduction are ONE kernel (the reference runs a
//     kernel plus two thrust::reduce calls that return to the host, src/bvh.cu:13-51);
//   * the sort key is the 30-bit Morton code alone.  The reference sorts the 64-bit key
//     (code << 32) | tid (src/bvh.cu:95,103); since the input is in ascending tid order, a STABLE
//     sort on the code gives exactly that order in 4 radix passes of 8 bits instead of 8;
//   * the radix sort is hand-written (per-tile digit counts -> per-digit row scan -> stable
//     warp-match scatter), no thrust/CUB;
//   * Karras hierarchy + leaf setup are one kernel; refit is the usual bottom-up walk with
//     arrival flags; a last kernel emits the 64-byte traversal nodes and inline triangles.
// min/max reductions and the Morton arithmetic contain no contractible a*b+c, so every array
// is reproducible bit for bit.
#include "lbvh.cuh"

namespace ptb {

static constexpr int kSortThreads = 256;
static constexpr int kSortItems = 16;
static constexpr int kSortTile = kSortThreads * kSortItems;   // 4096 keys per CTA

// ---- order-preserving float <-> uint (for atomicMin/atomicMax on floats) ---------------------
// Total order identical to PTX min.f32/max.f32 on non-NaN inputs, including -0 < +0.
__device__ __forceinline__ unsigned int enc_f(float f) {
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float dec_f(unsigned int u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void k_init(LbvhBuffers b) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int n = b.num_triangles;
    if (i < 2 * n) {
        b.ref.parent[i] = -1;             // cudaMemset(parent, -1)          src/bvh.cu:465
        b.ref.escape_indices[i] = -1;     // cudaMemset(escape_indices, -1)  src/bvh.cu:502
    }
    if (i < n) b.flags[i] = 0;
    if (i < 4) b.world_enc[i] = enc_f(FLT_MAX);         // thrust::reduce init, src/bvh.cu:43
    else if (i < 8) b.world_enc[i] = enc_f(-FLT_MAX);   // :44
    if (i == 8) *b.max_depth = 0;
}

// ComputeAABB (src/bvh.cu:13-35) fused with ComputeWorldAABB (:37-51).
// A fixed grid strides over the triangles and keeps its bounds in registers: 6 atomics per BLOCK at the end (one per warp
// per launch of one thread per triangle was 1.9 M atomics on six words at 10 M triangles: 1.36 ms of a 4.3 ms build).
__global__ void __launch_bounds__(256) k_prim_bounds(LbvhBuffers b, const float4* __restrict__ pos,
                                                     const int4* __restrict__ indices_matid) {
    const int n = b.num_triangles;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int tid = blockIdx.x * blockDim.x + threadIdx.x; tid < n; tid += gridDim.x * blockDim.x) {
        int4 im = __ldg(&indices_matid[tid]);
        float4 v0 = __ldg(&pos[im.x]), v1 = __ldg(&pos[im.y]), v2 = __ldg(&pos[im.z]);
        float4 mn = make_float4(fminf(fminf(v0.x, v1.x), v2.x), fminf(fminf(v0.y, v1.y), v2.y),
                                fminf(fminf(v0.z, v1.z), v2.z), 1.0f);
        float4 mx = make_float4(fmaxf(fmaxf(v0.x, v1.x), v2.x), fmaxf(fmaxf(v0.y, v1.y), v2.y),
                                fmaxf(fmaxf(v0.z, v1.z), v2.z), 1.0f);
        b.ref.aabb_min[tid] = mn;
        b.ref.aabb_max[tid] = mx;
        // (max + min) * 0.5, w = (1 + 1) * 0.5 = 1                                 src/bvh.cu:32
        b.ref.centroid[tid] = make_float4(__fmul_rn(__fadd_rn(mx.x, mn.x), 0.5f), __fmul_rn(__fadd_rn(mx.y, mn.y), 0.5f),
                                          __fmul_rn(__fadd_rn(mx.z, mn.z), 0.5f), 1.0f);
        lo[0] = fminf(lo[0], mn.x); lo[1] = fminf(lo[1], mn.y); lo[2] = fminf(lo[2], mn.z);
        hi[0] = fmaxf(hi[0], mx.x); hi[1] = fmaxf(hi[1], mx.y); hi[2] = fmaxf(hi[2], mx.z);
    }
    // warp reduce, block reduce through shared memory, then one atomic per block per component (min / max: order-free)
    __shared__ float s_lo[3][8], s_hi[3][8];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        float l = lo[c], h = hi[c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
            h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
        }
        if ((threadIdx.x & 31) == 0) { s_lo[c][threadIdx.x >> 5] = l; s_hi[c][threadIdx.x >> 5] = h; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int c = threadIdx.x;
        float l = s_lo[c][0], h = s_hi[c][0];
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) { l = fminf(l, s_lo[c][w]); h = fmaxf(h, s_hi[c][w]); }
        atomicMin(&b.world_enc[c], enc_f(l));
        atomicMax(&b.world_enc[4 + c], enc_f(h));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {   // w of every per-primitive box is 1 -> reduce(init=+-FLT_MAX) gives 1
        atomicMin(&b.world_enc[3], enc_f(1.0f));
        atomicMax(&b.world_enc[7], enc_f(1.0f));
    }
}

// ExpandBits / Morton3D (src/bvh.cu:53-71)
__device__ __forceinline__ unsigned int expand_bits10(unsigned int v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ unsigned int quantize10(float x) {
    x = fminf(fmaxf(__fmul_rn(x, 1024.0f), 0.0f), 1023.0f);
    return (unsigned int)x;
}

// ComputeMortonCode (src/bvh.cu:74-96); also decodes the world box to floats for the ABI view.
__global__ void __launch_bounds__(256) k_morton(LbvhBuffers b) {
    int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= b.num_triangles) return;
    float wmin[3], ext[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        wmin[c] = dec_f(b.world_enc[c]);
        ext[c] = __fsub_rn(dec_f(b.world_enc[4 + c]), wmin[c]);
        if (ext[c] < 1e-6f) ext[c] = 1.0f;
    }
    if (tid == 0) {
        b.ref.world_aabb[0] = make_float4(dec_f(b.world_enc[0]), dec_f(b.world_enc[1]), dec_f(b.world_enc[2]), dec_f(b.world_enc[3]));
        b.ref.world_aabb[1] = make_float4(dec_f(b.world_enc[4]), dec_f(b.world_enc[5]), dec_f(b.world_enc[6]), dec_f(b.world_enc[7]));
    }
    float4 c = b.ref.centroid[tid];
    float sx = __fdiv_rn(__fsub_rn(c.x, wmin[0]), ext[0]);
    float sy = __fdiv_rn(__fsub_rn(c.y, wmin[1]), ext[1]);
    float sz = __fdiv_rn(__fsub_rn(c.z, wmin[2]), ext[2]);
    unsigned int code = (expand_bits10(quantize10(sx)) << 2) + (expand_bits10(quantize10(sy)) << 1) + expand_bits10(quantize10(sz));
    b.key_a[tid] = code;
    b.ref.primitive_indices[tid] = tid;   // src/bvh.cu:34
}

// ---- radix sort, one 8-bit pass = count + rowscan + scatter ----------------------------------
__global__ void __launch_bounds__(kSortThreads) k_sort_count(const unsigned int* __restrict__ keys, int n, int shift,
                                                             int* __restrict__ counts, int num_tiles) {
    __shared__ int hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * kSortTile;
#pragma unroll
    for (int r = 0; r < kSortItems; r++) {
        int i = base + r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&hist[(__ldg(&keys[i]) >> shift) & 255u], 1);
    }
    __syncthreads();
    counts[threadIdx.x * num_tiles + blockIdx.x] = hist[threadIdx.x];   // digit-major
}

// One CTA per digit: exclusive scan of that digit's per-tile counts, in place; row total out.
__global__ void __launch_bounds__(256) k_sort_rowscan(int* __restrict__ counts, int num_tiles, int* __restrict__ totals) {
    __shared__ int warp_sum[8];
    __shared__ int carry_s;
    int* row = counts + (size_t)blockIdx.x * num_tiles;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < num_tiles; base += 256) {
        int i = base + threadIdx.x;
        int v = i < num_tiles ? row[i] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_sum[warp] = incl;
        __syncthreads();
        int wbase = 0;
        for (int w = 0; w < warp; w++) wbase += warp_sum[w];
        int carry = carry_s;
        if (i < num_tiles) row[i] = carry + wbase + incl - v;
        __syncthreads();
        if (threadIdx.x == 255) carry_s = carry + wbase + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry_s;
}

// Stable scatter: keys of a tile are ranked warp by warp (each warp owns a contiguous
// 512-key segment, visited in 16 rounds of 32 consecutive keys) with __match_any_sync.
__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const unsigned int* __restrict__ keys_in, const int* __restrict__ vals_in,
                                                               unsigned int* __restrict__ keys_out, int* __restrict__ vals_out,
                                                               int n, int shift, const int* __restrict__ counts, int num_tiles,
                                                               const int* __restrict__ totals) {
    __shared__ int warp_cnt[8][256];
    __shared__ int digit_base[256];
    __shared__ int scan_tmp[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int w = 0; w < 8; w++) warp_cnt[w][threadIdx.x] = 0;
    __syncthreads();

    const int seg = blockIdx.x * kSortTile + warp * (kSortItems * 32);
    unsigned int key[kSortItems];
    int off[kSortItems];
    const unsigned int lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kSortItems; r++) {
        int i = seg + r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? __ldg(&keys_in[i]) : 0xffffffffu;
        unsigned int digit = valid ? ((key[r] >> shift) & 255u) : 256u;
        unsigned int peers = __match_any_sync(0xffffffffu, digit);
        int prev = 0;
        if (valid) prev = warp_cnt[warp][digit];
        __syncwarp();
        if (valid && (peers & lt_mask) == 0) warp_cnt[warp][digit] = prev + __popc(peers);
        __syncwarp();
        off[r] = prev + __popc(peers & lt_mask);
    }
    __syncthreads();
    // exclusive scan across the 8 warps per digit (thread d owns digit d)
    {
        int d = threadIdx.x, run = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            int c = warp_cnt[w][d];
            warp_cnt[w][d] = run;
            run += c;
        }
        // global base of digit d = exclusive scan of totals + this tile's row prefix
        int t = __ldg(&totals[d]);
        int incl = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (lane == 31) scan_tmp[warp] = incl;
        __syncthreads();
        int wbase = 0;
        for (int w = 0; w < warp; w++) wbase += scan_tmp[w];
        digit_base[d] = wbase + incl - t + __ldg(&counts[(size_t)d * num_tiles + blockIdx.x]);
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortItems; r++) {
        int i = seg + r * 32 + lane;
        if (i < n) {
            unsigned int digit = (key[r] >> shift) & 255u;
            int p = digit_base[digit] + warp_cnt[warp][digit] + off[r];
            keys_out[p] = key[r];
            vals_out[p] = __ldg(&vals_in[i]);
        }
    }
}

// After the sort: rebuild the reference's 64-bit keys (src/bvh.cu:95) and gather the leaves
// to [N, 2N) (SetupLeafNodes, src/bvh.cu:106-124).
__global__ void __launch_bounds__(256) k_leaves(LbvhBuffers b) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int n = b.num_triangles;
    if (i >= n) return;
    int prim = b.ref.primitive_indices[i];
    b.ref.morton_codes[i] = ((unsigned long long)b.key_a[i] << 32) | (unsigned long long)(unsigned int)prim;
    int leaf = n + i;
    b.ref.aabb_min[leaf] = b.ref.aabb_min[prim];
    b.ref.aabb_max[leaf] = b.ref.aabb_max[prim];
    b.ref.centroid[leaf] = b.ref.centroid[prim];
    b.ref.primitive_indices[leaf] = prim;
}

// Karras 2012 binary radix tree (GenerateHierarchy, src/bvh.cu:128-203).
__device__ __forceinline__ int lcp(const unsigned long long* __restrict__ keys, int n, unsigned long long ki, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(ki ^ __ldg(&keys[j])));
}

__global__ void __launch_bounds__(256) k_hierarchy(LbvhBuffers b) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int n = b.num_triangles;
    if (i >= n - 1) return;
    const unsigned long long* keys = b.ref.morton_codes;
    unsigned long long ki = __ldg(&keys[i]);
    int dir = (lcp(keys, n, ki, i + 1) - lcp(keys, n, ki, i - 1)) >= 0 ? 1 : -1;
    int delta_min = lcp(keys, n, ki, i - dir);
    int l_max = 2;
    while (lcp(keys, n, ki, i + l_max * dir) > delta_min) l_max *= 2;
    int l = 0;
    for (int t = l_max / 2; t >= 1; t /= 2)
        if (lcp(keys, n, ki, i + (l + t) * dir) > delta_min) l += t;
    int j = i + l * dir;
    int delta_node = lcp(keys, n, ki, j);
    int s = 0;
    for (int t = (l + 1) >> 1; t >= 1; t = (t == 1) ? 0 : ((t + 1) >> 1))
        if (lcp(keys, n, ki, i + (s + t) * dir) > delta_node) s += t;
    int gamma = i + s * dir + min(dir, 0);
    int left, right;
    int2 ch;
    if (min(i, j) == gamma) { left = gamma + n; ch.x = ~left; } else { left = gamma; ch.x = left; }
    if (max(i, j) == gamma + 1) { right = gamma + 1 + n; ch.y = ~right; } else { right = gamma + 1; ch.y = right; }
    b.ref.child_nodes[i] = ch;
    b.ref.parent[left] = i;
    b.ref.parent[right] = i;
}

__device__ __forceinline__ int decode_node(int idx) { return idx < 0 ? ~idx : idx; }

// RefitAABB (src/bvh.cu:215-242): the second thread to arrive at a node merges its children.
// Also measures the tree depth (ComputeAndPrintMaxDepth, src/bvh.cu:416-454, is a host walk).
__global__ void __launch_bounds__(256) k_refit(LbvhBuffers b) {
    int tid = blockIdx.x * blockDim.x + threadIdx.x;
    int n = b.num_triangles;
    if (tid >= n) return;
    int cur = tid + n;
    // the climbing thread keeps the box of the node it came from in registers and reads only the sibling's
    float4 cmin = b.ref.aabb_min[cur], cmax = b.ref.aabb_max[cur];
    while (true) {
        int parent = b.ref.parent[cur];
        if (parent == -1) break;
        __threadfence();
        if (atomicAdd(&b.flags[parent], 1) == 0) return;
        int2 ch = b.ref.child_nodes[parent];
        int lc = decode_node(ch.x), rc = decode_node(ch.y);
        int sib = (lc == cur) ? rc : lc;
        float4 smin = __ldcg(&b.ref.aabb_min[sib]), smax = __ldcg(&b.ref.aabb_max[sib]);
        cmin = make_float4(fminf(cmin.x, smin.x), fminf(cmin.y, smin.y), fminf(cmin.z, smin.z), fminf(cmin.w, smin.w));
        cmax = make_float4(fmaxf(cmax.x, smax.x), fmaxf(cmax.y, smax.y), fmaxf(cmax.z, smax.z), fmaxf(cmax.w, smax.w));
        b.ref.aabb_min[parent] = cmin;
        b.ref.aabb_max[parent] = cmax;
        cur = parent;
    }
}

// BuildEscapeIdx (src/bvh.cu:244-284) + leaf depth (max over leaves of #ancestors).
__global__ void __launch_bounds__(256) k_escape_depth(LbvhBuffers b) {
    int cur = blockIdx.x * blockDim.x + threadIdx.x;
    int n = b.num_triangles;
    int depth = 0;
    if (cur < n - 1) {
        int2 ch = b.ref.child_nodes[cur];
        int lc = decode_node(ch.x), rc = decode_node(ch.y);
        b.ref.escape_indices[lc] = rc;
        int escape = -1;
        if (cur != 0) {
            int node = cur;
            int p = b.ref.parent[cur];
            while (p != -1) {
                int2 pc = b.ref.child_nodes[p];
                if (decode_node(pc.x) == node) { escape = decode_node(pc.y); break; }
                node = p;
                p = b.ref.parent[node];
            }
        }
        b.ref.escape_indices[rc] = escape;
    }
    if (cur < n) {
        int node = cur + n;
        while ((node = b.ref.parent[node]) != -1) depth++;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) depth = max(depth, __shfl_xor_sync(0xffffffffu, depth, o));
    if ((threadIdx.x & 31) == 0 && depth > 0) atomicMax(b.max_depth, depth);
}

// Traversal layout (BvhDev, common.cuh): one 64-byte record per internal node holding both
// children's boxes and links; leaves carry their vertices inline.
__global__ void __launch_bounds__(256) k_pack(LbvhBuffers b, const float4* __restrict__ pos, const int4* __restrict__ indices_matid) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int n = b.num_triangles;
    if (i < n - 1) {
        int2 ch = b.ref.child_nodes[i];
        int lc = decode_node(ch.x), rc = decode_node(ch.y);
        int llink = ch.x < 0 ? ~(lc - n) : lc;
        int rlink = ch.y < 0 ? ~(rc - n) : rc;
        float4 lmin = b.ref.aabb_min[lc], lmax = b.ref.aabb_max[lc];
        float4 rmin = b.ref.aabb_min[rc], rmax = b.ref.aabb_max[rc];
        b.node[4 * (size_t)i + 0] = make_float4(lmin.x, lmin.y, lmax.x, lmax.y);     // layout of node_test(), raycast.cuh
        b.node[4 * (size_t)i + 1] = make_float4(rmin.x, rmin.y, rmax.x, rmax.y);
        b.node[4 * (size_t)i + 2] = make_float4(lmin.z, lmax.z, rmin.z, rmax.z);
        b.node[4 * (size_t)i + 3] = make_float4(__int_as_float(llink), __int_as_float(rlink), 0.0f, 0.0f);
    }
    if (i < n) {
        int prim = b.ref.primitive_indices[n + i];
        int4 im = __ldg(&indices_matid[prim]);
        float4 v0 = __ldg(&pos[im.x]), v1 = __ldg(&pos[im.y]), v2 = __ldg(&pos[im.z]);
        b.tri[PTB_TRI_STRIDE * (size_t)i + 0] = make_float4(v0.x, v0.y, v0.z, __int_as_float(prim));
        b.tri[PTB_TRI_STRIDE * (size_t)i + 1] = make_float4(v1.x, v1.y, v1.z, __int_as_float(im.w));
        b.tri[PTB_TRI_STRIDE * (size_t)i + 2] = make_float4(v2.x, v2.y, v2.z, 0.0f);
        b.tri[PTB_TRI_STRIDE * (size_t)i + 3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

// 4-wide traversal nodes (BvhDev::wnode, wide.cuh).  Node i is binary node i with its two children replaced, greedily
// by surface area, by THEIR children until four are held (Wald et al.'s collapse; the tree, and therefore every leaf's
// ancestor chain, stays the LBVH's -- only intermediate boxes are skipped).  Child boxes are stored on a 15-bit grid
// over the scene box, rounded outwards with one extra step of margin on either side, so that the box the traversal
// reconstructs (one fma per plane, wide.cuh) always contains the fp32 box of the reference:
//   wnode[4*i+a] (a = x, y, z) = four words, one per child: lo_a | hi_a << 16 (15-bit steps)
//   wnode[4*i+3] = link[4]: >= 0 internal node, < 0 ~(sorted leaf position), PTB_WIDE_EMPTY unused slot (box inverted)
__device__ __forceinline__ float box_area(float4 mn, float4 mx) {
    float dx = mx.x - mn.x, dy = mx.y - mn.y, dz = mx.z - mn.z;
    return dx * dy + dy * dz + dz * dx;
}
__global__ void __launch_bounds__(256) k_pack_wide(LbvhBuffers b) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = b.num_triangles;
    if (i >= n - 1) return;
    int child[4];
    int count = 2;
    {
        int2 ch = b.ref.child_nodes[i];
        child[0] = decode_node(ch.x);
        child[1] = decode_node(ch.y);
    }
    while (count < 4) {
        int best = -1;
        float best_area = -1.0f;
        for (int k = 0; k < count; k++) {
            if (child[k] >= n) continue;                      // a leaf cannot be opened
            float a = box_area(b.ref.aabb_min[child[k]], b.ref.aabb_max[child[k]]);
            if (a > best_area) { best_area = a; best = k; }
        }
        if (best < 0) break;
        int2 ch = b.ref.child_nodes[child[best]];
        child[best] = decode_node(ch.x);
        child[count++] = decode_node(ch.y);
    }
    const float4 wmin = b.ref.aabb_min[0], wmax = b.ref.aabb_max[0];
    const float cell[3] = {__fdiv_rn(wmax.x - wmin.x, PTB_WIDE_GRID), __fdiv_rn(wmax.y - wmin.y, PTB_WIDE_GRID), __fdiv_rn(wmax.z - wmin.z, PTB_WIDE_GRID)};
    const float org[3] = {wmin.x, wmin.y, wmin.z};
    unsigned int lo[3][4], hi[3][4];
    int link[4];
    for (int k = 0; k < 4; k++) {
        if (k >= count) {
            link[k] = PTB_WIDE_EMPTY;
            for (int a = 0; a < 3; a++) { lo[a][k] = 32767u; hi[a][k] = 0u; }
            continue;
        }
        const int c = child[k];
        link[k] = c >= n ? ~(c - n) : c;
        float4 mn = b.ref.aabb_min[c], mx = b.ref.aabb_max[c];
        const float vmin[3] = {mn.x, mn.y, mn.z}, vmax[3] = {mx.x, mx.y, mx.z};
        for (int a = 0; a < 3; a++) {
            float inv_cell = cell[a] > 0.0f ? 1.0f / cell[a] : 0.0f;           // a flat scene box: every plane is step 0
            float ql = floorf((vmin[a] - org[a]) * inv_cell) - 2.0f;
            float qh = ceilf((vmax[a] - org[a]) * inv_cell) + 2.0f;
            lo[a][k] = (unsigned int)fminf(fmaxf(ql, 0.0f), 32767.0f);
            hi[a][k] = (unsigned int)fminf(fmaxf(qh, 0.0f), 32767.0f);
        }
    }
    uint4* out = b.wnode + 4 * (size_t)i;
    for (int a = 0; a < 3; a++)
        out[a] = make_uint4(lo[a][0] | (hi[a][0] << 16), lo[a][1] | (hi[a][1] << 16), lo[a][2] | (hi[a][2] << 16), lo[a][3] | (hi[a][3] << 16));
    out[3] = make_uint4((unsigned)link[0], (unsigned)link[1], (unsigned)link[2], (unsigned)link[3]);
}

// 32-byte binary traversal nodes (BvhDev::bnode, wide.cuh trace_closest_bq): node i keeps the LBVH's two children, their
// boxes on the 15-bit grid of k_pack_wide (same quantiser, same outward rounding and margin).  A node visit is one
// 32-byte sector instead of the two of the fp32 record: RayCast is co-limited by L1 wavefronts.
__device__ __forceinline__ void quantise_box(float4 mn, float4 mx, const float org[3], const float cell[3], unsigned int q[3]) {
    const float vmin[3] = {mn.x, mn.y, mn.z}, vmax[3] = {mx.x, mx.y, mx.z};
    for (int a = 0; a < 3; a++) {
        float inv_cell = cell[a] > 0.0f ? 1.0f / cell[a] : 0.0f;
        float ql = floorf((vmin[a] - org[a]) * inv_cell) - 2.0f;
        float qh = ceilf((vmax[a] - org[a]) * inv_cell) + 2.0f;
        unsigned int lo = (unsigned int)fminf(fmaxf(ql, 0.0f), 32767.0f);
        unsigned int hi = (unsigned int)fminf(fmaxf(qh, 0.0f), 32767.0f);
        q[a] = lo | (hi << 16);
    }
}
__global__ void __launch_bounds__(256) k_pack_bq(LbvhBuffers b) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = b.num_triangles;
    if (i >= n - 1) return;
    int2 ch = b.ref.child_nodes[i];
    int lc = decode_node(ch.x), rc = decode_node(ch.y);
    int llink = ch.x < 0 ? ~(lc - n) : lc;
    int rlink = ch.y < 0 ? ~(rc - n) : rc;
    const float4 wmin = b.ref.aabb_min[0], wmax = b.ref.aabb_max[0];
    const float cell[3] = {__fdiv_rn(wmax.x - wmin.x, PTB_WIDE_GRID), __fdiv_rn(wmax.y - wmin.y, PTB_WIDE_GRID), __fdiv_rn(wmax.z - wmin.z, PTB_WIDE_GRID)};
    const float org[3] = {wmin.x, wmin.y, wmin.z};
    unsigned int ql[3], qr[3];
    quantise_box(b.ref.aabb_min[lc], b.ref.aabb_max[lc], org, cell, ql);
    quantise_box(b.ref.aabb_min[rc], b.ref.aabb_max[rc], org, cell, qr);
    b.bnode[2 * (size_t)i + 0] = make_uint4(ql[0], qr[0], ql[1], qr[1]);
    b.bnode[2 * (size_t)i + 1] = make_uint4(ql[2], qr[2], (unsigned)llink, (unsigned)rlink);
}

// ---- host side -------------------------------------------------------------------------------
template <typename T>
static cudaError_t dalloc(T*& p, size_t count, size_t& total) {
    size_t bytes = (count ? count : 1) * sizeof(T);
    cudaError_t e = cudaMalloc((void**)&p, bytes);
    if (e == cudaSuccess) total += bytes;
    return e;
}

int lbvh_alloc(LbvhBuffers& b, int n) {
    b = LbvhBuffers{};
    b.num_triangles = n;
    b.num_tiles = (n + kSortTile - 1) / kSortTile;
    size_t t = 0;
    cudaError_t e;
#define A(x) if ((e = (x)) != cudaSuccess) { lbvh_free(b); return (int)e; }
    A(dalloc(b.ref.aabb_min, 2 * (size_t)n, t));
    A(dalloc(b.ref.aabb_max, 2 * (size_t)n, t));
    A(dalloc(b.ref.centroid, 2 * (size_t)n, t));
    A(dalloc(b.ref.primitive_indices, 2 * (size_t)n, t));
    A(dalloc(b.ref.morton_codes, (size_t)n, t));
    A(dalloc(b.ref.child_nodes, (size_t)n, t));
    A(dalloc(b.ref.parent, 2 * (size_t)n, t));
    A(dalloc(b.ref.escape_indices, 2 * (size_t)n, t));
    A(dalloc(b.ref.world_aabb, 2, t));
    A(dalloc(b.node, 4 * (size_t)n, t));
    A(dalloc(b.tri, PTB_TRI_STRIDE * (size_t)n, t));
    A(dalloc(b.wnode, 4 * (size_t)n, t));
    A(dalloc(b.bnode, 2 * (size_t)n, t));
    A(dalloc(b.key_a, (size_t)n, t));
    A(dalloc(b.key_b, (size_t)n, t));
    A(dalloc(b.val_b, (size_t)n, t));
    A(dalloc(b.counts, 256 * (size_t)b.num_tiles, t));
    A(dalloc(b.totals, 256, t));
    A(dalloc(b.flags, (size_t)n, t));
    A(dalloc(b.world_enc, 8, t));
    A(dalloc(b.max_depth, 1, t));
#undef A
    b.bytes = t;
    return 0;
}

void lbvh_free(LbvhBuffers& b) {
    void* ptrs[] = {b.ref.aabb_min, b.ref.aabb_max, b.ref.centroid, b.ref.primitive_indices, b.ref.morton_codes,
                    b.ref.child_nodes, b.ref.parent, b.ref.escape_indices, b.ref.world_aabb, b.node, b.tri, b.wnode, b.bnode,
                    b.key_a, b.key_b, b.val_b, b.counts, b.totals, b.flags, b.world_enc, b.max_depth};
    for (void* p : ptrs) if (p) cudaFree(p);
    b = LbvhBuffers{};
}

int lbvh_build(LbvhBuffers& b, const float4* pos, const int4* indices_matid, cudaStream_t s, int* launches) {
    const int n = b.num_triangles;
    const int bs = 256;
    const int nb = (n + bs - 1) / bs;
    const int nb2 = (2 * n + bs - 1) / bs;
    int count = 0;
    k_init<<<nb2 > 1 ? nb2 : 1, bs, 0, s>>>(b); count++;
    k_prim_bounds<<<nb < 148 * 8 ? nb : 148 * 8, bs, 0, s>>>(b, pos, indices_matid); count++;
    k_morton<<<nb, bs, 0, s>>>(b); count++;
    unsigned int* kin = b.key_a; unsigned int* kout = b.key_b;
    int* vin = b.ref.primitive_indices; int* vout = b.val_b;
    for (int pass = 0; pass < 4; pass++) {
        int shift = 8 * pass;
        k_sort_count<<<b.num_tiles, kSortThreads, 0, s>>>(kin, n, shift, b.counts, b.num_tiles); count++;
        k_sort_rowscan<<<256, 256, 0, s>>>(b.counts, b.num_tiles, b.totals); count++;
        k_sort_scatter<<<b.num_tiles, kSortThreads, 0, s>>>(kin, vin, kout, vout, n, shift, b.counts, b.num_tiles, b.totals); count++;
        unsigned int* tk = kin; kin = kout; kout = tk;
        int* tv = vin; vin = vout; vout = tv;
    }
    // 4 passes: results are back in key_a / ref.primitive_indices
    k_leaves<<<nb, bs, 0, s>>>(b); count++;
    k_hierarchy<<<nb, bs, 0, s>>>(b); count++;
    k_refit<<<nb, bs, 0, s>>>(b); count++;
    k_escape_depth<<<nb, bs, 0, s>>>(b); count++;
    k_pack<<<nb, bs, 0, s>>>(b, pos, indices_matid); count++;
    k_pack_wide<<<nb, bs, 0, s>>>(b); count++;
    k_pack_bq<<<nb, bs, 0, s>>>(b); count++;
    if (launches) *launches += count;
    return (int)cudaGetLastError();
}

}  // namespace ptb
