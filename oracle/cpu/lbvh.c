/* oracle/cpu/lbvh.c -- TEST INFRASTRUCTURE ONLY (never linked into or called by the product).
 *
 * Plain-C restatement of the reference's LBVH build, BuildLBVH (src/bvh.cu:456-516), step by step:
 *   ComputeAABB :13-35, ComputeWorldAABB :37-51, ExpandBits/Morton3D :53-71, ComputeMortonCode :74-96,
 *   SortMortonCode :98-104, SetupLeafNodes :106-124, GenerateHierarchy :128-203, RefitAABB :215-242,
 *   BuildEscapeIdx :244-284.
 * Integer / min-max / non-contractible float arithmetic only, so the arrays
 * are bit-comparable with the device build.  Pinned against the reference itself: tests/golden/ holds
 * arrays dumped from the unmodified reference build (oracle/_ref/refdump) on a B200.
 * Build with -ffp-contract=off.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float x, y, z, w; } f4;
typedef struct { int32_t x, y, z, w; } i4;

static uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

static uint32_t morton3d(float x, float y, float z) {
    x = fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f);
    y = fminf(fmaxf(y * 1024.0f, 0.0f), 1023.0f);
    z = fminf(fmaxf(z * 1024.0f, 0.0f), 1023.0f);
    return (expand_bits((uint32_t)x) << 2) + (expand_bits((uint32_t)y) << 1) + expand_bits((uint32_t)z);
}

static int cmp_u64(const void* a, const void* b) {
    uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

static int clz64(uint64_t v) { return v ? __builtin_clzll(v) : 64; }

static int lcp(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return clz64(keys[i] ^ keys[j]);
}

static int decode(int idx) { return idx < 0 ? ~idx : idx; }
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* All outputs use the reference's LBVHData layout (src/scene_structs.h:151-169):
 * aabb_min/aabb_max/centroid f4[2N], primitive_indices int[2N], morton u64[N], child int2[N-1],
 * parent int[2N], escape int[2N], world f4[2].  Returns 0, or -1 if N < 2. */
int ptb_oracle_build_lbvh(const f4* pos, int num_vertices, const i4* im, int n, f4* aabb_min, f4* aabb_max, f4* centroid,
                          int32_t* prim, uint64_t* morton, int32_t* child, int32_t* parent, int32_t* escape, f4* world) {
    (void)num_vertices;
    if (n < 2) return -1;
    for (int i = 0; i < 2 * n; i++) { parent[i] = -1; escape[i] = -1; }
    f4 wmin = {FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX}, wmax = {-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int t = 0; t < n; t++) {
        f4 v0 = pos[im[t].x], v1 = pos[im[t].y], v2 = pos[im[t].z];
        f4 mn = {fminf(fminf(v0.x, v1.x), v2.x), fminf(fminf(v0.y, v1.y), v2.y), fminf(fminf(v0.z, v1.z), v2.z), 1.0f};
        f4 mx = {fmaxf(fmaxf(v0.x, v1.x), v2.x), fmaxf(fmaxf(v0.y, v1.y), v2.y), fmaxf(fmaxf(v0.z, v1.z), v2.z), 1.0f};
        aabb_min[t] = mn;
        aabb_max[t] = mx;
        f4 c = {(mx.x + mn.x) * 0.5f, (mx.y + mn.y) * 0.5f, (mx.z + mn.z) * 0.5f, (mx.w + mn.w) * 0.5f};
        centroid[t] = c;
        wmin.x = fminf(wmin.x, mn.x); wmin.y = fminf(wmin.y, mn.y); wmin.z = fminf(wmin.z, mn.z); wmin.w = fminf(wmin.w, mn.w);
        wmax.x = fmaxf(wmax.x, mx.x); wmax.y = fmaxf(wmax.y, mx.y); wmax.z = fmaxf(wmax.z, mx.z); wmax.w = fmaxf(wmax.w, mx.w);
    }
    world[0] = wmin;
    world[1] = wmax;
    float ex = wmax.x - wmin.x, ey = wmax.y - wmin.y, ez = wmax.z - wmin.z;
    if (ex < 1e-6f) ex = 1.0f;
    if (ey < 1e-6f) ey = 1.0f;
    if (ez < 1e-6f) ez = 1.0f;
    for (int t = 0; t < n; t++) {
        float sx = (centroid[t].x - wmin.x) / ex, sy = (centroid[t].y - wmin.y) / ey, sz = (centroid[t].z - wmin.z) / ez;
        morton[t] = ((uint64_t)morton3d(sx, sy, sz) << 32) | (uint64_t)(uint32_t)t;
    }
    /* sort_by_key(u64 -> int): keys are unique (tid in the low word), so the order is unique */
    qsort(morton, (size_t)n, sizeof(uint64_t), cmp_u64);
    for (int i = 0; i < n; i++) prim[i] = (int32_t)(morton[i] & 0xffffffffu);
    /* Karras 2012 */
    for (int i = 0; i < n - 1; i++) {
        int dir = (lcp(morton, n, i, i + 1) - lcp(morton, n, i, i - 1)) >= 0 ? 1 : -1;
        int l_max = 2;
        int delta_min = lcp(morton, n, i, i - dir);
        while (lcp(morton, n, i, i + l_max * dir) > delta_min) l_max *= 2;
        int l = 0;
        for (int t = l_max / 2; t >= 1; t /= 2)
            if (lcp(morton, n, i, i + (l + t) * dir) > delta_min) l += t;
        int j = i + l * dir;
        int delta_node = lcp(morton, n, i, j);
        int s = 0;
        for (int t = (l + 1) >> 1; t >= 1; t = (t == 1) ? 0 : ((t + 1) >> 1))
            if (lcp(morton, n, i, i + (s + t) * dir) > delta_node) s += t;
        int gamma = i + s * dir + imin(dir, 0);
        int left, right;
        if (imin(i, j) == gamma) { left = gamma + n; child[2 * i] = ~left; } else { left = gamma; child[2 * i] = left; }
        if (imax(i, j) == gamma + 1) { right = gamma + 1 + n; child[2 * i + 1] = ~right; } else { right = gamma + 1; child[2 * i + 1] = right; }
        parent[left] = i;
        parent[right] = i;
    }
    /* leaves to [N, 2N) -- before the refit overwrites [0, N-2] */
    for (int i = 0; i < n; i++) {
        int p = prim[i];
        aabb_min[n + i] = aabb_min[p];
        aabb_max[n + i] = aabb_max[p];
        centroid[n + i] = centroid[p];
        prim[n + i] = p;
    }
    /* refit bottom-up: a node is merged once both children are done (arrival counter, like the flags) */
    int32_t* flags = (int32_t*)calloc((size_t)n, sizeof(int32_t));
    for (int t = 0; t < n; t++) {
        int cur = t + n;
        for (;;) {
            int p = parent[cur];
            if (p == -1) break;
            if (flags[p]++ == 0) break;
            int lc = decode(child[2 * p]), rc = decode(child[2 * p + 1]);
            f4 a = aabb_min[lc], b = aabb_min[rc], c = aabb_max[lc], d = aabb_max[rc];
            f4 mn = {fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z), fminf(a.w, b.w)};
            f4 mx = {fmaxf(c.x, d.x), fmaxf(c.y, d.y), fmaxf(c.z, d.z), fmaxf(c.w, d.w)};
            aabb_min[p] = mn;
            aabb_max[p] = mx;
            cur = p;
        }
    }
    free(flags);
    /* escape indices */
    for (int cur = 0; cur < n - 1; cur++) {
        int lc = decode(child[2 * cur]), rc = decode(child[2 * cur + 1]);
        escape[lc] = rc;
        int esc = -1;
        if (cur != 0) {
            int node = cur, p = parent[cur];
            while (p != -1) {
                if (decode(child[2 * p]) == node) { esc = decode(child[2 * p + 1]); break; }
                node = p;
                p = parent[node];
            }
        }
        escape[rc] = esc;
    }
    return 0;
}

/* TestHierarchyLogic (src/bvh.cu:291-414): root has no parent, parent<->child links agree, a DFS from
 * the root reaches exactly N-1 internal nodes and N leaves and never the gap index N-1.
 * Returns 0 if all hold, else a bit mask of the failed checks. */
int ptb_oracle_check_hierarchy(const int32_t* child, const int32_t* parent, int n) {
    int fail = 0;
    if (parent[0] != -1) fail |= 1;
    for (int i = 0; i < n - 1; i++) {
        int lc = decode(child[2 * i]), rc = decode(child[2 * i + 1]);
        if (parent[lc] != i || parent[rc] != i) fail |= 2;
    }
    char* seen = (char*)calloc((size_t)(2 * n), 1);
    int32_t* stack = (int32_t*)malloc(sizeof(int32_t) * (size_t)(2 * n));
    int sp = 0, internal = 0, leaves = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        int node = stack[--sp];
        if (node < 0 || node >= 2 * n || node == n - 1 || seen[node]) { fail |= 4; continue; }
        seen[node] = 1;
        if (node >= n) { leaves++; continue; }
        internal++;
        stack[sp++] = decode(child[2 * node]);
        stack[sp++] = decode(child[2 * node + 1]);
    }
    if (internal != n - 1 || leaves != n) fail |= 8;
    free(seen);
    free(stack);
    return fail;
}
