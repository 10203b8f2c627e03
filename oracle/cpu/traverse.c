/* oracle/cpu/traverse.c -- TEST INFRASTRUCTURE ONLY (never linked into or called by the product).
 *
 * Plain-C restatement of the reference's RayCast stage over the reference-layout LBVH arrays:
 *   TraceExtensionRayKernel  src/kernels/ray_cast.cu:12-122   (closest hit, 32-entry stack, near first)
 *   TraceShadowRayKernel     src/kernels/ray_cast.cu:125-201  (any hit, stackless escape-index walk)
 *   TriangleIntersectionTest src/intersections.cu:10-32
 *   BoudingboxIntersetionTest src/intersections.cu:34-61
 * The reference has no CPU path; this is also the timed "naive CPU traversal" baseline of bench.py
 * (north_star: reported baseline only).  The triangle test uses fmaf in the contraction the reference's
 * sm_100 SASS shows, so t/u/v are bit-comparable with the device.
 * Build with -ffp-contract=off.
 */
#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct { float x, y, z, w; } f4;
typedef struct { int32_t x, y, z, w; } i4;
typedef struct { float x, y, z; } v3;

#define EPSILON 0.001f

static float dot3(v3 a, v3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }
static v3 cross3(v3 a, v3 b) {
    v3 r = {fmaf(a.y, b.z, -(b.y * a.z)), fmaf(a.z, b.x, -(b.z * a.x)), fmaf(a.x, b.y, -(b.x * a.y))};
    return r;
}
static v3 sub3(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static v3 from4(f4 a) { v3 r = {a.x, a.y, a.z}; return r; }

static float tri_test(v3 v0, v3 v1, v3 v2, v3 o, v3 d, float* out_u, float* out_v) {
    v3 e1 = sub3(v1, v0), e2 = sub3(v2, v0);
    v3 pvec = cross3(d, e2);
    float det = dot3(e1, pvec);
    if (fabsf(det) < EPSILON) return -1.0f;
    float inv_det = 1.0f / det;
    v3 tvec = sub3(o, v0);
    *out_u = dot3(tvec, pvec) * inv_det;
    if (*out_u < 0.0f || *out_u > 1.0f) return -1.0f;
    v3 qvec = cross3(tvec, e1);
    *out_v = dot3(d, qvec) * inv_det;
    if (*out_v < 0.0f || (*out_u + *out_v) > 1.0f) return -1.0f;
    float t = dot3(e2, qvec) * inv_det;
    return (t > 1e-7f) ? t : -1.0f;
}

static float box_test(f4 mn, f4 mx, v3 o, v3 inv) {
    float t1x = (mn.x - o.x) * inv.x, t2x = (mx.x - o.x) * inv.x;
    float t_near = fminf(t1x, t2x), t_far = fmaxf(t1x, t2x);
    float t1y = (mn.y - o.y) * inv.y, t2y = (mx.y - o.y) * inv.y;
    t_near = fmaxf(t_near, fminf(t1y, t2y));
    t_far = fminf(t_far, fmaxf(t1y, t2y));
    float t1z = (mn.z - o.z) * inv.z, t2z = (mx.z - o.z) * inv.z;
    t_near = fmaxf(t_near, fminf(t1z, t2z));
    t_far = fminf(t_far, fmaxf(t1z, t2z));
    if (t_near <= t_far && t_far > 0.0f) return fmaxf(0.0f, t_near);
    return -1.0f;
}

typedef struct {
    const f4 *aabb_min, *aabb_max, *pos;
    const int32_t *prim, *child, *escape;
    const i4* im;
    int n;
} Bvh;

static int decode(int idx) { return idx < 0 ? ~idx : idx; }

static void closest_one(const Bvh* b, v3 o, v3 d, int32_t* hit, float* t_out, float* u_out, float* v_out, int* max_sp) {
    v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    int hit_id = -1;
    float t_min = FLT_MAX, hu = 0.0f, hv = 0.0f;
    int stack[32], sp = 0, node = 0;
    if (box_test(b->aabb_min[0], b->aabb_max[0], o, inv) == -1.0f) node = -1;
    while (node != -1 || sp > 0) {
        if (node == -1) node = stack[--sp];
        if (node >= b->n) {
            int tri = b->prim[node];
            i4 m = b->im[tri];
            float u, v;
            float t = tri_test(from4(b->pos[m.x]), from4(b->pos[m.y]), from4(b->pos[m.z]), o, d, &u, &v);
            if (t > EPSILON && t < t_min) { t_min = t; hit_id = tri; hu = u; hv = v; }
            node = -1;
        } else {
            int left = decode(b->child[2 * node]), right = decode(b->child[2 * node + 1]);
            float t_l = box_test(b->aabb_min[left], b->aabb_max[left], o, inv);
            float t_r = box_test(b->aabb_min[right], b->aabb_max[right], o, inv);
            int hit_l = (t_l != -1.0f && t_l < t_min), hit_r = (t_r != -1.0f && t_r < t_min);
            if (hit_l && hit_r) {
                int first = (t_l < t_r) ? left : right, second = (t_l < t_r) ? right : left;
                stack[sp++] = second;
                node = first;
            } else if (hit_l) node = left;
            else if (hit_r) node = right;
            else node = -1;
        }
        if (sp > *max_sp) *max_sp = sp;
        if (sp >= 32) break;
    }
    *hit = hit_id;
    *t_out = hit_id == -1 ? -1.0f : t_min;
    *u_out = hu;
    *v_out = hv;
}

static int occluded_one(const Bvh* b, v3 o, v3 d, float tmax) {
    v3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    int node = 0;
    while (node != -1) {
        float t_box = box_test(b->aabb_min[node], b->aabb_max[node], o, inv);
        if (t_box != -1.0f && tmax > t_box) {
            if (node >= b->n) {
                i4 m = b->im[b->prim[node]];
                float u, v;
                float t = tri_test(from4(b->pos[m.x]), from4(b->pos[m.y]), from4(b->pos[m.z]), o, d, &u, &v);
                if (t > EPSILON && t < tmax - EPSILON) return 1;
                node = b->escape[node];
            } else {
                node = decode(b->child[2 * node]);
            }
        } else {
            node = b->escape[node];
        }
    }
    return 0;
}

typedef struct {
    const Bvh* b;
    const f4 *ori, *dir;
    int begin, end, mode, max_sp;
    int32_t* hit;
    float *t, *u, *v;
} Job;

static void* worker(void* p) {
    Job* j = (Job*)p;
    for (int i = j->begin; i < j->end; i++) {
        if (j->mode == 0) closest_one(j->b, from4(j->ori[i]), from4(j->dir[i]), &j->hit[i], &j->t[i], &j->u[i], &j->v[i], &j->max_sp);
        else j->hit[i] = occluded_one(j->b, from4(j->ori[i]), from4(j->dir[i]), j->ori[i].w);
    }
    return 0;
}

static int run(const Bvh* b, const f4* ori, const f4* dir, int count, int mode, int32_t* hit, float* t, float* u, float* v, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    Job jobs[256];
    pthread_t tid[256];
    int chunk = (count + threads - 1) / threads, max_sp = 0;
    for (int k = 0; k < threads; k++) {
        Job j = {b, ori, dir, k * chunk, (k + 1) * chunk < count ? (k + 1) * chunk : count, mode, 0, hit, t, u, v};
        if (j.begin > count) j.begin = count;
        jobs[k] = j;
        if (threads == 1) worker(&jobs[k]);
        else pthread_create(&tid[k], 0, worker, &jobs[k]);
    }
    for (int k = 0; k < threads; k++) {
        if (threads > 1) pthread_join(tid[k], 0);
        if (jobs[k].max_sp > max_sp) max_sp = jobs[k].max_sp;
    }
    return max_sp;
}

/* Closest hit for `count` rays (ori/dir float4, xyz used).  Returns the deepest stack use seen
 * (32 means the reference's abort at src/kernels/ray_cast.cu:106 triggered for some ray). */
int ptb_oracle_trace_closest(const f4* aabb_min, const f4* aabb_max, const int32_t* prim, const int32_t* child, const f4* pos,
                             const i4* im, int n, const f4* ori, const f4* dir, int count, int32_t* hit, float* t, float* u,
                             float* v, int threads) {
    Bvh b = {aabb_min, aabb_max, pos, prim, child, 0, im, n};
    return run(&b, ori, dir, count, 0, hit, t, u, v, threads);
}

/* Visibility test for `count` shadow rays (ori.w = tmax): occluded[i] = 0/1. */
int ptb_oracle_trace_occluded(const f4* aabb_min, const f4* aabb_max, const int32_t* prim, const int32_t* child,
                              const int32_t* escape, const f4* pos, const i4* im, int n, const f4* ori_tmax, const f4* dir,
                              int count, int32_t* occluded, int threads) {
    Bvh b = {aabb_min, aabb_max, pos, prim, child, escape, im, n};
    run(&b, ori_tmax, dir, count, 1, occluded, 0, 0, 0, threads);
    return 0;
}
