// svgf.cu -- spatiotemporal variance-guided filter over the 1-spp frame (reference: src/kernels/svgf.cu, 11 launches
// per frame: demodulate, temporal, variance, gauss, 5 x a-trous, modulate, copy history).
//
// Same per-pixel arithmetic in the same order, 7 launches and less HBM traffic:
//   k_svgf_temporal             = KernelDemodulation + KernelTemporalFiltering (the demodulated value never leaves
//                                 registers); it also lists the pixels whose history is shorter than 4 frames
//   k_svgf_variance             = KernelVarianceEstimating, run over that list only and in place: the estimate reads
//                                 the neighbours' moments / depth / normal, never their variance, so the reference's
//                                 full-frame copy into the other ping-pong half is not needed
//   k_svgf_atrous<false, true>  = KernelVarianceGaussFilter + KernelAtrousFilter step 1
//   k_svgf_atrous<false, false> = KernelAtrousFilter, steps 2, 4, 8
//   k_svgf_atrous<true, false>  = KernelAtrousFilter step 16 + KernelModulation + KernelCopyHistory
// Buffer roles.  The reference swaps "current" (C) and "history" (H) every frame and ends with the NEXT frame's
// history being the output of the 4th a-trous level (svgf.cu:611-712: temporal H->C, variance C->H, gauss H->C,
// a-trous C->H->C->H->C->(H), then the indices swap).  Two hops fewer here, so the parity comes out the same:
// temporal H->C, variance on C, a-trous C->H->C->H->C->final, and this frame's C is the next frame's H.
// C = iteration & 1 is read on the device, so one captured graph serves every frame.
#include <cooperative_groups.h>
#include "render.cuh"

namespace cg = cooperative_groups;

namespace cg = cooperative_groups;

namespace ptb {

#define SVGF_SIGMA_Z 1.0f     /* src/kernels/svgf.h:6-8 */
#define SVGF_SIGMA_N 128.0f
#define SVGF_SIGMA_L 4.0f

namespace {

// Every multiply-add below is written with explicit intrinsics in the form the reference's sm_100 SASS has (which
// product of a sum is rounded on its own and which is fused): the filter feeds its own output back through the
// temporal history and through exp(-|dL| / sqrt(variance)), so a 1-ulp difference grows frame over frame.  With the
// forms pinned the final image is bit-identical to the reference's.
__device__ __forceinline__ float luminance(float x, float y, float z) {       // 0.2126 x + 0.7152 y + 0.0722 z
    return __fmaf_rn(z, 0.0722f, __fmaf_rn(x, 0.2126f, __fmul_rn(y, 0.7152f)));
}
__device__ __forceinline__ float safe_demodulate(float c, float a) { return a > 0.01f ? __fdiv_rn(c, a) : c; }
__device__ __forceinline__ float lerp1(float a, float b, float t) { return __fmaf_rn(t, __fsub_rn(b, a), a); }   // a + t (b - a)
__device__ __forceinline__ float4 lerp4(float4 a, float4 b, float t) {
    return make_float4(lerp1(a.x, b.x, t), lerp1(a.y, b.y, t), lerp1(a.z, b.z, t), lerp1(a.w, b.w, t));
}
__device__ __forceinline__ float4 mul4(float4 a, float b) { return make_float4(__fmul_rn(a.x, b), __fmul_rn(a.y, b), __fmul_rn(a.z, b), __fmul_rn(a.w, b)); }
__device__ __forceinline__ void fma4(float4& acc, float4 v, float w) {          // acc += v * w
    acc.x = __fmaf_rn(v.x, w, acc.x); acc.y = __fmaf_rn(v.y, w, acc.y); acc.z = __fmaf_rn(v.z, w, acc.z); acc.w = __fmaf_rn(v.w, w, acc.w);
}
__device__ __forceinline__ float4 div4(float4 a, float b) { return make_float4(__fdiv_rn(a.x, b), __fdiv_rn(a.y, b), __fdiv_rn(a.z, b), __fdiv_rn(a.w, b)); }
__device__ __forceinline__ float dot3f(float4 a, float4 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y))); }
__device__ __forceinline__ float variance_of(float m1, float m2) { return fmaxf(0.0f, __fmaf_rn(-m1, m1, m2)); }   // max(0, m2 - m1^2)

// The ping-pong halves are picked with a select, not an index: indexing a by-value kernel parameter with a run-time
// value makes the compiler copy the struct to local memory and fall back to generic loads.
template <class T> __device__ __forceinline__ T* pick(T* const (&a)[2], int i) { return i ? a[1] : a[0]; }

// pow(max(0, n.n'), sigma_n).  powf already returns early for an argument of exactly 1 (a neighbour on the same
// triangle, most of the time), so there is nothing to gain from a same-normal shortcut of our own: measured.
__device__ __forceinline__ float normal_weight(float4 cn, float4 nn) { return powf(fmaxf(0.0f, dot3f(cn, nn)), SVGF_SIGMA_N); }

// IsTapConsistent, src/kernels/svgf.cu:95-118
__device__ __forceinline__ bool tap_consistent(int xp, int yp, int w, int h, float4 n_matid, float depth, const float4* __restrict__ prev_normal,
                                               const float* __restrict__ prev_depth) {
    if (xp < 0 || xp >= w || yp < 0 || yp >= h) return false;
    int pi = xp + yp * w;
    float4 pn = __ldg(&prev_normal[pi]);
    float pd = __ldg(&prev_depth[pi]);
    bool cn = dot3f(n_matid, pn) > 0.95f;
    bool cd = fabsf(__fsub_rn(depth, pd)) < 2.0f;
    bool cm = ((int)n_matid.w == (int)pn.w);
    return cn && cd && cm;
}

}  // namespace

// ---- 1+2: demodulation and temporal accumulation (svgf.cu:55-92,121-254) ---------------------------------------
__global__ void __launch_bounds__(256, 6) k_svgf_temporal(SvgfDev s, GBufferDev g, const float* __restrict__ direct, const float* __restrict__ indirect,
                                                       const int* __restrict__ iteration, int w, int h, float alpha_color) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    int idx = x + y * w;
    const int C = (*iteration) & 1, H = C ^ 1;
    float4* __restrict__ dir_c = pick(s.direct, C);
    float4* __restrict__ ind_c = pick(s.indirect, C);
    float4* __restrict__ mom_c = pick(s.moments, C);
    const float4* __restrict__ dir_h = pick(s.direct, H);
    const float4* __restrict__ ind_h = pick(s.indirect, H);
    const float4* __restrict__ mom_h = pick(s.moments, H);
    float4 albedo = g.albedo[idx];
    float depth = g.depth[idx];
    float rdx = direct[3 * (size_t)idx], rdy = direct[3 * (size_t)idx + 1], rdz = direct[3 * (size_t)idx + 2];
    float rix = indirect[3 * (size_t)idx], riy = indirect[3 * (size_t)idx + 1], riz = indirect[3 * (size_t)idx + 2];
    if (depth < 0.0f) {   // sky: passes through, variance slot 1 (:146-153)
        s.history[idx] = 0;
        dir_c[idx] = make_float4(rdx, rdy, rdz, 1.0f);
        ind_c[idx] = make_float4(rix, riy, riz, 1.0f);
        return;
    }
    float4 dir = make_float4(safe_demodulate(rdx, albedo.x), safe_demodulate(rdy, albedo.y), safe_demodulate(rdz, albedo.z), 0.0f);
    float4 ind = make_float4(safe_demodulate(rix, albedo.x), safe_demodulate(riy, albedo.y), safe_demodulate(riz, albedo.z), 0.0f);
    float4 nm = g.normal_matid[idx];

    float lum_dir = luminance(dir.x, dir.y, dir.z);
    float lum_ind = luminance(ind.x, ind.y, ind.z);
    float4 moments;
    moments.x = lum_dir;
    moments.z = __fmul_rn(lum_dir, lum_dir);
    moments.y = lum_ind;
    moments.w = __fmul_rn(lum_ind, lum_ind);

    float2 mv = g.motion[idx];
    float s_prev = __fsub_rn((float)x, mv.x);
    float t_prev = __fsub_rn((float)y, mv.y);
    int x_p = (int)floorf(s_prev);
    int y_p = (int)floorf(t_prev);
    float fx = __fsub_rn(s_prev, (float)x_p);
    float fy = __fsub_rn(t_prev, (float)y_p);
    float gx1 = __fsub_rn(1.0f, fx), gy1 = __fsub_rn(1.0f, fy);
    float bw[4] = {__fmul_rn(gx1, gy1), __fmul_rn(fx, gy1), __fmul_rn(gx1, fy), __fmul_rn(fx, fy)};

    float4 pd = make_float4(0.0f, 0.0f, 0.0f, 0.0f), pi = pd, pm = pd;
    float wsum = 0.0f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int tx = x_p + i, ty = y_p + j;
            float wt = bw[i + j * 2];
            if (wt > 1e-6f) {
                if (tap_consistent(tx, ty, w, h, nm, depth, s.prev_normal_matid, s.prev_depth)) {
                    int pidx = tx + ty * w;
                    fma4(pd, dir_h[pidx], wt);
                    fma4(pi, ind_h[pidx], wt);
                    fma4(pm, mom_h[pidx], wt);
                    wsum = __fadd_rn(wsum, wt);
                }
            }
        }
    }
    int hist = 0;
    if (wsum > 1e-4f) {
        float inv_w = __frcp_rn(wsum);
        pd = mul4(pd, inv_w);
        pi = mul4(pi, inv_w);
        pm = mul4(pm, inv_w);
        hist = s.history[idx] + 1;
        s.history[idx] = hist;
        float alpha = fmaxf(alpha_color, __frcp_rn((float)hist));
        dir = lerp4(pd, dir, alpha);
        ind = lerp4(pi, ind, alpha);
        moments = lerp4(pm, moments, alpha);
        if (hist >= 4) {
            dir.w = variance_of(moments.x, moments.z);
            ind.w = variance_of(moments.y, moments.w);
        } else {
            dir.w = 1.0f;
            ind.w = 1.0f;
        }
    } else {
        s.history[idx] = 0;
        dir.w = 1.0f;
        ind.w = 1.0f;
    }
    dir_c[idx] = dir;
    ind_c[idx] = ind;
    mom_c[idx] = moments;
    if (hist < 4) {   // KernelVarianceEstimating's condition (svgf.cu:270): this pixel gets the spatial estimate
        cg::coalesced_group grp = cg::coalesced_threads();
        int base = 0;
        if (grp.thread_rank() == 0) base = atomicAdd(s.short_count, (int)grp.size());
        base = grp.shfl(base, 0);
        s.short_list[base + grp.thread_rank()] = idx;
    }
}

// ---- 3: spatial variance estimate where the history is short (svgf.cu:257-332) -----------------------------------
// After the first frames only disoccluded pixels and the jittered silhouettes need the 7x7 estimate: a few per cent
// of the pixels, scattered over a third of the warps.  One thread per listed pixel keeps the lanes full; per-pixel
// arithmetic and its order are untouched (the list order only decides which lane does which pixel).
__global__ void __launch_bounds__(256) k_svgf_variance(SvgfDev s, GBufferDev g, const int* __restrict__ iteration, int w, int h) {
    const int C = (*iteration) & 1;
    const float4* __restrict__ mom_c = pick(s.moments, C);
    const int count = *s.short_count;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const int idx = s.short_list[k];
        const int y = idx / w, x = idx - y * w;
        float center_depth = g.depth[idx];
        float4 cn = g.normal_matid[idx];
        const float EPS = 1e-6f;
        int idx_r = min(x + 1, w - 1) + y * w;
        int idx_d = x + min(y + 1, h - 1) * w;
        float gx = __fsub_rn(g.depth[idx_r], center_depth), gy = __fsub_rn(g.depth[idx_d], center_depth);
        float sum_w = 1.0f;
        float4 sum_m = mom_c[idx];
        for (int j = -3; j <= 3; ++j) {
            for (int i = -3; i <= 3; ++i) {
                if (i == 0 && j == 0) continue;
                int tx = x + i, ty = y + j;
                if (tx < 0 || tx >= w || ty < 0 || ty >= h) continue;
                int t = tx + ty * w;
                float4 nn = g.normal_matid[t];
                float nd = g.depth[t];
                float4 nm = mom_c[t];
                float w_n = normal_weight(cn, nn);
                float dz = fabsf(__fsub_rn(center_depth, nd));
                float thr = __fadd_rn(fabsf(__fmaf_rn(gy, (float)j, __fmul_rn(gx, (float)i))), EPS);   // |gx i + gy j| + eps, sigma_z = 1
                float w_z = expf(__fdiv_rn(-dz, thr));
                float wt = __fmul_rn(w_z, w_n);
                sum_w = __fadd_rn(sum_w, wt);
                fma4(sum_m, nm, wt);
            }
        }
        sum_w = fmaxf(sum_w, EPS);
        float4 avg = div4(sum_m, sum_w);
        pick(s.direct, C)[idx].w = variance_of(avg.x, avg.z);
        pick(s.indirect, C)[idx].w = variance_of(avg.y, avg.w);
    }
}

// ---- 4: 3x3 gaussian on the variance channel (KernelVarianceGaussFilter, svgf.cu:359-395), for one pixel ---------
__device__ __forceinline__ float2 gauss_variance(const float4* __restrict__ dir, const float4* __restrict__ ind, int x, int y, int w, int h) {
    const float k3[3][3] = {{0.0625f, 0.125f, 0.0625f}, {0.125f, 0.25f, 0.125f}, {0.0625f, 0.125f, 0.0625f}};
    float sd = 0.0f, si = 0.0f;
#pragma unroll
    for (int j = -1; j <= 1; ++j) {
        int ty = min(max(y + j, 0), h - 1);
#pragma unroll
        for (int i = -1; i <= 1; ++i) {
            int tx = min(max(x + i, 0), w - 1);
            int n = tx + ty * w;
            float kw = k3[j + 1][i + 1];
            sd = __fmaf_rn(dir[n].w, kw, sd);
            si = __fmaf_rn(ind[n].w, kw, si);
        }
    }
    return make_float2(sd, si);
}

// ---- 5 (+4, +6, +7): edge-stopping a-trous level ------------------------------------------------------------------
// src/kernels/svgf.cu:334-357 (weights), 398-494 (filter), 497-523 (modulation), 525-536 (history copy).
// GAUSS (level 1): the gaussian on the variance is folded in.  A step-1 level reads the variance of its own pixel and
// of the 8 adjacent ones, so the block filters the variance of its 32x8 tile plus a one-pixel rim into shared memory
// (each value with the reference's own 9-term sum) and the taps take it from there; the colour channels pass through
// the reference's gauss kernel unchanged, so they are read straight from the variance pass's buffer.
// LAST (level 5): the output is consumed only by the modulation of the same pixel, so it is multiplied by the albedo
// and written as the final image; the history G-buffer copy rides along.
#define SVGF_BX 32
#define SVGF_BY 8
template <bool LAST, bool GAUSS>
__global__ void __launch_bounds__(SVGF_BX * SVGF_BY, 5) k_svgf_atrous(SvgfDev s, GBufferDev g, const int* __restrict__ iteration, int w, int h, int step, int src_is_cur) {
    __shared__ float2 s_var[GAUSS ? (SVGF_BX + 2) * (SVGF_BY + 2) : 1];
    const int C = (*iteration) & 1;
    const int SRC = src_is_cur ? C : (C ^ 1), DST = SRC ^ 1;
    const float4* __restrict__ in_dir = pick(s.direct, SRC);
    const float4* __restrict__ in_ind = pick(s.indirect, SRC);
    float4* __restrict__ out_dir = pick(s.direct, DST);
    float4* __restrict__ out_ind = pick(s.indirect, DST);
    int x = blockIdx.x * SVGF_BX + threadIdx.x;
    int y = blockIdx.y * SVGF_BY + threadIdx.y;
    if (GAUSS) {
        if (x == 0 && y == 0) *s.short_count = 0;   // the variance pass has consumed the list
        const int x0 = blockIdx.x * SVGF_BX - 1, y0 = blockIdx.y * SVGF_BY - 1;
        for (int p = threadIdx.y * SVGF_BX + threadIdx.x; p < (SVGF_BX + 2) * (SVGF_BY + 2); p += SVGF_BX * SVGF_BY) {
            int ly = p / (SVGF_BX + 2), lx = p - ly * (SVGF_BX + 2);
            int qx = x0 + lx, qy = y0 + ly;
            if (qx >= 0 && qx < w && qy >= 0 && qy < h) s_var[p] = gauss_variance(in_dir, in_ind, qx, qy, w, h);
        }
        __syncthreads();
    }
    if (x >= w || y >= h) return;
    int idx = x + y * w;
    const float EPS = 1e-6f;
    float center_depth = g.depth[idx];
    float4 c_dir = in_dir[idx];
    float4 c_ind = in_ind[idx];
    if (GAUSS) {
        float2 v = s_var[(threadIdx.y + 1) * (SVGF_BX + 2) + threadIdx.x + 1];
        c_dir.w = v.x;
        c_ind.w = v.y;
    }
    float4 cn = g.normal_matid[idx];
    if (LAST) {   // KernelCopyHistory
        s.prev_depth[idx] = center_depth;
        s.prev_normal_matid[idx] = cn;
    }
    if (center_depth < 0.0f) {
        if (LAST) {   // sky: the indirect buffer holds the environment colour (:509-514)
            s.final_image[3 * (size_t)idx] = c_ind.x; s.final_image[3 * (size_t)idx + 1] = c_ind.y; s.final_image[3 * (size_t)idx + 2] = c_ind.z;
        } else {
            out_dir[idx] = c_dir;
            out_ind[idx] = c_ind;
        }
        return;
    }
    float c_lum_dir = luminance(c_dir.x, c_dir.y, c_dir.z);
    float lum_denom_dir = __frcp_rn(__fmaf_rn(sqrtf(fmaxf(0.0f, c_dir.w)), SVGF_SIGMA_L, EPS));
    float c_lum_ind = luminance(c_ind.x, c_ind.y, c_ind.z);
    float lum_denom_ind = __frcp_rn(__fmaf_rn(sqrtf(fmaxf(0.0f, c_ind.w)), SVGF_SIGMA_L, EPS));
    int idx_r = min(x + 1, w - 1) + y * w;
    int idx_d = x + min(y + 1, h - 1) * w;
    float gx = __fsub_rn(g.depth[idx_r], center_depth), gy = __fsub_rn(g.depth[idx_d], center_depth);
    if (fabsf(gx) < EPS) gx = EPS;
    if (fabsf(gy) < EPS) gy = EPS;
    float sw_dir = 1.0f, sw_ind = 1.0f;
    float4 sum_dir = c_dir, sum_ind = c_ind;
#pragma unroll
    for (int j = -1; j <= 1; ++j) {
#pragma unroll
        for (int i = -1; i <= 1; ++i) {
            if (i == 0 && j == 0) continue;
            int tx = x + i * step, ty = y + j * step;
            if (tx < 0 || tx >= w || ty < 0 || ty >= h) continue;
            int t = tx + ty * w;
            float4 nn = g.normal_matid[t];
            float nd = g.depth[t];
            float4 n_dir = in_dir[t];
            float4 n_ind = in_ind[t];
            if (GAUSS) {   // step == 1
                float2 v = s_var[(threadIdx.y + 1 + j) * (SVGF_BX + 2) + threadIdx.x + 1 + i];
                n_dir.w = v.x;
                n_ind.w = v.y;
            }
            float w_n = normal_weight(cn, nn);
            float n_lum_dir = luminance(n_dir.x, n_dir.y, n_dir.z);
            float n_lum_ind = luminance(n_ind.x, n_ind.y, n_ind.z);
            // EdgeStoppingWeightsWithDenom
            float d_approx = __fmaf_rn(gx, (float)(i * step), __fmul_rn(gy, (float)(j * step)));
            float w_z = expf(__fdiv_rn(-fabsf(__fsub_rn(center_depth, nd)), __fadd_rn(fabsf(d_approx), EPS)));   // sigma_z = 1
            float w_l_dir = __fmul_rn(w_n, expf(__fmul_rn(-fabsf(__fsub_rn(c_lum_dir, n_lum_dir)), lum_denom_dir)));
            float w_l_ind = __fmul_rn(w_n, expf(__fmul_rn(-fabsf(__fsub_rn(c_lum_ind, n_lum_ind)), lum_denom_ind)));
            float w_zn = __fmul_rn(w_z, w_n);
            float w_dir = __fmul_rn(w_zn, w_l_dir);
            float w_ind = __fmul_rn(w_zn, w_l_ind);
            sw_dir = __fadd_rn(sw_dir, w_dir);
            fma4(sum_dir, n_dir, w_dir);
            sw_ind = __fadd_rn(sw_ind, w_ind);
            fma4(sum_ind, n_ind, w_ind);
        }
    }
    float4 o_dir = div4(sum_dir, sw_dir);
    float4 o_ind = div4(sum_ind, sw_ind);
    if (LAST) {   // KernelModulation: (direct + indirect) * albedo, no gamma
        float4 a = g.albedo[idx];
        s.final_image[3 * (size_t)idx] = __fmul_rn(a.x, __fadd_rn(o_dir.x, o_ind.x));
        s.final_image[3 * (size_t)idx + 1] = __fmul_rn(a.y, __fadd_rn(o_dir.y, o_ind.y));
        s.final_image[3 * (size_t)idx + 2] = __fmul_rn(a.z, __fadd_rn(o_dir.z, o_ind.z));
    } else {
        out_dir[idx] = o_dir;
        out_ind[idx] = o_ind;
    }
}

// One frame of the filter: 7 launches.  `direct` / `indirect` are this frame's 1-spp accumulators.
int launch_svgf(const FrameDev& f, const SvgfDev& s, const GBufferDev& g, int num_sms, cudaStream_t st) {
    const int w = f.cam.res_x, h = f.cam.res_y;
    const dim3 bs(SVGF_BX, SVGF_BY), grid((w + SVGF_BX - 1) / SVGF_BX, (h + SVGF_BY - 1) / SVGF_BY);
    k_svgf_temporal<<<grid, bs, 0, st>>>(s, g, f.direct, f.indirect, f.iteration, w, h, 0.1f);
    k_svgf_variance<<<8 * num_sms, 256, 0, st>>>(s, g, f.iteration, w, h);
    k_svgf_atrous<false, true><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 1, 1);    // C -> H, gauss folded in
    k_svgf_atrous<false, false><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 2, 0);   // H -> C
    k_svgf_atrous<false, false><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 4, 1);   // C -> H
    k_svgf_atrous<false, false><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 8, 0);   // H -> C: the next frame's history
    k_svgf_atrous<true, false><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 16, 1);   // C -> final image
    return 7;
}

}  // namespace ptb
