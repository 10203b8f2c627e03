// svgf.cu -- spatiotemporal variance-guided filter over the 1-spp frame (reference: src/kernels/svgf.cu, 11 launches
// per frame: demodulate, temporal, variance, gauss, 5 x a-trous, modulate, copy history).
//
// Same per-pixel arithmetic in the same order, 8 launches and less HBM traffic:
//   k_svgf_temporal      = KernelDemodulation + KernelTemporalFiltering      (the demodulated value never leaves registers)
//   k_svgf_variance      = KernelVarianceEstimating
//   k_svgf_gauss         = KernelVarianceGaussFilter
//   k_svgf_atrous<false> = KernelAtrousFilter, steps 1, 2, 4, 8
//   k_svgf_atrous<true>  = KernelAtrousFilter step 16 + KernelModulation + KernelCopyHistory: the last level's output
//                          is consumed only by the modulation of the same pixel, so it is never written.
// Ping-pong roles follow the reference (svgf.cu:611-712): with C = current and H = history index, temporal writes C,
// variance C->H, gauss H->C, a-trous C->H->C->H->C->(H), and the NEXT frame's history is this frame's C -- i.e. the
// output of the 4th a-trous level -- because the indices swap.  Here C = iteration & 1, read on the device, so one
// captured graph serves every frame.
#include "render.cuh"

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
__global__ void __launch_bounds__(256) k_svgf_temporal(SvgfDev s, GBufferDev g, const float* __restrict__ direct, const float* __restrict__ indirect,
                                                       const int* __restrict__ iteration, int w, int h, float alpha_color) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    int idx = x + y * w;
    const int C = (*iteration) & 1, H = C ^ 1;
    float4 albedo = g.albedo[idx];
    float depth = g.depth[idx];
    float rdx = direct[3 * (size_t)idx], rdy = direct[3 * (size_t)idx + 1], rdz = direct[3 * (size_t)idx + 2];
    float rix = indirect[3 * (size_t)idx], riy = indirect[3 * (size_t)idx + 1], riz = indirect[3 * (size_t)idx + 2];
    if (depth < 0.0f) {   // sky: passes through, variance slot 1 (:146-153)
        s.history[idx] = 0;
        s.direct[C][idx] = make_float4(rdx, rdy, rdz, 1.0f);
        s.indirect[C][idx] = make_float4(rix, riy, riz, 1.0f);
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
                    fma4(pd, s.direct[H][pidx], wt);
                    fma4(pi, s.indirect[H][pidx], wt);
                    fma4(pm, s.moments[H][pidx], wt);
                    wsum = __fadd_rn(wsum, wt);
                }
            }
        }
    }
    if (wsum > 1e-4f) {
        float inv_w = __frcp_rn(wsum);
        pd = mul4(pd, inv_w);
        pi = mul4(pi, inv_w);
        pm = mul4(pm, inv_w);
        int hist = s.history[idx] + 1;
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
    s.direct[C][idx] = dir;
    s.indirect[C][idx] = ind;
    s.moments[C][idx] = moments;
}

// ---- 3: spatial variance estimate where the history is short (svgf.cu:257-332) -----------------------------------
__global__ void __launch_bounds__(256) k_svgf_variance(SvgfDev s, GBufferDev g, const int* __restrict__ iteration, int w, int h) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    int idx = x + y * w;
    const int C = (*iteration) & 1, H = C ^ 1;
    float4 out_dir = s.direct[C][idx];
    float4 out_ind = s.indirect[C][idx];
    float center_depth = g.depth[idx];
    if (s.history[idx] >= 4 || center_depth < 0.0f) {
        s.direct[H][idx] = out_dir;
        s.indirect[H][idx] = out_ind;
        return;
    }
    float4 cn = g.normal_matid[idx];
    const float EPS = 1e-6f;
    int idx_r = min(x + 1, w - 1) + y * w;
    int idx_d = x + min(y + 1, h - 1) * w;
    float gx = __fsub_rn(g.depth[idx_r], center_depth), gy = __fsub_rn(g.depth[idx_d], center_depth);
    float sum_w = 1.0f;
    float4 sum_m = s.moments[C][idx];
    for (int j = -3; j <= 3; ++j) {
        for (int i = -3; i <= 3; ++i) {
            if (i == 0 && j == 0) continue;
            int tx = x + i, ty = y + j;
            if (tx < 0 || tx >= w || ty < 0 || ty >= h) continue;
            int t = tx + ty * w;
            float nd = g.depth[t];
            float4 nn = g.normal_matid[t];
            float dz = fabsf(__fsub_rn(center_depth, nd));
            float thr = __fadd_rn(fabsf(__fmaf_rn(gy, (float)j, __fmul_rn(gx, (float)i))), EPS);   // |gx i + gy j| + eps, sigma_z = 1
            float w_z = expf(__fdiv_rn(-dz, thr));
            float w_n = powf(fmaxf(0.0f, dot3f(cn, nn)), SVGF_SIGMA_N);
            float wt = __fmul_rn(w_z, w_n);
            sum_w = __fadd_rn(sum_w, wt);
            fma4(sum_m, s.moments[C][t], wt);
        }
    }
    sum_w = fmaxf(sum_w, EPS);
    float4 avg = div4(sum_m, sum_w);
    out_dir.w = variance_of(avg.x, avg.z);
    out_ind.w = variance_of(avg.y, avg.w);
    s.direct[H][idx] = out_dir;
    s.indirect[H][idx] = out_ind;
}

// ---- 4: 3x3 gaussian on the variance channel (svgf.cu:359-395) ---------------------------------------------------
__global__ void __launch_bounds__(256) k_svgf_gauss(SvgfDev s, const int* __restrict__ iteration, int w, int h) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    int idx = x + y * w;
    const int C = (*iteration) & 1, H = C ^ 1;
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
            sd = __fmaf_rn(s.direct[H][n].w, kw, sd);
            si = __fmaf_rn(s.indirect[H][n].w, kw, si);
        }
    }
    float4 od = s.direct[H][idx];
    od.w = sd;
    s.direct[C][idx] = od;
    float4 oi = s.indirect[H][idx];
    oi.w = si;
    s.indirect[C][idx] = oi;
}

// ---- 5 (+6, +7): edge-stopping a-trous level; the last level modulates and saves the history G-buffer -------------
// src/kernels/svgf.cu:334-357 (weights), 398-494 (filter), 497-523 (modulation), 525-536 (history copy).
template <bool LAST>
__global__ void __launch_bounds__(256) k_svgf_atrous(SvgfDev s, GBufferDev g, const int* __restrict__ iteration, int w, int h, int step, int src_is_cur) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    int idx = x + y * w;
    const int C = (*iteration) & 1;
    const int SRC = src_is_cur ? C : (C ^ 1), DST = SRC ^ 1;
    const float4* __restrict__ in_dir = s.direct[SRC];
    const float4* __restrict__ in_ind = s.indirect[SRC];
    const float EPS = 1e-6f;
    float center_depth = g.depth[idx];
    float4 c_dir = in_dir[idx];
    float4 c_ind = in_ind[idx];
    float4 cn = g.normal_matid[idx];
    if (LAST) {   // KernelCopyHistory
        s.prev_depth[idx] = center_depth;
        s.prev_normal_matid[idx] = cn;
    }
    if (center_depth < 0.0f) {
        if (LAST) {   // sky: the indirect buffer holds the environment colour (:509-514)
            s.final_image[3 * (size_t)idx] = c_ind.x; s.final_image[3 * (size_t)idx + 1] = c_ind.y; s.final_image[3 * (size_t)idx + 2] = c_ind.z;
        } else {
            s.direct[DST][idx] = c_dir;
            s.indirect[DST][idx] = c_ind;
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
            float nd = g.depth[t];
            float4 nn = g.normal_matid[t];
            float4 n_dir = in_dir[t];
            float4 n_ind = in_ind[t];
            float n_lum_dir = luminance(n_dir.x, n_dir.y, n_dir.z);
            float n_lum_ind = luminance(n_ind.x, n_ind.y, n_ind.z);
            // EdgeStoppingWeightsWithDenom
            float d_approx = __fmaf_rn(gx, (float)(i * step), __fmul_rn(gy, (float)(j * step)));
            float w_z = expf(__fdiv_rn(-fabsf(__fsub_rn(center_depth, nd)), __fadd_rn(fabsf(d_approx), EPS)));   // sigma_z = 1
            float w_n = powf(fmaxf(0.0f, dot3f(cn, nn)), SVGF_SIGMA_N);
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
        s.direct[DST][idx] = o_dir;
        s.indirect[DST][idx] = o_ind;
    }
}

// One frame of the filter: 8 launches.  `direct` / `indirect` are this frame's 1-spp accumulators.
int launch_svgf(const FrameDev& f, const SvgfDev& s, const GBufferDev& g, cudaStream_t st) {
    const int w = f.cam.res_x, h = f.cam.res_y;
    const dim3 bs(32, 8), grid((w + 31) / 32, (h + 7) / 8);
    k_svgf_temporal<<<grid, bs, 0, st>>>(s, g, f.direct, f.indirect, f.iteration, w, h, 0.1f);
    k_svgf_variance<<<grid, bs, 0, st>>>(s, g, f.iteration, w, h);
    k_svgf_gauss<<<grid, bs, 0, st>>>(s, f.iteration, w, h);
    int src_is_cur = 1;
    for (int i = 0; i < 4; i++) {
        k_svgf_atrous<false><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 1 << i, src_is_cur);
        src_is_cur ^= 1;
    }
    k_svgf_atrous<true><<<grid, bs, 0, st>>>(s, g, f.iteration, w, h, 16, src_is_cur);
    return 8;
}

}  // namespace ptb
