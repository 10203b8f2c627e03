// views.cu -- what the reference shows through its GL preview, as headless device buffers:
//   * the first-bounce G-buffer the Logic stage writes (depth, geometric normal + material id, albedo, screen
//     motion; src/kernels/logic.cu:66-73,119-138) -- the inputs of the SVGF denoiser,
//   * the display modes of SendImageToPBOKernel (src/pathtrace_wavefront.cu:47-116): final / normals / depth /
//     albedo / motion as RGBA8,
//   * the LBVH traversal-cost heat map (VisualizeLBVHKernel, src/bvh.cu:518-620).
// None of this is on the benchmark path: the G-buffer kernel is enqueued only when ptb_config.gbuffer is set.
#include "logic.cuh"
#include "raycast.cuh"
#include "render.cuh"

namespace ptb {

// GetScreenUV, src/kernels/logic.cu:21-28: clip = viewProj * vec4(p, 1) in GLM's association
// (m[0]*x + m[1]*y) + (m[2]*z + m[3]*1) (glm/detail/type_mat4x4.inl:617-628), perspective divide, (0,0) = top left.
// Returns (u, v) in [0,1]; the caller scales by the resolution.  The two fused multiply-adds are the ones ptxas forms
// from the reference's plain mul/add pairs (read off its sm_100 SASS): u = fma(x, .5, .5), v = fma(y, -.5, .5).
__device__ __forceinline__ void screen_uv(const Mat4& vp, V3R p, float& u, float& v) {
    float cx = (vp.m[0][0] * p.x + vp.m[1][0] * p.y) + (vp.m[2][0] * p.z + vp.m[3][0]);
    float cy = (vp.m[0][1] * p.x + vp.m[1][1] * p.y) + (vp.m[2][1] * p.z + vp.m[3][1]);
    float cw = (vp.m[0][3] * p.x + vp.m[1][3] * p.y) + (vp.m[2][3] * p.z + vp.m[3][3]);
    float nx = cx / cw;
    float ny = cy / cw;
    u = __fmaf_rn(nx, 0.5f, 0.5f);
    v = __fmaf_rn(ny, -0.5f, 0.5f);
}

// Runs between TraceExtension(0) and Shade(0): every slot still holds its camera ray and first hit.
// The matrices live in device memory (f.view_proj: current, previous) because the frame is a replayed CUDA graph;
// k_frame_end copies current -> previous, as src/pathtrace_wavefront.cu:368 does on the host.
__global__ void __launch_bounds__(128) k_gbuffer(const __grid_constant__ FrameDev f, GBufferDev g) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= f.num_paths) return;
    const Mat4& cur = f.view_proj[0];
    const Mat4& prev = f.view_proj[1];
    const HitRecord hr = f.pool.load_hit(idx);
    int hit = hr.prim;
    if (hit == -1) {   // sky, src/kernels/logic.cu:66-73
        g.depth[idx] = -1000.0f;
        g.normal_matid[idx] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        g.albedo[idx] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
        g.motion[idx] = make_float2(0.0f, 0.0f);
        return;
    }
    float4 dd = f.pool.ray_dir_dist[idx];
    float t = hr.t;
    float4 huv = make_float4(hr.u, hr.v, 0.0f, 0.0f);
    int mat_id = hr.mat;
    Mat m = load_material(f.materials, mat_id);
    Vec3 N; float uvx, uvy;
    surface_properties(f, hit, huv.x, huv.y, m, N, uvx, uvy);
    Vec3 Ng = make_v3(__ldg(&f.mesh.nor_geom[hit]));
    g.depth[idx] = t;
    g.normal_matid[idx] = make_float4(Ng.x, Ng.y, Ng.z, (float)mat_id);
    if (m.diffuse_tex >= 0) {
        float4 tc = tex2D<float4>(__ldg(&f.tex.handles[m.diffuse_tex]), uvx, uvy);
        m.basecolor = m.basecolor * Vec3(powf(tc.x, 2.2f), powf(tc.y, 2.2f), powf(tc.z, 2.2f));
    }
    g.albedo[idx] = make_float4(m.basecolor.x, m.basecolor.y, m.basecolor.z, 1.0f);
    float4 oo = f.pool.ray_ori[idx];
    Vec3 ray_ori(oo.x, oo.y, oo.z);
    Vec3 ray_dir(dd.x, dd.y, dd.z);
    Vec3 hit_pos = ray_ori + ray_dir * t;
    float cu, cv, pu, pv;
    screen_uv(cur, hit_pos, cu, cv);
    screen_uv(prev, hit_pos, pu, pv);
    // motion = uv_curr * res - uv_prev * res: the previous product is rounded, the current one is fused (reference SASS)
    const float rx = (float)f.cam.res_x, ry = (float)f.cam.res_y;
    g.motion[idx] = make_float2(__fmaf_rn(cu, rx, -__fmul_rn(pu, rx)), __fmaf_rn(cv, ry, -__fmul_rn(pv, ry)));
}

// SendImageToPBOKernel modes 1-4 (normals, depth, albedo, motion) and the final clamp, :82-115.
__global__ void __launch_bounds__(256) k_view(GBufferDev g, int n, uchar4* dst, int mode) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Vec3 c(0.0f);
    if (mode == 1) {
        float4 nm = g.normal_matid[i];
        c = Vec3(nm.x, nm.y, nm.z) * 0.5f + 0.5f;
    } else if (mode == 2) {
        float d = g.depth[i];
        float view_d = clampf(1.0f - (d / 500.0f), 0.0f, 1.0f);
        c = Vec3(view_d);
    } else if (mode == 3) {
        float4 a = g.albedo[i];
        c = Vec3(a.x, a.y, a.z);
    } else if (mode == 4) {
        float2 mv = g.motion[i];
        float scale = 100.0f;
        c.x = clampf(mv.x * scale + 0.5f, 0.0f, 1.0f);
        c.y = clampf(mv.y * scale + 0.5f, 0.0f, 1.0f);
        c.z = 0.5f;
    }
    uchar4 o;
    o.x = (unsigned char)min(max((int)(c.x * 255.0f), 0), 255);
    o.y = (unsigned char)min(max((int)(c.y * 255.0f), 0), 255);
    o.z = (unsigned char)min(max((int)(c.z * 255.0f), 0), 255);
    o.w = 255;
    dst[i] = o;
}

// TemperatureColor, src/bvh.cu:518-534
__device__ __forceinline__ Vec3 temperature_color(float t) {
    t = clampf(t, 0.0f, 1.0f);
    Vec3 c(0.0f);
    if (t < 0.25f) c = mix3(Vec3(0.0f, 0.0f, 1.0f), Vec3(0.0f, 1.0f, 1.0f), t * 4.0f);
    else if (t < 0.5f) c = mix3(Vec3(0.0f, 1.0f, 1.0f), Vec3(0.0f, 1.0f, 0.0f), (t - 0.25f) * 4.0f);
    else if (t < 0.75f) c = mix3(Vec3(0.0f, 1.0f, 0.0f), Vec3(1.0f, 1.0f, 0.0f), (t - 0.5f) * 4.0f);
    else c = mix3(Vec3(1.0f, 1.0f, 0.0f), Vec3(1.0f, 0.0f, 0.0f), (t - 0.75f) * 4.0f);
    return c;
}

// VisualizeLBVHKernel, src/bvh.cu:536-620: un-jittered pinhole ray per pixel, count the nodes a near-first walk
// visits when nothing is culled by distance (boxes only, no triangles), colour = steps / 300.  64-entry stack with
// the reference's give-up rule.  Walks the packed traversal records: a leaf link is < 0.
__global__ void __launch_bounds__(128) k_bvh_heatmap(BvhDev bvh, CameraDev cam, TileMap tile, int n, float* out_rgb, int* out_steps) {
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    int pixel = tile.global_pixel(slot);
    int x = pixel % cam.res_x, y = pixel / cam.res_x;
    Vec3 view(cam.view[0], cam.view[1], cam.view[2]), right(cam.right[0], cam.right[1], cam.right[2]), up(cam.up[0], cam.up[1], cam.up[2]);
    Vec3 dir = normalize3(view
        + right * cam.pixel_length[0] * ((float)x - (float)cam.res_x * 0.5f)
        - up * cam.pixel_length[1] * ((float)y - (float)cam.res_y * 0.5f));
    Vec3 o(cam.pos[0], cam.pos[1], cam.pos[2]);
    Vec3 inv(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
    int steps = 0;
    int stack[64];
    int sp = 0;
    int node = (slab_test(bvh.root_min, bvh.root_max, o, inv) == -1.0f) ? PTB_NODE_NONE : 0;
    while (node != PTB_NODE_NONE || sp > 0) {
        if (node == PTB_NODE_NONE) node = stack[--sp];
        steps++;
        if (node < 0) {
            node = PTB_NODE_NONE;
        } else {
            const NodePair nt = node_test(bvh.node + 4 * (size_t)node, o, inv);
            const float t_l = nt.t_l, t_r = nt.t_r;
            const int left = nt.left, rightc = nt.right;
            bool hit_l = (t_l != -1.0f), hit_r = (t_r != -1.0f);
            if (hit_l && hit_r) {
                bool left_first = t_l < t_r;
                stack[sp++] = left_first ? rightc : left;
                node = left_first ? left : rightc;
            } else if (hit_l) node = left;
            else if (hit_r) node = rightc;
            else node = PTB_NODE_NONE;
        }
        if (sp >= 64) break;
    }
    const float MAX_STEPS = 300.0f;
    Vec3 c = temperature_color((float)steps / MAX_STEPS);
    if (out_rgb) { out_rgb[3 * (size_t)slot] = c.x; out_rgb[3 * (size_t)slot + 1] = c.y; out_rgb[3 * (size_t)slot + 2] = c.z; }
    if (out_steps) out_steps[slot] = steps;
}

int launch_gbuffer(const FrameDev& f, const GBufferDev& g, cudaStream_t s) {
    k_gbuffer<<<(f.num_paths + 127) / 128, 128, 0, s>>>(f, g);
    return 1;
}

int launch_view(const GBufferDev& g, int n, void* dst_rgba8, int mode, cudaStream_t s) {
    k_view<<<(n + 255) / 256, 256, 0, s>>>(g, n, (uchar4*)dst_rgba8, mode);
    return 1;
}

int launch_bvh_heatmap(const FrameDev& f, float* out_rgb, int* out_steps, cudaStream_t s) {
    k_bvh_heatmap<<<(f.num_paths + 127) / 128, 128, 0, s>>>(f.bvh, f.cam, f.tile, f.num_paths, out_rgb, out_steps);
    return 1;
}

}  // namespace ptb
