// dedup.cu -- vertex de-duplication and the AoS -> SoA pack of a parsed triangle soup, on the GPU.
//
// What the reference does on one host thread: Scene::loadFromJSON walks the triangle corners in file order through an
// unordered_map<Vertex,int> keyed on (pos, nor, uv) and numbers vertices by first occurrence (src/scene.cpp:330-539, the
// numbering at :525-532); WavefrontPathTracerState::initSceneGeometry then repacks the Vertex structs into float4 / float2 / int4
// arrays, normalising normals and tangents on the way (src/wavefront_internal.cu:118-167).
//
// The numbering only depends on, per corner, the EARLIEST corner with an equal key.  Here every corner inserts itself into one
// open-addressing table in device memory whose slots hold a corner index: an empty slot is claimed with atomicCAS, a slot whose
// occupant has an equal key is lowered with atomicMin -- a slot's key class never changes once claimed, so after the insert kernel
// each slot holds the minimum corner index of its class, whatever order the threads ran in.  An exclusive scan of the
// "I am my own first occurrence" flags in corner order gives the sequential ids; one more pass writes the vertex arrays in upload
// layout and the index quadruples.  Equality is float == on the eight floats, like Vertex::operator== (src/scene.h:26-35):
// -0 equals +0 (the hash folds them), a corner with a NaN anywhere equals nothing, itself included, and becomes a vertex of its own.
//
// The arrays leave this file bit-identical to the host path's (csrc/host/scene.cpp dedup_and_emit + the normalisation loop); the
// normalisation is written with round-to-nearest intrinsics (no contraction) and reproduces the NaNs an x86 host makes
// (0 * inf is the negative default NaN there, the positive canonical one on the GPU): tests/test_gpu_loader.py compares bits.
#include <cuda_runtime.h>

#include <chrono>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../host/dedup_gpu.hpp"
#include "../host/scene.hpp"

namespace ptb {
namespace {

constexpr uint32_t kEmpty = 0xffffffffu;
constexpr int kThreads = 256;
constexpr int kPerThread = 8;                      // consecutive corners per thread in the scan passes: one 32-byte sector of `first`
constexpr int kTile = kThreads * kPerThread;

// ---- keys ------------------------------------------------------------------------------------------------------------------------
struct Key {
    float4 a, b;   // (pos.x, pos.y, pos.z, nor.x) (nor.y, nor.z, uv.x, uv.y)
};
__device__ __forceinline__ Key load_key(const float4* __restrict__ keys, uint32_t c) {
    Key k;
    k.a = __ldg(keys + 2 * (size_t)c);
    k.b = __ldg(keys + 2 * (size_t)c + 1);
    return k;
}
__device__ __forceinline__ bool key_has_nan(const Key& k) {
    return k.a.x != k.a.x || k.a.y != k.a.y || k.a.z != k.a.z || k.a.w != k.a.w || k.b.x != k.b.x || k.b.y != k.b.y || k.b.z != k.b.z ||
           k.b.w != k.b.w;
}
__device__ __forceinline__ bool key_equal(const Key& p, const Key& q) {
    return p.a.x == q.a.x && p.a.y == q.a.y && p.a.z == q.a.z && p.a.w == q.a.w && p.b.x == q.b.x && p.b.y == q.b.y && p.b.z == q.b.z &&
           p.b.w == q.b.w;
}
__device__ __forceinline__ uint32_t rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__device__ __forceinline__ uint32_t mix_word(uint32_t h, float f) {
    uint32_t k = f == 0.0f ? 0u : __float_as_uint(f);   // -0 == +0 must hash alike
    k *= 0xcc9e2d51u;
    k = rotl(k, 15);
    k *= 0x1b873593u;
    h ^= k;
    h = rotl(h, 13);
    return h * 5u + 0xe6546b64u;
}
__device__ __forceinline__ uint32_t key_hash(const Key& k) {
    uint32_t h = 0x9747b28cu;
    h = mix_word(h, k.a.x); h = mix_word(h, k.a.y); h = mix_word(h, k.a.z); h = mix_word(h, k.a.w);
    h = mix_word(h, k.b.x); h = mix_word(h, k.b.y); h = mix_word(h, k.b.z); h = mix_word(h, k.b.w);
    h ^= 32u;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}

// ---- first occurrences -------------------------------------------------------------------------------------------------------------
// first[c] <- the slot corner c's class lives in (kEmpty for a NaN corner)
__global__ void __launch_bounds__(kThreads) k_dedup_insert(const float4* __restrict__ keys, uint32_t* table, uint32_t mask, uint32_t C,
                                                           uint32_t* __restrict__ first) {
    const uint32_t c = blockIdx.x * (uint32_t)kThreads + threadIdx.x;
    if (c >= C) return;
    const Key k = load_key(keys, c);
    if (key_has_nan(k)) { first[c] = kEmpty; return; }
    uint32_t slot = key_hash(k) & mask;
    for (;;) {
        uint32_t o = reinterpret_cast<volatile uint32_t*>(table)[slot];
        if (o == kEmpty) o = atomicCAS(table + slot, kEmpty, c);
        if (o == kEmpty) break;                                   // claimed
        // o may be stale (a later atomicMin lowered the slot) but it is a member of the slot's class all the same
        if (key_equal(load_key(keys, o), k)) {
            if (c < o) atomicMin(table + slot, c);
            break;
        }
        slot = (slot + 1) & mask;
    }
    first[c] = slot;
}
// first[c] <- the earliest corner with an equal key
__global__ void __launch_bounds__(kThreads) k_dedup_resolve(const uint32_t* __restrict__ table, uint32_t C, uint32_t* __restrict__ first) {
    const uint32_t c = blockIdx.x * (uint32_t)kThreads + threadIdx.x;
    if (c >= C) return;
    const uint32_t slot = first[c];
    first[c] = slot == kEmpty ? c : table[slot];
}

// ---- numbering ---------------------------------------------------------------------------------------------------------------------
// bit i of the result: corner c0 + i is its own first occurrence
__device__ __forceinline__ uint32_t own_flags(const uint32_t* __restrict__ first, uint32_t C, uint32_t c0) {
    uint32_t bits = 0;
    if (c0 + kPerThread <= C) {
        const uint4 lo = __ldg(reinterpret_cast<const uint4*>(first + c0));
        const uint4 hi = __ldg(reinterpret_cast<const uint4*>(first + c0) + 1);
        bits |= (uint32_t)(lo.x == c0 + 0) << 0; bits |= (uint32_t)(lo.y == c0 + 1) << 1;
        bits |= (uint32_t)(lo.z == c0 + 2) << 2; bits |= (uint32_t)(lo.w == c0 + 3) << 3;
        bits |= (uint32_t)(hi.x == c0 + 4) << 4; bits |= (uint32_t)(hi.y == c0 + 5) << 5;
        bits |= (uint32_t)(hi.z == c0 + 6) << 6; bits |= (uint32_t)(hi.w == c0 + 7) << 7;
    } else {
        for (int i = 0; i < kPerThread; i++)
            if (c0 + i < C && __ldg(first + c0 + i) == c0 + i) bits |= 1u << i;
    }
    return bits;
}
// exclusive prefix of v over the block's threads; *total = the block's sum (valid in every thread)
__device__ __forceinline__ uint32_t block_exclusive(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sum[kThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    uint32_t before = 0, all = 0;
    for (int w = 0; w < kThreads / 32; w++) {
        const uint32_t s = warp_sum[w];
        if (w < warp) before += s;
        all += s;
    }
    __syncthreads();
    *total = all;
    return before + incl - v;
}
__global__ void __launch_bounds__(kThreads) k_dedup_count(const uint32_t* __restrict__ first, uint32_t C, uint32_t* __restrict__ tile_sum) {
    const uint32_t c0 = blockIdx.x * (uint32_t)kTile + threadIdx.x * (uint32_t)kPerThread;
    uint32_t total;
    block_exclusive(__popc(own_flags(first, C, c0)), &total);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}
// tile_off[i] = sum of tile_sum[0, i), tile_off[n] = everything.  One block; n is corners / 2048.
__global__ void __launch_bounds__(1024) k_dedup_scan_tiles(const uint32_t* __restrict__ tile_sum, uint32_t n, uint32_t* __restrict__ tile_off) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n ? tile_sum[i] : 0u;
        uint32_t incl = v;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += up;
        }
        if (lane == 31) warp_sum[warp] = incl;
        __syncthreads();
        uint32_t before = 0;
        for (int w = 0; w < warp; w++) before += warp_sum[w];
        const uint32_t c = carry;
        if (i < n) tile_off[i] = c + before + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + before + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) tile_off[n] = carry;
}

// ---- the host's float arithmetic ------------------------------------------------------------------------------------------------
// Round-to-nearest, no contraction, and the NaN an x86 SSE unit would produce: an operand NaN comes back quieted with its payload,
// an invalid operation (0 * inf, sqrt of a negative) gives the NEGATIVE default NaN 0xffc00000 (the GPU's is 0x7fffffff).  With two
// NaN operands of different payloads SSE returns its first source operand and which one that is is the host compiler's choice;
// here it is the left one (such vertices need differently signed NaNs written out in the OBJ text).
__device__ __forceinline__ float quiet(float a) { return __uint_as_float(__float_as_uint(a) | 0x00400000u); }
__device__ __forceinline__ float host_nan(float a, float b, float r) {
    if (a != a) return quiet(a);
    if (b != b) return quiet(b);
    if (r != r) return __uint_as_float(0xffc00000u);
    return r;
}
__device__ __forceinline__ float host_mul(float a, float b) { return host_nan(a, b, __fmul_rn(a, b)); }
__device__ __forceinline__ float host_add(float a, float b) { return host_nan(a, b, __fadd_rn(a, b)); }
__device__ __forceinline__ float host_div(float a, float b) { return host_nan(a, b, __fdiv_rn(a, b)); }
__device__ __forceinline__ float host_sqrt(float a) { return host_nan(a, a, __fsqrt_rn(a)); }
// vecmath.hpp normalize: v * (1 / sqrt((x*x + y*y) + z*z))   (GLM detail/func_geometric.inl:154-159)
__device__ __forceinline__ float4 host_normalize(float x, float y, float z) {
    const float d = host_add(host_add(host_mul(x, x), host_mul(y, y)), host_mul(z, z));
    const float s = host_div(1.0f, host_sqrt(d));
    return make_float4(host_mul(x, s), host_mul(y, s), host_mul(z, s), 0.0f);
}

// ---- emit ---------------------------------------------------------------------------------------------------------------------------
struct PackOut {
    float4* pos;
    float4* nor;
    float4* tangent;
    float2* uv;
    float* raw_nor;       // 3 per vertex
    float* raw_tangent;   // 3 per vertex
};
// id[c] for every first occurrence c, and the vertex arrays; vertex v of this soup is written at [v], its id is vertex_base + v
__global__ void __launch_bounds__(kThreads) k_dedup_emit(const float4* __restrict__ keys, const float* __restrict__ tri_tangent,
                                                         const uint32_t* __restrict__ first, const uint32_t* __restrict__ tile_off, uint32_t C,
                                                         uint32_t vertex_base, uint32_t* __restrict__ id, PackOut out) {
    const uint32_t c0 = blockIdx.x * (uint32_t)kTile + threadIdx.x * (uint32_t)kPerThread;
    const uint32_t bits = own_flags(first, C, c0);
    uint32_t total;
    uint32_t v = tile_off[blockIdx.x] + block_exclusive(__popc(bits), &total);
    for (int i = 0; i < kPerThread; i++) {
        if (!(bits >> i & 1u)) continue;
        const uint32_t c = c0 + i;
        const Key k = load_key(keys, c);
        const uint32_t t = c / 3u;
        const float tx = __ldg(tri_tangent + 3 * (size_t)t), ty = __ldg(tri_tangent + 3 * (size_t)t + 1), tz = __ldg(tri_tangent + 3 * (size_t)t + 2);
        id[c] = vertex_base + v;
        out.pos[v] = make_float4(k.a.x, k.a.y, k.a.z, 1.0f);
        out.uv[v] = make_float2(k.b.z, k.b.w);
        out.raw_nor[3 * (size_t)v + 0] = k.a.w; out.raw_nor[3 * (size_t)v + 1] = k.b.x; out.raw_nor[3 * (size_t)v + 2] = k.b.y;
        out.raw_tangent[3 * (size_t)v + 0] = tx; out.raw_tangent[3 * (size_t)v + 1] = ty; out.raw_tangent[3 * (size_t)v + 2] = tz;
        out.nor[v] = host_normalize(k.a.w, k.b.x, k.b.y);         // upload-time normalisation (src/wavefront_internal.cu:136-144)
        out.tangent[v] = host_normalize(tx, ty, tz);
        v++;
    }
}
__global__ void __launch_bounds__(kThreads) k_dedup_indices(const uint32_t* __restrict__ first, const uint32_t* __restrict__ id,
                                                            const int32_t* __restrict__ material, uint32_t T, int4* __restrict__ indices_matid) {
    const uint32_t t = blockIdx.x * (uint32_t)kThreads + threadIdx.x;
    if (t >= T) return;
    const uint32_t a = __ldg(first + 3 * (size_t)t), b = __ldg(first + 3 * (size_t)t + 1), c = __ldg(first + 3 * (size_t)t + 2);
    indices_matid[t] = make_int4((int)__ldg(id + a), (int)__ldg(id + b), (int)__ldg(id + c), __ldg(material + t));
}

// ---- host side ----------------------------------------------------------------------------------------------------------------------
void check(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw std::runtime_error(std::string("cuda: ") + what + ": " + cudaGetErrorString(e));
}
struct DevBuf {
    void* p = nullptr;
    DevBuf() {}
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { if (p) cudaFree(p); }
    void alloc(size_t bytes, const char* what) { check(cudaMalloc(&p, bytes ? bytes : 1), what); }
    template <class T> T* as() const { return static_cast<T*>(p); }
};
struct DeviceGuard {
    int previous = -1;
    explicit DeviceGuard(int device) {
        check(cudaGetDevice(&previous), "cudaGetDevice");
        check(cudaSetDevice(device), "cudaSetDevice");
    }
    ~DeviceGuard() { if (previous >= 0) cudaSetDevice(previous); }
};
double ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

void gpu_dedup_pack(int device, const float* keys, const float* tangents, const int32_t* materials, size_t triangles, ptb_scene& s,
                    GpuDedupTimes* times) {
    const size_t T = triangles, C = 3 * T;
    if (T == 0) return;
    if (C > (1ull << 30)) throw std::runtime_error("scene too large for the GPU loader: more than 2^30 triangle corners");
    const size_t base = s.pos.size();
    if (base + C >= 0x7fffffffull) throw std::runtime_error("scene too large: vertex ids would not fit 31 bits");
    int count = 0;
    check(cudaGetDeviceCount(&count), "cudaGetDeviceCount");
    if (device < 0 || device >= count)
        throw std::runtime_error("cuda: device ordinal " + std::to_string(device) + " out of range (" + std::to_string(count) + " devices)");
    DeviceGuard guard(device);

    size_t cap = 1024;
    while (cap < 2 * C) cap <<= 1;
    const uint32_t tiles = (uint32_t)((C + kTile - 1) / kTile);

    auto t0 = std::chrono::steady_clock::now();
    DevBuf d_keys, d_tan, d_mat, d_table, d_first, d_id, d_tile_sum, d_tile_off;
    d_keys.alloc(C * 32, "cudaMalloc(keys)");
    d_tan.alloc(T * 12, "cudaMalloc(tangents)");
    d_mat.alloc(T * 4, "cudaMalloc(materials)");
    d_table.alloc(cap * 4, "cudaMalloc(table)");
    d_first.alloc(C * 4, "cudaMalloc(first)");
    d_id.alloc(C * 4, "cudaMalloc(id)");
    d_tile_sum.alloc((size_t)tiles * 4, "cudaMalloc(tile sums)");
    d_tile_off.alloc(((size_t)tiles + 1) * 4, "cudaMalloc(tile offsets)");
    check(cudaMemcpy(d_keys.p, keys, C * 32, cudaMemcpyHostToDevice), "upload keys");
    check(cudaMemcpy(d_tan.p, tangents, T * 12, cudaMemcpyHostToDevice), "upload tangents");
    check(cudaMemcpy(d_mat.p, materials, T * 4, cudaMemcpyHostToDevice), "upload materials");
    const double upload_ms = ms_since(t0);

    t0 = std::chrono::steady_clock::now();
    check(cudaMemset(d_table.p, 0xff, cap * 4), "clear table");
    const uint32_t corner_blocks = (uint32_t)((C + kThreads - 1) / kThreads);
    k_dedup_insert<<<corner_blocks, kThreads>>>(d_keys.as<float4>(), d_table.as<uint32_t>(), (uint32_t)(cap - 1), (uint32_t)C, d_first.as<uint32_t>());
    k_dedup_resolve<<<corner_blocks, kThreads>>>(d_table.as<uint32_t>(), (uint32_t)C, d_first.as<uint32_t>());
    k_dedup_count<<<tiles, kThreads>>>(d_first.as<uint32_t>(), (uint32_t)C, d_tile_sum.as<uint32_t>());
    k_dedup_scan_tiles<<<1, 1024>>>(d_tile_sum.as<uint32_t>(), tiles, d_tile_off.as<uint32_t>());
    check(cudaGetLastError(), "de-duplication kernels");
    uint32_t V32 = 0;
    check(cudaMemcpy(&V32, d_tile_off.as<uint32_t>() + tiles, 4, cudaMemcpyDeviceToHost), "read vertex count");
    const size_t V = V32;
    DevBuf d_pos, d_nor, d_tangent, d_uv, d_raw_nor, d_raw_tangent, d_idx;
    d_pos.alloc(V * 16, "cudaMalloc(pos)");
    d_nor.alloc(V * 16, "cudaMalloc(nor)");
    d_tangent.alloc(V * 16, "cudaMalloc(tangent)");
    d_uv.alloc(V * 8, "cudaMalloc(uv)");
    d_raw_nor.alloc(V * 12, "cudaMalloc(raw normals)");
    d_raw_tangent.alloc(V * 12, "cudaMalloc(raw tangents)");
    d_idx.alloc(T * 16, "cudaMalloc(indices)");
    PackOut out{d_pos.as<float4>(), d_nor.as<float4>(), d_tangent.as<float4>(), d_uv.as<float2>(), d_raw_nor.as<float>(), d_raw_tangent.as<float>()};
    k_dedup_emit<<<tiles, kThreads>>>(d_keys.as<float4>(), d_tan.as<float>(), d_first.as<uint32_t>(), d_tile_off.as<uint32_t>(), (uint32_t)C,
                                      (uint32_t)base, d_id.as<uint32_t>(), out);
    k_dedup_indices<<<(uint32_t)((T + kThreads - 1) / kThreads), kThreads>>>(d_first.as<uint32_t>(), d_id.as<uint32_t>(), d_mat.as<int32_t>(),
                                                                             (uint32_t)T, d_idx.as<int4>());
    check(cudaGetLastError(), "pack kernels");
    check(cudaDeviceSynchronize(), "de-duplication and pack");
    const double device_ms = ms_since(t0);

    t0 = std::chrono::steady_clock::now();
    static_assert(sizeof(f4) == 16 && sizeof(f2) == 8 && sizeof(i4) == 16 && sizeof(V3) == 12, "scene containers are in upload layout");
    const size_t tri_base = s.indices_matid.size();
    {   // growing a vector touches every new page once: one thread per array (half a gigabyte at 10 M triangles)
        std::vector<std::thread> grow;
        grow.emplace_back([&] { s.pos.resize(base + V); });
        grow.emplace_back([&] { s.nor.resize(base + V); });
        grow.emplace_back([&] { s.tangent.resize(base + V); });
        grow.emplace_back([&] { s.uv.resize(base + V); });
        grow.emplace_back([&] { s.raw_nor.resize(base + V); });
        grow.emplace_back([&] { s.raw_tangent.resize(base + V); });
        s.indices_matid.resize(tri_base + T);
        for (std::thread& t : grow) t.join();
    }
    if (V) {
        check(cudaMemcpy(&s.pos[base], d_pos.p, V * 16, cudaMemcpyDeviceToHost), "download pos");
        check(cudaMemcpy(&s.nor[base], d_nor.p, V * 16, cudaMemcpyDeviceToHost), "download nor");
        check(cudaMemcpy(&s.tangent[base], d_tangent.p, V * 16, cudaMemcpyDeviceToHost), "download tangent");
        check(cudaMemcpy(&s.uv[base], d_uv.p, V * 8, cudaMemcpyDeviceToHost), "download uv");
        check(cudaMemcpy(&s.raw_nor[base], d_raw_nor.p, V * 12, cudaMemcpyDeviceToHost), "download raw normals");
        check(cudaMemcpy(&s.raw_tangent[base], d_raw_tangent.p, V * 12, cudaMemcpyDeviceToHost), "download raw tangents");
    }
    check(cudaMemcpy(&s.indices_matid[tri_base], d_idx.p, T * 16, cudaMemcpyDeviceToHost), "download indices");
    if (times) {
        times->upload_ms = upload_ms;
        times->device_ms = device_ms;
        times->download_ms = ms_since(t0);
        times->corners = C;
        times->vertices = V;
        times->table_slots = cap;
    }
}

}  // namespace ptb
