// lbvh.cuh -- device LBVH build: interface between the kernels (lbvh.cu) and the C ABI glue.
// Replaces BuildLBVH (reference src/bvh.cu:456-516).
#pragma once
#include "common.cuh"

namespace ptb {

struct LbvhBuffers {
    // reference-layout arrays (LBVHData, src/scene_structs.h:151-169), all device pointers
    LbvhRef ref{};
    // traversal layout (see BvhDev in common.cuh)
    float4* node = nullptr;      // 4 * (N-1)
    float4* tri = nullptr;       // PTB_TRI_STRIDE * N
    uint4* wnode = nullptr;      // 4 * (N-1): 4-wide quantised nodes (BvhDev::wnode), indexed by the binary node they collapse
    uint4* bnode = nullptr;      // 2 * (N-1): 32-byte binary nodes with 15-bit boxes (BvhDev::bnode)
    // scratch kept for rebuilds
    unsigned int* key_a = nullptr;   // 30-bit Morton codes, ping
    unsigned int* key_b = nullptr;   // pong
    int* val_b = nullptr;            // pong for primitive indices
    int* counts = nullptr;           // 256 * num_tiles
    int* totals = nullptr;           // 256
    int* flags = nullptr;            // N refit arrival flags
    unsigned int* world_enc = nullptr;   // 8 order-preserving encoded floats (min xyzw, max xyzw)
    int* max_depth = nullptr;        // 1
    int num_triangles = 0;
    int num_tiles = 0;
    size_t bytes = 0;                // total device bytes allocated
};

// Allocation / release.  Return cudaError_t as int (0 = ok).
int lbvh_alloc(LbvhBuffers& b, int num_triangles);
void lbvh_free(LbvhBuffers& b);

// Enqueue the whole build on `stream`.  pos: float4[V], indices_matid: int4[N].
// No host synchronisation inside; launches are counted into *launches if non-null.
int lbvh_build(LbvhBuffers& b, const float4* pos, const int4* indices_matid, cudaStream_t stream, int* launches);

}  // namespace ptb
