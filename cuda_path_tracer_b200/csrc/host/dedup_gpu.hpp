// dedup_gpu.hpp -- vertex de-duplication and the AoS -> SoA pack of a parsed triangle soup on the GPU (csrc/cuda/dedup.cu).
#pragma once
#include <cstddef>
#include <cstdint>

struct ptb_scene;

namespace ptb {

struct GpuDedupTimes {
    double upload_ms = 0, device_ms = 0, download_ms = 0;
    size_t corners = 0, vertices = 0, table_slots = 0;
};

// The job of Scene::loadFromJSON's per-corner unordered_map walk (src/scene.cpp:330-539, the numbering of :525-532) and of
// WavefrontPathTracerState::initSceneGeometry's repack (src/wavefront_internal.cu:118-167), on GPU `device`:
//   keys      3 * triangles corner keys of 8 floats (pos.xyz, nor.xyz, uv.xy) in file order
//   tangents  one float3 per triangle, materials one id per triangle
// Appends to s.pos / nor / tangent / uv / raw_nor / raw_tangent / indices_matid exactly what the host path
// (dedup_and_emit + the normalisation loop in scene.cpp) appends.  Throws std::runtime_error("cuda: ...") when the device cannot
// be used -- there is no fall-back to the host path from here.
void gpu_dedup_pack(int device, const float* keys, const float* tangents, const int32_t* materials, size_t triangles, ptb_scene& s,
                    GpuDedupTimes* times);

}  // namespace ptb
