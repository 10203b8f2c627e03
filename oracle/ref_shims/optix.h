// Harness-side stand-in for the OptiX SDK header (test infrastructure only).
// The reference's wavefront front-end includes optix/optix_ray_cast.h unconditionally
// (pathtrace_wavefront.cu:24) although USE_OPTIX is off (:27).  The SDK is not on this
// image and B200 has no RT cores, so only the handful of type names the class
// declaration needs are provided here; no OptiX entry point is ever called.
#pragma once
#include <cuda.h>
typedef unsigned long long OptixTraversableHandle;
typedef struct OptixPipeline_t* OptixPipeline;
struct OptixShaderBindingTable { void* a[8]; unsigned int b[8]; };
