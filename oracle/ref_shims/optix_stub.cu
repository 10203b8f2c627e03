// No-op replacement for src/optix/optix_ray_cast.cu (test infrastructure only).
// PathtraceInit constructs and Init()s the caster unconditionally
// (pathtrace_wavefront.cu:156-160) but never traces with it unless USE_OPTIX is defined.
#include "optix/optix_ray_cast.h"
OptixRayCast::OptixRayCast() {}
OptixRayCast::~OptixRayCast() {}
void OptixRayCast::Init(const WavefrontPathTracerState*) { m_initialized = true; }
void OptixRayCast::TraceExtensionRay(int, const WavefrontPathTracerState*) {}
void OptixRayCast::TraceShadowRay(int, const WavefrontPathTracerState*) {}
