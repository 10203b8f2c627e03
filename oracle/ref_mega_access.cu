// oracle/ref_mega_access.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// The reference's megakernel front-end (pathtrace_megakernel::Pathtrace, src/pathtrace.cu:645-706) keeps its image
// in TU-local statics (`d_image`, `d_pixel_sample_count`, src/pathtrace.cu:479-480) and shows it only through the
// 8-bit PBO kernel.  Parity of the float sums needs the buffers themselves, so this TU is the reference's own
// pathtrace.cu by textual inclusion (resolved through -I$(REF)/src; not a copy) plus two accessors in its namespace --
// the same device as ref_harness.cu uses for the wavefront state.  It takes the place of pathtrace.o in the link.
#include "pathtrace.cu"

namespace pathtrace_megakernel {
    glm::vec3* OracleImage() { return d_image; }
    int* OracleSampleCount() { return d_pixel_sample_count; }
}
