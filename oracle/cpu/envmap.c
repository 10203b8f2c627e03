/* envmap.c -- CPU restatement of sampleEnvironmentMap (reference src/interactions.cu:447-488): alias-table lookup,
 * texel centre -> (u, v) -> spherical direction, pdf read from the solid-angle pdf map the host loader builds
 * (src/scene.cpp:270-299).  TEST INFRASTRUCTURE: imported only by tests/ (see oracle/Makefile).
 * The device reads the pdf with a linear-filtered tex2D at the texel centre, which returns the texel itself; here it
 * is the array element.  sinf / cosf are libm's, the device's are CUDA's: directions agree to a few ulp, not bitwise;
 * the chosen texel and the pdf must agree exactly. */
#include <math.h>
#include <stdint.h>

#define PT_PI 3.1415926535897932384626422832795028841971f   /* src/common/utilities.h:12 */

void ptb_oracle_sample_environment(const float* probs, const int32_t* aliases, const float* pdf_map, int width, int height,
                                   const float* rnd, int count, float* dir, float* pdf, int32_t* pixel) {
    const int N = width * height;
    for (int i = 0; i < count; i++) {
        float u = rnd[2 * i] * N;                       /* :455 */
        int idx = (int)u;
        if (idx > N - 1) idx = N - 1;                    /* :456 */
        float xi = u - idx;                              /* :457 */
        float prob = probs[idx];
        int p = (xi < prob) ? idx : aliases[idx];        /* :459-460 */
        int py = p / width, px = p % width;              /* :463-464 */
        float u_coord = (px + 0.5f) / width;             /* :467-468 */
        float v_coord = (py + 0.5f) / height;
        pdf[i] = pdf_map[p];                             /* :472-473 */
        float phi = u_coord * 2.0f * PT_PI;              /* :476-477 */
        float theta = v_coord * PT_PI;
        dir[3 * i + 0] = sinf(theta) * cosf(phi);        /* :481-484, Y up */
        dir[3 * i + 1] = cosf(theta);
        dir[3 * i + 2] = sinf(theta) * sinf(phi);
        pixel[i] = p;
    }
}
