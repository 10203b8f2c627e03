/* oracle/cpu/raygen.c -- TEST INFRASTRUCTURE ONLY.
 * Restatement of the integer part of GenerateCameraRaysKernel (src/kernels/ray_gen.cu:36-57,75) and of
 * the RNG (src/rng.h:6-27): per-pixel seed chain and jitter.  Integer arithmetic + one exact float
 * scaling, so bit-comparable with the device. */
#include <stdint.h>

static uint32_t wang_hash(uint32_t seed) {
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}

/* For pixel_idx in [first, first+count): rng_state after RayGen and the two jitters. */
void ptb_oracle_raygen_seeds(int first, int count, int iter, uint32_t* rng_state, float* jitter_x, float* jitter_y) {
    for (int k = 0; k < count; k++) {
        uint32_t pixel = (uint32_t)(first + k);
        uint32_t seed = wang_hash(pixel * 19990303u + (uint32_t)iter * 719393u);
        if (seed == 0) seed = 1;
        seed = wang_hash(seed);
        jitter_x[k] = ((float)seed / 4294967296.0f) - 0.5f;
        seed = wang_hash(seed);
        jitter_y[k] = ((float)seed / 4294967296.0f) - 0.5f;
        rng_state[k] = seed;
    }
}

/* xorshift32 rand_float (src/rng.h:18-27): advances *state, returns the float draw. */
float ptb_oracle_rand_float(uint32_t* state) {
    uint32_t s = *state;
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
    *state = s;
    return (float)s * 2.3283064365386963e-10f;
}
