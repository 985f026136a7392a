// Philox4x32-10 and the torch-CUDA exponential_ element layout.
// Spec: Salmon et al. SC'11; constants as torch ATen/core/PhiloxRNGEngine.h:229-232.
// Element layout: torch ATen/native/cuda/DistributionTemplates.h:50-90 (grid-stride
// kernel, block 256, unroll 4, thread idx = Philox subsequence, curand_uniform4).
// Transform: torch ATen/core/TransformationHelper.h:129-146.
#pragma once
#include <stdint.h>

namespace ocppo {

struct PhiloxLayout {
    uint64_t seed;
    uint64_t offset;     // torch generator offset (in 32-bit outputs; multiple of 4)
    uint64_t threads;    // G = grid * 256 of torch's launch for this numel
};

__host__ __device__ inline uint64_t torch_philox_grid(int64_t numel, int sm_count, int max_threads_per_sm) {
    uint64_t grid = ((uint64_t)numel + 255) / 256;
    uint64_t cap = (uint64_t)sm_count * (uint64_t)(max_threads_per_sm / 256);
    return grid < cap ? grid : cap;
}

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}

// The 32-bit word torch feeds element `li` of an exponential_/uniform_ call.
__device__ __forceinline__ uint32_t torch_philox_word(const PhiloxLayout &L, uint64_t li) {
    const uint64_t idx = li % L.threads;
    const uint64_t slot = li / L.threads;            // ii + 4*call
    const uint64_t c64 = L.offset / 4 + slot / 4;
    const uint4 ctr = make_uint4((uint32_t)c64, (uint32_t)(c64 >> 32), (uint32_t)idx, (uint32_t)(idx >> 32));
    const uint2 key = make_uint2((uint32_t)L.seed, (uint32_t)(L.seed >> 32));
    const uint4 o = philox4x32_10(ctr, key);
    const uint32_t w[4] = {o.x, o.y, o.z, o.w};
    return w[slot & 3];
}

// curand_uniform: (0, 1]
__device__ __forceinline__ float curand_uniform_from_word(uint32_t x) {
    return __fmaf_rn((float)x, 2.3283064e-10f, 2.3283064e-10f / 2.0f);
}

// q ~ Exp(1) exactly as torch's CUDA exponential_ for fp32
__device__ __forceinline__ float torch_exponential_from_uniform(float u) {
    const float eps_half = 1.1920928955078125e-07f / 2.0f;
    // at::log on CUDA is the fast __logf (torch ATen/NumericUtils.h), not logf
    const float lg = (u >= 1.0f - eps_half) ? -eps_half : __logf(u);
    return -1.0f * lg;
}

}  // namespace ocppo
