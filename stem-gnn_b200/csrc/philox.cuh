// philox.cuh -- counter-based random numbers shared by K7 (dropout), K9 (edge dropout / feature
// masking) and K10 (neighbour sampling): every draw is a pure function of (seed, counter), so a
// backward pass or a test can recompute it and nothing is stored.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace stemgnn {

// Philox4x32 with 7 rounds (Random123's "crush-resistant" minimum; curand's default is 10): the
// K7 forward apply pass with dropout measured 83 us against 65 us without -- the rounds are visible
__device__ __forceinline__ uint4 philox4x32_7(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}

// four 32-bit draws for 64-bit counter `ctr` in stream `sub` of `seed`
__device__ __forceinline__ uint4 philox_draw4(uint64_t seed, uint64_t ctr, uint32_t sub) {
    return philox4x32_7(make_uint4((uint32_t)ctr, (uint32_t)(ctr >> 32), sub, 0u),
                        make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
}

// uniform in [0, 1) with 24 random bits (what torch.rand produces for fp32)
__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * (1.f / 16777216.f); }

// uniform float of element `i` of a stream: elements 4q..4q+3 share one Philox block
__device__ __forceinline__ float philox_u01_at(uint64_t seed, uint64_t i, uint32_t sub) {
    const uint4 r = philox_draw4(seed, i >> 2, sub);
    const uint32_t k = (uint32_t)i & 3u;
    return u01(k == 0 ? r.x : k == 1 ? r.y : k == 2 ? r.z : r.w);
}

}  // namespace stemgnn
