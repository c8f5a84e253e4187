// usv_philox.cuh -- counter-based RNG for in-kernel env reset.
//
// Philox4x32-10, key = 64-bit job seed, counter = (global env id lo, hi, episode, block).  A
// stream of uniforms for one (env, episode) therefore does not depend on which GPU owns the env
// or on how many GPUs share the job.  The reference draws from numpy's PCG64 (simple_env.py:229);
// matching that bit stream on a GPU is not a goal -- the DISTRIBUTIONS are matched
// (tests/test_oracle_golden.py::test_reset_distributions_match_reference_samples) and the CPU oracle uses this same generator so reset states
// can be compared value for value.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace usv {

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

// uniform double in [0,1) from 53 random bits
__device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}

struct DrawStream {
    uint2 key;
    uint32_t gid_lo, gid_hi, episode, block;
    bool have;
    double spare;

    __device__ DrawStream(uint64_t seed, uint64_t gid, uint32_t ep)
        : key(make_uint2((uint32_t)seed, (uint32_t)(seed >> 32))), gid_lo((uint32_t)gid), gid_hi((uint32_t)(gid >> 32)),
          episode(ep), block(0), have(false), spare(0.0) {}

    __device__ double u01() {
        if (have) {
            have = false;
            return spare;
        }
        const uint4 o = philox4x32_10(make_uint4(gid_lo, gid_hi, episode, block++), key);
        spare = u53(o.z, o.w);
        have = true;
        return u53(o.x, o.y);
    }
    __device__ double uniform(double lo, double hi) { return lo + (hi - lo) * u01(); }
    // Box-Muller pair
    __device__ void normal2(double& z0, double& z1) {
        const double u1 = u01(), u2 = u01();
        const double rad = sqrt(-2.0 * log(1.0 - u1));
        double s, c;
        sincos(6.283185307179586 * u2, &s, &c);
        z0 = rad * c;
        z1 = rad * s;
    }
};

// Random access to the same stream: draw number k of (env, episode) is half (k & 1) of Philox block
// (k >> 1).  The warp-cooperative reset of the sensor kernel uses this so that 32 lanes produce the
// draws of one env in parallel and still reproduce DrawStream's sequence value for value.
struct DrawKey {
    uint2 key;
    uint32_t gid_lo, gid_hi, episode;
    __device__ DrawKey(uint64_t seed, uint64_t gid, uint32_t ep)
        : key(make_uint2((uint32_t)seed, (uint32_t)(seed >> 32))), gid_lo((uint32_t)gid), gid_hi((uint32_t)(gid >> 32)),
          episode(ep) {}
    // both uniforms of one block: draws 2*b and 2*b + 1
    __device__ __forceinline__ void pair(uint32_t b, double& first, double& second) const {
        const uint4 o = philox4x32_10(make_uint4(gid_lo, gid_hi, episode, b), key);
        first = u53(o.x, o.y);
        second = u53(o.z, o.w);
    }
};

}  // namespace usv
