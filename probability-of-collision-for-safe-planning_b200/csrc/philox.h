// philox.h -- counter-based Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as
// 1, 2, 3", SC'11) and the Box-Muller transform used for every noise draw of the Philox mode.
// Host + device; the CPU oracle port (oracle/port/) has its own independent restatement.
//
// Replaces, for Philox mode, the reference's global Mersenne-Twister stream behind arma::randn()
// (armadillo-8.400.0/include/armadillo_bits/arma_rng.hpp:320-436) with one independent subsequence
// per rollout:   counter = (rollout id lo, rollout id hi, block, purpose),  key = seed.
#ifndef MCSIM_PHILOX_H
#define MCSIM_PHILOX_H

#include <stdint.h>

#ifdef __CUDACC__
#define MCSIM_PHD __host__ __device__ __forceinline__
#else
#define MCSIM_PHD inline
#endif

namespace mcsim_rng {

struct u32x4 { uint32_t x, y, z, w; };

MCSIM_PHD void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#ifdef __CUDA_ARCH__
    lo = a * b;
    hi = __umulhi(a, b);
#else
    const uint64_t p = uint64_t(a) * uint64_t(b);
    lo = uint32_t(p);
    hi = uint32_t(p >> 32);
#endif
}

MCSIM_PHD u32x4 philox4x32_10(u32x4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        mulhilo(0xD2511F53u, c.x, hi0, lo0);
        mulhilo(0xCD9E8D57u, c.z, hi1, lo1);
        u32x4 n;
        n.x = hi1 ^ c.y ^ k0;
        n.y = lo1;
        n.z = hi0 ^ c.w ^ k1;
        n.w = lo0;
        c = n;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}

// purposes (counter word w)
enum { PURPOSE_STEP = 0, PURPOSE_INIT = 1, PURPOSE_GMM_SAMPLE = 2, PURPOSE_GMM_HOST = 3 };

// uniform in (0, 1]: x * 2^-32 + 2^-33
MCSIM_PHD float u01(uint32_t x) { return float(x) * 2.3283064365386963e-10f + 1.1641532182693481e-10f; }

#ifdef __CUDACC__
// two standard normals from two 32-bit words (Box-Muller); fast-math intrinsics: the absolute
// error of __sincosf on [0, 2 pi] and the relative error of __logf are ~1e-6, far below the
// Monte-Carlo resolution this mode is used at.
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& n0, float& n1) {
    const float r = sqrtf(-2.0f * __logf(u01(a)));
    float s, c;
    __sincosf(6.283185307179586f * u01(b), &s, &c);
    n0 = r * c;
    n1 = r * s;
}
#endif

}  // namespace mcsim_rng
#endif
