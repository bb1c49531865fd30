// philox.h -- counter-based Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as
// 1, 2, 3", SC'11) and the Box-Muller transform used for every noise draw of the Philox mode.
// Host + device.
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

MCSIM_PHD void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {  // @phase philox
#ifdef __CUDA_ARCH__
    lo = a * b;
    hi = __umulhi(a, b);
#else
    const uint64_t p = uint64_t(a) * uint64_t(b);
    lo = uint32_t(p);
    hi = uint32_t(p >> 32);
#endif
}

MCSIM_PHD u32x4 philox4x32_10(u32x4 c, uint32_t k0, uint32_t k1) {  // @phase philox
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

// ---- step draws: 12 standard normals of one rollout-step from TWO Philox blocks ------------------
// A rollout-step needs 3 odometry + L observation normals (11 at the reference's 8 landmarks).  Six
// Box-Muller pairs at 24 radius bits + 16 angle bits each use 240 of the 256 bits of two blocks
// (a, b) instead of three blocks at 32 + 32 bits: the FP32 uniform keeps 24 bits either way, the
// radius reaches sqrt(2 * 25 ln 2) = 5.89 sigma (P(|z| > 5.89) = 4e-9 per draw) and the angle has
// 65,536 directions, both far below the Monte-Carlo resolution of a run.  Layout (shared by the packed
// kernel, the scalar audit kernel and the scoring kernel, so their rollouts are the same experiments):
//   radius of pair j = 0..3: a.{x,y,z,w} >> 8;  pair 4: b.x >> 8;  pair 5: low bytes of a.x, a.y, a.z
//   angle  of pair j: half (j & 1) of b.{y,z,w}[j >> 1]
MCSIM_PHD void step_bits(const u32x4& a, const u32x4& b, uint32_t rad[6], uint32_t ang[6]) {  // @phase draw_bits
    rad[0] = a.x >> 8; rad[1] = a.y >> 8; rad[2] = a.z >> 8; rad[3] = a.w >> 8; rad[4] = b.x >> 8;
#ifdef __CUDA_ARCH__
    rad[5] = __byte_perm(__byte_perm(a.x, a.y, 0x0400), a.z, 0x4210) >> 8;
#else
    rad[5] = (a.x & 0xffu) | ((a.y & 0xffu) << 8) | ((a.z & 0xffu) << 16);
#endif
    ang[0] = b.y & 0xffffu; ang[1] = b.y >> 16;
    ang[2] = b.z & 0xffffu; ang[3] = b.z >> 16;
    ang[4] = b.w & 0xffffu; ang[5] = b.w >> 16;
}
// uniform in (0, 1] from 24 bits: (r + 1/2) 2^-24;  angle in (0, 2 pi) from 16 bits: (a + 1/2) 2 pi 2^-16
#define MCSIM_U24_SCALE 5.9604644775390625e-8f
#define MCSIM_U24_HALF 2.98023223876953125e-8f
#define MCSIM_A16_SCALE 9.587379924285257e-5f
#define MCSIM_A16_HALF 4.7936899621426287e-5f

// ---- GMM sample draws: the 3 normals of TWO samples from ONE Philox block ---------------------------
// Samples 2 j and 2 j + 1 of a waypoint share the block with counter (j, stream, 2 t, PURPOSE_GMM_SAMPLE): three
// Box-Muller pairs at 24 + 16 bits use 120 of its 128 bits (a block per sample used 80 and threw a fourth normal away;
// Philox was 22 % of the batch kernel's instructions).  Layout (shared by the packed and the scalar sample kernels):
//   pair A -> z0, z1 of the even sample: radius c.x >> 8, angle c.z & 0xffff
//   pair B -> z0, z1 of the odd sample:  radius c.y >> 8, angle c.z >> 16
//   pair C -> z2 of both (cos: even, sin: odd): radius c.w >> 8, angle (c.x & 0xff) << 8 | (c.y & 0xff)
MCSIM_PHD uint32_t gmm_pair_c_angle(const u32x4& c) { return ((c.x & 0xffu) << 8) | (c.y & 0xffu); }

#ifdef __CUDACC__
// two standard normals from the 24 + 16 bits of one step pair (scalar kernels)
__device__ __forceinline__ void box_muller_bits(uint32_t rad, uint32_t ang, float& n0, float& n1) {
    const float r = sqrtf(-2.0f * __logf(fmaf(float(rad), MCSIM_U24_SCALE, MCSIM_U24_HALF)));
    float s, c;
    __sincosf(fmaf(float(ang), MCSIM_A16_SCALE, MCSIM_A16_HALF), &s, &c);
    n0 = r * c;
    n1 = r * s;
}
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
