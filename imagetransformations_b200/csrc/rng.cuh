// On-chip random variates for the throughput mode of the noise corruptions (Philox4x32-10 streams).
// These never feed parity tests; they are validated statistically (tests/test_gpu_noise_stats.py).
#pragma once
#include "common.cuh"

namespace imgtx {
#ifdef __CUDACC__

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
  const float u1 = u01(a);                                   // (0,1]
  const float u2 = (float)(b >> 8) * (1.0f / 16777216.0f);   // [0,1)
  const float r = sqrtf(-2.0f * __logf(u1));
  float s, c;
  sincospif(2.0f * u2, &s, &c);
  z0 = r * c;
  z1 = r * s;
}

// The same transform on the special-function unit (MUFU.LG2 / RSQ / SIN / COS): absolute error ~1e-6, a few instructions
// per pair instead of ~25.  For the throughput kernels, whose draws no oracle replays.
__device__ __forceinline__ void box_muller_fast(uint32_t a, uint32_t b, float& z0, float& z1) {
  const float u1 = u01(a);                                   // (0,1]
  const float ang = (float)(b >> 8) * (6.283185307179586f / 16777216.0f);
  const float r = sqrtf(-1.3862943611198906f * __log2f(u1));
  z0 = r * __cosf(ang);
  z1 = r * __sinf(ang);
}

// A small buffered uniform stream on top of Philox: counter = (c0, c1, c2, block++).
struct UniformStream {
  const Philox& rng;
  uint32_t c0, c1, c2, blk;
  uint4 buf;
  int left;
  __device__ __forceinline__ UniformStream(const Philox& r, uint32_t a, uint32_t b, uint32_t c, uint32_t start)
      : rng(r), c0(a), c1(b), c2(c), blk(start), left(0) {}
  __device__ __forceinline__ float next() {
    if (left == 0) { buf = rng(c0, c1, c2, blk++); left = 4; }
    uint32_t v = buf.x;
    buf.x = buf.y; buf.y = buf.z; buf.z = buf.w;
    --left;
    return u01(v);
  }
};

// Poisson(lam): Knuth's product method below 10, Hoermann's PTRS transformed rejection above
// (the same split numpy's legacy generator uses).
__device__ inline int poisson_sample(const Philox& rng, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t& ctr, float lam) {
  if (!(lam > 0.0f)) return 0;
  UniformStream us_(rng, c0, c1, c2, ctr);
  int result;
  if (lam < 10.0f) {
    const float enlam = __expf(-lam);
    int k = 0;
    float prod = us_.next();
    while (prod > enlam) { ++k; prod *= us_.next(); }
    result = k;
  } else {
    const float slam = sqrtf(lam), loglam = __logf(lam);
    const float b = 0.931f + 2.53f * slam;
    const float a = -0.059f + 0.02483f * b;
    const float invalpha = 1.1239f + 1.1328f / (b - 3.4f);
    const float vr = 0.9277f - 3.6224f / (b - 2.0f);
    for (;;) {
      const float U = us_.next() - 0.5f;
      const float V = us_.next();
      const float us = 0.5f - fabsf(U);
      const int k = (int)floorf((2.0f * a / us + b) * U + lam + 0.43f);
      if (us >= 0.07f && V <= vr) { result = k; break; }
      if (k < 0 || (us < 0.013f && V > us)) continue;
      if ((__logf(V) + __logf(invalpha) - __logf(a / (us * us) + b)) <= (-lam + k * loglam - lgammaf((float)k + 1.0f))) {
        result = k;
        break;
      }
    }
  }
  ctr = us_.blk;
  return result;
}

#endif
}  // namespace imgtx
