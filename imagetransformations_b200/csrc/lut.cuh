// Exact per-value formulas of the pointwise uint8 ops (shared by the pointwise kernels and the fused
// AugMix / TTA chains).  SURVEY.md App. A.4, A.4b, A.5.
#pragma once
#include "common.cuh"

namespace imgtx {
#ifdef __CUDACC__

// Pillow ImagingBlend(degenerate d, image p, alpha f) for one byte (libImaging/Blend.c semantics):
// f32 arithmetic, truncation inside [0,1], clip-then-truncate outside.
__device__ __forceinline__ uint8_t pil_blend(int d, int p, float f) {
  if (f == 0.0f) return (uint8_t)d;
  if (f == 1.0f) return (uint8_t)p;
  float t = __fadd_rn((float)d, __fmul_rn(f, (float)(p - d)));
  if (f >= 0.0f && f <= 1.0f) return (uint8_t)__float2int_rz(t);
  if (t <= 0.0f) return 0;
  if (t >= 255.0f) return 255;
  return (uint8_t)__float2int_rz(t);
}

__device__ __forceinline__ uint8_t lut_entry(int op, float param, int e) {
  switch (op) {
    case IMGTX_PW_BRIGHTNESS:
      return pil_blend(0, e, param);
    case IMGTX_PW_SCALE_ABS: {
      float v = fabsf(__fmul_rn((float)e, param));
      return sat_rint8(v);
    }
    case IMGTX_PW_POSTERIZE: {
      int bits = (int)param;
      int mask = ~((1 << (8 - bits)) - 1);
      return (uint8_t)(e & mask);
    }
    case IMGTX_PW_SOLARIZE: {
      int thr = (int)param;
      return (uint8_t)(e < thr ? e : 255 - e);
    }
    default:
      return (uint8_t)e;
  }
}

// ImageOps.equalize LUT of one channel from its 256-bin histogram, computed by ONE WARP:
// step = (sum(h) - h[last non-empty bin]) / 255; identity if <= 1 non-empty bin or step == 0; else
// lut[i] = min((step/2 + sum_{k<i} h[k]) / step, 255).
__device__ __forceinline__ void equalize_lut_warp(const uint32_t* hj, uint8_t* lj, int lane) {
  uint32_t h8[8];
  uint32_t tot = 0, nz = 0;
  int last = -1;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    h8[k] = hj[lane * 8 + k];
    tot += h8[k];
    if (h8[k]) { ++nz; last = lane * 8 + k; }
  }
  uint32_t incl = tot;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const uint32_t excl = incl - tot;
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  uint32_t nzt = nz;
  int lastm = last;
  for (int o = 16; o > 0; o >>= 1) {
    nzt += __shfl_xor_sync(0xffffffffu, nzt, o);
    lastm = max(lastm, __shfl_xor_sync(0xffffffffu, lastm, o));
  }
  const uint32_t step = lastm >= 0 ? (total - hj[lastm]) / 255u : 0u;
  uint32_t run = excl;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int e = lane * 8 + k;
    uint32_t v = e;
    if (nzt > 1 && step != 0) {
      v = (step / 2 + run) / step;
      v = v > 255u ? 255u : v;
    }
    lj[e] = (uint8_t)v;
    run += h8[k];
  }
}

#endif
}  // namespace imgtx
