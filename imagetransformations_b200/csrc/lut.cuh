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

#endif
}  // namespace imgtx
