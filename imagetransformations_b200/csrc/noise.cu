// Noise corruptions on uint8 images.
// Parity mode ("injected"): the reference's RNG draws are kernel inputs, results are bit-exact.
// Throughput mode ("philox"): draws come from an on-chip counter RNG (Philox4x32-10 keyed by
// (seed, element index)); validated statistically, never bit-wise.
//   IMGTX_NOISE_GAUSS_F32  transformation.apply_gaussian_noise (transformation.py:272-281):
//                          clip(f32(p) + f32 noise, 0, 255) -> trunc; draws = normal(0, std*255) cast to f32
//   IMGTX_NOISE_GAUSS_F64  TransformationPool.gaussian_noise (cifar_image_transformations.py:39-48): f64 sum
//   IMGTX_NOISE_IMPULSE    TransformationPool.impulse_noise (:50-59): one uniform per PIXEL, shared by channels
//   IMGTX_NOISE_SHOT       TransformationPool.shot_noise (:61-70): k ~ Poisson(f32(p)/255*lam); k/lam*255.0
// Roofline: HBM; parity mode additionally reads the draws (4-8 B per element).
#include "common.cuh"
#include "rng.cuh"

namespace imgtx {

__device__ __forceinline__ uint8_t clip_trunc_f64(double v) {
  v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);
  return (uint8_t)(int)v;
}
__device__ __forceinline__ uint8_t clip_trunc_f32(float v) {
  v = v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v);
  return (uint8_t)__float2int_rz(v);
}

__global__ void __launch_bounds__(256)
noise_injected_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t total, int c, int kind,
                      const void* __restrict__ draws, double param) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int pv = in[i];
    uint8_t o;
    switch (kind) {
      case IMGTX_NOISE_GAUSS_F32:
        o = clip_trunc_f32(__fadd_rn((float)pv, reinterpret_cast<const float*>(draws)[i]));
        break;
      case IMGTX_NOISE_GAUSS_F64:
        o = clip_trunc_f64(__dadd_rn((double)(float)pv, reinterpret_cast<const double*>(draws)[i]));
        break;
      case IMGTX_NOISE_IMPULSE: {
        const double m = reinterpret_cast<const double*>(draws)[i / c];
        o = (uint8_t)pv;
        if (m < param / 2) o = 0;
        if (m > 1 - param / 2) o = 255;
        break;
      }
      default: {  // IMGTX_NOISE_SHOT
        const long long k = reinterpret_cast<const long long*>(draws)[i];
        o = clip_trunc_f64(__dmul_rn(__ddiv_rn((double)k, param), 255.0));
        break;
      }
    }
    out[i] = o;
  }
}

// Throughput mode: one thread per 4 consecutive bytes, one Philox call per thread.
__global__ void __launch_bounds__(256)
noise_philox_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int64_t n_words, int c, int kind,
                    double param, uint64_t seed) {
  const Philox rng(seed);
  const float fparam = (float)param;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t w = in[i];
    uint32_t res = 0;
    if (kind == IMGTX_NOISE_GAUSS_F32 || kind == IMGTX_NOISE_GAUSS_F64) {
      const uint4 r = rng((uint32_t)i, (uint32_t)(i >> 32), 0u, 0u);
      float z[4];
      box_muller(r.x, r.y, z[0], z[1]);
      box_muller(r.z, r.w, z[2], z[3]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float v = (float)((w >> (8 * k)) & 255u) + z[k] * fparam;  // param = std*255
        res |= (uint32_t)clip_trunc_f32(v) << (8 * k);
      }
    } else if (kind == IMGTX_NOISE_IMPULSE) {
      // the mask is per pixel: key the draw by the pixel index so channels share it
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t px = (i * 4 + k) / c;
        const uint4 r = rng((uint32_t)px, (uint32_t)(px >> 32), 1u, 0u);
        const float m = u01(r.x);
        uint32_t v = (w >> (8 * k)) & 255u;
        if (m < fparam * 0.5f) v = 0;
        if (m > 1.0f - fparam * 0.5f) v = 255;
        res |= v << (8 * k);
      }
    } else {  // shot
      uint32_t ctr = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float lam = (float)((w >> (8 * k)) & 255u) / 255.0f * fparam;
        const int kk = poisson_sample(rng, (uint32_t)i, (uint32_t)(i >> 32), (uint32_t)k, ctr, lam);
        const float v = (float)kk / fparam * 255.0f;
        res |= (uint32_t)clip_trunc_f32(v) << (8 * k);
      }
    }
    out[i] = res;
  }
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_noise_injected_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                                       const void* draws_dev, double param, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(kind >= IMGTX_NOISE_GAUSS_F32 && kind <= IMGTX_NOISE_SHOT, "unknown noise kind %d", kind);
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && n >= 0, "bad shape");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out && draws_dev, "null pointer");
  const int64_t total = n * h * w * c;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  noise_injected_kernel<<<(unsigned)(blocks > cap ? cap : blocks), 256, 0, (cudaStream_t)stream>>>(in, out, total, c, kind,
                                                                                                 draws_dev, param);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

extern "C" int imgtx_noise_philox_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                                     double param, uint64_t seed, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(kind >= IMGTX_NOISE_GAUSS_F32 && kind <= IMGTX_NOISE_SHOT, "unknown noise kind %d", kind);
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && n >= 0, "bad shape");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null pointer");
  const int64_t total = n * h * w * c;
  IMGTX_CHECK_ARG(total % 4 == 0 && aligned4(in) && aligned4(out), "philox mode needs 4-byte aligned batches");
  const int64_t words = total / 4;
  int64_t blocks = (words + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  noise_philox_kernel<<<(unsigned)(blocks > cap ? cap : blocks), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint32_t*>(in), reinterpret_cast<uint32_t*>(out), words, c, kind, param, seed);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
