// Pointwise uint8 -> uint8 maps: every op is a 256-entry LUT evaluated with the library's exact f32 formula.
//   IMGTX_PW_BRIGHTNESS  Pillow ImageEnhance.Brightness  (reference: transformation.py:261-269)
//   IMGTX_PW_SCALE_ABS   cv2.convertScaleAbs(alpha,0)    (reference: transformation.py:203-210)
//   IMGTX_PW_POSTERIZE   ImageOps.posterize              (reference: fall_2025/AugMix.py:31)
//   IMGTX_PW_SOLARIZE    ImageOps.solarize               (reference: fall_2025/AugMix.py:37)
// Roofline: HBM, 1 B read + 1 B written per element.
#include "common.cuh"
#include "lut.cuh"
#include <stdlib.h>

namespace imgtx {

// ---- uniform parameter: one LUT for the whole batch, replicated per shared-memory bank ----------
// lut layout: word e*32 + lane holds entry e for that lane => lane L only ever touches bank L, so the 16
// dependent lookups of a thread never conflict with its warp mates; the byte index of a lookup is
// ((value << 7) | lane*4), i.e. one shift and one 3-input logic op.
constexpr int kPwThreads = 256;    // per-image kernel
constexpr int kPwUniThreads = 512;
constexpr int kPwLutBytes = 256 * 32 * 4;

__device__ __forceinline__ uint32_t lut4(const uint8_t* lut, uint32_t lane4, uint32_t w) {
  const uint32_t i0 = ((w << 7) & 0x7F80u) | lane4;
  const uint32_t i1 = ((w >> 1) & 0x7F80u) | lane4;
  const uint32_t i2 = ((w >> 9) & 0x7F80u) | lane4;
  const uint32_t i3 = ((w >> 17) & 0x7F80u) | lane4;
  const uint32_t v0 = *reinterpret_cast<const uint32_t*>(lut + i0);
  const uint32_t v1 = *reinterpret_cast<const uint32_t*>(lut + i1);
  const uint32_t v2 = *reinterpret_cast<const uint32_t*>(lut + i2);
  const uint32_t v3 = *reinterpret_cast<const uint32_t*>(lut + i3);
  return __byte_perm(__byte_perm(v0, v1, 0x0040), __byte_perm(v2, v3, 0x0040), 0x5410);
}
__device__ __forceinline__ uint4 lut16(const uint8_t* lut, uint32_t lane4, uint4 a) {
  a.x = lut4(lut, lane4, a.x); a.y = lut4(lut, lane4, a.y);
  a.z = lut4(lut, lane4, a.z); a.w = lut4(lut, lane4, a.w);
  return a;
}

__global__ void __launch_bounds__(kPwUniThreads)
pointwise_uniform_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int64_t n_vec,
                         const uint8_t* __restrict__ in_tail, uint8_t* __restrict__ out_tail,
                         int n_tail, int op, float param) {
  extern __shared__ __align__(16) uint8_t lut[];
  {
    uint32_t* l32 = reinterpret_cast<uint32_t*>(lut);
    // thread t fills entries t/32 + 16k for its own lane column: 16 formula evaluations per thread
    for (int i = threadIdx.x; i < 256 * 32; i += kPwUniThreads) l32[i] = lut_entry(op, param, i >> 5);
  }
  __syncthreads();
  const uint32_t lane4 = (threadIdx.x & 31) * 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // four independent 16-byte loads in flight per thread
  for (; i + 3 * stride < n_vec; i += 4 * stride) {
    uint4 a = ldg_stream(in + i);
    uint4 b = ldg_stream(in + i + stride);
    uint4 c = ldg_stream(in + i + 2 * stride);
    uint4 d = ldg_stream(in + i + 3 * stride);
    stg_stream(out + i, lut16(lut, lane4, a));
    stg_stream(out + i + stride, lut16(lut, lane4, b));
    stg_stream(out + i + 2 * stride, lut16(lut, lane4, c));
    stg_stream(out + i + 3 * stride, lut16(lut, lane4, d));
  }
  for (; i < n_vec; i += stride) stg_stream(out + i, lut16(lut, lane4, ldg_stream(in + i)));
  if (blockIdx.x == 0 && (int)threadIdx.x < n_tail) {
    const uint32_t e = in_tail[threadIdx.x];
    out_tail[threadIdx.x] = (uint8_t) * reinterpret_cast<const uint32_t*>(lut + ((e << 7) | lane4));
  }
}

// ---- per-image parameter: one CTA per image, private 256-entry LUT ------------------------------
__global__ void __launch_bounds__(kPwThreads)
pointwise_per_image_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n,
                           int64_t image_bytes, int op, const float* __restrict__ params, int vec_ok) {
  __shared__ uint8_t lut[256];
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    __syncthreads();
    lut[threadIdx.x] = lut_entry(op, params[img], threadIdx.x);
    __syncthreads();
    const uint8_t* src = in + img * image_bytes;
    uint8_t* dst = out + img * image_bytes;
    if (vec_ok) {
      const uint4* s4 = reinterpret_cast<const uint4*>(src);
      uint4* d4 = reinterpret_cast<uint4*>(dst);
      const int nv = (int)(image_bytes >> 4);
      for (int i = threadIdx.x; i < nv; i += blockDim.x) {
        uint4 a = ldg_stream(s4 + i);
        uint32_t* w = reinterpret_cast<uint32_t*>(&a);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t v = w[k];
          w[k] = lut[v & 255u] | (lut[(v >> 8) & 255u] << 8) | (lut[(v >> 16) & 255u] << 16) |
                 (lut[v >> 24] << 24);
        }
        stg_stream(d4 + i, a);
      }
    } else {
      for (int64_t i = threadIdx.x; i < image_bytes; i += blockDim.x) dst[i] = lut[src[i]];
    }
  }
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_pointwise_u8(const uint8_t* in, uint8_t* out, int64_t n, int64_t image_bytes,
                                  int op, float param, const float* param_per_image_dev,
                                  imgtx_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  IMGTX_CHECK_ARG(n >= 0 && image_bytes >= 0, "negative size");
  IMGTX_CHECK_ARG(op >= IMGTX_PW_BRIGHTNESS && op <= IMGTX_PW_IDENTITY, "unknown pointwise op %d", op);
  if (n == 0 || image_bytes == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null image pointer");
  if (param_per_image_dev == nullptr) {
    const int64_t total = n * image_bytes;
    const bool vec = aligned16(in) && aligned16(out);
    const int64_t n_vec = vec ? (total >> 4) : 0;
    const int64_t tail0 = n_vec << 4;
    int64_t rem = total - tail0;
    // the tail handled inside the main kernel is < 16 bytes when vectorised; otherwise fall back to the
    // per-image kernel with a broadcast parameter
    if (!vec) {
      float* tmp = nullptr;
      IMGTX_CUDA_OK(workspace_alloc((void**)&tmp, sizeof(float), stream));
      IMGTX_CUDA_OK(cudaMemcpyAsync(tmp, &param, sizeof(float), cudaMemcpyHostToDevice, stream));
      pointwise_per_image_kernel<<<1, kPwThreads, 0, stream>>>(in, out, 1, total, op, tmp, 0);
      IMGTX_LAUNCH_OK();
      IMGTX_CUDA_OK(cudaFreeAsync(tmp, stream));
      return IMGTX_OK;
    }
    static const int kCtasPerSm = getenv("IMGTX_PW_CTAS") ? atoi(getenv("IMGTX_PW_CTAS")) : 4;
    int64_t want = (n_vec + 4 * kPwUniThreads - 1) / (4 * kPwUniThreads);
    int64_t cap = (int64_t)sm_count() * kCtasPerSm;
    int64_t grid = want < 1 ? 1 : (want > cap ? cap : want);
    IMGTX_CUDA_OK(cudaFuncSetAttribute(pointwise_uniform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kPwLutBytes));
    pointwise_uniform_kernel<<<(unsigned)grid, kPwUniThreads, kPwLutBytes, stream>>>(
        reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), n_vec, in + tail0,
        out + tail0, (int)rem, op, param);
    IMGTX_LAUNCH_OK();
    return IMGTX_OK;
  }
  const int vec_ok = (image_bytes % 16 == 0) && aligned16(in) && aligned16(out);
  int64_t grid = n < (int64_t)sm_count() * 8 ? n : (int64_t)sm_count() * 8;
  pointwise_per_image_kernel<<<(unsigned)grid, kPwThreads, 0, stream>>>(in, out, n, image_bytes, op,
                                                                        param_per_image_dev, vec_ok);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
