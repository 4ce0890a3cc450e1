// Per-image Shannon entropy of the 256-bin value histogram (the TENT / adaptive-transform gate).
//
//   reference: fall_2025/TENT_Pipelines.py:99-121 (normalised tensor: x*std+mean per channel, clamp[0,1]) and
//              fall_2025/Initial_Experiments.py:95-113 ([0,1] tensor as is): np.histogram(bins=256, range=(0,1),
//              density=True) over all c*h*w values, zero bins dropped, scipy.stats.entropy(base=2).
//
// One CTA per image: the image is read from HBM once (12 288 B for 3x32x32), binned into per-warp histograms in shared
// memory (bin = min(int(256*v), 255); values outside [0,1] and NaN are not counted, as np.histogram does), the 256
// counts are integer-exact, and the entropy -sum(p ln p)/ln 2 is summed in f64 in a fixed order (deterministic;
// within a few ulp of SciPy's pairwise sum).  Roofline: HBM reads, c*h*w*4 bytes per image, 8 (+1024) bytes out.
#include "common.cuh"
#include <algorithm>
#include <string.h>

namespace imgtx {

constexpr int kEntThreads = 256;
constexpr int kEntWarps = kEntThreads / 32;

struct EntropyParams {
  float mean[4], std[4];
  int denorm;
};

// mean/std of the channel that element index e (0 <= e < c*hw, c <= 4) belongs to, without an indexed parameter load
struct ChanSel {
  float m[4], s[4];
  int hw;
  __device__ __forceinline__ ChanSel(const EntropyParams& p, int hw_) : hw(hw_) {
#pragma unroll
    for (int k = 0; k < 4; ++k) { m[k] = p.mean[k]; s[k] = p.std[k]; }
  }
  __device__ __forceinline__ void at(int e, float& mn, float& sd) const {
    const int ch = (e >= hw) + (e >= 2 * hw) + (e >= 3 * hw);
    mn = ch == 0 ? m[0] : ch == 1 ? m[1] : ch == 2 ? m[2] : m[3];
    sd = ch == 0 ? s[0] : ch == 1 ? s[1] : ch == 2 ? s[2] : s[3];
  }
};

__device__ __forceinline__ void ent_count(uint32_t* wh, float v, bool denorm, float mn, float sd) {
  if (denorm) {
    v = __fadd_rn(__fmul_rn(v, sd), mn);
    if (v != v) return;                       // torch.clamp keeps NaN, np.histogram drops it
    v = fminf(fmaxf(v, 0.0f), 1.0f);
  } else if (!(v >= 0.0f && v <= 1.0f)) {
    return;
  }
  const int b = min(__float2int_rz(__fmul_rn(v, 256.0f)), 255);
  atomicAdd(wh + b, 1u);
}

__global__ void __launch_bounds__(kEntThreads)
entropy_kernel(const float* __restrict__ x, int64_t n, int c, int hw, EntropyParams p, double* __restrict__ ent,
               uint32_t* __restrict__ counts) {
  __shared__ uint32_t hist[kEntWarps][256];
  __shared__ double red[kEntWarps];
  __shared__ uint32_t redn[kEntWarps];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int chw = c * hw;
  const bool vec = (hw & 3) == 0, dn = p.denorm != 0;
  const ChanSel sel(p, hw);
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    for (int i = tid; i < kEntWarps * 256; i += kEntThreads) (&hist[0][0])[i] = 0u;
    __syncthreads();
    const float* xi = x + img * chw;
    uint32_t* wh = hist[warp];
    if (vec) {
      const float4* x4 = reinterpret_cast<const float4*>(xi);
      for (int i = tid; i < (chw >> 2); i += kEntThreads) {
        const float4 f = ldg_stream_f4(x4 + i);
        float mn, sd;
        sel.at(i * 4, mn, sd);
        ent_count(wh, f.x, dn, mn, sd); ent_count(wh, f.y, dn, mn, sd);
        ent_count(wh, f.z, dn, mn, sd); ent_count(wh, f.w, dn, mn, sd);
      }
    } else {
      for (int i = tid; i < chw; i += kEntThreads) {
        float mn, sd;
        sel.at(i, mn, sd);
        ent_count(wh, xi[i], dn, mn, sd);
      }
    }
    __syncthreads();
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < kEntWarps; ++k) cnt += hist[k][tid];
    if (counts) counts[img * 256 + tid] = cnt;
    uint32_t tot = cnt;
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (lane == 0) redn[warp] = tot;
    __syncthreads();
    tot = 0;
#pragma unroll
    for (int k = 0; k < kEntWarps; ++k) tot += redn[k];
    double term = 0.0;
    if (cnt > 0) {
      const double pk = (double)cnt / (double)tot;
      term = pk * log(pk);
    }
    for (int o = 16; o > 0; o >>= 1) term += __shfl_xor_sync(0xffffffffu, term, o);
    if (lane == 0) red[warp] = term;
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < kEntWarps; ++k) s += red[k];
      ent[img] = tot > 0 ? -s / 0.69314718055994530942 : __longlong_as_double(0x7ff8000000000000ll);
    }
    __syncthreads();
  }
}

// Small images (CIFAR): one WARP per image, eight images in flight per CTA, no block-wide barrier on the path.
__global__ void __launch_bounds__(kEntThreads)
entropy_warp_kernel(const float* __restrict__ x, int64_t n, int c, int hw, EntropyParams p, double* __restrict__ ent,
                    uint32_t* __restrict__ counts) {
  __shared__ uint32_t hist[kEntWarps][256];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int chw = c * hw;
  const bool vec = (hw & 3) == 0, dn = p.denorm != 0;
  const ChanSel sel(p, hw);
  uint32_t* wh = hist[warp];
  const int64_t stride = (int64_t)gridDim.x * kEntWarps;
  for (int64_t img = (int64_t)blockIdx.x * kEntWarps + warp; img < n; img += stride) {
#pragma unroll
    for (int k = 0; k < 8; ++k) wh[k * 32 + lane] = 0u;
    __syncwarp();
    const float* xi = x + img * chw;
    if (vec) {
      const float4* x4 = reinterpret_cast<const float4*>(xi);
      const int n4 = chw >> 2;
      for (int i0 = 0; i0 < n4; i0 += 32 * 8) {
        float4 f[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * 32 + lane;
          f[k] = i < n4 ? ldg_stream_f4(x4 + i) : make_float4(-1.f, -1.f, -1.f, -1.f);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * 32 + lane;
          if (i < n4) {
            float mn, sd;
            sel.at(i * 4, mn, sd);
            ent_count(wh, f[k].x, dn, mn, sd); ent_count(wh, f[k].y, dn, mn, sd);
            ent_count(wh, f[k].z, dn, mn, sd); ent_count(wh, f[k].w, dn, mn, sd);
          }
        }
      }
    } else {
      for (int i = lane; i < chw; i += 32) {
        float mn, sd;
        sel.at(i, mn, sd);
        ent_count(wh, xi[i], dn, mn, sd);
      }
    }
    __syncwarp();
    uint32_t cnt[8], tot = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      cnt[k] = wh[k * 32 + lane];
      tot += cnt[k];
      if (counts) counts[img * 256 + k * 32 + lane] = cnt[k];
    }
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    double term = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (cnt[k] > 0) {
        const double pk = (double)cnt[k] / (double)tot;
        term += pk * log(pk);
      }
    for (int o = 16; o > 0; o >>= 1) term += __shfl_xor_sync(0xffffffffu, term, o);
    if (lane == 0) ent[img] = tot > 0 ? -term / 0.69314718055994530942 : __longlong_as_double(0x7ff8000000000000ll);
    __syncwarp();
  }
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_shannon_entropy_f32(const float* x, int64_t n, int c, int h, int w, const float* mean_host,
                                         const float* std_host, double* entropy_dev, uint32_t* counts_dev,
                                         imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(n >= 0 && c >= 1 && h > 0 && w > 0, "bad tensor shape");
  IMGTX_CHECK_ARG((mean_host != nullptr) == (std_host != nullptr), "mean and std go together");
  IMGTX_CHECK_ARG(!mean_host || c <= 4, "per-channel denormalisation takes at most 4 channels");
  IMGTX_CHECK_ARG((int64_t)c * h * w < (1ll << 31), "image too large");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(x && entropy_dev, "null pointer");
  const int hw = h * w;
  IMGTX_CHECK_ARG((hw & 3) != 0 || aligned16(x), "f32 tensors must be 16-byte aligned");
  EntropyParams p;
  memset(&p, 0, sizeof(p));
  p.denorm = mean_host != nullptr;
  for (int i = 0; i < c && p.denorm; ++i) {
    p.mean[i] = mean_host[i];
    p.std[i] = std_host[i];
  }
  if ((int64_t)c * hw <= 16384) {
    const int64_t grid = std::min<int64_t>((n + kEntWarps - 1) / kEntWarps, (int64_t)sm_count() * 8);
    entropy_warp_kernel<<<(unsigned)grid, kEntThreads, 0, (cudaStream_t)stream>>>(x, n, c, hw, p, entropy_dev,
                                                                                 counts_dev);
    IMGTX_LAUNCH_OK();
    return IMGTX_OK;
  }
  const int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * 8);
  entropy_kernel<<<(unsigned)grid, kEntThreads, 0, (cudaStream_t)stream>>>(x, n, c, hw, p, entropy_dev, counts_dev);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
