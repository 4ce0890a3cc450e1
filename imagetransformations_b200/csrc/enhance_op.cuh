// OpEnhance: the per-image "statistics then map" operator of enhance.cu, in a header so that fused chains (the
// TransformationPool chain of stencil.cu) can run it on an image that is already in shared memory.
#pragma once
#include "pipeline.cuh"
#include "lut.cuh"

namespace imgtx {
#ifdef __CUDACC__

struct EnhanceParams {
  int kind;
  float factor;
  const float* factor_per_image;  // device, nullable
};

__device__ __forceinline__ int pil_luma(int r, int g, int b) {
  return (19595 * r + 38470 * g + 7471 * b + 0x8000) >> 16;
}
__device__ __forceinline__ int d14(int x) { return (x + 8192) >> 14; }

struct OpEnhance : NoSpill {
  using Params = EnhanceParams;
  static constexpr int kThreads = 256;
  // per image: 768 B LUT + 3 KB histogram + 16 B sums
  static constexpr int kPerImage = 768 + 3 * 256 * 4 + 16;
  static size_t scratch_bytes(const Shape&, int G) { return (size_t)G * kPerImage + 16; }
  static size_t scratch_bytes_direct(const Shape&) { return kPerImage + 16; }
  __device__ static void prepare(const Shape&, const Params&, uint8_t*, bool) {}

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s,
                             const Params& p, uint8_t* scratch, uint8_t*) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int npix = s.h * s.w;
    const int c = s.c;
    unsigned long long* sums = reinterpret_cast<unsigned long long*>(scratch);            // [g] (2 per image)
    uint32_t* hist = reinterpret_cast<uint32_t*>(scratch + align_up((size_t)g * 16, 16));  // [g][3][256]
    uint8_t* lut = reinterpret_cast<uint8_t*>(hist + (size_t)g * 768);                     // [g][3][256]
    const int kind = p.kind;

    // ---------------- phase 1: per-image statistics ----------------
    if (kind == IMGTX_ENH_CONTRAST) {
      for (int j = tid; j < g; j += kThreads) sums[2 * j] = 0ull;
      __syncthreads();
      for (int j = 0; j < g; ++j) {
        const uint8_t* im = src + (size_t)j * s.in_bytes;
        unsigned int acc = 0;
        for (int px = tid; px < npix; px += kThreads) acc += pil_luma(im[px * 3], im[px * 3 + 1], im[px * 3 + 2]);
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) atomicAdd(&sums[2 * j], (unsigned long long)acc);
      }
      __syncthreads();
    } else if (kind == IMGTX_ENH_EQUALIZE || kind == IMGTX_ENH_HISTEQ_YUV) {
      for (int i = tid; i < g * 768; i += kThreads) hist[i] = 0u;
      __syncthreads();
      for (int j = 0; j < g; ++j) {
        const uint8_t* im = src + (size_t)j * s.in_bytes;
        uint32_t* hj = hist + j * 768;
        if (kind == IMGTX_ENH_EQUALIZE) {
          for (int i = tid; i < npix * c; i += kThreads) atomicAdd(&hj[(i % c) * 256 + im[i]], 1u);
        } else {
          for (int px = tid; px < npix; px += kThreads) {
            const int r = im[px * 3], gg = im[px * 3 + 1], b = im[px * 3 + 2];
            int Y = d14(4899 * r + 9617 * gg + 1868 * b);
            Y = Y > 255 ? 255 : Y;
            atomicAdd(&hj[Y], 1u);
          }
        }
      }
      __syncthreads();
    }

    // ---------------- phase 2: LUTs ----------------
    if (kind == IMGTX_ENH_BRIGHTNESS || kind == IMGTX_ENH_CONTRAST) {
      for (int i = tid; i < g * 256; i += kThreads) {
        const int j = i >> 8, e = i & 255;
        const float f = p.factor_per_image ? p.factor_per_image[img0 + j] : p.factor;
        int d = 0;
        if (kind == IMGTX_ENH_CONTRAST) d = (int)((double)sums[2 * j] / (double)npix + 0.5);
        lut[j * 768 + e] = pil_blend(d, e, f);
      }
      __syncthreads();
    } else if (kind == IMGTX_ENH_EQUALIZE) {
      // one warp per (image, channel): ImageOps.equalize LUT via a warp prefix sum over 256 bins
      for (int job = warp; job < g * c; job += kThreads >> 5) {
        const int j = job / c, ch = job - j * c;
        equalize_lut_warp(hist + j * 768 + ch * 256, lut + j * 768 + ch * 256, lane);
      }
      __syncthreads();
    } else if (kind == IMGTX_ENH_HISTEQ_YUV) {
      // cv2.equalizeHist LUT on Y, one warp per image
      for (int j = warp; j < g; j += kThreads >> 5) {
        const uint32_t* hj = hist + j * 768;
        uint32_t h8[8];
        uint32_t tot = 0;
        int first = 256;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          h8[k] = hj[lane * 8 + k];
          tot += h8[k];
          if (h8[k] && first == 256) first = lane * 8 + k;
        }
        uint32_t incl = tot;
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        uint32_t run = incl - tot;  // sum of bins below lane*8
        for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
        const uint32_t hfirst = hj[first];
        uint8_t* lj = lut + j * 768;
        if (hfirst == (uint32_t)npix) {
#pragma unroll
          for (int k = 0; k < 8; ++k) lj[lane * 8 + k] = (uint8_t)first;  // constant image
        } else {
          const float scale = __fdiv_rn(255.0f, (float)((uint32_t)npix - hfirst));
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int e = lane * 8 + k;
            run += h8[k];  // inclusive sum through bin e
            uint8_t v = 0;
            if (e > first) v = sat_rint8(__fmul_rn((float)(run - hfirst), scale));
            lj[e] = v;
          }
        }
      }
      __syncthreads();
    }

    // ---------------- phase 3: map ----------------
    for (int j = 0; j < g; ++j) {
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      const uint8_t* lj = lut + j * 768;
      const float f = p.factor_per_image ? p.factor_per_image[img0 + j] : p.factor;
      switch (kind) {
        case IMGTX_ENH_BRIGHTNESS:
        case IMGTX_ENH_CONTRAST:
          if (STAGED) {
            const uint32_t* i32 = reinterpret_cast<const uint32_t*>(im);
            uint32_t* o32 = reinterpret_cast<uint32_t*>(om);
            for (int i = tid; i < (s.in_bytes >> 2); i += kThreads) {
              const uint32_t v = i32[i];
              o32[i] = lj[v & 255u] | (lj[(v >> 8) & 255u] << 8) | (lj[(v >> 16) & 255u] << 16) | (lj[v >> 24] << 24);
            }
          } else {
            for (int i = tid; i < s.in_bytes; i += kThreads) om[i] = lj[im[i]];
          }
          break;
        case IMGTX_ENH_EQUALIZE:
          for (int i = tid; i < s.in_bytes; i += kThreads) om[i] = lj[(i % c) * 256 + im[i]];
          break;
        case IMGTX_ENH_COLOR:
          for (int px = tid; px < npix; px += kThreads) {
            const int r = im[px * 3], gg = im[px * 3 + 1], b = im[px * 3 + 2];
            const int L = pil_luma(r, gg, b);
            om[px * 3] = pil_blend(L, r, f);
            om[px * 3 + 1] = pil_blend(L, gg, f);
            om[px * 3 + 2] = pil_blend(L, b, f);
          }
          break;
        case IMGTX_ENH_SHARPNESS: {
          const float k1 = __fdiv_rn(1.0f, 13.0f), k5 = __fdiv_rn(5.0f, 13.0f);
          const int rowb = s.w * c;
          for (int i = tid; i < s.in_bytes; i += kThreads) {
            const int px = i / c;
            const int y = px / s.w, x = px - y * s.w;
            const int pv = im[i];
            int d = pv;
            if (x > 0 && x < s.w - 1 && y > 0 && y < s.h - 1) {
              // libImaging ImagingFilter3x3: ss = 0.5 + MAC3(row y+1) + MAC3(row y) + MAC3(row y-1)
              const uint8_t* r2 = im + i + rowb;
              const uint8_t* r1 = im + i;
              const uint8_t* r0 = im + i - rowb;
              float ss = 0.5f;
              ss = __fadd_rn(ss, __fadd_rn(__fadd_rn(__fmul_rn((float)r2[-c], k1), __fmul_rn((float)r2[0], k1)), __fmul_rn((float)r2[c], k1)));
              ss = __fadd_rn(ss, __fadd_rn(__fadd_rn(__fmul_rn((float)r1[-c], k1), __fmul_rn((float)r1[0], k5)), __fmul_rn((float)r1[c], k1)));
              ss = __fadd_rn(ss, __fadd_rn(__fadd_rn(__fmul_rn((float)r0[-c], k1), __fmul_rn((float)r0[0], k1)), __fmul_rn((float)r0[c], k1)));
              d = clip8((int)floorf(ss));
            }
            om[i] = pil_blend(d, pv, f);
          }
          break;
        }
        case IMGTX_ENH_HISTEQ_YUV:
          for (int px = tid; px < npix; px += kThreads) {
            const int r = im[px * 3], gg = im[px * 3 + 1], b = im[px * 3 + 2];
            int Y = d14(4899 * r + 9617 * gg + 1868 * b);
            const int U = clip8(d14((b - Y) * 8061 + (128 << 14)));
            const int V = clip8(d14((r - Y) * 14369 + (128 << 14)));
            Y = Y > 255 ? 255 : Y;
            const int ye = lj[Y];
            const int u = U - 128, v = V - 128;
            om[px * 3] = clip8(ye + d14(v * 18678));
            om[px * 3 + 1] = clip8(ye + d14(v * (-9519) + u * (-6472)));
            om[px * 3 + 2] = clip8(ye + d14(u * 33292));
          }
          break;
      }
    }
  }
};

#endif  // __CUDACC__
}  // namespace imgtx
