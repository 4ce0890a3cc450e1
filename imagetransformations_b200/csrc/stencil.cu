// Stencils on uint8 HWC images held whole in shared memory (staged mode) — each image is read from HBM
// once and written once; all intermediate planes live on chip.
//   imgtx_cv_gaussian_blur_u8   cv2.GaussianBlur(ksize, sigma) for CV_8U: Q8 separable kernel with error
//                               diffusion, REFLECT_101, u16 intermediate saturated at 65535, (v+2^15)>>16.
//                               reference: transformation.py:228-257 (apply_blur)
//   imgtx_pil_gaussian_blur_u8  PIL ImageFilter.GaussianBlur(radius): 3 extended-box passes per axis
//                               (libImaging/BoxBlur.c), u32 fixed point, edge replicate.
//                               reference: pipenline/cifar_image_transformations.py:72-77 (defocus_blur)
//   imgtx_box_filter_h_u8       cv2.filter2D with a size x size kernel whose middle row is 1/size:
//                               horizontal box, REFLECT_101, f32 accumulate, round-half-even.
//                               reference: pipenline/cifar_image_transformations.py:109-119 (motion_blur)
// Roofline: nominally HBM (2 B/element); wide kernels (31 taps x 2 passes) are integer-ALU bound.
#include "pipeline.cuh"
#include <math.h>
#include <string.h>

namespace imgtx {

// ---------------------------------------------------------------------------------------------
struct CvGaussParams {
  int ksize;
  int k[32];
};

struct OpCvGauss : NoSpill {
  using Params = CvGaussParams;
  static constexpr int kThreads = 256;
  static size_t scratch_bytes(const Shape& s, int G) { return align_up((size_t)G * s.in_bytes * 2, 16); }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  static size_t gtmp_bytes(const Shape& s) { return (size_t)s.in_bytes * 2; }
  static size_t scratch_bytes_spilled(const Shape&) { return 16; }
  __device__ static void prepare(const Shape&, const Params&, uint8_t*, bool) {}

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t* gtmp) {
    uint16_t* tmp = reinterpret_cast<uint16_t*>(gtmp ? gtmp : scratch);
    const int tid = threadIdx.x;
    const int c = s.c, w = s.w, h = s.h, n = p.ksize, r = n >> 1;
    const int npix = h * w;
    // horizontal
    for (int i = tid; i < g * npix; i += kThreads) {
      const int j = i / npix, px = i - j * npix;
      const int y = px / w, x = px - y * w;
      const uint8_t* row = src + (size_t)j * s.in_bytes + (size_t)y * w * c;
      uint16_t* t = tmp + ((size_t)j * npix + px) * c;
      if (c == 3) {
        int a0 = 0, a1 = 0, a2 = 0;
        if (x >= r && x + r < w) {
          const uint8_t* sp = row + (x - r) * 3;
#pragma unroll 4
          for (int k = 0; k < n; ++k) {
            const int kk = p.k[k];
            a0 += kk * sp[k * 3]; a1 += kk * sp[k * 3 + 1]; a2 += kk * sp[k * 3 + 2];
          }
        } else {
          for (int k = 0; k < n; ++k) {
            const int kk = p.k[k];
            const uint8_t* sp = row + reflect101(x + k - r, w) * 3;
            a0 += kk * sp[0]; a1 += kk * sp[1]; a2 += kk * sp[2];
          }
        }
        t[0] = (uint16_t)min(a0, 65535); t[1] = (uint16_t)min(a1, 65535); t[2] = (uint16_t)min(a2, 65535);
      } else {
        for (int ch = 0; ch < c; ++ch) {
          int a = 0;
          for (int k = 0; k < n; ++k) a += p.k[k] * row[reflect101(x + k - r, w) * c + ch];
          t[ch] = (uint16_t)min(a, 65535);
        }
      }
    }
    __syncthreads();
    // vertical
    for (int i = tid; i < g * npix; i += kThreads) {
      const int j = i / npix, px = i - j * npix;
      const int y = px / w, x = px - y * w;
      const uint16_t* tj = tmp + (size_t)j * npix * c;
      uint8_t* o = dst + (size_t)j * s.out_bytes + (size_t)px * c;
      if (c == 3) {
        unsigned a0 = 1u << 15, a1 = 1u << 15, a2 = 1u << 15;
        for (int k = 0; k < n; ++k) {
          const unsigned kk = (unsigned)p.k[k];
          const uint16_t* sp = tj + ((size_t)reflect101(y + k - r, h) * w + x) * 3;
          a0 += kk * sp[0]; a1 += kk * sp[1]; a2 += kk * sp[2];
        }
        o[0] = (uint8_t)min(a0 >> 16, 255u); o[1] = (uint8_t)min(a1 >> 16, 255u); o[2] = (uint8_t)min(a2 >> 16, 255u);
      } else {
        for (int ch = 0; ch < c; ++ch) {
          unsigned a = 1u << 15;
          for (int k = 0; k < n; ++k) a += (unsigned)p.k[k] * tj[((size_t)reflect101(y + k - r, h) * w + x) * c + ch];
          o[ch] = (uint8_t)min(a >> 16, 255u);
        }
      }
    }
    __syncthreads();
  }
};

// ---------------------------------------------------------------------------------------------
struct PilGaussParams {
  int ri;
  uint32_t ww, fw;
};

struct OpPilGauss : NoSpill {
  using Params = PilGaussParams;
  static constexpr int kThreads = 256;
  static size_t scratch_bytes(const Shape& s, int G) { return align_up((size_t)G * s.in_bytes, 16); }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  static size_t gtmp_bytes(const Shape& s) { return (size_t)s.in_bytes; }
  static size_t scratch_bytes_spilled(const Shape&) { return 16; }
  __device__ static void prepare(const Shape&, const Params&, uint8_t*, bool) {}

  // One extended-box pass over `len` samples with element stride `st` (running sum, edge replicate).
  __device__ static void line_pass(const uint8_t* in, uint8_t* out, int len, int st, const Params& p) {
    const int ri = p.ri;
    // window sum for x = 0: sum_{i=-ri..ri} in[clamp(i)]
    uint32_t acc = 0;
    for (int i = -ri; i <= ri; ++i) acc += in[clampi(i, 0, len - 1) * st];
    for (int x = 0; x < len; ++x) {
      const uint32_t far = (uint32_t)in[clampi(x - ri - 1, 0, len - 1) * st] + in[clampi(x + ri + 1, 0, len - 1) * st];
      const uint32_t v = acc * p.ww + far * p.fw + (1u << 23);
      out[x * st] = (uint8_t)(v >> 24);
      acc += in[clampi(x + ri + 1, 0, len - 1) * st];
      acc -= in[clampi(x - ri, 0, len - 1) * st];
    }
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t* gtmp) {
    uint8_t* tmp = gtmp ? gtmp : scratch;
    const int tid = threadIdx.x;
    const int c = s.c, w = s.w, h = s.h;
    const int rowb = w * c;
    // 6 passes: src->tmp, tmp->dst, dst->tmp (horizontal); tmp->dst, dst->tmp, tmp->dst (vertical)
    for (int pass = 0; pass < 6; ++pass) {
      const uint8_t* in = pass == 0 ? src : ((pass & 1) ? tmp : dst);
      uint8_t* out = (pass & 1) ? dst : tmp;
      const size_t in_stride = (pass == 0) ? (size_t)s.in_bytes : (size_t)s.in_bytes;
      if (pass < 3) {
        const int lines = h * c;  // (y, ch)
        for (int i = tid; i < g * lines; i += kThreads) {
          const int j = i / lines, l = i - j * lines;
          const int y = l / c, ch = l - y * c;
          line_pass(in + j * in_stride + (size_t)y * rowb + ch, out + (size_t)j * s.in_bytes + (size_t)y * rowb + ch, w, c, p);
        }
      } else {
        const int lines = w * c;  // (x, ch) contiguous bytes of a row
        for (int i = tid; i < g * lines; i += kThreads) {
          const int j = i / lines, l = i - j * lines;
          line_pass(in + j * in_stride + l, out + (size_t)j * s.in_bytes + l, h, rowb, p);
        }
      }
      __syncthreads();
    }
  }
};

// ---------------------------------------------------------------------------------------------
struct BoxHParams {
  int size;
  float kv;
};

struct OpBoxH : NoSpill {
  using Params = BoxHParams;
  static constexpr int kThreads = 256;
  static size_t scratch_bytes(const Shape&, int) { return 16; }
  static size_t scratch_bytes_direct(const Shape&) { return 16; }
  __device__ static void prepare(const Shape&, const Params&, uint8_t*, bool) {}

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t*, uint8_t*) {
    const int c = s.c, w = s.w, r = (p.size - 1) / 2;
    const int rowb = w * c;
    for (int i = threadIdx.x; i < g * s.in_bytes; i += kThreads) {
      const int j = i / s.in_bytes, b = i - j * s.in_bytes;
      const int y = b / rowb, xb = b - y * rowb;
      const int x = xb / c, ch = xb - x * c;
      const uint8_t* row = src + (size_t)j * s.in_bytes + (size_t)y * rowb;
      float acc = 0.0f;
      for (int k = 0; k < p.size; ++k)
        acc = __fadd_rn(acc, __fmul_rn((float)row[reflect101(x + k - r, w) * c + ch], p.kv));
      dst[(size_t)j * s.out_bytes + b] = sat_rint8(acc);
    }
  }
};

}  // namespace imgtx

using namespace imgtx;

// OpenCV getGaussianKernelBitExact for CV_8U (softdouble maths reproduced in IEEE double; the products are
// verified bit-exact against cv2 for sigma in {0.5,...,5} by tests/test_oracle_golden.py via the oracle).
extern "C" int imgtx_cv_gaussian_kernel_host(int ksize, double sigma, int32_t* k_host) {
  IMGTX_CHECK_ARG(ksize >= 1 && ksize <= 31 && (ksize & 1) && sigma > 0 && k_host, "bad Gaussian kernel request");
  double g[32], sum = 0.0;
  const double scale2x = -0.5 / (sigma * sigma);
  for (int i = 0; i < ksize; ++i) {
    const double x = i - (ksize - 1) * 0.5;
    g[i] = exp(scale2x * x * x);
    sum += g[i];
  }
  double err = 0.0;
  int acc = 0;
  for (int i = 0; i < ksize / 2; ++i) {
    const double adj = g[i] / sum * 256 + err;
    const double v = nearbyint(adj);
    err = adj - v;
    k_host[i] = k_host[ksize - 1 - i] = (int)v;
    acc += 2 * (int)v;
  }
  k_host[ksize / 2] = 256 - acc;
  return IMGTX_OK;
}

extern "C" int imgtx_cv_gaussian_blur_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ksize,
                                         double sigma, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4, "bad image shape");
  CvGaussParams p;
  memset(&p, 0, sizeof(p));
  p.ksize = ksize;
  int rc = imgtx_cv_gaussian_kernel_host(ksize, sigma, p.k);
  if (rc) return rc;
  return launch_image_op<OpCvGauss>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream);
}

extern "C" int imgtx_pil_gaussian_blur_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                          float radius, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4, "bad image shape");
  IMGTX_CHECK_ARG(radius > 0, "radius must be positive");
  // libImaging/BoxBlur.c _gaussian_blur_radius (passes = 3), C float / double mix reproduced literally
  volatile float sigma2, L, l, a;
  sigma2 = radius * radius / 3;
  L = sqrt(12.0 * sigma2 + 1.0);
  l = floor((L - 1.0) / 2.0);
  a = (2 * l + 1) * (l * (l + 1) - 3 * sigma2);
  a /= 6 * (sigma2 - (l + 1) * (l + 1));
  volatile float R = l + a;
  PilGaussParams p;
  p.ri = (int)R;
  p.ww = (uint32_t)((uint32_t)(1 << 24) / (R * 2 + 1));
  p.fw = ((1u << 24) - (uint32_t)(p.ri * 2 + 1) * p.ww) / 2;
  return launch_image_op<OpPilGauss>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream);
}

extern "C" int imgtx_box_filter_h_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int size,
                                     imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4, "bad image shape");
  IMGTX_CHECK_ARG(size >= 1 && (size & 1), "box size must be odd");
  BoxHParams p;
  p.size = size;
  p.kv = (float)(1.0 / size);
  return launch_image_op<OpBoxH>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream);
}
