// CIFAR-10-C-style corruptions, part 3: elastic_transform and jpeg_compression (specification: oracle/cifar_c.py,
// both pinned bit-for-bit against the libraries they restate: cv2.warpAffine + scipy.ndimage gaussian_filter /
// map_coordinates, and Pillow's libjpeg-turbo round trip).
//
//   imgtx_cc_elastic_u8  random affine (cv2.warpAffine: inverse map in f64, 1/1024 fixed-point coordinates snapped to 1/32
//                        sub-pixels, f32 table weights, REFLECT_101) -> Gaussian-smoothed random displacement field
//                        (scipy symmetric-pair f64 order, half-sample 'reflect') -> bilinear map_coordinates (f64 weights,
//                        'reflect').
//   imgtx_cc_jpeg_u8     baseline JPEG round trip without the (lossless) entropy coder: RGB->YCbCr, 4:2:0 h2v2 downsample,
//                        islow forward DCT, quantise / dequantise, islow inverse DCT, h2v2 fancy upsample, YCbCr->RGB —
//                        all in libjpeg's integer arithmetic.  8 threads per 8x8 block (one per row / column).
// Roofline: nominally HBM (2 B/element); both are integer/f64-ALU bound.
#include "pipeline.cuh"
#include "rng.cuh"
#include "cifar_c_team2.cuh"
#include <math.h>
#include <string.h>
#include <algorithm>
#include <vector>

namespace imgtx {

// ------------------------------------------------------------------------------------------------
// elastic_transform
// ------------------------------------------------------------------------------------------------
constexpr int kMaxLw = 20;
struct ElasticParams {
  double alpha;
  double w[kMaxLw + 1];
  int lw;                 // -1: sigma == 0 (no smoothing)
  const double* M;        // device (n,6) forward affine matrices, or nullptr (jitter drawn on chip, amplitude c2)
  const double* ux;       // device (n,h,w) uniform(-1,1) fields or nullptr
  const double* uy;
  double c2;
  uint64_t seed;
  int force_exact;        // IMGTX_CC_REPLAY_ALL
  const double* wdev;     // lw > kMaxLw (large images): the lw + 1 Gaussian weights in device memory instead of w[]
};

__device__ __forceinline__ int reflect_half(int i, int n) {  // scipy 'reflect': d c b a | a b c d | d c b a
  const int period = 2 * n;
  int m = i % period;
  if (m < 0) m += period;
  return m >= n ? period - 1 - m : m;
}

struct OpCcElastic {
  using Params = ElasticParams;
  static constexpr int kThreads = 256;
  // large images (224x224: 1.8 MB of planes): everything but the byte table lives in a global workspace
  static size_t gtmp_bytes(const Shape& s) { return (size_t)s.h * s.w * (12 + 16 + 8) + 64; }
  static size_t scratch_bytes_spilled(const Shape&) { return 1024 + 16; }
  // T32 (1 KB) + warped image f32 (hw*3*4) + two f64 field planes (hw*8 each) + dx, dy f32 (hw*4 each) + M (64 B)
  static size_t scratch_bytes(const Shape& s, int) { return 1024 + (size_t)s.h * s.w * (12 + 16 + 8) + 64 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params&, uint8_t* scratch, bool) {
    float* T = reinterpret_cast<float*>(scratch);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = __fdiv_rn((float)i, 255.0f);
    __syncthreads();
  }

  // gaussian_filter(field, sigma, mode='reflect') * alpha -> f32; F0 holds the field, F1 is scratch
  __device__ static void smooth(double* F0, double* F1, float* out, const Shape& s, const Params& p) {
    const int hw = s.h * s.w, tid = threadIdx.x;
    if (p.lw < 0) {
      for (int i = tid; i < hw; i += kThreads) out[i] = (float)(F0[i] * p.alpha);
      __syncthreads();
      return;
    }
    const double* wv = p.wdev ? p.wdev : p.w;
    for (int i = tid; i < hw; i += kThreads) {  // axis y
      const int y = i / s.w, x = i - y * s.w;
      double acc = F0[i] * wv[0];
      for (int k = p.lw; k >= 1; --k)
        acc = acc + (F0[reflect_half(y - k, s.h) * s.w + x] + F0[reflect_half(y + k, s.h) * s.w + x]) * wv[k];
      F1[i] = acc;
    }
    __syncthreads();
    for (int i = tid; i < hw; i += kThreads) {  // axis x
      const int y = i / s.w, x = i - y * s.w;
      const double* row = F1 + y * s.w;
      double acc = row[x] * wv[0];
      for (int k = p.lw; k >= 1; --k) acc = acc + (row[reflect_half(x - k, s.w)] + row[reflect_half(x + k, s.w)]) * wv[k];
      out[i] = (float)(acc * p.alpha);
    }
    __syncthreads();
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t* gtmp) {
    const int tid = threadIdx.x;
    const int hw = s.h * s.w;
    const float* T = reinterpret_cast<const float*>(scratch);
    float* W = reinterpret_cast<float*>(gtmp ? gtmp : scratch + 1024);        // warped image, HWC f32
    double* F0 = reinterpret_cast<double*>(W + hw * 3);
    double* F1 = F0 + hw;
    float* DX = reinterpret_cast<float*>(F1 + hw);
    float* DY = DX + hw;
    double* Ms = reinterpret_cast<double*>(DY + hw);            // inverse matrix (6)
    const Philox rng(p.seed);
    for (int j = 0; j < g; ++j) {
      const int64_t img = img0 + j;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      __syncthreads();
      if (tid == 0) {
        double M[6];
        if (p.M) {
          for (int k = 0; k < 6; ++k) M[k] = p.M[img * 6 + k];
        } else {
          // pts1 = [(c+q, c+q), (c+q, c-q), (c-q, c-q)], pts2 = pts1 + U(-c2, c2): closed-form affine through 3 points
          const float cs = (float)(s.h / 2), q = (float)(s.h / 3);
          const float p1[3][2] = {{cs + q, cs + q}, {cs + q, cs - q}, {cs - q, cs - q}};
          double p2[3][2];
          const uint4 r0 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x80u, 0u);
          const uint4 r1 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x80u, 1u);
          const uint32_t rr[6] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y};
          for (int k = 0; k < 6; ++k)
            p2[k / 2][k % 2] = (double)(p1[k / 2][k % 2] + (float)(-p.c2 + 2.0 * p.c2 * (double)u01(rr[k])));
          const double x0 = p1[0][0], y0 = p1[0][1], x1 = p1[1][0], y1 = p1[1][1], x2 = p1[2][0], y2 = p1[2][1];
          const double det = x0 * (y1 - y2) - y0 * (x1 - x2) + (x1 * y2 - x2 * y1);
          for (int rrow = 0; rrow < 2; ++rrow) {
            const double u0 = p2[0][rrow], u1 = p2[1][rrow], u2 = p2[2][rrow];
            M[rrow * 3 + 0] = (u0 * (y1 - y2) - y0 * (u1 - u2) + (u1 * y2 - u2 * y1)) / det;
            M[rrow * 3 + 1] = (x0 * (u1 - u2) - u0 * (x1 - x2) + (x1 * u2 - x2 * u1)) / det;
            M[rrow * 3 + 2] = (x0 * (y1 * u2 - y2 * u1) - y0 * (x1 * u2 - x2 * u1) + u0 * (x1 * y2 - x2 * y1)) / det;
          }
        }
        // cv2.warpAffine: invert (no WARP_INVERSE_MAP)
        double D = M[0] * M[4] - M[1] * M[3];
        D = D != 0.0 ? 1.0 / D : 0.0;
        const double A11 = M[4] * D, A22 = M[0] * D;
        M[0] = A11; M[1] *= -D; M[3] *= -D; M[4] = A22;
        const double b1 = -M[0] * M[2] - M[1] * M[5];
        const double b2 = -M[3] * M[2] - M[4] * M[5];
        M[2] = b1; M[5] = b2;
        for (int k = 0; k < 6; ++k) Ms[k] = M[k];
      }
      // displacement-field draws
      for (int i = tid; i < hw; i += kThreads) {
        if (p.ux) {
          F0[i] = p.ux[img * hw + i];
        } else {
          const uint64_t e = (uint64_t)(img * hw + i);
          const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x81u, 0u);
          F0[i] = -1.0 + 2.0 * (double)u01(r.x);
        }
      }
      __syncthreads();
      smooth(F0, F1, DX, s, p);
      for (int i = tid; i < hw; i += kThreads) {
        if (p.uy) {
          F0[i] = p.uy[img * hw + i];
        } else {
          const uint64_t e = (uint64_t)(img * hw + i);
          const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x82u, 0u);
          F0[i] = -1.0 + 2.0 * (double)u01(r.x);
        }
      }
      __syncthreads();
      smooth(F0, F1, DY, s, p);
      // affine warp into W
      for (int px = tid; px < hw; px += kThreads) {
        const int y = px / s.w, x = px - y * s.w;
        const int adelta = (int)rint(Ms[0] * (double)x * 1024), bdelta = (int)rint(Ms[3] * (double)x * 1024);
        const int X0 = (int)rint((Ms[1] * (double)y + Ms[2]) * 1024) + 16;
        const int Y0 = (int)rint((Ms[4] * (double)y + Ms[5]) * 1024) + 16;
        const int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
        const int sx = X >> 5, sy = Y >> 5;
        const float fx = __fmul_rn((float)(X & 31), 0.03125f), fy = __fmul_rn((float)(Y & 31), 0.03125f);
        const float wx0 = __fsub_rn(1.0f, fx), wy0 = __fsub_rn(1.0f, fy);
        const float w00 = __fmul_rn(wy0, wx0), w01 = __fmul_rn(wy0, fx), w10 = __fmul_rn(fy, wx0), w11 = __fmul_rn(fy, fx);
        const int ya = reflect101(sy, s.h), yb = reflect101(sy + 1, s.h), xa = reflect101(sx, s.w), xb = reflect101(sx + 1, s.w);
        const uint8_t* q00 = im + (ya * s.w + xa) * 3;
        const uint8_t* q01 = im + (ya * s.w + xb) * 3;
        const uint8_t* q10 = im + (yb * s.w + xa) * 3;
        const uint8_t* q11 = im + (yb * s.w + xb) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float acc = __fmul_rn(T[q00[c]], w00);
          acc = __fadd_rn(acc, __fmul_rn(T[q01[c]], w01));
          acc = __fadd_rn(acc, __fmul_rn(T[q10[c]], w10));
          acc = __fadd_rn(acc, __fmul_rn(T[q11[c]], w11));
          W[px * 3 + c] = acc;
        }
      }
      __syncthreads();
      // bilinear resampling at (y + dy, x + dx)
      for (int px = tid; px < hw; px += kThreads) {
        const int y = px / s.w, x = px - y * s.w;
        const double yy = (double)__fadd_rn((float)y, DY[px]), xx = (double)__fadd_rn((float)x, DX[px]);
        const double fy0 = floor(yy), fx0 = floor(xx);
        const double ty = yy - fy0, tx = xx - fx0;
        const int y0 = (int)fy0, x0 = (int)fx0;
        const int ya = reflect_half(y0, s.h), yb = reflect_half(y0 + 1, s.h), xa = reflect_half(x0, s.w), xb = reflect_half(x0 + 1, s.w);
        const float* q00 = W + (ya * s.w + xa) * 3;
        const float* q01 = W + (ya * s.w + xb) * 3;
        const float* q10 = W + (yb * s.w + xa) * 3;
        const float* q11 = W + (yb * s.w + xb) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          double acc = ((double)q00[c] * (1 - ty)) * (1 - tx);
          acc = acc + ((double)q01[c] * (1 - ty)) * tx;
          acc = acc + ((double)q10[c] * ty) * (1 - tx);
          acc = acc + ((double)q11[c] * ty) * tx;
          float v = (float)acc;
          v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
          om[px * 3 + c] = (uint8_t)__float2int_rz(__fmul_rn(v, 255.0f));
        }
      }
    }
  }
};

// Throughput form.  The displacement field only reaches the image through its float32 cast, so the smoothing may be
// evaluated in any float64 order: each thread produces a strip of 8 outputs of one line from a sliding register window
// (one shared-memory load per tap instead of eight) with fused multiply-adds, and the literal scipy order is replayed only
// if the result lies within 4e-14*max(1,|alpha|) of a float32 rounding boundary (probability ~1e-5 per value).  The
// intermediate plane is stored transposed with an odd row pitch so that both passes walk shared memory conflict-free.
// cv2.warpAffine's per-row / per-column fixed-point terms are tabulated once per image (as OpenCV does), the warped image
// stays float32, and the bilinear stage runs in float32 with a literal-f64 replay for the pixels whose value lies within
// 1e-3 of an integer after scaling by 255 (the float32 evaluation is accurate to 1e-4 there; about 0.2 % of the values).
struct ElasticCommon {
  __device__ static __forceinline__ int refl_half(int i, int n) {  // one fold is enough when lw <= n
    return i < 0 ? -1 - i : (i >= n ? 2 * n - 1 - i : i);
  }
  // in range / one fold: branch-free; further out (possible only with injected matrices or fields): the general routine
  __device__ static __forceinline__ int refl_half_any(int i, int n) {
    if (i < -n || i >= 2 * n) return reflect_half(i, n);
    return i < 0 ? -1 - i : (i >= n ? 2 * n - 1 - i : i);
  }
  __device__ static __forceinline__ int refl101_any(int i, int n) {
    if (i <= -n || i >= 2 * n - 1) return reflect101(i, n);
    return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
  }
  // inverse affine matrix of image `img` (cv2.getAffineTransform of the jittered triangle, then cv2.warpAffine's inversion)
  __device__ static void inverse_matrix(const ElasticParams& p, const Philox& rng, int64_t img, int h, double* Mo) {
    double M[6];
    if (p.M) {
      for (int k = 0; k < 6; ++k) M[k] = p.M[img * 6 + k];
    } else {
      // pts1 = [(c+q, c+q), (c+q, c-q), (c-q, c-q)], pts2 = pts1 + U(-c2, c2): closed-form affine through 3 points
      const float cs = (float)(h / 2), q = (float)(h / 3);
      const float p1[3][2] = {{cs + q, cs + q}, {cs + q, cs - q}, {cs - q, cs - q}};
      double p2[3][2];
      const uint4 r0 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x80u, 0u);
      const uint4 r1 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x80u, 1u);
      const uint32_t rr[6] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y};
      for (int k = 0; k < 6; ++k)
        p2[k / 2][k % 2] = (double)(p1[k / 2][k % 2] + (float)(-p.c2 + 2.0 * p.c2 * (double)u01(rr[k])));
      const double x0 = p1[0][0], y0 = p1[0][1], x1 = p1[1][0], y1 = p1[1][1], x2 = p1[2][0], y2 = p1[2][1];
      const double det = x0 * (y1 - y2) - y0 * (x1 - x2) + (x1 * y2 - x2 * y1);
      for (int rrow = 0; rrow < 2; ++rrow) {
        const double u0 = p2[0][rrow], u1 = p2[1][rrow], u2 = p2[2][rrow];
        M[rrow * 3 + 0] = (u0 * (y1 - y2) - y0 * (u1 - u2) + (u1 * y2 - u2 * y1)) / det;
        M[rrow * 3 + 1] = (x0 * (u1 - u2) - u0 * (x1 - x2) + (x1 * u2 - x2 * u1)) / det;
        M[rrow * 3 + 2] = (x0 * (y1 * u2 - y2 * u1) - y0 * (x1 * u2 - x2 * u1) + u0 * (x1 * y2 - x2 * y1)) / det;
      }
    }
    double D = M[0] * M[4] - M[1] * M[3];
    D = D != 0.0 ? 1.0 / D : 0.0;
    const double A11 = M[4] * D, A22 = M[0] * D;
    M[0] = A11; M[1] *= -D; M[3] *= -D; M[4] = A22;
    const double b1 = -M[0] * M[2] - M[1] * M[5];
    const double b2 = -M[3] * M[2] - M[4] * M[5];
    M[2] = b1; M[5] = b2;
    for (int k = 0; k < 6; ++k) Mo[k] = M[k];
  }
  // cv2.warpAffine's tables for one image: tab[2x], tab[2x+1] = adelta, bdelta; tab[2(w+y)], tab[2(w+y)+1] = X0, Y0
  __device__ static __forceinline__ void affine_tables(const double* Mi, int h, int w, int* tab, int i) {
    if (i < w) {
      tab[2 * i] = (int)rint(Mi[0] * (double)i * 1024);
      tab[2 * i + 1] = (int)rint(Mi[3] * (double)i * 1024);
    } else if (i < w + h) {
      const int y = i - w;
      tab[2 * i] = (int)rint((Mi[1] * (double)y + Mi[2]) * 1024) + 16;
      tab[2 * i + 1] = (int)rint((Mi[4] * (double)y + Mi[5]) * 1024) + 16;
    }
  }
  // one pixel of cv2.warpAffine(INTER_LINEAR, BORDER_REFLECT_101) of the image / 255: three float32 channel values
  __device__ static __forceinline__ void warp_pixel(const uint8_t* im, const float* T, const int* tab, int h, int w, int y, int x,
                                                    float& o0, float& o1, float& o2) {
    const int X = (tab[2 * (w + y)] + tab[2 * x]) >> 5, Y = (tab[2 * (w + y) + 1] + tab[2 * x + 1]) >> 5;
    const int sx = X >> 5, sy = Y >> 5;
    const float fx = __fmul_rn((float)(X & 31), 0.03125f), fy = __fmul_rn((float)(Y & 31), 0.03125f);
    const float wx0 = __fsub_rn(1.0f, fx), wy0 = __fsub_rn(1.0f, fy);
    const float w00 = __fmul_rn(wy0, wx0), w01 = __fmul_rn(wy0, fx), w10 = __fmul_rn(fy, wx0), w11 = __fmul_rn(fy, fx);
    const int ya = refl101_any(sy, h), yb = refl101_any(sy + 1, h), xa = refl101_any(sx, w), xb = refl101_any(sx + 1, w);
    const uint32_t c00 = load_rgb(im, (ya * w + xa) * 3), c01 = load_rgb(im, (ya * w + xb) * 3);
    const uint32_t c10 = load_rgb(im, (yb * w + xa) * 3), c11 = load_rgb(im, (yb * w + xb) * 3);
    float o[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float acc = __fmul_rn(T[(c00 >> (8 * c)) & 255u], w00);
      acc = __fadd_rn(acc, __fmul_rn(T[(c01 >> (8 * c)) & 255u], w01));
      acc = __fadd_rn(acc, __fmul_rn(T[(c10 >> (8 * c)) & 255u], w10));
      acc = __fadd_rn(acc, __fmul_rn(T[(c11 >> (8 * c)) & 255u], w11));
      o[c] = acc;
    }
    o0 = o[0]; o1 = o[1]; o2 = o[2];
  }
  __device__ static __forceinline__ uint32_t q255_f32(float v) {   // np.uint8(np.clip(v, 0, 1) * f32(255))
    v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    return (uint32_t)__float2int_rz(__fmul_rn(v, 255.0f));
  }
};

// alpha == 0 (severity 1 of the CIFAR table): the displacement is (float)(field * 0) = +-0 everywhere, so the bilinear stage
// samples at integer coordinates with ty = tx = 0 and returns the warped image itself — (q00*1)*1 + (q01*1)*0 + (q10*0)*1 +
// (q11*0)*0 == q00 for finite fields.  Only the affine warp is left; whole groups of images are processed per barrier.
struct OpCcAffineWarp : NoSpill, ElasticCommon {
  using Params = ElasticParams;
  static constexpr bool kCifarCopy = true;
  static constexpr int kThreads = 256;
  // T32 | per image: inverse matrix (6 f64) + tables (w + h) x 2 int
  static size_t scratch_bytes(const Shape& s, int G) { return 1024 + (size_t)G * (48 + (s.h + s.w) * 8) + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params&, uint8_t* scratch, bool) {
    float* T = reinterpret_cast<float*>(scratch);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = __fdiv_rn((float)i, 255.0f);
    __syncthreads();
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const int tid = threadIdx.x, hw = s.h * s.w, per = 48 + (s.h + s.w) * 8;
    const float* T = reinterpret_cast<const float*>(scratch);
    uint8_t* base = scratch + 1024;
    const Philox rng(p.seed);
    __syncthreads();
    for (int j = tid; j < g; j += kThreads) inverse_matrix(p, rng, img0 + j, s.h, reinterpret_cast<double*>(base + (size_t)j * per));
    __syncthreads();
    for (int i = tid; i < g * (s.w + s.h); i += kThreads) {
      const int j = i / (s.w + s.h), k = i - j * (s.w + s.h);
      affine_tables(reinterpret_cast<const double*>(base + (size_t)j * per), s.h, s.w,
                    reinterpret_cast<int*>(base + (size_t)j * per + 48), k);
    }
    __syncthreads();
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / s.w, x = px - y * s.w;
      float a, b, c;
      warp_pixel(src + (size_t)j * s.in_bytes, T, reinterpret_cast<const int*>(base + (size_t)j * per + 48), s.h, s.w, y, x, a, b, c);
      const uint32_t o = q255_f32(a) | (q255_f32(b) << 8) | (q255_f32(c) << 16);
      uint8_t* op = dst + (size_t)j * s.out_bytes + px * 3;
      op[0] = (uint8_t)o; op[1] = (uint8_t)(o >> 8); op[2] = (uint8_t)(o >> 16);
    }
  }
};

struct OpCcElasticFast : NoSpill, ElasticCommon {
  using Params = ElasticParams;
  static constexpr int kStages = 2, kOutBufs = 1;
  static constexpr int kMinBlocks = 3;   // 256 threads x 3 CTAs: <= 80 registers
  static constexpr bool kCifarCopy = true;
  static constexpr int kThreads = 256;
  static constexpr int kStrip = 8;
  static constexpr int kQueue = 63;      // smoothing replays (per image)
  static constexpr int kQueue2 = 255;    // bilinear replays (per image)
  __host__ __device__ static constexpr int pitch(int n) { return n | 1; }   // odd pitch of the transposed planes
  // 32x32 images: the two smoothing passes are dense 32x32x32 float64 products with the folded Gaussian operator
  // (A[o][i] = sum of the taps that the reflecting border maps onto input i), run on the tensor cores (mma.m8n8k4.f64).
  // Row pitches are chosen so that every fragment load is conflict-free: 36 doubles for row-fragment operands, 40 for the
  // column-fragment operand (the raw field).
  static constexpr int kN = 32, kPA = 36, kPF = 40;
  __host__ __device__ static bool use_mma(const Shape& s) { return s.h == kN && s.w == kN; }
  __host__ __device__ static int field_pitch(const Shape& s) { return use_mma(s) ? kPF : s.w; }
  __host__ __device__ static size_t mid_bytes(const Shape& s) {   // plane pair between the two passes
    return use_mma(s) ? (size_t)2 * kN * kPA * 8 : (size_t)2 * s.w * pitch(s.h) * 8;
  }
  // T32 | Wg | Mi | affine tables | F0 [2][h][field_pitch] f64 | mid planes f64 | DXYT [2][w][pitch(h)] f32 | queues | columns |
  // folded operator A [32][36] f64 (mma path)
  __host__ __device__ static size_t scratch_bytes(const Shape& s, int) {
    return 1024 + 256 + 48 + (size_t)(s.h + s.w) * 8 + (size_t)2 * s.h * field_pitch(s) * 8 + mid_bytes(s) +
           (size_t)2 * s.w * pitch(s.h) * 4 + (kQueue + 1) * 4 + (kQueue2 + 1) * 4 + (kThreads / 32) * (2 * kMaxLw + 1) * 8 +
           (use_mma(s) ? (size_t)kN * kPA * 8 : 0) + 32;
  }
  __device__ static __forceinline__ double* operator_matrix(const Shape& s, uint8_t* scratch) {
    return reinterpret_cast<double*>(scratch + scratch_bytes(s, 1) - 32 - (size_t)kN * kPA * 8);
  }
  __device__ static __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
  }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape& s, const Params& p, uint8_t* scratch, bool) {
    float* T = reinterpret_cast<float*>(scratch);
    double* Wg = reinterpret_cast<double*>(scratch + 1024);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = __fdiv_rn((float)i, 255.0f);
    for (int i = threadIdx.x; i <= p.lw; i += kThreads) Wg[i] = p.w[i];
    if (use_mma(s) && p.lw >= 0) {
      double* A = operator_matrix(s, scratch);
      for (int e = threadIdx.x; e < kN * kN; e += kThreads) {
        const int o = e / kN, i = e - o * kN;
        double a = 0.0;
        for (int k = -p.lw; k <= p.lw; ++k)
          if (refl_half(o + k, kN) == i) a += p.w[k < 0 ? -k : k];
        A[o * kPA + i] = a;
      }
    }
    __syncthreads();
  }

  // 8 consecutive outputs of one line: out[o] = sum_k w[|k|] * in[refl(s0 + o + k)].  The 8-entry input window rotates
  // through registers: at step u (unrolled by 8) logical entry o lives in physical slot (o + u) % 8, so no moves are needed.
  __device__ static __forceinline__ void strip(const double* line, int stride, int n, int s0, const double* Wg, int lw,
                                               double* acc) {
    double win[kStrip];
#pragma unroll
    for (int o = 0; o < kStrip; ++o) { acc[o] = 0.0; win[o] = line[refl_half(s0 + o - lw, n) * stride]; }
    for (int k0 = -lw; k0 <= lw; k0 += kStrip) {
#pragma unroll
      for (int u = 0; u < kStrip; ++u) {
        const int k = k0 + u;
        if (k <= lw) {
          const double wk = Wg[k < 0 ? -k : k];
#pragma unroll
          for (int o = 0; o < kStrip; ++o) acc[o] = __fma_rn(wk, win[(o + u) % kStrip], acc[o]);
          win[u % kStrip] = line[refl_half(s0 + kStrip + k, n) * stride];
        }
      }
    }
  }

  // literal scipy map_coordinates(order=1, mode='reflect') of one pixel of the float32 warped image W at float32 (yy, xx)
  __device__ static __noinline__ uint32_t bilinear_exact(const float* W, int h, int w, float yyf, float xxf) {
    const double yy = (double)yyf, xx = (double)xxf;
    const double fy0 = floor(yy), fx0 = floor(xx);
    const double ty = yy - fy0, tx = xx - fx0, uy = 1 - ty, ux = 1 - tx;
    const int y0 = (int)fy0, x0 = (int)fx0;
    const int ya = refl_half_any(y0, h), yb = refl_half_any(y0 + 1, h), xa = refl_half_any(x0, w), xb = refl_half_any(x0 + 1, w);
    const float* q00 = W + (ya * w + xa) * 3;
    const float* q01 = W + (ya * w + xb) * 3;
    const float* q10 = W + (yb * w + xa) * 3;
    const float* q11 = W + (yb * w + xb) * 3;
    uint32_t o = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      double acc = ((double)q00[c] * uy) * ux;
      acc = acc + ((double)q01[c] * uy) * tx;
      acc = acc + ((double)q10[c] * ty) * ux;
      acc = acc + ((double)q11[c] * ty) * tx;
      o |= q255_f32((float)acc) << (8 * c);
    }
    return o;
  }

  // literal scipy order at one point of the field F (h x w doubles)
  __device__ static __noinline__ double literal(const double* F, int fp, const double* Wg, int lw, int h, int w, int y, int x) {
    double col[2 * kMaxLw + 1];
    for (int d = -lw; d <= lw; ++d) {
      const int xx = refl_half(x + d, w);
      double a = F[y * fp + xx] * Wg[0];
      for (int k = lw; k >= 1; --k) a = a + (F[refl_half(y - k, h) * fp + xx] + F[refl_half(y + k, h) * fp + xx]) * Wg[k];
      col[d + lw] = a;
    }
    double a = col[lw] * Wg[0];
    for (int k = lw; k >= 1; --k) a = a + (col[lw - k] + col[lw + k]) * Wg[k];
    return a;
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const int tid = threadIdx.x;
    const int hw = s.h * s.w, ph = pitch(s.h), plane = s.w * ph;
    const float* T = reinterpret_cast<const float*>(scratch);
    const double* Wg = reinterpret_cast<const double*>(scratch + 1024);
    double* Mi = reinterpret_cast<double*>(scratch + 1024 + 256);
    int* tab = reinterpret_cast<int*>(Mi + 6);
    const bool mma = use_mma(s);
    const int fp = field_pitch(s), fplane = s.h * fp;
    double* F0 = reinterpret_cast<double*>(tab + 2 * (s.h + s.w));     // [2][h][fp] raw fields (x, y); (h + w) * 8 keeps alignment
    double* F1T = F0 + 2 * fplane;                                      // strips: [2][w][ph] after the y pass, transposed; mma: G [2][32][36]
    float* DXYT = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(F1T) + mid_bytes(s));   // [2][w][ph] f32 displacements, transposed
    float* Wf = reinterpret_cast<float*>(F0);                           // warped image, HWC f32 (aliases F0: 12 <= 16 B/px)
    int* qn = reinterpret_cast<int*>(DXYT + 2 * plane);                 // smoothing replay queue
    int* queue = qn + 1;
    int* q2n = queue + kQueue;                                          // bilinear replay queue
    int* queue2 = q2n + 1;
    double* qcols = reinterpret_cast<double*>(queue2 + kQueue2);        // (64 + 256) ints: stays 8-byte aligned
    const Philox rng(p.seed);
    const double tol = 4e-14 * fmax(1.0, fabs(p.alpha));
    for (int j = 0; j < g; ++j) {
      const int64_t img = img0 + j;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      __syncthreads();  // previous image's bilinear stage has finished reading Wf / DXYT / queues
      if (tid == 0) { *qn = 0; *q2n = 0; inverse_matrix(p, rng, img, s.h, Mi); }
      // ---- raw fields: F0[0] = ux, F0[1] = uy ----
      if (p.ux) {
        for (int i = tid; i < 2 * hw; i += kThreads) {
          const int f = i >= hw, px = i - f * hw, y = px / s.w, x = px - y * s.w;
          F0[f * fplane + y * fp + x] = f ? p.uy[img * hw + px] : p.ux[img * hw + px];
        }
      } else {
        for (int i = tid; i < (2 * hw) >> 2; i += kThreads) {   // hw % 4 == 0 (checked at launch)
          const int f = (4 * i) >= hw, e4 = 4 * i - f * hw;
          const uint64_t e = (uint64_t)(img * hw + e4);
          const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x81u + f, 1u);
          const int y4 = e4 / s.w, x4 = e4 - y4 * s.w;          // w % 4 == 0: the four values share a row
          double* d = F0 + f * fplane + y4 * fp + x4;
          d[0] = -1.0 + 2.0 * (double)u01(r.x); d[1] = -1.0 + 2.0 * (double)u01(r.y);
          d[2] = -1.0 + 2.0 * (double)u01(r.z); d[3] = -1.0 + 2.0 * (double)u01(r.w);
        }
      }
      __syncthreads();
      if (tid < s.w + s.h) affine_tables(Mi, s.h, s.w, tab, tid);   // h + w <= 128 < kThreads
      if (p.lw < 0) {
        for (int i = tid; i < 2 * hw; i += kThreads) {
          const int f = i >= hw, px = i - f * hw, y = px / s.w, x = px - y * s.w;
          DXYT[f * plane + x * ph + y] = (float)(F0[f * fplane + y * fp + x] * p.alpha);
        }
      } else if (mma) {
        // tensor-core passes: warp = (field, 8 output rows).  Fragment owner: r = lane / 4, c = lane % 4.
        const double* A = operator_matrix(s, scratch);
        const int lane = tid & 31, wq = tid >> 5, f = wq >> 2, m0 = (wq & 3) * 8, r = lane >> 2, c = lane & 3;
        {   // G = A * F   (axis y):  G[o][x] = sum_i A[o][i] F[i][x]
          const double* Ff = F0 + f * fplane;
          double acc[4][2] = {{0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}};
          // the operator is banded (|o - i| <= lw after the reflection is folded in): k-steps outside the band of this
          // warp's 8 output rows multiply exact zeros and are skipped (warp-uniform branch)
          const int klo = m0 - p.lw - 3, khi = m0 + 7 + p.lw;
#pragma unroll
          for (int k0 = 0; k0 < kN; k0 += 4) {
            if (k0 < klo || k0 > khi) continue;
            const double a = A[(m0 + r) * kPA + k0 + c];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) dmma(acc[nt][0], acc[nt][1], a, Ff[(k0 + c) * kPF + nt * 8 + r]);
          }
          double* G = F1T + f * kN * kPA + (m0 + r) * kPA + 2 * c;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) *reinterpret_cast<double2*>(G + nt * 8) = make_double2(acc[nt][0], acc[nt][1]);
        }
        __syncthreads();
        {   // S = G * A^T (axis x):  S[y][o] = sum_i G[y][i] A[o][i]
          const double* Gf = F1T + f * kN * kPA;
          double acc[4][2] = {{0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
          for (int k0 = 0; k0 < kN; k0 += 4) {
            const double a = Gf[(m0 + r) * kPA + k0 + c];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
              if (k0 < nt * 8 - p.lw - 3 || k0 > nt * 8 + 7 + p.lw) continue;   // outside the band of outputs nt*8 .. nt*8+7
              dmma(acc[nt][0], acc[nt][1], a, A[(nt * 8 + r) * kPA + k0 + c]);
            }
          }
          const int y = m0 + r;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int x = nt * 8 + 2 * c + u;
              const double t = acc[nt][u] * p.alpha;
              DXYT[f * plane + x * ph + y] = (float)t;
              if ((float)(t - tol) != (float)(t + tol) || p.force_exact) {     // too close to a float32 rounding boundary
                const int slot = atomicAdd(qn, 1);
                if (slot < kQueue) queue[slot] = (f << 24) | (y << 12) | x;
                else DXYT[f * plane + x * ph + y] = (float)(literal(F0 + f * fplane, fp, Wg, p.lw, s.h, s.w, y, x) * p.alpha);
              }
            }
          }
        }
      } else {
        // y pass: item = (field, strip of 8 rows, column x), x fastest -> F1T[f][x][y]
        const int sy = (s.h + kStrip - 1) / kStrip, sx = (s.w + kStrip - 1) / kStrip;
        for (int item = tid; item < 2 * s.w * sy; item += kThreads) {
          const int f = item / (s.w * sy), r = item - f * s.w * sy, st = r / s.w, x = r - st * s.w;
          double acc[kStrip];
          strip(F0 + f * fplane + x, fp, s.h, st * kStrip, Wg, p.lw, acc);
          double* o = F1T + f * plane + x * ph + st * kStrip;
#pragma unroll
          for (int k = 0; k < kStrip; ++k)
            if (st * kStrip + k < s.h) o[k] = acc[k];
        }
        __syncthreads();
        // x pass: item = (field, strip of 8 columns, row y), y fastest -> f32 displacement DXYT[f][x][y]
        for (int item = tid; item < 2 * s.h * sx; item += kThreads) {
          const int f = item / (s.h * sx), r = item - f * s.h * sx, st = r / s.h, y = r - st * s.h;
          double acc[kStrip];
          strip(F1T + f * plane + y, ph, s.w, st * kStrip, Wg, p.lw, acc);
#pragma unroll
          for (int k = 0; k < kStrip; ++k) {
            const int x = st * kStrip + k;
            if (x < s.w) {
              const double t = acc[k] * p.alpha;
              DXYT[f * plane + x * ph + y] = (float)t;
              if ((float)(t - tol) != (float)(t + tol) || p.force_exact) {     // too close to a float32 rounding boundary: replay literally
                const int slot = atomicAdd(qn, 1);
                if (slot < kQueue) queue[slot] = (f << 24) | (y << 12) | x;
                else DXYT[f * plane + x * ph + y] = (float)(literal(F0 + f * fplane, fp, Wg, p.lw, s.h, s.w, y, x) * p.alpha);
              }
            }
          }
        }
      }
      if (p.lw >= 0) {
        __syncthreads();
        // replay queue: one warp per point, lanes over the 2*lw+1 columns (each in the literal order), lane 0 finishes
        const int nq = min(*qn, kQueue);
        for (int e = tid >> 5; e < nq; e += kThreads >> 5) {
          const int f = queue[e] >> 24, y = (queue[e] >> 12) & 4095, x = queue[e] & 4095;
          const double* F = F0 + f * fplane;
          double* col = qcols + (tid >> 5) * (2 * kMaxLw + 1);
          for (int d = (tid & 31) - p.lw; d <= p.lw; d += 32) {
            const int xx = refl_half(x + d, s.w);
            double a = F[y * fp + xx] * Wg[0];
            for (int k = p.lw; k >= 1; --k)
              a = a + (F[refl_half(y - k, s.h) * fp + xx] + F[refl_half(y + k, s.h) * fp + xx]) * Wg[k];
            col[d + p.lw] = a;
          }
          __syncwarp();
          if ((tid & 31) == 0) {
            double a = col[p.lw] * Wg[0];
            for (int k = p.lw; k >= 1; --k) a = a + (col[p.lw - k] + col[p.lw + k]) * Wg[k];
            DXYT[f * plane + x * ph + y] = (float)(a * p.alpha);
          }
          __syncwarp();
        }
      }
      __syncthreads();
      // ---- affine warp (cv2.warpAffine arithmetic) into Wf (F0 is dead from here on) ----
      for (int px = tid; px < hw; px += kThreads) {
        const int y = px / s.w, x = px - y * s.w;
        float a, b, c;
        warp_pixel(im, T, tab, s.h, s.w, y, x, a, b, c);
        Wf[px * 3] = a; Wf[px * 3 + 1] = b; Wf[px * 3 + 2] = c;
      }
      __syncthreads();
      // ---- bilinear resampling at (y + dy, x + dx): float32 evaluation, literal replay near integers ----
      for (int px = tid; px < hw; px += kThreads) {
        const int y = px / s.w, x = px - y * s.w;
        const float yy = __fadd_rn((float)y, DXYT[plane + x * ph + y]), xx = __fadd_rn((float)x, DXYT[x * ph + y]);
        const float fy0 = floorf(yy), fx0 = floorf(xx);
        const float ty = __fsub_rn(yy, fy0), tx = __fsub_rn(xx, fx0), uy = __fsub_rn(1.0f, ty), ux = __fsub_rn(1.0f, tx);
        const int y0 = (int)fy0, x0 = (int)fx0;
        const int ya = refl_half_any(y0, s.h), yb = refl_half_any(y0 + 1, s.h), xa = refl_half_any(x0, s.w), xb = refl_half_any(x0 + 1, s.w);
        const float* q00 = Wf + (ya * s.w + xa) * 3;
        const float* q01 = Wf + (ya * s.w + xb) * 3;
        const float* q10 = Wf + (yb * s.w + xa) * 3;
        const float* q11 = Wf + (yb * s.w + xb) * 3;
        const float w00 = uy * ux, w01 = uy * tx, w10 = ty * ux, w11 = ty * tx;
        uint32_t o = 0;
        float dist = 1.0f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float acc = fmaf(q11[c], w11, fmaf(q10[c], w10, fmaf(q01[c], w01, q00[c] * w00)));
          const float r = acc * 255.0f;
          dist = fminf(dist, fabsf(r - rintf(r)));
          o |= (uint32_t)min(max(__float2int_rz(r), 0), 255) << (8 * c);
        }
        if (!(dist >= 1e-3f) || p.force_exact) {    // also catches NaN (injected fields)
          const int slot = atomicAdd(q2n, 1);
          if (slot < kQueue2) queue2[slot] = px;
          else o = bilinear_exact(Wf, s.h, s.w, yy, xx);
        }
        uint8_t* op = om + px * 3;
        op[0] = (uint8_t)o; op[1] = (uint8_t)(o >> 8); op[2] = (uint8_t)(o >> 16);
      }
      __syncthreads();
      const int n2 = min(*q2n, kQueue2);
      for (int e = tid; e < n2; e += kThreads) {
        const int px = queue2[e], y = px / s.w, x = px - y * s.w;
        const float yy = __fadd_rn((float)y, DXYT[plane + x * ph + y]), xx = __fadd_rn((float)x, DXYT[x * ph + y]);
        const uint32_t o = bilinear_exact(Wf, s.h, s.w, yy, xx);
        uint8_t* op = om + px * 3;
        op[0] = (uint8_t)o; op[1] = (uint8_t)(o >> 8); op[2] = (uint8_t)(o >> 16);
      }
    }
  }
};

// ------------------------------------------------------------------------------------------------
// elastic_transform, throughput form (IMGTX_CC_FAST) on the team pipeline: a team of four warps owns one 32x32 image.
//   fields     on-chip draws are uniform on the 1024-point grid (2m + 1 - 1024)/1024 of (-1, 1), built directly as fp16 from
//              Philox bits (exactly representable, so the tensor-core products see the true draw); injected float64 fields
//              are split into fp16 hi + lo
//   smoothing  S = A F A^T with the folded 32x32 Gaussian operator: mma.sync.m16n8k16 (fp16 in, float32 accumulate), A and
//              the intermediate G = A F carried as hi + lo halves (22 significant bits): |error| < 1e-6 of the field range.
//              The accumulators of the first product are re-used in registers as the A operand of the second one.
//   warp       cv2.warpAffine's integer coordinate arithmetic kept exactly (tables in float64 as OpenCV builds them); the
//              bilinear weights are 10-bit products, so sum w*b is EXACT in float32 on the 0..255 scale; source pixels come
//              from an fp16 RGBx canvas, REFLECT_101 through a shared index table
//   remap      bilinear at (y + dy, x + dx) in float32 (floor by a round-down add), scipy 'reflect' through an index table
// Everything stays on the 0..255 scale, so the output is floor(v) with no multiply.  Differences from the exact form:
// float32 instead of float64 interpolation weights and the ~1e-6 smoothing error: <= 1 LSB after floor.
struct ElasticFastParams {
  ElasticParams e;
  float alpha;
};

template <bool WARP_ONLY>
struct OpTElastic : ElasticCommon {
  using Params = ElasticFastParams;
  // the affine-only form (alpha = 0, CIFAR severity 1) is a latency chain per image (one thread inverts the matrix, 64 build
  // cv2's tables, then four rows per lane): four-warp teams, twice as many of them, hide it twice as well
  static constexpr int kTeamWarps = WARP_ONLY ? 4 : 8, kInBufs = 3, kOutBufs = 0, kMaxTeams = WARP_ONLY ? 4 : 2, kCtasPerSm = 2, kMinBlocks = 2;
  static constexpr int kTT = kTeamWarps * 32;
  static constexpr int kPA = 40;            // pitch (halves) of the fp16 operator / field rows: conflict-free fragment loads
  static constexpr int kLutLo = 64, kLut = 160;   // index tables cover [-64, 96)
  // Shared-memory traffic bounds this kernel (ncu: 125 M wavefronts per 50 000 images before the layouts below), so:
  //   * index tables hold the byte offsets of i AND i + 1 in one word (one load per axis, not two);
  //   * the fp16 canvas skips one 8-byte slot after every 16 pixels (row pitch 272 B): a lane expands four pixels (32 B), and
  //     with the skew the 16 lanes of a store wavefront fall into 16 different bank pairs (was a 4-way conflict);
  //   * the displacement rows have a pitch of 40 floats: the accumulator fragments store without bank conflicts;
  //   * the warped image W is kept as three Q7 half-words per pixel (8 bytes, not a float4): 0x8000 | round(128 v) planted
  //     under the exponent byte 0x43 IS the float 256 + v (one PRMT per tap and channel), and since the bilinear weights
  //     sum to 1 the blend comes out as 256 + result, whose integer part has floor(result) in its low byte.  The Q7
  //     rounding (<= 2^-8 of an LSB) is far inside the fast contract.
  // CTA: Ah, Al [32][40] half | refl101 row / col offset pairs for the canvas | reflect row / col offset pairs for W
  static size_t cta_scratch(const Shape&, const Params&) { return 2 * 32 * kPA * 2 + 4 * kLut * 4; }
  // team: Mi (6 f64) + tables (128 int) | canvas [32][34] x 8 B | DXY [2][32][40] f32 | W [1024] x 8 B | FT hi (, lo) [2][32][40] half
  // (the canvas has its own region so that it is built WHILE the tensor-core warps smooth the fields)
  static constexpr int kPD = 40;
  static constexpr int kOffTab = 64, kOffCanvas = 64 + 512, kOffDXY = kOffCanvas + 32 * 272, kOffW = kOffDXY + 2 * 32 * kPD * 4,
                       kOffFT = kOffW + 8192;
  static size_t team_scratch(const Shape&, const Params& p) {
    return WARP_ONLY ? kOffCanvas + 32 * 272 : kOffFT + (p.e.ux ? 10240 : 5120);
  }

  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    __half* Ah = reinterpret_cast<__half*>(cta);
    __half* Al = Ah + 32 * kPA;
    int* lut = reinterpret_cast<int*>(cta + 2 * 32 * kPA * 2);
    if (!WARP_ONLY) {
      for (int e = threadIdx.x; e < 32 * 32; e += blockDim.x) {
        const int o = e >> 5, i = e & 31;
        double a = 0.0;
        for (int k = -p.e.lw; k <= p.e.lw; ++k)
          if (refl_half(o + k, 32) == i) a += p.e.w[k < 0 ? -k : k];
        const __half hi = __double2half(a);
        Ah[o * kPA + i] = hi;
        Al[o * kPA + i] = __double2half(a - (double)__half2float(hi));
      }
    }
    for (int e = threadIdx.x; e < kLut; e += blockDim.x) {
      const int i = e - kLutLo;
      const int r0 = reflect101(i, 32), r1 = reflect101(i + 1, 32), h0 = reflect_half(i, 32), h1 = reflect_half(i + 1, 32);
      auto ccol = [](int c) { return c * 8 + (c >> 4) * 8; };
      lut[e] = (r0 * 272) | ((r1 * 272) << 16);
      lut[kLut + e] = ccol(r0) | (ccol(r1) << 16);
      lut[2 * kLut + e] = (h0 * 256) | ((h1 * 256) << 16);
      lut[3 * kLut + e] = (h0 * 8) | ((h1 * 8) << 16);
    }
  }

  // 1 / x to ~1e-16 without the float64 division sequence (one thread does this per image: latency matters)
  __device__ static __forceinline__ double rcp_fast(double x) {
    double r = (double)__frcp_rn((float)x);
    r = r * (2.0 - x * r);
    r = r * (2.0 - x * r);
    r = r * (2.0 - x * r);
    return r;
  }
  // ElasticCommon::inverse_matrix with the divisions replaced by rcp_fast (on-chip jitter: two Philox calls)
  __device__ static void inverse_matrix_fast(const ElasticParams& p, const Philox& rng, int64_t img, double* Mo) {
    double M[6];
    if (p.M) {
#pragma unroll
      for (int k = 0; k < 6; ++k) M[k] = p.M[img * 6 + k];
    } else {
      // on-chip jitter: the forward matrix through the three jittered corners in float32 (the draw itself is float32; no
      // oracle sees this matrix), then the float64 inversion and tables below as cv2 does for any matrix
      const float cs = 16.0f, q = 10.0f;
      const float x0 = cs + q, y0 = cs + q, x1 = cs + q, y1 = cs - q, x2 = cs - q, y2 = cs - q;
      const uint4 r0 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x80u, 0u);
      const uint4 r1 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x80u, 1u);
      const float c2 = (float)p.c2, two = 2.0f * c2;
      const float u[2][3] = {{x0 + (two * u01(r0.x) - c2), x1 + (two * u01(r0.z) - c2), x2 + (two * u01(r1.x) - c2)},
                             {y0 + (two * u01(r0.y) - c2), y1 + (two * u01(r0.w) - c2), y2 + (two * u01(r1.y) - c2)}};
      const float idet = 1.0f / (x0 * (y1 - y2) - y0 * (x1 - x2) + (x1 * y2 - x2 * y1));
#pragma unroll
      for (int rrow = 0; rrow < 2; ++rrow) {
        const float u0 = u[rrow][0], u1 = u[rrow][1], u2 = u[rrow][2];
        M[rrow * 3 + 0] = (double)((u0 * (y1 - y2) - y0 * (u1 - u2) + (u1 * y2 - u2 * y1)) * idet);
        M[rrow * 3 + 1] = (double)((x0 * (u1 - u2) - u0 * (x1 - x2) + (x1 * u2 - x2 * u1)) * idet);
        M[rrow * 3 + 2] = (double)((x0 * (y1 * u2 - y2 * u1) - y0 * (x1 * u2 - x2 * u1) + u0 * (x1 * y2 - x2 * y1)) * idet);
      }
    }
    double D = M[0] * M[4] - M[1] * M[3];
    D = D != 0.0 ? rcp_fast(D) : 0.0;
    const double A11 = M[4] * D, A22 = M[0] * D;
    M[0] = A11; M[1] *= -D; M[3] *= -D; M[4] = A22;
    const double b1 = -M[0] * M[2] - M[1] * M[5];
    const double b2 = -M[3] * M[2] - M[4] * M[5];
    M[2] = b1; M[5] = b2;
#pragma unroll
    for (int k = 0; k < 6; ++k) Mo[k] = M[k];
  }

  __device__ static __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  // (x, y) -> fp16 pair hi and the residual pair lo
  __device__ static __forceinline__ void split2(float x, float y, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(x, y);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(__fsub_rn(x, hf.x), __fsub_rn(y, hf.y));
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
  }

  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape&, const Params& p, uint8_t* scr,
                             const uint8_t* cta, int tt, int team) {
    const __half* Ah = reinterpret_cast<const __half*>(cta);
    const __half* Al = Ah + 32 * kPA;
    const int* lut = reinterpret_cast<const int*>(cta + 2 * 32 * kPA * 2);
    double* Mi = reinterpret_cast<double*>(scr);
    int* tab = reinterpret_cast<int*>(scr + kOffTab);
    __half* FT = reinterpret_cast<__half*>(scr + kOffFT);     // [hi/lo][f][x][kPA]  (element (y, x) of field f at x*kPA + y)
    uint8_t* canvas = scr + kOffCanvas;
    float* DXY = reinterpret_cast<float*>(scr + kOffDXY);      // [f][y][x]
    uint8_t* Wb = scr + kOffW;                                 // [y][x] float4
    const int lane = tt & 31, wq = tt >> 5;
    const Philox rng(p.e.seed);
    const bool has_lo = p.e.ux != nullptr;

    if (WARP_ONLY) {
      if (tt == 0) inverse_matrix_fast(p.e, rng, img, Mi);
      const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
      uint2* c2 = reinterpret_cast<uint2*>(canvas);
      for (int q = tt; q < 256; q += kTT) quad_to_half4(s32[3 * q], s32[3 * q + 1], s32[3 * q + 2], c2 + 4 * q + (q >> 2));
      team_sync<kTeamWarps>(team);
      if (tt < 64) affine_tables(Mi, 32, 32, tab, tt);
    }
    if (!WARP_ONLY) {
      // ---- fields -> FT (fp16, transposed) ----
      if (p.e.ux) {
        for (int i = tt; i < 2048; i += kTT) {
          const int f = i >> 10, px = i & 1023, y = px >> 5, x = px & 31;
          float v = (float)(f ? p.e.uy[img * 1024 + px] : p.e.ux[img * 1024 + px]);
          v = fminf(fmaxf(v, -1.0f), 1.0f);
          const __half hi = __float2half_rn(v);
          FT[(f * 32 + x) * kPA + y] = hi;
          FT[((2 + f) * 32 + x) * kPA + y] = __float2half_rn(__fsub_rn(v, __half2float(hi)));
        }
      } else {
        const __half2 two = __float2half2_rn(2.0f), m3 = __float2half2_rn(-3.0f), ulp = __float2half2_rn(0.0009765625f);
        for (int i = tt; i < 256; i += kTT) {   // (f, x, group of 8 rows)
          const int f = i >> 7, x = (i >> 2) & 31, y0 = (i & 3) * 8;
          const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x81u + f, (uint32_t)(i & 127));
          uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t hb = (rr[e] & 0x03FF03FFu) | 0x3C003C00u;    // two halves in [1, 2)
            const __half2 u = __hadd2(__hfma2(*reinterpret_cast<const __half2*>(&hb), two, m3), ulp);
            rr[e] = *reinterpret_cast<const uint32_t*>(&u);
          }
          *reinterpret_cast<uint4*>(FT + (f * 32 + x) * kPA + y0) = make_uint4(rr[0], rr[1], rr[2], rr[3]);
        }
      }
    }
    if (!WARP_ONLY) team_sync<kTeamWarps>(team);
    // ---- three things at once: warps 0-3 smooth the fields on the tensor cores, warp 4 inverts the matrix and builds cv2's
    // coordinate tables (float64, one lane then 64 entries), warps 5-7 expand the staged image into the fp16 canvas ----
    if (!WARP_ONLY && wq == 4) {
      if (lane == 0) inverse_matrix_fast(p.e, rng, img, Mi);
      __syncwarp();
      affine_tables(Mi, 32, 32, tab, lane);
      affine_tables(Mi, 32, 32, tab, lane + 32);
    }
    if (!WARP_ONLY && wq >= 5) {
      const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
      uint2* c2 = reinterpret_cast<uint2*>(canvas);
      for (int q = tt - 160; q < 256; q += 96) quad_to_half4(s32[3 * q], s32[3 * q + 1], s32[3 * q + 2], c2 + 4 * q + (q >> 2));
    }
    if (!WARP_ONLY && wq < 4) {
      // ---- smoothing on the tensor cores: warp = (field, 16 output rows); warps 4-7 go straight to the barrier ----
      const int f = wq >> 1, mt = wq & 1, g = lane >> 2, t = lane & 3;
      const uint32_t* Ah32 = reinterpret_cast<const uint32_t*>(Ah);
      const uint32_t* Al32 = reinterpret_cast<const uint32_t*>(Al);
      const uint32_t* Fh32 = reinterpret_cast<const uint32_t*>(FT + f * 32 * kPA);
      const uint32_t* Fl32 = reinterpret_cast<const uint32_t*>(FT + (2 + f) * 32 * kPA);
      constexpr int PW = kPA / 2;   // pitch in 32-bit words
      float c1[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) c1[nt][0] = c1[nt][1] = c1[nt][2] = c1[nt][3] = 0.0f;
#pragma unroll
      for (int kt = 0; kt < 2; ++kt) {
        uint32_t ah[4], al[4];
        const int r0 = (16 * mt + g) * PW + 8 * kt + t, r1 = r0 + 8 * PW;
        ah[0] = Ah32[r0]; ah[1] = Ah32[r1]; ah[2] = Ah32[r0 + 4]; ah[3] = Ah32[r1 + 4];
        al[0] = Al32[r0]; al[1] = Al32[r1]; al[2] = Al32[r0 + 4]; al[3] = Al32[r1 + 4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const int b = (8 * nt + g) * PW + 8 * kt + t;
          const uint32_t b0 = Fh32[b], b1 = Fh32[b + 4];
          mma16816(c1[nt], ah, b0, b1);
          mma16816(c1[nt], al, b0, b1);
          if (has_lo) mma16816(c1[nt], ah, Fl32[b], Fl32[b + 4]);
        }
      }
      float c2[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) c2[nt][0] = c2[nt][1] = c2[nt][2] = c2[nt][3] = 0.0f;
#pragma unroll
      for (int kt = 0; kt < 2; ++kt) {
        uint32_t gh[4], gl[4];
        split2(c1[2 * kt][0], c1[2 * kt][1], gh[0], gl[0]);
        split2(c1[2 * kt][2], c1[2 * kt][3], gh[1], gl[1]);
        split2(c1[2 * kt + 1][0], c1[2 * kt + 1][1], gh[2], gl[2]);
        split2(c1[2 * kt + 1][2], c1[2 * kt + 1][3], gh[3], gl[3]);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const int b = (8 * nt + g) * PW + 8 * kt + t;
          const uint32_t bh0 = Ah32[b], bh1 = Ah32[b + 4];
          mma16816(c2[nt], gh, bh0, bh1);
          mma16816(c2[nt], gl, bh0, bh1);
          mma16816(c2[nt], gh, Al32[b], Al32[b + 4]);
        }
      }
      float* D = DXY + (f * 32 + 16 * mt + g) * kPD + 2 * t;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        *reinterpret_cast<float2*>(D + 8 * nt) = make_float2(__fmul_rn(c2[nt][0], p.alpha), __fmul_rn(c2[nt][1], p.alpha));
        *reinterpret_cast<float2*>(D + 8 * nt + 8 * kPD) = make_float2(__fmul_rn(c2[nt][2], p.alpha), __fmul_rn(c2[nt][3], p.alpha));
      }
    }
    team_sync<kTeamWarps>(team);   // displacements, tables and canvas are ready
    // ---- affine warp (cv2.warpAffine coordinates), thread = column `lane`, rows wq, wq + 4, ... ----
    const int* lrow = lut + kLutLo;
    const int* lcol = lut + kLut + kLutLo;
    const int adelta = tab[2 * lane], bdelta = tab[2 * lane + 1];
#pragma unroll 2
    for (int y = wq; y < 32; y += kTeamWarps) {
      const int X = (tab[2 * (32 + y)] + adelta) >> 5, Y = (tab[2 * (32 + y) + 1] + bdelta) >> 5;
      const int sx = min(max(X >> 5, -kLutLo), kLut - kLutLo - 2), sy = min(max(Y >> 5, -kLutLo), kLut - kLutLo - 2);
      const float fx = __uint_as_float(0x4B000000u | (uint32_t)(X & 31)) - 8388608.0f,
                  fy = __uint_as_float(0x4B000000u | (uint32_t)(Y & 31)) - 8388608.0f;
      // weights in 1/1024 units as exact floats, scaled by 2^-10: w11 = fx fy, w10 = fy (32 - fx), ...
      const float w11 = fx * fy * 0.0009765625f, w10 = fy * 0.03125f - w11, w01 = fx * 0.03125f - w11,
                  w00 = 1.0f - fx * 0.03125f - w10;
      const int rr2 = lrow[sy], cc2 = lcol[sx];
      const int ra = rr2 & 0xFFFF, rb = rr2 >> 16, ca = cc2 & 0xFFFF, cb = cc2 >> 16;
      const uint2 p00 = *reinterpret_cast<const uint2*>(canvas + ra + ca), p01 = *reinterpret_cast<const uint2*>(canvas + ra + cb);
      const uint2 p10 = *reinterpret_cast<const uint2*>(canvas + rb + ca), p11 = *reinterpret_cast<const uint2*>(canvas + rb + cb);
      const float2 a00 = __half22float2(*reinterpret_cast<const __half2*>(&p00.x)), a01 = __half22float2(*reinterpret_cast<const __half2*>(&p01.x));
      const float2 a10 = __half22float2(*reinterpret_cast<const __half2*>(&p10.x)), a11 = __half22float2(*reinterpret_cast<const __half2*>(&p11.x));
      const float b00 = __low2float(*reinterpret_cast<const __half2*>(&p00.y)), b01 = __low2float(*reinterpret_cast<const __half2*>(&p01.y));
      const float b10 = __low2float(*reinterpret_cast<const __half2*>(&p10.y)), b11 = __low2float(*reinterpret_cast<const __half2*>(&p11.y));
      const float r = __fmaf_rn(a11.x, w11, __fmaf_rn(a10.x, w10, __fmaf_rn(a01.x, w01, a00.x * w00)));
      const float gg = __fmaf_rn(a11.y, w11, __fmaf_rn(a10.y, w10, __fmaf_rn(a01.y, w01, a00.y * w00)));
      const float b = __fmaf_rn(b11, w11, __fmaf_rn(b10, w10, __fmaf_rn(b01, w01, b00 * w00)));
      if (WARP_ONLY) {
        uint8_t* o = dst + (y * 32 + lane) * 3;
        o[0] = (uint8_t)floor_bits(r); o[1] = (uint8_t)floor_bits(gg); o[2] = (uint8_t)floor_bits(b);
      } else {
        const float kq = 8388608.0f + 32768.0f;   // low half-word of the sum: 0x8000 | round(128 v)
        const uint32_t qr = __float_as_uint(__fmaf_rn(r, 128.0f, kq)), qg = __float_as_uint(__fmaf_rn(gg, 128.0f, kq)),
                       qb = __float_as_uint(__fmaf_rn(b, 128.0f, kq));
        *reinterpret_cast<uint2*>(Wb + (y * 32 + lane) * 8) = make_uint2(__byte_perm(qr, qg, 0x5410u), qb);
      }
    }
    if (WARP_ONLY) return;
    team_sync<kTeamWarps>(team);
    // ---- bilinear remap at (y + dy, x + dx), scipy 'reflect' ----
    const int* hrow = lut + 2 * kLut;   // biased by kLutLo: index = floor(coordinate + 64)
    const int* hcol = lut + 3 * kLut;
    const float xb = (float)(lane + kLutLo);
#pragma unroll 2
    for (int y = wq; y < 32; y += kTeamWarps) {
      const float yy = (float)(y + kLutLo) + DXY[(32 + y) * kPD + lane], xx = xb + DXY[y * kPD + lane];
      const float ry = __fadd_rd(yy, 8388608.0f), rx = __fadd_rd(xx, 8388608.0f);
      const float ty = yy - (ry - 8388608.0f), tx = xx - (rx - 8388608.0f);
      const int iy = min((int)(__float_as_uint(ry) & 255u), kLut - 2), ix = min((int)(__float_as_uint(rx) & 255u), kLut - 2);
      const int rr2 = hrow[iy], cc2 = hcol[ix];
      const int ra = rr2 & 0xFFFF, rb = rr2 >> 16, ca = cc2 & 0xFFFF, cb = cc2 >> 16;
      const uint2 q00 = *reinterpret_cast<const uint2*>(Wb + ra + ca), q01 = *reinterpret_cast<const uint2*>(Wb + ra + cb);
      const uint2 q10 = *reinterpret_cast<const uint2*>(Wb + rb + ca), q11 = *reinterpret_cast<const uint2*>(Wb + rb + cb);
      const float w11 = ty * tx, w10 = ty - w11, w01 = tx - w11, w00 = 1.0f - ty - w01;
      // half-word h under the exponent byte 0x43: the float 256 + (h & 0x7FFF) / 128
      auto lo = [](uint32_t w) { return __uint_as_float(__byte_perm(w, 0x43000000u, 0x7104u)); };
      auto hi = [](uint32_t w) { return __uint_as_float(__byte_perm(w, 0x43000000u, 0x7324u)); };
      const float r = __fmaf_rn(lo(q11.x), w11, __fmaf_rn(lo(q10.x), w10, __fmaf_rn(lo(q01.x), w01, lo(q00.x) * w00)));
      const float gg = __fmaf_rn(hi(q11.x), w11, __fmaf_rn(hi(q10.x), w10, __fmaf_rn(hi(q01.x), w01, hi(q00.x) * w00)));
      const float b = __fmaf_rn(lo(q11.y), w11, __fmaf_rn(lo(q10.y), w10, __fmaf_rn(lo(q01.y), w01, lo(q00.y) * w00)));
      // the sums are 256 + value: the low byte of the integer part is floor(value)
      uint8_t* o = dst + (y * 32 + lane) * 3;
      o[0] = (uint8_t)floor_bits(fmaxf(r, 256.0f)); o[1] = (uint8_t)floor_bits(fmaxf(gg, 256.0f)); o[2] = (uint8_t)floor_bits(fmaxf(b, 256.0f));
    }
  }
};

// ------------------------------------------------------------------------------------------------
// elastic_transform, throughput form for images that do not fit one team's shared memory (224 x 224).  Same stages as
// OpTElastic, each as its own kernel over a chunk of the batch, intermediates (two float32 fields, the Q7 warped image:
// 24 bytes per pixel) in a stream-ordered workspace that stays in L2:
//   el_fields     U(-1, 1) fields (on-chip draws, or the injected float64 fields) and the inverse affine matrices
//   el_smooth<A>  Gaussian along one axis, 'reflect' border, float32: a CTA owns a strip of 32 lines, holds it with its
//                 reflected margins in shared memory ([sample][line], pitch 33) and every thread slides a window of eight
//                 accumulators along its line: per new sample one load of the sample, one of a weight, eight FFMA
//                 (269 taps at 224 x 224 severity 2: 54 M multiply-adds per image, ~1.3 instructions each)
//   el_warp       cv2.warpAffine's fixed-point bilinear warp of the uint8 image -> Q7 RGB (exact integers / 1024)
//   el_remap      bilinear resampling at (y + dy, x + dx), scipy 'reflect', float32 -> uint8
// float32 instead of the exact form's float64 smoothing and interpolation: displacement error < 1e-4 pixel, <= 1 LSB.
struct ElasticLargeParams {
  ElasticParams e;
  const float* wf;     // device, 2 * lw + 1 + 14 floats: the full symmetric kernel with 7 zeros on both sides
  int h, w;
};

__global__ void __launch_bounds__(256)
el_matrix_kernel(double* __restrict__ Minv, int64_t img0, int64_t cn, ElasticLargeParams p) {
  const Philox rng(p.e.seed);
  for (int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; li < cn; li += (int64_t)gridDim.x * blockDim.x)
    ElasticCommon::inverse_matrix(p.e, rng, img0 + li, p.h, Minv + li * 6);
}

// blockIdx.y walks (image, field) planes, blockIdx.x the plane in groups of four elements (one Philox call each)
__global__ void __launch_bounds__(256)
el_fields_kernel(float* __restrict__ F, int64_t img0, int64_t cn, ElasticLargeParams p) {
  const int hw = p.h * p.w, nq = (hw + 3) / 4;
  const Philox rng(p.e.seed);
  for (int64_t plane = blockIdx.y; plane < cn * 2; plane += gridDim.y) {
    const int64_t li = plane >> 1, img = img0 + li;
    const int f = (int)(plane & 1);
    float* dst = F + plane * (int64_t)hw;
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += gridDim.x * blockDim.x) {
      float v[4];
      if (p.e.ux) {
        const double* src = (f ? p.e.uy : p.e.ux) + img * hw;
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = 4 * q + k < hw ? fminf(fmaxf((float)src[4 * q + k], -1.0f), 1.0f) : 0.0f;
      } else {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x81u + f, (uint32_t)q);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (float)(rr[k] >> 8) * (2.0f / 16777216.0f) - (1.0f - 1.0f / 16777216.0f);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (4 * q + k < hw) dst[4 * q + k] = v[k];
    }
  }
}

// AXIS 0: along y (strip = 32 columns), AXIS 1: along x (strip = 32 rows).  src, dst: (planes, h, w) float32.
template <int AXIS>
__global__ void __launch_bounds__(256)
el_smooth_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t planes, ElasticLargeParams p, float scale) {
  extern __shared__ __align__(16) float esm[];
  const int L = AXIS == 0 ? p.h : p.w, other = AXIS == 0 ? p.w : p.h, lw = p.e.lw;
  const int nk = 2 * lw + 1, E = L + 2 * lw;
  float* wp = esm;                        // nk + 14 padded weights
  float* ext = esm + ((nk + 14 + 3) & ~3);   // [E][33]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int strips = (other + 31) / 32;
  for (int i = tid; i < nk + 14; i += 256) wp[i] = p.wf[i];
  for (int64_t job = blockIdx.x; job < planes * strips; job += gridDim.x) {
    const int64_t plane = job / strips;
    const int s0 = (int)(job - plane * strips) * 32;
    const float* sp = src + plane * (int64_t)p.h * p.w;
    float* dp = dst + plane * (int64_t)p.h * p.w;
    __syncthreads();
    // interior samples [lw, lw + L)
    if (AXIS == 0) {
      for (int e = tid; e < L * 32; e += 256) {
        const int j = e >> 5, c = e & 31;
        ext[(lw + j) * 33 + c] = s0 + c < other ? sp[(int64_t)j * p.w + s0 + c] : 0.0f;
      }
    } else {
      for (int r = warp; r < 32; r += 8) {
        const bool ok = s0 + r < other;
        for (int xx = lane; xx < L; xx += 32) ext[(lw + xx) * 33 + r] = ok ? sp[(int64_t)(s0 + r) * p.w + xx] : 0.0f;
      }
    }
    __syncthreads();
    // reflected margins ('reflect': d c b a | a b c d | d c b a), from the interior
    for (int e = tid; e < 2 * lw * 32; e += 256) {
      const int m = e >> 5, c = e & 31;
      const int j = m < lw ? m : L + m;                       // ext index of the margin sample
      const int period = 2 * L;
      int q = (j - lw) % period;
      if (q < 0) q += period;
      q = q >= L ? period - 1 - q : q;
      ext[j * 33 + c] = ext[(lw + q) * 33 + c];
    }
    __syncthreads();
    // out[o + j] = sum_t ext[o + t] * wpad[t - j + 7], t = 0 .. nk + 6  (wpad: 7 zeros, the kernel, 7 zeros)
    const int nblk = (L + 7) / 8;
    for (int b = warp; b < nblk; b += 8) {
      const int o = min(b * 8, L - 8 >= 0 ? L - 8 : 0);
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
      float wr[15];                                   // wp[8 tb + 0 .. 8 tb + 14]
#pragma unroll
      for (int u = 0; u < 7; ++u) wr[u] = wp[u];
      const float* col = ext + o * 33 + lane;
      const int nt = nk + 7;                          // samples that touch the eight outputs
      for (int tb = 0; tb * 8 < nt; ++tb) {
#pragma unroll
        for (int u = 0; u < 8; ++u) wr[7 + u] = 8 * tb + 7 + u < nk + 14 ? wp[8 * tb + 7 + u] : 0.0f;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int t = 8 * tb + u;
          const float v = t < nt && o + t < E ? col[t * 33] : 0.0f;
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = __fmaf_rn(v, wr[u + 7 - j], acc[j]);
        }
#pragma unroll
        for (int u = 0; u < 7; ++u) wr[u] = wr[8 + u];
      }
      if (AXIS == 0) {
        if (s0 + lane < other) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (o + j < L) dp[(int64_t)(o + j) * p.w + s0 + lane] = acc[j] * scale;
        }
      } else if (s0 + lane < other) {
        float* row = dp + (int64_t)(s0 + lane) * p.w + o;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (o + j < L) row[j] = acc[j] * scale;
      }
    }
  }
}

// cv2.warpAffine's per-column / per-row fixed-point terms, tabulated once per image as OpenCV does:
// tab[li] = adelta[w], bdelta[w], X0[h], Y0[h]
__global__ void __launch_bounds__(256)
el_tables_kernel(const double* __restrict__ Minv, int* __restrict__ tab, int64_t cn, int h, int w) {
  const int per = 2 * w + 2 * h;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cn * per; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t li = i / per;
    const int k = (int)(i - li * per);
    const double* Ms = Minv + li * 6;
    int v;
    if (k < w) v = (int)rint(Ms[0] * (double)k * 1024);
    else if (k < 2 * w) v = (int)rint(Ms[3] * (double)(k - w) * 1024);
    else if (k < 2 * w + h) v = (int)rint((Ms[1] * (double)(k - 2 * w) + Ms[2]) * 1024) + 16;
    else v = (int)rint((Ms[4] * (double)(k - 2 * w - h) + Ms[5]) * 1024) + 16;
    tab[i] = v;
  }
}

// one reflection is enough while the coordinate is within one image size of the image; further out: the general routine
__device__ __forceinline__ int el_refl101(int i, int n) {
  if (i <= -n || i >= 2 * n - 1) return reflect101(i, n);
  return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
}
__device__ __forceinline__ int el_refl_half(int i, int n) {
  if (i < -n || i >= 2 * n) return reflect_half(i, n);
  return i < 0 ? -1 - i : (i >= n ? 2 * n - 1 - i : i);
}

// cv2.warpAffine (fixed-point bilinear, REFLECT_101) -> Q7 RGB; alpha == 0: the finished uint8 image
__global__ void __launch_bounds__(256)
el_warp_kernel(const uint8_t* __restrict__ in, ushort4* __restrict__ Wq, uint8_t* __restrict__ out, const int* __restrict__ tab,
               int64_t img0, int64_t cn, ElasticLargeParams p, int direct) {
  const int hw = p.h * p.w;
  for (int64_t li = blockIdx.y; li < cn; li += gridDim.y)
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < hw; px += gridDim.x * blockDim.x) {
    const int64_t i = li * hw + px;
    const int y = px / p.w, x = px - y * p.w;
    const int* tb = tab + li * (2 * p.w + 2 * p.h);
    const uint8_t* im = in + (img0 + li) * (int64_t)hw * 3;
    const int X = (tb[2 * p.w + y] + tb[x]) >> 5, Y = (tb[2 * p.w + p.h + y] + tb[p.w + x]) >> 5;
    const int sx = X >> 5, sy = Y >> 5, fx = X & 31, fy = Y & 31;
    const int w11 = fx * fy, w10 = fy * 32 - w11, w01 = fx * 32 - w11, w00 = 1024 - fx * 32 - w10;   // sum 1024
    const int ya = el_refl101(sy, p.h), yb = el_refl101(sy + 1, p.h), xa = el_refl101(sx, p.w), xb = el_refl101(sx + 1, p.w);
    const uint8_t* q00 = im + ((int64_t)ya * p.w + xa) * 3;
    const uint8_t* q01 = im + ((int64_t)ya * p.w + xb) * 3;
    const uint8_t* q10 = im + ((int64_t)yb * p.w + xa) * 3;
    const uint8_t* q11 = im + ((int64_t)yb * p.w + xb) * 3;
    int v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) v[c] = q00[c] * w00 + q01[c] * w01 + q10[c] * w10 + q11[c] * w11;   // value * 1024, exact
    if (direct) {
      uint8_t* o = out + ((img0 + li) * (int64_t)hw + px) * 3;
      o[0] = (uint8_t)(v[0] >> 10); o[1] = (uint8_t)(v[1] >> 10); o[2] = (uint8_t)(v[2] >> 10);
    } else {
      Wq[i] = make_ushort4((unsigned short)((v[0] + 4) >> 3), (unsigned short)((v[1] + 4) >> 3), (unsigned short)((v[2] + 4) >> 3), 0);
    }
  }
}

__global__ void __launch_bounds__(256)
el_remap_kernel(const ushort4* __restrict__ Wq, const float* __restrict__ D, uint8_t* __restrict__ out, int64_t img0, int64_t cn,
                ElasticLargeParams p) {
  const int hw = p.h * p.w;
  for (int64_t li = blockIdx.y; li < cn; li += gridDim.y)
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < hw; px += gridDim.x * blockDim.x) {
    const int y = px / p.w, x = px - y * p.w;
    const float yy = (float)y + D[(li * 2 + 1) * (int64_t)hw + px], xx = (float)x + D[(li * 2) * (int64_t)hw + px];
    const float fy0 = floorf(yy), fx0 = floorf(xx);
    const float ty = yy - fy0, tx = xx - fx0;
    const int y0 = (int)fy0, x0 = (int)fx0;
    const int ya = el_refl_half(y0, p.h), yb = el_refl_half(y0 + 1, p.h), xa = el_refl_half(x0, p.w), xb = el_refl_half(x0 + 1, p.w);
    const ushort4* W = Wq + li * (int64_t)hw;
    const ushort4 q00 = W[ya * p.w + xa], q01 = W[ya * p.w + xb], q10 = W[yb * p.w + xa], q11 = W[yb * p.w + xb];
    const float w11 = ty * tx, w10 = ty - w11, w01 = tx - w11, w00 = 1.0f - ty - w01;
    const float k = 0.0078125f;   // Q7
    const float r = (__fmaf_rn((float)q11.x, w11, __fmaf_rn((float)q10.x, w10, __fmaf_rn((float)q01.x, w01, (float)q00.x * w00)))) * k;
    const float g = (__fmaf_rn((float)q11.y, w11, __fmaf_rn((float)q10.y, w10, __fmaf_rn((float)q01.y, w01, (float)q00.y * w00)))) * k;
    const float b = (__fmaf_rn((float)q11.z, w11, __fmaf_rn((float)q10.z, w10, __fmaf_rn((float)q01.z, w01, (float)q00.z * w00)))) * k;
    uint8_t* o = out + ((img0 + li) * (int64_t)hw + px) * 3;
    o[0] = (uint8_t)(int)fminf(fmaxf(r, 0.0f), 255.0f);
    o[1] = (uint8_t)(int)fminf(fmaxf(g, 0.0f), 255.0f);
    o[2] = (uint8_t)(int)fminf(fmaxf(b, 0.0f), 255.0f);
  }
}

static bool elastic_large_fast_ok(int h, int w, int lw) {
  const int L = std::max(h, w);
  return lw <= 2048 && ((size_t)(2 * lw + 1 + 14 + 3) + (size_t)(L + 2 * std::max(lw, 0)) * 33) * 4 <= (size_t)max_smem_optin() - 1024;
}

static int elastic_large_fast(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, const ElasticParams& e,
                              const double* weights_host, cudaStream_t st) {
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null image pointer");
  const int hw = h * w, lw = e.lw;
  ElasticLargeParams p;
  memset(&p, 0, sizeof(p));
  p.e = e; p.h = h; p.w = w;
  const bool direct = e.alpha == 0.0;   // zero displacement: the warp alone
  const size_t per = (size_t)hw * 24 + 64 + (size_t)(2 * w + 2 * h) * 4;
  const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(n, (int64_t)((768u << 20) / per)));   // workspace <= 768 MB
  uint8_t* ws = nullptr;
  const size_t wbytes = lw >= 0 ? (size_t)(2 * lw + 1 + 14) * 4 : 16;
  IMGTX_CUDA_OK(workspace_alloc((void**)&ws, per * chunk + wbytes + 256, st));
  float* F = reinterpret_cast<float*>(ws);                                   // [chunk][2][hw]
  float* G = F + chunk * 2 * (size_t)hw;                                     // [chunk][2][hw]
  ushort4* Wq = reinterpret_cast<ushort4*>(G + chunk * 2 * (size_t)hw);      // [chunk][hw]
  double* Minv = reinterpret_cast<double*>(Wq + chunk * (size_t)hw);         // [chunk][8]
  int* tab = reinterpret_cast<int*>(Minv + chunk * 8);                       // [chunk][2w + 2h]
  float* wf = reinterpret_cast<float*>(ws + align_up(per * chunk, 256));
  if (lw >= 0) {
    std::vector<float> wh((size_t)2 * lw + 1 + 14, 0.0f);
    for (int k = -lw; k <= lw; ++k) wh[7 + lw + k] = (float)weights_host[k < 0 ? -k : k];
    // pageable source: the call returns once the buffer has been copied to the driver's staging memory
    IMGTX_CUDA_OK(cudaMemcpyAsync(wf, wh.data(), wh.size() * 4, cudaMemcpyHostToDevice, st));
  }
  p.wf = wf;
  const int grid = sm_count() * 8;
  const size_t smem = lw >= 0 ? ((size_t)((2 * lw + 1 + 14 + 3) & ~3) + (size_t)(std::max(h, w) + 2 * lw) * 33) * 4 : 0;
  if (lw >= 0) {
    IMGTX_CUDA_OK(cudaFuncSetAttribute(el_smooth_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    IMGTX_CUDA_OK(cudaFuncSetAttribute(el_smooth_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  for (int64_t c0 = 0; c0 < n; c0 += chunk) {
    const int64_t cn = std::min<int64_t>(chunk, n - c0);
    el_matrix_kernel<<<(unsigned)((cn + 255) / 256), 256, 0, st>>>(Minv, c0, cn, p);
    IMGTX_LAUNCH_OK();
    const dim3 g2((unsigned)std::min(64, (hw + 255) / 256), (unsigned)std::min<int64_t>(cn, 4096));
    if (!direct) {
      el_fields_kernel<<<dim3(g2.x, (unsigned)std::min<int64_t>(2 * cn, 8192)), 256, 0, st>>>(F, c0, cn, p);
      IMGTX_LAUNCH_OK();
    }
    const float* D = F;
    if (!direct && lw >= 0) {
      const int64_t planes = cn * 2;
      const int sg = (int)std::min<int64_t>(planes * ((std::max(h, w) + 31) / 32), (int64_t)sm_count() * 3);
      el_smooth_kernel<0><<<sg, 256, smem, st>>>(F, G, planes, p, 1.0f);
      IMGTX_LAUNCH_OK();
      el_smooth_kernel<1><<<sg, 256, smem, st>>>(G, F, planes, p, (float)e.alpha);
      IMGTX_LAUNCH_OK();
    }
    el_tables_kernel<<<(unsigned)std::min<int64_t>(grid, (cn * (2 * w + 2 * h) + 255) / 256), 256, 0, st>>>(Minv, tab, cn, h, w);
    IMGTX_LAUNCH_OK();
    el_warp_kernel<<<g2, 256, 0, st>>>(in, Wq, out, tab, c0, cn, p, direct ? 1 : 0);
    IMGTX_LAUNCH_OK();
    if (!direct) {
      el_remap_kernel<<<g2, 256, 0, st>>>(Wq, D, out, c0, cn, p);
      IMGTX_LAUNCH_OK();
    }
  }
  IMGTX_CUDA_OK(cudaFreeAsync(ws, st));
  return IMGTX_OK;
}

// ------------------------------------------------------------------------------------------------
// jpeg_compression
// ------------------------------------------------------------------------------------------------
struct JpegParams {
  int16_t q[2][64];   // luma, chroma quantisation tables (natural order)
  uint32_t rq[2][64]; // floor(2^32 / (8 q)) + 1: (n * rq) >> 32 == n / (8 q) exactly for n < 2^17 (excess < n * 2^-32 < 1 / (8 q))
  uint2* metrics;     // device (n) per-image {sum, weighted sum} of the output bytes, or nullptr
};

#define JFIX(x) ((int)((x) * 8192 + 0.5))
constexpr int F_0_298631336 = JFIX(0.298631336), F_0_390180644 = JFIX(0.390180644), F_0_541196100 = JFIX(0.541196100),
              F_0_765366865 = JFIX(0.765366865), F_0_899976223 = JFIX(0.899976223), F_1_175875602 = JFIX(1.175875602),
              F_1_501321110 = JFIX(1.501321110), F_1_847759065 = JFIX(1.847759065), F_1_961570560 = JFIX(1.961570560),
              F_2_053119869 = JFIX(2.053119869), F_2_562915447 = JFIX(2.562915447), F_3_072711026 = JFIX(3.072711026);

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// jfdctint.c, one 1-D pass.  PASS 1: even outputs << 2, odd DESCALE(.., 11); PASS 2: even DESCALE(.., 2), odd DESCALE(.., 15)
template <int PASS>
__device__ __forceinline__ void fdct8(int* d) {
  const int tmp0 = d[0] + d[7], tmp7 = d[0] - d[7], tmp1 = d[1] + d[6], tmp6 = d[1] - d[6];
  const int tmp2 = d[2] + d[5], tmp5 = d[2] - d[5], tmp3 = d[3] + d[4], tmp4 = d[3] - d[4];
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  constexpr int SH = PASS == 1 ? 11 : 15;
  d[0] = PASS == 1 ? (tmp10 + tmp11) << 2 : descale(tmp10 + tmp11, 2);
  d[4] = PASS == 1 ? (tmp10 - tmp11) << 2 : descale(tmp10 - tmp11, 2);
  int z1 = (tmp12 + tmp13) * F_0_541196100;
  d[2] = descale(z1 + tmp13 * F_0_765366865, SH);
  d[6] = descale(z1 + tmp12 * (-F_1_847759065), SH);
  z1 = tmp4 + tmp7;
  int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
  const int z5 = (z3 + z4) * F_1_175875602;
  const int t4 = tmp4 * F_0_298631336, t5 = tmp5 * F_2_053119869, t6 = tmp6 * F_3_072711026, t7 = tmp7 * F_1_501321110;
  z1 *= -F_0_899976223; z2 *= -F_2_562915447;
  z3 = z3 * (-F_1_961570560) + z5; z4 = z4 * (-F_0_390180644) + z5;
  d[7] = descale(t4 + z1 + z3, SH); d[5] = descale(t5 + z2 + z4, SH);
  d[3] = descale(t6 + z2 + z3, SH); d[1] = descale(t7 + z1 + z4, SH);
}

// jidctint.c, one 1-D pass.  PASS 1 (columns): DESCALE(.., 11); PASS 2 (rows): DESCALE(.., 18)
template <int PASS>
__device__ __forceinline__ void idct8(int* d) {
  constexpr int SH = PASS == 1 ? 11 : 18;
  int z2 = d[2], z3 = d[6];
  int z1 = (z2 + z3) * F_0_541196100;
  const int tmp2 = z1 + z3 * (-F_1_847759065), tmp3 = z1 + z2 * F_0_765366865;
  z2 = d[0]; z3 = d[4];
  const int tmp0 = (z2 + z3) << 13, tmp1 = (z2 - z3) << 13;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  int t0 = d[7], t1 = d[5], t2 = d[3], t3 = d[1];
  z1 = t0 + t3; z2 = t1 + t2; z3 = t0 + t2;
  int z4 = t1 + t3;
  const int z5 = (z3 + z4) * F_1_175875602;
  t0 *= F_0_298631336; t1 *= F_2_053119869; t2 *= F_3_072711026; t3 *= F_1_501321110;
  z1 *= -F_0_899976223; z2 *= -F_2_562915447;
  z3 = z3 * (-F_1_961570560) + z5; z4 = z4 * (-F_0_390180644) + z5;
  t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;
  d[0] = descale(tmp10 + t3, SH); d[7] = descale(tmp10 - t3, SH);
  d[1] = descale(tmp11 + t2, SH); d[6] = descale(tmp11 - t2, SH);
  d[2] = descale(tmp12 + t1, SH); d[5] = descale(tmp12 - t1, SH);
  d[3] = descale(tmp13 + t0, SH); d[4] = descale(tmp13 - t0, SH);
}

struct OpCcJpeg {
  using Params = JpegParams;
  static constexpr bool kCifarCopy = true;
  static constexpr int kThreads = 256;
  // per image: three planes as int32 (Y: h x (w + 1), Cb/Cr: h/2 x (w/2 + 1) each).  The odd row pitch keeps the row passes,
  // where the eight threads of a block sit on eight consecutive rows, off a single bank (ncu r02m: 58 % of the shared-memory
  // wavefronts of the unpadded layout were bank-conflict replays)
  static size_t per_image(const Shape& s) { return ((size_t)s.h * (s.w + 1) + (size_t)s.h * (s.w / 2 + 1)) * 4; }
  static constexpr int kHeader = 2 * 64 * 8;   // quantiser and reciprocal per (table, coefficient), interleaved
  static size_t scratch_bytes(const Shape& s, int G) { return kHeader + (size_t)G * per_image(s) + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  // large images (224x224: 303 KB of planes): the planes of the CTA's image live in a global workspace (L2 resident)
  static size_t gtmp_bytes(const Shape& s) { return per_image(s); }
  static size_t scratch_bytes_spilled(const Shape&) { return kHeader + 16; }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    uint2* qt = reinterpret_cast<uint2*>(scratch);
    for (int i = threadIdx.x; i < 128; i += kThreads) qt[i] = make_uint2((uint32_t)p.q[i >> 6][i & 63], p.rq[i >> 6][i & 63]);
    __syncthreads();
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t* gtmp) {
    const int tid = threadIdx.x;
    const int h = s.h, w = s.w, hw = h * w, ch = h >> 1, cw = w >> 1, pw = w + 1, pcw = cw + 1;
    const int ysz = h * pw, csz = ch * pcw, plane_words = ysz + 2 * csz;
    const uint2* qtab = reinterpret_cast<const uint2*>(scratch);
    int* planes = gtmp ? reinterpret_cast<int*>(gtmp) : reinterpret_cast<int*>(scratch + kHeader);
    __syncthreads();
    // ---- colour conversion + 4:2:0 downsample (one thread per 2x2 quad) ----
    for (int i = tid; i < g * ch * cw; i += kThreads) {
      const int j = i / (ch * cw), qd = i - j * ch * cw;
      const int yc = qd / cw, xc = qd - yc * cw;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      int* Yp = planes + (size_t)j * plane_words;
      int* Cb = Yp + ysz;
      int* Cr = Cb + csz;
      int sb = 0, sr = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int y = yc * 2 + (k >> 1), x = xc * 2 + (k & 1);
        const uint8_t* q = im + (y * w + x) * 3;
        const int r = q[0], gg = q[1], b = q[2];
        Yp[y * pw + x] = (19595 * r + 38470 * gg + 7471 * b + 32768) >> 16;
        sb += (-11059 * r - 21709 * gg + 32768 * b + (128 << 16) + 32767) >> 16;
        sr += (32768 * r - 27439 * gg - 5329 * b + (128 << 16) + 32767) >> 16;
      }
      const int bias = 1 + (xc & 1);
      Cb[yc * pcw + xc] = (sb + bias) >> 2;
      Cr[yc * pcw + xc] = (sr + bias) >> 2;
    }
    __syncthreads();
    // ---- DCT round trip: 8 threads per block ----
    const int ybx = w >> 3, yblocks = (h >> 3) * ybx, cbx = cw >> 3, cblocks = (ch >> 3) * cbx;
    const int blocks_per_img = yblocks + 2 * cblocks;
    const int lines = g * blocks_per_img * 8;
    // pass A: forward rows
    for (int i = tid; i < lines; i += kThreads) {
      const int l = i & 7, bi = i >> 3;
      const int j = bi / blocks_per_img;
      int b = bi - j * blocks_per_img;
      int* base; int stride;
      int* Yp = planes + (size_t)j * plane_words;
      if (b < yblocks) { base = Yp + ((b / ybx) * 8) * pw + (b % ybx) * 8; stride = pw; }
      else { b -= yblocks; const int pl = b / cblocks; b -= pl * cblocks; base = Yp + ysz + pl * csz + ((b / cbx) * 8) * pcw + (b % cbx) * 8; stride = pcw; }
      int d[8];
      int* row = base + l * stride;
#pragma unroll
      for (int k = 0; k < 8; ++k) d[k] = row[k] - 128;
      fdct8<1>(d);
#pragma unroll
      for (int k = 0; k < 8; ++k) row[k] = d[k];
    }
    __syncthreads();
    // pass B: forward columns, quantise, dequantise, inverse columns
    for (int i = tid; i < lines; i += kThreads) {
      const int l = i & 7, bi = i >> 3;
      const int j = bi / blocks_per_img;
      int b = bi - j * blocks_per_img;
      int* base; int stride; int tq = 0;
      int* Yp = planes + (size_t)j * plane_words;
      if (b < yblocks) { base = Yp + ((b / ybx) * 8) * pw + (b % ybx) * 8; stride = pw; }
      else { tq = 1; b -= yblocks; const int pl = b / cblocks; b -= pl * cblocks; base = Yp + ysz + pl * csz + ((b / cbx) * 8) * pcw + (b % cbx) * 8; stride = pcw; }
      int d[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) d[k] = base[k * stride + l];
      fdct8<2>(d);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint2 qr = qtab[tq * 64 + k * 8 + l];     // lanes l = 0..7 read consecutive entries
        const int qt = (int)qr.x, t = d[k];
        const int qn = (int)__umulhi((uint32_t)(abs(t) + (qt << 2)), qr.y);   // (|t| + 8q/2) / (8q), exact (see JpegParams)
        d[k] = (t < 0 ? -qn : qn) * qt;
      }
      idct8<1>(d);
#pragma unroll
      for (int k = 0; k < 8; ++k) base[k * stride + l] = d[k];
    }
    __syncthreads();
    // pass C: inverse rows, range limit
    for (int i = tid; i < lines; i += kThreads) {
      const int l = i & 7, bi = i >> 3;
      const int j = bi / blocks_per_img;
      int b = bi - j * blocks_per_img;
      int* base; int stride;
      int* Yp = planes + (size_t)j * plane_words;
      if (b < yblocks) { base = Yp + ((b / ybx) * 8) * pw + (b % ybx) * 8; stride = pw; }
      else { b -= yblocks; const int pl = b / cblocks; b -= pl * cblocks; base = Yp + ysz + pl * csz + ((b / cbx) * 8) * pcw + (b % cbx) * 8; stride = pcw; }
      int d[8];
      int* row = base + l * stride;
#pragma unroll
      for (int k = 0; k < 8; ++k) d[k] = row[k];
      idct8<2>(d);
#pragma unroll
      for (int k = 0; k < 8; ++k) { int v = d[k] + 128; row[k] = v < 0 ? 0 : (v > 255 ? 255 : v); }
    }
    __syncthreads();
    // ---- fancy upsample + YCbCr -> RGB ----
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / w, x = px - y * w;
      const int* Yp = planes + (size_t)j * plane_words;
      const int* Cb = Yp + ysz;
      const int* Cr = Cb + csz;
      const int yc = y >> 1, xc = x >> 1;
      const int yn = clampi((y & 1) ? yc + 1 : yc - 1, 0, ch - 1);
      const int xn = clampi((x & 1) ? xc + 1 : xc - 1, 0, cw - 1);
      const int bias = (x & 1) ? 7 : 8;
      int cbv, crv;
      {
        const int a = Cb[yc * pcw + xc] * 3 + Cb[yn * pcw + xc], bq = Cb[yc * pcw + xn] * 3 + Cb[yn * pcw + xn];
        cbv = ((a * 3 + bq + bias) >> 4) - 128;
        const int a2 = Cr[yc * pcw + xc] * 3 + Cr[yn * pcw + xc], b2 = Cr[yc * pcw + xn] * 3 + Cr[yn * pcw + xn];
        crv = ((a2 * 3 + b2 + bias) >> 4) - 128;
      }
      const int Y = Yp[y * pw + x];
      uint8_t* o = dst + (size_t)j * s.out_bytes + px * 3;
      o[0] = clip8(Y + ((91881 * crv + 32768) >> 16));
      o[1] = clip8(Y + ((-22554 * cbv - 46802 * crv + 32768) >> 16));
      o[2] = clip8(Y + ((116130 * cbv + 32768) >> 16));
    }
    if (p.metrics) {
      // per-image metric from the finished group (still in shared memory when staged): one warp per image
      __syncthreads();
      const int lane = tid & 31;
      for (int j = tid >> 5; j < g; j += kThreads / 32) {
        const uint32_t* d32 = reinterpret_cast<const uint32_t*>(dst + (size_t)j * s.out_bytes);
        uint32_t c0 = 0, c1 = 0;
        for (int i = lane; i < (s.out_bytes >> 2); i += 32) checksum_word(d32[i], (uint32_t)i, c0, c1);
        c0 = __reduce_add_sync(0xffffffffu, c0);
        c1 = __reduce_add_sync(0xffffffffu, c1);
        if (lane == 0) p.metrics[img0 + j] = make_uint2(c0, c1);
      }
    }
  }
};

}  // namespace imgtx

using namespace imgtx;

static int impl_cc_elastic_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, double alpha,
                                   const double* weights_host, int lw, const double* M_dev, const double* ux_dev,
                                   const double* uy_dev, double c2, uint64_t seed, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 1 && w > 1 && c == 3 && n >= 0, "elastic_transform takes (n,h,w,3) images");
  IMGTX_CHECK_ARG(lw >= -1 && lw <= 4096 && (lw < 0 || weights_host), "Gaussian radius must be <= 4096");
  const bool large = h > 64 || w > 64 || lw > kMaxLw;   // the literal form below: any size, any radius
  IMGTX_CHECK_ARG((ux_dev == nullptr) == (uy_dev == nullptr), "give both displacement fields or none");
  ElasticParams p;
  memset(&p, 0, sizeof(p));
  p.alpha = alpha; p.lw = lw; p.M = M_dev; p.ux = ux_dev; p.uy = uy_dev; p.c2 = c2; p.seed = seed; p.force_exact = (flags & IMGTX_CC_REPLAY_ALL) ? 1 : 0;
  for (int i = 0; i <= lw && i <= kMaxLw; ++i) p.w[i] = weights_host[i];
  const Shape shp = make_shape(h, w, c, h, w);
  if (large) {
    // large images or long Gaussians: literal form (planes in a global workspace when they do not fit on chip, weights beyond
    // kMaxLw read from device memory)
    if (n == 0) return IMGTX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if ((flags & IMGTX_CC_FAST) && (alpha == 0.0 || lw >= 0) && fabs(alpha) <= 4096.0 && elastic_large_fast_ok(h, w, lw))
      return elastic_large_fast(in, out, n, h, w, p, weights_host, st);
    double* wd = nullptr;
    if (lw > kMaxLw) {
      IMGTX_CUDA_OK(workspace_alloc((void**)&wd, (size_t)(lw + 1) * 8, st));
      IMGTX_CUDA_OK(cudaMemcpyAsync(wd, weights_host, (size_t)(lw + 1) * 8, cudaMemcpyHostToDevice, st));
      p.wdev = wd;
    }
    const int rc = launch_image_op<OpCcElastic>(in, out, n, shp, p, st);
    if (wd) IMGTX_CUDA_OK(cudaFreeAsync(wd, st));
    return rc;
  }
  if ((flags & IMGTX_CC_FAST) && h == 32 && w == 32 && (alpha == 0.0 || (lw >= 0 && fabs(alpha) <= 16.0))) {
    ElasticFastParams f;
    f.e = p;
    f.alpha = (float)alpha;
    const int rc = alpha == 0.0 ? launch_team_op<OpTElastic<true>>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev)
                                : launch_team_op<OpTElastic<false>>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev);
    if (rc != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc; }
  }
  if (alpha == 0.0 && shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out))
    return launch_image_op<OpCcAffineWarp>(in, out, n, shp, p, (cudaStream_t)stream, 0, 96 * 1024);   // zero displacement: the warp alone (96 KB groups: 476 -> 438 us)
  if (lw <= std::min(h, w) && (h * w) % 4 == 0 && shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out)) {
    // four CTAs per SM
    const int target = 128 + 3 * 1 * shp.in_bytes + (int)OpCcElasticFast::scratch_bytes(shp, 1) + 512;  // one image per group
    return launch_image_op<OpCcElasticFast>(in, out, n, shp, p, (cudaStream_t)stream, 0, target);
  }
  return launch_image_op<OpCcElastic>(in, out, n, shp, p, (cudaStream_t)stream, 0, 72 * 1024);
}

static int impl_cc_jpeg_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                const int32_t* qtables_host /* 2 x 64, natural order */, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && (h % 16) == 0 && (w % 16) == 0 && c == 3 && n >= 0,
                  "jpeg_compression takes (n,h,w,3) images with h, w multiples of 16");
  IMGTX_CHECK_ARG(qtables_host, "missing quantisation tables");
  JpegParams p;
  for (int t = 0; t < 2; ++t)
    for (int i = 0; i < 64; ++i) {
      const int v = qtables_host[t * 64 + i];
      IMGTX_CHECK_ARG(v >= 1 && v <= 255, "quantiser out of baseline range");
      p.q[t][i] = (int16_t)v;
      p.rq[t][i] = (uint32_t)((1ull << 32) / (uint64_t)(8 * v)) + 1u;
    }
  (void)flags;   // integer arithmetic: one form (a warp- or team-per-image port measured slower: 610-640 us vs 472 us per pass)
  p.metrics = reinterpret_cast<uint2*>(metrics_dev);
  *metrics_done = true;
  return launch_image_op<OpCcJpeg>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream, 0, 80 * 1024);
}

extern "C" int imgtx_cc_elastic_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, double alpha,
                                   const double* weights_host, int lw, const double* M_dev, const double* ux_dev,
                                   const double* uy_dev, double c2, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_elastic_u8(in, out, n, h, w, c, alpha, weights_host, lw, M_dev, ux_dev, uy_dev, c2, seed, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}

extern "C" int imgtx_cc_jpeg_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                const int32_t* qtables_host /* 2 x 64, natural order */, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_jpeg_u8(in, out, n, h, w, c, qtables_host, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}
