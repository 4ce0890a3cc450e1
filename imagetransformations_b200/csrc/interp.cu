// Interpolating warps and the small float stencil:
//   imgtx_pil_transform_u8     Pillow Image.transform(size, AFFINE|PERSPECTIVE, coeffs, BILINEAR|BICUBIC,
//                              fillcolor) — libImaging/Geometry.c ImagingGenericTransform with
//                              bilinear_filter8 / bicubic_filter8, evaluated in f64 exactly like the C code.
//                              reference: transformation.py:212-226 (apply_shear, BICUBIC, white fill),
//                              fall_2025/Initial_Experiments.py:88-92 (RandomPerspective on PIL, BILINEAR)
//   imgtx_perspective_tensor_u8  transforms.RandomPerspective(p=1) applied to ToTensor(img) and converted back
//                              with ToPILImage — torchvision _perspective_grid + grid_sample(bilinear, zeros,
//                              align_corners=False) + fill mask, all f32.
//                              reference: fall_2025/transformations_code:50-66 (apply_perspective_warp)
//   imgtx_gaussian_blur_f32    TF.gaussian_blur(img, kernel_size) default sigma, reflect padding, f32 NCHW.
//                              reference: fall_2025/angellic_with_corrupt.py:54-56
// Roofline: HBM for the data, but the f64 coordinate/tap arithmetic of the Pillow path makes it FP64-bound on
// B200; none of the three is on the C2 headline path.
#include "common.cuh"
#include <math.h>
#include <algorithm>
#include <string.h>

namespace imgtx {

struct XformParams {
  double a[8];
  const double* per_image;  // device (n,8) or nullptr
  int perspective, bicubic;
  uint8_t fill[4];
};

__device__ __forceinline__ double bicubic_poly(double v1, double v2, double v3, double v4, double d) {
  const double p1 = v2;
  const double p2 = -v1 + v3;
  const double p3 = 2 * (v1 - v2) + v3 - v4;
  const double p4 = -v1 + v2 - v3 + v4;
  return p1 + d * (p2 + d * (p3 + d * p4));
}

__global__ void __launch_bounds__(256)
pil_transform_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w, int c,
                     int oh, int ow, XformParams p) {
  const int64_t total = n * oh * ow;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t img = i / ((int64_t)oh * ow);
    const int rem = (int)(i - img * oh * ow);
    const int y = rem / ow, x = rem - y * ow;
    const double* a = p.per_image ? p.per_image + img * 8 : p.a;
    const double xin = x + 0.5, yin = y + 0.5;
    double sx = a[0] * xin + a[1] * yin + a[2];
    double sy = a[3] * xin + a[4] * yin + a[5];
    if (p.perspective) {
      const double den = a[6] * xin + a[7] * yin + 1;
      sx = sx / den;
      sy = sy / den;
    }
    uint8_t* o = out + i * c;
    if (sx < 0.0 || sx >= (double)w || sy < 0.0 || sy >= (double)h) {
      for (int ch = 0; ch < c; ++ch) o[ch] = p.fill[ch];
      continue;
    }
    sx -= 0.5;
    sy -= 0.5;
    const int x0 = (int)floor(sx), y0 = (int)floor(sy);
    const double dx = sx - x0, dy = sy - y0;
    const uint8_t* im = in + img * (int64_t)h * w * c;
    if (!p.bicubic) {
      const int xl = clampi(x0, 0, w - 1), xr = clampi(x0 + 1, 0, w - 1);
      const int ya = clampi(y0, 0, h - 1);
      const bool ok2 = (y0 + 1 >= 0) && (y0 + 1 < h);
      for (int ch = 0; ch < c; ++ch) {
        const double l1 = im[((int64_t)ya * w + xl) * c + ch], r1 = im[((int64_t)ya * w + xr) * c + ch];
        const double v1 = l1 + (r1 - l1) * dx;
        double v2 = v1;
        if (ok2) {
          const double l2 = im[((int64_t)(y0 + 1) * w + xl) * c + ch], r2 = im[((int64_t)(y0 + 1) * w + xr) * c + ch];
          v2 = l2 + (r2 - l2) * dx;
        }
        const double v = v1 + (v2 - v1) * dy;
        o[ch] = (uint8_t)(int)v;
      }
    } else {
      int xs[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) xs[k] = clampi(x0 - 1 + k, 0, w - 1);
      for (int ch = 0; ch < c; ++ch) {
        double rows[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int yy = y0 - 1 + r;
          if (r == 0 || (yy >= 0 && yy < h)) {
            const uint8_t* row = im + (int64_t)clampi(yy, 0, h - 1) * w * c + ch;
            rows[r] = bicubic_poly(row[xs[0] * c], row[xs[1] * c], row[xs[2] * c], row[xs[3] * c], dx);
          } else {
            rows[r] = rows[r - 1];
          }
        }
        const double v = bicubic_poly(rows[0], rows[1], rows[2], rows[3], dy);
        o[ch] = v <= 0.0 ? 0 : (v >= 255.0 ? 255 : (uint8_t)(int)v);
      }
    }
  }
}

// Entropy-gated adaptive chain of fall_2025/Initial_Experiments.py:88-92 on PIL images, one launch:
//   RandomHorizontalFlip -> RandomAffine (Pillow AFFINE, NEAREST: libImaging affine_fixed, 16.16 fixed point, fill 0) ->
//   RandomPerspective (Pillow PERSPECTIVE, BILINEAR: float64 bilinear_filter8, fill 0).
// The intermediate images are never materialised: a tap of the bilinear stage at integer pixel (x, y) of the affine output is
// fetched THROUGH the affine's fixed-point inverse map and the flip.  Per-image record: imgtx_adaptive_t (flags bit 0 = flip,
// bit 1 = affine fired, bit 2 = perspective fired; a[6] = the FIX'ed coefficients; persp[8]).
__device__ __forceinline__ uint32_t adaptive_fetch(const uint8_t* im, const imgtx_adaptive_t& r, int h, int w, int c, int x, int y,
                                                   int ch) {
  int xin = x, yin = y;
  if (r.flags & 2) {
    xin = (r.a[2] + y * r.a[1] + x * r.a[0]) >> 16;
    yin = (r.a[5] + y * r.a[4] + x * r.a[3]) >> 16;
    if (xin < 0 || xin >= w || yin < 0 || yin >= h) return 0u;
  }
  if (r.flags & 1) xin = w - 1 - xin;
  return im[((int64_t)yin * w + xin) * c + ch];
}

__global__ void __launch_bounds__(256)
adaptive_chain_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w, int c,
                      const imgtx_adaptive_t* __restrict__ recs) {
  const int64_t total = n * h * w;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t img = i / ((int64_t)h * w);
    const int rem = (int)(i - img * h * w);
    const int y = rem / w, x = rem - y * w;
    const imgtx_adaptive_t r = recs[img];
    const uint8_t* im = in + img * (int64_t)h * w * c;
    uint8_t* o = out + i * c;
    if (!(r.flags & 4)) {
      for (int ch = 0; ch < c; ++ch) o[ch] = (uint8_t)adaptive_fetch(im, r, h, w, c, x, y, ch);
      continue;
    }
    const double xin = x + 0.5, yin = y + 0.5;
    const double den = r.persp[6] * xin + r.persp[7] * yin + 1;
    double sx = (r.persp[0] * xin + r.persp[1] * yin + r.persp[2]) / den;
    double sy = (r.persp[3] * xin + r.persp[4] * yin + r.persp[5]) / den;
    if (sx < 0.0 || sx >= (double)w || sy < 0.0 || sy >= (double)h) {
      for (int ch = 0; ch < c; ++ch) o[ch] = 0;
      continue;
    }
    sx -= 0.5;
    sy -= 0.5;
    const int x0 = (int)floor(sx), y0 = (int)floor(sy);
    const double dx = sx - x0, dy = sy - y0;
    const int xl = clampi(x0, 0, w - 1), xr = clampi(x0 + 1, 0, w - 1), ya = clampi(y0, 0, h - 1);
    const bool ok2 = (y0 + 1 >= 0) && (y0 + 1 < h);
    for (int ch = 0; ch < c; ++ch) {
      const double l1 = adaptive_fetch(im, r, h, w, c, xl, ya, ch), r1 = adaptive_fetch(im, r, h, w, c, xr, ya, ch);
      const double v1 = l1 + (r1 - l1) * dx;
      double v2 = v1;
      if (ok2) {
        const double l2 = adaptive_fetch(im, r, h, w, c, xl, y0 + 1, ch), r2 = adaptive_fetch(im, r, h, w, c, xr, y0 + 1, ch);
        v2 = l2 + (r2 - l2) * dx;
      }
      o[ch] = (uint8_t)(int)(v1 + (v2 - v1) * dy);
    }
  }
}

// torchvision tensor-backend perspective on u8 images (ToTensor -> perspective -> ToPILImage).
__global__ void __launch_bounds__(256)
perspective_tensor_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w, int c,
                          const float* __restrict__ coeffs /* (n,8) */, float fill) {
  const int64_t total = n * h * w;
  const float hw = __fmul_rn(0.5f, (float)w), hh = __fmul_rn(0.5f, (float)h);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t img = i / ((int64_t)h * w);
    const int rem = (int)(i - img * h * w);
    const int y = rem / w, x = rem - y * w;
    const float* co = coeffs + img * 8;
    // linspace(0.5, w - 0.5, w)[x] == x + 0.5 exactly for these sizes
    const float X = (float)x + 0.5f, Y = (float)y + 0.5f;
    const float t0 = __fdiv_rn(co[0], hw), t1 = __fdiv_rn(co[1], hw), t2 = __fdiv_rn(co[2], hw);
    const float t3 = __fdiv_rn(co[3], hh), t4 = __fdiv_rn(co[4], hh), t5 = __fdiv_rn(co[5], hh);
    const float g1x = __fadd_rn(__fmaf_rn(Y, t1, __fmul_rn(X, t0)), t2);   // torch.bmm's order on the CPU reference
    const float g1y = __fadd_rn(__fmaf_rn(Y, t4, __fmul_rn(X, t3)), t5);
    const float g2 = __fadd_rn(__fmaf_rn(Y, co[7], __fmul_rn(X, co[6])), 1.0f);
    const float gx = __fsub_rn(__fdiv_rn(g1x, g2), 1.0f), gy = __fsub_rn(__fdiv_rn(g1y, g2), 1.0f);
    const float ix = __fmaf_rn(__fadd_rn(gx, 1.0f), __fmul_rn(0.5f, (float)w), -0.5f);   // ATen: fma(g + 1, size/2, -0.5)
    const float iy = __fmaf_rn(__fadd_rn(gy, 1.0f), __fmul_rn(0.5f, (float)h), -0.5f);
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const float wx1 = __fsub_rn(ix, fx0), wy1 = __fsub_rn(iy, fy0);
    const float wx0 = __fsub_rn(1.0f, wx1), wy0 = __fsub_rn(1.0f, wy1);
    // guard the int conversion against wild coordinates
    const int x0 = (fx0 < -2.0f || fx0 > (float)w + 1.0f) ? -4 : (int)fx0;
    const int y0 = (fy0 < -2.0f || fy0 > (float)h + 1.0f) ? -4 : (int)fy0;
    const float w00 = __fmul_rn(wx0, wy0), w01 = __fmul_rn(wx1, wy0), w10 = __fmul_rn(wx0, wy1), w11 = __fmul_rn(wx1, wy1);
    const bool okx0 = x0 >= 0 && x0 < w, okx1 = x0 + 1 >= 0 && x0 + 1 < w;
    const bool oky0 = y0 >= 0 && y0 < h, oky1 = y0 + 1 >= 0 && y0 + 1 < h;
    const float m00 = (okx0 && oky0) ? 1.f : 0.f, m01 = (okx1 && oky0) ? 1.f : 0.f;
    const float m10 = (okx0 && oky1) ? 1.f : 0.f, m11 = (okx1 && oky1) ? 1.f : 0.f;
    const float mask = __fmaf_rn(m11, w11, __fmaf_rn(m10, w10, __fmaf_rn(m01, w01, __fmul_rn(m00, w00))));
    const uint8_t* im = in + img * (int64_t)h * w * c;
    const int cx0 = clampi(x0, 0, w - 1), cx1 = clampi(x0 + 1, 0, w - 1), cy0 = clampi(y0, 0, h - 1), cy1 = clampi(y0 + 1, 0, h - 1);
    for (int ch = 0; ch < c; ++ch) {
      const float v00 = m00 * __fdiv_rn((float)im[((int64_t)cy0 * w + cx0) * c + ch], 255.0f);
      const float v01 = m01 * __fdiv_rn((float)im[((int64_t)cy0 * w + cx1) * c + ch], 255.0f);
      const float v10 = m10 * __fdiv_rn((float)im[((int64_t)cy1 * w + cx0) * c + ch], 255.0f);
      const float v11 = m11 * __fdiv_rn((float)im[((int64_t)cy1 * w + cx1) * c + ch], 255.0f);
      float v = __fmaf_rn(v11, w11, __fmaf_rn(v10, w10, __fmaf_rn(v01, w01, __fmul_rn(v00, w00))));
      v = __fadd_rn(__fmul_rn(v, mask), __fmul_rn(__fsub_rn(1.0f, mask), fill));
      out[i * c + ch] = (uint8_t)(__float2int_rz(__fmul_rn(v, 255.0f)) & 255);
    }
  }
}

struct BlurK {
  float k[7][7];
  int ks;
};

__global__ void __launch_bounds__(256)
gaussian_blur_f32_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t planes, int h, int w, BlurK K) {
  const int64_t total = planes * h * w;
  const int r = K.ks >> 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pl = i / ((int64_t)h * w);
    const int rem = (int)(i - pl * h * w);
    const int y = rem / w, x = rem - y * w;
    const float* im = in + pl * (int64_t)h * w;
    float acc = 0.0f;
    for (int dy = 0; dy < K.ks; ++dy) {
      const int yy = reflect101(y + dy - r, h);
      for (int dx = 0; dx < K.ks; ++dx) {
        const int xx = reflect101(x + dx - r, w);
        acc = __fadd_rn(acc, __fmul_rn(im[(int64_t)yy * w + xx], K.k[dy][dx]));
      }
    }
    out[i] = acc;
  }
}

static inline unsigned grid_for(int64_t items, int threads = 256, int per_sm = 16) {
  int64_t want = (items + threads - 1) / threads;
  int64_t cap = (int64_t)sm_count() * per_sm;
  return (unsigned)(want < 1 ? 1 : (want > cap ? cap : want));
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_pil_transform_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int oh, int ow,
                                      int perspective, int bicubic, const double* coeffs_host,
                                      const double* coeffs_per_image_dev, const uint8_t* fill_host,
                                      imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && oh > 0 && ow > 0 && c >= 1 && c <= 4 && n >= 0, "bad image shape");
  IMGTX_CHECK_ARG(coeffs_host || coeffs_per_image_dev, "no transform coefficients");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null image pointer");
  XformParams p;
  memset(&p, 0, sizeof(p));
  if (coeffs_host) memcpy(p.a, coeffs_host, sizeof(double) * (perspective ? 8 : 6));
  p.per_image = coeffs_per_image_dev;
  p.perspective = perspective;
  p.bicubic = bicubic;
  for (int i = 0; i < 4; ++i) p.fill[i] = fill_host ? fill_host[i < c ? i : 0] : 0;
  pil_transform_kernel<<<grid_for(n * oh * ow), 256, 0, (cudaStream_t)stream>>>(in, out, n, h, w, c, oh, ow, p);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

extern "C" int imgtx_perspective_tensor_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                           const float* coeffs_per_image_dev, float fill, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4 && n >= 0, "bad image shape");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out && coeffs_per_image_dev, "null pointer");
  perspective_tensor_kernel<<<grid_for(n * h * w), 256, 0, (cudaStream_t)stream>>>(in, out, n, h, w, c,
                                                                                   coeffs_per_image_dev, fill);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

extern "C" int imgtx_gaussian_blur_f32(const float* in, float* out, int64_t planes, int h, int w, int kernel_size,
                                       float sigma, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && planes >= 0, "bad shape");
  IMGTX_CHECK_ARG(kernel_size >= 1 && kernel_size <= 7 && (kernel_size & 1), "kernel_size must be odd and <= 7");
  IMGTX_CHECK_ARG(kernel_size / 2 < h && kernel_size / 2 < w, "kernel larger than the image");
  if (planes == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null pointer");
  // torchvision _get_gaussian_kernel1d in f32: linspace(-half, half, k); exp(-0.5*(x/sigma)^2); normalise
  BlurK K;
  memset(&K, 0, sizeof(K));
  K.ks = kernel_size;
  float k1[7], sum = 0.0f;
  const float half = (kernel_size - 1) * 0.5f;
  for (int i = 0; i < kernel_size; ++i) {
    const float x = -half + (float)i;
    const float q = x / sigma;
    k1[i] = expf(-0.5f * (q * q));
    sum += k1[i];
  }
  for (int i = 0; i < kernel_size; ++i) k1[i] = k1[i] / sum;
  for (int i = 0; i < kernel_size; ++i)
    for (int j = 0; j < kernel_size; ++j) K.k[i][j] = k1[i] * k1[j];
  gaussian_blur_f32_kernel<<<grid_for(planes * h * w), 256, 0, (cudaStream_t)stream>>>(in, out, planes, h, w, K);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

extern "C" int imgtx_adaptive_chain_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                       const imgtx_adaptive_t* records_dev, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4 && n >= 0, "bad image shape");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out && records_dev, "null pointer");
  IMGTX_CHECK_ARG(in != out, "the chain gathers: it cannot run in place");
  const int64_t total = n * h * w;
  const int64_t grid = std::min<int64_t>((total + 255) / 256, (int64_t)imgtx::sm_count() * 16);
  imgtx::adaptive_chain_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(in, out, n, h, w, c, records_dev);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
