// CIFAR-10-C-style corruptions, part 2: stencil types (specification: oracle/cifar_c.py).
//
//   imgtx_cc_defocus_u8  cv2.filter2D(x/255, disk kernel) per channel, REFLECT_101: sparse taps, float64 accumulation in
//                        row-major tap order (bit-identical to OpenCV for these kernels).
//   imgtx_cc_csr_u8      any fixed linear map out[p] = sum_k W[k] * (x/255)[idx[k]] given in CSR form, f32 accumulation in
//                        stored order — zoom_blur (the per-severity merged operator (I + sum_z Zoom_z)/(n+1)).
//   imgtx_cc_glass_u8    Gaussian (scipy symmetric-pair f64 order, edge replicate) -> uint8 -> sequential local pixel
//                        swaps -> Gaussian.  The swap sweep is a 2-D recurrence; it runs as a skew-2 wavefront, one warp
//                        per image (lane = row), 88 steps instead of 900.
//   imgtx_cc_motion_u8   one-sided Gaussian line kernel along a per-image angle (ImageMagick MotionBlurImage restated),
//                        edge virtual pixels, f32, rounded to nearest.
//   imgtx_cc_snow_u8     noise layer -> centre zoom -> threshold -> 8-bit -> motion blur -> blend with the whitened image
//                        plus the layer and its 180-degree rotation.
// All images of a group stay in shared memory (staged TMA pipeline): HBM sees 1 read + 1 write per byte (+ draws).
// Roofline: nominally HBM (2 B/element); the wide float stencils are ALU-bound (DESIGN.md has the per-pass numbers).
#include "pipeline.cuh"
#include "rng.cuh"
#include "sweep_stencil.cuh"
#include "cifar_c_team2.cuh"
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>

namespace imgtx {

__device__ __forceinline__ uint8_t q255d(double v) {
  v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
  return (uint8_t)(int)(v * 255.0);
}
__device__ __forceinline__ uint8_t q255f(float v) {
  v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
  return (uint8_t)__float2int_rz(__fmul_rn(v, 255.0f));
}

// ------------------------------------------------------------------------------------------------
// defocus_blur
// ------------------------------------------------------------------------------------------------
constexpr int kMaxTaps = 32;
struct DefocusParams {
  int ntaps;
  int8_t dy[kMaxTaps], dx[kMaxTaps];
  double w[kMaxTaps];
};

struct OpCcDefocus : NoSpill {
  using Params = DefocusParams;
  static constexpr int kThreads = 256;
  static size_t scratch_bytes(const Shape&, int) { return 2048 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params&, uint8_t* scratch, bool) {
    double* T = reinterpret_cast<double*>(scratch);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = (double)i / 255.0;
    __syncthreads();
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const double* T = reinterpret_cast<const double*>(scratch);
    const int hw = s.h * s.w, r = 2;  // |dy|,|dx| <= 2 for every severity; interior pixels skip the reflection
    for (int i = threadIdx.x; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / s.w, x = px - y * s.w;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      double a0 = 0.0, a1 = 0.0, a2 = 0.0;
      const bool interior = (y >= r && y < s.h - r && x >= r && x < s.w - r);
      for (int k = 0; k < p.ntaps; ++k) {
        int yy = y + p.dy[k], xx = x + p.dx[k];
        if (!interior) { yy = reflect101(yy, s.h); xx = reflect101(xx, s.w); }
        const uint8_t* q = im + (yy * s.w + xx) * 3;
        const double wt = p.w[k];
        a0 = a0 + T[q[0]] * wt; a1 = a1 + T[q[1]] * wt; a2 = a2 + T[q[2]] * wt;
      }
      uint8_t* o = dst + (size_t)j * s.out_bytes + px * 3;
      o[0] = q255d(a0); o[1] = q255d(a1); o[2] = q255d(a2);
    }
  }
};

// Fast form for mirror-symmetric kernels (every canonical severity): fixed-point column sweep, exact f64 taps only for the
// elements whose fixed-point value lies within 2^-11 LSB of an integer (sweep_stencil.cuh).
struct DefocusFastParams {
  DefocusParams exact;
  uint32_t Wt[18];  // (R+1)^2 hi words then (R+1)^2 lo words, [|dy|][|dx|] at stride R+1
  int R;
};

template <int R>
struct OpCcDefocusFast : NoSpill {
  using Params = DefocusFastParams;
  static constexpr int kStages = 2, kOutBufs = 1;
  static constexpr int kThreads = 192;
  static constexpr int kMinBlocks = R == 2 ? 2 : 3;
  static constexpr int kCtasPerSm = R == 2 ? 4 : 6;   // R = 1: 6 resident CTAs measured 5 % faster than the default cap of 4
  static constexpr int kSeg = 2;  // row segments per column item
  // scratch: T64 (2 KB) | tap weights f64 [32] | tap offsets (dy, dx) int8 [32][2]
  static size_t scratch_bytes(const Shape&, int) { return 2048 + 256 + 64 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    double* T = reinterpret_cast<double*>(scratch);
    double* tw = T + 256;
    int8_t* td = reinterpret_cast<int8_t*>(tw + kMaxTaps);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = (double)i / 255.0;
    for (int i = threadIdx.x; i < p.exact.ntaps; i += kThreads) {
      tw[i] = p.exact.w[i];
      td[2 * i] = p.exact.dy[i];
      td[2 * i + 1] = p.exact.dx[i];
    }
    __syncthreads();
  }
  __device__ static __noinline__ uint32_t exact_px(const uint8_t* im, const double* T, int ntaps, int h, int w, int y, int x,
                                                   int ch) {
    const double* tw = T + 256;
    const int8_t* td = reinterpret_cast<const int8_t*>(tw + kMaxTaps);
    double a = 0.0;
    for (int k = 0; k < ntaps; ++k) {
      const int yy = reflect101(y + td[2 * k], h), xx = reflect101(x + td[2 * k + 1], w);
      a = a + T[im[(yy * w + xx) * 3 + ch]] * tw[k];
    }
    return q255d(a);
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const double* T = reinterpret_cast<const double*>(scratch);
    const int pgs = s.w >> 2, per_img = 3 * pgs * kSeg;
    const int seg_rows = (s.h + kSeg - 1) / kSeg;
    for (int item = threadIdx.x; item < g * per_img; item += kThreads) {
      const int j = item / per_img, r = item - j * per_img;
      const int seg = r / (3 * pgs), q = r - seg * 3 * pgs;
      const int pg = q / 3, ch = q - pg * 3;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      const int y0 = seg * seg_rows, y1 = min(s.h, y0 + seg_rows);
      auto ex = [&](int y, int x, int c) { return exact_px(im, T, p.exact.ntaps, s.h, s.w, y, x, c); };
      sweep_sym_stencil<R, BORDER_REFLECT101, true>(im, om, s.h, s.w, p.Wt, ch, pg, y0, y1, ex);
    }
  }
};

// ------------------------------------------------------------------------------------------------
// fixed sparse linear operator (zoom_blur)
// ------------------------------------------------------------------------------------------------
struct CsrParams {
  const int32_t* indptr;   // device, P+1
  const int32_t* indices;  // device, nnz
  const float* weights;    // device, nnz
  int nnz;
};

struct OpCcCsr : NoSpill {
  using Params = CsrParams;
  static constexpr bool kCifarCopy = true;
  static constexpr int kThreads = 1024;  // the operator table (<= 100 KB) allows one CTA per SM: fill it with warps
  // scratch: T32 (1 KB) | indptr (P+1) i32 | packed taps nnz x {u16 idx... kept as int32 idx, f32 w} (extra_scratch)
  static size_t scratch_bytes(const Shape&, int) { return 1024 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  static size_t table_bytes(int P, int nnz) { return align_up((size_t)(P + 1) * 4, 16) + (size_t)nnz * 8 + 16; }
  __device__ static void prepare(const Shape& s, const Params& p, uint8_t* scratch, bool) {
    float* T = reinterpret_cast<float*>(scratch);
    const int P = s.h * s.w;
    int32_t* ip = reinterpret_cast<int32_t*>(scratch + 1024 + 16);
    int2* taps = reinterpret_cast<int2*>(scratch + 1024 + 16 + align_up((size_t)(P + 1) * 4, 16));
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = __fdiv_rn((float)i, 255.0f);
    for (int i = threadIdx.x; i <= P; i += kThreads) ip[i] = p.indptr[i];
    for (int i = threadIdx.x; i < p.nnz; i += kThreads) taps[i] = make_int2(p.indices[i] * 3, __float_as_int(p.weights[i]));
    __syncthreads();
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const float* T = reinterpret_cast<const float*>(scratch);
    const int P = s.h * s.w;
    const int32_t* ip = reinterpret_cast<const int32_t*>(scratch + 1024 + 16);
    const int2* taps = reinterpret_cast<const int2*>(scratch + 1024 + 16 + align_up((size_t)(P + 1) * 4, 16));
    for (int i = threadIdx.x; i < g * P; i += kThreads) {
      const int j = i / P, px = i - j * P;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
      const int k1 = ip[px + 1];
      for (int k = ip[px]; k < k1; ++k) {
        const int2 t = taps[k];
        const float wt = __int_as_float(t.y);
        const uint32_t rgb = load_rgb(im, t.x);
        a0 = __fadd_rn(a0, __fmul_rn(T[rgb & 255u], wt));
        a1 = __fadd_rn(a1, __fmul_rn(T[(rgb >> 8) & 255u], wt));
        a2 = __fadd_rn(a2, __fmul_rn(T[(rgb >> 16) & 255u], wt));
      }
      uint8_t* o = dst + (size_t)j * s.out_bytes + px * 3;
      o[0] = q255f(a0); o[1] = q255f(a1); o[2] = q255f(a2);
    }
  }
};

// Operators too large for shared memory (224x224: 0.7-1.8 M taps): the same ordered float32 sums with the operator and the
// image read through L2, one thread per output pixel.
__global__ void __launch_bounds__(256)
csr_global_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int P, CsrParams p) {
  const int64_t total = n * P;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = i / P;
    const int px = (int)(i - j * P);
    const uint8_t* im = in + j * (int64_t)P * 3;
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
    const int k1 = __ldg(p.indptr + px + 1);
    for (int k = __ldg(p.indptr + px); k < k1; ++k) {
      const float wt = __ldg(p.weights + k);
      const uint8_t* q = im + (int64_t)__ldg(p.indices + k) * 3;
      a0 = __fadd_rn(a0, __fmul_rn(__fdiv_rn((float)q[0], 255.0f), wt));
      a1 = __fadd_rn(a1, __fmul_rn(__fdiv_rn((float)q[1], 255.0f), wt));
      a2 = __fadd_rn(a2, __fmul_rn(__fdiv_rn((float)q[2], 255.0f), wt));
    }
    uint8_t* o = out + i * 3;
    o[0] = q255f(a0); o[1] = q255f(a1); o[2] = q255f(a2);
  }
}

// Throughput form.  The gather is shared-memory-bandwidth bound, so everything is laid out for conflict-free access: the
// operator is re-packed per CTA into ELL form [tap][pixel] (u16 source index, f32 weight; short rows padded with zero
// weights, which leave the ordered float32 sum bit-identical), the group is converted once into PLANAR float images (one
// LUT lookup per byte instead of one per tap), and a warp owns one output row with lanes over x, so tap loads are
// consecutive words and source reads are near-consecutive.
struct CsrEllParams {
  CsrParams csr;
  int max_row;   // taps in the longest row (ELL height)
};

struct OpCcCsrEll : NoSpill {
  using Params = CsrEllParams;
  static constexpr int kThreads = 1024;
  static constexpr int kStages = 2, kOutBufs = 1;
  static constexpr bool kCifarCopy = true;
  // scratch: T32 (1 KB) | ELL idx u16 [max_row][P] | ELL w f32 [max_row][P] | per image 3 float planes   (aux = table bytes)
  __host__ __device__ static size_t table_bytes(int P, int max_row) {
    return 1024 + align_up((size_t)max_row * P * 2, 16) + align_up((size_t)max_row * P * 4, 16);
  }
  static size_t scratch_bytes(const Shape& s, int G) { return (size_t)s.aux + (size_t)G * s.in_bytes * 4 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape& s, const Params& p, uint8_t* scratch, bool) {
    const int P = s.h * s.w, tid = threadIdx.x;
    float* T = reinterpret_cast<float*>(scratch);
    uint16_t* eidx = reinterpret_cast<uint16_t*>(scratch + 1024);
    float* ew = reinterpret_cast<float*>(scratch + 1024 + align_up((size_t)p.max_row * P * 2, 16));
    for (int i = tid; i < 256; i += kThreads) T[i] = __fdiv_rn((float)i, 255.0f);
    for (int px = tid; px < P; px += kThreads) {
      const int k0 = p.csr.indptr[px], n = min(p.csr.indptr[px + 1] - k0, p.max_row);
      for (int k = 0; k < p.max_row; ++k) {
        eidx[k * P + px] = k < n ? (uint16_t)p.csr.indices[k0 + k] : (uint16_t)px;
        ew[k * P + px] = k < n ? p.csr.weights[k0 + k] : 0.0f;
      }
    }
    __syncthreads();
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const int P = s.h * s.w, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* T = reinterpret_cast<const float*>(scratch);
    const uint16_t* eidx = reinterpret_cast<const uint16_t*>(scratch + 1024);
    const float* ew = reinterpret_cast<const float*>(scratch + 1024 + align_up((size_t)p.max_row * P * 2, 16));
    float* F = reinterpret_cast<float*>(scratch + s.aux);   // [g][3][P]
    __syncthreads();
    // HWC bytes -> planar floats, 4 pixels (3 words) per step
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
    const int quads = P >> 2;   // P % 4 == 0 (checked at launch)
    for (int i = tid; i < g * quads; i += kThreads) {
      const int j = i / quads, q = i - j * quads;
      const uint32_t* w = s32 + (size_t)j * (s.in_bytes >> 2) + 3 * q;
      const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
      float* f = F + (size_t)j * 3 * P + 4 * q;
      *reinterpret_cast<float4*>(f) = make_float4(T[w0 & 255u], T[w0 >> 24], T[(w1 >> 16) & 255u], T[(w2 >> 8) & 255u]);
      *reinterpret_cast<float4*>(f + P) = make_float4(T[(w0 >> 8) & 255u], T[w1 & 255u], T[w1 >> 24], T[(w2 >> 16) & 255u]);
      *reinterpret_cast<float4*>(f + 2 * P) = make_float4(T[(w0 >> 16) & 255u], T[(w1 >> 8) & 255u], T[w2 & 255u], T[w2 >> 24]);
    }
    __syncthreads();
    // one warp per output row
    for (int r = warp; r < g * s.h; r += kThreads >> 5) {
      const int j = r / s.h, y = r - j * s.h;
      const float* f0 = F + (size_t)j * 3 * P;
      uint8_t* orow = dst + (size_t)j * s.out_bytes + (size_t)y * s.w * 3;
      for (int x = lane; x < s.w; x += 32) {
        const int px = y * s.w + x;
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
        for (int k = 0; k < p.max_row; ++k) {
          const int si = eidx[k * P + px];
          const float wt = ew[k * P + px];
          a0 = __fadd_rn(a0, __fmul_rn(f0[si], wt));
          a1 = __fadd_rn(a1, __fmul_rn(f0[P + si], wt));
          a2 = __fadd_rn(a2, __fmul_rn(f0[2 * P + si], wt));
        }
        uint8_t* o = orow + x * 3;
        o[0] = q255f(a0); o[1] = q255f(a1); o[2] = q255f(a2);
      }
    }
  }
};

// ------------------------------------------------------------------------------------------------
// glass_blur
// ------------------------------------------------------------------------------------------------
struct GlassParams {
  double w[4];  // centre, +-1, +-2 (lw <= 2 for every severity)
  int lw, iters;
  const int8_t* deltas;  // device (n, iters, h-2, w-2, 2) or nullptr (on-chip draws)
  uint64_t seed;
};

struct OpCcGlass {
  using Params = GlassParams;
  static constexpr int kThreads = 256;
  // T64 (2 KB) + one f64 plane (h*w*3*8)
  static size_t scratch_bytes(const Shape& s, int) { return 2048 + (size_t)s.in_bytes * 8 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  // large images (224x224: a 1.2 MB plane): the f64 plane of the CTA's image lives in a global workspace
  static size_t gtmp_bytes(const Shape& s) { return (size_t)s.in_bytes * 8; }
  static size_t scratch_bytes_spilled(const Shape&) { return 2048 + 16; }
  __device__ static void prepare(const Shape&, const Params&, uint8_t* scratch, bool) {
    double* T = reinterpret_cast<double*>(scratch);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = (double)i / 255.0;
    __syncthreads();
  }

  // scipy gaussian_filter (axis y first, then x), mode 'nearest'; in/out are u8 images, tmp is the f64 plane
  template <bool FINAL>
  __device__ static void gaussian(const uint8_t* in, uint8_t* out, double* tmp, const double* T, const Shape& s,
                                  const Params& p) {
    const int ne = s.in_bytes, rowb = s.w * 3;
    for (int i = threadIdx.x; i < ne; i += kThreads) {
      const int y = i / rowb, xb = i - y * rowb;
      double acc = T[in[i]] * p.w[0];
      for (int k = p.lw; k >= 1; --k)
        acc = acc + (T[in[clampi(y - k, 0, s.h - 1) * rowb + xb]] + T[in[clampi(y + k, 0, s.h - 1) * rowb + xb]]) * p.w[k];
      tmp[i] = acc;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ne; i += kThreads) {
      const int y = i / rowb, xb = i - y * rowb;
      const int x = xb / 3, ch = xb - x * 3;
      const double* row = tmp + y * rowb + ch;
      double acc = row[x * 3] * p.w[0];
      for (int k = p.lw; k >= 1; --k)
        acc = acc + (row[clampi(x - k, 0, s.w - 1) * 3] + row[clampi(x + k, 0, s.w - 1) * 3]) * p.w[k];
      out[i] = FINAL ? q255d(acc) : (uint8_t)(int)(acc * 255.0);
    }
    __syncthreads();
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t* gtmp) {
    const double* T = reinterpret_cast<const double*>(scratch);
    double* tmp = gtmp ? reinterpret_cast<double*>(gtmp) : reinterpret_cast<double*>(scratch + 2048);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int j = 0; j < g; ++j)
      gaussian<false>(src + (size_t)j * s.in_bytes, dst + (size_t)j * s.out_bytes, tmp, T, s, p);
    const int R = s.h - 2, Ccols = s.w - 2;
    const Philox rng(p.seed);
    if (R > 32) {
      // ---- swaps, images taller than 34 rows (direct mode: g == 1, the image is the caller's output batch in global memory):
      // the same skew-2 wavefront with the rows dealt to the threads of the CTA and a block barrier per step.  On-chip
      // draws: one Philox call per (image, iteration, row, block of 64 columns); block 0 is the small-image convention ----
      for (int j = 0; j < g; ++j) {
        uint8_t* im = dst + (size_t)j * s.out_bytes;
        const int64_t img = img0 + j;
        for (int it = 0; it < p.iters; ++it) {
          uint4 rb = make_uint4(0u, 0u, 0u, 0u);
          int rb_key = -1;   // (row, column block) the cached draw belongs to
          for (int t = 0; t < 2 * (R - 1) + Ccols; ++t) {
            for (int r = tid; r < R; r += kThreads) {
              const int ci = t - 2 * r;
              if (ci < 0 || ci >= Ccols) continue;
              const int h = s.h - 1 - r, w = s.w - 1 - ci;
              int dx, dy;
              if (p.deltas) {
                const int8_t* d = p.deltas + ((((img * p.iters + it) * R + r) * Ccols + ci) << 1);
                dx = d[0]; dy = d[1];
              } else {
                const int key = r | ((ci >> 6) << 16);
                if (key != rb_key) {
                  rb = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)r + ((uint32_t)(ci >> 6) << 16));
                  rb_key = key;
                }
                const int cb = ci & 63;
                dx = -(int)(((cb < 32 ? rb.x : rb.z) >> (cb & 31)) & 1u);
                dy = -(int)(((cb < 32 ? rb.y : rb.w) >> (cb & 31)) & 1u);
              }
              uint8_t* a = im + (h * s.w + w) * 3;
              uint8_t* b = im + ((h + dy) * s.w + (w + dx)) * 3;
              const uint8_t t0 = a[0], t1 = a[1], t2 = a[2];
              a[0] = b[0]; a[1] = b[1]; a[2] = b[2];
              b[0] = t0; b[1] = t1; b[2] = t2;
            }
            __syncthreads();
          }
        }
      }
    }
    // ---- swaps: one warp per image, rows on lanes, skew-2 wavefront (h - 2 <= 32) ----
    for (int j = warp; j < g && R <= 32; j += kThreads >> 5) {
      uint8_t* im = dst + (size_t)j * s.out_bytes;
      const int64_t img = img0 + j;
      for (int it = 0; it < p.iters; ++it) {
        const int h = s.h - 1 - lane;
        // on-chip draws: one bit per swap from one Philox call per (image, iteration, row) — the convention of OpCcGlassFast
        uint4 rb = make_uint4(0u, 0u, 0u, 0u);
        if (!p.deltas && lane < R) rb = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)lane);
        for (int t = 0; t < 2 * (R - 1) + Ccols; ++t) {
          const int ci = t - 2 * lane;
          if (lane < R && ci >= 0 && ci < Ccols) {
            const int w = s.w - 1 - ci;
            int dx, dy;
            if (p.deltas) {
              const int8_t* d = p.deltas + ((((img * p.iters + it) * R + lane) * Ccols + ci) << 1);
              dx = d[0]; dy = d[1];
            } else {
              dx = -(int)(((ci < 32 ? rb.x : rb.z) >> (ci & 31)) & 1u);
              dy = -(int)(((ci < 32 ? rb.y : rb.w) >> (ci & 31)) & 1u);
            }
            uint8_t* a = im + (h * s.w + w) * 3;
            uint8_t* b = im + ((h + dy) * s.w + (w + dx)) * 3;
            const uint8_t t0 = a[0], t1 = a[1], t2 = a[2];
            a[0] = b[0]; a[1] = b[1]; a[2] = b[2];
            b[0] = t0; b[1] = t1; b[2] = t2;
          }
          __syncwarp();
        }
      }
    }
    __syncthreads();
    for (int j = 0; j < g; ++j) {
      uint8_t* im = dst + (size_t)j * s.out_bytes;
      gaussian<true>(im, im, tmp, T, s, p);
    }
  }
};

// Fast form (lw <= 2, w % 4 == 0): both Gaussians run as fixed-point column sweeps with the exact f64 two-pass evaluation
// only inside the tie band (sweep_stencil.cuh); lw == 0 degenerates to a byte LUT.  The swap sweep permutes pixels without
// looking at them, so it runs on a u16 index map (one warp per image, skew-2 wavefront, one Philox call per row and
// iteration) and is applied afterwards as one gather.
struct GlassFastParams {
  GlassParams g;
  uint32_t Wt[18];
};

template <int R>
struct OpCcGlassFast : NoSpill {
  using Params = GlassFastParams;
  static constexpr int kStages = R == 1 ? 3 : 2, kOutBufs = R == 1 ? 2 : 1;   // measured per radius
  static constexpr bool kCifarCopy = R == 0;
  static constexpr int kThreads = 192;
  static constexpr int kMinBlocks = R == 2 ? 2 : 3;   // (3 CTAs at R == 2 spill: measured 25 % slower)
  static constexpr int kSeg = 2;
  // scratch: T64 (2 KB) | flat-window LUTs 2 x 256 B + their composition 256 B | per image: u16 index map
  static size_t scratch_bytes(const Shape& s, int G) { return 2048 + 768 + (size_t)G * s.h * s.w * 2 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }

  // literal scipy two-pass evaluation at one element (axis y first, then x; mode 'nearest')
  __device__ static __noinline__ uint32_t exact_px(const uint8_t* im, const double* T, const double* wt, int lw, int h,
                                                   int w, int y, int x, int ch, bool FINAL) {
    const int rowb = w * 3;
    // flat window: the answer depends on the centre byte only
    const uint8_t* flat = reinterpret_cast<const uint8_t*>(T + 256) + (FINAL ? 256 : 0);
    const int c0 = im[y * rowb + x * 3 + ch];
    bool is_flat = true;
    for (int dy = -lw; dy <= lw && is_flat; ++dy)
      for (int dx = -lw; dx <= lw; ++dx)
        if (im[clampi(y + dy, 0, h - 1) * rowb + clampi(x + dx, 0, w - 1) * 3 + ch] != c0) { is_flat = false; break; }
    if (is_flat) return flat[c0];
    double col[5];
    for (int dx = -lw; dx <= lw; ++dx) {
      const int xb = clampi(x + dx, 0, w - 1) * 3 + ch;
      double acc = T[im[y * rowb + xb]] * wt[0];
      for (int k = lw; k >= 1; --k)
        acc = acc + (T[im[clampi(y - k, 0, h - 1) * rowb + xb]] + T[im[clampi(y + k, 0, h - 1) * rowb + xb]]) * wt[k];
      col[dx + lw] = acc;
    }
    double acc = col[lw] * wt[0];
    for (int k = lw; k >= 1; --k) acc = acc + (col[lw - k] + col[lw + k]) * wt[k];
    return FINAL ? q255d(acc) : (uint32_t)(int)(acc * 255.0);
  }
  __device__ static double flat_value(double t, const double* wt, int lw) {
    double acc = t * wt[0];
    for (int k = lw; k >= 1; --k) acc = acc + (t + t) * wt[k];
    double a2 = acc * wt[0];
    for (int k = lw; k >= 1; --k) a2 = a2 + (acc + acc) * wt[k];
    return a2;
  }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    double* T = reinterpret_cast<double*>(scratch);
    uint8_t* flat = reinterpret_cast<uint8_t*>(T + 256);
    for (int i = threadIdx.x; i < 256; i += kThreads) {
      const double t = (double)i / 255.0;
      T[i] = t;
      const double a = flat_value(t, p.g.w, p.g.lw);
      flat[i] = (uint8_t)(int)(a * 255.0);
      flat[256 + i] = q255d(a);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += kThreads) flat[512 + i] = flat[256 + flat[i]];   // second Gaussian of the first
    __syncthreads();
  }

  __device__ static __forceinline__ void blur(const uint8_t* src, uint8_t* dst, int g, const Shape& s, const Params& p,
                                             const double* T, bool FINAL) {
    const uint8_t* flat = reinterpret_cast<const uint8_t*>(T + 256) + (FINAL ? 256 : 0);
    if (R == 0) {
      const int total = g * s.in_bytes;
      for (int i = threadIdx.x; i < total; i += kThreads) dst[i] = flat[src[i]];
      return;
    }
    const int pgs = s.w >> 2, per_img = 3 * pgs * kSeg;
    const int seg_rows = (s.h + kSeg - 1) / kSeg;
    for (int item = threadIdx.x; item < g * per_img; item += kThreads) {
      const int j = item / per_img, r = item - j * per_img;
      const int seg = r / (3 * pgs), q = r - seg * 3 * pgs;
      const int pg = q / 3, ch = q - pg * 3;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      const int y0 = seg * seg_rows, y1 = min(s.h, y0 + seg_rows);
      auto ex = [&](int y, int x, int c) { return exact_px(im, T, p.g.w, p.g.lw, s.h, s.w, y, x, c, FINAL); };
      sweep_sym_stencil<(R > 0 ? R : 1), BORDER_REPLICATE, false>(im, om, s.h, s.w, p.Wt, ch, pg, y0, y1, ex);
    }
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const double* T = reinterpret_cast<const double*>(scratch);
    uint16_t* maps = reinterpret_cast<uint16_t*>(scratch + 2048 + 768);
    uint8_t* work = const_cast<uint8_t*>(src);  // the staged input buffer is ours until the next group is loaded
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int hw = s.h * s.w;
    const int Rr = s.h - 2, Cc = s.w - 2;
    const Philox rng(p.g.seed);
    __syncthreads();
    if (R == 0) {
    // index map of the swap sweep (independent of the pixels)
    for (int j = 0; j < g; ++j)
      for (int px = tid; px < hw; px += kThreads) maps[j * hw + px] = (uint16_t)px;
    __syncthreads();
    for (int j = warp; j < g; j += kThreads >> 5) {
      uint16_t* m = maps + (size_t)j * hw;
      const int64_t img = img0 + j;
      for (int it = 0; it < p.g.iters; ++it) {
        const int h = s.h - 1 - lane;
        uint32_t bx = 0, by = 0, bx2 = 0, by2 = 0;  // on-chip draws: one bit per swap, 64 swaps per row at most
        if (!p.g.deltas && lane < Rr) {
          const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)lane);
          bx = r.x; by = r.y; bx2 = r.z; by2 = r.w;
        }
        if (p.g.deltas) {
          for (int t = 0; t < 2 * (Rr - 1) + Cc; ++t) {
            const int ci = t - 2 * lane;
            if (lane < Rr && ci >= 0 && ci < Cc) {
              const int w = s.w - 1 - ci;
              const int8_t* d = p.g.deltas + ((((img * p.g.iters + it) * Rr + lane) * Cc + ci) << 1);
              const int dx = d[0], dy = d[1];
              const int a = h * s.w + w, b = (h + dy) * s.w + (w + dx);
              const uint16_t ta = m[a], tb = m[b];
              m[a] = tb; m[b] = ta;
            }
            __syncwarp();
          }
        } else {
          // on-chip draws: the bits are consumed from two 64-bit shift registers, the position walks down by one
          uint64_t BX = (uint64_t)bx | ((uint64_t)bx2 << 32), BY = (uint64_t)by | ((uint64_t)by2 << 32);
          uint16_t* ma = m + h * s.w + (s.w - 1);
          const uint32_t first = 2u * (uint32_t)lane, cnt = lane < Rr ? (uint32_t)Cc : 0u;
          for (int t = 0; t < 2 * (Rr - 1) + Cc; ++t) {
            if ((uint32_t)t - first < cnt) {
              uint16_t* mb = ma - ((int)((uint32_t)BY & 1u) * s.w + (int)((uint32_t)BX & 1u));
              BX >>= 1; BY >>= 1;
              const uint16_t ta = *ma, tb = *mb;
              *ma = tb; *mb = ta;
              --ma;
            }
            __syncwarp();
          }
        }
      }
    }
    __syncthreads();
    const int quads = hw >> 2;   // w % 4 == 0 (checked at launch)
      // sigma < 0.125: both Gaussians are byte maps, so the whole corruption is one gather: dst[p] = flat2[flat1[src[map[p]]]]
      const uint8_t* lut = reinterpret_cast<const uint8_t*>(T + 256) + 512;
      for (int j = 0; j < g; ++j) {
        const uint8_t* im = src + (size_t)j * s.in_bytes;
        const uint16_t* m = maps + (size_t)j * hw;
        uint32_t* o = reinterpret_cast<uint32_t*>(dst + (size_t)j * s.out_bytes);
        for (int q = tid; q < quads; q += kThreads) {
          uint32_t px[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t v = load_rgb(im, (int)m[4 * q + e] * 3);
            px[e] = (uint32_t)lut[v & 255u] | ((uint32_t)lut[(v >> 8) & 255u] << 8) | ((uint32_t)lut[v >> 16] << 16);
          }
          uint32_t a, b, c;
          pack_rgb4(px[0], px[1], px[2], px[3], a, b, c);
          o[3 * q] = a; o[3 * q + 1] = b; o[3 * q + 2] = c;
        }
      }
      return;
    }
    for (int pass = 0; pass < 2; ++pass) {   // one call site of the sweep: one copy of its code
    // Gaussian: pass 0 src -> dst as uint8(g * 255); pass 1 work -> dst, clipped
    blur(pass == 0 ? src : work, dst, g, s, p, T, pass == 1);
    if (pass == 1) break;
    // index map of the swap sweep (independent of the pixels)
    for (int j = 0; j < g; ++j)
      for (int px = tid; px < hw; px += kThreads) maps[j * hw + px] = (uint16_t)px;
    __syncthreads();
    for (int j = warp; j < g; j += kThreads >> 5) {
      uint16_t* m = maps + (size_t)j * hw;
      const int64_t img = img0 + j;
      for (int it = 0; it < p.g.iters; ++it) {
        const int h = s.h - 1 - lane;
        uint32_t bx = 0, by = 0, bx2 = 0, by2 = 0;  // on-chip draws: one bit per swap, 64 swaps per row at most
        if (!p.g.deltas && lane < Rr) {
          const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)lane);
          bx = r.x; by = r.y; bx2 = r.z; by2 = r.w;
        }
        if (p.g.deltas) {
          for (int t = 0; t < 2 * (Rr - 1) + Cc; ++t) {
            const int ci = t - 2 * lane;
            if (lane < Rr && ci >= 0 && ci < Cc) {
              const int w = s.w - 1 - ci;
              const int8_t* d = p.g.deltas + ((((img * p.g.iters + it) * Rr + lane) * Cc + ci) << 1);
              const int dx = d[0], dy = d[1];
              const int a = h * s.w + w, b = (h + dy) * s.w + (w + dx);
              const uint16_t ta = m[a], tb = m[b];
              m[a] = tb; m[b] = ta;
            }
            __syncwarp();
          }
        } else {
          // on-chip draws: the bits are consumed from two 64-bit shift registers, the position walks down by one
          uint64_t BX = (uint64_t)bx | ((uint64_t)bx2 << 32), BY = (uint64_t)by | ((uint64_t)by2 << 32);
          uint16_t* ma = m + h * s.w + (s.w - 1);
          const uint32_t first = 2u * (uint32_t)lane, cnt = lane < Rr ? (uint32_t)Cc : 0u;
          for (int t = 0; t < 2 * (Rr - 1) + Cc; ++t) {
            if ((uint32_t)t - first < cnt) {
              uint16_t* mb = ma - ((int)((uint32_t)BY & 1u) * s.w + (int)((uint32_t)BX & 1u));
              BX >>= 1; BY >>= 1;
              const uint16_t ta = *ma, tb = *mb;
              *ma = tb; *mb = ta;
              --ma;
            }
            __syncwarp();
          }
        }
      }
    }
    __syncthreads();
    const int quads = hw >> 2;   // w % 4 == 0 (checked at launch)
    // apply the map: work[p] = dst[map[p]], four pixels per thread
    for (int j = 0; j < g; ++j) {
      const uint8_t* im = dst + (size_t)j * s.out_bytes;
      const uint16_t* m = maps + (size_t)j * hw;
      uint32_t* o = reinterpret_cast<uint32_t*>(work + (size_t)j * s.in_bytes);
      for (int q = tid; q < quads; q += kThreads) {
        uint32_t a, b, c;
        pack_rgb4(load_rgb(im, (int)m[4 * q] * 3), load_rgb(im, (int)m[4 * q + 1] * 3), load_rgb(im, (int)m[4 * q + 2] * 3),
                  load_rgb(im, (int)m[4 * q + 3] * 3), a, b, c);
        o[3 * q] = a; o[3 * q + 1] = b; o[3 * q + 2] = c;
      }
    }
    __syncthreads();
    }
  }
};

// ------------------------------------------------------------------------------------------------
// motion_blur (also the snow layer's blur)
// ------------------------------------------------------------------------------------------------
constexpr int kMaxLine = 32;
struct MotionParams {
  int width;
  int force_exact;         // test knob (imgtx_set_force_exact)
  int n_eff;               // taps 0..n_eff-1 carry all but 2^-20 LSB of the weight (throughput forms; set by fill_motion)
  float k[kMaxLine];
  const int32_t* offsets;  // device (n, width, 2) (dx, dy) or nullptr: angle drawn on chip in [lo, hi] degrees
  double lo, hi;
  uint64_t seed;
};

// ImageMagick MotionBlurImage offsets for one angle (float64, like the host code path)
__device__ inline void motion_offsets(double angle_deg, int width, int* off /* width*2 */) {
  const double a = angle_deg * 0.017453292519943295;
  const double px = (double)width * sin(a), py = (double)width * cos(a);
  const double hyp = hypot(px, py);
  for (int i = threadIdx.x & 31; i < width; i += 32) {
    off[i * 2] = (int)ceil((double)i * py / hyp - 0.5);
    off[i * 2 + 1] = (int)ceil((double)i * px / hyp - 0.5);
  }
}

struct OpCcMotion : NoSpill {
  using Params = MotionParams;
  static constexpr int kThreads = 256;
  static size_t scratch_bytes(const Shape&, int G) { return (size_t)G * kMaxLine * 2 * 4 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params&, uint8_t*, bool) {}
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    int* off = reinterpret_cast<int*>(scratch);
    const int tid = threadIdx.x, warp = tid >> 5;
    const Philox rng(p.seed);
    __syncthreads();
    if (p.offsets) {
      for (int i = tid; i < g * p.width * 2; i += kThreads) {
        const int j = i / (p.width * 2), k = i - j * p.width * 2;
        off[j * kMaxLine * 2 + k] = p.offsets[(img0 + j) * p.width * 2 + k];
      }
    } else {
      for (int j = warp; j < g; j += kThreads >> 5) {
        const int64_t img = img0 + j;
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x60u, 0u);
        motion_offsets(p.lo + (p.hi - p.lo) * (double)u01(r.x), p.width, off + j * kMaxLine * 2);
      }
    }
    __syncthreads();
    const int hw = s.h * s.w;
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / s.w, x = px - y * s.w;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      const int* oj = off + j * kMaxLine * 2;
      float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
      for (int k = 0; k < p.width; ++k) {
        const int xx = clampi(x + oj[k * 2], 0, s.w - 1), yy = clampi(y + oj[k * 2 + 1], 0, s.h - 1);
        const uint32_t rgb = load_rgb(im, (yy * s.w + xx) * 3);
        const float wt = p.k[k];
        a0 = __fadd_rn(a0, __fmul_rn((float)(rgb & 255u), wt));
        a1 = __fadd_rn(a1, __fmul_rn((float)((rgb >> 8) & 255u), wt));
        a2 = __fadd_rn(a2, __fmul_rn((float)((rgb >> 16) & 255u), wt));
      }
      uint8_t* o = dst + (size_t)j * s.out_bytes + px * 3;
      o[0] = clip8((int)floorf(__fadd_rn(a0, 0.5f)));
      o[1] = clip8((int)floorf(__fadd_rn(a1, 0.5f)));
      o[2] = clip8((int)floorf(__fadd_rn(a2, 0.5f)));
    }
  }
};

// ------------------------------------------------------------------------------------------------
// snow
// ------------------------------------------------------------------------------------------------
constexpr int kSnowMax = 64;  // max image side
struct SnowParams {
  float thresh, c6, one_minus_c6;
  double loc, scale;  // on-chip layer draws
  int8_t y0[kSnowMax], y1[kSnowMax], x0[kSnowMax], x1[kSnowMax];  // centre-zoom tables of the layer
  float ty[kSnowMax], tx[kSnowMax];
  MotionParams m;          // layer blur
  const double* layer;     // device (n,h,w) f64 or nullptr
  const int32_t* zidx;     // images larger than kSnowMax: the zoom tables in device memory (y0, y1 (h each), x0, x1 (w each))
  const float* zt;         //                              ty (h), tx (w)
};

struct OpCcSnow {
  // large images: the layers of the CTA's image live in a global workspace
  static size_t gtmp_bytes(const Shape& s) { return per_image(s); }
  static size_t scratch_bytes_spilled(const Shape&) { return 16; }
  using Params = SnowParams;
  static constexpr int kThreads = 256;
  // per image: L0 f32[hw] + L1 u8[hw] + L2 u8[hw] + offsets
  static size_t per_image(const Shape& s) { return (size_t)s.h * s.w * 6 + kMaxLine * 2 * 4; }
  static size_t scratch_bytes(const Shape& s, int G) { return align_up((size_t)G * per_image(s), 16) + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params&, uint8_t*, bool) {}
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t* gtmp) {
    const int tid = threadIdx.x, warp = tid >> 5;
    const int hw = s.h * s.w;
    float* L0 = reinterpret_cast<float*>(gtmp ? gtmp : scratch);            // [g][hw]
    int* off = reinterpret_cast<int*>(L0 + (size_t)g * hw);                  // [g][kMaxLine*2]
    uint8_t* L1 = reinterpret_cast<uint8_t*>(off + (size_t)g * kMaxLine * 2);  // [g][hw]
    uint8_t* L2 = L1 + (size_t)g * hw;
    const Philox rng(p.m.seed);
    __syncthreads();
    // noise layer (f64 draws cast to f32) and blur offsets
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int64_t img = img0 + j;
      if (p.layer) {
        L0[i] = (float)p.layer[img * hw + px];
      } else {
        const uint64_t e = (uint64_t)(img * hw + px);
        const uint4 r = rng((uint32_t)(e >> 1), (uint32_t)(e >> 33), 0x61u, 0u);
        float z0, z1;
        box_muller(r.x, r.y, z0, z1);
        L0[i] = (float)(p.loc + p.scale * (double)((e & 1) ? z1 : z0));
      }
    }
    if (p.m.offsets) {
      for (int i = tid; i < g * p.m.width * 2; i += kThreads) {
        const int j = i / (p.m.width * 2), k = i - j * p.m.width * 2;
        off[j * kMaxLine * 2 + k] = p.m.offsets[(img0 + j) * p.m.width * 2 + k];
      }
    } else {
      for (int j = warp; j < g; j += kThreads >> 5) {
        const int64_t img = img0 + j;
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x62u, 0u);
        motion_offsets(p.m.lo + (p.m.hi - p.m.lo) * (double)u01(r.x), p.m.width, off + j * kMaxLine * 2);
      }
    }
    __syncthreads();
    // centre zoom (order-1), threshold, 8-bit
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / s.w, x = px - y * s.w;
      const float* L = L0 + (size_t)j * hw;
      const bool big = p.zidx != nullptr;
      const int zy0 = big ? p.zidx[y] : p.y0[y], zy1 = big ? p.zidx[s.h + y] : p.y1[y];
      const int zx0 = big ? p.zidx[2 * s.h + x] : p.x0[x], zx1 = big ? p.zidx[2 * s.h + s.w + x] : p.x1[x];
      const float ty = big ? p.zt[y] : p.ty[y], tx = big ? p.zt[s.h + x] : p.tx[x], uy = __fsub_rn(1.0f, ty), ux = __fsub_rn(1.0f, tx);
      float acc = __fmul_rn(__fmul_rn(L[zy0 * s.w + zx0], uy), ux);
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L[zy0 * s.w + zx1], uy), tx));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L[zy1 * s.w + zx0], ty), ux));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L[zy1 * s.w + zx1], ty), tx));
      if (acc < p.thresh) acc = 0.0f;
      L1[i] = q255f(acc);
    }
    __syncthreads();
    // motion blur of the 8-bit layer, rounded back to 8 bits
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / s.w, x = px - y * s.w;
      const uint8_t* L = L1 + (size_t)j * hw;
      const int* oj = off + j * kMaxLine * 2;
      float acc = 0.0f;
      for (int k = 0; k < p.m.width; ++k) {
        const int xx = clampi(x + oj[k * 2], 0, s.w - 1), yy = clampi(y + oj[k * 2 + 1], 0, s.h - 1);
        acc = __fadd_rn(acc, __fmul_rn((float)L[yy * s.w + xx], p.m.k[k]));
      }
      L2[i] = clip8((int)floorf(__fadd_rn(acc, 0.5f)));
    }
    __syncthreads();
    // blend
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const uint8_t* im = src + (size_t)j * s.in_bytes + px * 3;
      const uint8_t* L = L2 + (size_t)j * hw;
      const float sl = __fdiv_rn((float)L[px], 255.0f), sr = __fdiv_rn((float)L[hw - 1 - px], 255.0f);
      const float r = __fdiv_rn((float)im[0], 255.0f), gg = __fdiv_rn((float)im[1], 255.0f), b = __fdiv_rn((float)im[2], 255.0f);
      float gray = __fadd_rn(__fmul_rn(r, 0.299f), __fmul_rn(gg, 0.587f));
      gray = __fadd_rn(gray, __fmul_rn(b, 0.114f));
      const float whit = __fadd_rn(__fmul_rn(gray, 1.5f), 0.5f);
      uint8_t* o = dst + (size_t)j * s.out_bytes + px * 3;
      const float ch[3] = {r, gg, b};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float mix = __fadd_rn(__fmul_rn(p.c6, ch[c]), __fmul_rn(p.one_minus_c6, fmaxf(ch[c], whit)));
        o[c] = q255f(__fadd_rn(__fadd_rn(mix, sl), sr));
      }
    }
  }
};

// ------------------------------------------------------------------------------------------------
// motion_blur / snow, throughput forms
// ------------------------------------------------------------------------------------------------
// The ordered float32 sum is kept literally (it IS the specification), but organised around words: the image is copied once
// into an edge-replicated canvas that covers every tap offset of that image's line, so a tap is one unaligned 4-byte fetch
// (two LDS + one funnel shift) serving four outputs, with no clamping and no index arithmetic in the inner loop; bytes
// become floats with a byte-permute into the 2^23 mantissa instead of a conversion instruction.
struct LineGeom {   // per image, in shared memory
  int dxmin, dymin, pw, ph;   // canvas = image extended by the bounding box of the tap offsets
  int off[kMaxLine];          // byte offset of tap i relative to the output position inside the canvas
};

// offsets (dx, dy) of one image -> LineGeom; one warp, CH = bytes per pixel
template <int CH>
__device__ __forceinline__ void make_line_geom(LineGeom* gm, const int* offs /* (width,2) dx,dy */, int width, int w, int h) {
  const int lane = threadIdx.x & 31;
  int x0 = 0, x1 = 0, y0 = 0, y1 = 0;
  for (int i = lane; i < width; i += 32) {
    x0 = min(x0, offs[2 * i]); x1 = max(x1, offs[2 * i]);
    y0 = min(y0, offs[2 * i + 1]); y1 = max(y1, offs[2 * i + 1]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    x0 = min(x0, __shfl_xor_sync(0xffffffffu, x0, o)); x1 = max(x1, __shfl_xor_sync(0xffffffffu, x1, o));
    y0 = min(y0, __shfl_xor_sync(0xffffffffu, y0, o)); y1 = max(y1, __shfl_xor_sync(0xffffffffu, y1, o));
  }
  x0 = -((-x0 + 3) & ~3);                                        // left border: a multiple of 4 pixels, so image words stay words
  const int pw = (w + (x1 - x0) + 3) & ~3, ph = h + (y1 - y0);   // row pitch in bytes is a multiple of 4 for any CH
  if (lane == 0) { gm->dxmin = x0; gm->dymin = y0; gm->pw = pw; gm->ph = ph; }
  for (int i = lane; i < width; i += 32) gm->off[i] = ((offs[2 * i + 1] - y0) * pw + (offs[2 * i] - x0)) * CH;
}

// edge-replicated canvas of one image (all threads of the CTA: `tid` of `nt`): one warp per canvas row, one lane per canvas
// word.  The left border is a multiple of 4 pixels and w % 4 == 0, so the image part of a row is a plain word copy and the
// borders are the 3-word (CH = 3) or 1-word (CH = 1) pattern of the replicated edge pixel.  `im` must be 4-byte aligned.
template <int CH>
__device__ __forceinline__ void fill_canvas(uint8_t* canvas, const uint8_t* im, const LineGeom& gm, int w, int h, int tid, int nt) {
  static_assert(CH == 1 || CH == 3, "pattern words are written for 1 or 3 channels");
  const int lane = tid & 31, warp = tid >> 5, nw = nt >> 5;
  const int wpr = (gm.pw * CH) >> 2, lw = (-gm.dxmin * CH) >> 2, iw = (w * CH) >> 2;
  const uint32_t* im32 = reinterpret_cast<const uint32_t*>(im);
  uint32_t* c32 = reinterpret_cast<uint32_t*>(canvas);
  // narrow canvases (one channel): several rows per warp, lpr lanes per row
  const int lpr = wpr <= 8 ? 8 : (wpr <= 16 ? 16 : 32), rpi = 32 / lpr, lr = lane / lpr, lq = lane - lr * lpr;
  for (int py0 = warp * rpi; py0 < gm.ph; py0 += nw * rpi) {
    const int py = py0 + lr;
    if (py >= gm.ph) continue;
    const int sy = min(max(py + gm.dymin, 0), h - 1);
    const uint32_t* row = im32 + sy * iw;
    uint32_t L[3], R[3];
    if (CH == 3) {
      const uint32_t a = row[0] & 0x00FFFFFFu, b = row[iw - 1] >> 8;
      pack_rgb4(a, a, a, a, L[0], L[1], L[2]);
      pack_rgb4(b, b, b, b, R[0], R[1], R[2]);
    } else {
      L[0] = L[1] = L[2] = (row[0] & 255u) * 0x01010101u;
      R[0] = R[1] = R[2] = (row[iw - 1] >> 24) * 0x01010101u;
    }
    for (int q = lq; q < wpr; q += lpr) {
      uint32_t v;
      if (q < lw) {
        const int m = CH == 3 ? q % 3 : 0;
        v = m == 0 ? L[0] : (m == 1 ? L[1] : L[2]);
      } else if (q < lw + iw) {
        v = row[q - lw];
      } else {
        const int m = CH == 3 ? (q - lw - iw) % 3 : 0;
        v = m == 0 ? R[0] : (m == 1 ? R[1] : R[2]);
      }
      c32[py * wpr + q] = v;
    }
  }
}

// four consecutive output bytes starting at byte `ob` of an image with `rowb` bytes per row (rowb % 4 == 0)
template <int CH>
__device__ __forceinline__ uint32_t line_blur_word(const uint8_t* canvas, const LineGeom& gm, const float* k, int width,
                                                   int rowb, int ob) {
  const int y = ob / rowb, xb = ob - y * rowb;
  const uint32_t* c32 = reinterpret_cast<const uint32_t*>(canvas);   // canvas is 16-byte aligned
  const int base = y * gm.pw * CH + xb;
  float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
  for (int i = 0; i < width; ++i) {
    const int bo = base + gm.off[i];
    const uint32_t* wp = c32 + (bo >> 2);
    const uint32_t v = __funnelshift_r(wp[0], wp[1], (uint32_t)bo << 3);
    const float wt = k[i];
    a0 = __fadd_rn(a0, __fmul_rn(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7440u)), 8388608.0f), wt));
    a1 = __fadd_rn(a1, __fmul_rn(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7441u)), 8388608.0f), wt));
    a2 = __fadd_rn(a2, __fmul_rn(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7442u)), 8388608.0f), wt));
    a3 = __fadd_rn(a3, __fmul_rn(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7443u)), 8388608.0f), wt));
  }
  const int r0 = min(max(__float2int_rd(__fadd_rn(a0, 0.5f)), 0), 255), r1 = min(max(__float2int_rd(__fadd_rn(a1, 0.5f)), 0), 255);
  const int r2 = min(max(__float2int_rd(__fadd_rn(a2, 0.5f)), 0), 255), r3 = min(max(__float2int_rd(__fadd_rn(a3, 0.5f)), 0), 255);
  return (uint32_t)r0 | ((uint32_t)r1 << 8) | ((uint32_t)r2 << 16) | ((uint32_t)r3 << 24);
}

// Throughput evaluation of the same sum.  The specification is the float32 sum in tap order; here the taps that carry
// weight (the one-sided Gaussian underflows quickly: 7 of 21 taps at sigma = 1) are accumulated with fused multiply-adds in
// float32, which differs from the ordered sum by less than 4e-4 (every partial sum is below 256, so each of the <= 25 + 19 roundings
// is at most 2^-17, plus the products, the dropped tail and the rounding of acc + 0.5), and an output is only trusted when
// acc + 0.5 is at least 5e-4 away from an integer; the other words (about 0.4 %) are queued and replayed literally.
template <int CH, int NW>
__device__ __forceinline__ void line_blur_words_fast(const uint8_t* canvas, const LineGeom& gm, const float* k, int n_eff,
                                                     int rowb, int ob, uint32_t* out /* NW words */, bool& amb) {
  const int y = ob / rowb, xb = ob - y * rowb;
  const uint32_t* c32 = reinterpret_cast<const uint32_t*>(canvas);   // canvas is 16-byte aligned
  const int base = y * gm.pw * CH + xb;
  float a[4 * NW];
#pragma unroll
  for (int e = 0; e < 4 * NW; ++e) a[e] = 0.0f;
  for (int i = 0; i < n_eff; ++i) {
    const int bo = base + gm.off[i];
    const uint32_t* wp = c32 + (bo >> 2);
    const float wt = k[i];
    uint32_t w[NW + 1];
#pragma unroll
    for (int q = 0; q <= NW; ++q) w[q] = wp[q];
#pragma unroll
    for (int q = 0; q < NW; ++q) {
      const uint32_t v = __funnelshift_r(w[q], w[q + 1], (uint32_t)bo << 3);
      a[4 * q + 0] = fmaf(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7440u)), 8388608.0f), wt, a[4 * q + 0]);
      a[4 * q + 1] = fmaf(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7441u)), 8388608.0f), wt, a[4 * q + 1]);
      a[4 * q + 2] = fmaf(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7442u)), 8388608.0f), wt, a[4 * q + 2]);
      a[4 * q + 3] = fmaf(__fsub_rn(__uint_as_float(__byte_perm(v, 0x4B000000u, 0x7443u)), 8388608.0f), wt, a[4 * q + 3]);
    }
  }
  // round to nearest with the 1.5 * 2^23 magic add: n = rint(acc), frac = acc - n in [-1/2, 1/2]; floor(acc + 1/2) = n unless
  // acc sits on a tie, which is inside the replay band anyway.  0 <= acc < 255.5, so no clamping is needed.
  float dmax = 0.0f;
#pragma unroll
  for (int q = 0; q < NW; ++q) {
    uint32_t word = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float t = __fadd_rn(a[4 * q + e], 12582912.0f);
      const float n = __fsub_rn(t, 12582912.0f);
      dmax = fmaxf(dmax, fabsf(__fsub_rn(a[4 * q + e], n)));
      word |= (__float_as_uint(t) & 255u) << (8 * e);
    }
    out[q] = word;
  }
  amb = !(dmax <= 0.4995f);   // within 5e-4 of a rounding boundary of floor(acc + 1/2) (also true for NaN)
}

// literal evaluation of output byte `b` from the (unpadded) image with clamped coordinates.  The float32 additions are a
// dependent chain by definition; the loads and products of eight taps at a time are independent and issued together.
template <int CH>
__device__ __forceinline__ uint32_t line_blur_elem_exact(const uint8_t* im, const int* offs /* (width,2) dx,dy */, const float* k,
                                                         int width, int w, int h, int b) {
  const int px = b / CH, ch = b - px * CH, y = px / w, x = px - y * w;
  float a = 0.0f;
  for (int i0 = 0; i0 < width; i0 += 8) {
    float pr[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = min(i0 + u, width - 1);
      const int xx = min(max(x + offs[2 * i], 0), w - 1), yy = min(max(y + offs[2 * i + 1], 0), h - 1);
      pr[u] = __fmul_rn((float)im[(yy * w + xx) * CH + ch], k[i]);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (i0 + u < width) a = __fadd_rn(a, pr[u]);
  }
  return (uint32_t)min(max(__float2int_rd(__fadd_rn(a, 0.5f)), 0), 255);
}
template <int CH>
__device__ __noinline__ uint32_t line_blur_word_exact(const uint8_t* im, const int* offs, const float* k, int width, int w, int h,
                                                      int ob) {
  uint32_t out = 0;
  for (int e = 0; e < 4; ++e) out |= line_blur_elem_exact<CH>(im, offs, k, width, w, h, ob + e) << (8 * e);
  return out;
}

struct OpCcMotionFast : NoSpill {
  using Params = MotionParams;
  static constexpr bool kCifarCopy = true;
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 3;
  static constexpr int kStages = 2, kOutBufs = 1;   // compute-bound: a shallow ring frees shared memory for a third CTA
  static constexpr int kQueue = 255;
  static constexpr int kHeader = 128 + (kQueue + 1) * 4;   // weights f32[32] | replay queue
  // aux = n_eff: the canvas only has to cover the taps the fast path reads
  __host__ __device__ static size_t canvas_bytes(const Shape& s) { return align_up((size_t)(s.h + s.aux) * (s.w + s.aux + 6) * 3 + 8, 16); }
  // per image: LineGeom + raw offsets (2 x 32 int) + canvas
  __host__ __device__ static size_t per_image(const Shape& s) { return sizeof(LineGeom) + kMaxLine * 2 * 4 + canvas_bytes(s); }
  static size_t scratch_bytes(const Shape& s, int G) { return kHeader + (size_t)G * per_image(s) + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    float* k = reinterpret_cast<float*>(scratch);
    for (int i = threadIdx.x; i < kMaxLine; i += kThreads) k[i] = i < p.width ? p.k[i] : 0.0f;
    __syncthreads();
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const float* k = reinterpret_cast<const float*>(scratch);
    int* qn = reinterpret_cast<int*>(scratch + 128);
    int* queue = qn + 1;
    const size_t pi = per_image(s);
    const int tid = threadIdx.x, warp = tid >> 5;
    const Philox rng(p.seed);
    __syncthreads();
    if (tid == 0) *qn = 0;
    for (int j = warp; j < g; j += kThreads >> 5) {
      uint8_t* base = scratch + kHeader + (size_t)j * pi;
      LineGeom* gm = reinterpret_cast<LineGeom*>(base);
      int* raw = reinterpret_cast<int*>(base + sizeof(LineGeom));
      const int64_t img = img0 + j;
      if (p.offsets) {
        for (int i = tid & 31; i < p.width * 2; i += 32) raw[i] = p.offsets[img * p.width * 2 + i];
      } else {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x60u, 0u);
        motion_offsets(p.lo + (p.hi - p.lo) * (double)u01(r.x), p.width, raw);
      }
      __syncwarp();
      make_line_geom<3>(gm, raw, p.n_eff, s.w, s.h);
    }
    __syncthreads();
    for (int j = 0; j < g; ++j) {
      uint8_t* base = scratch + kHeader + (size_t)j * pi;
      fill_canvas<3>(base + sizeof(LineGeom) + kMaxLine * 2 * 4, src + (size_t)j * s.in_bytes,
                     *reinterpret_cast<const LineGeom*>(base), s.w, s.h, tid, kThreads);
    }
    __syncthreads();
    constexpr int NW = 3;                                  // 12 bytes = 4 pixels per thread (w % 4 == 0)
    const int ipi = s.in_bytes / (4 * NW), rowb = s.w * 3;
    uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
    for (int i = tid; i < g * ipi; i += kThreads) {
      const int j = i / ipi, ti = i - j * ipi;
      const uint8_t* base = scratch + kHeader + (size_t)j * pi;
      const LineGeom& gm = *reinterpret_cast<const LineGeom*>(base);
      const uint8_t* canvas = base + sizeof(LineGeom) + kMaxLine * 2 * 4;
      // canvas origin: output (0,0) sits at canvas (-dymin, -dxmin); tap offsets are relative to (dymin, dxmin)
      bool amb;
      uint32_t v[NW];
      line_blur_words_fast<3, NW>(canvas, gm, k, p.n_eff, rowb, ti * 4 * NW, v, amb);
      if (amb || p.force_exact) {
        const int slot = atomicAdd(qn, 1);
        if (slot < kQueue) queue[slot] = i;
        else
          for (int q = 0; q < NW; ++q)
            v[q] = line_blur_word_exact<3>(src + (size_t)j * s.in_bytes, reinterpret_cast<const int*>(base + sizeof(LineGeom)),
                                           k, p.width, s.w, s.h, (ti * NW + q) * 4);
      }
      uint32_t* o = d32 + (size_t)j * (s.out_bytes >> 2) + ti * NW;
#pragma unroll
      for (int q = 0; q < NW; ++q) o[q] = v[q];
    }
    __syncthreads();
    const int nq = min(*qn, kQueue);
    for (int e = tid; e < 4 * NW * nq; e += kThreads) {   // one thread per byte of the queued items
      const int i = queue[e / (4 * NW)], b = e % (4 * NW), j = i / ipi, ti = i - j * ipi;
      const uint8_t* base = scratch + kHeader + (size_t)j * pi;
      dst[(size_t)j * s.out_bytes + ti * 4 * NW + b] = (uint8_t)line_blur_elem_exact<3>(
          src + (size_t)j * s.in_bytes, reinterpret_cast<const int*>(base + sizeof(LineGeom)), k, p.width, s.w, s.h, ti * 4 * NW + b);
    }
  }
};

struct OpCcSnowFast : NoSpill {
  using Params = SnowParams;
  static constexpr bool kCifarCopy = true;
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 3;
  static constexpr int kStages = 2, kOutBufs = 1;   // compute-bound: a shallow ring frees shared memory for a third CTA
  // header: T32[256] | k[32] | ty[64] tx[64] | y0 y1 x0 x1 [64 each, int]        = 1024 + 128 + 512 + 1024
  static constexpr int kQueue = 255;
  static constexpr int kHeader = 1024 + 128 + 512 + 1024 + (kQueue + 1) * 4;   // ... | replay queue
  __host__ __device__ static size_t canvas_bytes(const Shape& s) { return align_up((size_t)(s.h + s.aux) * (s.w + s.aux + 6) + 8, 16); }  // aux = n_eff
  // per image: LineGeom + raw offsets + L0 f32[hw] + L1 u8[hw] + L2 u8[hw] + canvas
  __host__ __device__ static size_t per_image(const Shape& s) {
    return sizeof(LineGeom) + kMaxLine * 2 * 4 + align_up((size_t)s.h * s.w * 6, 16) + canvas_bytes(s);
  }
  static size_t scratch_bytes(const Shape& s, int G) { return kHeader + (size_t)G * per_image(s) + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape& s, const Params& p, uint8_t* scratch, bool) {
    float* T = reinterpret_cast<float*>(scratch);
    float* k = T + 256;
    float* ty = k + 32;
    float* tx = ty + 64;
    int* idx = reinterpret_cast<int*>(tx + 64);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = __fdiv_rn((float)i, 255.0f);
    for (int i = threadIdx.x; i < kMaxLine; i += kThreads) k[i] = i < p.m.width ? p.m.k[i] : 0.0f;
    for (int i = threadIdx.x; i < kSnowMax; i += kThreads) {
      ty[i] = p.ty[i]; tx[i] = p.tx[i];
      idx[i] = p.y0[i]; idx[64 + i] = p.y1[i]; idx[128 + i] = p.x0[i]; idx[192 + i] = p.x1[i];
    }
    __syncthreads();
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const float* T = reinterpret_cast<const float*>(scratch);
    const float* k = T + 256;
    const float* tyv = k + 32;
    const float* txv = tyv + 64;
    const int* idx = reinterpret_cast<const int*>(txv + 64);
    int* qn = const_cast<int*>(idx) + 256;
    int* queue = qn + 1;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int hw = s.h * s.w;
    const size_t pi = per_image(s);
    const size_t o_l0 = sizeof(LineGeom) + kMaxLine * 2 * 4, o_l1 = o_l0 + (size_t)hw * 4, o_l2 = o_l1 + hw,
                 o_cv = o_l0 + align_up((size_t)hw * 6, 16);
    const Philox rng(p.m.seed);
    __syncthreads();
    if (threadIdx.x == 0) *qn = 0;
    // A. line geometry, one warp per image
    for (int j = warp; j < g; j += kThreads >> 5) {
      uint8_t* base = scratch + kHeader + (size_t)j * pi;
      int* raw = reinterpret_cast<int*>(base + sizeof(LineGeom));
      const int64_t img = img0 + j;
      if (p.m.offsets) {
        for (int i = tid & 31; i < p.m.width * 2; i += 32) raw[i] = p.m.offsets[img * p.m.width * 2 + i];
      } else {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x62u, 0u);
        motion_offsets(p.m.lo + (p.m.hi - p.m.lo) * (double)u01(r.x), p.m.width, raw);
      }
      __syncwarp();
      make_line_geom<1>(reinterpret_cast<LineGeom*>(base), raw, p.m.n_eff, s.w, s.h);
    }
    // B. noise layer (f64 draws cast to f32)
    if (p.layer) {
      for (int i = tid; i < g * hw; i += kThreads) {
        const int j = i / hw, px = i - j * hw;
        reinterpret_cast<float*>(scratch + kHeader + (size_t)j * pi + o_l0)[px] = (float)p.layer[(img0 + j) * hw + px];
      }
    } else {
      for (int i = tid; i < (g * hw) >> 2; i += kThreads) {   // hw % 4 == 0 (checked at launch)
        const int j = (4 * i) / hw, px = 4 * i - j * hw;
        const uint64_t e = (uint64_t)((img0 + j) * hw + px) >> 2;
        const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x61u, 1u);
        float z0, z1, z2, z3;
        box_muller(r.x, r.y, z0, z1);
        box_muller(r.z, r.w, z2, z3);
        float* L0 = reinterpret_cast<float*>(scratch + kHeader + (size_t)j * pi + o_l0) + px;
        L0[0] = (float)(p.loc + p.scale * (double)z0); L0[1] = (float)(p.loc + p.scale * (double)z1);
        L0[2] = (float)(p.loc + p.scale * (double)z2); L0[3] = (float)(p.loc + p.scale * (double)z3);
      }
    }
    __syncthreads();
    // C. centre zoom (order 1), threshold, 8 bits
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const int y = px / s.w, x = px - y * s.w;
      const uint8_t* base = scratch + kHeader + (size_t)j * pi;
      const float* L = reinterpret_cast<const float*>(base + o_l0);
      const float ty = tyv[y], tx = txv[x], uy = __fsub_rn(1.0f, ty), ux = __fsub_rn(1.0f, tx);
      const int ya = idx[y] * s.w, yb = idx[64 + y] * s.w, xa = idx[128 + x], xb = idx[192 + x];
      float acc = __fmul_rn(__fmul_rn(L[ya + xa], uy), ux);
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L[ya + xb], uy), tx));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L[yb + xa], ty), ux));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L[yb + xb], ty), tx));
      if (acc < p.thresh) acc = 0.0f;
      const_cast<uint8_t*>(base)[o_l1 + px] = q255f(acc);
    }
    __syncthreads();
    // D. edge-replicated canvas of the 8-bit layer
    for (int j = 0; j < g; ++j) {
      uint8_t* base = scratch + kHeader + (size_t)j * pi;
      fill_canvas<1>(base + o_cv, base + o_l1, *reinterpret_cast<const LineGeom*>(base), s.w, s.h, tid, kThreads);
    }
    __syncthreads();
    // E. line blur, rounded back to 8 bits (float32 fused evaluation; words near a rounding boundary are replayed literally)
    constexpr int NW = 2;                                  // 8 layer bytes per thread (w % 8 == 0, else one word)
    const int ipl = hw / (4 * NW);
    for (int i = tid; i < g * ipl; i += kThreads) {
      const int j = i / ipl, ti = i - j * ipl;
      uint8_t* base = scratch + kHeader + (size_t)j * pi;
      bool amb;
      uint32_t v[NW];
      line_blur_words_fast<1, NW>(base + o_cv, *reinterpret_cast<const LineGeom*>(base), k, p.m.n_eff, s.w, ti * 4 * NW, v, amb);
      if (amb || p.m.force_exact) {
        const int slot = atomicAdd(qn, 1);
        if (slot < kQueue) queue[slot] = i;
        else
          for (int q = 0; q < NW; ++q)
            v[q] = line_blur_word_exact<1>(base + o_l1, reinterpret_cast<const int*>(base + sizeof(LineGeom)), k, p.m.width, s.w, s.h,
                                           (ti * NW + q) * 4);
      }
#pragma unroll
      for (int q = 0; q < NW; ++q) reinterpret_cast<uint32_t*>(base + o_l2)[ti * NW + q] = v[q];
    }
    __syncthreads();
    {
      const int nq = min(*qn, kQueue);
      for (int e = tid; e < 4 * NW * nq; e += kThreads) {   // one thread per byte of the queued items
        const int i = queue[e / (4 * NW)], b = e % (4 * NW), j = i / ipl, ti = i - j * ipl;
        uint8_t* base = scratch + kHeader + (size_t)j * pi;
        base[o_l2 + ti * 4 * NW + b] = (uint8_t)line_blur_elem_exact<1>(
            base + o_l1, reinterpret_cast<const int*>(base + sizeof(LineGeom)), k, p.m.width, s.w, s.h, ti * 4 * NW + b);
      }
    }
    __syncthreads();
    // F. blend with the whitened image, the layer and its 180-degree rotation
    for (int i = tid; i < g * hw; i += kThreads) {
      const int j = i / hw, px = i - j * hw;
      const uint8_t* L = scratch + kHeader + (size_t)j * pi + o_l2;
      const uint32_t rgb = load_rgb(src + (size_t)j * s.in_bytes, px * 3);
      const float sl = T[L[px]], sr = T[L[hw - 1 - px]];
      const float ch[3] = {T[rgb & 255u], T[(rgb >> 8) & 255u], T[(rgb >> 16) & 255u]};
      float gray = __fadd_rn(__fmul_rn(ch[0], 0.299f), __fmul_rn(ch[1], 0.587f));
      gray = __fadd_rn(gray, __fmul_rn(ch[2], 0.114f));
      const float whit = __fadd_rn(__fmul_rn(gray, 1.5f), 0.5f);
      uint8_t* o = dst + (size_t)j * s.out_bytes + px * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float mix = __fadd_rn(__fmul_rn(p.c6, ch[c]), __fmul_rn(p.one_minus_c6, fmaxf(ch[c], whit)));
        o[c] = q255f(__fadd_rn(__fadd_rn(mix, sl), sr));
      }
    }
  }
};

}  // namespace imgtx

// Images per CTA group for glass_blur.  The swap wavefront runs one warp per image, so the group size sets how many
// wavefront warps an SM holds (CTAs/SM x group): measured best 6 (R = 0: 445 -> 292 us), 4 (R = 1), 8 (R = 2: -6 %);
// IMGTX_GLASS_G overrides it for sweeps.
static int glass_group_images(int lw) {
  const int e = imgtx::staged_tuning().glass_g;
  if (e > 0) return e;
  return lw == 0 ? 6 : (lw == 1 ? 4 : 8);
}

using namespace imgtx;

namespace imgtx {
int zoom_fast(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, const int32_t* indptr, const int32_t* indices,
              const float* weights, int max_row_taps, uint32_t* metrics, cudaStream_t stream);   // cifar_c_zoom.cu
}

static int impl_cc_defocus_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ntaps,
                                   const int32_t* dydx_host, const double* weights_host, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 4 && w > 4 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  IMGTX_CHECK_ARG(ntaps >= 1 && ntaps <= kMaxTaps && dydx_host && weights_host, "1..%d taps", kMaxTaps);
  DefocusParams p;
  memset(&p, 0, sizeof(p));
  p.ntaps = ntaps;
  for (int i = 0; i < ntaps; ++i) {
    IMGTX_CHECK_ARG(abs(dydx_host[2 * i]) <= 2 && abs(dydx_host[2 * i + 1]) <= 2, "tap offset outside +-2");
    p.dy[i] = (int8_t)dydx_host[2 * i];
    p.dx[i] = (int8_t)dydx_host[2 * i + 1];
    p.w[i] = weights_host[i];
  }
  // mirror-symmetric kernel (K[dy][dx] == K[|dy|][|dx|], all weights >= 0) on a width that is a multiple of 4: fast form
  {
    double K[5][5];
    memset(K, 0, sizeof(K));
    int R = 0;
    for (int i = 0; i < ntaps; ++i) {
      K[p.dy[i] + 2][p.dx[i] + 2] += p.w[i];
      R = std::max(R, std::max(abs((int)p.dy[i]), abs((int)p.dx[i])));
    }
    if ((flags & IMGTX_CC_FAST) && h == 32 && w == 32) {
      // throughput form: the kernel restricted to 3x3 when what lies outside carries < 2e-3 LSB of weight
      bool ok = true;
      double outside = 0.0;
      for (int a = 0; a < 5; ++a)
        for (int b = 0; b < 5; ++b) {
          if (K[a][b] != K[4 - a][b] || K[a][b] != K[a][4 - b] || K[a][b] < 0.0) ok = false;
          if (a == 0 || a == 4 || b == 0 || b == 4) outside += K[a][b];
        }
      if (ok && outside * 255.0 < 2e-3) {
        Sym3x3 k3;
        const double tot = K[2][2] + 2.0 * (K[2][3] + K[3][2]) + 4.0 * K[3][3];   // renormalised: flat regions keep their value
        k3.k00 = (float)(K[2][2] / tot); k3.k01 = (float)(K[2][3] / tot); k3.k10 = (float)(K[3][2] / tot); k3.k11 = (float)(K[3][3] / tot);
        const int rc = launch_team_op<OpTDefocus>(in, out, n, make_shape(h, w, c, h, w), k3, (cudaStream_t)stream, metrics_dev);
        if (rc != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc; }
      }
    }
    bool sym = (w % 4 == 0) && R >= 1 && !(flags & IMGTX_CC_REPLAY_ALL);   // test knob: the literal kernel instead of the fixed-point sweep
    for (int a = 0; a < 5 && sym; ++a)
      for (int b = 0; b < 5; ++b)
        if (K[a][b] != K[4 - a][b] || K[a][b] != K[a][4 - b]) sym = false;
    if (sym) {
      DefocusFastParams f;
      memset(&f, 0, sizeof(f));
      f.exact = p;
      f.R = R;
      double cw[9];
      int mult[9];
      memset(cw, 0, sizeof(cw));
      for (int i = 0; i <= R; ++i)
        for (int k = 0; k <= R; ++k) {
          cw[i * (R + 1) + k] = K[2 + i][2 + k];
          mult[i * (R + 1) + k] = (i ? 2 : 1) * (k ? 2 : 1);
        }
      if (fix_weights40(cw, (R + 1) * (R + 1), mult, f.Wt)) {
        // 4 images per group (192 items = 192 threads), three CTAs per SM
        const Shape shp = make_shape(h, w, c, h, w);
        const int target = 128 + 3 * 4 * shp.in_bytes + 4096;
        if (R == 1) return launch_image_op<OpCcDefocusFast<1>>(in, out, n, shp, f, (cudaStream_t)stream, 0, target);
        return launch_image_op<OpCcDefocusFast<2>>(in, out, n, shp, f, (cudaStream_t)stream, 0, target);
      }
    }
  }
  return launch_image_op<OpCcDefocus>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream);
}

extern "C" int imgtx_cc_csr_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const int32_t* indptr_dev,
                               const int32_t* indices_dev, const float* weights_dev, int nnz, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  IMGTX_CHECK_ARG(indptr_dev && indices_dev && weights_dev && nnz > 0, "missing operator");
  CsrParams p;
  p.indptr = indptr_dev; p.indices = indices_dev; p.weights = weights_dev; p.nnz = nnz;
  const size_t extra = OpCcCsr::table_bytes(h * w, nnz);
  if (extra + 64 * 1024 > (size_t)max_smem_optin()) {
    if (n == 0) return IMGTX_OK;
    IMGTX_CHECK_ARG(in && out, "null image pointer");
    const int64_t want = (n * h * w + 255) / 256;
    const int64_t grid = std::min<int64_t>(want, (int64_t)sm_count() * 16);
    csr_global_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(in, out, n, h * w, p);
    IMGTX_LAUNCH_OK();
    return IMGTX_OK;
  }
  return launch_image_op<OpCcCsr>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream, extra,
                                  (int)extra + 64 * 1024);
}

static int impl_cc_csr_ell_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const int32_t* indptr_dev,
                                   const int32_t* indices_dev, const float* weights_dev, int nnz, int max_row_taps,
                                   int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  IMGTX_CHECK_ARG(indptr_dev && indices_dev && weights_dev && nnz > 0 && max_row_taps > 0, "missing operator");
  if (flags & IMGTX_CC_FAST) {
    const int rc = zoom_fast(in, out, n, h, w, indptr_dev, indices_dev, weights_dev, max_row_taps, metrics_dev, (cudaStream_t)stream);
    if (rc != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc; }
  }
  Shape shp = make_shape(h, w, c, h, w);
  const size_t tb = OpCcCsrEll::table_bytes(h * w, max_row_taps);
  shp.aux = (int)tb;
  int G = 6;   // float planes of G images + shallow ring next to the ELL table: one CTA of 32 warps per SM (6 measured 5-8 % faster than 4 where the table leaves room)
  auto need = [&](int g) { return 128 + 3 * (size_t)g * shp.in_bytes + OpCcCsrEll::scratch_bytes(shp, g) + 512; };
  while (G > 1 && need(G) > (size_t)max_smem_optin() - 1024) --G;
  if ((h * w) % 4 != 0 || h * w > 65535 || shp.in_bytes % 16 != 0 || !aligned16(in) || !aligned16(out) ||
      need(G) > (size_t)max_smem_optin() - 1024)
    return imgtx_cc_csr_u8(in, out, n, h, w, c, indptr_dev, indices_dev, weights_dev, nnz, stream);
  CsrEllParams p;
  p.csr.indptr = indptr_dev; p.csr.indices = indices_dev; p.csr.weights = weights_dev; p.csr.nnz = nnz;
  p.max_row = max_row_taps;
  return launch_image_op<OpCcCsrEll>(in, out, n, shp, p, (cudaStream_t)stream, 0, (int)need(G));
}

// glass_blur, throughput form for images larger than 32 x 32 (up to 65536 pixels; 224 x 224): one CTA per image.
//   1. Gaussian (3x3 float32, 'nearest' border; sigma <= 0.4: what lies outside carries < 5e-3 LSB) src -> B1 (workspace)
//   2. the swap sweeps on a u16 index map in SHARED memory (100 KB at 224 x 224): rows dealt to the threads, skew-2 wavefront,
//      one block barrier per step — the exact form does the same swaps on the pixels in global memory (~20x the latency)
//   3. gather B2[p] = B1[map[p]]      4. Gaussian B2 -> out.      sigma < 0.125: both Gaussians are one composed byte LUT.
// Same draw convention as OpCcGlass; <= 1 LSB from it (tests/test_gpu_cifar_c_large.py).
namespace imgtx {
__device__ __forceinline__ void glass_blur3_large(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int h, int w,
                                                  const Sym3x3& k) {
  // byte -> float without the conversion pipe: the byte planted in the mantissa of 2^23
  auto bf = [](uint8_t b) { return __uint_as_float(0x4B000000u | (uint32_t)b) - 8388608.0f; };
  for (int px = threadIdx.x; px < h * w; px += blockDim.x) {
    const int y = px / w, x = px - y * w;
    const int yu = max(y - 1, 0), yd = min(y + 1, h - 1), xl = max(x - 1, 0), xr = min(x + 1, w - 1);
    const uint8_t *ru = src + (size_t)yu * w * 3, *rc = src + (size_t)y * w * 3, *rd = src + (size_t)yd * w * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float ctr = bf(rc[x * 3 + c]);
      const float edge_h = bf(rc[xl * 3 + c]) + bf(rc[xr * 3 + c]);
      const float edge_v = bf(ru[x * 3 + c]) + bf(rd[x * 3 + c]);
      const float diag = (bf(ru[xl * 3 + c]) + bf(ru[xr * 3 + c])) + (bf(rd[xl * 3 + c]) + bf(rd[xr * 3 + c]));
      const float v = __fmaf_rn(k.k11, diag, __fmaf_rn(k.k10, edge_v, __fmaf_rn(k.k01, edge_h, __fmaf_rn(k.k00, ctr, 0.000244140625f))));
      dst[(size_t)px * 3 + c] = (uint8_t)(int)fminf(v, 255.0f);
    }
  }
}

__global__ void __launch_bounds__(256)
glass_large_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w, GlassTeamParams p,
                   uint8_t* __restrict__ work) {
  extern __shared__ __align__(16) uint8_t gsm[];
  uint8_t* lut = gsm;                                        // 256: composed byte map (lw == 0)
  uint16_t* map = reinterpret_cast<uint16_t*>(gsm + 256);   // h * w
  const int tid = threadIdx.x, hw = h * w, R = h - 2, Ccols = w - 2;
  uint8_t* B1 = work + (size_t)blockIdx.x * 2 * hw * 3;
  uint8_t* B2 = B1 + (size_t)hw * 3;
  if (p.lw == 0) {
    for (int i = tid; i < 256; i += 256) {
      const double t = (double)i / 255.0;
      const double a = (t * p.w0) * p.w0;
      const int u = (int)(a * 255.0);
      const double t2 = (double)u / 255.0;
      double b = (t2 * p.w0) * p.w0;
      b = b < 0.0 ? 0.0 : (b > 1.0 ? 1.0 : b);
      lut[i] = (uint8_t)(int)(b * 255.0);
    }
  }
  const Philox rng(p.seed);
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    const uint8_t* src = in + img * (int64_t)hw * 3;
    uint8_t* dst = out + img * (int64_t)hw * 3;
    __syncthreads();
    if (p.lw > 0) glass_blur3_large(src, B1, h, w, p.k);
    for (int i = tid; i < hw; i += 256) map[i] = (uint16_t)i;
    __syncthreads();
    for (int it = 0; it < p.iters; ++it) {
      uint4 rb = make_uint4(0u, 0u, 0u, 0u);
      int rb_key = -1;
      for (int t = 0; t < 2 * (R - 1) + Ccols; ++t) {
        for (int r = tid; r < R; r += 256) {
          const int ci = t - 2 * r;
          if (ci < 0 || ci >= Ccols) continue;
          int dx, dy;
          if (p.deltas) {
            const int8_t* d = p.deltas + ((((img * p.iters + it) * R + r) * Ccols + ci) << 1);
            dx = d[0]; dy = d[1];
          } else {
            const int key = r | ((ci >> 6) << 16);
            if (key != rb_key) {
              rb = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)r + ((uint32_t)(ci >> 6) << 16));
              rb_key = key;
            }
            const int cb = ci & 63;
            dx = -(int)(((cb < 32 ? rb.x : rb.z) >> (cb & 31)) & 1u);
            dy = -(int)(((cb < 32 ? rb.y : rb.w) >> (cb & 31)) & 1u);
          }
          const int a = (h - 1 - r) * w + (w - 1 - ci), b = a + dy * w + dx;
          const uint16_t ta = map[a], tb = map[b];
          map[a] = tb; map[b] = ta;
        }
        __syncthreads();
      }
    }
    if (p.lw > 0) {
      for (int i = tid; i < hw; i += 256) {
        const uint8_t* q = B1 + (size_t)map[i] * 3;
        B2[(size_t)i * 3] = q[0]; B2[(size_t)i * 3 + 1] = q[1]; B2[(size_t)i * 3 + 2] = q[2];
      }
      __syncthreads();
      glass_blur3_large(B2, dst, h, w, p.k);
    } else {
      for (int i = tid; i < hw; i += 256) {
        const uint8_t* q = src + (size_t)map[i] * 3;
        dst[(size_t)i * 3] = lut[q[0]]; dst[(size_t)i * 3 + 1] = lut[q[1]]; dst[(size_t)i * 3 + 2] = lut[q[2]];
      }
    }
  }
}
}  // namespace imgtx

static int glass_large_fast(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, const imgtx::GlassTeamParams& f,
                            cudaStream_t st) {
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null image pointer");
  const size_t smem = 256 + (size_t)h * w * 2 + 16;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(imgtx::glass_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)imgtx::smem_per_sm() / (smem + 1024)));
  const int64_t grid = std::min<int64_t>(n, (int64_t)imgtx::sm_count() * per_sm);
  uint8_t* work = nullptr;
  IMGTX_CUDA_OK(imgtx::workspace_alloc((void**)&work, (size_t)grid * 2 * h * w * 3, st));
  imgtx::glass_large_kernel<<<(unsigned)grid, 256, smem, st>>>(in, out, n, h, w, f, work);
  IMGTX_LAUNCH_OK();
  IMGTX_CUDA_OK(cudaFreeAsync(work, st));
  return IMGTX_OK;
}

static int impl_cc_glass_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const double* weights_host,
                                 int lw, int iters, const int8_t* deltas_dev, uint64_t seed, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 2 && w > 2 && c == 3 && n >= 0, "glass_blur takes (n,h>2,w>2,3) images");
  IMGTX_CHECK_ARG(lw >= 0 && lw <= 3 && weights_host && iters >= 0, "bad Gaussian");
  GlassParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i <= lw; ++i) p.w[i] = weights_host[i];
  p.lw = lw; p.iters = iters; p.deltas = deltas_dev; p.seed = seed;
  const Shape shp = make_shape(h, w, c, h, w);
  const bool cifar = h == 32 && w == 32;
  if ((flags & IMGTX_CC_FAST) && lw <= 2 && (cifar || (h * w <= 65536 && (size_t)h * w * 2 + 2048 <= (size_t)max_smem_optin()))) {
    GlassTeamParams f;
    memset(&f, 0, sizeof(f));
    f.lw = lw; f.iters = iters; f.deltas = deltas_dev; f.seed = seed; f.w0 = p.w[0]; f.n = n;
    static const int kSweeper = getenv("IMGTX_GLASS_SWEEPER") ? atoi(getenv("IMGTX_GLASS_SWEEPER")) : 1;   // tuning, read once
    f.sweeper = kSweeper;
    const double inside = (p.w[0] + 2.0 * p.w[1]) * (p.w[0] + 2.0 * p.w[1]);   // the 3x3 part, renormalised to sum to 1
    f.k.k00 = (float)(p.w[0] * p.w[0] / inside); f.k.k01 = f.k.k10 = (float)(p.w[0] * p.w[1] / inside);
    f.k.k11 = (float)(p.w[1] * p.w[1] / inside);
    if (lw == 0 || (1.0 - inside) * 255.0 < 5e-3) {   // what the 3x3 form drops
      if (!cifar) return glass_large_fast(in, out, n, h, w, f, (cudaStream_t)stream);
      const int rc = launch_team_op<OpTGlass>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev);
      if (rc != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc; }
    }
  }
  if (lw <= 2 && h <= 34 && w % 4 == 0 && w - 2 <= 64 && h * w <= 65536 && shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out) &&
      !(flags & IMGTX_CC_REPLAY_ALL)) {   // test knob: the literal kernel instead of the fixed-point sweeps
    GlassFastParams f;
    memset(&f, 0, sizeof(f));
    f.g = p;
    bool ok = true;
    if (lw >= 1) {
      double cw[9];
      int mult[9];
      for (int i = 0; i <= lw; ++i)
        for (int k = 0; k <= lw; ++k) {
          cw[i * (lw + 1) + k] = p.w[i] * p.w[k];
          mult[i * (lw + 1) + k] = (i ? 2 : 1) * (k ? 2 : 1);
        }
      ok = fix_weights(cw, (lw + 1) * (lw + 1), mult, f.Wt);
    }
    if (ok) {
      const int gi = glass_group_images(lw);
      const int target = 128 + (lw == 1 ? 5 : 3) * gi * shp.in_bytes + gi * h * w * 2 + 4096;  // gi images per group
      cudaStream_t st = (cudaStream_t)stream;
      if (lw == 0) return launch_image_op<OpCcGlassFast<0>>(in, out, n, shp, f, st, 0, target);
      if (lw == 1) return launch_image_op<OpCcGlassFast<1>>(in, out, n, shp, f, st, 0, target);
      return launch_image_op<OpCcGlassFast<2>>(in, out, n, shp, f, st, 0, target);
    }
  }
  return launch_image_op<OpCcGlass>(in, out, n, shp, p, (cudaStream_t)stream);
}

static int fill_motion(MotionParams& m, int width, const float* k_host, const int32_t* offsets_dev, double lo, double hi,
                       uint64_t seed, int flags) {
  IMGTX_CHECK_ARG(width >= 1 && width <= kMaxLine && k_host, "line kernel width must be in [1,%d]", kMaxLine);
  memset(&m, 0, sizeof(m));
  m.width = width;
  for (int i = 0; i < width; ++i) m.k[i] = k_host[i];
  // taps the throughput forms evaluate: drop the tail whose total weight is below 2^-20 LSB (|k| * 255 summed)
  m.n_eff = width;
  double tail = 0.0;
  while (m.n_eff > 1 && tail + fabs((double)k_host[m.n_eff - 1]) * 255.0 <= 9.5e-7) tail += fabs((double)k_host[--m.n_eff]) * 255.0;
  m.offsets = offsets_dev; m.lo = lo; m.hi = hi; m.seed = seed;
  m.force_exact = (flags & IMGTX_CC_REPLAY_ALL) ? 1 : 0;
  return IMGTX_OK;
}

// Throughput form: taps whose total weight is below 1e-3 LSB are dropped; the canvas is padded by the largest row offset the
// declared angle range can produce.
static void fill_motion_fast(MotionFastParams& f, const MotionParams& m, double lo, double hi) {
  f.width = m.width;
  f.n_eff = m.width;
  double tail = 0.0;
  while (f.n_eff > 1 && tail + fabs((double)m.k[f.n_eff - 1]) * 255.0 <= 1e-3) tail += fabs((double)m.k[--f.n_eff]) * 255.0;
  for (int i = 0; i < f.n_eff; ++i) f.k[i] = m.k[i];
  double smax = std::max(fabs(sin(lo * 0.017453292519943295)), fabs(sin(hi * 0.017453292519943295)));
  if (lo <= -90.0 && hi >= -90.0) smax = 1.0;
  if (lo <= 90.0 && hi >= 90.0) smax = 1.0;
  f.padv = std::min(f.n_eff - 1, (int)ceil((f.n_eff - 1) * smax + 0.5));
  f.offsets = m.offsets; f.lo = (float)lo; f.hi = (float)hi; f.seed = m.seed;
}

static int impl_cc_motion_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int width,
                                  const float* k_host, const int32_t* offsets_dev, double angle_lo, double angle_hi,
                                  uint64_t seed, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  MotionParams p;
  int rc = fill_motion(p, width, k_host, offsets_dev, angle_lo, angle_hi, seed, flags);
  if (rc) return rc;
  Shape shp = make_shape(h, w, c, h, w);
  if ((flags & IMGTX_CC_FAST) && h == 32 && w == 32) {
    MotionFastParams f;
    memset(&f, 0, sizeof(f));
    fill_motion_fast(f, p, angle_lo, angle_hi);
    const int rc2 = launch_team_op<OpTMotion>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev);
    if (rc2 != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc2; }
  }
  shp.aux = p.n_eff;
  if (w % 4 == 0 && shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out) &&
      OpCcMotionFast::scratch_bytes(shp, 1) + 6 * (size_t)shp.in_bytes < (size_t)max_smem_optin()) {
    const int target = 128 + 3 * 4 * shp.in_bytes + (int)OpCcMotionFast::scratch_bytes(shp, 4) + 512;  // 4 images per group
    return launch_image_op<OpCcMotionFast>(in, out, n, shp, p, (cudaStream_t)stream, 0, std::min(target, max_smem_optin() - 2048));
  }
  return launch_image_op<OpCcMotion>(in, out, n, shp, p, (cudaStream_t)stream);
}

// snow, throughput form for large images: one CTA of 1024 threads per image, the three layers (float32 noise, its zoomed and
// thresholded 8-bit form, the blurred 8-bit form: 6 bytes per pixel) in a workspace.  Up to the 8-bit layer the arithmetic is
// the exact form's, operation for operation (a threshold decision that flipped would move a byte by far more than 1 LSB);
// the line blur and the blend are float32 FFMA on the 0..255 scale.
namespace imgtx {
__global__ void __launch_bounds__(1024)
snow_large_fast_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w, SnowParams p,
                       uint8_t* __restrict__ work) {
  __shared__ int off[kMaxLine * 2];
  const int tid = threadIdx.x, nt = blockDim.x, hw = h * w;
  uint8_t* base = work + (size_t)blockIdx.x * align_up((size_t)hw * 6, 256);
  float* L0 = reinterpret_cast<float*>(base);
  uint8_t* L1 = base + (size_t)hw * 4;
  uint8_t* L2 = L1 + hw;
  const Philox rng(p.m.seed);
  const int width = p.m.n_eff;
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    __syncthreads();
    if (p.layer) {
      for (int i = tid; i < hw; i += nt) L0[i] = (float)p.layer[img * hw + i];
    } else {
      const float loc = (float)p.loc, sc = (float)p.scale;
      for (int i = tid; 2 * i < hw; i += nt) {
        const uint64_t e = (uint64_t)img * (uint64_t)((hw + 1) / 2) + (uint64_t)i;
        const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x61u, 0u);
        float z0, z1;
        box_muller_fast(r.x, r.y, z0, z1);
        L0[2 * i] = __fmaf_rn(sc, z0, loc);
        if (2 * i + 1 < hw) L0[2 * i + 1] = __fmaf_rn(sc, z1, loc);
      }
    }
    if (tid < 32) {
      if (p.m.offsets) {
        for (int k = tid; k < p.m.width * 2; k += 32) off[k] = p.m.offsets[img * p.m.width * 2 + k];
      } else {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x62u, 0u);
        motion_offsets(p.m.lo + (p.m.hi - p.m.lo) * (double)u01(r.x), p.m.width, off);
      }
    }
    __syncthreads();
    for (int i = tid; i < hw; i += nt) {   // centre zoom (order 1), threshold, 8 bits: the exact form's float32 operations
      const int y = i / w, x = i - y * w;
      const int zy0 = p.zidx[y], zy1 = p.zidx[h + y], zx0 = p.zidx[2 * h + x], zx1 = p.zidx[2 * h + w + x];
      const float ty = p.zt[y], tx = p.zt[h + x], uy = __fsub_rn(1.0f, ty), ux = __fsub_rn(1.0f, tx);
      float acc = __fmul_rn(__fmul_rn(L0[zy0 * w + zx0], uy), ux);
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L0[zy0 * w + zx1], uy), tx));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L0[zy1 * w + zx0], ty), ux));
      acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L0[zy1 * w + zx1], ty), tx));
      if (acc < p.thresh) acc = 0.0f;
      L1[i] = q255f(acc);
    }
    __syncthreads();
    for (int i = tid; i < hw; i += nt) {   // line blur of the 8-bit layer
      const int y = i / w, x = i - y * w;
      float acc = 0.5f;
      for (int k = 0; k < width; ++k) {
        const int xx = min(max(x + off[2 * k], 0), w - 1), yy = min(max(y + off[2 * k + 1], 0), h - 1);
        acc = __fmaf_rn(__uint_as_float(0x4B000000u | (uint32_t)L1[yy * w + xx]) - 8388608.0f, p.m.k[k], acc);
      }
      L2[i] = (uint8_t)(int)fminf(acc, 255.0f);
    }
    __syncthreads();
    const uint8_t* im = in + img * (int64_t)hw * 3;
    uint8_t* om = out + img * (int64_t)hw * 3;
    const float c6 = p.c6, d6 = p.one_minus_c6;
    for (int i = tid; i < hw; i += nt) {   // blend on the 0..255 scale
      const float r = (float)im[i * 3], g = (float)im[i * 3 + 1], b = (float)im[i * 3 + 2];
      const float whit = __fmaf_rn(__fmaf_rn(b, 0.114f, __fmaf_rn(g, 0.587f, r * 0.299f)), 1.5f, 127.5f);
      const float sn = (float)L2[i] + (float)L2[hw - 1 - i] + 0.0009765625f;
      const float v0 = __fmaf_rn(d6, fmaxf(r, whit), __fmaf_rn(c6, r, sn));
      const float v1 = __fmaf_rn(d6, fmaxf(g, whit), __fmaf_rn(c6, g, sn));
      const float v2 = __fmaf_rn(d6, fmaxf(b, whit), __fmaf_rn(c6, b, sn));
      om[i * 3] = (uint8_t)(int)fminf(fmaxf(v0, 0.0f), 255.0f);
      om[i * 3 + 1] = (uint8_t)(int)fminf(fmaxf(v1, 0.0f), 255.0f);
      om[i * 3 + 2] = (uint8_t)(int)fminf(fmaxf(v2, 0.0f), 255.0f);
    }
  }
}
}  // namespace imgtx

static int impl_cc_snow_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, float thresh, float c6,
                                float one_minus_c6, const int32_t* zoom_idx_host /* y0,y1 (h each), x0,x1 (w each) */,
                                const float* zoom_t_host /* ty (h), tx (w) */, int width, const float* k_host,
                                const int32_t* offsets_dev, const double* layer_dev, double loc, double scale,
                                double angle_lo, double angle_hi, uint64_t seed, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "snow takes (n,h,w,3) images");
  IMGTX_CHECK_ARG(zoom_idx_host && zoom_t_host, "missing zoom tables");
  SnowParams p;
  memset(&p, 0, sizeof(p));
  int rc = fill_motion(p.m, width, k_host, offsets_dev, angle_lo, angle_hi, seed, flags);
  if (rc) return rc;
  p.thresh = thresh; p.c6 = c6; p.one_minus_c6 = one_minus_c6; p.loc = loc; p.scale = scale; p.layer = layer_dev;
  if (h > kSnowMax || w > kSnowMax) {
    // large images: literal form, one CTA per image, layers in a global workspace, zoom tables in device memory
    if (n == 0) return IMGTX_OK;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* tab = nullptr;
    const size_t ib = (size_t)(2 * h + 2 * w) * 4, tb = (size_t)(h + w) * 4;
    IMGTX_CUDA_OK(workspace_alloc((void**)&tab, ib + tb, st));
    IMGTX_CUDA_OK(cudaMemcpyAsync(tab, zoom_idx_host, ib, cudaMemcpyHostToDevice, st));
    IMGTX_CUDA_OK(cudaMemcpyAsync(tab + ib, zoom_t_host, tb, cudaMemcpyHostToDevice, st));
    p.zidx = reinterpret_cast<const int32_t*>(tab);
    p.zt = reinterpret_cast<const float*>(tab + ib);
    if (flags & IMGTX_CC_FAST) {
      const int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * 2);
      uint8_t* work = nullptr;
      IMGTX_CUDA_OK(workspace_alloc((void**)&work, (size_t)grid * align_up((size_t)h * w * 6, 256), st));
      snow_large_fast_kernel<<<(unsigned)grid, 1024, 0, st>>>(in, out, n, h, w, p, work);
      IMGTX_LAUNCH_OK();
      IMGTX_CUDA_OK(cudaFreeAsync(work, st));
      rc = IMGTX_OK;
    } else {
      rc = launch_image_op<OpCcSnow>(in, out, n, make_shape(h, w, c, h, w), p, st);
    }
    IMGTX_CUDA_OK(cudaFreeAsync(tab, st));
    return rc;
  }
  for (int i = 0; i < h; ++i) { p.y0[i] = (int8_t)zoom_idx_host[i]; p.y1[i] = (int8_t)zoom_idx_host[h + i]; p.ty[i] = zoom_t_host[i]; }
  for (int i = 0; i < w; ++i) { p.x0[i] = (int8_t)zoom_idx_host[2 * h + i]; p.x1[i] = (int8_t)zoom_idx_host[2 * h + w + i]; p.tx[i] = zoom_t_host[h + i]; }
  Shape shp = make_shape(h, w, c, h, w);
  if ((flags & IMGTX_CC_FAST) && h == 32 && w == 32) {
    SnowFastParams f;
    memset(&f, 0, sizeof(f));
    fill_motion_fast(f.m, p.m, angle_lo, angle_hi);
    // rows the line can reach: dy_i = ceil(i*sin(a) - 1/2) over the declared angle range
    double smin = std::min(sin(angle_lo * 0.017453292519943295), sin(angle_hi * 0.017453292519943295)),
           smax = std::max(sin(angle_lo * 0.017453292519943295), sin(angle_hi * 0.017453292519943295));
    if (angle_lo <= -90.0 && angle_hi >= -90.0) smin = -1.0;
    if (angle_lo <= 90.0 && angle_hi >= 90.0) smax = 1.0;
    f.pad_top = std::min(f.m.n_eff - 1, std::max(0, (int)ceil(-(f.m.n_eff - 1) * smin + 0.5)));
    f.pad_bot = std::min(f.m.n_eff - 1, std::max(0, (int)ceil((f.m.n_eff - 1) * smax + 0.5)));
    f.thresh = thresh; f.c6 = c6; f.one_minus_c6 = one_minus_c6; f.loc = (float)loc; f.scale = (float)scale; f.layer = layer_dev;
    for (int i = 0; i < 32; ++i) {
      f.y0[i] = p.y0[i]; f.y1[i] = p.y1[i]; f.x0[i] = p.x0[i]; f.x1[i] = p.x1[i]; f.ty[i] = p.ty[i]; f.tx[i] = p.tx[i];
    }
    const int rc2 = launch_team_op<OpTSnow>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev);
    if (rc2 != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc2; }
  }
  shp.aux = p.m.n_eff;
  if (w % 8 == 0 && shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out) &&
      OpCcSnowFast::scratch_bytes(shp, 1) + 6 * (size_t)shp.in_bytes < (size_t)max_smem_optin()) {
    const int target = 128 + 3 * 4 * shp.in_bytes + (int)OpCcSnowFast::scratch_bytes(shp, 4) + 512;  // 4 images per group
    return launch_image_op<OpCcSnowFast>(in, out, n, shp, p, (cudaStream_t)stream, 0, std::min(target, max_smem_optin() - 2048));
  }
  return launch_image_op<OpCcSnow>(in, out, n, shp, p, (cudaStream_t)stream);
}

extern "C" int imgtx_cc_defocus_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ntaps,
                                   const int32_t* dydx_host, const double* weights_host, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_defocus_u8(in, out, n, h, w, c, ntaps, dydx_host, weights_host, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}

extern "C" int imgtx_cc_csr_ell_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const int32_t* indptr_dev,
                                   const int32_t* indices_dev, const float* weights_dev, int nnz, int max_row_taps,
                                   int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_csr_ell_u8(in, out, n, h, w, c, indptr_dev, indices_dev, weights_dev, nnz, max_row_taps, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}

extern "C" int imgtx_cc_glass_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const double* weights_host,
                                 int lw, int iters, const int8_t* deltas_dev, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_glass_u8(in, out, n, h, w, c, weights_host, lw, iters, deltas_dev, seed, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}

extern "C" int imgtx_cc_motion_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int width,
                                  const float* k_host, const int32_t* offsets_dev, double angle_lo, double angle_hi,
                                  uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_motion_u8(in, out, n, h, w, c, width, k_host, offsets_dev, angle_lo, angle_hi, seed, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}

extern "C" int imgtx_cc_snow_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, float thresh, float c6,
                                float one_minus_c6, const int32_t* zoom_idx_host /* y0,y1 (h each), x0,x1 (w each) */,
                                const float* zoom_t_host /* ty (h), tx (w) */, int width, const float* k_host,
                                const int32_t* offsets_dev, const double* layer_dev, double loc, double scale,
                                double angle_lo, double angle_hi, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_snow_u8(in, out, n, h, w, c, thresh, c6, one_minus_c6, zoom_idx_host, zoom_t_host, width, k_host, offsets_dev, layer_dev, loc, scale, angle_lo, angle_hi, seed, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}
