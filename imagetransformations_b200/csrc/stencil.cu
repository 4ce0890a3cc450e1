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
#include "sep_mma.cuh"
#include "enhance_op.cuh"
#include "rng.cuh"
#include <math.h>
#include <mutex>
#include <vector>
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
// The same blur with four multiply-adds per instruction (small images, every Q8 weight <= 255, radius < min(h, w)):
//   * the image is de-interleaved into per-channel byte rows with the REFLECT_101 border materialised (a canvas whose
//     left pad is a multiple of four), so a run of taps is a run of bytes;
//   * a thread produces four adjacent outputs; output i needs the canvas bytes from 4q + (pad - r) + i on, i.e. the
//     SAME aligned words for every thread with the kernel pre-shifted by (pad - r + i) bytes: four packed weight rows,
//     one __dp4a per word and output;
//   * the u16 intermediate (<= 65280) is stored transposed as a high-byte and a low-byte plane, border rows reflected,
//     so the vertical pass is the same routine on each plane: v = (sum_hi << 8) + sum_lo, out = (v + 2^15) >> 16.
// Integer arithmetic throughout: identical to the literal kernel above (tests compare both with the oracle).
struct OpCvGaussDp : NoSpill {
  using Params = CvGaussParams;
  static constexpr int kThreads = 256;
  static constexpr bool kCifarCopy = true;
  static constexpr int kHeader = 4 * 12 * 4;   // Wt[4][12]
  __host__ __device__ static int pad_of(int r) { return (r + 3) & ~3; }
  __host__ __device__ static int row_bytes(int len, int r) {
    int b = len + 2 * pad_of(r);
    return (b & 31) == 0 ? b + 4 : b;          // keep the word stride off a multiple of 8 banks
  }
  __host__ __device__ static size_t per_image(const Shape& s) {
    const int r = s.aux;
    return (size_t)s.c * s.h * row_bytes(s.w, r) + 2 * (size_t)s.c * s.w * row_bytes(s.h, r);
  }
  // +64: the zero-weight tail words of the last column may lie past the last plane
  static size_t scratch_bytes(const Shape& s, int G) { return align_up(kHeader + (size_t)G * per_image(s) + 64, 16); }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }

  __device__ static void prepare(const Shape& s, const Params& p, uint8_t* scratch, bool) {
    uint32_t* Wt = reinterpret_cast<uint32_t*>(scratch);
    const int r = s.aux, n = p.ksize, t0 = pad_of(r) - r;
    for (int e = threadIdx.x; e < 48; e += kThreads) {
      const int i = e / 12, m = e - i * 12;
      uint32_t wv = 0;
      for (int lane = 0; lane < 4; ++lane) {
        const int j = 4 * m + lane - (t0 + i);
        if (j >= 0 && j < n) wv |= (uint32_t)p.k[j] << (8 * lane);
      }
      Wt[e] = wv;
    }
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const int tid = threadIdx.x;
    const int c = s.c, w = s.w, h = s.h, n = p.ksize, r = s.aux;
    const int PL = pad_of(r), W2 = row_bytes(w, r), H2 = row_bytes(h, r);
    const int MW = ((PL - r + 3 + n - 1) >> 2) + 1;
    const uint32_t* Wt = reinterpret_cast<const uint32_t*>(scratch);
    const size_t per = per_image(s);
    const int wq = w >> 2, hq = h >> 2, W2w = W2 >> 2;
    const FastDiv dW2w(W2w), dh(h), dc(c), dwq(wq), dhq(hq), dw(w), dr(r), dcol(2 * c * w);
    // ---- canvas: planar rows with reflected borders ----
    if (c == 3) {
      // interior: four RGB pixels = three aligned words -> one word per channel plane (byte permutes)
      for (int it = tid; it < g * h * wq; it += kThreads) {
        int t, q, y, j;
        dwq.divmod(it, t, q);
        dh.divmod(t, j, y);
        const uint32_t* sp = reinterpret_cast<const uint32_t*>(src + (size_t)j * s.in_bytes + (size_t)y * w * 3) + q * 3;
        const uint32_t w0 = sp[0], w1 = sp[1], w2 = sp[2];
        uint32_t* cv = reinterpret_cast<uint32_t*>(scratch + kHeader + j * per) + y * W2w + (PL >> 2) + q;
        cv[0] = __byte_perm(__byte_perm(w0, w1, 0x6630), w2, 0x5210);            // bytes 0, 3, 6, 9
        cv[h * W2w] = __byte_perm(__byte_perm(w0, w1, 0x7741), w2, 0x6210);      // bytes 1, 4, 7, 10
        cv[2 * h * W2w] = __byte_perm(__byte_perm(w0, w1, 0x5552), w2, 0x7410);  // bytes 2, 5, 8, 11
      }
    }
    {
      // border words (all words when c != 3): byte gathers through the reflection
      const int PLw = PL >> 2, edge = c == 3 ? 2 * PLw : W2w;
      const FastDiv de(edge);
      for (int it = tid; it < g * c * h * edge; it += kThreads) {
        int t, e, y, ch, j;
        de.divmod(it, t, e);
        dh.divmod(t, t, y);
        dc.divmod(t, j, ch);
        const int xw = c == 3 ? (e < PLw ? e : e - PLw + PLw + wq) : e;
        const uint8_t* sj = src + (size_t)j * s.in_bytes + (size_t)y * w * c + ch;
        uint32_t word = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int cx = 4 * xw + k - PL;
          const int rx = cx < 0 ? -cx : (cx >= w ? 2 * w - 2 - cx : cx);     // r < w: one reflection
          if (cx >= -r && cx < w + r) word |= (uint32_t)sj[rx * c] << (8 * k);
        }
        if (xw < W2w) reinterpret_cast<uint32_t*>(scratch + kHeader + j * per)[(ch * h + y) * W2w + xw] = word;
      }
    }
    __syncthreads();
    // ---- horizontal: four outputs per item, transposed hi / lo planes out ----
    const size_t plane = (size_t)c * w * H2;
    for (int it = tid; it < g * c * h * wq; it += kThreads) {
      int t, q, y, ch, j;
      dwq.divmod(it, t, q);
      dh.divmod(t, t, y);
      dc.divmod(t, j, ch);
      uint8_t* base = scratch + kHeader + j * per;
      const uint32_t* row = reinterpret_cast<const uint32_t*>(base) + (ch * h + y) * W2w + q;
      uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
      for (int m = 0; m < MW; ++m) {
        const uint32_t v = row[m];
        a0 = __dp4a(v, Wt[m], a0); a1 = __dp4a(v, Wt[12 + m], a1);
        a2 = __dp4a(v, Wt[24 + m], a2); a3 = __dp4a(v, Wt[36 + m], a3);
      }
      uint8_t* th = base + (size_t)c * h * W2 + (size_t)(ch * w + 4 * q) * H2 + PL + y;
      uint8_t* tl = th + plane;
      th[0] = (uint8_t)(a0 >> 8); th[H2] = (uint8_t)(a1 >> 8); th[2 * H2] = (uint8_t)(a2 >> 8); th[3 * H2] = (uint8_t)(a3 >> 8);
      tl[0] = (uint8_t)a0; tl[H2] = (uint8_t)a1; tl[2 * H2] = (uint8_t)a2; tl[3 * H2] = (uint8_t)a3;
    }
    __syncthreads();
    // ---- reflect the border rows of both planes ----
    for (int it = tid; it < g * 2 * c * w * r; it += kThreads) {
      int t, d, col, j;
      dr.divmod(it, t, d);
      dcol.divmod(t, j, col);
      uint8_t* colp = scratch + kHeader + j * per + (size_t)c * h * W2 + (size_t)col * H2 + PL;
      colp[-1 - d] = colp[1 + d];
      colp[h + d] = colp[h - 2 - d];
    }
    __syncthreads();
    // ---- vertical: the same routine down each column of both planes ----
    for (int it = tid; it < g * c * w * hq; it += kThreads) {
      int t, q, x, ch, j;
      dhq.divmod(it, t, q);
      dw.divmod(t, t, x);
      dc.divmod(t, j, ch);
      const uint8_t* base = scratch + kHeader + j * per + (size_t)c * h * W2;
      const uint32_t* ch_ = reinterpret_cast<const uint32_t*>(base + (size_t)(ch * w + x) * H2) + q;
      const uint32_t* cl_ = reinterpret_cast<const uint32_t*>(base + plane + (size_t)(ch * w + x) * H2) + q;
      uint32_t hi[4] = {0, 0, 0, 0}, lo[4] = {0, 0, 0, 0};
      for (int m = 0; m < MW; ++m) {
        const uint32_t vh = ch_[m], vl = cl_[m];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t wv = Wt[12 * i + m];
          hi[i] = __dp4a(vh, wv, hi[i]);
          lo[i] = __dp4a(vl, wv, lo[i]);
        }
      }
      uint8_t* o = dst + (size_t)j * s.out_bytes + ((size_t)(4 * q) * w + x) * c + ch;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t v = (hi[i] << 8) + lo[i] + (1u << 15);
        o[(size_t)i * w * c] = (uint8_t)min(v >> 16, 255u);
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

// ---------------------------------------------------------------------------------------------
// TransformationPool chains (pipenline/cifar_image_transformations.py:141-152 applied to the population of
// :203-226): tensor_to_pil -> up to max_ops TransformationPool ops -> pil_to_tensor as ONE launch.  One CTA owns one
// individual: the uint8 image never leaves shared memory between the ops, every op is the operator the single-op entry
// points run (OpPilGauss, OpEnhance, OpBoxH and the noise formulas of noise.cu), so the result is the composition of
// those, byte for byte.  Random draws are injected per op (the reference's numpy draws: parity) or made on chip (Philox
// keyed by image, op position and element; shot noise draws Poisson(rate) from the image it finds, which is what removes
// the host round trip of the op-by-op path).
struct PoolOp {  // == imgtx_pool_op_t
  int32_t kind, iparam;
  float fparam;
  int32_t draw;   // index of this op's block in its draw array, < 0: on-chip draws
};
struct PoolChainParams {
  float mean[3], std[3];
  PilGaussParams blur[5];          // defocus_blur severities (:72-77: GaussianBlur radius 3, 4, 6, 8, 10)
  const PoolOp* ops;               // device (n, max_ops)
  int max_ops;
  const double* gauss;             // device draw arrays (blocks of h*w*3, h*w, h*w*3 values)
  const double* impulse;
  const long long* shot;
  uint64_t seed;
  int64_t x_stride;                // floats between input images; 0: every individual starts from the same image
};
// the reference's severity tables (:43, :54, :65)
__constant__ double kPoolGaussStd[5] = {0.08, 0.12, 0.18, 0.26, 0.38};
__constant__ double kPoolImpulse[5] = {0.03, 0.06, 0.09, 0.17, 0.27};
__constant__ int kPoolShot[5] = {60, 25, 12, 5, 3};
// On-chip shot noise samples its Poisson count by inverting the CDF with one 32-bit uniform: per severity and input byte b,
// cdf[k] = floor(P(K <= k) * 2^32) for k = 0 .. lam - 1 with K ~ Poisson(f32(b) / 255 * lam) (every k >= lam clips to 255),
// built on the host in float64 and uploaded once per device (fill_pool_shot_cdf).  Offsets of the five tables below.
constexpr int kPoolShotRows = 60 + 25 + 12 + 5 + 3;
static __device__ uint32_t g_pool_shot_cdf[kPoolShotRows * 256];   // [severity offset + k][b]
__constant__ int kPoolShotOff[5] = {0, 60, 85, 97, 102};

__global__ void __launch_bounds__(256)
pool_chain_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t n, Shape s, PoolChainParams p, int buf_bytes) {
  extern __shared__ __align__(16) uint8_t psm[];
  uint8_t* cur = psm;
  uint8_t* oth = psm + buf_bytes;
  uint8_t* scratch = psm + 2 * buf_bytes;                         // max(OpEnhance scratch, one image for OpPilGauss)
  PoolOp* ops = reinterpret_cast<PoolOp*>(scratch + (buf_bytes > OpEnhance::kPerImage + 16 ? buf_bytes : OpEnhance::kPerImage + 16));
  const int tid = threadIdx.x, hw = s.h * s.w, ne = hw * 3;
  const Philox rng(p.seed);
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    __syncthreads();   // the previous image is out of the buffers
    for (int i = tid; i < p.max_ops; i += 256) ops[i] = p.ops[img * p.max_ops + i];
    // tensor_to_pil (:27-30): x * std + mean, clamp [0, 1], * 255, truncate
    const float* xi = x + img * p.x_stride;
    for (int i = tid; i < ne; i += 256) {
      const int ch = i / hw, px = i - ch * hw;
      float v = __fadd_rn(__fmul_rn(xi[i], p.std[ch]), p.mean[ch]);
      v = fminf(fmaxf(v, 0.0f), 1.0f);
      cur[px * 3 + ch] = (uint8_t)(__float2int_rz(__fmul_rn(v, 255.0f)) & 255);
    }
    __syncthreads();
    for (int k = 0; k < p.max_ops; ++k) {
      const PoolOp op = ops[k];
      if (op.kind < 0) break;
      const int sev = min(max(op.iparam, 1), 5) - 1;
      bool swapped = true;
      switch (op.kind) {
        case IMGTX_POOL_GAUSSIAN_NOISE: {   // :39-48  clip(f32(p) + normal(0, std*255), 0, 255) in float64, truncate
          if (op.draw >= 0) {
            const double* d = p.gauss + (int64_t)op.draw * ne;
            for (int i = tid; i < ne; i += 256) {
              double v = __dadd_rn((double)(float)cur[i], d[i]);
              v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);
              cur[i] = (uint8_t)(int)v;
            }
          } else {
            const float sd = (float)(kPoolGaussStd[sev] * 255.0);
            for (int i = tid; 2 * i < ne; i += 256) {
              const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x90u + k, (uint32_t)i);
              float z[2];
              box_muller_fast(r.x, r.y, z[0], z[1]);
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                if (2 * i + e >= ne) break;
                float v = (float)cur[2 * i + e] + z[e] * sd;
                v = v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v);
                cur[2 * i + e] = (uint8_t)__float2int_rz(v);
              }
            }
          }
          swapped = false;
          break;
        }
        case IMGTX_POOL_IMPULSE_NOISE: {    // :50-59  one uniform per pixel: < p/2 -> 0, > 1 - p/2 -> 255
          const double pr = kPoolImpulse[sev];
          const double* d = op.draw >= 0 ? p.impulse + (int64_t)op.draw * hw : nullptr;
          for (int px = tid; px < hw; px += 256) {
            double m;
            if (d) m = d[px];
            else m = (double)u01(rng((uint32_t)img, (uint32_t)(img >> 32), 0x90u + k, (uint32_t)px).x);
            if (m < pr / 2) { cur[px * 3] = 0; cur[px * 3 + 1] = 0; cur[px * 3 + 2] = 0; }
            if (m > 1 - pr / 2) { cur[px * 3] = 255; cur[px * 3 + 1] = 255; cur[px * 3 + 2] = 255; }
          }
          swapped = false;
          break;
        }
        case IMGTX_POOL_SHOT_NOISE: {       // :61-70  k ~ Poisson(f32(p) / 255 * lam);  clip(k / lam * 255, 0, 255)
          const int lam = kPoolShot[sev];
          const long long* d = op.draw >= 0 ? p.shot + (int64_t)op.draw * ne : nullptr;
          // count -> byte once per op (the float64 division per element costs more than everything else here): counts of
          // lam or more clip to 255
          uint8_t* klut = scratch;
          if (tid < lam) klut[tid] = (uint8_t)(int)__dmul_rn(__ddiv_rn((double)tid, (double)lam), 255.0);
          __syncthreads();
          for (int i = tid; i < ne; i += 256) {
            long long cnt;
            if (d) cnt = d[i];
            else {
              // one Philox call serves four elements; count = number of CDF entries the uniform has passed (binary search)
              const uint4 r4 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x90u + k, (uint32_t)(i >> 2));
              const uint32_t r = (i & 3) == 0 ? r4.x : ((i & 3) == 1 ? r4.y : ((i & 3) == 2 ? r4.z : r4.w));
              const uint32_t* cdf = g_pool_shot_cdf + kPoolShotOff[sev] * 256 + cur[i];
              int lo = 0, hi = lam;            // first k in [0, lam] with r < cdf[k]  (cdf[lam] = 2^32)
              while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (r < cdf[mid * 256]) hi = mid; else lo = mid + 1;
              }
              cnt = lo;
            }
            cur[i] = cnt >= lam ? (uint8_t)255 : (cnt < 0 ? (uint8_t)0 : klut[cnt]);
          }
          swapped = false;
          break;
        }
        case IMGTX_POOL_DEFOCUS_BLUR:       // :72-77
          OpPilGauss::run<true>(cur, oth, 1, 0, s, p.blur[sev], scratch, nullptr);
          break;
        case IMGTX_POOL_MOTION_BLUR: {      // :109-119
          BoxHParams b;
          b.size = max(op.iparam, 1);
          b.kv = (float)(1.0 / b.size);
          OpBoxH::run<true>(cur, oth, 1, 0, s, b, scratch, nullptr);
          break;
        }
        default: {                          // ImageEnhance ops :81-106, histogram_equalization :122-129
          EnhanceParams e;
          e.kind = op.kind == IMGTX_POOL_ENHANCE_CONTRAST ? IMGTX_ENH_CONTRAST
                   : op.kind == IMGTX_POOL_ENHANCE_BRIGHTNESS ? IMGTX_ENH_BRIGHTNESS
                   : op.kind == IMGTX_POOL_ENHANCE_SHARPNESS ? IMGTX_ENH_SHARPNESS
                   : op.kind == IMGTX_POOL_ENHANCE_COLOR ? IMGTX_ENH_COLOR : IMGTX_ENH_HISTEQ_YUV;
          e.factor = op.fparam;
          e.factor_per_image = nullptr;
          if ((ne & 3) == 0) OpEnhance::run<true>(cur, oth, 1, 0, s, e, scratch, nullptr);
          else OpEnhance::run<false>(cur, oth, 1, 0, s, e, scratch, nullptr);
          break;
        }
      }
      __syncthreads();
      if (swapped) { uint8_t* t = cur; cur = oth; oth = t; }
    }
    // pil_to_tensor (:32-34): ToTensor, Normalize
    float* oi = out + img * (int64_t)ne;
    for (int i = tid; i < ne; i += 256) {
      const int ch = i / hw, px = i - ch * hw;
      const float t = __fdiv_rn((float)cur[px * 3 + ch], 255.0f);
      oi[i] = __fdiv_rn(__fsub_rn(t, p.mean[ch]), p.std[ch]);
    }
  }
}

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

// ---------------------------------------------------------------------------------------------
// 32x32x3 images: with the REFLECT_101 border folded in, the Q8 kernel is a 32x32 byte operator per axis and both passes
// run as integer MMAs in registers (sep_mma.cuh), whatever the tap count.  Valid when every folded weight is <= 255 and
// the weights sum to 256 (checked on the host); same integers as the literal loop.
struct OpCvGaussMma {
  using Params = CvGaussParams;
  static constexpr int kTeamWarps = 1, kInBufs = 3, kOutBufs = 0, kMaxTeams = 12, kCtasPerSm = 1;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  static size_t cta_scratch(const Shape&, const Params&) { return kSepOpBytes + 2 * kSepFragBytes; }
  __host__ __device__ static void fold_row(const Params& p, int o, int* row /* 32 */) {
    const int r = p.ksize >> 1;
    for (int i = 0; i < 32; ++i) row[i] = 0;
    for (int j = 0; j < p.ksize; ++j) {
      int i = o + j - r;
      while (i < 0 || i >= 32) i = i < 0 ? -i : 62 - i;   // REFLECT_101
      row[i] += p.k[j];
    }
  }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    int32_t* op = reinterpret_cast<int32_t*>(cta);
    uint4* frag = reinterpret_cast<uint4*>(cta + kSepOpBytes);
    for (int o = threadIdx.x; o < 32; o += blockDim.x) fold_row(p, o, op + o * 32);
    __syncthreads();
    sep_build_frags(op, frag, 0, 1, false);
    sep_build_frags(op, frag + 64, 0, 1, true);
  }
  __device__ static void run(const uint8_t* in, uint8_t* out, int64_t, const Shape&, const Params&, uint8_t*,
                             const uint8_t* cta, int tt, int) {
    sep_cvq8_chain(in, out, reinterpret_cast<const uint4*>(cta + kSepOpBytes), tt);
  }
};

static bool cv_mma_ok(const CvGaussParams& p) {
  int sum = 0;
  for (int j = 0; j < p.ksize; ++j) {
    if (p.k[j] < 0) return false;
    sum += p.k[j];
  }
  if (sum != 256 || (p.ksize >> 1) > 31) return false;
  int row[32];
  for (int o = 0; o < 32; ++o) {
    OpCvGaussMma::fold_row(p, o, row);
    for (int i = 0; i < 32; ++i)
      if (row[i] > 255) return false;
  }
  return true;
}

extern "C" int imgtx_cv_gaussian_blur_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ksize,
                                         double sigma, int flags, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4, "bad image shape");
  CvGaussParams p;
  memset(&p, 0, sizeof(p));
  p.ksize = ksize;
  int rc = imgtx_cv_gaussian_kernel_host(ksize, sigma, p.k);
  if (rc) return rc;
  Shape s = make_shape(h, w, c, h, w);
  const int r = ksize >> 1;
  bool small_weights = true;
  for (int i = 0; i < ksize; ++i) small_weights = small_weights && p.k[i] >= 0 && p.k[i] <= 255;
  s.aux = r;
  if (!(flags & IMGTX_CC_REPLAY_ALL) && h == 32 && w == 32 && c == 3 && r >= 1 && cv_mma_ok(p)) {
    const int rc2 = launch_team_op<OpCvGaussMma>(in, out, n, s, p, (cudaStream_t)stream);
    if (rc2 != IMGTX_ERR_UNSUPPORTED) return rc2;
  }
  if (!(flags & IMGTX_CC_REPLAY_ALL) && small_weights && r >= 1 && r < std::min(h, w) && (w & 3) == 0 && (h & 3) == 0 &&
      OpCvGaussDp::per_image(s) <= 40 * 1024 && (OpCvGaussDp::row_bytes(h, r) & 3) == 0 && (OpCvGaussDp::row_bytes(w, r) & 3) == 0)
    return launch_image_op<OpCvGaussDp>(in, out, n, s, p, (cudaStream_t)stream);
  s.aux = 0;
  return launch_image_op<OpCvGauss>(in, out, n, s, p, (cudaStream_t)stream);
}

// libImaging/BoxBlur.c _gaussian_blur_radius (passes = 3), C float / double mix reproduced literally
static PilGaussParams pil_gauss_params(float radius) {
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
  return p;
}

extern "C" int imgtx_pil_gaussian_blur_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                          float radius, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4, "bad image shape");
  IMGTX_CHECK_ARG(radius > 0, "radius must be positive");
  const PilGaussParams p = pil_gauss_params(radius);
  return launch_image_op<OpPilGauss>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream);
}

// Poisson CDF tables of the on-chip shot noise, uploaded once per device (stream-ordered before the first launch using them).
static int fill_pool_shot_cdf(cudaStream_t stream) {
  static std::mutex mu;
  constexpr int kPoolMaxDevices = 64;
  static bool done[kPoolMaxDevices] = {false};
  int dev = -1;
  IMGTX_CUDA_OK(cudaGetDevice(&dev));
  IMGTX_CHECK_ARG(dev >= 0 && dev < kPoolMaxDevices, "device ordinal %d outside the table cache", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (done[dev]) return IMGTX_OK;
  static std::vector<uint32_t> host;
  if (host.empty()) {
    host.resize((size_t)kPoolShotRows * 256);
    const int lams[5] = {60, 25, 12, 5, 3};
    int off = 0;
    for (int sv = 0; sv < 5; ++sv) {
      for (int b = 0; b < 256; ++b) {
        // the reference's rate: float32 array / 255.0 * lam stays float32 (pipenline/cifar_image_transformations.py:66)
        volatile float t = (float)b / 255.0f;
        volatile float rf = t * (float)lams[sv];
        const double r = (double)rf;
        double pk = exp(-r), cum = 0.0;
        for (int k = 0; k < lams[sv]; ++k) {
          cum += pk;
          const double v = floor(std::min(cum, 1.0) * 4294967296.0);
          host[(size_t)(off + k) * 256 + b] = v >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)v;
          pk = pk * r / (double)(k + 1);
        }
      }
      off += lams[sv];
    }
  }
  IMGTX_CUDA_OK(cudaMemcpyToSymbolAsync(g_pool_shot_cdf, host.data(), host.size() * 4, 0, cudaMemcpyHostToDevice, stream));
  IMGTX_CUDA_OK(cudaStreamSynchronize(stream));   // once per device: later launches on any stream may use the table
  done[dev] = true;
  return IMGTX_OK;
}

extern "C" int imgtx_pool_chain_f32(const float* x, int64_t x_image_stride, float* out, int64_t n, int h, int w,
                                    const imgtx_pool_op_t* ops_dev, int max_ops, const float* mean_host,
                                    const float* std_host, const double* gauss_draws_dev, const double* impulse_draws_dev,
                                    const int64_t* shot_draws_dev, uint64_t seed, imgtx_stream_t stream) {
  static_assert(sizeof(imgtx_pool_op_t) == sizeof(PoolOp) && sizeof(PoolOp) == 16, "op record layout");
  IMGTX_CHECK_ARG(h > 0 && w > 0 && n >= 0, "bad image shape");
  IMGTX_CHECK_ARG(max_ops >= 1 && max_ops <= 64, "max_ops must be in [1,64]");
  IMGTX_CHECK_ARG(x_image_stride == 0 || x_image_stride >= (int64_t)3 * h * w, "input images overlap");
  IMGTX_CHECK_ARG(mean_host && std_host, "mean / std missing");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(x && out && ops_dev, "null pointer");
  const Shape s = make_shape(h, w, 3, h, w);
  const int buf = (int)align_up((size_t)s.in_bytes + 4, 16);
  const size_t smem = 2 * (size_t)buf + std::max<size_t>(buf, OpEnhance::kPerImage + 16) + (size_t)max_ops * sizeof(PoolOp) + 16;
  IMGTX_CHECK_ARG(smem <= (size_t)max_smem_optin(), "image %dx%d does not fit in shared memory for the fused chain "
                  "(three copies are kept on chip); apply the ops one by one", h, w);
  PoolChainParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < 3; ++i) { p.mean[i] = mean_host[i]; p.std[i] = std_host[i]; }
  const float radii[5] = {3, 4, 6, 8, 10};
  for (int i = 0; i < 5; ++i) p.blur[i] = pil_gauss_params(radii[i]);
  p.ops = reinterpret_cast<const PoolOp*>(ops_dev); p.max_ops = max_ops;
  p.gauss = gauss_draws_dev; p.impulse = impulse_draws_dev; p.shot = reinterpret_cast<const long long*>(shot_draws_dev);
  p.seed = seed; p.x_stride = x_image_stride;
  const int rc_tab = fill_pool_shot_cdf((cudaStream_t)stream);
  if (rc_tab) return rc_tab;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(pool_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)smem_per_sm() / (smem + 1024)));
  const int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * per_sm);
  pool_chain_kernel<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(x, out, n, s, p, buf);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
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

// ---------------------------------------------------------------------------------------------
// imgtx_cv_gaussian_blur_noalloc_u8: the literal cv2.GaussianBlur path (OpCvGauss through direct_kernel, u16 intermediate
// in global memory) for callers that may not allocate: CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  No counterpart in the reference.  The u16 intermediate planes
// (2*h*w*c bytes per resident CTA) live in the file-scope device array below instead of a cudaMallocAsync workspace.
//   * The array is kNoAllocScratchBytes (32 MiB) on EVERY GPU that loads the library, until the process exits.
//   * Nothing in it has to start at zero: the row pass writes every element of a CTA's slice before the column pass reads
//     it, so the previous call's contents are never observed.
//   * Every call on a device shares the array: calls that overlap on the GPU (different streams, or graph branches that
//     run side by side) race.  Serialise them on one stream.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress,
// looked up once per device with that device current.
constexpr size_t kNoAllocScratchBytes = 32u << 20;
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) uint8_t g_noalloc_scratch[kNoAllocScratchBytes];

#include <mutex>
static int noalloc_scratch_ptr(uint8_t** ptr) {
  static std::mutex mu;
  static uint8_t* cache[kNoAllocMaxDevices] = {nullptr};
  int dev = -1;
  IMGTX_CUDA_OK(cudaGetDevice(&dev));
  IMGTX_CHECK_ARG(dev >= 0 && dev < kNoAllocMaxDevices, "device ordinal %d outside the scratch table (%d)", dev, kNoAllocMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!cache[dev]) {
    void* p = nullptr;
    IMGTX_CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));   // the current device's instance of the array
    IMGTX_CHECK_ARG(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    cache[dev] = static_cast<uint8_t*>(p);
  }
  *ptr = cache[dev];
  return IMGTX_OK;
}

extern "C" int64_t imgtx_noalloc_scratch_bytes(void) { return (int64_t)kNoAllocScratchBytes; }

extern "C" int imgtx_cv_gaussian_blur_noalloc_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ksize,
                                                 double sigma, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4 && n >= 0, "bad image shape");
  CvGaussParams p;
  memset(&p, 0, sizeof(p));
  p.ksize = ksize;
  int rc = imgtx_cv_gaussian_kernel_host(ksize, sigma, p.k);
  if (rc) return rc;
  const Shape s = make_shape(h, w, c, h, w);
  const size_t per_cta = align_up(OpCvGauss::gtmp_bytes(s), 256);
  if (per_cta > kNoAllocScratchBytes) {
    set_error("image %dx%dx%d needs %zu bytes of scratch for its 16-bit intermediate; the allocation-free entry point "
              "holds %zu bytes (use imgtx_cv_gaussian_blur_u8 for larger images)", h, w, c, per_cta, kNoAllocScratchBytes);
    return IMGTX_ERR_INVALID_ARGUMENT;
  }
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in != nullptr && out != nullptr, "null image pointer");
  uint8_t* scratch = nullptr;
  rc = noalloc_scratch_ptr(&scratch);
  if (rc) return rc;
  int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * 8);
  grid = std::min<int64_t>(grid, (int64_t)(kNoAllocScratchBytes / per_cta));
  const size_t smem = OpCvGauss::scratch_bytes_spilled(s);
  direct_kernel<OpCvGauss><<<(unsigned)grid, OpCvGauss::kThreads, smem, (cudaStream_t)stream>>>(in, out, n, s, p, scratch, per_cta);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
