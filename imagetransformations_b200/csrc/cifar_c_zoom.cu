// zoom_blur, throughput form (IMGTX_CC_FAST of imgtx_cc_csr_ell_u8).
//
// zoom_blur is ONE fixed sparse linear operator per severity (oracle/cifar_c.py:zoom_operator: 5.5-12.3 taps per output
// pixel on average, 7-19 at most, inside an 8x8 window).  The exact form (cifar_c_stencil.cu:OpCcCsrEll) keeps the operator in
// shared memory and is bound by shared-memory traffic: 18 bytes per tap and pixel (index + weight + three floats).  Here
//   * a thread OWNS one output pixel for the whole launch and keeps its taps (byte offset, float weight) in REGISTERS: the
//     operator is read from global memory once per CTA, never again;
//   * a warp is one image row, so its lanes read neighbouring canvas pixels (conflict-free 8-byte loads) and run the
//     longest tap count of that row, rounded up to a multiple of 4 (taps are issued four loads at a time);
//   * the staged image is expanded once into an fp16 RGBx canvas (u8 is exact in fp16): a tap is ONE 8-byte shared-memory
//     load, three conversions and three FFMA (float32 accumulation on the 0..255 scale);
//   * persistent CTAs (one per SM, one thread per pixel), images arrive by 1-D bulk TMA into a 3-deep ring, results leave by bulk
//     TMA stores; the canvas of image i+1 is built while image i is evaluated, so there is one CTA barrier per image.
// Arithmetic: sum_k w_k * x_k in float32 FFMA order instead of the oracle's ordered float32 sum of rounded products on the
// 0..1 scale: |difference| < 1e-4 LSB, i.e. <= 1 LSB after floor (tests/test_gpu_fast_mode.py).
#include "sep_mma.cuh"
#include <cuda_fp16.h>
#include <type_traits>

namespace imgtx {

constexpr int kZoomThreads = 1024;
constexpr int kZoomRing = 6;
constexpr int kZoomCanvas = 1024 * 8;   // bytes per canvas (fixed pitch: the second canvas is an immediate offset away)

// 8-byte shared-memory load at a register address plus a compile-time offset
template <int IMM>
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2+%3];" : "=r"(v.x), "=r"(v.y) : "r"(addr), "n"(IMM));
  return v;
}

template <int T>
__global__ void __launch_bounds__(kZoomThreads, 1)
zoom_fast_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int P /* pixels, <= 1024, % 4 == 0 */,
                 const int32_t* __restrict__ indptr, const int32_t* __restrict__ indices, const float* __restrict__ weights,
                 uint2* __restrict__ metrics) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int img_bytes = P * 3, ib = (int)align_up((size_t)img_bytes, 128);
  constexpr int cb = kZoomCanvas;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint2* msum = reinterpret_cast<uint2*>(smem + 256);        // per-warp metric partials (2 parities x 32 x 8 B)
  uint8_t* in_buf = smem + 768;
  uint8_t* canvas = in_buf + kZoomRing * ib;                   // 2 x P x half4
  uint8_t* out_buf = canvas + 2 * cb;                          // 2 x ib
  const int tid = threadIdx.x;

  if (tid == 0) {
    for (int i = 0; i < kZoomRing; ++i) mbar_init(&full[i], 1);
    fence_mbar_init();
  }
  __syncthreads();

  // my pixel and its taps (padding taps read canvas pixel 0 with weight 0)
  const int pix = tid < P ? tid : -1;
  int off[T];
  float wt[T];
  int cnt = 0;
  {
    const int k0 = pix >= 0 ? indptr[pix] : 0;
    cnt = pix >= 0 ? min(indptr[pix + 1] - k0, T) : 0;
#pragma unroll
    for (int k = 0; k < T; ++k) {
      const bool ok = k < cnt;
      off[k] = (int)smem_u32(canvas) + (ok ? indices[k0 + k] * 8 : 0);   // absolute shared address in canvas 0
      wt[k] = ok ? weights[k0 + k] : 0.0f;
    }
  }
  int wmax = cnt;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));

  const int64_t stride = gridDim.x, first = blockIdx.x;
  if (tid == 0) {
    for (int i = 0; i < kZoomRing; ++i) {
      const int64_t img = first + (int64_t)i * stride;
      if (img < n) {
        mbar_expect_tx(&full[i], (uint32_t)img_bytes);
        bulk_g2s(in_buf + i * ib, in + img * img_bytes, (uint32_t)img_bytes, &full[i]);
      }
    }
  }
  // u8 HWC -> half4 canvas, 4 pixels (3 words) per thread-step
  auto build = [&](int it) {
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(in_buf + (it % kZoomRing) * ib);
    uint2* cv = reinterpret_cast<uint2*>(canvas + (it & 1) * cb);
    const __half2 k1024 = __float2half2_rn(1024.0f);
    for (int q = tid; q < (P >> 2); q += kZoomThreads) {
      const uint32_t w0 = s32[3 * q], w1 = s32[3 * q + 1], w2 = s32[3 * q + 2];
      // half(1024 + b) has the bit pattern 0x6400 | b: plant two bytes, subtract 1024 (exact)
      auto h2 = [&](uint32_t a, uint32_t b, uint32_t sel) {
        const uint32_t v = __byte_perm(__byte_perm(a, b, sel), 0x64646464u, 0x7160u);   // (b0, 0x64, b1, 0x64)
        const __half2 r = __hsub2(*reinterpret_cast<const __half2*>(&v), k1024);
        return *reinterpret_cast<const uint32_t*>(&r);
      };
      // pixels: (w0.b0 b1 b2) (w0.b3 w1.b0 b1) (w1.b2 b3 w2.b0) (w2.b1 b2 b3); second half2 of a pixel = (blue, 0)
      const uint32_t zero_hi = 0xFFFF0000u;
      cv[4 * q + 0] = make_uint2(h2(w0, w0, 0x0010u), h2(w0, w0, 0x0002u) & ~zero_hi);
      cv[4 * q + 1] = make_uint2(h2(w0, w1, 0x0043u), h2(w1, w1, 0x0001u) & ~zero_hi);
      cv[4 * q + 2] = make_uint2(h2(w1, w1, 0x0032u), h2(w2, w2, 0x0000u) & ~zero_hi);
      cv[4 * q + 3] = make_uint2(h2(w2, w2, 0x0021u), h2(w2, w2, 0x0003u) & ~zero_hi);
    }
  };

  int64_t img = first;
  if (img < n) {
    mbar_wait(&full[0], 0u);
    build(0);
  }
  __syncthreads();
  // one image: build the next canvas, evaluate this one.  PAR (canvas / output buffer parity) is a compile-time constant so
  // that `canvas + PAR * cb + off[k]` is one register plus an immediate: no address arithmetic per tap.
  auto step = [&](auto par, int it, int64_t img) {
    constexpr int PAR = decltype(par)::value;
    if (img + stride < n) {
      mbar_wait(&full[(it + 1) % kZoomRing], (uint32_t)((it + 1) / kZoomRing) & 1u);
      build(it + 1);
    }
    uint8_t* ob = out_buf + PAR * ib;
    float a0 = 0.0009765625f, a1 = a0, a2 = a0;   // 2^-10: an exactly integral sum floors to itself
#pragma unroll
    for (int k0 = 0; k0 < T; k0 += 4) {
      if (k0 < wmax) {
        uint2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = lds64<PAR * cb>((uint32_t)off[k0 + u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&v[u].x));
          const float b = __low2float(*reinterpret_cast<const __half2*>(&v[u].y));
          a0 = __fmaf_rn(rg.x, wt[k0 + u], a0);
          a1 = __fmaf_rn(rg.y, wt[k0 + u], a1);
          a2 = __fmaf_rn(b, wt[k0 + u], a2);
        }
      }
    }
    uint32_t b0 = 0, b1 = 0, b2 = 0;
    if (pix >= 0) {
      uint8_t* o = ob + pix * 3;
      b0 = floor_bits(fminf(fmaxf(a0, 0.0f), 255.0f)) & 255u;
      b1 = floor_bits(fminf(fmaxf(a1, 0.0f), 255.0f)) & 255u;
      b2 = floor_bits(fminf(fmaxf(a2, 0.0f), 255.0f)) & 255u;
      o[0] = (uint8_t)b0; o[1] = (uint8_t)b1; o[2] = (uint8_t)b2;
    }
    if (metrics) {   // per-image metric (see team_pipeline.cuh:checksum_word) from the bytes this thread just produced
      uint32_t c0 = b0 + b1 + b2, c1 = (3u * (uint32_t)tid + 1u) * c0 + b1 + 2u * b2;
      c0 = __reduce_add_sync(0xffffffffu, c0);
      c1 = __reduce_add_sync(0xffffffffu, c1);
      if ((tid & 31) == 0) msum[PAR * 32 + (tid >> 5)] = make_uint2(c0, c1);
    }
    fence_async_smem();
    if (tid == 0) bulk_wait_read<0>();   // the store of the previous image has finished reading the other output buffer
    __syncthreads();
    if (metrics && (tid >> 5) == 1) {   // warp 1: warp 0 has the copies to issue; the partials are double-buffered by parity
      const uint2 m = msum[PAR * 32 + (tid & 31)];
      const uint32_t mx = __reduce_add_sync(0xffffffffu, m.x), my = __reduce_add_sync(0xffffffffu, m.y);
      if ((tid & 31) == 0) metrics[img] = make_uint2(mx, my);
    }
    if (tid == 0) {
      bulk_s2g(out + img * img_bytes, ob, (uint32_t)img_bytes);
      bulk_commit();
      const int64_t nxt = img + (int64_t)kZoomRing * stride;
      if (nxt < n) {
        const int st = it % kZoomRing;
        mbar_expect_tx(&full[st], (uint32_t)img_bytes);
        bulk_g2s(in_buf + st * ib, in + nxt * img_bytes, (uint32_t)img_bytes, &full[st]);
      }
    }
  };
  for (int it = 0; img < n; it += 2) {
    step(std::integral_constant<int, 0>{}, it, img);
    img += stride;
    if (img < n) step(std::integral_constant<int, 1>{}, it + 1, img);
    img += stride;
  }
  if (tid == 0) bulk_wait_all<0>();
}

template <int T>
static int launch_zoom_fast(const uint8_t* in, uint8_t* out, int64_t n, int P, const int32_t* indptr, const int32_t* indices,
                            const float* weights, uint32_t* metrics, cudaStream_t stream) {
  const size_t ib = align_up((size_t)P * 3, 128);
  const size_t smem = 768 + kZoomRing * ib + 2 * (size_t)kZoomCanvas + 2 * ib;
  auto kern = zoom_fast_kernel<T>;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t grid = std::min<int64_t>(n, sm_count());
  kern<<<(unsigned)grid, kZoomThreads, smem, stream>>>(in, out, n, P, indptr, indices, weights, reinterpret_cast<uint2*>(metrics));
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

// IMGTX_ERR_UNSUPPORTED when the shape is outside the fast form (the caller falls back to the exact ELL kernel)
int zoom_fast(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, const int32_t* indptr, const int32_t* indices,
              const float* weights, int max_row_taps, uint32_t* metrics, cudaStream_t stream) {
  const int P = h * w;
  if (P > 1024 || P % 4 != 0 || max_row_taps > 20 || (P * 3) % 16 != 0 || !aligned16(in) || !aligned16(out))
    return IMGTX_ERR_UNSUPPORTED;
  if (n <= 0) return IMGTX_OK;
  if (max_row_taps <= 8) return launch_zoom_fast<8>(in, out, n, P, indptr, indices, weights, metrics, stream);
  if (max_row_taps <= 12) return launch_zoom_fast<12>(in, out, n, P, indptr, indices, weights, metrics, stream);
  if (max_row_taps <= 16) return launch_zoom_fast<16>(in, out, n, P, indptr, indices, weights, metrics, stream);
  return launch_zoom_fast<20>(in, out, n, P, indptr, indices, weights, metrics, stream);
}


// ---------------------------------------------------------------------------------------------------------------
// zoom_blur on images too large for the register-resident operator (224 x 224: 0.7-2.2 M merged taps), throughput form.
// The merged operator is what the exact form streams through L2 (17 MB per image at severity 5); here the sum is evaluated
// level by level instead, out = (x + sum_z bilinear_z(x)) / (nz + 1), from per-axis tables (i0, i1, t per zoom level and
// output index: oracle/cifar_c.py:zoom_axis_table) that are tiny.  A CTA owns a 32 x 32 output tile: it copies the tables of
// its rows and columns and the source window all levels touch (the tile shrunk towards the centre, at most ~80 x 80) into
// shared memory, then every thread accumulates its four pixels in float32 on the 0..255 scale.
// Same real-valued expression as the merged operator (whose weights are these products summed in float64): <= 1 LSB.
struct ZoomLevelsParams {
  const int32_t* yidx;   // device (nz, 2, h): i0, i1 per level
  const float* yt;       // device (nz, h)
  const int32_t* xidx;   // device (nz, 2, w)
  const float* xt;       // device (nz, w)
  int nz, h, w;
  float inv;             // 1 / (nz + 1)
  uint32_t magic;        // 0x4b000000 (2^23 as float bits)
};
constexpr int kZoomTile = 32;

// byte K of w as a float: PRMT it into the mantissa of 2^23 (selector immediate, 2^23 in a register), subtract 2^23
template <int K>
__device__ __forceinline__ float zbyte(uint32_t w, uint32_t magic) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(magic), "n"(0x7440 + K));
  return __uint_as_float(r) - 8388608.0f;
}
constexpr int kZoomMaxLevels = 32;

__global__ void __launch_bounds__(256)
zoom_levels_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, ZoomLevelsParams p, int win_cap) {
  extern __shared__ __align__(16) uint8_t zsm[];
  // [nz][32] row tables (i0, i1, t) | [nz][32] column tables | source window (rows x cols x 3 bytes)
  int4* rowt = reinterpret_cast<int4*>(zsm);          // per (level, row): i0, i1, t -> then window byte offsets of the two rows, t
  int4* colt = rowt + p.nz * kZoomTile;               // per (level, column): the same for columns
  __shared__ int bbox[4];
  uint8_t* win = reinterpret_cast<uint8_t*>(colt + p.nz * kZoomTile);
  const int tid = threadIdx.x, lane = tid & 31, wy = tid >> 5;
  const int tiles_x = (p.w + kZoomTile - 1) / kZoomTile, tiles_y = (p.h + kZoomTile - 1) / kZoomTile;
  const int64_t total = n * tiles_x * tiles_y;
  for (int64_t job = blockIdx.x; job < total; job += gridDim.x) {
    const int64_t img = job / (tiles_x * tiles_y);
    const int tj = (int)(job - img * tiles_x * tiles_y), ty0 = (tj / tiles_x) * kZoomTile, tx0 = (tj % tiles_x) * kZoomTile;
    __syncthreads();   // the previous tile is done with the tables and the window
    if (tid < 4) bbox[tid] = tid < 2 ? 1 << 30 : -1;
    __syncthreads();
    int ymin = 1 << 30, ymax = -1, xmin = 1 << 30, xmax = -1;
    for (int e = tid; e < p.nz * kZoomTile; e += 256) {
      const int z = e >> 5, k = e & 31;
      const int y = min(ty0 + k, p.h - 1), x = min(tx0 + k, p.w - 1);
      const int a0 = p.yidx[(z * 2) * p.h + y], a1 = p.yidx[(z * 2 + 1) * p.h + y];
      const int b0 = p.xidx[(z * 2) * p.w + x], b1 = p.xidx[(z * 2 + 1) * p.w + x];
      rowt[e] = make_int4(a0, a1, __float_as_int(p.yt[z * p.h + y]), 0);
      colt[e] = make_int4(b0, b1, __float_as_int(p.xt[z * p.w + x]), 0);
      ymin = min(ymin, min(a0, a1)); ymax = max(ymax, max(a0, a1));
      xmin = min(xmin, min(b0, b1)); xmax = max(xmax, max(b0, b1));
    }
    ymin = min(ymin, ty0); ymax = max(ymax, min(ty0 + kZoomTile, p.h) - 1);      // the identity term reads the tile itself
    xmin = min(xmin, tx0); xmax = max(xmax, min(tx0 + kZoomTile, p.w) - 1);
    atomicMin(&bbox[0], ymin); atomicMin(&bbox[1], xmin); atomicMax(&bbox[2], ymax); atomicMax(&bbox[3], xmax);
    __syncthreads();
    const int wy0 = bbox[0], wx0 = bbox[1], wh = bbox[2] - bbox[0] + 1, ww = bbox[3] - bbox[1] + 1;
    const uint8_t* im = in + img * (int64_t)p.h * p.w * 3;
    const bool in_smem = wh * ww * 4 <= win_cap;   // always, for the zoom ranges of the severity table; else read through L1
    uint32_t* win32 = reinterpret_cast<uint32_t*>(win);
    if (in_smem) {
      // the window as one RGBx word per pixel: a tap is ONE shared-memory load
      // a warp per window row, lanes along it: 96 contiguous bytes per load instruction, several rows in flight
#pragma unroll 4
      for (int r = wy; r < wh; r += 8) {
        const uint8_t* q = im + ((int64_t)(wy0 + r) * p.w + wx0) * 3;
        for (int cc = lane; cc < ww; cc += 32)
          win32[r * ww + cc] = (uint32_t)q[cc * 3] | ((uint32_t)q[cc * 3 + 1] << 8) | ((uint32_t)q[cc * 3 + 2] << 16);
      }
    }
    __syncthreads();
    const int x = tx0 + lane;
    if (in_smem) {
      // tables of the inner loop, one 16-byte record per (level, row) and per (level, column): byte offsets into the window
      // (row records carry the window origin, so a tap's address is one add) and the interpolation weight
      for (int e = tid; e < p.nz * kZoomTile; e += 256) {
        const int4 r = rowt[e], c = colt[e];
        rowt[e] = make_int4(((r.x - wy0) * ww - wx0) * 4, ((r.y - wy0) * ww - wx0) * 4, r.z, 0);
        colt[e] = make_int4(c.x * 4, c.y * 4, c.z, 0);
      }
      __syncthreads();
      const uint32_t magic = p.magic;   // 0x4b000000 from the parameters: kept in a register, the selectors stay immediates
      const char* wb = reinterpret_cast<const char*>(win32);
      if (x < p.w) {
        for (int k = wy; k < kZoomTile; k += 8) {
          const int y = ty0 + k;
          if (y >= p.h) break;
          float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
#pragma unroll 2
          for (int z = 0; z < p.nz; ++z) {
            const int4 rr = rowt[z * kZoomTile + k], cc = colt[z * kZoomTile + lane];
            const float ty = __int_as_float(rr.z), tx = __int_as_float(cc.z);
            const float w11 = ty * tx, w10 = ty - w11, w01 = tx - w11, w00 = 1.0f - ty - w01;
            const uint32_t p00 = *reinterpret_cast<const uint32_t*>(wb + (rr.x + cc.x));
            const uint32_t p01 = *reinterpret_cast<const uint32_t*>(wb + (rr.x + cc.y));
            const uint32_t p10 = *reinterpret_cast<const uint32_t*>(wb + (rr.y + cc.x));
            const uint32_t p11 = *reinterpret_cast<const uint32_t*>(wb + (rr.y + cc.y));
            a0 = __fmaf_rn(zbyte<0>(p11, magic), w11, __fmaf_rn(zbyte<0>(p10, magic), w10, __fmaf_rn(zbyte<0>(p01, magic), w01, __fmaf_rn(zbyte<0>(p00, magic), w00, a0))));
            a1 = __fmaf_rn(zbyte<1>(p11, magic), w11, __fmaf_rn(zbyte<1>(p10, magic), w10, __fmaf_rn(zbyte<1>(p01, magic), w01, __fmaf_rn(zbyte<1>(p00, magic), w00, a1))));
            a2 = __fmaf_rn(zbyte<2>(p11, magic), w11, __fmaf_rn(zbyte<2>(p10, magic), w10, __fmaf_rn(zbyte<2>(p01, magic), w01, __fmaf_rn(zbyte<2>(p00, magic), w00, a2))));
          }
          const uint32_t self = win32[(y - wy0) * ww + (x - wx0)];
          uint8_t* o = out + (img * (int64_t)p.h * p.w + (int64_t)y * p.w + x) * 3;
          // 2^-10: a sum that is an integer in real arithmetic floors to itself
          o[0] = (uint8_t)floor_bits(fminf(fmaxf(__fmaf_rn(a0 + zbyte<0>(self, magic), p.inv, 0.0009765625f), 0.0f), 255.0f));
          o[1] = (uint8_t)floor_bits(fminf(fmaxf(__fmaf_rn(a1 + zbyte<1>(self, magic), p.inv, 0.0009765625f), 0.0f), 255.0f));
          o[2] = (uint8_t)floor_bits(fminf(fmaxf(__fmaf_rn(a2 + zbyte<2>(self, magic), p.inv, 0.0009765625f), 0.0f), 255.0f));
        }
      }
    }
    const uint8_t* base = im;
    const int pitch = p.w * 3, oy = 0, ox = 0;
    if (x < p.w && !in_smem) {
      for (int k = wy; k < kZoomTile; k += 8) {
        const int y = ty0 + k;
        if (y >= p.h) break;
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
        for (int z = 0; z < p.nz; ++z) {
          const int e = z * kZoomTile;
          const int4 rr = rowt[e + k], cr = colt[e + lane];
          const float ty = __int_as_float(rr.z), tx = __int_as_float(cr.z);
          const float w11 = ty * tx, w10 = ty - w11, w01 = tx - w11, w00 = 1.0f - ty - w01;
          const uint8_t* r0 = base + (rr.x - oy) * pitch;
          const uint8_t* r1 = base + (rr.y - oy) * pitch;
          const int c0 = (cr.x - ox) * 3, c1 = (cr.y - ox) * 3;
          // byte -> float without the conversion pipe: the byte planted in the mantissa of 2^23 (exact)
          auto bf = [](uint8_t b) { return __uint_as_float(0x4B000000u | (uint32_t)b) - 8388608.0f; };
          a0 = __fmaf_rn(bf(r1[c1]), w11, __fmaf_rn(bf(r1[c0]), w10, __fmaf_rn(bf(r0[c1]), w01, __fmaf_rn(bf(r0[c0]), w00, a0))));
          a1 = __fmaf_rn(bf(r1[c1 + 1]), w11, __fmaf_rn(bf(r1[c0 + 1]), w10, __fmaf_rn(bf(r0[c1 + 1]), w01, __fmaf_rn(bf(r0[c0 + 1]), w00, a1))));
          a2 = __fmaf_rn(bf(r1[c1 + 2]), w11, __fmaf_rn(bf(r1[c0 + 2]), w10, __fmaf_rn(bf(r0[c1 + 2]), w01, __fmaf_rn(bf(r0[c0 + 2]), w00, a2))));
        }
        const uint8_t* self = base + (y - oy) * pitch + (x - ox) * 3;
        uint8_t* o = out + (img * (int64_t)p.h * p.w + (int64_t)y * p.w + x) * 3;
        // 2^-10: a sum that is an integer in real arithmetic floors to itself
        o[0] = (uint8_t)floor_bits(fminf(fmaxf(__fmaf_rn(a0 + (float)self[0], p.inv, 0.0009765625f), 0.0f), 255.0f));
        o[1] = (uint8_t)floor_bits(fminf(fmaxf(__fmaf_rn(a1 + (float)self[1], p.inv, 0.0009765625f), 0.0f), 255.0f));
        o[2] = (uint8_t)floor_bits(fminf(fmaxf(__fmaf_rn(a2 + (float)self[2], p.inv, 0.0009765625f), 0.0f), 255.0f));
      }
    }
  }
}

int zoom_levels(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int nz, const int32_t* yidx, const float* yt,
                const int32_t* xidx, const float* xt, cudaStream_t stream) {
  IMGTX_CHECK_ARG(nz >= 1 && nz <= kZoomMaxLevels, "zoom levels must be in [1,%d]", kZoomMaxLevels);
  IMGTX_CHECK_ARG(yidx && yt && xidx && xt, "missing zoom tables");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null image pointer");
  ZoomLevelsParams p;
  p.yidx = yidx; p.yt = yt; p.xidx = xidx; p.xt = xt; p.nz = nz; p.h = h; p.w = w; p.inv = (float)(1.0 / (nz + 1));
  p.magic = 0x4b000000u;
  const size_t tab = (size_t)nz * kZoomTile * 32;
  // window side: the tile plus the drift of its source towards the centre over the levels (zoom <= 1.34 in the severity
  // table); a table with larger factors falls back to reading the source through L1
  const int side = std::min(96, kZoomTile + (int)(std::max(h, w) * 0.13f) + 6);
  const int win_cap = side * side * 4;
  const size_t smem = tab + win_cap + 16;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(zoom_levels_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t tiles = n * ((w + kZoomTile - 1) / kZoomTile) * ((h + kZoomTile - 1) / kZoomTile);
  const int64_t grid = std::min<int64_t>(tiles, (int64_t)sm_count() * 4);
  zoom_levels_kernel<<<(unsigned)grid, 256, smem, stream>>>(in, out, n, p, win_cap);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// The merged zoom_blur operator of 32 x 32 images on the integer tensor cores (imgtx_cc_op32_mma_u8).
//
// zoom_blur is ONE linear operator per severity with non-negative weights that sum to 1 per output pixel, the same for every
// image: out[p][c] = floor(sum_k W[p][k] x[k][c]).  zoom_fast_kernel above spends 7 issue slots per tap and pixel on it.
// As a matrix product the images are the columns: C[16 output pixels][8 images] += A[16][32 source pixels] B[32][8 images]
// is one `mma.sync.m16n8k32.u8.u8.s32` (IMMA.16832, 2.1 clk per SM), and W is sparse but local: the taps of a 4 x 4 tile of
// output pixels touch 14-24 aligned 4-pixel groups ("quads") of the source rows, i.e. 2-3 MMA steps of 8 quads each
// (113-144 steps per image for the five severities instead of 1024 x 5.5-12.3 taps).
//   * The host compiles W (imagetransformations_b200/cifar_c.py:op32_mma_tables): per tile its quad list and the A fragments
//     of the Q16 weights, low and high byte limb; per row a rounding bias.  Q16 rounded to nearest with the row sum kept at
//     2^16: flat regions are exact, the result differs from the real-valued sum by < 0.01 LSB before the floor.
//   * A CTA (16 warps) works on a GROUP of 8 consecutive images (24 576 contiguous bytes: one bulk TMA load into a ring slot,
//     one bulk store from the same slot).  The group is de-interleaved once into byte planes (8 images x 3 channels, row
//     pitch 10 words, image pitch = 4 mod 32 words: bank = 4 image + 2 row + column, and the host orders the quads of a step
//     so that the four a load touches differ in 2 row + column mod 4 -> conflict-free); a B fragment is then two 32-bit
//     shared-memory loads at (image = lane / 4, quad = table[lane % 4]).  The planes are double-buffered: group i+1 is
//     de-interleaved by the warps as they finish group i, so there is ONE CTA barrier per group.
//   * Warp w owns tiles 4w .. 4w+3 for the whole launch: the high-limb A fragments stay in 48 REGISTERS, the low-limb ones in
//     shared memory (56-72 KB).  out = (((bias + sum lo x) >> 8) + sum hi x) >> 8 is exactly (bias + sum w x) >> 16.
//   * Accumulator rows m and m + 8 are horizontal neighbours, so a lane packs 6 contiguous output bytes per image.
constexpr int kOpGroup = 8;
constexpr int kOpImg = 3072;
constexpr int kOpGroupBytes = kOpGroup * kOpImg;
constexpr int kOpRowWords = 10;                               // plane row pitch: 8 quads + 2 (bank spread, see op32_mma_tables)
constexpr int kOpPlane = 32 * kOpRowWords * 4;                 // one channel plane
constexpr int kOpPlanePitch = 3 * kOpPlane + 16;               // image pitch = 4 (mod 32) words
constexpr int kOpThreads = 512;
constexpr int kOpFixed = 384 + 6144 + 2048 + 1024;   // barriers + metric partials | quad table | row biases | tile meta

struct Op32Tables {
  const int32_t* meta;    // [64][4]: steps, first low-limb fragment
  const uint32_t* bias;   // [64][8]: bias of row gid | bias of row gid + 8 << 16
  const int2* qtab;       // [64][3][4]: plane byte offsets of quad slots tig and tig + 4
  const uint4* lo;        // [total_steps][32]
  const uint4* hi;        // [64][3][32]
  int total_steps;
};

__global__ void __launch_bounds__(kOpThreads, 1)
op32_mma_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, Op32Tables tb, int ring,
                uint2* __restrict__ metrics) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint2* msum = reinterpret_cast<uint2*>(smem + 128);
  int2* qt = reinterpret_cast<int2*>(smem + 384);
  uint32_t* biasw = reinterpret_cast<uint32_t*>(smem + 384 + 6144);
  int32_t* meta = reinterpret_cast<int32_t*>(smem + 384 + 6144 + 2048);
  uint4* alo = reinterpret_cast<uint4*>(smem + kOpFixed);
  uint8_t* planar = smem + kOpFixed + (size_t)tb.total_steps * 512;
  uint8_t* slots = planar + 2 * align_up((size_t)kOpGroup * kOpPlanePitch, 128);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, tig = lane & 3;

  if (tid == 0) {
    for (int i = 0; i < ring; ++i) mbar_init(&full[i], 1);
    fence_mbar_init();
  }
  for (int i = tid; i < 64 * 3 * 4; i += kOpThreads) qt[i] = tb.qtab[i];
  for (int i = tid; i < 64 * 8; i += kOpThreads) biasw[i] = tb.bias[i];
  for (int i = tid; i < 64 * 4; i += kOpThreads) meta[i] = tb.meta[i];
  for (int i = tid; i < tb.total_steps * 32; i += kOpThreads) alo[i] = tb.lo[i];
  uint4 ahi[4][3];
#pragma unroll
  for (int tl = 0; tl < 4; ++tl)
#pragma unroll
    for (int s2 = 0; s2 < 3; ++s2) ahi[tl][s2] = tb.hi[((warp * 4 + tl) * 3 + s2) * 32 + lane];
  __syncthreads();

  const int64_t ngroups = (n + kOpGroup - 1) / kOpGroup, stride = gridDim.x, first = blockIdx.x;
  auto group_bytes = [&](int64_t g) { return (uint32_t)min((int64_t)kOpGroup, n - g * kOpGroup) * (uint32_t)kOpImg; };
  if (tid == 0) {
    for (int i = 0; i < ring; ++i) {
      const int64_t g = first + (int64_t)i * stride;
      if (g < ngroups) {
        mbar_expect_tx(&full[i], group_bytes(g));
        bulk_g2s(slots + (size_t)i * kOpGroupBytes, in + g * kOpGroupBytes, group_bytes(g), &full[i]);
      }
    }
  }
  // metric of the group of iteration `it` from the per-warp partials (8 lanes of warp 0, one image each)
  auto flush_metrics = [&](int it, int64_t g) {
    if (warp == 0 && lane < kOpGroup && g * kOpGroup + lane < n) {
      const uint2 a = msum[(it & 1) * 16 + 2 * lane], b = msum[(it & 1) * 16 + 2 * lane + 1];
      metrics[g * kOpGroup + lane] = make_uint2(a.x + b.x, a.y + b.y);
    }
  };

  constexpr int kPlanes = (kOpGroup * kOpPlanePitch + 127) / 128 * 128;
  // RGB bytes of a ring slot -> byte planes, one quad (4 pixels, 3 words) per thread-step
  auto convert = [&](const uint8_t* sb, uint8_t* dst) {
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(sb);
    uint32_t* p32 = reinterpret_cast<uint32_t*>(dst);
#pragma unroll
    for (int task = tid; task < kOpGroup * 256; task += kOpThreads) {
      const int img = task >> 8, q = task & 255;
      const uint32_t* w = s32 + img * 768 + q * 3;
      const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
      uint32_t* d = p32 + img * (kOpPlanePitch / 4) + (q >> 3) * kOpRowWords + (q & 7);
      d[0] = __byte_perm(__byte_perm(w0, w1, 0x6630), w2, 0x5210);                  // bytes 0, 3, 6, 9
      d[kOpPlane / 4] = __byte_perm(__byte_perm(w0, w1, 0x7741), w2, 0x6210);       // bytes 1, 4, 7, 10
      d[2 * kOpPlane / 4] = __byte_perm(__byte_perm(w0, w1, 0x5552), w2, 0x7410);   // bytes 2, 5, 8, 11
    }
  };
  int it = 0, slot = 0, prev_slot = 0;
  uint32_t phase = 0;
  int64_t g = first;
  if (g < ngroups) {
    mbar_wait(&full[0], 0u);
    convert(slots, planar);
  }
  __syncthreads();
  for (; g < ngroups; ++it, g += stride) {
    uint8_t* sb = slots + (size_t)slot * kOpGroupBytes;
    int nslot = slot + 1;
    uint32_t nphase = phase;
    if (nslot == ring) { nslot = 0; nphase ^= 1u; }

    const uint8_t* pl = planar + (it & 1) * kPlanes + gid * kOpPlanePitch;
#pragma unroll
    for (int tl = 0; tl < 4; ++tl) {
      const int tile = warp * 4 + tl;
      const int ns = meta[tile * 4], ab = meta[tile * 4 + 1];
      const uint32_t bw = biasw[tile * 8 + gid];
      int acc[3][4];
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        acc[ch][0] = acc[ch][1] = (int)(bw & 0xffffu);
        acc[ch][2] = acc[ch][3] = (int)(bw >> 16);
      }
      uint32_t b[3][3][2];
#pragma unroll
      for (int s2 = 0; s2 < 3; ++s2) {
        if (s2 < ns) {
          const int2 qo = qt[(tile * 3 + s2) * 4 + tig];   // quad index * 4
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) {
            b[s2][ch][0] = *reinterpret_cast<const uint32_t*>(pl + ch * kOpPlane + qo.x);
            b[s2][ch][1] = *reinterpret_cast<const uint32_t*>(pl + ch * kOpPlane + qo.y);
          }
          const uint4 a = alo[(ab + s2) * 32 + lane];
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) imma_uu(acc[ch], a, b[s2][ch][0], b[s2][ch][1]);
        }
      }
#pragma unroll
      for (int ch = 0; ch < 3; ++ch)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[ch][i] >>= 8;
#pragma unroll
      for (int s2 = 0; s2 < 3; ++s2) {
        if (s2 < ns) {
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) imma_uu(acc[ch], ahi[tl][s2], b[s2][ch][0], b[s2][ch][1]);
        }
      }
      // rows gid and gid + 8 are the pixel pair (x, x + 1): six contiguous bytes per image, images 2 tig and 2 tig + 1
      // the weights of a row sum to 2^16 at most: the result is byte 1 of the accumulator, no saturation needed
      const int p = meta[tile * 4 + 2] + (gid >> 1) * 32 + (gid & 1) * 2;
      uint8_t* o = sb + (2 * tig) * kOpImg + p * 3;
#pragma unroll
      for (int im = 0; im < 2; ++im) {
        uint16_t* o16 = reinterpret_cast<uint16_t*>(o + im * kOpImg);
        o16[0] = (uint16_t)__byte_perm(acc[0][im], acc[1][im], 0x0051);
        o16[1] = (uint16_t)__byte_perm(acc[2][im], acc[0][2 + im], 0x0051);
        o16[2] = (uint16_t)__byte_perm(acc[1][2 + im], acc[2][2 + im], 0x0051);
      }
    }
    if (g + stride < ngroups) {   // the next group's planes, while slower warps are still on this group's tiles
      mbar_wait(&full[nslot], nphase);
      convert(slots + (size_t)nslot * kOpGroupBytes, planar + ((it + 1) & 1) * kPlanes);
    }
    fence_async_smem();
    __syncthreads();   // the slot holds the group's output, the other plane buffer the next group's input
    if (metrics && it > 0) flush_metrics(it - 1, g - stride);
    if (metrics) {     // two warps per image
      // lane owns words [12 l, 12 l + 12) of its half image: byte weights 1..48 fit a dp4a operand (team_pipeline.cuh)
      const int w0i = (warp & 1) * 384 + lane * 12;
      const uint4* o128 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(sb) + (warp >> 1) * 768 + w0i);
      uint32_t c0 = 0, c1 = 0, cw = 0;
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const uint4 v = o128[q];
        const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          c0 = __dp4a(wv[e], 0x01010101u, c0);
          cw = __dp4a(wv[e], 0x04030201u + (uint32_t)(q * 4 + e) * 0x04040404u, cw);
        }
      }
      c1 = cw + 4u * (uint32_t)w0i * c0;
      c0 = __reduce_add_sync(0xffffffffu, c0);
      c1 = __reduce_add_sync(0xffffffffu, c1);
      if (lane == 0) msum[(it & 1) * 16 + warp] = make_uint2(c0, c1);
    }
    if (tid == 0) {
      bulk_s2g(out + g * kOpGroupBytes, sb, group_bytes(g));
      bulk_commit();
      if (it >= 1) {
        bulk_wait_read<1>();   // the store of the previous iteration has finished reading its slot: refill it
        const int64_t gn = g + (int64_t)(ring - 1) * stride;
        if (gn < ngroups) {
          mbar_expect_tx(&full[prev_slot], group_bytes(gn));
          bulk_g2s(slots + (size_t)prev_slot * kOpGroupBytes, in + gn * kOpGroupBytes, group_bytes(gn), &full[prev_slot]);
        }
      }
    }
    prev_slot = slot;
    slot = nslot;
    phase = nphase;
  }
  __syncthreads();
  if (metrics && it > 0) flush_metrics(it - 1, g - stride);
  if (tid == 0) bulk_wait_all<0>();
}

int op32_mma(const uint8_t* in, uint8_t* out, int64_t n, const int32_t* tables, int64_t table_words, int total_steps,
             uint32_t* metrics, cudaStream_t stream) {
  IMGTX_CHECK_ARG(total_steps >= 64 && total_steps <= 192, "operator tables: %d steps (64..192)", total_steps);
  const int64_t off_meta = 16, off_bias = off_meta + 256, off_qtab = off_bias + 512, off_lo = off_qtab + 1536,
                off_hi = off_lo + (int64_t)total_steps * 128, words = off_hi + 64 * 3 * 128;
  IMGTX_CHECK_ARG(tables != nullptr && table_words == words && aligned16(tables),
                  "operator tables: %lld words, %lld expected for %d steps", (long long)table_words, (long long)words, total_steps);
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out && aligned16(in) && aligned16(out), "image buffers must be 16-byte aligned");
  Op32Tables tb;
  tb.meta = tables + off_meta;
  tb.bias = reinterpret_cast<const uint32_t*>(tables + off_bias);
  tb.qtab = reinterpret_cast<const int2*>(tables + off_qtab);
  tb.lo = reinterpret_cast<const uint4*>(tables + off_lo);
  tb.hi = reinterpret_cast<const uint4*>(tables + off_hi);
  tb.total_steps = total_steps;
  const size_t base = kOpFixed + (size_t)total_steps * 512 + 2 * align_up((size_t)kOpGroup * kOpPlanePitch, 128);
  const int ring = (int)std::min<size_t>(6, ((size_t)max_smem_optin() - base) / kOpGroupBytes);
  IMGTX_CHECK_ARG(ring >= 3, "operator tables too large for shared memory");
  const size_t smem = base + (size_t)ring * kOpGroupBytes;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(op32_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t ngroups = (n + kOpGroup - 1) / kOpGroup;
  const int64_t grid = std::min<int64_t>(ngroups, sm_count());
  op32_mma_kernel<<<(unsigned)grid, kOpThreads, smem, stream>>>(in, out, n, tb, ring, reinterpret_cast<uint2*>(metrics));
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_cc_op32_mma_u8(const uint8_t* in, uint8_t* out, int64_t n, const int32_t* tables_dev, int64_t table_words,
                                    int total_steps, uint32_t* metrics_dev, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(n >= 0, "negative batch size");
  return op32_mma(in, out, n, tables_dev, table_words, total_steps, metrics_dev, (cudaStream_t)stream);
}

extern "C" int imgtx_cc_zoom_levels_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int n_levels,
                                       const int32_t* yidx_dev, const float* yt_dev, const int32_t* xidx_dev,
                                       const float* xt_dev, uint32_t* metrics_dev, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  int rc = zoom_levels(in, out, n, h, w, n_levels, yidx_dev, yt_dev, xidx_dev, xt_dev, (cudaStream_t)stream);
  if (rc == IMGTX_OK && metrics_dev) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}
