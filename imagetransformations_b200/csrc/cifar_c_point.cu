// CIFAR-10-C-style corruptions, part 1: pointwise types (float64 arithmetic, exactly the canonical numpy expressions;
// specification: oracle/cifar_c.py).  The reference names the 15 types
// (attacks/fgsm/cifar-10/reverse_fgsm_cifar10_corrupted.py:18-23) and consumes (50000,32,32,3) uint8 files of them
// (fall_2025/TTA_transforms.py:16-26); SURVEY.md App. B gives the provisional spec.
//
//   IMGTX_CC_GAUSSIAN    clip(x/255 + n, 0, 1)*255                 n: injected f64 draws or on-chip N(0, p0)
//   IMGTX_CC_SHOT        clip(k/p0, 0, 1)*255                      k: injected Poisson counts or on-chip Poisson(x/255*p0)
//   IMGTX_CC_IMPULSE     salt & pepper per element                 u1 <= p0 flips; u2 <= 0.5 salts
//   IMGTX_CC_SPECKLE     clip(v + v*n, 0, 1)*255, v = x/255
//   IMGTX_CC_BRIGHTNESS  rgb -> hsv, V = clip(V + p0), hsv -> rgb (skimage formulas)
//   IMGTX_CC_CONTRAST    clip((v - mean_c)*p0 + mean_c)            per-image per-channel mean (warp-shuffle reduction)
//   IMGTX_CC_FOG         (v + p0*plasma) * max/(max + p0)          per-image max; plasma injected or generated on chip
//                                                                  (diamond-square, p1 = wibbledecay)
//   IMGTX_CC_FROST       clip(p0*x + p1*frost, 0, 255)             frost: injected crops, or random crops of a texture bank
// Every byte is read once and written once (staged TMA pipeline); injected draws are streamed from HBM.
// Roofline: HBM, 2 B/element (+ 8 B/element of draws in parity mode); on-chip RNG modes are ALU-bound.
#include "pipeline.cuh"
#include "rng.cuh"
#include "cifar_c_fast.cuh"
#include "cifar_c_team.cuh"
#include <math.h>
#include <string.h>
#include <vector>
#include <algorithm>

namespace imgtx {

struct CcPointParams {
  int kind;
  double p0, p1;
  const void* a;   // draws (see header), device
  const void* b;
  uint64_t seed;
  int tex_n, tex_h, tex_w;
  int plasma_pitch;   // fog, maps generated on the device for large images: rows of `plasma_pitch` doubles, pitch^2 per image
};

__device__ __forceinline__ uint8_t q255(double v) {  // np.uint8(np.clip(v, 0, 1) * 255)
  v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
  return (uint8_t)(int)(v * 255.0);
}

// On-chip plasma fractal (make_cifar_c.plasma_fractal) for one image: map is msz x msz doubles in shared memory.
__device__ void plasma_device(double* map, int msz, double decay, const Philox& rng, uint32_t img_lo, uint32_t img_hi,
                              double* red /* 2 doubles */) {
  const int tid = threadIdx.x, nt = blockDim.x;
  if (tid == 0) map[0] = 0.0;
  __syncthreads();
  double wibble = 100.0;
  uint32_t draw = 0;
  for (int s = msz; s >= 2; s >>= 1) {
    const int half = s >> 1, m = msz / s;
    // square centres
    for (int i = tid; i < m * m; i += nt) {
      const int a = i / m, b = i - a * m, a1 = (a + 1) % m, b1 = (b + 1) % m;
      const double acc = map[(a * s) * msz + b * s] + map[(a1 * s) * msz + b * s] + map[(a * s) * msz + b1 * s] +
                         map[(a1 * s) * msz + b1 * s];
      const uint4 r = rng(img_lo, img_hi, 0x70u + draw, (uint32_t)i);
      const double u = -wibble + 2.0 * wibble * (double)u01(r.x);
      map[(half + a * s) * msz + half + b * s] = acc / 4 + wibble * u;
    }
    __syncthreads();
    // edge midpoints (both families read only corners and centres: no hazard between them)
    for (int i = tid; i < 2 * m * m; i += nt) {
      const int fam = i / (m * m), k = i - fam * m * m;
      const int a = k / m, b = k - a * m;
      const uint4 r = rng(img_lo, img_hi, 0x71u + draw + fam, (uint32_t)k);
      const double u = -wibble + 2.0 * wibble * (double)u01(r.x);
      auto dr = [&](int ii, int jj) { return map[(half + ((ii + m) % m) * s) * msz + half + ((jj + m) % m) * s]; };
      auto ul = [&](int ii, int jj) { return map[(((ii + m) % m) * s) * msz + ((jj + m) % m) * s]; };
      if (fam == 0) {
        const double acc = dr(a, b) + dr(a - 1, b) + ul(a, b) + ul(a, b + 1);
        map[(a * s) * msz + half + b * s] = acc / 4 + wibble * u;
      } else {
        const double acc = dr(a, b) + dr(a, b - 1) + ul(a, b) + ul(a + 1, b);
        map[(half + a * s) * msz + b * s] = acc / 4 + wibble * u;
      }
    }
    __syncthreads();
    draw += 3;
    wibble /= decay;
  }
  // normalise to [0,1]
  double mn = 1e300, mx = -1e300;
  for (int i = tid; i < msz * msz; i += nt) { mn = fmin(mn, map[i]); mx = fmax(mx, map[i]); }
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (tid == 0) { red[0] = 1e300; red[1] = -1e300; }
  __syncthreads();
  if ((tid & 31) == 0) {
    // doubles have no native atomic min/max: serialise the few warps through a CAS loop on the bit pattern
    unsigned long long* p0 = reinterpret_cast<unsigned long long*>(&red[0]);
    unsigned long long* p1 = reinterpret_cast<unsigned long long*>(&red[1]);
    unsigned long long old = *p0, assumed;
    do { assumed = old; if (__longlong_as_double(assumed) <= mn) break; old = atomicCAS(p0, assumed, __double_as_longlong(mn)); } while (assumed != old);
    old = *p1;
    do { assumed = old; if (__longlong_as_double(assumed) >= mx) break; old = atomicCAS(p1, assumed, __double_as_longlong(mx)); } while (assumed != old);
  }
  __syncthreads();
  const double lo = red[0], span = red[1] - red[0];
  for (int i = tid; i < msz * msz; i += nt) map[i] = (map[i] - lo) / span;
  __syncthreads();
}

// Plasma maps for images that are not a power-of-two square of at most 64 pixels: one CTA per image builds the msz x msz
// map in global memory (msz = next power of two, as the canonical ImageNet-C generator draws 256 x 256 for 224 x 224); the
// fog pass then reads its top-left h x w corner.  Same draw convention as the shared-memory form.
__global__ void __launch_bounds__(256)
plasma_map_kernel(double* maps, int msz, double decay, uint64_t seed, int64_t img0) {
  __shared__ double red[2];
  const Philox rng(seed);
  const int64_t img = img0 + blockIdx.x;
  plasma_device(maps + (size_t)blockIdx.x * msz * msz, msz, decay, rng, (uint32_t)img, (uint32_t)(img >> 32), red);
}

// fog with on-chip draws on large images, throughput form: one CTA of 1024 threads per image draws the msz x msz
// diamond-square map in float32 in a global workspace (power-of-two index arithmetic, four uniforms per Philox call),
// reduces its range and the image's maximum, and writes q = x * S + plasma * C with S = m / (m + c0), C = 255 * c0 * S
// (m = max / 255) — the float32 evaluation of the exact form's expression, as OpTFog does for 32 x 32.
__global__ void __launch_bounds__(1024)
fog_large_fast_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w, int lg, float c0,
                      float decay, uint64_t seed, float* __restrict__ work) {
  __shared__ float red[3][32];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const int msz = 1 << lg, MASK = msz - 1;
  float* map = work + (size_t)blockIdx.x * msz * msz;
  const Philox rng(seed);
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    __syncthreads();
    if (tid == 0) map[0] = 0.0f;
    __syncthreads();
    float wib = 100.0f;
    uint32_t draw = 0;
    for (int ls = lg; ls >= 1; --ls) {
      const int s = 1 << ls, half = s >> 1, lm = lg - ls, m = 1 << lm, mm = m * m;
      const float w2 = wib * wib;   // wibble * uniform(-wibble, wibble)
      // square centres, four per thread and Philox call
      for (int g4 = tid; 4 * g4 < mm; g4 += nt) {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x70u + draw, (uint32_t)g4);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = 4 * g4 + e;
          if (i >= mm) break;
          const int a = i >> lm, b = i & (m - 1);
          const int y0 = a << ls, x0 = b << ls, y1 = (y0 + s) & MASK, x1 = (x0 + s) & MASK;
          const float acc = map[y0 * msz + x0] + map[y1 * msz + x0] + map[y0 * msz + x1] + map[y1 * msz + x1];
          const float u = (float)(rr[e] >> 8) * (2.0f / 16777216.0f) - 1.0f + (1.0f / 16777216.0f);
          map[(y0 + half) * msz + x0 + half] = acc * 0.25f + w2 * u;
        }
      }
      __syncthreads();
      // edge midpoints: family 0 at (a s, b s + half), family 1 at (a s + half, b s); both read corners and centres only
      for (int g4 = tid; 4 * g4 < 2 * mm; g4 += nt) {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x71u + draw, (uint32_t)g4);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = 4 * g4 + e;
          if (i >= 2 * mm) break;
          const int fam = i >= mm, k = i - fam * mm, a = k >> lm, b = k & (m - 1);
          const int y = (a << ls) + (fam ? half : 0), x = (b << ls) + (fam ? 0 : half);
          const float acc = map[((y - half) & MASK) * msz + x] + map[((y + half) & MASK) * msz + x] +
                            map[y * msz + ((x - half) & MASK)] + map[y * msz + ((x + half) & MASK)];
          const float u = (float)(rr[e] >> 8) * (2.0f / 16777216.0f) - 1.0f + (1.0f / 16777216.0f);
          map[y * msz + x] = acc * 0.25f + w2 * u;
        }
      }
      __syncthreads();
      draw += 2;
      wib /= decay;
    }
    // range of the whole map (the canonical generator normalises before cropping) and the image's maximum
    float mn = 3.4e38f, mx = -3.4e38f;
    for (int i = tid; i < msz * msz; i += nt) { const float v = map[i]; mn = fminf(mn, v); mx = fmaxf(mx, v); }
    const uint8_t* im = in + img * (int64_t)h * w * 3;
    unsigned bm = 0;
    for (int i = tid; i < h * w * 3; i += nt) bm = max(bm, (unsigned)im[i]);
    for (int o = 16; o > 0; o >>= 1) {
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      bm = max(bm, __shfl_xor_sync(0xffffffffu, bm, o));
    }
    if (lane == 0) { red[0][warp] = mn; red[1][warp] = mx; red[2][warp] = (float)bm; }
    __syncthreads();
    mn = red[0][0]; mx = red[1][0];
    float bmf = red[2][0];
    for (int k = 1; k < (nt >> 5); ++k) { mn = fminf(mn, red[0][k]); mx = fmaxf(mx, red[1][k]); bmf = fmaxf(bmf, red[2][k]); }
    const float mval = bmf / 255.0f, S = mval / (mval + c0), C = 255.0f * c0 * S / (mx - mn);
    uint8_t* om = out + img * (int64_t)h * w * 3;
    for (int px = tid; px < h * w; px += nt) {
      const int y = px / w, x = px - y * w;
      const float pl = (map[y * msz + x] - mn) * C;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float q = __fmaf_rn((float)im[px * 3 + c], S, pl);
        om[px * 3 + c] = (uint8_t)(int)fminf(fmaxf(q, 0.0f), 255.0f);
      }
    }
  }
}

struct OpCcPoint : NoSpill {
  using Params = CcPointParams;
  static constexpr int kThreads = 256;
  // T64 LUT (2 KB) + reduction cells (64 B) + plasma map (fog on-chip mode; sized for 32x32, see entry point)
  static size_t scratch_bytes(const Shape&, int) { return 2048 + 64 + 16; }
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }

  __device__ static void prepare(const Shape&, const Params&, uint8_t* scratch, bool) {
    double* T = reinterpret_cast<double*>(scratch);
    for (int i = threadIdx.x; i < 256; i += kThreads) T[i] = (double)i / 255.0;
    __syncthreads();
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const double* T = reinterpret_cast<const double*>(scratch);
    unsigned long long* cells = reinterpret_cast<unsigned long long*>(scratch + 2048);  // 8 cells
    double* plasma_map = reinterpret_cast<double*>(scratch + 2048 + 64 + 16);
    const int tid = threadIdx.x, lane = tid & 31;
    const int hw = s.h * s.w, ne = hw * s.c;
    const Philox rng(p.seed);
    for (int j = 0; j < g; ++j) {
      const int64_t img = img0 + j;
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      const int64_t ebase = img * ne;
      switch (p.kind) {
        case IMGTX_CC_GAUSSIAN:
        case IMGTX_CC_SPECKLE: {
          const double* nz = reinterpret_cast<const double*>(p.a);
          for (int i = tid; i < ne; i += kThreads) {
            double z;
            if (nz) {
              z = nz[ebase + i];
            } else {
              const uint64_t e = (uint64_t)(ebase + i);
              const uint4 r = rng((uint32_t)(e >> 1), (uint32_t)(e >> 33), 0x10u, 0u);
              float z0, z1;
              box_muller(r.x, r.y, z0, z1);
              z = (double)((e & 1) ? z1 : z0) * p.p0;
            }
            const double v = T[im[i]];
            om[i] = q255(p.kind == IMGTX_CC_GAUSSIAN ? v + z : v + v * z);
          }
          break;
        }
        case IMGTX_CC_SHOT: {
          const long long* cnt = reinterpret_cast<const long long*>(p.a);
          for (int i = tid; i < ne; i += kThreads) {
            double k;
            if (cnt) {
              k = (double)cnt[ebase + i];
            } else {
              const uint64_t e = (uint64_t)(ebase + i);
              uint32_t ctr = 0;
              k = (double)poisson_sample(rng, (uint32_t)e, (uint32_t)(e >> 32), 0x20u, ctr, (float)(T[im[i]] * p.p0));
            }
            om[i] = q255(k / p.p0);
          }
          break;
        }
        case IMGTX_CC_IMPULSE: {
          const double* u1 = reinterpret_cast<const double*>(p.a);
          const double* u2 = reinterpret_cast<const double*>(p.b);
          for (int i = tid; i < ne; i += kThreads) {
            double a, b;
            if (u1) {
              a = u1[ebase + i];
              b = u2[ebase + i];
            } else {
              const uint64_t e = (uint64_t)(ebase + i);
              const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x30u, 0u);
              a = (double)u01(r.x);
              b = (double)u01(r.y);
            }
            uint8_t o = im[i];
            if (a <= p.p0) o = (b <= 0.5) ? 255 : 0;
            om[i] = o;
          }
          break;
        }
        case IMGTX_CC_BRIGHTNESS: {
          for (int px = tid; px < hw; px += kThreads) {
            const double r = T[im[px * 3]], gg = T[im[px * 3 + 1]], b = T[im[px * 3 + 2]];
            const double v = fmax(r, fmax(gg, b)), mn = fmin(r, fmin(gg, b));
            const double delta = v - mn;
            double sat = 0.0, h = 0.0;
            if (delta != 0.0) {
              sat = delta / v;
              if (r == v) h = (gg - b) / delta;
              if (gg == v) h = 2.0 + (b - r) / delta;
              if (b == v) h = 4.0 + (r - gg) / delta;
              h = h / 6.0;
              h = h - trunc(h);          // fmod(h, 1.0), exact
              if (h < 0.0) h = h + 1.0;   // numpy's % takes the divisor's sign
            }
            double v2 = v + p.p0;
            v2 = v2 < 0.0 ? 0.0 : (v2 > 1.0 ? 1.0 : v2);
            const double h6 = h * 6;
            const double hif = floor(h6);
            const double f = h6 - hif;
            const double pp = v2 * (1 - sat), q = v2 * (1 - f * sat), t = v2 * (1 - (1 - f) * sat);
            const int hi = ((int)hif & 255) % 6;
            double o0, o1, o2;
            switch (hi) {
              case 0: o0 = v2; o1 = t; o2 = pp; break;
              case 1: o0 = q; o1 = v2; o2 = pp; break;
              case 2: o0 = pp; o1 = v2; o2 = t; break;
              case 3: o0 = pp; o1 = q; o2 = v2; break;
              case 4: o0 = t; o1 = pp; o2 = v2; break;
              default: o0 = v2; o1 = pp; o2 = q; break;
            }
            om[px * 3] = q255(o0); om[px * 3 + 1] = q255(o1); om[px * 3 + 2] = q255(o2);
          }
          break;
        }
        case IMGTX_CC_CONTRAST: {
          __syncthreads();
          if (tid < 4) cells[tid] = 0ull;
          __syncthreads();
          unsigned a0 = 0, a1 = 0, a2 = 0;
          for (int px = tid; px < hw; px += kThreads) { a0 += im[px * 3]; a1 += im[px * 3 + 1]; a2 += im[px * 3 + 2]; }
          for (int o = 16; o > 0; o >>= 1) {
            a0 += __shfl_xor_sync(0xffffffffu, a0, o);
            a1 += __shfl_xor_sync(0xffffffffu, a1, o);
            a2 += __shfl_xor_sync(0xffffffffu, a2, o);
          }
          if (lane == 0) { atomicAdd(&cells[0], a0); atomicAdd(&cells[1], a1); atomicAdd(&cells[2], a2); }
          __syncthreads();
          double mean[3];
          for (int ch = 0; ch < 3; ++ch) mean[ch] = (double)cells[ch] / (double)hw / 255.0;
          for (int i = tid; i < ne; i += kThreads) {
            const int ch = i % 3;
            om[i] = q255((T[im[i]] - mean[ch]) * p.p0 + mean[ch]);
          }
          break;
        }
        case IMGTX_CC_FOG: {
          __syncthreads();
          if (tid == 0) cells[4] = 0ull;
          __syncthreads();
          unsigned mx = 0;
          for (int i = tid; i < ne; i += kThreads) mx = max(mx, (unsigned)im[i]);
          for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
          if (lane == 0) atomicMax(&cells[4], (unsigned long long)mx);
          __syncthreads();
          const double mval = T[(int)cells[4]];
          const double* pl = reinterpret_cast<const double*>(p.a);
          if (pl && p.plasma_pitch) {
            // maps made by plasma_map_kernel (power-of-two side, top-left h x w corner used): own loop, then next image
            pl += img * (int64_t)p.plasma_pitch * p.plasma_pitch;
            const double den2 = mval + p.p0;
            for (int i = tid; i < ne; i += kThreads) {
              const int px = i / 3, y = px / s.w, x = px - y * s.w;
              const double v = T[im[i]] + p.p0 * pl[(size_t)y * p.plasma_pitch + x];
              om[i] = q255(v * mval / den2);
            }
            break;
          }
          if (pl) {
            pl += img * hw;
          } else {
            plasma_device(plasma_map, s.h, p.p1, rng, (uint32_t)img, (uint32_t)(img >> 32),
                          reinterpret_cast<double*>(&cells[5]));
            pl = plasma_map;
          }
          const double den = mval + p.p0;
          for (int i = tid; i < ne; i += kThreads) {
            const double v = T[im[i]] + p.p0 * pl[i / 3];
            om[i] = q255(v * mval / den);
          }
          break;
        }
        default: {  // IMGTX_CC_FROST
          const uint8_t* fr = reinterpret_cast<const uint8_t*>(p.a);
          if (p.tex_n == 0) {
            fr += ebase;
            for (int i = tid; i < ne; i += kThreads) {
              double v = p.p0 * (double)im[i] + p.p1 * (double)fr[i];
              v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);
              om[i] = (uint8_t)(int)v;
            }
          } else {
            // random crop of a random texture: idx = randint(n_tex), (y0, x0) = randint(th - h), randint(tw - w)
            const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x50u, 0u);
            const int ti = (int)(r.x % (uint32_t)p.tex_n);
            const int y0 = (int)(r.y % (uint32_t)(p.tex_h - s.h)), x0 = (int)(r.z % (uint32_t)(p.tex_w - s.w));
            const uint8_t* tex = fr + ((size_t)ti * p.tex_h + y0) * p.tex_w * 3 + (size_t)x0 * 3;
            const int rowb = s.w * 3;
            for (int i = tid; i < ne; i += kThreads) {
              const int y = i / rowb, xb = i - y * rowb;
              double v = p.p0 * (double)im[i] + p.p1 * (double)__ldg(tex + (size_t)y * p.tex_w * 3 + xb);
              v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);
              om[i] = (uint8_t)(int)v;
            }
          }
          break;
        }
      }
    }
  }
};

}  // namespace imgtx

using namespace imgtx;

static int impl_cc_pointwise_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind, double p0,
                                     double p1, const void* draws_a_dev, const void* draws_b_dev, uint64_t seed,
                                     int tex_n, int tex_h, int tex_w, int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  IMGTX_CHECK_ARG(kind >= IMGTX_CC_GAUSSIAN && kind <= IMGTX_CC_FROST, "unknown corruption kind %d", kind);
  if (kind == IMGTX_CC_IMPULSE) IMGTX_CHECK_ARG((draws_a_dev == nullptr) == (draws_b_dev == nullptr), "impulse needs both draw arrays or none");
  if (kind == IMGTX_CC_SHOT || kind == IMGTX_CC_GAUSSIAN) IMGTX_CHECK_ARG(p0 > 0 || draws_a_dev, "scale must be positive");
  size_t extra = 0;
  if (kind == IMGTX_CC_FOG && !draws_a_dev) {
    if (!(h == w && (h & (h - 1)) == 0 && h <= 64)) {
      // other sizes: maps of the next power-of-two side made in global memory, in chunks of at most 256 MB, then the fog pass
      // with those maps as its draws
      IMGTX_CHECK_ARG(p1 > 0, "plasma decay must be positive");
      int msz = 4;
      while (msz < std::max(h, w)) msz <<= 1;
      IMGTX_CHECK_ARG(msz <= 4096, "image too large for the on-chip plasma (side <= 4096)");
      if (n == 0) return IMGTX_OK;
      IMGTX_CHECK_ARG(in && out, "null image pointer");
      cudaStream_t st = (cudaStream_t)stream;
      if ((flags & IMGTX_CC_FAST) && p0 > 0 && p0 < 100) {
        int lg = 2;
        while ((1 << lg) < msz) ++lg;
        const int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * 2);
        float* work = nullptr;
        IMGTX_CUDA_OK(workspace_alloc((void**)&work, (size_t)grid * msz * msz * sizeof(float), st));
        fog_large_fast_kernel<<<(unsigned)grid, 1024, 0, st>>>(in, out, n, h, w, lg, (float)p0, (float)p1, seed, work);
        IMGTX_LAUNCH_OK();
        IMGTX_CUDA_OK(cudaFreeAsync(work, st));
        return IMGTX_OK;
      }
      const size_t per = (size_t)msz * msz * sizeof(double);
      const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(n, (int64_t)((256u << 20) / per)));
      double* maps = nullptr;
      IMGTX_CUDA_OK(workspace_alloc((void**)&maps, per * chunk, st));
      CcPointParams q;
      memset(&q, 0, sizeof(q));
      q.kind = kind; q.p0 = p0; q.p1 = p1; q.a = maps; q.seed = seed; q.plasma_pitch = msz;
      const Shape shl = make_shape(h, w, c, h, w);
      int rc = IMGTX_OK;
      for (int64_t c0 = 0; c0 < n && rc == IMGTX_OK; c0 += chunk) {
        const int64_t cn = std::min<int64_t>(chunk, n - c0);
        plasma_map_kernel<<<(unsigned)cn, 256, 0, st>>>(maps, msz, p1, seed, c0);
        IMGTX_LAUNCH_OK();
        rc = launch_image_op<OpCcPoint>(in + c0 * shl.in_bytes, out + c0 * shl.out_bytes, cn, shl, q, st);
      }
      IMGTX_CUDA_OK(cudaFreeAsync(maps, st));
      return rc;
    }
    extra = (size_t)h * w * sizeof(double);
  }
  if (kind == IMGTX_CC_FROST) {
    IMGTX_CHECK_ARG(draws_a_dev, "frost needs crops or a texture bank");
    if (tex_n > 0) IMGTX_CHECK_ARG(tex_h > h && tex_w > w, "texture smaller than the image");
  }
  const Shape shp = make_shape(h, w, c, h, w);
  const bool wordable = shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out);
  if (wordable) {
    CcFastParams f;
    memset(&f, 0, sizeof(f));
    f.kind = kind; f.p0 = p0; f.p1 = p1; f.a = draws_a_dev; f.b = draws_b_dev; f.seed = seed;
    f.tex_n = kind == IMGTX_CC_FROST ? tex_n : 0; f.tex_h = tex_h; f.tex_w = tex_w;
    f.force_exact = (flags & IMGTX_CC_REPLAY_ALL) ? 1 : 0;
    if (flags & IMGTX_CC_FAST) {
      // throughput forms (cifar_c_team.cuh): one warp per image, float32 / fixed point, <= 1 LSB from the forms below
      int rc = IMGTX_ERR_UNSUPPORTED;
      cudaStream_t st = (cudaStream_t)stream;
      if (kind == IMGTX_CC_CONTRAST && (h * w) % 4 == 0 && p0 >= 0.0 && p0 <= 1.0 && (int64_t)h * w * 255 < (1ll << 24))
        rc = launch_team_op<OpTContrast>(in, out, n, shp, f, st, metrics_dev);
      else if (kind == IMGTX_CC_BRIGHTNESS && (h * w) % 4 == 0 && p0 > -10 && p0 < 10)
        rc = launch_team_op<OpTBrightness>(in, out, n, shp, f, st, metrics_dev);
      else if (kind == IMGTX_CC_IMPULSE && !draws_a_dev && p0 > 0.0 && p0 < 1.0) {
        f.p1 = 1.0 / log2(1.0 - p0);
        rc = launch_team_op<OpTImpulse>(in, out, n, shp, f, st, metrics_dev);
        f.p1 = p1;
      } else if (kind == IMGTX_CC_FROST && w % 4 == 0 && p0 >= 0 && p1 >= 0 && p0 < 1.99 && p1 < 1.99 &&
                 ((reinterpret_cast<uintptr_t>(draws_a_dev) & 3u) == 0))
        rc = launch_team_op<OpTFrost>(in, out, n, shp, f, st, metrics_dev);
      else if (kind == IMGTX_CC_FOG && (h * w) % 4 == 0 && p0 > 0 && p0 < 100 &&
               (draws_a_dev || (h == w && (h & (h - 1)) == 0 && h >= 4 && h <= 64 && p1 > 0)))
        rc = launch_team_op<OpTFog>(in, out, n, shp, f, st, metrics_dev);
      if (rc != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc; }
    }
    if (kind == IMGTX_CC_IMPULSE && !draws_a_dev) return launch_image_op<OpCcImpulseFast>(in, out, n, shp, f, (cudaStream_t)stream);
    if (kind == IMGTX_CC_CONTRAST && (h * w) % 4 == 0 && (int64_t)h * w * 255 < (1ll << 31))
      return launch_image_op<OpCcContrastFast>(in, out, n, shp, f, (cudaStream_t)stream, 0,
                                               128 + 3 * 8 * shp.in_bytes + (int)OpCcContrastFast::scratch_bytes(shp, 8) + 512);   // 8 images per group
    if (kind == IMGTX_CC_BRIGHTNESS && (h * w) % 4 == 0 && p0 > -10 && p0 < 10)
      return launch_image_op<OpCcBrightnessFast>(in, out, n, shp, f, (cudaStream_t)stream);
    if (kind == IMGTX_CC_FOG && !draws_a_dev && h == w && (h & (h - 1)) == 0 && h >= 4 && h <= 64 && p0 > 0 && p0 < 100 && p1 > 0) {
      const int target = 128 + 3 * 4 * shp.in_bytes + (int)OpCcFogFast::scratch_bytes(shp, 4) + 512;  // one image per warp
      return launch_image_op<OpCcFogFast>(in, out, n, shp, f, (cudaStream_t)stream, 0, std::min(target, max_smem_optin() - 2048));
    }
    if (kind == IMGTX_CC_FROST && w % 4 == 0 && p0 >= 0 && p1 >= 0 && p0 <= 1000 && p1 <= 1000 &&
        (tex_n == 0 || ((reinterpret_cast<uintptr_t>(draws_a_dev) & 3u) == 0)))
      return launch_image_op<OpCcFrostFast>(in, out, n, shp, f, (cudaStream_t)stream);
  }
  CcPointParams p;
  memset(&p, 0, sizeof(p));
  p.kind = kind; p.p0 = p0; p.p1 = p1; p.a = draws_a_dev; p.b = draws_b_dev; p.seed = seed;
  p.tex_n = kind == IMGTX_CC_FROST ? tex_n : 0; p.tex_h = tex_h; p.tex_w = tex_w;
  return launch_image_op<OpCcPoint>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream, extra);
}

// ---------------------------------------------------------------------------------------------------------------
// alias tables for the on-chip noise (host, f64)
// ---------------------------------------------------------------------------------------------------------------
namespace {

// Walker/Vose alias table of a discrete distribution; entry = (threshold23 << 9) | alias9
void vose(const std::vector<double>& pr, std::vector<uint32_t>& entries) {
  const int W = (int)pr.size();
  std::vector<double> q(W);
  std::vector<int> alias(W), small, large;
  double tot = 0;
  for (double v : pr) tot += v;
  for (int i = 0; i < W; ++i) {
    q[i] = pr[i] / tot * W;
    alias[i] = i;
    (q[i] < 1.0 ? small : large).push_back(i);
  }
  while (!small.empty() && !large.empty()) {
    const int s = small.back(); small.pop_back();
    const int l = large.back();
    alias[s] = l;
    q[l] = (q[l] + q[s]) - 1.0;
    if (q[l] < 1.0) { large.pop_back(); small.push_back(l); }
  }
  for (int i : large) { q[i] = 1.0; alias[i] = i; }
  for (int i : small) { q[i] = 1.0; alias[i] = i; }   // round-off leftovers
  for (int i = 0; i < W; ++i) {
    double t = floor(q[i] * 8388608.0 + 0.5);
    uint32_t thr = t >= 8388607.0 ? 8388607u : (t <= 0 ? 0u : (uint32_t)t);
    int a = alias[i];
    if (q[i] >= 1.0) a = i;  // "always keep": both branches give i
    entries.push_back((thr << 9) | (uint32_t)a);
  }
}

struct AliasTable {
  std::vector<uint32_t> rows, entries;
  std::vector<uint8_t> out;
};

// gaussian_noise: the increment of the output byte is floor(255*sigma*Z): P(d) = Phi((d+1)/s) - Phi(d/s), |d| <= 6.5 s
bool build_gaussian(double sigma, AliasTable& t) {
  const double s = sigma * 255.0;
  if (!(s > 0)) return false;
  const int K = (int)ceil(6.5 * s) + 1;
  if (2 * K > 512) return false;
  std::vector<double> pr(2 * K);
  const double inv = 1.0 / (s * sqrt(2.0));
  for (int d = 0; d < K; ++d) {
    const double pd = 0.5 * (erfc(d * inv) - erfc((d + 1) * inv));
    pr[K + d] = pd;        // delta d
    pr[K - 1 - d] = pd;    // delta -d-1
  }
  // padded with zero-probability outcomes to 512 slots: the device picks the slot with a shift instead of a 32 x 32 -> 64
  // multiply (IMAD.WIDE is the scarcest pipe of the noise kernels)
  pr.resize(512, 0.0);
  t.rows = {0u, (512u << 16) | (uint32_t)(uint16_t)(int16_t)(-K)};
  vose(pr, t.entries);
  return true;
}

// shot_noise: row x is the law of uint8(clip(Poisson(x/255*c)/c, 0, 1)*255); Poisson mass below 1e-10 is dropped
bool build_shot(double c, AliasTable& t) {
  if (!(c > 0)) return false;
  for (int x = 0; x < 256; ++x) {
    const double lam = (double)x / 255.0 * c;
    std::vector<double> pr;
    std::vector<uint8_t> ov;
    auto outv = [&](long long k) {
      double v = (double)k / c;
      v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
      return (uint8_t)(int)(v * 255.0);
    };
    if (lam <= 0) {
      pr.push_back(1.0); ov.push_back(outv(0));
    } else {
      const double ll = log(lam);
      auto pmf = [&](long long k) { return exp((double)k * ll - lam - lgamma((double)k + 1.0)); };
      long long mode = (long long)floor(lam), klo = mode, khi = mode;
      while (klo > 0 && pmf(klo - 1) > 1e-10) --klo;
      while (pmf(khi + 1) > 1e-10) ++khi;
      for (long long k = klo; k <= khi; ++k) {
        const uint8_t o = outv(k);
        if (!ov.empty() && ov.back() == o) pr.back() += pmf(k);
        else { ov.push_back(o); pr.push_back(pmf(k)); }
      }
    }
    if (pr.size() > 512 || t.out.size() + ov.size() > 65535) return false;
    t.rows.push_back((uint32_t)t.entries.size());
    t.rows.push_back(((uint32_t)pr.size() << 16) | (uint32_t)t.out.size());
    vose(pr, t.entries);
    t.out.insert(t.out.end(), ov.begin(), ov.end());
  }
  return true;
}

// shot_noise, flat layout (IMGTX_CC_SHOT_FLAT): the same rows, each padded with zero-probability outcomes to K slots
bool build_shot_flat(double c, std::vector<uint32_t>& words) {
  if (!(c > 0)) return false;
  std::vector<std::vector<double>> prs(256);
  std::vector<std::vector<uint8_t>> ovs(256);
  size_t widest = 1;
  for (int x = 0; x < 256; ++x) {
    const double lam = (double)x / 255.0 * c;
    std::vector<double>& pr = prs[x];
    std::vector<uint8_t>& ov = ovs[x];
    auto outv = [&](long long k) {
      double v = (double)k / c;
      v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
      return (uint8_t)(int)(v * 255.0);
    };
    if (lam <= 0) {
      pr.push_back(1.0); ov.push_back(outv(0));
    } else {
      const double ll = log(lam);
      auto pmf = [&](long long k) { return exp((double)k * ll - lam - lgamma((double)k + 1.0)); };
      long long mode = (long long)floor(lam), klo = mode, khi = mode;
      while (klo > 0 && pmf(klo - 1) > 1e-10) --klo;
      while (pmf(khi + 1) > 1e-10) ++khi;
      for (long long k = klo; k <= khi; ++k) {
        const uint8_t o = outv(k);
        if (!ov.empty() && ov.back() == o) pr.back() += pmf(k);
        else { ov.push_back(o); pr.push_back(pmf(k)); }
      }
    }
    widest = std::max(widest, pr.size());
  }
  int lg = 1;
  while ((size_t)1 << lg < widest) ++lg;
  if (lg > 7) return false;   // 256 x 128 words = 128 KB is the most that leaves room for the images
  const int K = 1 << lg;
  words.assign(kAliasHeaderWords, 0u);
  words[0] = kAliasMagicFlat; words[1] = (uint32_t)IMGTX_CC_SHOT; words[2] = 256u; words[3] = 256u * (uint32_t)K; words[5] = (uint32_t)lg;
  for (int x = 0; x < 256; ++x) {
    std::vector<double> pr = prs[x];
    pr.resize(K, 0.0);
    std::vector<uint32_t> e;
    vose(pr, e);   // (threshold23 << 9) | alias9
    for (int i = 0; i < K; ++i) {
      const uint32_t a = e[i] & 511u, t23 = e[i] >> 9;
      uint32_t t16 = t23 >= 8388607u ? 65535u : (t23 + 64u) >> 7;     // own outcome iff (16 random bits) < t16
      if (t16 > 65535u) t16 = 65535u;
      const bool real = (size_t)i < ovs[x].size(), areal = (size_t)a < ovs[x].size();
      // a padded slot has probability 0: its alias is a real outcome (vose pairs it with a large one) and t16 = 0
      uint32_t own = real ? ovs[x][i] : (areal ? ovs[x][a] : ovs[x][0]);
      uint32_t ali = areal ? ovs[x][a] : own;
      if (a == (uint32_t)i) ali = own;                                  // "always keep"
      if (!real) t16 = 0;
      words.push_back((t16 << 16) | (ali << 8) | own);
    }
  }
  return true;
}

bool build_table(int kind, double p0, std::vector<uint32_t>& words) {
  if (kind == IMGTX_CC_SHOT_FLAT) return build_shot_flat(p0, words);
  AliasTable t;
  const bool ok = kind == IMGTX_CC_GAUSSIAN ? build_gaussian(p0, t) : (kind == IMGTX_CC_SHOT ? build_shot(p0, t) : false);
  if (!ok) return false;
  while (t.out.size() % 4) t.out.push_back(0);
  words.assign(kAliasHeaderWords, 0u);
  words[0] = kAliasMagic; words[1] = (uint32_t)kind; words[2] = (uint32_t)(t.rows.size() / 2);
  words[3] = (uint32_t)t.entries.size(); words[4] = (uint32_t)t.out.size();
  words.insert(words.end(), t.rows.begin(), t.rows.end());
  words.insert(words.end(), t.entries.begin(), t.entries.end());
  const size_t base = words.size();
  words.resize(base + t.out.size() / 4);
  if (!t.out.empty()) memcpy(&words[base], t.out.data(), t.out.size());
  return true;
}

}  // namespace

extern "C" int64_t imgtx_cc_noise_table_words(int kind, double p0) {
  std::vector<uint32_t> w;
  return build_table(kind, p0, w) ? (int64_t)w.size() : 0;
}

extern "C" int imgtx_cc_noise_table_host(int kind, double p0, int32_t* table_host, int64_t capacity_words) {
  std::vector<uint32_t> w;
  IMGTX_CHECK_ARG(table_host != nullptr, "null table");
  if (!build_table(kind, p0, w)) {
    set_error("no alias table for kind %d with parameter %g (gaussian: 255*sigma <= 39; shot: <= 512 outcomes per value)", kind, p0);
    return IMGTX_ERR_UNSUPPORTED;
  }
  IMGTX_CHECK_ARG((int64_t)w.size() <= capacity_words, "table needs %lld words", (long long)w.size());
  memcpy(table_host, w.data(), w.size() * 4);
  return IMGTX_OK;
}

static int impl_cc_noise_onchip_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                                        const int32_t* table_dev, int64_t table_words, uint64_t seed,
                                        uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  // (on-chip draws only: this entry point IS a throughput form; it runs on the team pipeline when the table fits beside it)
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "CIFAR-C corruptions take (n,h,w,3) images");
  IMGTX_CHECK_ARG(kind == IMGTX_CC_GAUSSIAN || kind == IMGTX_CC_SHOT || kind == IMGTX_CC_SHOT_FLAT, "alias-table noise is gaussian or shot");
  IMGTX_CHECK_ARG(table_dev != nullptr && table_words > kAliasHeaderWords && aligned16(table_dev), "bad alias table");
  const Shape shp = make_shape(h, w, c, h, w);
  if (kind == IMGTX_CC_SHOT_FLAT) {
    IMGTX_CHECK_ARG(table_words >= kAliasHeaderWords + 512 && ((table_words - kAliasHeaderWords) & (table_words - kAliasHeaderWords - 1)) == 0,
                    "a flat shot table is 8 + 256 K words, K a power of two");
    CcFastParams f;
    memset(&f, 0, sizeof(f));
    f.kind = IMGTX_CC_SHOT; f.b = table_dev; f.seed = seed; f.table_words = (int)table_words;
    const int rc = launch_team_op<OpTShotFlat>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev);
    if (rc == IMGTX_OK) *metrics_done = true;
    return rc;   // IMGTX_ERR_UNSUPPORTED (no error string): the caller falls back to the IMGTX_CC_SHOT table
  }
  IMGTX_CHECK_ARG(shp.in_bytes % 16 == 0 && aligned16(in) && aligned16(out),
                  "image bytes must be a multiple of 16 and the buffers 16-byte aligned");
  const size_t tbytes = (size_t)table_words * 4;
  IMGTX_CHECK_ARG(tbytes + 32 * 1024 <= (size_t)max_smem_optin(), "alias table does not fit in shared memory");
  CcFastParams f;
  memset(&f, 0, sizeof(f));
  f.kind = kind; f.b = table_dev; f.seed = seed; f.table_words = (int)table_words;
  {
    const int rc = tbytes > 48 * 1024 ? launch_team_op<OpTNoiseAlias<2>>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev)
                                      : launch_team_op<OpTNoiseAlias<1>>(in, out, n, shp, f, (cudaStream_t)stream, metrics_dev);
    if (rc != IMGTX_ERR_UNSUPPORTED) { *metrics_done = true; return rc; }
  }
  // small table: default occupancy; large table (shot): one CTA per SM, the rest of shared memory stages images
  int target = 0;
  if (tbytes > 16 * 1024) target = max_smem_optin() - 1024;   // big table (shot noise): one CTA per SM anyway, give it every byte for images
  else target = 96 * 1024;   // small table (gaussian): larger groups measured 5 % faster than the default 56 KB target
  return launch_image_op<OpCcNoiseAlias>(in, out, n, shp, f, (cudaStream_t)stream, tbytes, target);
}

extern "C" int imgtx_cc_pointwise_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind, double p0,
                                     double p1, const void* draws_a_dev, const void* draws_b_dev, uint64_t seed,
                                     int tex_n, int tex_h, int tex_w, int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_cc_pointwise_u8(in, out, n, h, w, c, kind, p0, p1, draws_a_dev, draws_b_dev, seed, tex_n, tex_h, tex_w, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}

extern "C" int imgtx_cc_noise_onchip_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                                        const int32_t* table_dev, int64_t table_words, uint64_t seed,
                                        int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  (void)flags;
  int rc = impl_cc_noise_onchip_u8(in, out, n, h, w, c, kind, table_dev, table_words, seed, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}
