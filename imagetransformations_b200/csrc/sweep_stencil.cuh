// Column-sweep evaluation of a mirror-symmetric (2R+1)x(2R+1) stencil on interleaved u8 images in shared memory, with an
// exact-arithmetic fallback.
//
// The CIFAR-10-C-style blurs are specified in float64 with a fixed summation order (oracle/cifar_c.py), but the result is
// only observed after floor(255*v).  The fast path evaluates the same real-valued sum  t = sum_k w_k * x_k  in 8.32 fixed
// point: weights rounded to 2^-32, one 32 x 32 + 64-bit multiply-add per folded tap, so |t_fixed - t| <= taps*255*2^-33
// < 7.5e-7 output LSB for 5x5 taps (the f64 reference itself is within 1e-12 LSB of t), and accepts floor(t_fixed) whenever
// t_fixed is farther than kBand = 2^-20 LSB from an integer, or below kBand (all terms are non-negative, so the result is
// 0).  The elements inside the band (2e-6 of random data, plus results that are integers by construction) are recomputed by
// the caller's exact functor (the literal f64 loop).  Bit-exactness therefore never depends on the fast path.
// Kernels whose weights are (nearly) reciprocals of small integers — the 1/5 "plus" and 1/9 box of defocus_blur severities 4
// and 5 — put 10-20 % of all results within 1e-6 of an integer; for those the FIX40 variant keeps 2^-40 weights in two 32-bit
// accumulators (error < 2^-28 LSB, band 2^-28 LSB) at twice the multiplies.
//
// Work decomposition: one ITEM = (channel, group of 4 adjacent pixels, row segment).  An item walks down its rows keeping
// the horizontally folded taps of the last 2R+1 rows in registers (5 loads per 4 outputs per row for R = 2), so every input
// byte is read (2R+4)/4 times instead of (2R+1)^2 times.
#pragma once
#include "common.cuh"

namespace imgtx {
#ifdef __CUDACC__

constexpr int kFixShift = 32;              // W = round(w * 2^32)
constexpr uint32_t kBand = 1u << 12;       // 2^-20 LSB  >  25 * 255 * 2^-33 (worst-case weight rounding, 5x5 taps)
constexpr int kFix40HiShift = 24;          // FIX40: Whi = floor(w * 2^24)
constexpr int kFix40LoShift = 16;          //        Wlo = round((w * 2^24 - Whi) * 2^16)
constexpr int kFix40Shift = kFix40HiShift + kFix40LoShift;
constexpr uint64_t kFix40One = 1ull << kFix40Shift;
constexpr uint64_t kBand40 = 1ull << 12;   // 2^-28 LSB  >  25 * 255 * 2^-41

enum { BORDER_REFLECT101 = 0, BORDER_REPLICATE = 1 };

// valid for -n < i < 2n - 1 (one fold), which covers every stencil here (R <= 3 < n)
template <int MODE>
__device__ __forceinline__ int border_index(int i, int n) {
  if (MODE == BORDER_REFLECT101) return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
  return min(max(i, 0), n - 1);
}

// Fold(k): horizontally folded tap k of output o from the 2R+4 loaded bytes v[]: k = 0 centre, k >= 1 the pair at +-k.
template <int R>
__device__ __forceinline__ int hfold(const int* v, int o, int k) {
  return k == 0 ? v[o + R] : v[o + R - k] + v[o + R + k];
}

// im/om: one image, (h, w, 3) interleaved u8 in shared memory (om may not alias im).  Wt[2*(R+1)*(R+1)]: fixed-point weights
// indexed [|dy|][|dx|] (2^-32 words, or for FIX40 hi words then lo words).  Rows [y_begin, y_end) of channel ch, pixels 4*pg .. 4*pg+3.  exact(y, x, ch) -> byte.
template <int R, int BORDER, bool FIX40, class Exact>
__device__ __forceinline__ void sweep_sym_stencil(const uint8_t* im, uint8_t* om, int h, int w,
                                                  const uint32_t* Wt, int ch, int pg, int y_begin, int y_end,
                                                  Exact exact) {
  constexpr int N = 2 * R + 1;   // ring depth
  constexpr int L = 2 * R + 4;   // loaded bytes per row
  const int rowb = w * 3;
  int cidx[L];
#pragma unroll
  for (int i = 0; i < L; ++i) cidx[i] = border_index<BORDER>(4 * pg - R + i, w) * 3 + ch;
  constexpr int NW = (R + 1) * (R + 1);
  uint32_t Wf[NW], Wl[FIX40 ? NW : 1];
#pragma unroll
  for (int i = 0; i < NW; ++i) { Wf[i] = Wt[i]; if (FIX40) Wl[i] = Wt[NW + i]; }

  int ring[N][R + 1][4];
  for (int yb = y_begin - R; yb < y_end + R; yb += N) {
#pragma unroll
    for (int s = 0; s < N; ++s) {
      const int row = yb + s;
      if (row < y_end + R) {
        const uint8_t* rp = im + border_index<BORDER>(row, h) * rowb;
        int v[L];
#pragma unroll
        for (int i = 0; i < L; ++i) v[i] = rp[cidx[i]];
#pragma unroll
        for (int k = 0; k <= R; ++k)
#pragma unroll
          for (int o = 0; o < 4; ++o) ring[s][k][o] = hfold<R>(v, o, k);
        const int y = row - R;
        if (y >= y_begin) {
          // slots: row y+d lives in slot (s - R + d) mod N, all static
          uint32_t pack = 0, need = 0;
#pragma unroll
          for (int o = 0; o < 4; ++o) {
            uint32_t ip;
            if (FIX40) {
              uint32_t th = 0, tl = 0;
#pragma unroll
              for (int i = 0; i <= R; ++i) {
#pragma unroll
                for (int k = 0; k <= R; ++k) {
                  const int sa = ((s - R + i) % N + N) % N, sb = ((s - R - i) % N + N) % N;
                  const uint32_t val = (uint32_t)(i == 0 ? ring[sa][k][o] : ring[sa][k][o] + ring[sb][k][o]);
                  th += Wf[i * (R + 1) + k] * val;
                  tl += Wl[i * (R + 1) + k] * val;
                }
              }
              const uint64_t t = ((uint64_t)th << kFix40LoShift) + tl;
              const uint64_t frac = t & (kFix40One - 1);
              if ((frac < kBand40 && t >= kBand40) || frac > kFix40One - kBand40) need |= 1u << o;
              ip = (uint32_t)(t >> kFix40Shift);
            } else {
              uint64_t t = 0;
#pragma unroll
              for (int i = 0; i <= R; ++i) {
#pragma unroll
                for (int k = 0; k <= R; ++k) {
                  const int sa = ((s - R + i) % N + N) % N, sb = ((s - R - i) % N + N) % N;
                  const uint32_t val = (uint32_t)(i == 0 ? ring[sa][k][o] : ring[sa][k][o] + ring[sb][k][o]);
                  t += (uint64_t)Wf[i * (R + 1) + k] * val;     // IMAD.WIDE.U32 with 64-bit accumulate
                }
              }
              const uint32_t frac = (uint32_t)t;
              ip = (uint32_t)(t >> kFixShift);
              if ((frac < kBand && ip != 0u) || frac > 0u - kBand) need |= 1u << o;
            }
            pack |= min(ip, 255u) << (8 * o);
          }
          if (need) {  // rare: the literal f64 evaluation decides
            for (int o = 0; o < 4; ++o)
              if ((need >> o) & 1u) pack = (pack & ~(255u << (8 * o))) | ((uint32_t)exact(y, 4 * pg + o, ch) << (8 * o));
          }
          uint8_t* op = om + y * rowb + (4 * pg) * 3 + ch;
          op[0] = (uint8_t)pack; op[3] = (uint8_t)(pack >> 8); op[6] = (uint8_t)(pack >> 16); op[9] = (uint8_t)(pack >> 24);
        }
      }
    }
  }
}

// Host: quantise weights to 2^-32, out[0..count); false when they cannot be used (negative, or not below 1).
static inline bool fix_weights(const double* w, int count, const int* multiplicity, uint32_t* out) {
  double total = 0;
  for (int i = 0; i < count; ++i) {
    if (!(w[i] >= 0.0) || !(w[i] < 1.0)) return false;
    const double q = floor(w[i] * 4294967296.0 + 0.5);
    if (q > 4294967295.0) return false;
    out[i] = (uint32_t)q;
    out[count + i] = 0u;
    total += q * multiplicity[i];
  }
  return total * 255.0 < 1.8e19;   // the 64-bit accumulator cannot overflow
}

// Host: quantise weights to 2^-40 as (hi, lo) words, out[0..count) = hi, out[count..2*count) = lo; false when they cannot
// be used (negative, or an accumulator could overflow 32 bits).
static inline bool fix_weights40(const double* w, int count, const int* multiplicity, uint32_t* out) {
  double total_hi = 0, total_lo = 0;
  for (int i = 0; i < count; ++i) {
    if (!(w[i] >= 0.0) || w[i] > 1.0) return false;
    const double s = w[i] * (double)(1u << kFix40HiShift);
    const double hi = floor(s);
    double lo = floor((s - hi) * (double)(1u << kFix40LoShift) + 0.5);
    out[i] = (uint32_t)hi;
    out[count + i] = (uint32_t)lo;
    total_hi += hi * multiplicity[i];
    total_lo += lo * multiplicity[i];
  }
  return total_hi * 255.0 < 4294967295.0 && total_lo * 255.0 < 4294967295.0;
}

#endif
}  // namespace imgtx
