// CIFAR-10-C-style pointwise corruptions, throughput ("fast") forms on the team pipeline (team_pipeline.cuh).
//
// Contract of the fast mode (IMGTX_CC_FAST): float32 / fixed-point evaluation of the SAME real-valued expression the exact
// mode evaluates in float64, so every byte is within 1 LSB of the exact mode on the same inputs and draws (the north-star
// tolerance for uint8 pixels; tests/test_gpu_fast_mode.py measures the differing fraction per type).  No float64, no replay
// queues, no per-image LUT builds: one warp owns one image, reads it once from its staged copy and overwrites it in place.
//
//   contrast    q = x*c + (1-c)*mean_ch (0..255 scale), one FFMA per byte; floor by a round-down add of 2^23
//   brightness  HSV round trip == x * K[max(r,g,b)], K = 255*clip(max/255 + c)/max, one table lookup per pixel
//   impulse     geometric skip sampling: only the ~p*3072 flipped bytes are touched (on-chip draws only)
//   frost       (C0*x + C1*f) >> 15 with Q15 constants: IDP.2A, two bytes per instruction pair
//   gaussian / shot (on-chip)  Walker alias tables in shared memory (see cifar_c_fast.cuh), one warp per image
//   fog         warp-serial diamond-square plasma in float32 + one FFMA per byte
#pragma once
#include "team_pipeline.cuh"
#include "cifar_c_fast.cuh"

namespace imgtx {

// ---------------------------------------------------------------------------------------------------------------
struct OpTContrast : NoTeamPrepare {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = 1, kInBufs = 4, kOutBufs = 0, kMaxTeams = 16, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }

  __device__ static __forceinline__ uint32_t map4(uint32_t w, float c, float ba, float bb, float bc, float bd) {
    return pack4(floor_bits(__fmaf_rn(byte_f(w, 0), c, ba)), floor_bits(__fmaf_rn(byte_f(w, 1), c, bb)),
                 floor_bits(__fmaf_rn(byte_f(w, 2), c, bc)), floor_bits(__fmaf_rn(byte_f(w, 3), c, bd)));
  }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t, const Shape& s, const Params& p, uint8_t*,
                             const uint8_t*, int lane, int) {
    const uint32_t* im = reinterpret_cast<const uint32_t*>(src);
    uint32_t* om = reinterpret_cast<uint32_t*>(dst);
    const int triples = s.in_bytes / 12;   // 4 pixels = 3 words (h*w % 4 == 0 checked at launch)
    // byte sums per position of the 12-byte period, packed 2 x 16 bit (<= 255 * 2 per step: flushed every 128 steps)
    uint32_t c0 = 0, c1 = 0, c2 = 0;
    uint32_t e0 = 0, o0 = 0, e1 = 0, o1 = 0, e2 = 0, o2 = 0;
    int since = 0;
    for (int t = lane; t < triples; t += 32) {
      const uint32_t w0 = im[3 * t], w1 = im[3 * t + 1], w2 = im[3 * t + 2];
      e0 += w0 & 0x00FF00FFu; o0 += (w0 >> 8) & 0x00FF00FFu;
      e1 += w1 & 0x00FF00FFu; o1 += (w1 >> 8) & 0x00FF00FFu;
      e2 += w2 & 0x00FF00FFu; o2 += (w2 >> 8) & 0x00FF00FFu;
      if (++since == 128 || t + 32 >= triples) {
        // bytes of the period: w0 = r g b r | w1 = g b r g | w2 = b r g b
        c0 += (e0 & 0xFFFFu) + (o0 >> 16) + (e1 >> 16) + (o2 & 0xFFFFu);
        c1 += (o0 & 0xFFFFu) + (e1 & 0xFFFFu) + (o1 >> 16) + (e2 >> 16);
        c2 += (e0 >> 16) + (o1 & 0xFFFFu) + (e2 & 0xFFFFu) + (o2 >> 16);
        e0 = o0 = e1 = o1 = e2 = o2 = 0;
        since = 0;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      c0 += __shfl_xor_sync(0xffffffffu, c0, o);
      c1 += __shfl_xor_sync(0xffffffffu, c1, o);
      c2 += __shfl_xor_sync(0xffffffffu, c2, o);
    }
    // (x/255 - m)*c + m, times 255  ==  x*c + (1 - c)*M with M = sum/hw: a convex combination, so 0 <= q <= 255
    const float c = (float)p.p0, k = __fdiv_rn(1.0f - c, (float)(s.h * s.w));
    const float br = __fmul_rn((float)c0, k), bg = __fmul_rn((float)c1, k), bb = __fmul_rn((float)c2, k);
    __syncwarp();
    for (int t = lane; t < triples; t += 32) {
      const uint32_t w0 = im[3 * t], w1 = im[3 * t + 1], w2 = im[3 * t + 2];
      om[3 * t] = map4(w0, c, br, bg, bb, br);
      om[3 * t + 1] = map4(w1, c, bg, bb, br, bg);
      om[3 * t + 2] = map4(w2, c, bb, br, bg, bb);
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
struct OpTBrightness {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = 1, kInBufs = 4, kOutBufs = 0, kMaxTeams = 16, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  static size_t cta_scratch(const Shape&, const Params&) { return 256 * 8; }
  // {K, E}[M]: out = x*K + E.  K = 255*v2/M with v2 = clip(M/255 + c, 0, 1); E = 2^-14 keeps x == M on floor(255*v2)
  // when x*K rounds just below it; M == 0 (a black pixel becomes grey 255*clip(c)) is carried entirely by E.
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    float2* KE = reinterpret_cast<float2*>(cta);
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
      double v2 = (double)i / 255.0 + p.p0;
      v2 = v2 < 0.0 ? 0.0 : (v2 > 1.0 ? 1.0 : v2);
      const float eps = 6.103515625e-05f;
      KE[i] = i == 0 ? make_float2(0.0f, (float)(255.0 * v2) + eps) : make_float2((float)(255.0 * v2 / (double)i), eps);
    }
  }
  __device__ static __forceinline__ uint32_t pixel(uint32_t px, const float2* KE) {
    const float r = byte_f(px, 0), g = byte_f(px, 1), b = byte_f(px, 2);
    const float M = fmaxf(fmaxf(r, g), b);
    const float2 ke = KE[__float_as_uint(__fadd_rn(M, 8388608.0f)) & 255u];
    const uint32_t a0 = floor_bits(__fmaf_rn(r, ke.x, ke.y)), a1 = floor_bits(__fmaf_rn(g, ke.x, ke.y)),
                   a2 = floor_bits(__fmaf_rn(b, ke.x, ke.y));
    return __byte_perm(__byte_perm(a0, a1, 0x0040u), a2, 0x7410u) & 0x00FFFFFFu;
  }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t, const Shape& s, const Params&, uint8_t*,
                             const uint8_t* cta, int lane, int) {
    const float2* KE = reinterpret_cast<const float2*>(cta);
    const uint32_t* im = reinterpret_cast<const uint32_t*>(src);
    uint32_t* om = reinterpret_cast<uint32_t*>(dst);
    const int quads = s.in_bytes / 12;
    for (int t = lane; t < quads; t += 32) {
      const uint32_t w0 = im[3 * t], w1 = im[3 * t + 1], w2 = im[3 * t + 2];
      const uint32_t o0 = pixel(w0, KE), o1 = pixel(__funnelshift_r(w0, w1, 24), KE), o2 = pixel(__funnelshift_r(w1, w2, 16), KE),
                     o3 = pixel(w2 >> 8, KE);
      uint32_t a, b, c;
      pack_rgb4(o0, o1, o2, o3, a, b, c);
      om[3 * t] = a; om[3 * t + 1] = b; om[3 * t + 2] = c;
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// impulse_noise, on-chip draws: every element flips independently with probability p, so the gaps between flipped
// elements are geometric: gap = floor(log(u) / log(1 - p)).  Each lane walks its own 1/32 of the image and touches only the
// bytes that flip (salt or pepper from an independent bit); the image is otherwise left as staged.
struct OpTImpulse : NoTeamPrepare {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = 1, kInBufs = 4, kOutBufs = 0, kMaxTeams = 16, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  __device__ static void run(const uint8_t*, uint8_t* dst, int64_t img, const Shape& s, const Params& p, uint8_t*,
                             const uint8_t*, int lane, int) {
    const Philox rng(p.seed);
    const int seg = (s.in_bytes + 31) >> 5;
    const int end = min(s.in_bytes, (lane + 1) * seg);
    int pos = lane * seg - 1;
    const float inv = (float)p.p1;   // 1 / log2(1 - p), set by the entry point
    uint32_t blk = 0;
    while (pos < end) {
      const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x32u + (uint32_t)lane, blk++);
      const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        // u in (0, 1]: log2(u) <= 0; gap >= 0
        const float gap = floorf(__log2f(u01(rr[e])) * inv);
        pos += 1 + (int)fminf(gap, 1.0e6f);
        if (pos < end) dst[pos] = (rr[e] & 1u) ? 255 : 0;
      }
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// frost: floor(clip(c0*x + c1*f, 0, 255)) with c0, c1 in Q15 (0 <= error < 0.016 LSB): one IDP.2A per byte.
struct OpTFrost : NoTeamPrepare {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = 1, kInBufs = 4, kOutBufs = 0, kMaxTeams = 16, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  __device__ static __forceinline__ uint32_t mix4(uint32_t w, uint32_t f, uint32_t cc) {
    const uint32_t lo = __byte_perm(w, f, 0x5140u), hi = __byte_perm(w, f, 0x7362u);
    const uint32_t a0 = min(__dp2a_lo(cc, lo, 0u) >> 15, 255u), a1 = min(__dp2a_hi(cc, lo, 0u) >> 15, 255u);
    const uint32_t a2 = min(__dp2a_lo(cc, hi, 0u) >> 15, 255u), a3 = min(__dp2a_hi(cc, hi, 0u) >> 15, 255u);
    return pack4(a0, a1, a2, a3);
  }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape& s, const Params& p, uint8_t*,
                             const uint8_t*, int lane, int) {
    const uint32_t* im = reinterpret_cast<const uint32_t*>(src);
    uint32_t* om = reinterpret_cast<uint32_t*>(dst);
    const int wpi = s.in_bytes >> 2;
    // Q15 constants rounded UP: sums that are exact integers in real arithmetic (c1 = 0.3 with f = 10, 20, ...) stay at or above
    // the integer, where the float64 expression of the exact mode puts them; the price is a one-sided error < 0.016 LSB
    const uint32_t cc = (uint32_t)__double2int_ru(p.p0 * 32768.0 - 1e-9) | ((uint32_t)__double2int_ru(p.p1 * 32768.0 - 1e-9) << 16);
    if (p.tex_n == 0) {
      const uint32_t* fr = reinterpret_cast<const uint32_t*>(p.a) + (size_t)img * wpi;
      for (int i = lane; i < wpi; i += 32) om[i] = mix4(im[i], __ldg(fr + i), cc);
      return;
    }
    const Philox rng(p.seed);
    const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x50u, 0u);
    const int ti = (int)(r.x % (uint32_t)p.tex_n);
    const int y0 = (int)(r.y % (uint32_t)(p.tex_h - s.h)), x0 = (int)(r.z % (uint32_t)(p.tex_w - s.w));
    const uint8_t* tex = reinterpret_cast<const uint8_t*>(p.a) + ((size_t)ti * p.tex_h + y0) * p.tex_w * 3 + (size_t)x0 * 3;
    const int tex_pitch = p.tex_w * 3, row_words = (s.w * 3) >> 2;
    if (row_words == 24 && (tex_pitch & 3) == 0 && wpi == 768) {
      // 32 x 32 images, texture rows a whole number of words: the crop's misalignment is the same on every row, so a lane
      // owns three fixed word columns (j, j + 8, j + 16) of every fourth row and walks down with one pointer increment
      const int rr = lane >> 3, jj = lane & 7;
      const uintptr_t a0 = reinterpret_cast<uintptr_t>(tex + (size_t)rr * tex_pitch + (size_t)jj * 4);
      const uint32_t sh = (uint32_t)(a0 & 3) * 8;
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(a0 & ~(uintptr_t)3);
      const int step = tex_pitch;                 // four rows = 4 * pitch bytes = pitch words
      const uint32_t* xi = im + rr * 24 + jj;
      uint32_t* xo = om + rr * 24 + jj;
#pragma unroll 4
      for (int m = 0; m < 8; ++m) {
        const uint32_t f0 = __funnelshift_r(__ldg(wp), __ldg(wp + 1), sh);
        const uint32_t f1 = __funnelshift_r(__ldg(wp + 8), __ldg(wp + 9), sh);
        const uint32_t f2 = __funnelshift_r(__ldg(wp + 16), __ldg(wp + 17), sh);
        xo[0] = mix4(xi[0], f0, cc);
        xo[8] = mix4(xi[8], f1, cc);
        xo[16] = mix4(xi[16], f2, cc);
        wp += step; xi += 96; xo += 96;
      }
      return;
    }
    const FastDiv rd(row_words);
    // the texture words come from L2: eight independent fetches are issued before the first one is consumed
    for (int i0 = lane; i0 < wpi; i0 += 256) {
      uint32_t f[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = min(i0 + 32 * u, wpi - 1);
        int y, xw;
        rd.divmod(i, y, xw);
        const uintptr_t a = reinterpret_cast<uintptr_t>(tex + (size_t)y * tex_pitch + (size_t)xw * 4);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        f[u] = __funnelshift_r(__ldg(wp), __ldg(wp + 1), (uint32_t)(a & 3) * 8);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + 32 * u;
        if (i < wpi) om[i] = mix4(im[i], f[u], cc);
      }
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// gaussian_noise / shot_noise with on-chip draws: the alias tables of cifar_c_fast.cuh, shared by the teams of a CTA.
template <int TW>
struct OpTNoiseAlias {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = TW, kInBufs = 3, kOutBufs = 0, kMaxTeams = TW == 1 ? 16 : 12, kCtasPerSm = TW == 1 ? 2 : 1,
                       kMinBlocks = TW == 1 ? 2 : 1;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  static size_t cta_scratch(const Shape&, const Params& p) { return (size_t)p.table_words * 4; }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(p.b);
    uint32_t* dst = reinterpret_cast<uint32_t*>(cta);
    const uint4* s4 = reinterpret_cast<const uint4*>(src);   // the table is 16-byte aligned (checked at launch)
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    const int nv = p.table_words >> 2;
    for (int i = threadIdx.x; i < nv; i += blockDim.x) d4[i] = __ldg(s4 + i);
    for (int i = 4 * nv + threadIdx.x; i < p.table_words; i += blockDim.x) dst[i] = __ldg(src + i);
  }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape& s, const Params& p, uint8_t*,
                             const uint8_t* cta, int tt, int) {
    const AliasView av = alias_view(cta);
    const Philox rng(p.seed);
    const int wpi = s.in_bytes >> 2;
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
    uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
    const uint64_t word0 = (uint64_t)img * (uint64_t)wpi;
    if (p.kind == IMGTX_CC_GAUSSIAN) {
      const uint2 row = av.rows[0];
      const int dlo = (int)(int16_t)(row.y & 0xFFFFu);
      for (int i = tt; i < wpi; i += TW * 32) {
        const uint32_t w = s32[i];
        const uint64_t gw = word0 + (uint64_t)i;
        const uint4 r = rng((uint32_t)gw, (uint32_t)(gw >> 32), 0x11u, 0u);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
        uint32_t o = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int x = (int)((w >> (8 * b)) & 255u);
          const int d = dlo + (int)((row.y >> 16) == 512u ? alias_draw512(av.entries, row.x, rr[b]) : alias_draw(av.entries, row, rr[b]));
          o |= (uint32_t)min(max(x + d, 0), 255) << (8 * b);
        }
        d32[i] = o;
      }
    } else {
      for (int i = tt; i < wpi; i += TW * 32) {
        const uint32_t w = s32[i];
        const uint64_t gw = word0 + (uint64_t)i;
        const uint4 r = rng((uint32_t)gw, (uint32_t)(gw >> 32), 0x21u, 0u);
        const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
        uint32_t o = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const uint2 row = av.rows[(w >> (8 * b)) & 255u];
          o |= (uint32_t)av.out[(row.y & 0xFFFFu) + alias_draw(av.entries, row, rr[b])] << (8 * b);
        }
        d32[i] = o;
      }
    }
  }
};

// shot_noise with on-chip draws from the FLAT table (cifar_c_fast.cuh): one lookup per element.  The slot is the top log2 K
// bits of the element's Philox word, the keep / alias decision its low 16 bits; the entry carries both output bytes.
// Table in shared memory once per CTA (one CTA per SM: 64 KB at c = 60), two-warp teams beside it.
struct OpTShotFlat {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = 2, kInBufs = 3, kOutBufs = 0, kMaxTeams = 15, kCtasPerSm = 1, kMinBlocks = 1;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  static size_t cta_scratch(const Shape&, const Params& p) { return (size_t)(p.table_words - kAliasHeaderWords) * 4 + 16; }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(p.b);
    uint32_t* dst = reinterpret_cast<uint32_t*>(cta);
    const int ne = p.table_words - kAliasHeaderWords;   // 256 K words, after a 32-byte header: 16-byte vectors
    // (every CTA reads the same 64-128 KB from L2: with 4-byte loads the copy took ~10 % of the pass, ncu r02zzz3)
    const uint4* s4 = reinterpret_cast<const uint4*>(src + kAliasHeaderWords);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    for (int i = threadIdx.x; i < (ne >> 2); i += blockDim.x) d4[i] = __ldg(s4 + i);
    if (threadIdx.x == 0) dst[ne] = __ldg(src + 5);   // log2 K
  }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape& s, const Params& p, uint8_t*,
                             const uint8_t* cta, int tt, int) {
    const int ne = p.table_words - kAliasHeaderWords;
    const uint32_t lg = reinterpret_cast<const uint32_t*>(cta)[ne];
    const uint32_t tab = (uint32_t)__cvta_generic_to_shared(cta);
    const Philox rng(p.seed);
    const int wpi = s.in_bytes >> 2;
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
    uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
    const uint64_t word0 = (uint64_t)img * (uint64_t)wpi;
    const uint32_t rsh = 30u - lg, smask = ((1u << lg) - 1u) << 2, xsh = lg + 2u;   // byte offsets: (x << lg | slot) * 4
    for (int i = tt; i < wpi; i += kTeamWarps * 32) {
      const uint32_t w = s32[i];
      const uint64_t gw = word0 + (uint64_t)i;
      const uint4 r = rng((uint32_t)gw, (uint32_t)(gw >> 32), 0x21u, 0u);
      const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
      uint32_t e[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t x = (w >> (8 * b)) & 255u;
        const uint32_t a = tab + ((x << xsh) | ((rr[b] >> rsh) & smask));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e[b]) : "r"(a));
      }
      uint32_t o[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) o[b] = (rr[b] & 0xFFFFu) < (e[b] >> 16) ? e[b] : (e[b] >> 8);
      d32[i] = pack4(o[0], o[1], o[2], o[3]);
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// fog: per-image byte maximum, plasma map (diamond-square drawn on chip, or injected float64 maps read from HBM), then
// q = x*S + plasma*C in float32 with S = m/(m + c0), C = 255*c0*S (m = max/255).  |error| < 2e-4 LSB.
//
// On-chip plasma for 2^LG x 2^LG maps, one warp, every level constant compile-time: the points of a step are dealt flat to
// the lanes (point -> (a, b) by shifts), both diamond families are the same "average of the four axis neighbours at distance
// s/2" with wrap-around, and all 2^(2 LG) - 1 uniforms of a lane's points come from five Philox calls made up front (16 bits
// per uniform, planted in a float mantissa and folded into the noise FFMA).  Leaves the UN-normalised map and returns its
// minimum and 1 / (max - min).
constexpr int plasma_iters(int LG, int ls_from) {   // lane-iterations of levels ls_from .. 1
  int t = 0;
  for (int ls = ls_from; ls >= 1; --ls) { const int mm = 1 << (2 * (LG - ls)); t += (mm + 31) / 32 + (2 * mm + 31) / 32; }
  return t;
}

// one level (step 2^LS) of the diamond-square recursion; ITN0 = index of this level's first 16-bit uniform in the lane's pool
template <int LG, int LS, int ITN0, int POOL>
__device__ __forceinline__ void plasma_level(float* map, const uint32_t (&pool)[POOL], float wib2, int lane) {
  constexpr int N = 1 << LG, MASK = N - 1, s = 1 << LS, half = s >> 1, lm = LG - LS, m = 1 << lm, mm = m * m;
  constexpr int T1 = (mm + 31) / 32, T2 = (2 * mm + 31) / 32;
  // uniform in (-1, 1) from 16 bits h: (h + 1/2)/32768 - 1; with hf = 2^23 + h:  noise = wib2*u = hf*k1 + k0
  const float k1 = wib2 * (1.0f / 32768.0f), k0 = wib2 * (0.5f / 32768.0f - 1.0f) - 8388608.0f * k1;
#pragma unroll
  for (int t = 0; t < T1; ++t) {
    constexpr int dummy = 0; (void)dummy;
    const int itn = ITN0 + t, p = t * 32 + lane;
    const float hf = __uint_as_float(__byte_perm(pool[itn >> 1], 0x4B000000u, (itn & 1) ? 0x7432u : 0x7410u));
    if (p < mm) {
      const int a = p >> lm, b = p & (m - 1);
      const int y0 = a << LS, x0 = b << LS, y1 = (y0 + s) & MASK, x1 = (x0 + s) & MASK;
      const float acc = (map[y0 * N + x0] + map[y1 * N + x0]) + (map[y0 * N + x1] + map[y1 * N + x1]);
      map[(y0 + half) * N + x0 + half] = __fmaf_rn(acc, 0.25f, __fmaf_rn(hf, k1, k0));
    }
  }
  __syncwarp();
#pragma unroll
  for (int t = 0; t < T2; ++t) {
    const int itn = ITN0 + T1 + t, p = t * 32 + lane;
    const float hf = __uint_as_float(__byte_perm(pool[itn >> 1], 0x4B000000u, (itn & 1) ? 0x7432u : 0x7410u));
    if (p < 2 * mm) {
      const int fam = p >> (2 * lm), k = p & (mm - 1), a = k >> lm, b = k & (m - 1);
      const int y = (a << LS) + (fam ? half : 0), x = (b << LS) + (fam ? 0 : half);
      const float acc = (map[((y - half) & MASK) * N + x] + map[((y + half) & MASK) * N + x]) +
                        (map[y * N + ((x - half) & MASK)] + map[y * N + ((x + half) & MASK)]);
      map[y * N + x] = __fmaf_rn(acc, 0.25f, __fmaf_rn(hf, k1, k0));
    }
  }
  __syncwarp();
}

template <int LG, int LS, int POOL>
__device__ __forceinline__ void plasma_levels(float* map, const uint32_t (&pool)[POOL], float wib2, float idec2, int lane) {
  plasma_level<LG, LS, plasma_iters(LG, LG) - plasma_iters(LG, LS), POOL>(map, pool, wib2, lane);
  if constexpr (LS > 1) plasma_levels<LG, LS - 1, POOL>(map, pool, wib2 * idec2, idec2, lane);
}

template <int LG>
__device__ __forceinline__ void plasma_flat(float* map, float decay, const Philox& rng, uint32_t img_lo, uint32_t img_hi,
                                            float& mn_out, float& inv_span_out) {
  constexpr int N = 1 << LG;
  const int lane = threadIdx.x & 31;
  constexpr int kCalls = (plasma_iters(LG, LG) + 7) / 8;
  uint32_t pool[kCalls * 4];
#pragma unroll
  for (int c = 0; c < kCalls; ++c) {
    const uint4 r = rng(img_lo, img_hi, 0x78u, (uint32_t)(c * 32 + lane));
    pool[4 * c] = r.x; pool[4 * c + 1] = r.y; pool[4 * c + 2] = r.z; pool[4 * c + 3] = r.w;
  }
  if (lane == 0) map[0] = 0.0f;
  __syncwarp();
  plasma_levels<LG, LG, kCalls * 4>(map, pool, 10000.0f, __fdiv_rn(1.0f, decay * decay), lane);
  float mn = 3.0e38f, mx = -3.0e38f;
#pragma unroll 4
  for (int i = lane; i < N * N / 4; i += 32) {
    const float4 v = *reinterpret_cast<const float4*>(map + 4 * i);
    mn = fminf(fminf(mn, v.x), fminf(v.y, fminf(v.z, v.w)));
    mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  mn_out = mn;
  inv_span_out = __fdiv_rn(1.0f, mx - mn);
}

struct OpTFog : NoTeamPrepare {
  using Params = CcFastParams;
  static constexpr int kTeamWarps = 1, kInBufs = 3, kOutBufs = 0, kMaxTeams = 8, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t team_scratch(const Shape& s, const Params&) { return (size_t)s.h * s.w * 4; }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape& s, const Params& p, uint8_t* scr,
                             const uint8_t*, int lane, int) {
    float* map = reinterpret_cast<float*>(scr);
    const uint32_t* im = reinterpret_cast<const uint32_t*>(src);
    uint32_t* om = reinterpret_cast<uint32_t*>(dst);
    const int hw = s.h * s.w, triples = s.in_bytes / 12;
    uint32_t mx4 = 0;
    for (int i = lane; i < (s.in_bytes >> 2); i += 32) mx4 = __vmaxu4(mx4, im[i]);
    uint32_t mxb = max(max(mx4 & 255u, (mx4 >> 8) & 255u), max((mx4 >> 16) & 255u, mx4 >> 24));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mxb = max(mxb, __shfl_xor_sync(0xffffffffu, mxb, o));
    const float mval = __fdiv_rn((float)mxb, 255.0f), c0 = (float)p.p0;
    const float S = __fdiv_rn(mval, mval + c0);
    float C = 255.0f * c0 * S, bias = 0.0f;
    if (p.a) {
      const double* pl = reinterpret_cast<const double*>(p.a) + (size_t)img * hw;
      for (int i = lane; i < hw; i += 32) map[i] = (float)pl[i];
      __syncwarp();
    } else {
      const Philox rng(p.seed);
      float mn, inv;
      if (s.h == 32) {
        plasma_flat<5>(map, (float)p.p1, rng, (uint32_t)img, (uint32_t)(img >> 32), mn, inv);
      } else {
        OpCcFogFast::plasma_warp(map, 31 - __clz(s.h), (float)p.p1, rng, (uint32_t)img, (uint32_t)(img >> 32));
        mn = 0.0f; inv = 1.0f;
      }
      // (map - mn) * inv * C  ==  map * C' + bias
      C *= inv;
      bias = -mn * C;
    }
    for (int t = lane; t < triples; t += 32) {   // 4 pixels = 3 words
      const uint32_t w0 = im[3 * t], w1 = im[3 * t + 1], w2 = im[3 * t + 2];
      const float4 pl = *reinterpret_cast<const float4*>(map + 4 * t);
      const float a = __fmaf_rn(pl.x, C, bias), b = __fmaf_rn(pl.y, C, bias), c = __fmaf_rn(pl.z, C, bias), d = __fmaf_rn(pl.w, C, bias);
      // no clip: 0 <= q <= 255*m*(1 + 2^-20) (x <= 255 m, plasma <= 1), so floor(q) always fits the byte
      auto q = [S](uint32_t w, int k, float add) { return floor_bits(fmaxf(__fmaf_rn(byte_f(w, k), S, add), 0.0f)); };
      om[3 * t] = pack4(q(w0, 0, a), q(w0, 1, a), q(w0, 2, a), q(w0, 3, b));
      om[3 * t + 1] = pack4(q(w1, 0, b), q(w1, 1, b), q(w1, 2, c), q(w1, 3, c));
      om[3 * t + 2] = pack4(q(w2, 0, c), q(w2, 1, d), q(w2, 2, d), q(w2, 3, d));
    }
  }
};

}  // namespace imgtx
