// CIFAR-10-C-style stencil corruptions, throughput ("fast") forms on the team pipeline (team_pipeline.cuh).
// Contract: see cifar_c_team.cuh (float32 evaluation of the exact mode's real-valued expression, <= 1 LSB per byte).
#pragma once
#include "team_pipeline.cuh"
#include "rng.cuh"
#include <cuda_fp16.h>

namespace imgtx {

// ---- u8 HWC image -> fp16 RGBx canvas (8 bytes per pixel; u8 is exact in fp16) ------------------------------------
// half(1024 + b) has the bit pattern 0x6400 | b: plant two bytes, subtract 1024 (exact).
__device__ __forceinline__ uint32_t bytes_to_half2(uint32_t a, uint32_t b, uint32_t sel) {
  const uint32_t v = __byte_perm(__byte_perm(a, b, sel), 0x64646464u, 0x7160u);   // (b0, 0x64, b1, 0x64)
  const __half2 r = __hsub2(*reinterpret_cast<const __half2*>(&v), __float2half2_rn(1024.0f));
  return *reinterpret_cast<const uint32_t*>(&r);
}
// four pixels (three words) -> four uint2 {(r, g), (b, 0)}
__device__ __forceinline__ void quad_to_half4(uint32_t w0, uint32_t w1, uint32_t w2, uint2* o) {
  o[0] = make_uint2(bytes_to_half2(w0, w0, 0x0010u), bytes_to_half2(w0, w0, 0x0002u) & 0xFFFFu);
  o[1] = make_uint2(bytes_to_half2(w0, w1, 0x0043u), bytes_to_half2(w1, w1, 0x0001u) & 0xFFFFu);
  o[2] = make_uint2(bytes_to_half2(w1, w1, 0x0032u), bytes_to_half2(w2, w2, 0x0000u) & 0xFFFFu);
  o[3] = make_uint2(bytes_to_half2(w2, w2, 0x0021u), bytes_to_half2(w2, w2, 0x0003u) & 0xFFFFu);
}

// ---------------------------------------------------------------------------------------------------------------
// motion_blur: out = floor(sum_i k_i * src[clamp(y + dy_i)][clamp(x + dx_i)] + 1/2), the line offsets (dx_i, dy_i) shared by
// all pixels of an image.  32-pixel-wide images only (lane = column).  The image is expanded into an fp16 RGBx canvas with
// `padv` replicated rows above and below, so the vertical clamp disappears and the address of (tap, row) is
// base_tap[lane] + row*256: the inner loop is one 8-byte load, three conversions and three FFMA, no index arithmetic.
// A team is four warps (8 rows each); 24 float accumulators per lane.
struct MotionFastParams {
  int width, n_eff, padv;
  float k[32];
  const int32_t* offsets;   // device (n, width, 2) (dx, dy) or nullptr: angle drawn on chip in [lo, hi] degrees
  float lo, hi;
  uint64_t seed;
};

struct OpTMotion : NoTeamPrepare {
  using Params = MotionFastParams;
  static constexpr int kTeamWarps = 4, kInBufs = 3, kOutBufs = 0, kMaxTeams = 4, kCtasPerSm = 2, kMinBlocks = 2;
  static constexpr int kTT = kTeamWarps * 32, kRows = 32 / kTeamWarps;
  static size_t team_scratch(const Shape&, const Params& p) { return 512 + (size_t)(32 + 2 * p.padv) * 256; }

  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape&, const Params& p, uint8_t* scr,
                             const uint8_t*, int tt, int team) {
    int2* offs = reinterpret_cast<int2*>(scr);          // (dx, dy) per tap
    uint8_t* cv = scr + 512;                             // [32 + 2 padv][32] x 8 B
    const int lane = tt & 31, wq = tt >> 5;
    if (tt < p.n_eff) {
      if (p.offsets) {
        const int32_t* o = p.offsets + ((size_t)img * p.width + tt) * 2;
        offs[tt] = make_int2(o[0], o[1]);
      } else {
        const Philox rng(p.seed);
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x60u, 0u);
        const float a = (p.lo + (p.hi - p.lo) * u01(r.x)) * 0.017453292519943295f;
        float sn, cs;
        sincosf(a, &sn, &cs);
        offs[tt] = make_int2((int)ceilf((float)tt * cs - 0.5f), (int)ceilf((float)tt * sn - 0.5f));
      }
    }
    {
      const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
      uint2* c2 = reinterpret_cast<uint2*>(cv + p.padv * 256);
      for (int q = tt; q < 256; q += kTT) quad_to_half4(s32[3 * q], s32[3 * q + 1], s32[3 * q + 2], c2 + 4 * q);
    }
    team_sync<kTeamWarps>(team);
    {
      // (expanding the replicated rows straight from the source instead, to save this barrier, measured 7 % slower)
      uint4* c4 = reinterpret_cast<uint4*>(cv);
      for (int i = tt; i < 2 * p.padv * 16; i += kTT) {
        const int r = i >> 4, c = i & 15;
        const int drow = r < p.padv ? r : r + 32, srow = r < p.padv ? p.padv : p.padv + 31;
        c4[drow * 16 + c] = c4[srow * 16 + c];
      }
    }
    team_sync<kTeamWarps>(team);
    float acc[kRows][3];
#pragma unroll
    for (int r = 0; r < kRows; ++r) acc[r][0] = acc[r][1] = acc[r][2] = 0.5f;
    const uint8_t* base = cv + (p.padv + kRows * wq) * 256;
    int2 o_next = offs[0];
    float k_next = p.k[0];
    for (int i = 0; i < p.n_eff; ++i) {
      const int2 o = o_next;
      const float k = k_next;
      o_next = offs[min(i + 1, p.n_eff - 1)];   // the next tap's offsets are fetched while this one runs
      k_next = p.k[min(i + 1, p.n_eff - 1)];
      const uint8_t* a = base + min(max(o.y, -p.padv), p.padv) * 256 + min(max(lane + o.x, 0), 31) * 8;
#pragma unroll
      for (int r = 0; r < kRows; ++r) {
        const uint2 v = *reinterpret_cast<const uint2*>(a + r * 256);
        const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
        const float b = __low2float(*reinterpret_cast<const __half2*>(&v.y));
        acc[r][0] = __fmaf_rn(rg.x, k, acc[r][0]);
        acc[r][1] = __fmaf_rn(rg.y, k, acc[r][1]);
        acc[r][2] = __fmaf_rn(b, k, acc[r][2]);
      }
    }
    // floor(acc) (the 1/2 is already in): 0.5 <= acc < 256
    uint8_t* o = dst + (kRows * wq) * 96 + lane * 3;
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) o[r * 96 + c] = (uint8_t)floor_bits(acc[r][c]);
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// Mirror-symmetric 3x3 stencil on a 32x32x3 image in float32: out = K00*c + K01*(l + r) + K10*(u + d) + K11*(four diagonals)
// + 2^-12 (callers normalise the weights to sum to 1, so a flat region keeps its value instead of flooring one below it).
// Serves defocus_blur (every CIFAR severity's disk kernel is 3x3 up to < 2e-3 LSB of weight) and glass_blur's Gaussians.
// Work item = (word column c of 24, segment of 8 rows): the thread walks down its rows with a three-row register window;
// per row it loads three words, forms the four left+right byte sums with packed 16-bit adds, and converts 4 centres +
// 4 sums to float (byte / halfword planted in the mantissa of 2^23).  BORDER: 0 = REFLECT_101, 1 = replicate.
struct Sym3x3 { float k00, k01, k10, k11; };

template <int BORDER>
__device__ __forceinline__ void sym3x3_load_row(const uint32_t* row, int c, float* C, float* H) {
  const uint32_t wc = row[c], wl = row[max(c - 1, 0)], wr = row[min(c + 1, 23)];
  uint32_t L = __funnelshift_r(wl, wc, 8), R = __funnelshift_r(wc, wr, 24);
  if (c == 0) L = BORDER == 0 ? __byte_perm(wc, wr, 0x0543u) : __byte_perm(wc, wc, 0x0210u);      // bytes (3,4,5,0) / (0,1,2,0)
  if (c == 23) R = BORDER == 0 ? __byte_perm(wl, wc, 0x4327u) : __byte_perm(wc, wc, 0x3213u);     // bytes (95,90,91,92) / (95,93,94,95)
  const uint32_t he = (L & 0x00FF00FFu) + (R & 0x00FF00FFu), ho = ((L >> 8) & 0x00FF00FFu) + ((R >> 8) & 0x00FF00FFu);
  H[0] = __uint_as_float(__byte_perm(he, 0x4B000000u, 0x7410u)) - 8388608.0f;
  H[2] = __uint_as_float(__byte_perm(he, 0x4B000000u, 0x7432u)) - 8388608.0f;
  H[1] = __uint_as_float(__byte_perm(ho, 0x4B000000u, 0x7410u)) - 8388608.0f;
  H[3] = __uint_as_float(__byte_perm(ho, 0x4B000000u, 0x7432u)) - 8388608.0f;
#pragma unroll
  for (int j = 0; j < 4; ++j) C[j] = byte_f(wc, j);
}

// rows [y0, y0 + 8) of word column c: src, dst = 32 x 24 words
template <int BORDER>
__device__ __forceinline__ void sym3x3_segment(const uint32_t* src, uint32_t* dst, int c, int y0, const Sym3x3& k) {
  float C[3][4], H[3][4];
  auto brow = [](int y) { return BORDER == 0 ? (y < 0 ? -y : (y > 31 ? 62 - y : y)) : min(max(y, 0), 31); };
  sym3x3_load_row<BORDER>(src + brow(y0 - 1) * 24, c, C[0], H[0]);
  sym3x3_load_row<BORDER>(src + brow(y0) * 24, c, C[1], H[1]);
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    // window slots rotate statically: row y0 + r - 1 in slot r % 3, y0 + r in (r + 1) % 3, y0 + r + 1 in (r + 2) % 3
    const int a = r % 3, b = (r + 1) % 3, d = (r + 2) % 3;
    sym3x3_load_row<BORDER>(src + brow(y0 + r + 1) * 24, c, C[d], H[d]);
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float v = __fmaf_rn(k.k11, H[a][j] + H[d][j], __fmaf_rn(k.k10, C[a][j] + C[d][j], __fmaf_rn(k.k01, H[b][j], __fmaf_rn(k.k00, C[b][j], 0.000244140625f))));
      o[j] = floor_bits(v);
    }
    dst[(y0 + r) * 24 + c] = pack4(o[0], o[1], o[2], o[3]);
  }
}

// defocus_blur (fast): K from the disk taps, REFLECT_101
struct OpTDefocus : NoTeamPrepare {
  using Params = Sym3x3;
  static constexpr int kTeamWarps = 3, kInBufs = 3, kOutBufs = 2, kMaxTeams = 5, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t, const Shape&, const Params& p, uint8_t*, const uint8_t*,
                             int tt, int) {
    sym3x3_segment<0>(reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(dst), tt % 24, (tt / 24) * 8, p);
  }
};

// ---------------------------------------------------------------------------------------------------------------
// glass_blur (32x32, fast): Gaussian -> 8 bits -> sequential local pixel swaps -> Gaussian.  A team is four warps: warps 0-2
// (24 word columns x 4 row segments) run the two Gaussians as 3x3 float32 stencils (sigma <= 0.4: the taps at distance 2
// carry < 4e-3 LSB and are dropped), warp 3 meanwhile runs the swap sweep of the same image on a u16 index map (the swaps
// never look at pixel values: skew-2 wavefront, lane = row, 88 steps per iteration), then everybody applies the map as a
// gather.  Every fourth image all four warps sweep (the maps of the team's next four images).  sigma < 0.125 (severity 1): both Gaussians are byte maps, composed into one 256-entry LUT with the float64 formula.
struct GlassTeamParams {
  Sym3x3 k;
  int lw, iters;
  double w0;                 // lw == 0: the single Gaussian weight (LUT build)
  const int8_t* deltas;      // device (n, iters, 30, 30, 2) or nullptr (on-chip draws)
  uint64_t seed;
  int64_t n;                 // images in the batch (maps are made ahead: never past the end)
  int sweeper;               // on-chip draws: 1 = warp 3 of a team only sweeps (see OpTGlass::run), 0 = grouped sweeps
};

struct OpTGlass {
  using Params = GlassTeamParams;
  static constexpr int kTeamWarps = 4, kInBufs = 3, kOutBufs = 0, kMaxTeams = 4, kCtasPerSm = 2, kMinBlocks = 2;
  static size_t cta_scratch(const Shape&, const Params&) { return 256; }
  static size_t team_scratch(const Shape&, const Params&) { return 3072 + 3072 + 4 * 2052 + 256 + 16; }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    // lw == 0: x -> uint8(g*255) -> uint8(clip(g')*255) with g = (x/255*w0)*w0 evaluated as the exact mode does
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
      const double t = (double)i / 255.0;
      const double a = (t * p.w0) * p.w0;
      const int u = (int)(a * 255.0);
      const double t2 = (double)u / 255.0;
      double b = (t2 * p.w0) * p.w0;
      b = b < 0.0 ? 0.0 : (b > 1.0 ? 1.0 : b);
      cta[i] = (uint8_t)(int)(b * 255.0);
    }
  }
  // the swap sweeps of image `img` on the index map m (one warp; skew-2 wavefront, lane = row, 88 steps per iteration)
  __device__ static void sweep(uint16_t* m, int64_t img, const Params& p, int lane) {
    for (int px = lane; px < 1024; px += 32) m[px] = (uint16_t)px;
    __syncwarp();
    const Philox rng(p.seed);
    for (int it = 0; it < p.iters; ++it) {
      const int h = 31 - lane;
      if (p.deltas) {
        for (int t = 0; t < 2 * 29 + 30; ++t) {
          const int ci = t - 2 * lane;
          if (lane < 30 && ci >= 0 && ci < 30) {
            const int w = 31 - ci;
            const int8_t* d = p.deltas + ((((img * p.iters + it) * 30 + lane) * 30 + ci) << 1);
            const int a = h * 32 + w, b = (h + d[1]) * 32 + (w + d[0]);
            const uint16_t ta = m[a], tb = m[b];
            m[a] = tb; m[b] = ta;
          }
          __syncwarp();
        }
      } else {
        // on-chip draws: one bit per swap and axis, consumed from two 64-bit shift registers (one Philox call per row)
        uint64_t BX = 0, BY = 0;
        if (lane < 30) {
          const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)lane);
          BX = (uint64_t)r.x | ((uint64_t)r.z << 32); BY = (uint64_t)r.y | ((uint64_t)r.w << 32);
        }
        uint16_t* ma = m + h * 32 + 31;
        const uint32_t first = 2u * (uint32_t)lane, cnt = lane < 30 ? 30u : 0u;
        for (int t = 0; t < 2 * 29 + 30; ++t) {
          if ((uint32_t)t - first < cnt) {
            uint16_t* mb = ma - ((int)((uint32_t)BY & 1u) * 32 + (int)((uint32_t)BX & 1u));
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
  // On-chip draws, TWO images per warp.  In the skew-2 wavefront row r is busy during steps [2r, 2r + 30), so at most 15 rows
  // of an image are in flight at any step: 15 lanes per image are all the parallelism there is.  Lanes 0-14 sweep image `img0`
  // and lanes 16-30 image `img0 + stride`; lane l takes row l during steps [2l, 2l + 30) and row l + 15 during
  // [2l + 30, 2l + 60) (row 15 starts two steps after row 14, as the skew wants), 88 steps in all as before.  The loop body
  // is branch-free: a lane that has not started or has finished swaps inside a small dummy region.  Three phases of the
  // step counter have one kind of event each (start / switch to the second row / finish), so a step costs one compare and
  // three selects on top of the swap.  Same draws as `sweep` (one Philox call per image, iteration and row; column ci uses
  // bit ci of words x and y), hence the same maps.  About 17 instructions per step for two images instead of 25 for one.
  __device__ static __forceinline__ void swap_step(uint32_t& a, uint32_t& X, uint32_t& Y) {
    const uint32_t b = a - (X & 2u) - (Y & 64u);
    uint16_t ta, tb;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(ta) : "r"(a) : "memory");
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(tb) : "r"(b) : "memory");
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(tb) : "memory");
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(b), "h"(ta) : "memory");
    a -= 2u;
    X >>= 1;
    Y = __funnelshift_r(Y, Y, 1);
    __syncwarp();
  }
  __device__ static void sweep_pair(uint16_t* maps, uint16_t* dummy, int64_t img0, int64_t stride, const Params& p, int lane,
                                    int it_lo, int it_hi, bool init) {
    if (img0 >= p.n) return;
    const int half = lane >> 4, l = lane & 15;
    const int64_t img = img0 + half * stride;
    uint16_t* m = maps + half * 1026;   // the next ring slot (map_slot): one word past 2 KB, so the two halves use disjoint banks
    if (init) {
      for (int i = l; i < 512; i += 16) reinterpret_cast<uint32_t*>(m)[i] = (uint32_t)(2 * i) * 0x00010001u + 0x00010000u;   // identity
    }
    __syncwarp();
    const bool valid = img < p.n && l < 15;
    const uint32_t dm = (uint32_t)__cvta_generic_to_shared(dummy + 100);
    const uint32_t row1 = (uint32_t)__cvta_generic_to_shared(m + (31 - l) * 32 + 31), row2 = row1 - 15u * 64u;
    const int first = valid ? 2 * l : 1000;
    const Philox rng(p.seed);
    for (int it = it_lo; it < it_hi; ++it) {
      uint32_t X1 = 0, Y1 = 0, X2 = 0, Y2 = 0;
      if (valid) {
        const uint4 r1 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)l);
        const uint4 r2 = rng((uint32_t)img, (uint32_t)(img >> 32), 0x40u + it, (uint32_t)(l + 15));
        // X: bit ci at position 1 after ci shifts (byte offset 2 = one entry); Y: bit ci at position 6 after ci rotations
        // (byte offset 64 = one map row)
        X1 = (r1.x & 0x3FFFFFFFu) << 1; Y1 = __funnelshift_l(r1.y & 0x3FFFFFFFu, r1.y & 0x3FFFFFFFu, 6);
        X2 = (r2.x & 0x3FFFFFFFu) << 1; Y2 = __funnelshift_l(r2.y & 0x3FFFFFFFu, r2.y & 0x3FFFFFFFu, 6);
      }
      uint32_t a = dm, X = 0, Y = 0;
#pragma unroll 6
      for (int t = 0; t < 30; ++t) {
        const bool ev = t == first;
        a = ev ? row1 : a; X = ev ? X1 : X; Y = ev ? Y1 : Y;
        swap_step(a, X, Y);
      }
#pragma unroll 6
      for (int t = 30; t < 60; ++t) {
        const bool ev = t == first + 30;
        a = ev ? row2 : a; X = ev ? X2 : X; Y = ev ? Y2 : Y;
        swap_step(a, X, Y);
      }
#pragma unroll 7
      for (int t = 60; t < 88; ++t) {
        const bool ev = t == first + 60;
        a = ev ? dm : a;   // X has run out; whatever Y holds keeps the partner inside the dummy region
        swap_step(a, X, Y);
      }
    }
  }
  // ring slot s of the index maps; the slot pitch is 1024 entries + one 4-byte word, so that the lanes of the two images of a
  // sweep_pair (slots 2k, 2k + 1) fall into disjoint shared-memory banks (lane l: bank 15 l + const, resp. + 1)
  __device__ static __forceinline__ uint16_t* map_slot(uint16_t* maps, int s) { return maps + s * 1026; }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape&, const Params& p, uint8_t* scr,
                             const uint8_t* cta, int tt, int team) {
    uint8_t* B1 = scr;
    uint8_t* B2 = scr + 3072;
    uint16_t* maps = reinterpret_cast<uint16_t*>(scr + 6144);    // four index maps: this image and the team's next three
    uint16_t* dummy = maps + 4 * 1026 + 8;                       // 128 entries where idle lanes of sweep_pair swap
    const int lane = tt & 31, wq = tt >> 5;
    // The sweep is a latency chain (88 dependent steps per iteration) that does not look at the pixels: the map of an image
    // depends on its index and the seed only, so maps are made ahead of the images in a ring of four.
    //   injected deltas      every fourth image warp q makes the map of the team's image it + q
    //   on-chip, grouped     every fourth image warps 0 and 1 sweep two images each (sweep_pair), warps 2 and 3 wait
    //   on-chip, sweeper     warp 3 only sweeps: while warps 0-2 run the two Gaussians and the gather of images it, it + 1 (their
    //                        own 96-thread barrier), it makes the maps of images it + 2, it + 3 — half of the iterations during
    //                        each of the two images — and meets the others at the end-of-image barrier of the pipeline
    const int64_t teams_per_cta = blockDim.x >> 7, stride = (int64_t)gridDim.x * teams_per_cta;
    const int64_t done = img - ((int64_t)blockIdx.x * teams_per_cta + team);
    const int it = (uint64_t)p.n >> 31 ? (int)(done / stride) : (int)((uint32_t)done / (uint32_t)stride);
    int nt = 128;
    if (p.deltas) {
      if ((it & 3) == 0) {
        if (img + wq * stride < p.n) sweep(map_slot(maps, wq), img + wq * stride, p, lane);
        team_sync<4>(team);
      }
    } else if (!p.sweeper) {
      if ((it & 3) == 0) {
        if (wq < 2) sweep_pair(map_slot(maps, 2 * wq), dummy, img + 2 * wq * stride, stride, p, lane, 0, p.iters, true);
        team_sync<4>(team);
      }
    } else {
      // the warps of a team sit on the four schedulers of the SM in order, so the sweeper is warp (team mod 4): every
      // scheduler gets one latency-bound sweeper per four teams instead of one scheduler getting all of them
      const int h = (p.iters + 1) >> 1, sw = team & 3;
      if (it == 0) {
        if (wq == sw) sweep_pair(maps, dummy, img, stride, p, lane, 0, p.iters, true);
        team_sync<4>(team);
      }
      if (wq == sw) {
        if ((it & 1) == 0) sweep_pair(map_slot(maps, (it + 2) & 3), dummy, img + 2 * stride, stride, p, lane, 0, h, true);
        else sweep_pair(map_slot(maps, (it + 1) & 3), dummy, img + stride, stride, p, lane, h, p.iters, false);
        return;
      }
      nt = 96;
      if (wq > sw) tt -= 32;   // the 96 computing threads, numbered 0..95
    }
    auto csync = [&]() {   // the threads that compute this image
      if (nt == 128) team_sync<4>(team);
      else asm volatile("bar.sync %0, 96;" ::"r"(team + 9) : "memory");
    };
    const uint16_t* m = map_slot(maps, it & 3);
    if (p.lw > 0) {
      if (tt < 96) sym3x3_segment<1>(reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(B1), tt % 24, (tt / 24) * 8, p.k);
      csync();
    }
    // ---- apply the map: four pixels per thread ----
    {
      const uint8_t* from = p.lw > 0 ? B1 : src;
      uint32_t* o = reinterpret_cast<uint32_t*>(B2);
      for (int q = tt; q < 256; q += nt) {
        uint32_t px[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const uint32_t v = load_rgb(from, (int)m[4 * q + e] * 3);
          px[e] = p.lw > 0 ? v : ((uint32_t)cta[v & 255u] | ((uint32_t)cta[(v >> 8) & 255u] << 8) | ((uint32_t)cta[v >> 16] << 16));
        }
        uint32_t a, b, c;
        pack_rgb4(px[0], px[1], px[2], px[3], a, b, c);
        o[3 * q] = a; o[3 * q + 1] = b; o[3 * q + 2] = c;
      }
    }
    csync();
    if (p.lw > 0) {
      if (tt < 96) sym3x3_segment<1>(reinterpret_cast<const uint32_t*>(B2), reinterpret_cast<uint32_t*>(dst), tt % 24, (tt / 24) * 8, p.k);
    } else {
      const uint32_t* f = reinterpret_cast<const uint32_t*>(B2);
      uint32_t* o = reinterpret_cast<uint32_t*>(dst);
      for (int i = tt; i < 768; i += nt) o[i] = f[i];
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// snow (32x32): noise layer -> centre zoom -> threshold -> 8 bits -> line blur -> blend.  A team is two warps, lane = column,
// warp = 16 rows.  The layer up to its 8-bit quantisation follows the exact mode operation for operation (a threshold
// decision that flipped would move a byte by far more than 1 LSB); the blur and the blend are float32 FFMA on the 0..255
// scale.  The quantised layer lives in a float canvas with the rows the line can reach replicated above and below, so a
// (tap, row) is one 4-byte load and one FFMA.
struct SnowFastParams {
  MotionFastParams m;        // layer blur: weights, offsets / angle range, seed
  int pad_top, pad_bot;      // replicated canvas rows
  float thresh, c6, one_minus_c6;
  float loc, scale;          // on-chip layer draws: normal(loc, scale)
  const double* layer;       // device (n, 32, 32) f64 or nullptr
  int8_t y0[32], y1[32], x0[32], x1[32];
  float ty[32], tx[32];
};

struct OpTSnow {
  using Params = SnowFastParams;
  static constexpr int kTeamWarps = 2, kInBufs = 3, kOutBufs = 0, kMaxTeams = 6, kCtasPerSm = 2, kMinBlocks = 2;
  // CTA: ty[32] tx[32] f32 | y0 y1 x0 x1 [32] int
  static size_t cta_scratch(const Shape&, const Params&) { return 256 + 512; }
  // team: offsets (256 B) | L0 f32[1024] (later: blurred layer LB) | canvas f32 [32 + pads][32]
  static size_t team_scratch(const Shape&, const Params& p) { return 256 + 4096 + (size_t)(32 + p.pad_top + p.pad_bot) * 128; }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    float* t = reinterpret_cast<float*>(cta);
    int* idx = reinterpret_cast<int*>(cta + 256);
    for (int i = threadIdx.x; i < 32; i += blockDim.x) {
      t[i] = p.ty[i]; t[32 + i] = p.tx[i];
      idx[i] = p.y0[i]; idx[32 + i] = p.y1[i]; idx[64 + i] = p.x0[i]; idx[96 + i] = p.x1[i];
    }
  }
  __device__ static void run(const uint8_t* src, uint8_t* dst, int64_t img, const Shape&, const Params& p, uint8_t* scr,
                             const uint8_t* cta, int tt, int team) {
    const float* tyv = reinterpret_cast<const float*>(cta);
    const float* txv = tyv + 32;
    const int* idx = reinterpret_cast<const int*>(cta + 256);
    int2* offs = reinterpret_cast<int2*>(scr);
    float* L0 = reinterpret_cast<float*>(scr + 256);
    float* cv = L0 + 1024;
    const int lane = tt & 31, wq = tt >> 5;
    const Philox rng(p.m.seed);
    if (tt < p.m.n_eff) {
      if (p.m.offsets) {
        const int32_t* o = p.m.offsets + ((size_t)img * p.m.width + tt) * 2;
        offs[tt] = make_int2(o[0], o[1]);
      } else {
        const uint4 r = rng((uint32_t)img, (uint32_t)(img >> 32), 0x62u, 0u);
        const float a = (p.m.lo + (p.m.hi - p.m.lo) * u01(r.x)) * 0.017453292519943295f;
        float sn, cs;
        sincosf(a, &sn, &cs);
        offs[tt] = make_int2((int)ceilf((float)tt * cs - 0.5f), (int)ceilf((float)tt * sn - 0.5f));
      }
    }
    // A. noise layer (float64 draws cast to float32, like the exact mode)
    if (p.layer) {
      const double* l = p.layer + (size_t)img * 1024;
      for (int i = tt; i < 1024; i += 64) L0[i] = (float)l[i];
    } else {
      for (int i = tt; i < 256; i += 64) {
        const uint64_t e = (uint64_t)img * 256u + (uint64_t)i;
        const uint4 r = rng((uint32_t)e, (uint32_t)(e >> 32), 0x61u, 1u);
        float z0, z1, z2, z3;
        box_muller_fast(r.x, r.y, z0, z1);
        box_muller_fast(r.z, r.w, z2, z3);
        *reinterpret_cast<float4*>(L0 + 4 * i) = make_float4(__fmaf_rn(p.scale, z0, p.loc), __fmaf_rn(p.scale, z1, p.loc),
                                                             __fmaf_rn(p.scale, z2, p.loc), __fmaf_rn(p.scale, z3, p.loc));
      }
    }
    team_sync<2>(team);
    // B. centre zoom (order 1), threshold, 8 bits -> canvas rows pad_top .. pad_top + 31 (the exact mode's operation order)
    {
      const float tx = txv[lane], ux = __fsub_rn(1.0f, tx);
      const int xa = idx[64 + lane], xb = idx[96 + lane];
#pragma unroll 4
      for (int r = 0; r < 16; ++r) {
        const int y = 16 * wq + r;
        const float ty = tyv[y], uy = __fsub_rn(1.0f, ty);
        const int ya = idx[y] * 32, yb = idx[32 + y] * 32;
        float acc = __fmul_rn(__fmul_rn(L0[ya + xa], uy), ux);
        acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L0[ya + xb], uy), tx));
        acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L0[yb + xa], ty), ux));
        acc = __fadd_rn(acc, __fmul_rn(__fmul_rn(L0[yb + xb], ty), tx));
        if (acc < p.thresh) acc = 0.0f;
        acc = fminf(fmaxf(acc, 0.0f), 1.0f);   // (a NaN draw becomes 0, as the float -> int conversion made it)
        // trunc(255 acc) as a float, 0 <= 255 acc <= 255: floor by a round-down add of 2^23 (no F2I / I2F: 16 lanes/clk)
        const float q8 = __fsub_rn(__fadd_rd(__fmul_rn(acc, 255.0f), 8388608.0f), 8388608.0f);
        cv[(p.pad_top + y) * 32 + lane] = q8;
        // the first and the last row are replicated above / below by the lanes that hold them (no extra pass, no barrier)
        if (y == 0) for (int k = 0; k < p.pad_top; ++k) cv[k * 32 + lane] = q8;
        if (y == 31) for (int k = 0; k < p.pad_bot; ++k) cv[(p.pad_top + 32 + k) * 32 + lane] = q8;
      }
    }
    team_sync<2>(team);
    // C. line blur of the 8-bit layer, rounded back to 8 bits -> LB (aliases L0)
    {
      float acc[16];
#pragma unroll
      for (int r = 0; r < 16; ++r) acc[r] = 0.5f;
      const float* base = cv + (p.pad_top + 16 * wq) * 32;
      for (int i = 0; i < p.m.n_eff; ++i) {
        const int2 o = offs[i];
        const float k = p.m.k[i];
        const float* a = base + min(max(o.y, -p.pad_top), p.pad_bot) * 32 + min(max(lane + o.x, 0), 31);
#pragma unroll
        for (int r = 0; r < 16; ++r) acc[r] = __fmaf_rn(a[r * 32], k, acc[r]);
      }
#pragma unroll
      for (int r = 0; r < 16; ++r) L0[(16 * wq + r) * 32 + lane] = __fsub_rn(__fadd_rd(acc[r], 8388608.0f), 8388608.0f);
    }
    team_sync<2>(team);
    // D. blend: c6*x + (1 - c6)*max(x, 1.5*gray + 127.5) + layer + rotated layer, four pixels (three words) per step
    {
      const uint32_t* im = reinterpret_cast<const uint32_t*>(src);
      uint32_t* om = reinterpret_cast<uint32_t*>(dst);
      for (int q = tt; q < 256; q += 64) {
        const uint32_t w0 = im[3 * q], w1 = im[3 * q + 1], w2 = im[3 * q + 2];
        const float4 sl = *reinterpret_cast<const float4*>(L0 + 4 * q), sr = *reinterpret_cast<const float4*>(L0 + 1020 - 4 * q);
        const float add[4] = {sl.x + sr.w, sl.y + sr.z, sl.z + sr.y, sl.w + sr.x};
        const uint32_t px[4] = {w0, __funnelshift_r(w0, w1, 24), __funnelshift_r(w1, w2, 16), w2 >> 8};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float r = byte_f(px[e], 0), g = byte_f(px[e], 1), b = byte_f(px[e], 2);
          const float gray = __fmaf_rn(b, 0.114f, __fmaf_rn(g, 0.587f, r * 0.299f));
          const float whit = __fmaf_rn(gray, 1.5f, 127.5f);
          const float vr = fminf(__fmaf_rn(p.one_minus_c6, fmaxf(r, whit), __fmaf_rn(p.c6, r, add[e])), 255.0f);
          const float vg = fminf(__fmaf_rn(p.one_minus_c6, fmaxf(g, whit), __fmaf_rn(p.c6, g, add[e])), 255.0f);
          const float vb = fminf(__fmaf_rn(p.one_minus_c6, fmaxf(b, whit), __fmaf_rn(p.c6, b, add[e])), 255.0f);
          o[e] = __byte_perm(__byte_perm(floor_bits(vr), floor_bits(vg), 0x0040u), floor_bits(vb), 0x7410u) & 0x00FFFFFFu;
        }
        uint32_t a, b, c;
        pack_rgb4(o[0], o[1], o[2], o[3], a, b, c);
        om[3 * q] = a; om[3 * q + 1] = b; om[3 * q + 2] = c;
      }
    }
  }
};

}  // namespace imgtx
