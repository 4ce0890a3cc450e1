// Pillow ImagingResample (Image.resize) on uint8 HWC images, fused with the integer crop in front of it
// and the crop / zero-pad paste behind it.  Two passes (horizontal first), Q22 integer coefficients,
// uint8 intermediate — bit-exact with libImaging/Resample.c.
//   reference call sites: transformation.py:173-196 (apply_scale: LANCZOS + centre crop / pad),
//   fall_2025/tta_initial_claudeFixed.py:56-62 and fall_2025/transformations_code:43-48 (rand_crop:
//   crop 24x24 + BICUBIC resize to 32x32), fall_2025/TTA_transforms.py:104-113 (Resize(36|40) BILINEAR +
//   CenterCrop(32)), CIFAR-10-C pixelate (BOX down/up).
// Coefficient tables depend only on (in_size, out_size, filter): they are built on the host in double
// exactly as Pillow does (imgtx_resample_table_host) and uploaded once by the caller.
// Roofline: nominally HBM (in_bytes + out_bytes per image) but the 2x(up to 7-tap) integer stencil makes
// this pass instruction-bound; see DESIGN.md.
#include "sep_mma.cuh"
#include <math.h>
#include <string.h>
#include <vector>
#include <algorithm>

namespace imgtx {

// table layout (int32 words): [0]=out_size, [1]=ksize, then out_size records of (xmin, n, k[ksize])
struct ResampleParams {
  const int32_t* htab;  // device; nullptr = horizontal pass skipped (same width)
  const int32_t* vtab;  // device; nullptr = vertical pass skipped
  int htab_words, vtab_words;
  int crop_x, crop_y, crop_w, crop_h;
  const int32_t* crop_xy;  // device (n,2) per-image crop origin, nullable
  int rw, rh;              // resized size
  int place_x, place_y;    // where the resized image's origin lands in the output canvas
};

struct OpResample {
  using Params = ResampleParams;
  static constexpr int kThreads = 256;
  // needed part of the resized image: columns [cx0, cx0+ncols) x rows [ry0, ry0+nrows)
  __host__ __device__ static void needed(const Shape& s, const Params& p, int& cx0, int& ncols, int& ry0, int& nrows) {
    cx0 = p.place_x < 0 ? -p.place_x : 0;
    int cx1 = p.rw < s.ow - p.place_x ? p.rw : s.ow - p.place_x;
    ncols = cx1 > cx0 ? cx1 - cx0 : 0;
    ry0 = p.place_y < 0 ? -p.place_y : 0;
    int ry1 = p.rh < s.oh - p.place_y ? p.rh : s.oh - p.place_y;
    nrows = ry1 > ry0 ? ry1 - ry0 : 0;
  }
  // scratch: [tables (htab+vtab words)] [tmp: crop_h * ncols * c bytes]
  static size_t table_bytes(const Params& p) { return align_up((size_t)(p.htab_words + p.vtab_words) * 4, 16); }
  static size_t tmp_bytes(const Shape& s, const Params& p) {
    int cx0, ncols, ry0, nrows;
    needed(s, p, cx0, ncols, ry0, nrows);
    return align_up((size_t)p.crop_h * (ncols > 0 ? ncols : 1) * s.c + 4, 16);
  }
  // Params are not available to scratch_bytes(); the launcher below sizes scratch itself.
  static size_t gtmp_bytes(const Shape&) { return 0; }
  static size_t scratch_bytes_spilled(const Shape&) { return 0; }

  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    int32_t* t = reinterpret_cast<int32_t*>(scratch);
    for (int i = threadIdx.x; i < p.htab_words; i += kThreads) t[i] = p.htab[i];
    for (int i = threadIdx.x; i < p.vtab_words; i += kThreads) t[p.htab_words + i] = p.vtab[i];
    __syncthreads();
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s,
                             const Params& p, uint8_t* scratch, uint8_t* gtmp) {
    const int tid = threadIdx.x;
    const int c = s.c;
    const int32_t* ht = p.htab ? reinterpret_cast<const int32_t*>(scratch) : nullptr;
    const int32_t* vt = p.vtab ? reinterpret_cast<const int32_t*>(scratch) + p.htab_words : nullptr;
    uint8_t* tmp = gtmp ? gtmp : scratch + align_up((size_t)(p.htab_words + p.vtab_words) * 4, 16);
    int cx0, ncols, ry0, nrows;
    needed(s, p, cx0, ncols, ry0, nrows);
    const int hk = ht ? ht[1] : 0, vk = vt ? vt[1] : 0;
    for (int j = 0; j < g; ++j) {
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* om = dst + (size_t)j * s.out_bytes;
      int ox = p.crop_x, oy = p.crop_y;
      if (p.crop_xy) {
        // per-image origins live on the device and cannot be validated by the host without a sync: clamp the box into the
        // image (Pillow would refuse such a box; here it can at least never read outside the image)
        ox = min(max(p.crop_xy[(img0 + j) * 2], 0), s.w - p.crop_w);
        oy = min(max(p.crop_xy[(img0 + j) * 2 + 1], 0), s.h - p.crop_h);
      }
      // zero the canvas if the resized image does not cover it
      const bool covers = (p.place_x <= 0 && p.place_y <= 0 && p.place_x + p.rw >= s.ow && p.place_y + p.rh >= s.oh);
      if (!covers) {
        for (int i = tid; i < s.out_bytes; i += kThreads) om[i] = 0;
      }
      __syncthreads();
      // pass 1: horizontal, rows of the crop box, needed columns only -> tmp (crop_h x ncols)
      for (int i = tid; i < p.crop_h * ncols; i += kThreads) {
        const int y = i / ncols, xx = i - y * ncols;
        const uint8_t* row = im + ((size_t)(oy + y) * s.w + ox) * c;
        uint8_t* t = tmp + (size_t)i * c;
        if (ht) {
          const int32_t* rec = ht + 2 + (cx0 + xx) * (2 + hk);
          const int xmin = rec[0], n = rec[1];
          int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
          if (c == 3) {
            for (int k = 0; k < n; ++k) {
              const int kk = rec[2 + k];
              const uint8_t* sp = row + (xmin + k) * 3;
              a0 += sp[0] * kk; a1 += sp[1] * kk; a2 += sp[2] * kk;
            }
            t[0] = clip8(a0 >> 22); t[1] = clip8(a1 >> 22); t[2] = clip8(a2 >> 22);
          } else {
            for (int ch = 0; ch < c; ++ch) {
              int a = 1 << 21;
              for (int k = 0; k < n; ++k) a += row[(xmin + k) * c + ch] * rec[2 + k];
              t[ch] = clip8(a >> 22);
            }
          }
        } else {
          for (int ch = 0; ch < c; ++ch) t[ch] = row[(cx0 + xx) * c + ch];
        }
      }
      __syncthreads();
      // pass 2: vertical, needed rows -> canvas
      for (int i = tid; i < nrows * ncols; i += kThreads) {
        const int yy = i / ncols, xx = i - yy * ncols;
        uint8_t* o = om + ((size_t)(p.place_y + ry0 + yy) * s.ow + (p.place_x + cx0 + xx)) * c;
        if (vt) {
          const int32_t* rec = vt + 2 + (ry0 + yy) * (2 + vk);
          const int ymin = rec[0], n = rec[1];
          if (c == 3) {
            int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
            for (int k = 0; k < n; ++k) {
              const int kk = rec[2 + k];
              const uint8_t* sp = tmp + ((size_t)(ymin + k) * ncols + xx) * 3;
              a0 += sp[0] * kk; a1 += sp[1] * kk; a2 += sp[2] * kk;
            }
            o[0] = clip8(a0 >> 22); o[1] = clip8(a1 >> 22); o[2] = clip8(a2 >> 22);
          } else {
            for (int ch = 0; ch < c; ++ch) {
              int a = 1 << 21;
              for (int k = 0; k < n; ++k) a += tmp[((size_t)(ymin + k) * ncols + xx) * c + ch] * rec[2 + k];
              o[ch] = clip8(a >> 22);
            }
          }
        } else {
          const uint8_t* sp = tmp + ((size_t)(ry0 + yy) * ncols + xx) * c;
          for (int ch = 0; ch < c; ++ch) o[ch] = sp[ch];
        }
      }
      __syncthreads();
    }
  }
};

// ---- two chained resizes in one launch (CIFAR-10-C pixelate: BOX down to (sw, sh), BOX back up) -----------------------
// Four Q22 passes (h, v, h, v) with uint8 intermediates exactly like two Image.resize calls, but the intermediates never
// leave shared memory and every pass is flattened over all images of the group (4 barriers per group, not 6 per image).
struct ChainParams {
  const int32_t* tab[4];   // device tables: h1 (w -> sw), v1 (h -> sh), h2 (sw -> w), v2 (sh -> h)
  int words[4];
  int sw, sh;
};

struct OpResizeChain : NoSpill {
  using Params = ChainParams;
  static constexpr int kThreads = 256;
  static constexpr int kStages = 2, kOutBufs = 1;
  // scratch: tables | per image two ping-pong planes of max(h,sh) x max(w,sw) x 3 bytes
  __host__ __device__ static size_t plane_bytes(const Shape& s) { return align_up((size_t)s.h * s.w * 3 + 4, 16); }
  static size_t scratch_bytes(const Shape& s, int G) { return (size_t)s.aux + (size_t)G * 2 * plane_bytes(s) + 16; }  // aux = table bytes
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    int32_t* t = reinterpret_cast<int32_t*>(scratch);
    int off = 0;
    for (int q = 0; q < 4; ++q) {
      for (int i = threadIdx.x; i < p.words[q]; i += kThreads) t[off + i] = p.tab[q][i];
      off += p.words[q];
    }
    __syncthreads();
  }
  // one pass over all g images: out[j][y][x] = clip8((2^21 + sum_k in[j][..] * k) >> 22), along x (HORIZ) or y
  template <bool HORIZ>
  __device__ static __forceinline__ void pass(const uint8_t* in, size_t in_pitch_img, int in_w, uint8_t* out,
                                              size_t out_pitch_img, int out_w, int out_h, const int32_t* tab, int g) {
    // one warp per output row, lanes over x: no integer division anywhere (it costs ~35 instructions per pixel otherwise)
    const int ks = tab[1], lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int j = 0; j < g; ++j) {
      const uint8_t* ip = in + (size_t)j * in_pitch_img;
      uint8_t* op = out + (size_t)j * out_pitch_img;
      for (int y = warp; y < out_h; y += kThreads >> 5) {
        for (int x = lane; x < out_w; x += 32) {
          const int32_t* rec = tab + 2 + (HORIZ ? x : y) * (2 + ks);
          const int lo = rec[0], n = rec[1];
          int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
          for (int k = 0; k < n; ++k) {
            const int kk = rec[2 + k];
            const uint32_t px = load_rgb(ip, HORIZ ? (y * in_w + lo + k) * 3 : ((lo + k) * in_w + x) * 3);
            a0 += (int)(px & 255u) * kk; a1 += (int)((px >> 8) & 255u) * kk; a2 += (int)((px >> 16) & 255u) * kk;
          }
          uint8_t* o = op + (y * out_w + x) * 3;
          o[0] = clip8(a0 >> 22); o[1] = clip8(a1 >> 22); o[2] = clip8(a2 >> 22);
        }
      }
    }
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const int32_t* t0 = reinterpret_cast<const int32_t*>(scratch);
    const int32_t* t1 = t0 + p.words[0];
    const int32_t* t2 = t1 + p.words[1];
    const int32_t* t3 = t2 + p.words[2];
    const size_t pb = plane_bytes(s);
    uint8_t* A = scratch + s.aux;
    uint8_t* B = A + (size_t)g * pb;
    __syncthreads();
    pass<true>(src, s.in_bytes, s.w, A, pb, p.sw, s.h, t0, g);        // (h, w)  -> (h, sw)
    __syncthreads();
    pass<false>(A, pb, p.sw, B, pb, p.sw, p.sh, t1, g);               // (h, sw) -> (sh, sw)
    __syncthreads();
    pass<true>(B, pb, p.sw, A, pb, s.w, p.sh, t2, g);                 // (sh, sw) -> (sh, w)
    __syncthreads();
    pass<false>(A, pb, s.w, dst, s.out_bytes, s.w, s.h, t3, g);       // (sh, w) -> (h, w)
  }
};

// ---- BOX pixelate, throughput form ------------------------------------------------------------------------------------
// For the BOX filter the two up-sampling passes have exactly one tap of weight 2^22 per output (a pure gather), and the two
// down-sampling passes have at most two taps.  So only the small (sh x sw) image is ever computed — 32-bit pixels, two
// branch-free taps — and the output is one gather written as words.  The host verifies the table structure; anything else
// goes through OpResizeChain.  Same Q22 arithmetic and uint8 rounding between passes as Resample.c.
constexpr int kPixMax = 64;
struct BoxPixParams {
  int sw, sh;
  uint8_t hlo[kPixMax], hlo1[kPixMax], vlo[kPixMax], vlo1[kPixMax], xmap[kPixMax], ymap[kPixMax];
  int32_t hk0[kPixMax], hk1[kPixMax], vk0[kPixMax], vk1[kPixMax];
};

struct OpBoxPixelate : NoSpill {
  using Params = BoxPixParams;
  static constexpr int kThreads = 256;
  static constexpr bool kCifarCopy = true;
  static constexpr int kHeader = 6 * kPixMax * 4 + 4 * kPixMax * 4;   // index tables widened to int + weights
  // per image: A (h x sw) and B (sh x sw) as 32-bit pixels
  static size_t scratch_bytes(const Shape& s, int G) { return kHeader + (size_t)G * 2 * s.h * s.aux * 4 + 16; }   // aux = sw
  static size_t scratch_bytes_direct(const Shape& s) { return scratch_bytes(s, 1); }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* scratch, bool) {
    int* t = reinterpret_cast<int*>(scratch);
    for (int i = threadIdx.x; i < kPixMax; i += kThreads) {
      t[i] = p.hlo[i]; t[kPixMax + i] = p.hlo1[i]; t[2 * kPixMax + i] = p.vlo[i]; t[3 * kPixMax + i] = p.vlo1[i];
      t[4 * kPixMax + i] = p.xmap[i]; t[5 * kPixMax + i] = p.ymap[i];
      t[6 * kPixMax + i] = p.hk0[i]; t[7 * kPixMax + i] = p.hk1[i]; t[8 * kPixMax + i] = p.vk0[i]; t[9 * kPixMax + i] = p.vk1[i];
    }
    __syncthreads();
  }
  __device__ static __forceinline__ uint32_t blend2(uint32_t a, uint32_t b, int k0, int k1) {
    const int r0 = ((1 << 21) + (int)(a & 255u) * k0 + (int)(b & 255u) * k1) >> 22;
    const int r1 = ((1 << 21) + (int)((a >> 8) & 255u) * k0 + (int)((b >> 8) & 255u) * k1) >> 22;
    const int r2 = ((1 << 21) + (int)((a >> 16) & 255u) * k0 + (int)((b >> 16) & 255u) * k1) >> 22;
    return (uint32_t)min(r0, 255) | ((uint32_t)min(r1, 255) << 8) | ((uint32_t)min(r2, 255) << 16);
  }
  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t, const Shape& s, const Params& p,
                             uint8_t* scratch, uint8_t*) {
    const int* t = reinterpret_cast<const int*>(scratch);
    const int *hlo = t, *hlo1 = t + kPixMax, *vlo = t + 2 * kPixMax, *vlo1 = t + 3 * kPixMax, *xmap = t + 4 * kPixMax,
              *ymap = t + 5 * kPixMax, *hk0 = t + 6 * kPixMax, *hk1 = t + 7 * kPixMax, *vk0 = t + 8 * kPixMax, *vk1 = t + 9 * kPixMax;
    const int sw = p.sw, sh = p.sh, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = kThreads >> 5;
    uint32_t* A = reinterpret_cast<uint32_t*>(scratch + kHeader);   // [g][h][sw]
    uint32_t* B = A + (size_t)g * s.h * sw;                          // [g][sh][sw]
    __syncthreads();
    // 1. horizontal down-sampling: one warp per (image, row), lanes over the small width
    for (int r = warp; r < g * s.h; r += nw) {
      const int j = r / s.h, y = r - j * s.h;
      const uint8_t* row = src + (size_t)j * s.in_bytes + y * s.w * 3;
      for (int x = lane; x < sw; x += 32)
        A[(size_t)r * sw + x] = blend2(load_rgb(row, hlo[x] * 3), load_rgb(row, hlo1[x] * 3), hk0[x], hk1[x]);
    }
    __syncthreads();
    // 2. vertical down-sampling
    for (int r = warp; r < g * sh; r += nw) {
      const int j = r / sh, y = r - j * sh;
      const uint32_t* a0 = A + ((size_t)j * s.h + vlo[y]) * sw;
      const uint32_t* a1 = A + ((size_t)j * s.h + vlo1[y]) * sw;
      const int k0 = vk0[y], k1 = vk1[y];
      for (int x = lane; x < sw; x += 32) B[(size_t)r * sw + x] = blend2(a0[x], a1[x], k0, k1);
    }
    __syncthreads();
    // 3. both up-sampling passes: out[y][x] = B[ymap[y]][xmap[x]], four pixels (three words) per thread
    const int qpr = s.w >> 2;   // w % 4 == 0 (checked on the host)
    for (int i = threadIdx.x; i < g * s.h * qpr; i += kThreads) {
      const int r = i / qpr, q = i - r * qpr;
      const int j = r / s.h, y = r - j * s.h;
      const uint32_t* b = B + ((size_t)j * sh + ymap[y]) * sw;
      uint32_t* o = reinterpret_cast<uint32_t*>(dst + (size_t)j * s.out_bytes + y * s.w * 3);
      uint32_t w0, w1, w2;
      pack_rgb4(b[xmap[4 * q]], b[xmap[4 * q + 1]], b[xmap[4 * q + 2]], b[xmap[4 * q + 3]], w0, w1, w2);
      o[3 * q] = w0; o[3 * q + 1] = w1; o[3 * q + 2] = w2;
    }
  }
};

// ---- 32x32x3 images: the same passes as dense 32x32 operators on the tensor cores (sep_mma.cuh) ----------------------------
// Operator of one Pillow pass along one axis in canvas coordinates: row o = output index on the 32-pixel canvas, column =
// input index; rows outside the resized extent stay zero ((2^21 + 0) >> 22 = 0 is the zero canvas / the unused rows).
__device__ inline void sep_fill_axis(int32_t* op, const int32_t* tab, int crop, int place, int rsize) {
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) op[i] = 0;
  __syncthreads();
  for (int o = threadIdx.x; o < 32; o += blockDim.x) {
    const int r = o - place;
    if (r < 0 || r >= rsize) continue;
    if (tab) {
      const int32_t* rec = tab + 2 + r * (2 + tab[1]);
      for (int j = 0; j < rec[1]; ++j) op[o * 32 + crop + rec[0] + j] = rec[2 + j];
    } else {
      op[o * 32 + crop + r] = 1 << 22;
    }
  }
  __syncthreads();
}

struct OpResampleMma {
  using Params = ResampleParams;
  static constexpr int kTeamWarps = 1, kInBufs = 3, kOutBufs = 0, kMaxTeams = 12, kCtasPerSm = 1;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  static size_t cta_scratch(const Shape&, const Params&) { return kSepOpBytes + 6 * kSepFragBytes; }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    int32_t* op = reinterpret_cast<int32_t*>(cta);
    uint4* frag = reinterpret_cast<uint4*>(cta + kSepOpBytes);
    sep_fill_axis(op, p.htab, p.crop_x, p.place_x, p.rw);
    sep_build_frags(op, frag, 0, 3, false);
    __syncthreads();
    sep_fill_axis(op, p.vtab, p.crop_y, p.place_y, p.rh);
    sep_build_frags(op, frag + 3 * 64, 0, 3, true);
  }
  __device__ static void run(const uint8_t* in, uint8_t* out, int64_t, const Shape&, const Params&, uint8_t*,
                             const uint8_t* cta, int tt, int) {
    sep_q22_chain<3, 3, 0, 0>(in, out, reinterpret_cast<const uint4*>(cta + kSepOpBytes), tt);
  }
};

// pixelate: BOX down (two taps at most) in both axes, then the two unit-tap BOX up passes
struct OpBoxPixelateMma {
  using Params = BoxPixParams;
  static constexpr int kTeamWarps = 1, kInBufs = 3, kOutBufs = 0, kMaxTeams = 12, kCtasPerSm = 1;
  static size_t team_scratch(const Shape&, const Params&) { return 0; }
  static size_t cta_scratch(const Shape&, const Params&) { return kSepOpBytes + 6 * kSepFragBytes; }
  // The up-sampling passes are selections (unit taps), and a selection commutes with the rounding of the pass before it:
  // out[yo][xo] = B[ymap[yo]][xmap[xo]] with B = round(Dv round(X Dh^T)).  So row o of the composed operator is row map[o]
  // of the down-sampling operator, and the four Pillow passes are TWO 32x32 stages.
  __device__ static void fill(int32_t* op, const uint8_t* map, const uint8_t* lo, const uint8_t* lo1, const int32_t* k0,
                              const int32_t* k1) {
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) op[i] = 0;
    __syncthreads();
    for (int o = threadIdx.x; o < 32; o += blockDim.x) {
      const int m = map[o];
      op[o * 32 + lo[m]] = k0[m];
      if (lo1[m] != lo[m]) op[o * 32 + lo1[m]] = k1[m];
    }
    __syncthreads();
  }
  __device__ static void prepare(const Shape&, const Params& p, uint8_t* cta) {
    int32_t* op = reinterpret_cast<int32_t*>(cta);
    uint4* frag = reinterpret_cast<uint4*>(cta + kSepOpBytes);
    fill(op, p.xmap, p.hlo, p.hlo1, p.hk0, p.hk1);
    sep_build_frags(op, frag, 0, 3, false);
    __syncthreads();
    fill(op, p.ymap, p.vlo, p.vlo1, p.vk0, p.vk1);
    sep_build_frags(op, frag + 3 * 64, 0, 3, true);
  }
  __device__ static void run(const uint8_t* in, uint8_t* out, int64_t, const Shape&, const Params&, uint8_t*,
                             const uint8_t* cta, int tt, int) {
    sep_q22_chain<3, 3, 0, 0>(in, out, reinterpret_cast<const uint4*>(cta + kSepOpBytes), tt);
  }
};

// pixelate on images larger than the table block of BoxPixParams (224 x 224): one thread per output pixel evaluates its two
// BOX passes from the four source pixels they touch (the up-sampling passes are selections); tables in device memory:
// xt[x] = {3 * c0, 3 * c1, kh0, kh1} and yt[y] = {r0, r1, kv0, kv1} of the output pixel's column / row.  Integer arithmetic:
// the same bytes as the chained passes.
__global__ void __launch_bounds__(256)
box_pixelate_large_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int h, int w,
                          const int4* __restrict__ xt, const int4* __restrict__ yt) {
  const int hw = h * w;
  for (int64_t img = blockIdx.y; img < n; img += gridDim.y) {
    const uint8_t* im = in + img * (int64_t)hw * 3;
    uint8_t* om = out + img * (int64_t)hw * 3;
    for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < hw; px += gridDim.x * blockDim.x) {
      const int y = px / w, x = px - y * w;
      const int4 cx = __ldg(xt + x), ry = __ldg(yt + y);
      const uint8_t* r0 = im + (int64_t)ry.x * w * 3;
      const uint8_t* r1 = im + (int64_t)ry.y * w * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int a0 = min(((1 << 21) + r0[cx.x + c] * cx.z + r0[cx.y + c] * cx.w) >> 22, 255);
        const int a1 = min(((1 << 21) + r1[cx.x + c] * cx.z + r1[cx.y + c] * cx.w) >> 22, 255);
        om[(int64_t)px * 3 + c] = (uint8_t)min(((1 << 21) + a0 * ry.z + a1 * ry.w) >> 22, 255);
      }
    }
  }
}

// Specialised launcher: scratch depends on the parameters (table sizes, needed window).
static int launch_resample(const uint8_t* in, uint8_t* out, int64_t n, const Shape& s, const ResampleParams& p,
                           cudaStream_t stream) {
  if (n <= 0) return IMGTX_OK;
  const StagedTuning tn = staged_tuning();
  const int max_smem = max_smem_optin();
  const size_t op_scratch = OpResample::table_bytes(p) + OpResample::tmp_bytes(s, p);
  bool staged = (s.in_bytes % 16 == 0) && (s.out_bytes % 16 == 0) && aligned16(in) && aligned16(out);
  int G = 1, stages = tn.stages;
  size_t smem_bytes = 0;
  auto need = [&](int g, int st) {
    return (size_t)kBarBytes + align_up((size_t)g * s.in_bytes, 128) * st + align_up((size_t)g * s.out_bytes, 128) * 2 +
           op_scratch + 128;
  };
  if (staged) {
    while (stages > 2 && need(1, stages) > (size_t)tn.target_smem) --stages;
    if (need(1, stages) > (size_t)max_smem - 1024) staged = false;
    else {
      while (G < 64 && need(G + 1, stages) <= (size_t)tn.target_smem) ++G;
      smem_bytes = need(G, stages);
    }
  }
  if (staged) {
    auto kern = staged_kernel<OpResample>;
    IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int64_t n_groups = (n + G - 1) / G;
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(tn.ctas_per_sm, (size_t)max_smem / smem_bytes));
    int64_t grid = std::min<int64_t>(n_groups, (int64_t)sm_count() * per_sm);
    kern<<<(unsigned)grid, OpResample::kThreads, smem_bytes, stream>>>(in, out, n, s, p, G, stages);
    IMGTX_LAUNCH_OK();
    return IMGTX_OK;
  }
  // direct mode; intermediate plane in shared memory if it fits, else in a stream-ordered workspace
  size_t scratch = op_scratch;
  int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * 4);
  uint8_t* gtmp = nullptr;
  size_t per_cta = 0;
  if (scratch > (size_t)max_smem - 1024) {
    per_cta = align_up(OpResample::tmp_bytes(s, p), 256);
    scratch = OpResample::table_bytes(p) + 16;
    IMGTX_CHECK_ARG(scratch <= (size_t)max_smem - 1024, "resample tables do not fit in shared memory");
    IMGTX_CUDA_OK(workspace_alloc((void**)&gtmp, per_cta * grid, stream));
  }
  auto kern = direct_kernel<OpResample>;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scratch));
  kern<<<(unsigned)grid, OpResample::kThreads, scratch, stream>>>(in, out, n, s, p, gtmp, per_cta);
  IMGTX_LAUNCH_OK();
  if (gtmp) IMGTX_CUDA_OK(cudaFreeAsync(gtmp, stream));
  return IMGTX_OK;
}

// ---- host: Pillow coefficient tables ----------------------------------------------------------
static double filt_box(double x) { return (x > -0.5 && x <= 0.5) ? 1.0 : 0.0; }
static double filt_bilinear(double x) {
  if (x < 0.0) x = -x;
  return x < 1.0 ? 1.0 - x : 0.0;
}
static double filt_bicubic(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}
static double sinc(double x) {
  if (x == 0.0) return 1.0;
  x = x * M_PI;
  return sin(x) / x;
}
static double filt_lanczos(double x) { return (-3.0 <= x && x < 3.0) ? sinc(x) * sinc(x / 3) : 0.0; }

static bool filter_info(int filter, double (**fn)(double), double* support) {
  switch (filter) {
    case IMGTX_FILTER_BOX: *fn = filt_box; *support = 0.5; return true;
    case IMGTX_FILTER_BILINEAR: *fn = filt_bilinear; *support = 1.0; return true;
    case IMGTX_FILTER_BICUBIC: *fn = filt_bicubic; *support = 2.0; return true;
    case IMGTX_FILTER_LANCZOS: *fn = filt_lanczos; *support = 3.0; return true;
  }
  return false;
}

static int table_ksize(int in_size, int out_size, double support0) {
  double filterscale = (double)in_size / out_size;
  if (filterscale < 1.0) filterscale = 1.0;
  return (int)ceil(support0 * filterscale) * 2 + 1;
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int64_t imgtx_resample_table_words(int in_size, int out_size, int filter) {
  double (*fn)(double);
  double support0;
  if (in_size <= 0 || out_size <= 0 || !filter_info(filter, &fn, &support0)) return -1;
  return 2 + (int64_t)out_size * (2 + table_ksize(in_size, out_size, support0));
}

extern "C" int imgtx_resample_table_host(int in_size, int out_size, int filter, int32_t* table_host) {
  double (*fn)(double);
  double support0;
  IMGTX_CHECK_ARG(in_size > 0 && out_size > 0 && table_host, "bad table arguments");
  IMGTX_CHECK_ARG(filter_info(filter, &fn, &support0), "unknown filter %d", filter);
  // libImaging/Resample.c precompute_coeffs + normalize_coeffs_8bpc
  double scale, filterscale;
  scale = filterscale = (double)in_size / out_size;
  if (filterscale < 1.0) filterscale = 1.0;
  const double support = support0 * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  table_host[0] = out_size;
  table_host[1] = ksize;
  std::vector<double> k(ksize);
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = 0.0 + (xx + 0.5) * scale;
    double ww = 0.0;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    for (int x = 0; x < xmax; ++x) {
      double w = fn((x + xmin - center + 0.5) * ss);
      k[x] = w;
      ww += w;
    }
    for (int x = 0; x < xmax; ++x)
      if (ww != 0.0) k[x] /= ww;
    int32_t* rec = table_host + 2 + (size_t)xx * (2 + ksize);
    rec[0] = xmin;
    rec[1] = xmax;
    for (int x = 0; x < ksize; ++x) {
      double v = x < xmax ? k[x] : 0.0;
      rec[2 + x] = v < 0 ? (int)(-0.5 + v * (1 << 22)) : (int)(0.5 + v * (1 << 22));
    }
  }
  return IMGTX_OK;
}

extern "C" int imgtx_resample_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int crop_x,
                                 int crop_y, int crop_w, int crop_h, const int32_t* crop_xy_per_image_dev, int rw,
                                 int rh, const int32_t* htable_dev, int64_t htable_words, const int32_t* vtable_dev,
                                 int64_t vtable_words, int oh, int ow, int place_x, int place_y,
                                 imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4 && oh > 0 && ow > 0, "bad image shape");
  IMGTX_CHECK_ARG(crop_w > 0 && crop_h > 0 && crop_w <= w && crop_h <= h, "bad crop box");
  if (!crop_xy_per_image_dev)
    IMGTX_CHECK_ARG(crop_x >= 0 && crop_y >= 0 && crop_x + crop_w <= w && crop_y + crop_h <= h, "crop box outside image");
  IMGTX_CHECK_ARG(rw > 0 && rh > 0, "bad resized size");
  IMGTX_CHECK_ARG((htable_dev != nullptr) == (rw != crop_w), "horizontal table must be given iff the width changes");
  IMGTX_CHECK_ARG((vtable_dev != nullptr) == (rh != crop_h), "vertical table must be given iff the height changes");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null image pointer");
  IMGTX_CHECK_ARG(!htable_dev || htable_words >= 2, "horizontal table too short");
  IMGTX_CHECK_ARG(!vtable_dev || vtable_words >= 2, "vertical table too short");
  ResampleParams p;
  memset(&p, 0, sizeof(p));
  p.htab = htable_dev; p.vtab = vtable_dev;
  p.htab_words = htable_dev ? (int)htable_words : 0;
  p.vtab_words = vtable_dev ? (int)vtable_words : 0;
  p.crop_x = crop_x; p.crop_y = crop_y; p.crop_w = crop_w; p.crop_h = crop_h;
  p.crop_xy = crop_xy_per_image_dev;
  p.rw = rw; p.rh = rh; p.place_x = place_x; p.place_y = place_y;
  if (h == 32 && w == 32 && c == 3 && oh == 32 && ow == 32 && !crop_xy_per_image_dev && (htable_dev || vtable_dev)) {
    const int rc = launch_team_op<OpResampleMma>(in, out, n, make_shape(h, w, c, oh, ow), p, (cudaStream_t)stream);
    if (rc != IMGTX_ERR_UNSUPPORTED) return rc;
  }
  return launch_resample(in, out, n, make_shape(h, w, c, oh, ow), p, (cudaStream_t)stream);
}

/* Image.resize((sw, sh), filter) followed by Image.resize((w, h), filter) in one launch; tables from
   imgtx_resample_table_host for (w -> sw), (h -> sh), (sw -> w), (sh -> h).  CIFAR-10-C pixelate (BOX both ways). */
extern "C" int imgtx_resize_chain_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int sw, int sh,
                                     const int32_t* h1_dev, int64_t h1_words, const int32_t* v1_dev, int64_t v1_words,
                                     const int32_t* h2_dev, int64_t h2_words, const int32_t* v2_dev, int64_t v2_words,
                                     imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "resize chain takes (n,h,w,3) images");
  IMGTX_CHECK_ARG(sw > 0 && sh > 0 && sw <= w && sh <= h, "the intermediate size must not exceed the image");
  IMGTX_CHECK_ARG(h1_dev && v1_dev && h2_dev && v2_dev && h1_words > 2 && v1_words > 2 && h2_words > 2 && v2_words > 2,
                  "four coefficient tables are required");
  if (n == 0) return IMGTX_OK;
  ChainParams p;
  p.tab[0] = h1_dev; p.tab[1] = v1_dev; p.tab[2] = h2_dev; p.tab[3] = v2_dev;
  p.words[0] = (int)h1_words; p.words[1] = (int)v1_words; p.words[2] = (int)h2_words; p.words[3] = (int)v2_words;
  p.sw = sw; p.sh = sh;
  Shape shp = make_shape(h, w, c, h, w);
  shp.aux = (int)align_up((size_t)(h1_words + v1_words + h2_words + v2_words) * 4, 16);
  const int target = 128 + 3 * 4 * shp.in_bytes + (int)OpResizeChain::scratch_bytes(shp, 4) + 512;   // 4 images per group
  return launch_image_op<OpResizeChain>(in, out, n, shp, p, (cudaStream_t)stream, 0, std::min(target, max_smem_optin() - 2048));
}

/* CIFAR-10-C pixelate: Image.resize((sw, sh), BOX) then Image.resize((w, h), BOX); gather forms (shared-memory tables up to
   64x64, device tables above).
   Returns IMGTX_OK when launched, 1 when the BOX tables do not have the (<= 2 taps down, single unit tap up) structure or
   the shape is outside the fast path — the caller then uses imgtx_resize_chain_u8 (same results). */
static int impl_box_pixelate_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int sw, int sh,
                                     int flags, uint32_t* metrics_dev, bool* metrics_done, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c == 3 && n >= 0, "pixelate takes (n,h,w,3) images");
  IMGTX_CHECK_ARG(sw > 0 && sh > 0 && sw <= w && sh <= h, "the intermediate size must not exceed the image");
  if (sw == w || sh == h) return 1;
  if (h > kPixMax || w > kPixMax || (w & 3)) {
    // large images: per-pixel form with the tables in device memory (same structure test as below)
    const int kBoxL = IMGTX_FILTER_BOX;
    std::vector<int32_t> th((size_t)imgtx_resample_table_words(w, sw, kBoxL)), tv((size_t)imgtx_resample_table_words(h, sh, kBoxL)),
        uh((size_t)imgtx_resample_table_words(sw, w, kBoxL)), uv((size_t)imgtx_resample_table_words(sh, h, kBoxL));
    if (imgtx_resample_table_host(w, sw, kBoxL, th.data()) || imgtx_resample_table_host(h, sh, kBoxL, tv.data()) ||
        imgtx_resample_table_host(sw, w, kBoxL, uh.data()) || imgtx_resample_table_host(sh, h, kBoxL, uv.data()))
      return 1;
    std::vector<int32_t> host((size_t)(w + h) * 4);
    auto fill = [&](const std::vector<int32_t>& up, const std::vector<int32_t>& down, int out_size, int mul, int32_t* dst) -> bool {
      const int ku = up[1], kd = down[1];
      for (int i = 0; i < out_size; ++i) {
        const int32_t* ru = up.data() + 2 + (size_t)i * (2 + ku);
        if (ru[1] != 1 || ru[2] != (1 << 22)) return false;          // up-sampling: one unit tap
        const int32_t* rd = down.data() + 2 + (size_t)ru[0] * (2 + kd);
        if (rd[1] < 1 || rd[1] > 2 || rd[2] < 0 || (rd[1] == 2 && rd[3] < 0)) return false;   // down-sampling: <= 2 taps
        dst[4 * i] = rd[0] * mul;
        dst[4 * i + 1] = (rd[1] == 2 ? rd[0] + 1 : rd[0]) * mul;
        dst[4 * i + 2] = rd[2];
        dst[4 * i + 3] = rd[1] == 2 ? rd[3] : 0;
      }
      return true;
    };
    if (!fill(uh, th, w, 3, host.data()) || !fill(uv, tv, h, 1, host.data() + (size_t)w * 4)) return 1;
    if (n == 0) return IMGTX_OK;
    IMGTX_CHECK_ARG(in && out, "null image pointer");
    cudaStream_t st = (cudaStream_t)stream;
    int32_t* tab = nullptr;
    IMGTX_CUDA_OK(workspace_alloc((void**)&tab, host.size() * 4, st));
    // pageable source: the call returns once the buffer has been copied to the driver's staging memory
    IMGTX_CUDA_OK(cudaMemcpyAsync(tab, host.data(), host.size() * 4, cudaMemcpyHostToDevice, st));
    const dim3 grid((unsigned)std::min(64, (h * w + 255) / 256), (unsigned)std::min<int64_t>(n, 4096));
    box_pixelate_large_kernel<<<grid, 256, 0, st>>>(in, out, n, h, w, reinterpret_cast<const int4*>(tab),
                                                    reinterpret_cast<const int4*>(tab) + w);
    IMGTX_LAUNCH_OK();
    IMGTX_CUDA_OK(cudaFreeAsync(tab, st));
    return IMGTX_OK;
  }
  if (n == 0) return IMGTX_OK;
  BoxPixParams p;
  memset(&p, 0, sizeof(p));
  p.sw = sw; p.sh = sh;
  const int kBox = IMGTX_FILTER_BOX;
  auto down = [&](int in_size, int out_size, uint8_t* lo, uint8_t* lo1, int32_t* k0, int32_t* k1) -> bool {
    std::vector<int32_t> t((size_t)imgtx_resample_table_words(in_size, out_size, kBox));
    if (imgtx_resample_table_host(in_size, out_size, kBox, t.data()) != IMGTX_OK) return false;
    const int ks = t[1];
    for (int i = 0; i < out_size; ++i) {
      const int32_t* rec = t.data() + 2 + (size_t)i * (2 + ks);
      if (rec[1] < 1 || rec[1] > 2) return false;
      lo[i] = (uint8_t)rec[0];
      lo1[i] = (uint8_t)(rec[1] == 2 ? rec[0] + 1 : rec[0]);
      k0[i] = rec[2];
      k1[i] = rec[1] == 2 ? rec[3] : 0;
      if (k0[i] < 0 || k1[i] < 0) return false;
    }
    return true;
  };
  auto up = [&](int in_size, int out_size, uint8_t* map) -> bool {
    std::vector<int32_t> t((size_t)imgtx_resample_table_words(in_size, out_size, kBox));
    if (imgtx_resample_table_host(in_size, out_size, kBox, t.data()) != IMGTX_OK) return false;
    const int ks = t[1];
    for (int i = 0; i < out_size; ++i) {
      const int32_t* rec = t.data() + 2 + (size_t)i * (2 + ks);
      if (rec[1] != 1 || rec[2] != (1 << 22)) return false;
      map[i] = (uint8_t)rec[0];
    }
    return true;
  };
  if (!down(w, sw, p.hlo, p.hlo1, p.hk0, p.hk1) || !down(h, sh, p.vlo, p.vlo1, p.vk0, p.vk1) || !up(sw, w, p.xmap) ||
      !up(sh, h, p.ymap))
    return 1;
  Shape shp = make_shape(h, w, c, h, w);
  shp.aux = sw;
  // integer arithmetic: every form gives the same bytes.  32x32 images go through the tensor-core operator chain (sep_mma.cuh);
  // IMGTX_CC_REPLAY_ALL keeps the scalar form (a scalar warp- or team-per-image port measured slower: 360-545 us vs 295-348 us).
  if (h == 32 && w == 32 && !(flags & IMGTX_CC_REPLAY_ALL)) {
    const int rc = launch_team_op<OpBoxPixelateMma>(in, out, n, shp, p, (cudaStream_t)stream, metrics_dev);
    if (rc != IMGTX_ERR_UNSUPPORTED) {
      *metrics_done = metrics_dev != nullptr;
      return rc;
    }
  }
  return launch_image_op<OpBoxPixelate>(in, out, n, shp, p, (cudaStream_t)stream);
}

extern "C" int imgtx_box_pixelate_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int sw, int sh,
                                     int flags, uint32_t* metrics_dev, imgtx_stream_t stream) {
  bool metrics_done = false;
  int rc = impl_box_pixelate_u8(in, out, n, h, w, c, sw, sh, flags, metrics_dev, &metrics_done, stream);
  // operators without a metric epilogue: one extra read of the output (imgtx_image_checksums_u8)
  if (rc == IMGTX_OK && metrics_dev && !metrics_done) rc = imgtx::image_checksums(out, n, (int64_t)h * w * c, metrics_dev, (cudaStream_t)stream);
  return rc;
}
