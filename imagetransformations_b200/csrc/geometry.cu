// Integer geometry on uint8 HWC images: every op is a pure index gather (no interpolation).
//   IMGTX_GEOM_AFFINE        Pillow Image.transform(AFFINE, NEAREST) / Image.rotate: 16.16 fixed-point
//                            gather, or the double-accumulating scale/translate path when m1 == m3 == 0
//                            (libImaging/Geometry.c affine_fixed / ImagingScaleAffine).
//                            reference: transformation.py:198-201, fall_2025/AugMix.py:30-35,
//                            fall_2025/Initial_Experiments.py:88-92
//   IMGTX_GEOM_SHIFT         apply_translation (transformation.py:284-307)
//   IMGTX_GEOM_FLIP_LR       Image.transpose(FLIP_LEFT_RIGHT) / RandomHorizontalFlip(p=1)
//                            (fall_2025/transformations_code:39-41, TTA_transforms.py:75)
//   IMGTX_GEOM_CROP_REFLECT  T.RandomCrop(size, padding, 'reflect') (TTA_transforms.py:80)
//   IMGTX_GEOM_ROT90/180/270 the exact-transpose special cases of Image.rotate
// Roofline: HBM, in_bytes read + out_bytes written per image; the gather itself runs out of shared
// memory (staged mode) so HBM only sees sequential bulk copies.
#include "pipeline.cuh"
#include <string.h>

namespace imgtx {

struct GeomParams {
  imgtx_geom_t uni;                // used when per_image == nullptr
  const imgtx_geom_t* per_image;   // device, n entries, or nullptr
  uint8_t fill[4];
};

struct OpGeometry : NoSpill {
  using Params = GeomParams;
  static constexpr int kThreads = 256;
  // per image of the group: x table (ow ints) + y table (oh ints)
  static size_t scratch_bytes(const Shape& s, int G) { return (size_t)G * (s.ow + s.oh) * sizeof(int); }

  __device__ static bool is_fixed_affine(const imgtx_geom_t& q) {
    return q.kind == IMGTX_GEOM_AFFINE && !(q.m[1] == 0.0 && q.m[3] == 0.0);
  }

  // Separable index tables for everything except the general fixed-point affine and the 90/270 turns.
  __device__ static void build_tables(const imgtx_geom_t& q, const Shape& s, int* xt, int* yt, int t,
                                      int nt) {
    switch (q.kind) {
      case IMGTX_GEOM_AFFINE: {
        // ImagingScaleAffine: xo = m2 + m0*0.5, then xo += m0 per column, in double.
        if (t == 0) {
          double xo = q.m[2] + q.m[0] * 0.5;
          for (int x = 0; x < s.ow; ++x) {
            int xin = (xo < 0.0) ? -1 : (int)xo;
            xt[x] = (xin >= 0 && xin < s.w) ? xin : -1;
            xo += q.m[0];
          }
        } else if (t == 32 % nt) {
          double yo = q.m[5] + q.m[4] * 0.5;
          for (int y = 0; y < s.oh; ++y) {
            int yin = (yo < 0.0) ? -1 : (int)yo;
            yt[y] = (yin >= 0 && yin < s.h) ? yin : -1;
            yo += q.m[4];
          }
        }
        break;
      }
      case IMGTX_GEOM_SHIFT:
        for (int x = t; x < s.ow; x += nt) { int xi = x - q.dx; xt[x] = (xi >= 0 && xi < s.w) ? xi : -1; }
        for (int y = t; y < s.oh; y += nt) { int yi = y - q.dy; yt[y] = (yi >= 0 && yi < s.h) ? yi : -1; }
        break;
      case IMGTX_GEOM_FLIP_LR:
        for (int x = t; x < s.ow; x += nt) xt[x] = s.w - 1 - x;
        for (int y = t; y < s.oh; y += nt) yt[y] = y;
        break;
      case IMGTX_GEOM_ROT180:
        for (int x = t; x < s.ow; x += nt) xt[x] = s.w - 1 - x;
        for (int y = t; y < s.oh; y += nt) yt[y] = s.h - 1 - y;
        break;
      case IMGTX_GEOM_CROP_REFLECT:
        for (int x = t; x < s.ow; x += nt) xt[x] = reflect101(x + q.dx - q.pad, s.w);
        for (int y = t; y < s.oh; y += nt) yt[y] = reflect101(y + q.dy - q.pad, s.h);
        break;
      default:  // identity
        for (int x = t; x < s.ow; x += nt) xt[x] = x < s.w ? x : -1;
        for (int y = t; y < s.oh; y += nt) yt[y] = y < s.h ? y : -1;
        break;
    }
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s,
                             const Params& p, uint8_t* scratch, uint8_t*) {
    int* tabs = reinterpret_cast<int*>(scratch);
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5, n_warps = kThreads >> 5;
    const bool uniform = (p.per_image == nullptr);
    const int tab_stride = s.ow + s.oh;
    // 1) index tables (one set per image, or one shared set when parameters are uniform)
    const int n_sets = uniform ? 1 : g;
    for (int j = 0; j < n_sets; ++j) {
      const imgtx_geom_t& q = uniform ? p.uni : p.per_image[img0 + j];
      if (!is_fixed_affine(q) && q.kind != IMGTX_GEOM_ROT90 && q.kind != IMGTX_GEOM_ROT270)
        build_tables(q, s, tabs + j * tab_stride, tabs + j * tab_stride + s.ow, tid, kThreads);
    }
    __syncthreads();
    // 2) gather, one warp per output row
    const int c = s.c;
    for (int row = warp; row < g * s.oh; row += n_warps) {
      const int j = row / s.oh;
      const int y = row - j * s.oh;
      const imgtx_geom_t& q = uniform ? p.uni : p.per_image[img0 + j];
      const uint8_t* im = src + (size_t)j * s.in_bytes;
      uint8_t* orow = dst + (size_t)j * s.out_bytes + (size_t)y * s.ow * c;
      if (is_fixed_affine(q)) {
        // FIX(v) = floor(v*65536 + 0.5) in double, exactly as libImaging
        const int a0 = (int)floor(q.m[0] * 65536.0 + 0.5), a1 = (int)floor(q.m[1] * 65536.0 + 0.5);
        const int a3 = (int)floor(q.m[3] * 65536.0 + 0.5), a4 = (int)floor(q.m[4] * 65536.0 + 0.5);
        const int a2 = (int)floor((q.m[2] + q.m[0] * 0.5 + q.m[1] * 0.5) * 65536.0 + 0.5);
        const int a5 = (int)floor((q.m[5] + q.m[3] * 0.5 + q.m[4] * 0.5) * 65536.0 + 0.5);
        const int bx = a2 + y * a1, by = a5 + y * a4;
        for (int x = lane; x < s.ow; x += 32) {
          const int xin = (bx + x * a0) >> 16, yin = (by + x * a3) >> 16;
          const bool inside = xin >= 0 && xin < s.w && yin >= 0 && yin < s.h;
          const uint8_t* sp = im + ((size_t)yin * s.w + xin) * c;
          for (int ch = 0; ch < c; ++ch) orow[x * c + ch] = inside ? sp[ch] : p.fill[ch];
        }
      } else if (q.kind == IMGTX_GEOM_ROT90 || q.kind == IMGTX_GEOM_ROT270) {
        // Image.rotate(90) == transpose(ROTATE_90): out[y][x] = in[x][w-1-y]; 270: out[y][x] = in[h-1-x][y]
        for (int x = lane; x < s.ow; x += 32) {
          int yin, xin;
          if (q.kind == IMGTX_GEOM_ROT90) { yin = x; xin = s.w - 1 - y; }
          else { yin = s.h - 1 - x; xin = y; }
          const uint8_t* sp = im + ((size_t)yin * s.w + xin) * c;
          for (int ch = 0; ch < c; ++ch) orow[x * c + ch] = sp[ch];
        }
      } else {
        const int* xt = tabs + (uniform ? 0 : j) * tab_stride;
        const int yin = xt[s.ow + y];
        for (int x = lane; x < s.ow; x += 32) {
          const int xin = xt[x];
          const bool inside = (xin >= 0) && (yin >= 0);
          const uint8_t* sp = im + ((size_t)(inside ? yin : 0) * s.w + (inside ? xin : 0)) * c;
          for (int ch = 0; ch < c; ++ch) orow[x * c + ch] = inside ? sp[ch] : p.fill[ch];
        }
      }
    }
  }
};

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_geometry_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int oh,
                                 int ow, const imgtx_geom_t* uniform_host,
                                 const imgtx_geom_t* per_image_dev, const uint8_t* fill_host,
                                 imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && oh > 0 && ow > 0 && c >= 1 && c <= 4, "bad image shape");
  IMGTX_CHECK_ARG(uniform_host != nullptr || per_image_dev != nullptr, "no geometry parameters");
  GeomParams p;
  memset(&p, 0, sizeof(p));
  if (uniform_host) p.uni = *uniform_host;
  p.per_image = per_image_dev;
  for (int i = 0; i < 4; ++i) p.fill[i] = fill_host ? fill_host[i < c ? i : 0] : 0;
  if (uniform_host && !per_image_dev) {
    const int k = uniform_host->kind;
    IMGTX_CHECK_ARG(k >= IMGTX_GEOM_IDENTITY && k <= IMGTX_GEOM_ROT270, "unknown geometry kind %d", k);
    if (k == IMGTX_GEOM_ROT90 || k == IMGTX_GEOM_ROT270)
      IMGTX_CHECK_ARG(oh == w && ow == h, "90/270 degree turns need transposed output size");
  }
  return launch_image_op<OpGeometry>(in, out, n, make_shape(h, w, c, oh, ow), p, (cudaStream_t)stream);
}
