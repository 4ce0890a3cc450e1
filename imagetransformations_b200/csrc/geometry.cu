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

// A geometry record "compiled" to integers: src = ((a2 + y*a1 + x*a0) >> 16, (a5 + y*a4 + x*a3) >> 16),
// optionally reflected (REFLECT_101) instead of bounds-checked, or looked up in separable tables.
struct CompiledGeom {
  int a0, a1, a2, a3, a4, a5;
  int flags;  // 1: reflect, 2: separable x/y tables (scale-only affine)
};
constexpr int kGeomReflect = 1, kGeomTables = 2;

__device__ __forceinline__ int fix16(double v) { return (int)floor(v * 65536.0 + 0.5); }

__device__ inline CompiledGeom compile_geom(const imgtx_geom_t& q, const Shape& s) {
  CompiledGeom c;
  c.a0 = 65536; c.a1 = 0; c.a2 = 0; c.a3 = 0; c.a4 = 65536; c.a5 = 0; c.flags = 0;
  switch (q.kind) {
    case IMGTX_GEOM_AFFINE:
      if (q.m[1] == 0.0 && q.m[3] == 0.0) {
        if (q.m[0] == 1.0 && q.m[4] == 1.0) {
          // pure translation through ImagingScaleAffine: xin = COORD(m2 + 0.5 + x) == x + floor(m2 + 0.5)
          // wherever the result is inside the image
          c.a2 = (int)floor(q.m[2] + 0.5) * 65536;
          c.a5 = (int)floor(q.m[5] + 0.5) * 65536;
        } else {
          c.flags = kGeomTables;
        }
      } else {
        // libImaging affine_fixed: FIX(v) = floor(v*65536 + 0.5)
        c.a0 = fix16(q.m[0]); c.a1 = fix16(q.m[1]); c.a3 = fix16(q.m[3]); c.a4 = fix16(q.m[4]);
        c.a2 = fix16(q.m[2] + q.m[0] * 0.5 + q.m[1] * 0.5);
        c.a5 = fix16(q.m[5] + q.m[3] * 0.5 + q.m[4] * 0.5);
      }
      break;
    case IMGTX_GEOM_SHIFT:
      c.a2 = -q.dx * 65536; c.a5 = -q.dy * 65536;
      break;
    case IMGTX_GEOM_FLIP_LR:
      c.a0 = -65536; c.a2 = (s.w - 1) * 65536;
      break;
    case IMGTX_GEOM_ROT180:
      c.a0 = -65536; c.a2 = (s.w - 1) * 65536; c.a4 = -65536; c.a5 = (s.h - 1) * 65536;
      break;
    case IMGTX_GEOM_ROT90:  // out[y][x] = in[x][w-1-y]
      c.a0 = 0; c.a1 = -65536; c.a2 = (s.w - 1) * 65536; c.a3 = 65536; c.a4 = 0; c.a5 = 0;
      break;
    case IMGTX_GEOM_ROT270:  // out[y][x] = in[h-1-x][y]
      c.a0 = 0; c.a1 = 65536; c.a2 = 0; c.a3 = -65536; c.a4 = 0; c.a5 = (s.h - 1) * 65536;
      break;
    case IMGTX_GEOM_CROP_REFLECT:
      c.a2 = (q.dx - q.pad) * 65536; c.a5 = (q.dy - q.pad) * 65536; c.flags = kGeomReflect;
      break;
    default:
      break;
  }
  return c;
}

// source byte offset of output pixel (x, y), or -1 for fill
__device__ __forceinline__ int geom_src(const CompiledGeom& c, const Shape& s, int x, int y, const int* xt,
                                        const int* yt) {
  int xin, yin;
  if (c.flags & kGeomTables) {
    xin = xt[x]; yin = yt[y];
    if (xin < 0 || yin < 0) return -1;
  } else {
    xin = (c.a2 + y * c.a1 + x * c.a0) >> 16;
    yin = (c.a5 + y * c.a4 + x * c.a3) >> 16;
    if (c.flags & kGeomReflect) {
      xin = reflect101(xin, s.w); yin = reflect101(yin, s.h);
    } else if (xin < 0 || xin >= s.w || yin < 0 || yin >= s.h) {
      return -1;
    }
  }
  return (yin * s.w + xin) * s.c;
}

// ImagingScaleAffine index tables: xo = m2 + m0*0.5; xo += m0 per column, accumulated in double.
__device__ inline void build_scale_tables(const imgtx_geom_t& q, const Shape& s, int* xt, int* yt) {
  double xo = q.m[2] + q.m[0] * 0.5;
  for (int x = 0; x < s.ow; ++x) {
    int xin = (xo < 0.0) ? -1 : (int)xo;
    xt[x] = (xin >= 0 && xin < s.w) ? xin : -1;
    xo += q.m[0];
  }
  double yo = q.m[5] + q.m[4] * 0.5;
  for (int y = 0; y < s.oh; ++y) {
    int yin = (yo < 0.0) ? -1 : (int)yo;
    yt[y] = (yin >= 0 && yin < s.h) ? yin : -1;
    yo += q.m[4];
  }
}

struct OpGeometry : NoSpill {
  using Params = GeomParams;
  static constexpr int kThreads = 256;
  // layout: [CompiledGeom x 1][x table ow][y table oh][offset table oh*ow (staged only)]
  __host__ __device__ static size_t head_bytes(const Shape& s) { return align_up(sizeof(CompiledGeom) + (size_t)(s.ow + s.oh) * 4, 16); }
  static size_t scratch_bytes(const Shape& s, int G) {
    return head_bytes(s) + (size_t)s.oh * s.ow * 4 + 16;  // G-independent: one table, rebuilt per image if needed
  }
  static size_t scratch_bytes_direct(const Shape& s) { return head_bytes(s); }

  __device__ static void build_offsets(const imgtx_geom_t& q, const Shape& s, uint8_t* scratch, bool with_table) {
    CompiledGeom* cg = reinterpret_cast<CompiledGeom*>(scratch);
    int* xt = reinterpret_cast<int*>(scratch + sizeof(CompiledGeom));
    int* yt = xt + s.ow;
    if (threadIdx.x == 0) {
      *cg = compile_geom(q, s);
      if (cg->flags & kGeomTables) build_scale_tables(q, s, xt, yt);
    }
    __syncthreads();
    if (with_table) {
      int* offs = reinterpret_cast<int*>(scratch + head_bytes(s));
      const CompiledGeom c = *cg;
      for (int y = threadIdx.x >> 5; y < s.oh; y += kThreads >> 5)
        for (int x = threadIdx.x & 31; x < s.ow; x += 32) offs[y * s.ow + x] = geom_src(c, s, x, y, xt, yt);
      __syncthreads();
    }
  }

  __device__ static void prepare(const Shape& s, const Params& p, uint8_t* scratch, bool staged) {
    if (p.per_image == nullptr) build_offsets(p.uni, s, scratch, staged);
  }

  template <bool STAGED>
  __device__ static void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0, const Shape& s,
                             const Params& p, uint8_t* scratch, uint8_t*) {
    const int tid = threadIdx.x;
    const bool uniform = (p.per_image == nullptr);
    const int npix = s.oh * s.ow;
    if (STAGED) {
      const int* offs = reinterpret_cast<const int*>(scratch + head_bytes(s));
      const uint32_t fill = p.fill[0] | (p.fill[1] << 8) | (p.fill[2] << 16);
      for (int j = 0; j < g; ++j) {
        if (!uniform) {
          __syncthreads();
          build_offsets(p.per_image[img0 + j], s, scratch, true);
        }
        const uint8_t* im = src + (size_t)j * s.in_bytes;
        uint8_t* om = dst + (size_t)j * s.out_bytes;
        if (s.c == 3 && (npix & 3) == 0) {
          const int4* offs4 = reinterpret_cast<const int4*>(offs);
          uint32_t* ow32 = reinterpret_cast<uint32_t*>(om);
          for (int q4 = tid; q4 < (npix >> 2); q4 += kThreads) {
            const int4 o = offs4[q4];
            const uint32_t p0 = o.x >= 0 ? load_rgb(im, o.x) : fill;
            const uint32_t p1 = o.y >= 0 ? load_rgb(im, o.y) : fill;
            const uint32_t p2 = o.z >= 0 ? load_rgb(im, o.z) : fill;
            const uint32_t p3 = o.w >= 0 ? load_rgb(im, o.w) : fill;
            uint32_t w0, w1, w2;
            pack_rgb4(p0, p1, p2, p3, w0, w1, w2);
            ow32[q4 * 3] = w0; ow32[q4 * 3 + 1] = w1; ow32[q4 * 3 + 2] = w2;
          }
        } else {
          for (int px = tid; px < npix; px += kThreads) {
            const int o = offs[px];
            for (int ch = 0; ch < s.c; ++ch) om[px * s.c + ch] = o >= 0 ? im[o + ch] : p.fill[ch];
          }
        }
      }
    } else {
      // direct mode: one image per call, evaluate the map per pixel
      if (!uniform) {
        __syncthreads();
        build_offsets(p.per_image[img0], s, scratch, false);
      }
      const CompiledGeom c = *reinterpret_cast<const CompiledGeom*>(scratch);
      const int* xt = reinterpret_cast<const int*>(scratch + sizeof(CompiledGeom));
      const int* yt = xt + s.ow;
      for (int y = tid >> 5; y < s.oh; y += kThreads >> 5) {
        for (int x = tid & 31; x < s.ow; x += 32) {
          const int o = geom_src(c, s, x, y, xt, yt);
          uint8_t* op = dst + ((size_t)y * s.ow + x) * s.c;
          for (int ch = 0; ch < s.c; ++ch) op[ch] = o >= 0 ? src[o + ch] : p.fill[ch];
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
