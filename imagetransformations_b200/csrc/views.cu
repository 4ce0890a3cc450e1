// Fused test-time-augmentation view expansion: one image in, V normalised f32 views out, one pass.
//
//   reference: fall_2025/tta_initial_claudeFixed.py:37-78,218-230 (16 views: tensor_to_pil, then 1 flip, 5 crop(24)+
//              BICUBIC resize, 5 ImageEnhance.Brightness, 5 ImageEnhance.Contrast, each followed by pil_to_tensor),
//              fall_2025/TTA_transforms.py:69-138 (three strategies of 4 views: identity, RandomHorizontalFlip(p=1),
//              RandomCrop(32, 4, 'reflect'), flip+crop, Resize(36|40)+CenterCrop(32), ColorJitter brightness /
//              contrast; all end in ToTensor+Normalize).
//
// The image is read from HBM once (uint8 HWC, or the normalised f32 CHW tensor that is first quantised exactly like
// tensor_to_pil), lives in shared memory, and every view is produced from it: index maps for flips / reflect crops,
// Pillow's two-pass Q22 resample for the crop-resize views (uint8 intermediate in shared memory), Pillow's f32 blend
// for the enhancers, and a 768-entry table for ToTensor+Normalize ((u/255 - mean)/std with IEEE division, evaluated
// once per CTA).  Roofline: HBM writes — V * 12 * h * w bytes out per image versus 3 (or 12) * h * w in.
#include "common.cuh"
#include "lut.cuh"
#include <algorithm>
#include <string.h>

namespace imgtx {

constexpr int kMaxViews = 32;
constexpr int kViewThreads = 256;

struct ViewSet {
  imgtx_view_t v[kMaxViews];
  int n_views;
  int table_off[kMaxViews + 1];  // word offsets of the resample tables inside the table blob
  float in_mean[4], in_std[4], out_mean[4], out_std[4];
  int in_is_f32, any_contrast, tmp_rows;
};

__device__ __forceinline__ int pil_luma_v(int r, int g, int b) { return (19595 * r + 38470 * g + 7471 * b + 0x8000) >> 16; }

// byte c of a packed pixel as a byte offset into a float[256] table
__device__ __forceinline__ float lut_at(const float* t, uint32_t px, int c) {
  const uint32_t off = c == 0 ? (px << 2) & 0x3FCu : (px >> (8 * c - 2)) & 0x3FCu;
  return *reinterpret_cast<const float*>(reinterpret_cast<const char*>(t + c * 256) + off);
}
__device__ __forceinline__ uint32_t pack_q22(int a0, int a1, int a2) {
  return (uint32_t)clip8(a0 >> 22) | ((uint32_t)clip8(a1 >> 22) << 8) | ((uint32_t)clip8(a2 >> 22) << 16);
}

// Pillow's two-pass resample of one view on packed 32-bit pixels.  KS > 0: the table's tap count, fully unrolled
// (records are zero-padded to KS taps, and both pixel arrays are padded so the extra taps read harmless words);
// KS == 0: run-time tap count.
template <int KS>
__device__ __forceinline__ void resize_view(const uint32_t* u32, uint32_t* tmp32, float* ov, const float* NT,
                                            const int32_t* ht, const int32_t* vt, int h, int w, int ox, int oy, int bh,
                                            int left, int top, int c_rows, int c_ty, int c_tx) {
  const int hw = h * w;
  const int hk = KS ? KS : ht[1], vk = KS ? KS : vt[1];
  if (c_ty < c_rows) {
    // horizontal pass: thread = column of rows c_ty, c_ty + c_rows, ...; its coefficient record stays in registers
    const int32_t* rec = ht + 2 + (left + c_tx) * (2 + hk);
    const int xmin = rec[0];
    if (KS) {
      int kk[KS ? KS : 1];
#pragma unroll
      for (int k = 0; k < KS; ++k) kk[k] = rec[2 + k];
      for (int y = c_ty; y < bh; y += c_rows) {
        const uint32_t* row = u32 + (oy + y) * w + ox + xmin;
        int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
#pragma unroll
        for (int k = 0; k < KS; ++k) {
          const uint32_t px = row[k];
          a0 += (int)(px & 255u) * kk[k]; a1 += (int)((px >> 8) & 255u) * kk[k]; a2 += (int)(px >> 16) * kk[k];
        }
        tmp32[y * w + c_tx] = pack_q22(a0, a1, a2);
      }
    } else {
      const int nk = rec[1];
      for (int y = c_ty; y < bh; y += c_rows) {
        const uint32_t* row = u32 + (oy + y) * w + ox + xmin;
        int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
        for (int k = 0; k < nk; ++k) {
          const uint32_t px = row[k];
          const int kk = rec[2 + k];
          a0 += (int)(px & 255u) * kk; a1 += (int)((px >> 8) & 255u) * kk; a2 += (int)(px >> 16) * kk;
        }
        tmp32[y * w + c_tx] = pack_q22(a0, a1, a2);
      }
    }
  }
  __syncthreads();
  if (c_ty < c_rows) {
    // vertical pass: thread = pixel (y, x); a warp shares the row's record (broadcast loads)
    for (int y = c_ty; y < h; y += c_rows) {
      const int32_t* rec = vt + 2 + (top + y) * (2 + vk);
      const uint32_t* col = tmp32 + rec[0] * w + c_tx;
      int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
      if (KS) {
#pragma unroll
        for (int k = 0; k < KS; ++k) {
          const uint32_t px = col[k * w];
          const int kk = rec[2 + k];
          a0 += (int)(px & 255u) * kk; a1 += (int)((px >> 8) & 255u) * kk; a2 += (int)(px >> 16) * kk;
        }
      } else {
        const int nk = rec[1];
        for (int k = 0; k < nk; ++k) {
          const uint32_t px = col[k * w];
          const int kk = rec[2 + k];
          a0 += (int)(px & 255u) * kk; a1 += (int)((px >> 8) & 255u) * kk; a2 += (int)(px >> 16) * kk;
        }
      }
      float* o = ov + y * w + c_tx;
      __stcs(o, NT[clip8(a0 >> 22)]);
      __stcs(o + hw, NT[256 + clip8(a1 >> 22)]);
      __stcs(o + 2 * hw, NT[512 + clip8(a2 >> 22)]);
    }
  }
}

constexpr int kPixelPad = 8;   // words after the image, read (times a zero coefficient) by the unrolled taps
constexpr int kTmpPadRows = 6;

__global__ void __launch_bounds__(kViewThreads, 5)
views_kernel(const void* __restrict__ in, float* __restrict__ out, int64_t n, int h, int w, ViewSet vs,
             const float* __restrict__ params, const int32_t* __restrict__ tables_g, int table_words) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int hw = h * w;
  const int tid = threadIdx.x;
  const int V = vs.n_views;
  float* NT = reinterpret_cast<float*>(smem);                        // [3][256] normalise table
  float* VL = NT + 768;                                              // [3][256] the current view's byte -> float map
  int32_t* tables = reinterpret_cast<int32_t*>(VL + 768);            // table_words
  uint32_t* u32 = reinterpret_cast<uint32_t*>(tables + ((table_words + 3) & ~3));  // hw packed pixels r | g<<8 | b<<16
  uint32_t* tmp32 = u32 + ((hw + kPixelPad + 3) & ~3);               // (tmp_rows + pad) * w packed pixels
  // threads are laid out as (row, quad of 4 pixels) / (row, pixel) once, so no loop divides by a run-time size
  const bool quads_ok = (w & 3) == 0 && (w >> 2) <= kViewThreads;
  const int qpr = quads_ok ? (w >> 2) : 1, q_rows = kViewThreads / qpr, q_ty = tid / qpr, q_tx = tid - q_ty * qpr;
  const bool cols_ok = w <= kViewThreads;
  const int c_rows = cols_ok ? kViewThreads / w : 1, c_ty = cols_ok ? tid / w : 0, c_tx = cols_ok ? tid - c_ty * w : 0;
  __shared__ unsigned long long s_luma;
  __shared__ float s_par[kMaxViews * 2];

  for (int i = tid; i < 768; i += kViewThreads) {
    const int c = i >> 8;
    NT[i] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)(i & 255), 255.0f), vs.out_mean[c]), vs.out_std[c]);
  }
  for (int i = tid; i < table_words; i += kViewThreads) tables[i] = tables_g[i];
  for (int i = tid; i < kPixelPad; i += kViewThreads) u32[hw + i] = 0u;
  for (int i = tid; i < (vs.tmp_rows + kTmpPadRows) * w; i += kViewThreads) tmp32[i] = 0u;

  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    __syncthreads();
    if (tid == 0) s_luma = 0ull;
    for (int i = tid; i < V * 2; i += kViewThreads) s_par[i] = params[img * V * 2 + i];
    // ---- load / quantise into packed pixels ----
    if (vs.in_is_f32) {
      const float* xi = reinterpret_cast<const float*>(in) + img * 3 * (int64_t)hw;
      for (int p = tid; p < hw; p += kViewThreads) {
        uint32_t px = 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float v = __fadd_rn(__fmul_rn(__ldcs(xi + c * hw + p), vs.in_std[c]), vs.in_mean[c]);
          v = fminf(fmaxf(v, 0.0f), 1.0f);
          px |= (uint32_t)(__float2int_rz(__fmul_rn(v, 255.0f)) & 255) << (8 * c);
        }
        u32[p] = px;
      }
    } else {
      const uint8_t* xi = reinterpret_cast<const uint8_t*>(in) + img * 3 * (int64_t)hw;
      for (int p = tid; p < hw; p += kViewThreads)
        u32[p] = (uint32_t)xi[p * 3] | ((uint32_t)xi[p * 3 + 1] << 8) | ((uint32_t)xi[p * 3 + 2] << 16);
    }
    __syncthreads();
    int degenerate = 0;
    if (vs.any_contrast) {
      unsigned acc = 0;
      for (int p = tid; p < hw; p += kViewThreads) {
        const uint32_t px = u32[p];
        acc += pil_luma_v(px & 255u, (px >> 8) & 255u, px >> 16);
      }
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if ((tid & 31) == 0) atomicAdd(&s_luma, (unsigned long long)acc);
      __syncthreads();
      degenerate = (int)((double)s_luma / (double)hw + 0.5);
    }

    for (int vi = 0; vi < V; ++vi) {
      const imgtx_view_t vw = vs.v[vi];
      float* ov = out + ((img * V + vi) * 3) * (int64_t)hw;
      const float p0 = s_par[vi * 2], p1 = s_par[vi * 2 + 1];
      switch (vw.kind) {
        case IMGTX_VIEW_IDENTITY:
        case IMGTX_VIEW_FLIP:
        case IMGTX_VIEW_CROP_REFLECT:
        case IMGTX_VIEW_FLIP_CROP_REFLECT:
        case IMGTX_VIEW_BRIGHTNESS:
        case IMGTX_VIEW_CONTRAST: {
          const bool flip = vw.kind == IMGTX_VIEW_FLIP || vw.kind == IMGTX_VIEW_FLIP_CROP_REFLECT;
          const bool crop = vw.kind == IMGTX_VIEW_CROP_REFLECT || vw.kind == IMGTX_VIEW_FLIP_CROP_REFLECT;
          const bool blend = vw.kind == IMGTX_VIEW_BRIGHTNESS || vw.kind == IMGTX_VIEW_CONTRAST;
          const int d = vw.kind == IMGTX_VIEW_CONTRAST ? degenerate : 0;
          const int ci = (int)p0 - vw.i0, cj = (int)p1 - vw.i0;  // crop offsets minus padding (row, col)
          const float* lut = NT;
          if (blend) {
            // the enhancer is a byte map per image: compose it with ToTensor+Normalize once (256 blends instead of 3*h*w)
            __syncthreads();   // the previous view may still read VL
            for (int e = tid; e < 256; e += kViewThreads) {
              const int bv = pil_blend(d, e, p0);
              VL[e] = NT[bv]; VL[256 + e] = NT[256 + bv]; VL[512 + e] = NT[512 + bv];
            }
            __syncthreads();
            lut = VL;
          }
          if (quads_ok) {
            // thread = four adjacent pixels of a row: the pixels are fetched once for the three planes
            if (q_ty < q_rows) {
              for (int y = q_ty; y < h; y += q_rows) {
                uint32_t px[4];
                if (!crop) {
                  const uint4 v = *reinterpret_cast<const uint4*>(u32 + y * w + (flip ? w - 4 - q_tx * 4 : q_tx * 4));
                  px[0] = flip ? v.w : v.x; px[1] = flip ? v.z : v.y; px[2] = flip ? v.y : v.z; px[3] = flip ? v.x : v.w;
                } else {
                  const uint32_t* row = u32 + reflect101(y + ci, h) * w;
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    int x = reflect101(q_tx * 4 + e + cj, w);
                    if (flip) x = w - 1 - x;
                    px[e] = row[x];
                  }
                }
                float4* o4 = reinterpret_cast<float4*>(ov + y * w) + q_tx;
#pragma unroll
                for (int c = 0; c < 3; ++c)
                  __stcs(o4 + c * (hw >> 2),
                         make_float4(lut_at(lut, px[0], c), lut_at(lut, px[1], c), lut_at(lut, px[2], c), lut_at(lut, px[3], c)));
              }
            }
          } else {
            for (int i = tid; i < hw * 3; i += kViewThreads) {
              const int c = i / hw, p = i - c * hw;
              int y = p / w, x = p - y * w;
              if (crop) {
                y = reflect101(y + ci, h);
                x = reflect101(x + cj, w);
              }
              if (flip) x = w - 1 - x;
              ov[i] = lut[c * 256 + ((u32[y * w + x] >> (8 * c)) & 255u)];
            }
          }
          break;
        }
        case IMGTX_VIEW_CROP_RESIZE:
        case IMGTX_VIEW_RESIZE_CENTER_CROP: {
          // source box (ox, oy, bw, bh) -> resized (rw, rh) -> window (left, top, w, h) of the resized image
          const bool cr = vw.kind == IMGTX_VIEW_CROP_RESIZE;
          const int ox = cr ? (int)p0 : 0, oy = cr ? (int)p1 : 0;
          const int bh = cr ? vw.i0 : h;
          const int left = cr ? 0 : vw.i0, top = cr ? 0 : vw.i1;
          const int32_t* ht = tables + vs.table_off[vw.table_h];
          const int32_t* vt = tables + vs.table_off[vw.table_v];
          const int hk = ht[1], vk = vt[1];
          __syncthreads();  // tmp may still be read by the previous resample view
          if (cols_ok) {
            if (hk == 5 && vk == 5)
              resize_view<5>(u32, tmp32, ov, NT, ht, vt, h, w, ox, oy, bh, left, top, c_rows, c_ty, c_tx);
            else if (hk == 3 && vk == 3)
              resize_view<3>(u32, tmp32, ov, NT, ht, vt, h, w, ox, oy, bh, left, top, c_rows, c_ty, c_tx);
            else
              resize_view<0>(u32, tmp32, ov, NT, ht, vt, h, w, ox, oy, bh, left, top, c_rows, c_ty, c_tx);
          } else {
            for (int i = tid; i < bh * w; i += kViewThreads) {
              const int y = i / w, xx = i - y * w;
              const int32_t* rec = ht + 2 + (left + xx) * (2 + hk);
              const int xmin = rec[0], nk = rec[1];
              const uint32_t* row = u32 + (oy + y) * w + ox + xmin;
              int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
              for (int k = 0; k < nk; ++k) {
                const uint32_t px = row[k];
                const int kk = rec[2 + k];
                a0 += (int)(px & 255u) * kk; a1 += (int)((px >> 8) & 255u) * kk; a2 += (int)(px >> 16) * kk;
              }
              tmp32[i] = pack_q22(a0, a1, a2);
            }
            __syncthreads();
            for (int i = tid; i < hw * 3; i += kViewThreads) {
              const int c = i / hw, p = i - c * hw;
              const int y = p / w, x = p - y * w;
              const int32_t* rec = vt + 2 + (top + y) * (2 + vk);
              const int ymin = rec[0], nk = rec[1];
              int a = 1 << 21;
              for (int k = 0; k < nk; ++k) a += (int)((tmp32[(ymin + k) * w + x] >> (8 * c)) & 255u) * rec[2 + k];
              ov[i] = NT[c * 256 + clip8(a >> 22)];
            }
          }
          break;
        }
      }
    }
  }
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_views_f32(const void* in, int in_is_f32, float* out, int64_t n, int h, int w, int n_views,
                               const imgtx_view_t* views_host, const float* params_dev, const int32_t* tables_dev,
                               const int32_t* table_offsets_host, int n_tables, const float* in_mean_host,
                               const float* in_std_host, const float* out_mean_host, const float* out_std_host,
                               imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && n >= 0, "bad image shape");
  IMGTX_CHECK_ARG(n_views >= 1 && n_views <= kMaxViews && views_host, "1..%d views", kMaxViews);
  IMGTX_CHECK_ARG(n_tables >= 0 && n_tables <= kMaxViews, "too many resample tables");
  IMGTX_CHECK_ARG(out_mean_host && out_std_host, "output mean/std missing");
  if (in_is_f32) IMGTX_CHECK_ARG(in_mean_host && in_std_host, "input mean/std missing (tensor_to_pil prologue)");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out && params_dev, "null pointer");
  ViewSet vs;
  memset(&vs, 0, sizeof(vs));
  vs.n_views = n_views;
  vs.in_is_f32 = in_is_f32;
  int tmp_rows = 1;
  for (int i = 0; i < n_views; ++i) {
    vs.v[i] = views_host[i];
    const int k = views_host[i].kind;
    IMGTX_CHECK_ARG(k >= IMGTX_VIEW_IDENTITY && k <= IMGTX_VIEW_RESIZE_CENTER_CROP, "unknown view kind %d", k);
    if (k == IMGTX_VIEW_CONTRAST) vs.any_contrast = 1;
    if (k == IMGTX_VIEW_CROP_RESIZE || k == IMGTX_VIEW_RESIZE_CENTER_CROP) {
      IMGTX_CHECK_ARG(tables_dev && table_offsets_host && views_host[i].table_h >= 0 && views_host[i].table_h < n_tables &&
                          views_host[i].table_v >= 0 && views_host[i].table_v < n_tables, "bad resample table index");
      tmp_rows = std::max(tmp_rows, k == IMGTX_VIEW_CROP_RESIZE ? views_host[i].i0 : h);
      if (k == IMGTX_VIEW_CROP_RESIZE) IMGTX_CHECK_ARG(views_host[i].i0 > 0 && views_host[i].i0 <= std::min(h, w), "bad crop size");
    }
  }
  int table_words = 0;
  for (int i = 0; i <= n_tables; ++i) vs.table_off[i] = table_offsets_host ? table_offsets_host[i] : 0;
  if (n_tables > 0) table_words = table_offsets_host[n_tables];
  for (int i = 0; i < 3; ++i) {
    vs.in_mean[i] = in_mean_host ? in_mean_host[i] : 0.f;
    vs.in_std[i] = in_std_host ? in_std_host[i] : 1.f;
    vs.out_mean[i] = out_mean_host[i];
    vs.out_std[i] = out_std_host[i];
  }
  vs.tmp_rows = tmp_rows;
  const int hw = h * w;
  size_t smem = 2 * 768 * 4 + (size_t)((table_words + 3) & ~3) * 4 + (size_t)((hw + kPixelPad + 3) & ~3) * 4 +
                ((size_t)(tmp_rows + kTmpPadRows) * w * 4 + 15) / 16 * 16 + 16;
  IMGTX_CHECK_ARG(smem <= (size_t)max_smem_optin(), "image (%dx%d) does not fit in shared memory for fused views", h, w);
  IMGTX_CUDA_OK(cudaFuncSetAttribute(views_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)max_smem_optin() / (smem + 1024)));
  int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * per_sm);
  views_kernel<<<(unsigned)grid, kViewThreads, smem, (cudaStream_t)stream>>>(in, out, n, h, w, vs, params_dev, tables_dev,
                                                                            table_words);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
