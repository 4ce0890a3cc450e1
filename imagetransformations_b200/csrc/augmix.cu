// Fused AugMix (fall_2025/AugMix.py:45-62): one pass per image.
//
//   q   = TF.to_pil_image(x)            (trunc(x*255) mod 256, kept in shared memory as HWC uint8)
//   u_i = op_{i,d-1}( ... op_{i,0}(q))  for each of `width` chains, WITHOUT materialising intermediates:
//         NEAREST geometric ops (rotate / shear / translate: Pillow 16.16 fixed-point affine gathers) compose as
//         index maps applied stage by stage (each stage rounds separately, exactly like Pillow), point ops
//         (posterize / solarize / equalize) compose as per-channel 256-entry LUTs; a pixel that leaves the image at
//         geometric stage k takes the fill value 0 pushed through the point ops after k.
//         `equalize` needs the histogram of its input image: one extra sweep over the shared-memory image evaluating
//         the chain prefix.
//   mix = sum_i f32(w_i) * (u_i / 255)   sequential f32
//   out = f32(1-m) * x + f32(m) * mix
//
// HBM traffic per image: x read once (12*h*w B; the second read in the output phase hits L2) + out written once.
// Roofline: HBM, 24 B per pixel.
#include "common.cuh"
#include "lut.cuh"
#include <algorithm>

namespace imgtx {

// L2 residency hints for the two reads of the f32 image: the first read (quantise) asks L2 to keep the lines, the second
// (final blend, one image-time later) marks them first to go; the output bypasses L1.  Without them the second read of
// a 602 KB image misses L2 behind the 148 outputs written in between (round 1: 1.75x read traffic).
__device__ __forceinline__ uint64_t l2_policy_keep() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_drop() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ float4 ldg_hint_f4(const float* ptr, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(ptr), "l"(pol));
  return v;
}
__device__ __forceinline__ void stg_hint_f4(float* ptr, const float4& v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(ptr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}

struct AugOp {  // == imgtx_augmix_op_t
  int32_t kind;
  int32_t a[6];
  int32_t param;
};

__device__ __forceinline__ uint8_t augmix_point(const AugOp& op, int v) {
  if (op.kind == IMGTX_AUG_POSTERIZE) return (uint8_t)(v & op.param);
  return (uint8_t)(v < op.param ? v : 255 - v);  // SOLARIZE
}

// Walk the geometric stages [0, upto) backwards from output pixel (x, y).  Returns the source byte offset, or -1
// with fillk = the stage whose source fell outside the image.
__device__ __forceinline__ int trace_back(const AugOp* ops, int upto, int x, int y, int w, int h, int& fillk) {
  for (int k = upto - 1; k >= 0; --k) {
    if (ops[k].kind != IMGTX_AUG_GEOM) continue;
    const int nx = (ops[k].a[2] + y * ops[k].a[1] + x * ops[k].a[0]) >> 16;
    const int ny = (ops[k].a[5] + y * ops[k].a[4] + x * ops[k].a[3]) >> 16;
    if (nx < 0 || nx >= w || ny < 0 || ny >= h) {
      fillk = k;
      return -1;
    }
    x = nx;
    y = ny;
  }
  return (y * w + x) * 3;
}

// comp[c][v] = point ops of stages [0, upto) applied in order to v; fills[k][c] = 0 pushed through the point ops of
// stages (k, upto).  stage_lut holds one [3][256] table per point-op stage.
__device__ void build_composed(const AugOp* ops, const uint8_t* stage_lut, int upto, uint8_t* comp, uint8_t* fills) {
  for (int t = threadIdx.x; t < 768; t += blockDim.x) {
    const int c = t >> 8;
    int v = t & 255;
    for (int k = 0; k < upto; ++k)
      if (ops[k].kind >= IMGTX_AUG_POSTERIZE) v = stage_lut[k * 768 + c * 256 + v];
    comp[t] = (uint8_t)v;
  }
  for (int t = threadIdx.x; t < upto * 3; t += blockDim.x) {
    const int k = t / 3, c = t - k * 3;
    int v = 0;
    for (int kk = k + 1; kk < upto; ++kk)
      if (ops[kk].kind >= IMGTX_AUG_POSTERIZE) v = stage_lut[kk * 768 + c * 256 + v];
    fills[k * 4 + c] = (uint8_t)v;
  }
}

// Trace four horizontally adjacent output pixels back through the chain's geometric records [s0, ng) (stored in trace-back
// order).  A pixel whose source leaves the image at some stage stops there: ok = false, fk = that stage's index.
template <bool kNeedFk>
__device__ __forceinline__ void trace4(const int32_t* g, int s0, int ng, int w, int h, int (&xs)[4], int (&ys)[4],
                                       int (&fk)[4], bool (&ok)[4]) {
  for (int s2 = s0; s2 < ng; ++s2) {
    const int4 A = *reinterpret_cast<const int4*>(g + 4 + s2 * 8);
    const int4 Bv = *reinterpret_cast<const int4*>(g + 8 + s2 * 8);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int nx = (A.z + ys[k] * A.y + xs[k] * A.x) >> 16;
      const int ny = (Bv.y + ys[k] * Bv.x + xs[k] * A.w) >> 16;
      const bool in = ok[k] && (unsigned)nx < (unsigned)w && (unsigned)ny < (unsigned)h;
      if (kNeedFk && ok[k] && !in) fk[k] = Bv.z;
      ok[k] = in;
      xs[k] = in ? nx : xs[k];
      ys[k] = in ? ny : ys[k];
    }
  }
}

// Do all point-op stages [0, upto) of a chain map the byte 0 to 0 in every channel?  (posterize and equalize always do,
// solarize unless its threshold is <= 0.)  Then the fill value 0 that a geometric stage k puts in, pushed through the point
// ops after k, equals comp[c][0] for every k, and "outside" is simply a source pixel whose bytes are zero.
__device__ __forceinline__ bool zero_is_fixed(const AugOp* ops, const uint8_t* stage_lut, int upto) {
  bool zf = true;
  for (int k = 0; k < upto; ++k)
    if (ops[k].kind >= IMGTX_AUG_POSTERIZE)
      zf = zf && stage_lut[k * 768] == 0 && stage_lut[k * 768 + 256] == 0 && stage_lut[k * 768 + 512] == 0;
  return zf;
}

__global__ void __launch_bounds__(1024)
augmix_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t n, int h, int w, int width,
              int depth, const AugOp* __restrict__ ops_g, const float* __restrict__ wts_g, int use_stash) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int hw = h * w;
  const int nstage = width * depth;
  uint8_t* q = smem;                                                   // hw*3 (+pad)
  uint8_t* stage_lut = q + ((hw * 3 + 4 + 15) & ~15);                  // nstage * 768
  uint8_t* comp = stage_lut + (size_t)nstage * 768;                    // width * 768
  uint8_t* fills = comp + (size_t)width * 768;                         // nstage * 4
  uint32_t* hist = reinterpret_cast<uint32_t*>(fills + ((nstage * 4 + 15) & ~15));  // 768
  uint32_t* base = hist + 768;                                         // 768: histogram of q itself
  uint32_t* tmp = base + 768;                                          // 768: histogram of a gathered image before its LUT
  float* T = reinterpret_cast<float*>(tmp + 768);                      // 256: u/255
  float* FT = T + 256;                                                 // width * 768: w_i * T[comp_i[c][v]]
  float* FF = FT + (size_t)width * 768;                                // nstage * 4:  w_i * T[fill]
  int32_t* G = reinterpret_cast<int32_t*>(FF + (size_t)nstage * 4);    // per chain: [ng,-,-,-] + depth * [a0..a5,k,-]
  const int gstride = 4 + depth * 8;
  AugOp* ops = reinterpret_cast<AugOp*>(G + (size_t)width * gstride);  // nstage
  float* wts = reinterpret_cast<float*>(ops + nstage);                 // width + 2 (+pad)
  float* stash = use_stash ? wts + ((width + 2 + 3) & ~3) : nullptr;   // blockDim * 12 floats (when they fit)
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const bool vec = (hw & 3) == 0 && (w & 3) == 0;
  const int zoff = hw * 3;                                             // a pixel of zero bytes behind the image
  if (tid < 4) q[zoff + tid] = 0;

  for (int i = tid; i < 256; i += nt) T[i] = __fdiv_rn((float)i, 255.0f);
  const uint64_t keep = l2_policy_keep(), drop = l2_policy_drop();

  // the records and weights of the NEXT image ride in one register per thread (fetched while this image is gathered), so
  // an image does not start with a dependent global load that stalls the whole CTA
  const int nrec = nstage * 8, nhead = nrec + width + 2;
  const bool pre = nhead <= nt;
  auto fetch_head = [&](int64_t im) -> int32_t {
    if (im >= n || tid >= nhead) return 0;
    return tid < nrec ? reinterpret_cast<const int32_t*>(ops_g + im * nstage)[tid]
                      : __float_as_int(wts_g[im * (width + 2) + (tid - nrec)]);
  };
  int32_t head = pre ? fetch_head(blockIdx.x) : 0;

  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    const float* xi = x + img * 3 * (int64_t)hw;
    float* oi = out + img * 3 * (int64_t)hw;
    __syncthreads();  // previous image fully written before its tables are overwritten
    if (pre) {
      if (tid < nrec) reinterpret_cast<int32_t*>(ops)[tid] = head;
      else if (tid < nhead) wts[tid - nrec] = __int_as_float(head);
    } else {
      const int32_t* src = reinterpret_cast<const int32_t*>(ops_g + img * nstage);
      int32_t* dst = reinterpret_cast<int32_t*>(ops);
      for (int i = tid; i < nrec; i += nt) dst[i] = src[i];
      for (int i = tid; i < width + 2; i += nt) wts[i] = wts_g[img * (width + 2) + i];
    }
    for (int i = tid; i < 768; i += nt) base[i] = 0u;
    __syncthreads();
    // the chain's geometric stages compacted in trace-back order; does any equalize see the quantised image itself?
    if (tid < width) {
      int32_t* g = G + tid * gstride;
      int ng = 0;
      for (int k = depth - 1; k >= 0; --k) {
        const AugOp& op = ops[tid * depth + k];
        if (op.kind != IMGTX_AUG_GEOM) continue;
        int32_t* r = g + 4 + ng * 8;
        for (int e = 0; e < 6; ++e) r[e] = op.a[e];
        r[6] = k;
        ++ng;
      }
      g[0] = ng;
    }
    bool need_base = false;
    for (int ci = 0; ci < width; ++ci)
      for (int j = 0; j < depth; ++j) {
        const int kind = ops[ci * depth + j].kind;
        if (kind == IMGTX_AUG_GEOM) break;
        if (kind == IMGTX_AUG_EQUALIZE) { need_base = true; break; }
      }
    // ---- phase A: quantise to HWC uint8 in shared memory (and count its bytes when an equalize will ask) ----
    if ((hw & 3) == 0) {
      uint32_t* q32 = reinterpret_cast<uint32_t*>(q);
#pragma unroll 4
      for (int p4 = tid; p4 < (hw >> 2); p4 += nt) {
        uint32_t v[3][4];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float4 f = ldg_hint_f4(xi + (int64_t)c * hw + p4 * 4, keep);
          v[c][0] = (uint32_t)(__float2int_rz(__fmul_rn(f.x, 255.0f)) & 255);
          v[c][1] = (uint32_t)(__float2int_rz(__fmul_rn(f.y, 255.0f)) & 255);
          v[c][2] = (uint32_t)(__float2int_rz(__fmul_rn(f.z, 255.0f)) & 255);
          v[c][3] = (uint32_t)(__float2int_rz(__fmul_rn(f.w, 255.0f)) & 255);
        }
        q32[p4 * 3 + 0] = v[0][0] | (v[1][0] << 8) | (v[2][0] << 16) | (v[0][1] << 24);
        q32[p4 * 3 + 1] = v[1][1] | (v[2][1] << 8) | (v[0][2] << 16) | (v[1][2] << 24);
        q32[p4 * 3 + 2] = v[2][2] | (v[0][3] << 8) | (v[1][3] << 16) | (v[2][3] << 24);
        if (need_base) {
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int k = 0; k < 4; ++k) atomicAdd(&base[c * 256 + (int)v[c][k]], 1u);
        }
      }
    } else {
      for (int i = tid; i < hw * 3; i += nt) {
        const int c = i / hw, p = i - c * hw;
        const int v = __float2int_rz(__fmul_rn(xi[i], 255.0f)) & 255;
        q[p * 3 + c] = (uint8_t)v;
        if (need_base) atomicAdd(&base[c * 256 + v], 1u);
      }
    }
    __syncthreads();

    // ---- phase B: per-chain tables ----
    for (int ci = 0; ci < width; ++ci) {
      const AugOp* cops = ops + ci * depth;
      uint8_t* clut = stage_lut + (size_t)ci * depth * 768;
      uint8_t* ccomp = comp + ci * 768;
      uint8_t* cfill = fills + ci * depth * 4;
      const int32_t* g = G + ci * gstride;
      for (int j = 0; j < depth; ++j) {
        const int kind = cops[j].kind;
        if (kind == IMGTX_AUG_POSTERIZE || kind == IMGTX_AUG_SOLARIZE) {
          for (int t = tid; t < 768; t += nt) clut[j * 768 + t] = augmix_point(cops[j], t & 255);
          __syncthreads();
        } else if (kind == IMGTX_AUG_EQUALIZE) {
          build_composed(cops, clut, j, ccomp, cfill);
          for (int t = tid; t < 768; t += nt) hist[t] = 0u;
          __syncthreads();
          // geometric records of the prefix [0, j): those with stage index < j, i.e. the tail of the trace-back list
          const int ng = g[0];
          int s0 = 0;
          while (s0 < ng && g[4 + s0 * 8 + 6] >= j) ++s0;
          if (s0 == ng) {
            // no geometric stage before it: the histogram of LUT(q) is the histogram of q pushed through the LUT
            for (int t = tid; t < 768; t += nt)
              if (base[t]) atomicAdd(&hist[(t & ~255) + ccomp[t]], base[t]);
          } else if (vec && zero_is_fixed(cops, clut, j)) {
            // count the gathered SOURCE bytes (outside pixels read the zero pixel), then push the counts through the LUT
            for (int t = tid; t < 768; t += nt) tmp[t] = 0u;
            __syncthreads();
            const int wq = w >> 2, dy = nt / wq, dx = nt - dy * wq;
            int y = tid / wq, xq = tid - y * wq;
            for (int p4 = tid; p4 < (hw >> 2); p4 += nt, y += dy, xq += dx) {
              if (xq >= wq) { xq -= wq; ++y; }
              const int x0 = xq * 4;
              int xs[4], ys[4], fk[4];
              bool ok[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) { xs[k] = x0 + k; ys[k] = y; fk[k] = 0; ok[k] = true; }
              trace4<false>(g, s0, ng, w, h, xs, ys, fk, ok);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint8_t* sp = q + (ok[k] ? (ys[k] * w + xs[k]) * 3 : zoff);
#pragma unroll
                for (int c = 0; c < 3; ++c) atomicAdd(&tmp[c * 256 + sp[c]], 1u);
              }
            }
            __syncthreads();
            for (int t = tid; t < 768; t += nt)
              if (tmp[t]) atomicAdd(&hist[(t & ~255) + ccomp[t]], tmp[t]);
          } else {
            for (int p = tid; p < hw; p += nt) {
              const int y = p / w, xx = p - y * w;
              int fk = 0;
              const int o = trace_back(cops, j, xx, y, w, h, fk);
#pragma unroll
              for (int c = 0; c < 3; ++c) {
                const int v = o >= 0 ? ccomp[c * 256 + q[o + c]] : cfill[fk * 4 + c];
                atomicAdd(&hist[c * 256 + v], 1u);
              }
            }
          }
          __syncthreads();
          if (warp < 3) equalize_lut_warp(hist + warp * 256, clut + j * 768 + warp * 256, lane);
          __syncthreads();
        }
      }
      build_composed(cops, clut, depth, ccomp, cfill);
    }
    __syncthreads();

    // ---- phase B': per-chain float tables (weight folded in: fmul(w_i, T[u]) is evaluated once per table entry
    //      instead of once per pixel) ----
    for (int t = tid; t < width * 768; t += nt) FT[t] = __fmul_rn(wts[t / 768], T[comp[t]]);
    for (int t = tid; t < nstage * 4; t += nt) FF[t] = __fmul_rn(wts[(t >> 2) / depth], T[fills[t]]);
    __syncthreads();

    // ---- phase C: mix and write ----
    const float m = wts[width], one_m = wts[width + 1];
    if (pre) head = fetch_head(img + gridDim.x);
    uint32_t zfix = 0;   // chains whose fill is the zero pixel
    for (int ci = 0; ci < width && ci < 32; ++ci)
      if (zero_is_fixed(ops + ci * depth, stage_lut + (size_t)ci * depth * 768, depth)) zfix |= 1u << ci;
    if (vec) {
      const int wq = w >> 2, dy = nt / wq, dx = nt - dy * wq;
      int y = tid / wq, xq = tid - y * wq;
      for (int p4 = tid; p4 < (hw >> 2); p4 += nt, y += dy, xq += dx) {
        if (xq >= wq) { xq -= wq; ++y; }
        const int p0 = p4 * 4, x0 = xq * 4;
        if (stash) {
          // the blend's second read of x starts now and lands in this thread's 48-byte slot while the chains are traced
#pragma unroll
          for (int c = 0; c < 3; ++c)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stash + (tid * 3 + c) * 4)),
                         "l"(xi + (int64_t)c * hw + p0) : "memory");
          asm volatile("cp.async.commit_group;" ::: "memory");
        }
        float mix[3][4];
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int k = 0; k < 4; ++k) mix[c][k] = 0.0f;
        for (int ci = 0; ci < width; ++ci) {
          const int32_t* g = G + ci * gstride;
          const float* ft = FT + ci * 768;
          // four adjacent pixels walk the chain's geometric stages together: one coefficient fetch per stage
          int xs[4], ys[4], fk[4];
          bool ok[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) { xs[k] = x0 + k; ys[k] = y; fk[k] = 0; ok[k] = true; }
          if (zfix & (1u << ci)) {
            trace4<false>(g, 0, g[0], w, h, xs, ys, fk, ok);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint8_t* sp = q + (ok[k] ? (ys[k] * w + xs[k]) * 3 : zoff);
#pragma unroll
              for (int c = 0; c < 3; ++c) mix[c][k] = __fadd_rn(mix[c][k], ft[c * 256 + sp[c]]);
            }
          } else {
            trace4<true>(g, 0, g[0], w, h, xs, ys, fk, ok);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (ok[k]) {
                const uint8_t* sp = q + (ys[k] * w + xs[k]) * 3;
#pragma unroll
                for (int c = 0; c < 3; ++c) mix[c][k] = __fadd_rn(mix[c][k], ft[c * 256 + sp[c]]);
              } else {
                const float* ff = FF + (ci * depth + fk[k]) * 4;
#pragma unroll
                for (int c = 0; c < 3; ++c) mix[c][k] = __fadd_rn(mix[c][k], ff[c]);
              }
            }
          }
        }
        if (stash) asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float4 f = stash ? *reinterpret_cast<const float4*>(stash + (tid * 3 + c) * 4)
                                 : ldg_hint_f4(xi + (int64_t)c * hw + p0, drop);
          float4 r;
          r.x = __fadd_rn(__fmul_rn(one_m, f.x), __fmul_rn(m, mix[c][0]));
          r.y = __fadd_rn(__fmul_rn(one_m, f.y), __fmul_rn(m, mix[c][1]));
          r.z = __fadd_rn(__fmul_rn(one_m, f.z), __fmul_rn(m, mix[c][2]));
          r.w = __fadd_rn(__fmul_rn(one_m, f.w), __fmul_rn(m, mix[c][3]));
          stg_hint_f4(oi + (int64_t)c * hw + p0, r, drop);
        }
      }
    } else {
      for (int p = tid; p < hw; p += nt) {
        const int y = p / w, xx = p - y * w;
        float mix[3] = {0.0f, 0.0f, 0.0f};
        for (int ci = 0; ci < width; ++ci) {
          const AugOp* cops = ops + ci * depth;
          int fk = 0;
          const int o = trace_back(cops, depth, xx, y, w, h, fk);
          for (int c = 0; c < 3; ++c) {
            const int u = o >= 0 ? comp[ci * 768 + c * 256 + q[o + c]] : fills[(ci * depth + fk) * 4 + c];
            mix[c] = __fadd_rn(mix[c], __fmul_rn(wts[ci], T[u]));
          }
        }
        for (int c = 0; c < 3; ++c)
          oi[(int64_t)c * hw + p] = __fadd_rn(__fmul_rn(one_m, xi[(int64_t)c * hw + p]), __fmul_rn(m, mix[c]));
      }
    }
  }
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_augmix_f32(const float* x, float* out, int64_t n, int h, int w, int width, int depth,
                                const imgtx_augmix_op_t* ops_dev, const float* weights_dev, imgtx_stream_t stream) {
  static_assert(sizeof(imgtx_augmix_op_t) == sizeof(AugOp) && sizeof(AugOp) == 32, "op record layout");
  IMGTX_CHECK_ARG(h > 0 && w > 0 && n >= 0, "bad image shape");
  IMGTX_CHECK_ARG(width >= 1 && width <= 16 && depth >= 1 && depth <= 16, "width/depth must be in [1,16]");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(x && out && ops_dev && weights_dev, "null pointer");
  const int hw = h * w, nstage = width * depth;
  IMGTX_CHECK_ARG(((int64_t)hw % 4 != 0) || (aligned16(x) && aligned16(out)), "f32 tensors must be 16-byte aligned");
  const int threads = hw >= 128 * 128 ? 1024 : 256;
  size_t smem = ((size_t)hw * 3 + 4 + 15) / 16 * 16 + (size_t)nstage * 768 + (size_t)width * 768 +
                ((size_t)nstage * 4 + 15) / 16 * 16 + 3 * 768 * 4 + 256 * 4 + (size_t)width * 768 * 4 + (size_t)nstage * 16 +
                (size_t)width * (4 + depth * 8) * 4 + (size_t)nstage * 32 + (size_t)(width + 2) * 4 + 16;
  IMGTX_CHECK_ARG(smem <= (size_t)max_smem_optin(), "image (%dx%d) does not fit in shared memory for fused AugMix", h, w);
  // 48 bytes per thread for the blend's second read of x (cp.async issued before the chains are traced) when they fit
  const size_t stash_bytes = (size_t)threads * 48 + 16;
  const int use_stash = (hw % 4 == 0 && w % 4 == 0 && smem + stash_bytes <= (size_t)max_smem_optin()) ? 1 : 0;
  if (use_stash) smem += stash_bytes;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(augmix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = (int)std::max<size_t>(1, std::min<size_t>(2048 / threads, (size_t)max_smem_optin() / (smem + 1024)));
  int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * per_sm);
  augmix_kernel<<<(unsigned)grid, threads, smem, (cudaStream_t)stream>>>(x, out, n, h, w, width, depth,
                                                                        reinterpret_cast<const AugOp*>(ops_dev), weights_dev, use_stash);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
