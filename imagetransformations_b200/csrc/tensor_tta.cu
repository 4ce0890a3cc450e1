// Tensor-backend test-time-augmentation chain and the mean over views that follows it.
//
//   reference: fall_2025/post_hoc_aaryaa.py:24-51 — TestTimeAugmentation = Compose([RandomHorizontalFlip(0.5),
//              RandomRotation(10), RandomAffine(0, scale=(1.0, 1.1)), ColorJitter(brightness=0.1),
//              RandomPerspective(0.2, p=0.5)]) applied to f32 CHW tensors, i.e. torchvision's tensor backend
//              (_functional_tensor.py: _gen_affine_grid / _perspective_grid + grid_sample(zeros, align_corners=False),
//              NEAREST for rotate / affine, BILINEAR with a ones-mask channel for perspective, _blend for brightness);
//              :317-342 (mean of T+1 logit vectors), fall_2025/TTA_transforms.py:208-209 (mean of view probabilities).
//
// One CTA per OUTPUT image: the source image is read once into shared memory (flip folded into the read), the two
// NEAREST resamples ping-pong between two shared buffers, brightness rides on the second gather, and the perspective
// stage (or a plain copy) writes the result.  The affine grids reproduce torch.bmm's f32 evaluation order
// (x*a rounded, then fma(y, b, .), then + c), so the NEAREST picks are the ones the CPU reference makes.
// Roofline: HBM, c*h*w*4 bytes in (L2 hits when several views share an image) + c*h*w*4 bytes out per view.
#include "common.cuh"
#include <algorithm>

namespace imgtx {

struct TtaChain {  // == imgtx_tta_chain_t
  int32_t flip, perspective;
  float brightness, reserved;
  float rot[6], aff[6], persp[8];
};

// _gen_affine_grid + grid_sample(nearest) unnormalise for output pixel (x, y): source column / row (rounded half-even)
// r[] = theta rescaled: th[0..2] / (w/2), th[3..5] / (h/2) (rescaled_theta of _gen_affine_grid), once per image
__device__ __forceinline__ void rescale_theta(const float* th, int w, int h, float* r) {
  const float hw2 = __fmul_rn(0.5f, (float)w), hh2 = __fmul_rn(0.5f, (float)h);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    r[k] = __fdiv_rn(th[k], hw2);
    r[3 + k] = __fdiv_rn(th[3 + k], hh2);
  }
}
__device__ __forceinline__ void affine_pick(const float* r, int x, int y, int w, int h, float& rx, float& ry) {
  // linspace(-w/2 + .5, w/2 - .5, w)[x]: exact for these sizes (half-integers)
  const float X = __fadd_rn(__fadd_rn(__fmul_rn(-(float)w, 0.5f), 0.5f), (float)x);
  const float Y = __fadd_rn(__fadd_rn(__fmul_rn(-(float)h, 0.5f), 0.5f), (float)y);
  const float gx = __fadd_rn(__fmaf_rn(Y, r[1], __fmul_rn(X, r[0])), r[2]);
  const float gy = __fadd_rn(__fmaf_rn(Y, r[4], __fmul_rn(X, r[3])), r[5]);
  const float ix = __fmaf_rn(__fadd_rn(gx, 1.0f), __fmul_rn(0.5f, (float)w), -0.5f);   // ATen: fma(g + 1, size/2, -0.5)
  const float iy = __fmaf_rn(__fadd_rn(gy, 1.0f), __fmul_rn(0.5f, (float)h), -0.5f);
  rx = rintf(ix);
  ry = rintf(iy);
}

// CT: compile-time channel count (3 = RGB: the channel loops unroll), 0 = run-time c_rt
template <int CT>
__global__ void __launch_bounds__(256)
tensor_tta_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t n_out, int views, int c_rt, int h, int w,
                  const TtaChain* __restrict__ chains) {
  const int c = CT ? CT : c_rt;
  extern __shared__ __align__(16) float sm[];
  const int hw = h * w, chw = c * hw;
  float* A = sm;
  float* B = sm + chw;
  __shared__ TtaChain ch;
  const int tid = threadIdx.x;
  // pixel p = tid + 256*k walks (yy, xx) without a division: step (dy, dx) with one carry
  const int ty0 = tid / w, tx0 = tid - ty0 * w, dy = 256 / w, dx = 256 - dy * w;
  for (int64_t o = blockIdx.x; o < n_out; o += gridDim.x) {
    __syncthreads();
    if (tid < (int)(sizeof(TtaChain) / 4)) reinterpret_cast<int32_t*>(&ch)[tid] = reinterpret_cast<const int32_t*>(chains + o)[tid];
    __syncthreads();
    const float* xi = x + (o / views) * (int64_t)chw;
    float* oi = out + o * (int64_t)chw;
    // stage 0: load (RandomHorizontalFlip folded in)
    if (ch.flip) {
      for (int i = tid; i < chw; i += 256) {
        const int row = i / w, xx = i - row * w;
        A[i] = xi[row * w + (w - 1 - xx)];
      }
    } else {
      for (int i = tid; i < chw; i += 256) A[i] = xi[i];
    }
    __syncthreads();
    // stage 1: RandomRotation -> F.rotate(NEAREST): A -> B
    float r1[6], r2[6];
    rescale_theta(ch.rot, w, h, r1);
    rescale_theta(ch.aff, w, h, r2);
    for (int p = tid, yy = ty0, xx = tx0; p < hw; p += 256, yy += dy, xx += dx) {
      if (xx >= w) { xx -= w; ++yy; }
      float rx, ry;
      affine_pick(r1, xx, yy, w, h, rx, ry);
      const bool ok = rx >= 0.0f && rx < (float)w && ry >= 0.0f && ry < (float)h;
      const int s = ok ? (int)ry * w + (int)rx : 0;
#pragma unroll
      for (int k = 0; k < c; ++k) B[k * hw + p] = ok ? A[k * hw + s] : 0.0f;
    }
    __syncthreads();
    // stage 2: RandomAffine(scale) -> F.affine(NEAREST), then ColorJitter brightness: (b*v + 0).clamp(0,1): B -> A / out
    const bool persp = ch.perspective != 0;
    const float bright = ch.brightness;
    for (int p = tid, yy = ty0, xx = tx0; p < hw; p += 256, yy += dy, xx += dx) {
      if (xx >= w) { xx -= w; ++yy; }
      float rx, ry;
      affine_pick(r2, xx, yy, w, h, rx, ry);
      const bool ok = rx >= 0.0f && rx < (float)w && ry >= 0.0f && ry < (float)h;
      const int s = ok ? (int)ry * w + (int)rx : 0;
#pragma unroll
      for (int k = 0; k < c; ++k) {
        float v = ok ? B[k * hw + s] : 0.0f;
        v = __fadd_rn(__fmul_rn(bright, v), 0.0f);
        v = v != v ? v : fminf(fmaxf(v, 0.0f), 1.0f);
        if (persp) A[k * hw + p] = v;
        else oi[k * hw + p] = v;
      }
    }
    if (!persp) continue;
    __syncthreads();
    // stage 3: RandomPerspective -> F.perspective(BILINEAR, fill 0): sampled image times sampled ones-mask
    const float hw2 = __fmul_rn(0.5f, (float)w), hh2 = __fmul_rn(0.5f, (float)h);
    const float* co = ch.persp;
    const float t0 = __fdiv_rn(co[0], hw2), t1 = __fdiv_rn(co[1], hw2), t2 = __fdiv_rn(co[2], hw2);
    const float t3 = __fdiv_rn(co[3], hh2), t4 = __fdiv_rn(co[4], hh2), t5 = __fdiv_rn(co[5], hh2);
    for (int p = tid, yy = ty0, xx = tx0; p < hw; p += 256, yy += dy, xx += dx) {
      if (xx >= w) { xx -= w; ++yy; }
      const float X = (float)xx + 0.5f, Y = (float)yy + 0.5f;
      const float g1x = __fadd_rn(__fmaf_rn(Y, t1, __fmul_rn(X, t0)), t2);   // torch.bmm's order on the CPU reference
      const float g1y = __fadd_rn(__fmaf_rn(Y, t4, __fmul_rn(X, t3)), t5);
      const float g2 = __fadd_rn(__fmaf_rn(Y, co[7], __fmul_rn(X, co[6])), 1.0f);
      const float gx = __fsub_rn(__fdiv_rn(g1x, g2), 1.0f), gy = __fsub_rn(__fdiv_rn(g1y, g2), 1.0f);
      const float ix = __fmaf_rn(__fadd_rn(gx, 1.0f), __fmul_rn(0.5f, (float)w), -0.5f);   // ATen: fma(g + 1, size/2, -0.5)
      const float iy = __fmaf_rn(__fadd_rn(gy, 1.0f), __fmul_rn(0.5f, (float)h), -0.5f);
      const float fx0 = floorf(ix), fy0 = floorf(iy);
      const float wx1 = __fsub_rn(ix, fx0), wy1 = __fsub_rn(iy, fy0);
      const float wx0 = __fsub_rn(1.0f, wx1), wy0 = __fsub_rn(1.0f, wy1);
      const int x0 = (fx0 < -2.0f || fx0 > (float)w + 1.0f) ? -4 : (int)fx0;   // also catches NaN-free wild coordinates
      const int y0 = (fy0 < -2.0f || fy0 > (float)h + 1.0f) ? -4 : (int)fy0;
      const float w00 = __fmul_rn(wx0, wy0), w01 = __fmul_rn(wx1, wy0), w10 = __fmul_rn(wx0, wy1), w11 = __fmul_rn(wx1, wy1);
      const bool okx0 = x0 >= 0 && x0 < w, okx1 = x0 + 1 >= 0 && x0 + 1 < w;
      const bool oky0 = y0 >= 0 && y0 < h, oky1 = y0 + 1 >= 0 && y0 + 1 < h;
      const float m00 = (okx0 && oky0) ? 1.f : 0.f, m01 = (okx1 && oky0) ? 1.f : 0.f;
      const float m10 = (okx0 && oky1) ? 1.f : 0.f, m11 = (okx1 && oky1) ? 1.f : 0.f;
      const float mask = __fmaf_rn(m11, w11, __fmaf_rn(m10, w10, __fmaf_rn(m01, w01, __fmul_rn(m00, w00))));
      const int cx0 = clampi(x0, 0, w - 1), cx1 = clampi(x0 + 1, 0, w - 1), cy0 = clampi(y0, 0, h - 1), cy1 = clampi(y0 + 1, 0, h - 1);
#pragma unroll
      for (int k = 0; k < c; ++k) {
        const float* a = A + k * hw;
        const float v00 = (okx0 && oky0) ? a[cy0 * w + cx0] : 0.f, v01 = (okx1 && oky0) ? a[cy0 * w + cx1] : 0.f;
        const float v10 = (okx0 && oky1) ? a[cy1 * w + cx0] : 0.f, v11 = (okx1 && oky1) ? a[cy1 * w + cx1] : 0.f;
        float v = __fmaf_rn(v11, w11, __fmaf_rn(v10, w10, __fmaf_rn(v01, w01, __fmul_rn(v00, w00))));
        oi[k * hw + p] = __fadd_rn(__fmul_rn(v, mask), __fmul_rn(__fsub_rn(1.0f, mask), 0.0f));
      }
    }
  }
}

// out[i] = (((0 + in[0][i]) + in[1][i]) + ...) / V — the accumulation order of post_hoc_aaryaa.py:334-342
__global__ void __launch_bounds__(256)
mean_views_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t count, int views) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < count; i += (int64_t)gridDim.x * 256) {
    float s = 0.0f;
    for (int v = 0; v < views; ++v) s = __fadd_rn(s, in[(int64_t)v * count + i]);
    out[i] = __fdiv_rn(s, (float)views);
  }
}

}  // namespace imgtx

using namespace imgtx;

extern "C" int imgtx_tensor_tta_f32(const float* x, float* out, int64_t n, int views, int c, int h, int w,
                                    const imgtx_tta_chain_t* chains_dev, imgtx_stream_t stream) {
  static_assert(sizeof(imgtx_tta_chain_t) == sizeof(TtaChain) && sizeof(TtaChain) == 96, "chain record layout");
  IMGTX_CHECK_ARG(n >= 0 && views >= 1 && c >= 1 && h > 0 && w > 0, "bad tensor shape");
  if (n == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(x && out && chains_dev, "null pointer");
  const size_t smem = (size_t)2 * c * h * w * sizeof(float);
  IMGTX_CHECK_ARG(smem + 1024 <= (size_t)max_smem_optin(), "image (%dx%dx%d) does not fit in shared memory for the fused TTA chain", c, h, w);
  auto kern = c == 3 ? tensor_tta_kernel<3> : tensor_tta_kernel<0>;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)max_smem_optin() / (smem + 1024)));
  const int64_t n_out = n * views;
  const int64_t grid = std::min<int64_t>(n_out, (int64_t)sm_count() * per_sm);
  kern<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(x, out, n_out, views, c, h, w,
                                                          reinterpret_cast<const TtaChain*>(chains_dev));
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

extern "C" int imgtx_mean_views_f32(const float* in, float* out, int64_t count, int views, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(count >= 0 && views >= 1, "bad shape");
  if (count == 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in && out, "null pointer");
  const int64_t grid = std::min<int64_t>((count + 255) / 256, (int64_t)sm_count() * 8);
  mean_views_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(in, out, count, views);
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
