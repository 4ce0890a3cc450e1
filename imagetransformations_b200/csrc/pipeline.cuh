// Image-group pipeline: the one execution skeleton shared by every u8->u8 image operator.
//
// An operator ("Op") is a functor that transforms a *group* of whole images with all threads of a CTA:
//
//   struct Op {
//     struct Params { ... };                       // POD, passed by value
//     static constexpr int kThreads = 256;
//     static size_t scratch_bytes(const Shape&, int G);   // extra shared memory per CTA
//     static __device__ void prepare(const Shape&, const Params&, uint8_t* scratch, bool staged);
//                                                          // once per CTA (e.g. batch-uniform tables)
//     template <bool STAGED>
//     static __device__ void run(const uint8_t* src, uint8_t* dst, int g, int64_t img0,
//                                const Shape& s, const Params& p, uint8_t* scratch, uint8_t* gtmp);
//   };
//
// Two launch modes, chosen on the host:
//  * STAGED  (small images, e.g. CIFAR 32x32x3 = 3072 B): persistent CTAs; groups of G images are
//    moved HBM -> shared memory with 1-D bulk TMA copies (cp.async.bulk + mbarrier, a `stages`-deep
//    ring), transformed shared -> shared, and written back with bulk TMA stores.  HBM sees only
//    full-line sequential traffic no matter how irregular the operator's gather pattern is.
//  * DIRECT  (large or oddly sized images): one CTA per image, src/dst point at global memory and the
//    gather goes through L1/L2.
#pragma once
#include "common.cuh"
#include <stdlib.h>
#include <algorithm>

namespace imgtx {

struct Shape {
  int h, w, c;        // input height, width, channels (HWC interleaved u8)
  int oh, ow;         // output height, width (same c)
  int in_bytes;       // h*w*c
  int out_bytes;      // oh*ow*c
  int aux;            // operator-specific size hint for scratch_bytes (e.g. line-kernel width); 0 otherwise
};

static inline Shape make_shape(int h, int w, int c, int oh, int ow) {
  Shape s;
  s.h = h; s.w = w; s.c = c; s.oh = oh; s.ow = ow;
  s.in_bytes = h * w * c;
  s.out_bytes = oh * ow * c;
  s.aux = 0;
  return s;
}

#ifdef __CUDACC__

constexpr int kBarBytes = 128;  // mbarrier block at the front of dynamic smem
constexpr int kMaxStages = 8;

__host__ __device__ static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Ops may declare `static constexpr int kMinBlocks` (occupancy target handed to ptxas via __launch_bounds__).
template <class Op, class = void>
struct MinBlocks { static constexpr int value = 1; };
template <class Op>
struct MinBlocks<Op, decltype((void)Op::kMinBlocks)> { static constexpr int value = Op::kMinBlocks; };

// Ops may declare `static constexpr bool kCifarCopy = true` to get a second copy of run() with compile-time 32x32x3
// dimensions (measured per operator: it helps index-heavy kernels and hurts the large switch / sweep kernels).
template <class Op, class = void>
struct CifarCopy { static constexpr bool value = false; };
template <class Op>
struct CifarCopy<Op, decltype((void)Op::kCifarCopy)> { static constexpr bool value = Op::kCifarCopy; };

// Compute-heavy Ops may declare `static constexpr int kStages` / `kOutBufs` (input ring depth / output buffers, default:
// tuned depth and 2): their launches last several times the DMA time, so a shallow ring costs nothing and the shared
// memory it frees buys occupancy.
template <class Op, class = void>
struct RingDepth { static constexpr int stages = 0, out_bufs = 2; };
template <class Op>
struct RingDepth<Op, decltype((void)Op::kStages)> { static constexpr int stages = Op::kStages, out_bufs = Op::kOutBufs; };

// Ops may declare `static constexpr int kCtasPerSm = N` to replace the tuned CTAs-per-SM cap (grid size) for their launches.
template <class Op, class = void>
struct CtasPerSm { static constexpr int value = 0; };
template <class Op>
struct CtasPerSm<Op, decltype((void)Op::kCtasPerSm)> { static constexpr int value = Op::kCtasPerSm; };

template <class Op>
__global__ void __launch_bounds__(Op::kThreads, MinBlocks<Op>::value)
staged_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, Shape s,
              typename Op::Params p, int G, int stages) {
  constexpr int kOut = RingDepth<Op>::out_bufs;
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  const size_t in_stage = align_up((size_t)G * s.in_bytes, 128);
  const size_t out_stage = align_up((size_t)G * s.out_bytes, 128);
  uint8_t* in_buf = smem + kBarBytes;
  uint8_t* out_buf = in_buf + in_stage * stages;
  uint8_t* scratch = out_buf + out_stage * kOut;

  const int64_t n_groups = (n + G - 1) / G;
  const int tid = threadIdx.x;
  const bool cifar = CifarCopy<Op>::value && s.h == 32 && s.w == 32 && s.c == 3 && s.oh == 32 && s.ow == 32;

  if (tid == 0) {
    for (int i = 0; i < stages; ++i) mbar_init(&full[i], 1);
    fence_mbar_init();
  }
  __syncthreads();

  Op::prepare(s, p, scratch, true);

  // prologue: fill the ring
  if (tid == 0) {
    for (int i = 0; i < stages; ++i) {
      int64_t grp = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      if (grp < n_groups) {
        int64_t img0 = grp * G;
        int g = (int)min((int64_t)G, n - img0);
        uint32_t bytes = (uint32_t)g * s.in_bytes;
        mbar_expect_tx(&full[i], bytes);
        bulk_g2s(in_buf + in_stage * i, in + img0 * s.in_bytes, bytes, &full[i]);
      }
    }
  }

  int it = 0;
  for (int64_t grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++it) {
    const int st = it % stages;
    const uint32_t parity = (uint32_t)(it / stages) & 1u;
    const int ob = kOut == 2 ? (it & 1) : 0;
    const int64_t img0 = grp * G;
    const int g = (int)min((int64_t)G, n - img0);

    // the bulk store issued two iterations ago must have finished reading out_buf[ob]
    if (tid == 0) { if (kOut == 2) bulk_wait_read<1>(); else bulk_wait_read<0>(); }
    mbar_wait(&full[st], parity);
    __syncthreads();

    // CIFAR geometry (the headline workload) gets its own copy of the operator with compile-time dimensions, so index
    // arithmetic (px / w, i / (h*w), ...) strength-reduces to shifts instead of runtime integer division
    if (CifarCopy<Op>::value && cifar) {
      const Shape k32 = {32, 32, 3, 32, 32, 3072, 3072, s.aux};
      Op::template run<true>(in_buf + in_stage * st, out_buf + out_stage * ob, g, img0, k32, p, scratch, nullptr);
    } else {
      Op::template run<true>(in_buf + in_stage * st, out_buf + out_stage * ob, g, img0, s, p, scratch, nullptr);
    }

    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(out + img0 * s.out_bytes, out_buf + out_stage * ob, (uint32_t)g * s.out_bytes);
      bulk_commit();
      int64_t nxt = grp + (int64_t)stages * gridDim.x;
      if (nxt < n_groups) {
        int64_t nimg0 = nxt * G;
        int ng = (int)min((int64_t)G, n - nimg0);
        uint32_t bytes = (uint32_t)ng * s.in_bytes;
        mbar_expect_tx(&full[st], bytes);
        bulk_g2s(in_buf + in_stage * st, in + nimg0 * s.in_bytes, bytes, &full[st]);
      }
    }
  }
  if (tid == 0) bulk_wait_all<0>();
}

template <class Op>
__global__ void __launch_bounds__(Op::kThreads)
direct_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, Shape s,
              typename Op::Params p, uint8_t* gtmp, size_t gtmp_per_cta) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* my_tmp = gtmp ? gtmp + (size_t)blockIdx.x * gtmp_per_cta : nullptr;
  Op::prepare(s, p, smem, false);
  __syncthreads();
  for (int64_t img = blockIdx.x; img < n; img += gridDim.x) {
    Op::template run<false>(in + img * s.in_bytes, out + img * s.out_bytes, 1, img, s, p, smem, my_tmp);
    __syncthreads();
  }
}

// Tunables (overridable from the environment for the on-GPU sweeps in bench.py).
struct StagedTuning {
  int target_smem;  // dynamic smem per CTA we aim for (bytes)
  int stages;
  int ctas_per_sm;
  int ctas_from_env;   // IMGTX_CTAS_PER_SM was given: it overrides the operators' own caps
  int target_scale;    // IMGTX_TARGET_SCALE (percent of an operator's own target), 0 = off
  int glass_g;         // IMGTX_GLASS_G
  int pw_ctas;         // IMGTX_PW_CTAS
};
StagedTuning staged_tuning();

// extra_scratch: parameter-dependent shared memory on top of Op::scratch_bytes (tables sized at run time);
// target_override > 0 replaces the tuned per-CTA shared-memory target (operators with large fixed tables).
template <class Op>
int launch_image_op(const uint8_t* in, uint8_t* out, int64_t n, const Shape& s,
                    const typename Op::Params& p, cudaStream_t stream, size_t extra_scratch = 0,
                    int target_override = 0) {
  if (n <= 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in != nullptr && out != nullptr, "null image pointer");
  StagedTuning tn = staged_tuning();
  if (target_override > 0) {
    tn.target_smem = target_override;
    // sweep knob: scale an operator's own target (percent), e.g. IMGTX_TARGET_SCALE=150
    if (tn.target_scale > 0) tn.target_smem = (int)((int64_t)target_override * tn.target_scale / 100);
  }
  if (CtasPerSm<Op>::value > 0 && !tn.ctas_from_env) tn.ctas_per_sm = CtasPerSm<Op>::value;
  const int max_smem = max_smem_optin();
  if (tn.target_smem > max_smem - 1024) tn.target_smem = max_smem - 1024;   // a target can ask for more than the SM has
  bool staged = (s.in_bytes % 16 == 0) && (s.out_bytes % 16 == 0) && aligned16(in) && aligned16(out);
  int G = 1, stages = RingDepth<Op>::stages > 0 ? RingDepth<Op>::stages : tn.stages;
  constexpr int kOut = RingDepth<Op>::out_bufs;
  size_t smem_bytes = 0;
  if (staged) {
    auto need = [&](int g, int st) {
      return (size_t)kBarBytes + align_up((size_t)g * s.in_bytes, 128) * st +
             align_up((size_t)g * s.out_bytes, 128) * kOut + Op::scratch_bytes(s, g) + extra_scratch + 128;
    };
    while (stages > 2 && need(1, stages) > (size_t)tn.target_smem) --stages;
    if (need(1, stages) > (size_t)max_smem - 1024) {
      staged = false;
    } else {
      G = 1;
      while (G < 64 && need(G + 1, stages) <= (size_t)tn.target_smem) ++G;
      smem_bytes = need(G, stages);
    }
  }
  if (staged) {
    auto kern = staged_kernel<Op>;
    IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int64_t n_groups = (n + G - 1) / G;
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(tn.ctas_per_sm, (size_t)max_smem / smem_bytes));
    int64_t grid = std::min<int64_t>(n_groups, (int64_t)sm_count() * per_sm);
    kern<<<(unsigned)grid, Op::kThreads, smem_bytes, stream>>>(in, out, n, s, p, G, stages);
    IMGTX_LAUNCH_OK();
    return IMGTX_OK;
  }
  // direct mode
  size_t scratch = Op::scratch_bytes_direct(s) + extra_scratch;
  uint8_t* gtmp = nullptr;
  size_t gtmp_per_cta = 0;
  int64_t grid = std::min<int64_t>(n, (int64_t)sm_count() * 8);
  if (scratch > (size_t)max_smem - 1024) {
    // intermediate planes do not fit on chip: spill them to a stream-ordered global workspace.  Operators without a spilled
    // form (NoSpill) cannot run on images this large: refuse instead of launching with shared memory they would overrun.
    if (Op::gtmp_bytes(s) == 0) {
      set_error("image %dx%dx%d needs %zu bytes of shared memory per CTA (limit %d): this operator has no large-image form",
                s.h, s.w, s.c, scratch, max_smem - 1024);
      return IMGTX_ERR_INVALID_ARGUMENT;
    }
    gtmp_per_cta = align_up(Op::gtmp_bytes(s), 256);
    scratch = Op::scratch_bytes_spilled(s) + extra_scratch;
    IMGTX_CUDA_OK(workspace_alloc((void**)&gtmp, gtmp_per_cta * grid, stream));
  }
  auto kern = direct_kernel<Op>;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scratch));
  kern<<<(unsigned)grid, Op::kThreads, scratch, stream>>>(in, out, n, s, p, gtmp, gtmp_per_cta);
  IMGTX_LAUNCH_OK();
  if (gtmp) IMGTX_CUDA_OK(cudaFreeAsync(gtmp, stream));
  return IMGTX_OK;
}

// Default "no spill" hooks for operators without intermediate planes.
// Per-image metric of an output batch (runtime.cu): metrics = device (n, 2) uint32 {byte sum, position-weighted sum}.
int image_checksums(const uint8_t* x, int64_t n, int64_t image_bytes, uint32_t* metrics, cudaStream_t stream);

struct NoSpill {
  static size_t gtmp_bytes(const Shape&) { return 0; }
  static size_t scratch_bytes_spilled(const Shape&) { return 0; }
};
struct NoPrepare {
  template <class S, class P>
  __device__ static void prepare(const S&, const P&, uint8_t*, bool) {}
};

// ---- shared-memory byte helpers used by the gather-style operators -----------------------------
// 3 bytes at byte offset `off` from `base` (reads the two enclosing aligned words; staged buffers
// are padded so the second word is always readable).
__device__ __forceinline__ uint32_t load_rgb(const uint8_t* base, int off) {
  // offsets, not pointers: integer arithmetic on a generic pointer makes the compiler forget that `base` is shared memory
  // (LD.E + 64-bit address maths instead of LDS)
  const int mis = (int)(reinterpret_cast<uintptr_t>(base) & 3);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(base - mis) + ((off + mis) >> 2);
  // the second word is read only when the three bytes straddle it (direct mode reads caller memory: never past its end)
  const uint32_t hi = ((off + mis) & 3) > 1 ? w[1] : 0u;
  return __funnelshift_r(w[0], hi, (uint32_t)(off + mis) << 3) & 0x00FFFFFFu;
}
// pack four 24-bit pixels into three words (little endian byte order r0 g0 b0 r1 ...)
__device__ __forceinline__ void pack_rgb4(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3, uint32_t& w0,
                                          uint32_t& w1, uint32_t& w2) {
  w0 = p0 | (p1 << 24);
  w1 = (p1 >> 8) | (p2 << 16);
  w2 = (p2 >> 16) | (p3 << 8);
}

#endif  // __CUDACC__
}  // namespace imgtx
