// Team pipeline: the execution skeleton of the throughput ("fast") forms of the CIFAR-10-C-style suite.
//
// The staged pipeline of pipeline.cuh moves GROUPS of images per CTA and separates load / compute / store with CTA-wide
// barriers, which leaves most warps waiting when the per-image work is short (ncu: 2-7 barrier stalls per issue, 18-50 % of
// the warps active).  Here the unit is a TEAM of kTeamWarps warps (1, 2 or 4) that owns ONE image at a time from start to
// finish:
//   * every team has its own input ring (kInBufs whole images) fed by 1-D bulk TMA copies completing on the team's own
//     mbarriers, its own output buffers drained by bulk TMA stores, and its own scratch;
//   * the team leader (one lane) issues the copies; the only synchronisation inside the loop is the team's own barrier
//     (`__syncwarp` for one warp, a named `bar.sync id, 32*kTeamWarps` otherwise) — no `__syncthreads` after start-up;
//   * teams of one CTA share read-only tables (built once in `prepare`), nothing else, so a stalled team never holds up
//     another one and the SM always has many independent instruction streams.
// kOutBufs == 0 declares an IN-PLACE operator (the result overwrites the staged input and is stored from there).
//
//   struct Op {
//     using Params = ...;                                   // POD, by value
//     static constexpr int kTeamWarps = 1, kInBufs = 3, kOutBufs = 0, kMaxTeams = 16, kCtasPerSm = 2;
//     static size_t team_scratch(const Shape&, const Params&);   // bytes per team
//     static size_t cta_scratch(const Shape&, const Params&);    // bytes per CTA (shared tables)
//     static __device__ void prepare(const Shape&, const Params&, uint8_t* cta);          // all threads of the CTA, once
//     static __device__ void run(const uint8_t* in, uint8_t* out, int64_t img, const Shape&, const Params&,
//                                uint8_t* team_scratch, const uint8_t* cta, int tt /* thread in team */, int team);
//   };
#pragma once
#include "pipeline.cuh"

namespace imgtx {
#ifdef __CUDACC__

template <int TW>
__device__ __forceinline__ void team_sync(int team) {
  if (TW == 1) {
    __syncwarp();
  } else {
    asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(TW * 32) : "memory");
  }
}

template <class Op, class = void>
struct TeamMinBlocks { static constexpr int value = 1; };
template <class Op>
struct TeamMinBlocks<Op, decltype((void)Op::kMinBlocks)> { static constexpr int value = Op::kMinBlocks; };

// Per-image metric of a pass: {sum of the output bytes, sum of (index + 1) * byte mod 2^32} — order-sensitive inside an image,
// so a sharded run and a single-GPU run can be compared without moving the images.  Partial sums of 4-byte
// words: (4 i + 1) * (b0 + b1 + b2 + b3) + (b1 + 2 b2 + 3 b3), two IDP.4A per word.
__device__ __forceinline__ void checksum_word(uint32_t w, uint32_t i, uint32_t& c0, uint32_t& c1) {
  const uint32_t s = __dp4a(w, 0x01010101u, 0u);
  c0 += s;
  c1 += (4u * i + 1u) * s + __dp4a(w, 0x03020100u, 0u);
}

template <class Op>
__global__ void __launch_bounds__(Op::kTeamWarps * 32 * Op::kMaxTeams, TeamMinBlocks<Op>::value)
team_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, Shape s, typename Op::Params p,
            int cta_scratch_bytes, int team_bytes, uint2* __restrict__ metrics) {
  constexpr int TW = Op::kTeamWarps, TT = TW * 32, KI = Op::kInBufs, KO = Op::kOutBufs;
  static_assert(KI >= 2 && KI <= 8 && KO != 1 && TW <= 8, "ring depths / header layout (8 mbarriers + 8 metric partials)");
  extern __shared__ __align__(128) uint8_t smem[];
  const int team = threadIdx.x / TT, tt = threadIdx.x - team * TT;
  const int teams_per_cta = blockDim.x / TT;
  uint8_t* base = smem + cta_scratch_bytes + (size_t)team * team_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(base);
  const int ib = (int)align_up((size_t)s.in_bytes, 128), ob = (int)align_up((size_t)s.out_bytes, 128);
  uint8_t* in_buf = base + 128;
  uint8_t* out_buf = in_buf + KI * ib;
  uint8_t* scr = out_buf + KO * ob;

  if (tt == 0) {
    for (int i = 0; i < KI; ++i) mbar_init(&full[i], 1);
    fence_mbar_init();
  }
  Op::prepare(s, p, smem);
  __syncthreads();   // the only CTA-wide barrier

  const int64_t stride = (int64_t)gridDim.x * teams_per_cta;
  const int64_t first = (int64_t)blockIdx.x * teams_per_cta + team;
  if (tt == 0) {
    for (int i = 0; i < KI; ++i) {
      const int64_t img = first + (int64_t)i * stride;
      if (img < n) {
        mbar_expect_tx(&full[i], (uint32_t)s.in_bytes);
        bulk_g2s(in_buf + i * ib, in + img * s.in_bytes, (uint32_t)s.in_bytes, &full[i]);
      }
    }
  }
  int it = 0;
  for (int64_t img = first; img < n; img += stride, ++it) {
    const int st = it % KI;
    mbar_wait(&full[st], (uint32_t)(it / KI) & 1u);
    uint8_t* src = in_buf + st * ib;
    uint8_t* dst = KO == 0 ? src : out_buf + (it % (KO > 0 ? KO : 1)) * ob;
    Op::run(src, dst, img, s, p, scr, smem, tt, team);
    if (metrics) {
      // epilogue: the pass's per-image metric straight from the staged output (no second trip to HBM); each warp leaves its
      // partial sums in the team header, the leader adds them after the barrier below
      if (TW > 1) team_sync<TW>(team);   // every warp's part of dst is written
      else __syncwarp();
      const uint32_t* d32 = reinterpret_cast<const uint32_t*>(dst);
      uint32_t c0 = 0, c1 = 0;
      if (768 % TT == 0 && s.out_bytes == 3072) {
        // CIFAR images: thread tt owns words [tt WPL, (tt + 1) WPL), so the weight of a byte inside the thread's run is
        // 4 k + t + 1 <= 96 and fits a dp4a weight: two IDP.4A per word, vector loads
        constexpr int WPL = 768 % TT == 0 ? 768 / TT : 4;
        constexpr int V = WPL % 4 == 0 ? 4 : (WPL % 2 == 0 ? 2 : 1);
        const uint32_t* base = d32 + tt * WPL;
        uint32_t Cw = 0;
#pragma unroll
        for (int q = 0; q < WPL / V; ++q) {
          uint32_t w[4];
          if (V == 4) { const uint4 v = reinterpret_cast<const uint4*>(base)[q]; w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w; }
          else if (V == 2) { const uint2 v = reinterpret_cast<const uint2*>(base)[q]; w[0] = v.x; w[1] = v.y; }
          else w[0] = base[q];
#pragma unroll
          for (int e = 0; e < V; ++e) {
            c0 = __dp4a(w[e], 0x01010101u, c0);
            Cw = __dp4a(w[e], 0x04030201u + (uint32_t)(q * V + e) * 0x04040404u, Cw);
          }
        }
        c1 = Cw + 4u * (uint32_t)(tt * WPL) * c0;
      } else {
        for (int i = tt; i < (s.out_bytes >> 2); i += TT) checksum_word(d32[i], (uint32_t)i, c0, c1);
      }
      c0 = __reduce_add_sync(0xffffffffu, c0);   // REDUX: one instruction per sum instead of a five-level shuffle tree
      c1 = __reduce_add_sync(0xffffffffu, c1);
      if ((tt & 31) == 0) reinterpret_cast<uint2*>(base + 64)[tt >> 5] = make_uint2(c0, c1);
    }
    fence_async_smem();
    // the output buffer of the NEXT iteration must be free before anybody passes the barrier below: the store issued
    // KO - 1 iterations ago has to have finished reading it
    if (KO >= 2 && tt == 0) bulk_wait_read<(KO >= 2 ? KO - 2 : 0)>();
    team_sync<TW>(team);
    if (tt == 0) {
      if (metrics) {
        uint2 m = make_uint2(0u, 0u);
#pragma unroll
        for (int w = 0; w < TW; ++w) {
          const uint2 q = reinterpret_cast<const uint2*>(base + 64)[w];
          m.x += q.x; m.y += q.y;
        }
        metrics[img] = m;
      }
      bulk_s2g(out + img * s.out_bytes, dst, (uint32_t)s.out_bytes);
      bulk_commit();
      if (KO == 0) {
        // in place: the buffer just handed to the store is reloaded one iteration later, when the store has read it
        if (it >= 1) {
          bulk_wait_read<1>();
          const int64_t nxt = img - stride + (int64_t)KI * stride;
          if (nxt < n) {
            const int pb = (it - 1) % KI;
            mbar_expect_tx(&full[pb], (uint32_t)s.in_bytes);
            bulk_g2s(in_buf + pb * ib, in + nxt * s.in_bytes, (uint32_t)s.in_bytes, &full[pb]);
          }
        }
      } else {
        const int64_t nxt = img + (int64_t)KI * stride;
        if (nxt < n) {
          mbar_expect_tx(&full[st], (uint32_t)s.in_bytes);
          bulk_g2s(in_buf + st * ib, in + nxt * s.in_bytes, (uint32_t)s.in_bytes, &full[st]);
        }
      }
    }
  }
  if (tt == 0) bulk_wait_all<0>();
}

// Shared memory of one SM that CTAs can use (B200: 228 KB, 1 KB reserved per resident CTA).
static inline int smem_per_sm() { return 228 * 1024; }

// teams_override > 0 caps the teams per CTA (tuning).  Returns IMGTX_ERR_UNSUPPORTED (without an error string) when the
// shape cannot be staged, so callers fall through to the general pipeline.  metrics: device (n, 2) uint32 or nullptr.
template <class Op>
int launch_team_op(const uint8_t* in, uint8_t* out, int64_t n, const Shape& s, const typename Op::Params& p,
                   cudaStream_t stream, uint32_t* metrics = nullptr) {
  if (n <= 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(in != nullptr && out != nullptr, "null image pointer");
  if (s.in_bytes % 16 || s.out_bytes % 16 || !aligned16(in) || !aligned16(out)) return IMGTX_ERR_UNSUPPORTED;
  constexpr int TT = Op::kTeamWarps * 32, KI = Op::kInBufs, KO = Op::kOutBufs;
  const size_t team_bytes = align_up(128 + KI * align_up((size_t)s.in_bytes, 128) + KO * align_up((size_t)s.out_bytes, 128) +
                                         Op::team_scratch(s, p), 128);
  const size_t cta_scr = align_up(Op::cta_scratch(s, p), 128);
  const int max_smem = max_smem_optin();
  int ctas = Op::kCtasPerSm;
  int teams = 0;
  for (; ctas >= 1; --ctas) {
    const size_t budget = std::min<size_t>((size_t)max_smem, (size_t)smem_per_sm() / ctas - 1024);
    if (budget <= cta_scr + team_bytes) continue;
    teams = (int)std::min<size_t>((size_t)Op::kMaxTeams, (budget - cta_scr) / team_bytes);
    if (teams >= 1) break;
  }
  if (teams < 1) return IMGTX_ERR_UNSUPPORTED;
  if (Op::kTeamWarps > 1 && teams > 15) teams = 15;   // named barriers 1..15
  const size_t smem_bytes = cta_scr + (size_t)teams * team_bytes;
  auto kern = team_kernel<Op>;
  IMGTX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  int64_t grid = (int64_t)sm_count() * ctas;
  const int64_t need = (n + teams - 1) / teams;
  if (grid > need) grid = need;
  kern<<<(unsigned)grid, teams * TT, smem_bytes, stream>>>(in, out, n, s, p, (int)cta_scr, (int)team_bytes,
                                                           reinterpret_cast<uint2*>(metrics));
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}

struct NoTeamPrepare {
  template <class S, class P>
  __device__ static void prepare(const S&, const P&, uint8_t*) {}
  template <class S, class P>
  static size_t cta_scratch(const S&, const P&) { return 0; }
};

// ---- byte <-> float helpers of the fast forms ---------------------------------------------------------------------
// byte `k` (0..3) of word w as a float, exact: the byte is planted in the mantissa of 2^23 and the bias subtracted
__device__ __forceinline__ float byte_f(uint32_t w, int k) {
  return __fsub_rn(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7440u + k)), 8388608.0f);
}
// floor(v) for 0 <= v < 2^22 in the low bits of the returned word (one round-down add of 2^23); callers mask / permute
__device__ __forceinline__ uint32_t floor_bits(float v) { return __float_as_uint(__fadd_rd(v, 8388608.0f)); }
// bytes 0 of four words -> one word
__device__ __forceinline__ uint32_t pack4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  return __byte_perm(__byte_perm(a, b, 0x0040u), __byte_perm(c, d, 0x0040u), 0x5410u);
}

#endif  // __CUDACC__
}  // namespace imgtx
