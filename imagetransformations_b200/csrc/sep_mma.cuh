// Separable integer operators on 32x32x3 images, evaluated on the tensor cores without leaving the register file.
//
// A wide separable filter on a CIFAR image is not a stencil: cv2.GaussianBlur with 13..31 taps, Pillow's LANCZOS / BICUBIC /
// BOX resample (7-9 taps out of 32, reflected or clipped at the border) are dense or nearly dense 32x32 operators per axis,
//      out = round( V * round( X * H^T ) ),
// and the scalar forms spend 30-50 issue slots per byte on them (C1 blur 6 %, LANCZOS 7 %, pixelate 13 % of the HBM peak).
// Integer MMA does the same sums exactly: `mma.sync.m16n8k32` u8 x u8/s8 -> s32 (SASS IMMA.16832, measured 2.1 clk per
// warp-MMA per SM = 1920 MAC/clk/SM; tools/ubench/pipes.cu).  One WARP owns one image:
//   * stage 1 computes T^T[xo][y] = sum_x H[xo][x] X[y][x]: A = the operator (fragments prebuilt once per CTA), B = the image
//     rows, de-interleaved from RGB with byte permutes straight into B fragments;
//   * the accumulator fragment of a stage IS the B fragment of the next stage up to a fixed permutation of the contraction
//     index (a lane holds C[m][2t], C[m][2t+1] of every 8-column tile; packing two tiles gives it four k-slots of column
//     n' = m), so the next operator is stored with its columns permuted the same way and the intermediate never goes to shared
//     memory (the back-to-back GEMM trick).  Stages alternate horizontal / vertical, any number of them;
//   * Q22 coefficients (Pillow) are split into three byte limbs k = k2 2^16 + k1 2^8 + k0 (k2 signed): three IMMAs per tile
//     chained through the accumulator, c = ((((2^21 + S0) >> 8) + S1) >> 8) + S2) >> 6 — nested floors of integers, i.e.
//     exactly (2^21 + sum x k) >> 22 — then one saturating pack (I2IP) gives clip8 and the next stage's bytes;
//   * cv2's Q8 kernel has byte weights: one IMMA for the row pass, two for the column pass (high and low byte plane of the
//     u16 intermediate), out = byte 2 of (256 Shi + Slo + 2^15).
// Everything is integer arithmetic: results are bit-identical to the scalar kernels (tests compare both with the oracle).
#pragma once
#include "team_pipeline.cuh"

namespace imgtx {
#ifdef __CUDACC__

__device__ __forceinline__ void imma_uu(int (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}
// signed operator limb (A), unsigned data (B)
__device__ __forceinline__ void imma_su(int (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}
// bytes [sat8(v0), sat8(v1), sat8(v2), sat8(v3)]
__device__ __forceinline__ uint32_t sat_pack4(int v0, int v1, int v2, int v3) {
  uint32_t t, r;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v3), "r"(v2), "r"(0));
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(v1), "r"(v0), "r"(t));
  return r;
}

// Contraction slot k of a stage fed from the previous accumulators holds previous column sep_perm(k).
__host__ __device__ __forceinline__ int sep_perm(int k) {
  return (k & 16) | (((k >> 1) & 1) << 3) | (((k >> 2) & 3) << 1) | (k & 1);
}

constexpr int kSepOpBytes = 32 * 32 * 4;       // one operator as int32 [out][in]
constexpr int kSepFragBytes = 2 * 32 * 16;     // A fragments of one limb: [m-tile][lane] uint4

// All threads of the CTA: A fragments of byte limbs first_limb .. first_limb + limbs - 1 of op (int32 [32][32], row = output
// index) -> frag[limb][mt][lane].
__device__ inline void sep_build_frags(const int32_t* op, uint4* frag, int first_limb, int limbs, bool permuted) {
  uint32_t* f = reinterpret_cast<uint32_t*>(frag);
  for (int e = threadIdx.x; e < limbs * 256; e += blockDim.x) {
    const int reg = e & 3, lane = (e >> 2) & 31, mt = (e >> 7) & 1, limb = first_limb + (e >> 8);
    const int gid = lane >> 2, tig = lane & 3;
    const int row = mt * 16 + gid + (reg & 1) * 8, kb = (reg >> 1) * 16 + tig * 4;
    uint32_t w = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = kb + i;
      const int v = op[row * 32 + (permuted ? sep_perm(k) : k)];
      w |= (uint32_t)((v >> (8 * limb)) & 255) << (8 * i);
    }
    f[e] = w;
  }
}

// B fragments of the three channel planes of an interleaved 32x32x3 image: B[ch][nt][half] = X[y = 8 nt + gid][x = 16 half + 4 tig ..+3]
__device__ __forceinline__ void sep_load_image(const uint8_t* img, int lane, uint32_t (&B)[3][4][2]) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(img) + (lane >> 2) * 24 + (lane & 3) * 3;
#pragma unroll
  for (int nt = 0; nt < 4; ++nt)
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const uint32_t* q = w + nt * 192 + hf * 12;
      const uint32_t w0 = q[0], w1 = q[1], w2 = q[2];
      B[0][nt][hf] = __byte_perm(__byte_perm(w0, w1, 0x6630), w2, 0x5210);   // bytes 0, 3, 6, 9
      B[1][nt][hf] = __byte_perm(__byte_perm(w0, w1, 0x7741), w2, 0x6210);   // bytes 1, 4, 7, 10
      B[2][nt][hf] = __byte_perm(__byte_perm(w0, w1, 0x5552), w2, 0x7410);   // bytes 2, 5, 8, 11
    }
}

// P[ch][rs][half]: bytes of row yo = 8 rs + gid at xo = 16 half + 2 tig, +1 (low half-word) and 16 half + 8 + 2 tig, +1
__device__ __forceinline__ void sep_store_image(uint8_t* out, int lane, const uint32_t (&P)[3][4][2]) {
  uint16_t* o = reinterpret_cast<uint16_t*>(out) + (lane >> 2) * 48 + (lane & 3) * 3;
#pragma unroll
  for (int rs = 0; rs < 4; ++rs)
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const uint32_t R = P[0][rs][hf], G = P[1][rs][hf], Bc = P[2][rs][hf];
      uint16_t* q = o + rs * 384 + hf * 24;
      q[0] = (uint16_t)__byte_perm(R, G, 0x0040);
      q[1] = (uint16_t)__byte_perm(Bc, R, 0x0050);
      q[2] = (uint16_t)__byte_perm(G, Bc, 0x0051);
      q[12] = (uint16_t)__byte_perm(R, G, 0x0062);
      q[13] = (uint16_t)__byte_perm(Bc, R, 0x0072);
      q[14] = (uint16_t)__byte_perm(G, Bc, 0x0073);
    }
}

// One Q22 stage for one channel plane.  LIMBS == 3: general coefficients; LIMBS == 1: coefficients that are multiples of 2^16.
// frag points at this stage's first limb.  B -> Bn (the next stage's B fragments / the final packed bytes).
template <int LIMBS>
__device__ __forceinline__ void sep_q22_stage(const uint4 (&A)[LIMBS][2], const uint32_t (&B)[4][2], uint32_t (&Bn)[4][2]) {
  int c[2][4][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) c[mt][nt][i] = LIMBS == 3 ? (1 << 21) : (1 << 5);
  if (LIMBS == 3) {
#pragma unroll
    for (int l = 0; l < 2; ++l) {
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) imma_uu(c[mt][nt], A[l][mt], B[nt][0], B[nt][1]);
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
          for (int i = 0; i < 4; ++i) c[mt][nt][i] >>= 8;
    }
  }
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) imma_su(c[mt][nt], A[LIMBS - 1][mt], B[nt][0], B[nt][1]);
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int hf = 0; hf < 2; ++hf)
#pragma unroll
      for (int h = 0; h < 2; ++h)
        Bn[2 * mt + hf][h] = sat_pack4(c[mt][2 * h][2 * hf] >> 6, c[mt][2 * h][2 * hf + 1] >> 6, c[mt][2 * h + 1][2 * hf] >> 6,
                                       c[mt][2 * h + 1][2 * hf + 1] >> 6);
}

template <int LIMBS>
__device__ __forceinline__ void sep_q22_stage3(const uint4* frag, int lane, uint32_t (&B)[3][4][2]) {
  uint4 A[LIMBS][2];
#pragma unroll
  for (int l = 0; l < LIMBS; ++l) {
    A[l][0] = frag[(l * 2) * 32 + lane];
    A[l][1] = frag[(l * 2 + 1) * 32 + lane];
  }
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    uint32_t Bn[4][2];
    sep_q22_stage<LIMBS>(A, B[ch], Bn);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) { B[ch][nt][0] = Bn[nt][0]; B[ch][nt][1] = Bn[nt][1]; }
  }
}

// Stage list as limb counts; L == 0 ends the chain.  frag: the stages' limbs one after the other.
template <int L0, int L1, int L2, int L3>
__device__ __forceinline__ void sep_q22_chain(const uint8_t* src, uint8_t* dst, const uint4* frag, int lane) {
  uint32_t B[3][4][2];
  sep_load_image(src, lane, B);
  sep_q22_stage3<L0>(frag, lane, B);
  if (L1) sep_q22_stage3<(L1 ? L1 : 1)>(frag + L0 * 64, lane, B);
  if (L2) sep_q22_stage3<(L2 ? L2 : 1)>(frag + (L0 + L1) * 64, lane, B);
  if (L3) sep_q22_stage3<(L3 ? L3 : 1)>(frag + (L0 + L1 + L2) * 64, lane, B);
  __syncwarp();   // in place: every lane has read its part of the image
  sep_store_image(dst, lane, B);
}

// cv2's Q8 separable kernel (weights <= 255 after the border fold, sum 256): frag[0] = row operator, frag[1] = column
// operator with permuted contraction slots (64 uint4 each).
__device__ __forceinline__ void sep_cvq8_chain(const uint8_t* src, uint8_t* dst, const uint4* frag, int lane) {
  uint32_t B[3][4][2];
  sep_load_image(src, lane, B);
  const uint4 A1[2] = {frag[lane], frag[32 + lane]}, A2[2] = {frag[64 + lane], frag[96 + lane]};
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    int c[2][4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) c[mt][nt][i] = 0;
        imma_uu(c[mt][nt], A1[mt], B[ch][nt][0], B[ch][nt][1]);
      }
    uint32_t lo[4][2], hi[4][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int hf = 0; hf < 2; ++hf)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t t0 = __byte_perm(c[mt][2 * h][2 * hf], c[mt][2 * h][2 * hf + 1], 0x5140);
          const uint32_t t1 = __byte_perm(c[mt][2 * h + 1][2 * hf], c[mt][2 * h + 1][2 * hf + 1], 0x5140);
          lo[2 * mt + hf][h] = __byte_perm(t0, t1, 0x5410);
          hi[2 * mt + hf][h] = __byte_perm(t0, t1, 0x7632);
        }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) c[mt][nt][i] = 0;
        imma_uu(c[mt][nt], A2[mt], hi[nt][0], hi[nt][1]);
#pragma unroll
        for (int i = 0; i < 4; ++i) c[mt][nt][i] = (c[mt][nt][i] << 8) + 32768;
        imma_uu(c[mt][nt], A2[mt], lo[nt][0], lo[nt][1]);
      }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int hf = 0; hf < 2; ++hf)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          B[ch][2 * mt + hf][h] = __byte_perm(__byte_perm(c[mt][2 * h][2 * hf], c[mt][2 * h][2 * hf + 1], 0x0062),
                                              __byte_perm(c[mt][2 * h + 1][2 * hf], c[mt][2 * h + 1][2 * hf + 1], 0x0062), 0x5410);
  }
  __syncwarp();
  sep_store_image(dst, lane, B);
}

#endif  // __CUDACC__
}  // namespace imgtx
