// Pipe-throughput microbenchmark (development tool): lanes/clk/SM of a few instruction classes on this GPU.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  float f0 = threadIdx.x, f1 = f0 + 1, f2 = f0 + 2, f3 = f0 + 3, f4 = f0 + 4, f5 = f0 + 5, f6 = f0 + 6, f7 = f0 + 7;
  float fa = (float)a, fb = (float)b;
  unsigned u0 = threadIdx.x, u1 = u0 + 1, u2 = u0 + 2, u3 = u0 + 3, u4 = u0 + 4, u5 = u0 + 5, u6 = u0 + 6, u7 = u0 + 7, ua = (unsigned)a | 3u;
  unsigned v0 = u0 * 3, v1 = u1 * 3, v2 = u2 * 3, v3 = u3 * 3, v4 = u4 * 3, v5 = u5 * 3, v6 = u6 * 3, v7 = u7 * 3;
  for (int i = 0; i < iters; ++i) {
    if (OP == 0) { x0 = __fma_rn(x0, a, b); x1 = __fma_rn(x1, a, b); x2 = __fma_rn(x2, a, b); x3 = __fma_rn(x3, a, b); x4 = __fma_rn(x4, a, b); x5 = __fma_rn(x5, a, b); x6 = __fma_rn(x6, a, b); x7 = __fma_rn(x7, a, b); }
    if (OP == 1) { x0 = __dadd_rn(x0, a); x1 = __dadd_rn(x1, a); x2 = __dadd_rn(x2, a); x3 = __dadd_rn(x3, a); x4 = __dadd_rn(x4, a); x5 = __dadd_rn(x5, a); x6 = __dadd_rn(x6, a); x7 = __dadd_rn(x7, a); }
    if (OP == 2) { f0 = __fmaf_rn(f0, fa, fb); f1 = __fmaf_rn(f1, fa, fb); f2 = __fmaf_rn(f2, fa, fb); f3 = __fmaf_rn(f3, fa, fb); f4 = __fmaf_rn(f4, fa, fb); f5 = __fmaf_rn(f5, fa, fb); f6 = __fmaf_rn(f6, fa, fb); f7 = __fmaf_rn(f7, fa, fb); }
    if (OP == 3) { u0 = u0 * ua + 1; u1 = u1 * ua + 1; u2 = u2 * ua + 1; u3 = u3 * ua + 1; u4 = u4 * ua + 1; u5 = u5 * ua + 1; u6 = u6 * ua + 1; u7 = u7 * ua + 1; }
    if (OP == 4) { u0 = __umulhi(u0, ua) + 1; u1 = __umulhi(u1, ua) + 1; u2 = __umulhi(u2, ua) + 1; u3 = __umulhi(u3, ua) + 1; u4 = __umulhi(u4, ua) + 1; u5 = __umulhi(u5, ua) + 1; u6 = __umulhi(u6, ua) + 1; u7 = __umulhi(u7, ua) + 1; }
    if (OP == 5) { f0 = (float)(int)f0 + fa; f1 = (float)(int)f1 + fa; f2 = (float)(int)f2 + fa; f3 = (float)(int)f3 + fa; f4 = (float)(int)f4 + fa; f5 = (float)(int)f5 + fa; f6 = (float)(int)f6 + fa; f7 = (float)(int)f7 + fa; }  // F2I + I2F + FADD
    if (OP == 6) { u0 = (u0 ^ ua) + (u0 >> 3); u1 = (u1 ^ ua) + (u1 >> 3); u2 = (u2 ^ ua) + (u2 >> 3); u3 = (u3 ^ ua) + (u3 >> 3); u4 = (u4 ^ ua) + (u4 >> 3); u5 = (u5 ^ ua) + (u5 >> 3); u6 = (u6 ^ ua) + (u6 >> 3); u7 = (u7 ^ ua) + (u7 >> 3); }  // LOP3+SHF+IADD
    if (OP == 8) { u0 = __dp2a_lo(u0, ua, u0); u1 = __dp2a_lo(u1, ua, u1); u2 = __dp2a_hi(u2, ua, u2); u3 = __dp2a_hi(u3, ua, u3); u4 = __dp4a(u4, ua, u4); u5 = __dp4a(u5, ua, u5); u6 = __dp2a_lo(u6, ua, u6); u7 = __dp2a_lo(u7, ua, u7); }  // IDP
    if (OP == 9) { unsigned long long t0 = (unsigned long long)u0 * ua + u1, t1 = (unsigned long long)u2 * ua + u3, t2 = (unsigned long long)u4 * ua + u5, t3 = (unsigned long long)u6 * ua + u7; u0 = (unsigned)t0; u1 = (unsigned)(t0 >> 32); u2 = (unsigned)t1; u3 = (unsigned)(t1 >> 32); u4 = (unsigned)t2; u5 = (unsigned)(t2 >> 32); u6 = (unsigned)t3; u7 = (unsigned)(t3 >> 32); }  // 4 x IMAD.WIDE
    if (OP == 10) { f0 = __fadd_rd(f0, fa); f1 = __fadd_rd(f1, fa); f2 = __fadd_rd(f2, fa); f3 = __fadd_rd(f3, fa); f4 = __fadd_rd(f4, fa); f5 = __fadd_rd(f5, fa); f6 = __fadd_rd(f6, fa); f7 = __fadd_rd(f7, fa); }
    if (OP == 11) { u0 = __byte_perm(u0, ua, 0x5140); u1 = __byte_perm(u1, ua, 0x7362); u2 = __byte_perm(u2, ua, 0x5140); u3 = __byte_perm(u3, ua, 0x7362); u4 = __byte_perm(u4, ua, 0x5140); u5 = __byte_perm(u5, ua, 0x7362); u6 = __byte_perm(u6, ua, 0x5140); u7 = __byte_perm(u7, ua, 0x7362); }
    if (OP == 12) {   // 4 independent HMMA m16n8k16 f16 -> f32 per iteration
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(f0), "+f"(f1), "+f"(f2), "+f"(f3) : "r"(u0), "r"(u1), "r"(u2), "r"(u3), "r"(u4), "r"(u5));
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(f4), "+f"(f5), "+f"(f6), "+f"(f7) : "r"(u0), "r"(u1), "r"(u2), "r"(u3), "r"(u6), "r"(u7));
      float g0 = (float)x0, g1 = (float)x1, g2 = (float)x2, g3 = (float)x3, g4 = (float)x4, g5 = (float)x5, g6 = (float)x6, g7 = (float)x7;
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(g0), "+f"(g1), "+f"(g2), "+f"(g3) : "r"(u0), "r"(u1), "r"(u2), "r"(u3), "r"(u4), "r"(u5));
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(g4), "+f"(g5), "+f"(g6), "+f"(g7) : "r"(u0), "r"(u1), "r"(u2), "r"(u3), "r"(u6), "r"(u7));
      x0 = g0; x1 = g1; x2 = g2; x3 = g3; x4 = g4; x5 = g5; x6 = g6; x7 = g7;
    }
    if (OP == 13) {   // 4 independent IMMA m16n8k32 u8 x s8 -> s32 per iteration
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+r"(u0), "+r"(u1), "+r"(u2), "+r"(u3) : "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua));
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+r"(u4), "+r"(u5), "+r"(u6), "+r"(u7) : "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua));
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+r"(v0), "+r"(v1), "+r"(v2), "+r"(v3) : "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua));
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+r"(v4), "+r"(v5), "+r"(v6), "+r"(v7) : "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua));
    }
    if (OP == 7) { f0 = __log2f(f0) + fa; f1 = __log2f(f1) + fa; f2 = __log2f(f2) + fa; f3 = __log2f(f3) + fa; f4 = __log2f(f4) + fa; f5 = __log2f(f5) + fa; f6 = __log2f(f6) + fa; f7 = __log2f(f7) + fa; }  // MUFU + FADD
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + f0 + f1 + f2 + f3 + f4 + f5 + f6 + f7 + u0 + u1 + u2 + u3 + u4 + u5 + u6 + u7 + v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
}
template <int OP>
void run(const char* name, int ops_per_iter, double* out) {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const int iters = 4096, blocks = sms * 4, threads = 512;
  k<OP><<<blocks, threads>>>(out, 16, 1.0000001, 0.5);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k<OP><<<blocks, threads>>>(out, iters, 1.0000001, 0.5);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double lane_ops = (double)blocks * threads * iters * 8 * ops_per_iter;
  printf("%-28s %8.3f ms  %7.1f lane-ops/clk/SM (at %d MHz nominal)  %8.2f Tops/s\n", name, ms, lane_ops / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000, lane_ops / (ms * 1e-3) / 1e12);
}
int main() {
  double* out; cudaMalloc(&out, 148 * 4 * 512 * 8 * 2);
  run<0>("DFMA", 1, out); run<1>("DADD", 1, out); run<2>("FFMA", 1, out); run<3>("IMAD", 1, out); run<4>("IMAD.HI(+IADD)", 1, out);
  run<8>("IDP.2A/4A", 1, out); run<9>("IMAD.WIDE (x4 per 8)", 1, out); run<10>("FADD.RD", 1, out); run<11>("PRMT", 1, out);
  run<12>("HMMA.16816.F32 (4/iter: lane-ops = warp-MMAs*32/8*...)", 1, out);
  run<13>("IMMA.16832.S32 u8xs8 (4/iter)", 1, out);
  run<5>("F2I+I2F+FADD (3 ops)", 3, out); run<6>("LOP3+SHF+IADD (3 ops)", 3, out); run<7>("MUFU.LG2+FADD (2 ops)", 2, out);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
