// Library runtime: error string, cached device attributes, staged-pipeline tuning.
#include "common.cuh"
#include "pipeline.cuh"
#include "team_pipeline.cuh"
#include <algorithm>
#include <atomic>
#include <mutex>
#include <stdlib.h>

namespace imgtx {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

struct DevInfo {
  int sms = 0;
  int smem = 0;
};
static DevInfo g_dev;
static std::once_flag g_dev_once;

static void init_dev() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  cudaDeviceGetAttribute(&g_dev.sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&g_dev.smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
}
// Stream-ordered workspaces (operators whose intermediates do not fit on chip) come from the device's default memory pool.
// Its release threshold is 0 out of the box: every synchronisation hands the freed workspace back to the driver and the
// next call maps it again (measured: ~1.6 ms for 768 MB).  Raised once per device to 1 GiB, so a workspace up to that size
// stays mapped between calls.
cudaError_t workspace_alloc(void** ptr, size_t bytes, cudaStream_t stream) {
  static std::mutex mu;
  static bool done[64] = {false};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64) {
    std::lock_guard<std::mutex> lock(mu);
    if (!done[dev]) {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        uint64_t cur = 0, want = 1ull << 30;
        if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &cur) == cudaSuccess && cur < want)
          cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &want);
      }
      done[dev] = true;
    }
  }
  return cudaMallocAsync(ptr, bytes, stream);
}

int sm_count() {
  std::call_once(g_dev_once, init_dev);
  return g_dev.sms > 0 ? g_dev.sms : 148;
}
int max_smem_optin() {
  std::call_once(g_dev_once, init_dev);
  return g_dev.smem > 0 ? g_dev.smem : 232448;
}

// Development sweeps: IMGTX_* environment variables, read once (std::call_once) and immutable afterwards.
static int g_target_smem = 56 * 1024, g_stages = 3, g_ctas_per_sm = 4, g_ctas_env = 0, g_target_scale = 0, g_glass_g = 0,
           g_pw_ctas = 4;
static std::once_flag g_env_once;
static void init_env() {
  if (const char* e = getenv("IMGTX_TARGET_SMEM")) g_target_smem = atoi(e);
  if (const char* e = getenv("IMGTX_STAGES")) g_stages = atoi(e);
  if (const char* e = getenv("IMGTX_CTAS_PER_SM")) { g_ctas_per_sm = atoi(e); g_ctas_env = 1; }
  if (const char* e = getenv("IMGTX_TARGET_SCALE")) g_target_scale = atoi(e);
  if (const char* e = getenv("IMGTX_GLASS_G")) g_glass_g = atoi(e);
  if (const char* e = getenv("IMGTX_PW_CTAS")) g_pw_ctas = atoi(e);
}
StagedTuning staged_tuning() {
  std::call_once(g_env_once, init_env);
  StagedTuning t;
  t.target_smem = g_target_smem;
  t.stages = g_stages < 2 ? 2 : (g_stages > kMaxStages ? kMaxStages : g_stages);
  t.ctas_per_sm = g_ctas_per_sm < 1 ? 1 : g_ctas_per_sm;
  t.ctas_from_env = g_ctas_env;
  t.target_scale = g_target_scale;
  t.glass_g = g_glass_g;
  t.pw_ctas = g_pw_ctas < 1 ? 1 : g_pw_ctas;
  return t;
}

}  // namespace imgtx

namespace imgtx {
// one warp per image; images of any size that is a multiple of 4 bytes (others: byte tail handled by lane 0)
__global__ void __launch_bounds__(256) checksum_kernel(const uint8_t* __restrict__ x, int64_t n, int64_t image_bytes,
                                                       uint2* __restrict__ metrics) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t img = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); img < n; img += warps) {
    const uint8_t* p = x + img * image_bytes;
    uint32_t c0 = 0, c1 = 0;
    if ((reinterpret_cast<uintptr_t>(p) & 3) == 0) {
      const uint32_t* p32 = reinterpret_cast<const uint32_t*>(p);
      for (int64_t i = lane; i < (image_bytes >> 2); i += 32) checksum_word(__ldg(p32 + i), (uint32_t)i, c0, c1);
      if (lane == 0)
        for (int64_t b = image_bytes & ~(int64_t)3; b < image_bytes; ++b) { c0 += p[b]; c1 += ((uint32_t)b + 1u) * p[b]; }
    } else {
      for (int64_t b = lane; b < image_bytes; b += 32) { c0 += p[b]; c1 += ((uint32_t)b + 1u) * p[b]; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      c0 += __shfl_xor_sync(0xffffffffu, c0, o);
      c1 += __shfl_xor_sync(0xffffffffu, c1, o);
    }
    if (lane == 0) metrics[img] = make_uint2(c0, c1);
  }
}

int image_checksums(const uint8_t* x, int64_t n, int64_t image_bytes, uint32_t* metrics, cudaStream_t stream) {
  if (n <= 0) return IMGTX_OK;
  IMGTX_CHECK_ARG(x && metrics && image_bytes > 0, "null pointer");
  const int64_t grid = std::min<int64_t>((n + 7) / 8, (int64_t)sm_count() * 8);
  checksum_kernel<<<(unsigned)grid, 256, 0, stream>>>(x, n, image_bytes, reinterpret_cast<uint2*>(metrics));
  IMGTX_LAUNCH_OK();
  return IMGTX_OK;
}
}  // namespace imgtx

extern "C" int imgtx_image_checksums_u8(const uint8_t* x, int64_t n, int64_t image_bytes, uint32_t* metrics_dev,
                                        imgtx_stream_t stream) {
  return imgtx::image_checksums(x, n, image_bytes, metrics_dev, (cudaStream_t)stream);
}

extern "C" int imgtx_version(void) { return IMGTX_VERSION; }
extern "C" const char* imgtx_last_error(void) { return imgtx::g_err; }
extern "C" int imgtx_device_info(int* sms, int* smem) {
  int dev = 0;
  IMGTX_CUDA_OK(cudaGetDevice(&dev));
  if (sms) *sms = imgtx::sm_count();
  if (smem) *smem = imgtx::max_smem_optin();
  return IMGTX_OK;
}
