// Library runtime: error string, cached device attributes, staged-pipeline tuning.
#include "common.cuh"
#include "pipeline.cuh"
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

extern "C" int imgtx_version(void) { return IMGTX_VERSION; }
extern "C" const char* imgtx_last_error(void) { return imgtx::g_err; }
extern "C" int imgtx_device_info(int* sms, int* smem) {
  int dev = 0;
  IMGTX_CUDA_OK(cudaGetDevice(&dev));
  if (sms) *sms = imgtx::sm_count();
  if (smem) *smem = imgtx::max_smem_optin();
  return IMGTX_OK;
}
