// Library-level entry points: version, error string, device facts.
#include <stdarg.h>
#include <stdio.h>

#include "common.cuh"

namespace b200attn {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  (void)cudaGetLastError();  // clear the (non-sticky) error so it is not reported again by a later, unrelated call
  set_error("CUDA error %d (%s) in %s", static_cast<int>(e), cudaGetErrorString(e), what);
  return B200ATTN_ERR_CUDA;
}

struct DevFacts {
  int dev = -1, sms = 0, major = 0, minor = 0, smem = 0;
};
static thread_local DevFacts g_facts;

static const DevFacts& facts() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return g_facts;
  if (g_facts.dev != dev) {
    DevFacts f;
    f.dev = dev;
    cudaDeviceGetAttribute(&f.sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&f.major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&f.minor, cudaDevAttrComputeCapabilityMinor, dev);
    cudaDeviceGetAttribute(&f.smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    g_facts = f;
  }
  return g_facts;
}

int sm_count() {
  const int n = facts().sms;
  return n > 0 ? n : 148;
}
int max_smem_optin() { return facts().smem; }

}  // namespace b200attn

using namespace b200attn;

extern "C" B2_EXPORT int b200attn_version(void) { return B200ATTN_VERSION; }

extern "C" B2_EXPORT const char* b200attn_last_error(void) { return g_err; }

extern "C" B2_EXPORT int b200attn_device_info(int* sms, int* cc_major, int* cc_minor, int* max_smem) {
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  const DevFacts& f = facts();
  if (sms) *sms = f.sms;
  if (cc_major) *cc_major = f.major;
  if (cc_minor) *cc_minor = f.minor;
  if (max_smem) *max_smem = f.smem;
  return B200ATTN_OK;
}
