// Error plumbing and launch accounting for liblipvq.so.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace lipvq {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return LIPVQ_ERR_CUDA;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

bool pdl_enabled() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("LIPVQ_PDL");
    cached = (e && e[0] == '0') ? 0 : 1;
  }
  return cached == 1;
}

int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return 0;
  return dev < kMaxDevices ? dev : kMaxDevices - 1;
}

int sm_count() {
  static int cached[kMaxDevices] = {};  // per device: a process may drive several GPUs
  const int dev = current_device();
  if (cached[dev] > 0) return cached[dev];
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  cached[dev] = n;
  return n;
}

}  // namespace lipvq

extern "C" {
int lipvq_abi_version(void) { return LIPVQ_ABI_VERSION; }
const char* lipvq_last_error(void) { return lipvq::g_err; }
uint64_t lipvq_launch_count(void) { return lipvq::g_launches.load(std::memory_order_relaxed); }
}
