// Error reporting and device queries shared by every entry point.
#include <stdarg.h>
#include <stdlib.h>

#include <atomic>

#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

static thread_local std::string g_last_error;

void set_error(const std::string& msg) { g_last_error = msg; }

long long launches();

int fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return 1;
}

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
long long launches() { return g_launches.load(std::memory_order_relaxed); }

int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

int current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return -1;
  return dev;
}

bool pdl_enabled() {
  static const int on = env_int("JM_PDL", 0);  // measured on B200: no gain without the early trigger, -2 % with it
  return on != 0;
}

int sm_count() {
  static int cached[64];
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace jm

extern "C" const char* jm_last_error(void) { return jm::g_last_error.c_str(); }

extern "C" int jm_version(void) { return 100; }

extern "C" long long jm_launch_count(void) { return jm::launches(); }

extern "C" int jm_device_ok(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

extern "C" int jm_zero(void* ptr, size_t bytes, void* stream) {
  if (bytes == 0) return 0;
  JM_CHECK(ptr != nullptr, "jm_zero: null pointer");
  JM_CUDA(cudaMemsetAsync(ptr, 0, bytes, static_cast<cudaStream_t>(stream)));
  return 0;
}
