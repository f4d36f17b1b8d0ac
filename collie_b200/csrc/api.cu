// Library-wide state: last-error string, launch counter, device properties.
#include "common.cuh"

namespace cmf {
thread_local std::string g_last_error;
std::atomic<long long> g_launches{0};

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}
}  // namespace cmf

extern "C" const char* cmf_last_error(void) { return cmf::g_last_error.c_str(); }
extern "C" int cmf_version(void) { return 100; }
extern "C" int64_t cmf_launch_count(void) { return cmf::g_launches.load(); }
