// Library-wide state of libb200attn.so: error string, launch counter, device queries.
#include "common.cuh"

namespace b200 {
thread_local char g_err[512] = {0};
std::atomic<int64_t> g_launches{0};

int cached_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
      (void)cudaGetLastError();
      n = 148;
    }
    sms[dev] = n;
  }
  return sms[dev];
}
}  // namespace b200

extern "C" int b200_abi_version(void) { return B200ATTN_ABI_VERSION; }
extern "C" const char* b200_last_error(void) { return b200::g_err; }
extern "C" int b200_num_sms(void) {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return n;
}
extern "C" int64_t b200_launch_count(void) { return b200::g_launches.load(); }
extern "C" void b200_reset_launch_count(void) { b200::g_launches.store(0); }
