// Error text, ABI version and small host helpers behind include/aphrodite_b200.h.
#include <cstring>

#include <cstdlib>

#include "common.cuh"

namespace aphb200 {

char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("APH_PDL");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

bool dev_env_live() {
  static const bool on = getenv("APH_DEV_ENV") != nullptr;
  return on;
}

}  // namespace aphb200

extern "C" int aph_b200_abi_version(void) { return APH_B200_ABI_VERSION; }

extern "C" const char* aph_b200_last_error(void) { return aphb200::error_buffer(); }

extern "C" int aph_kv_dtype_from_string(const char* s) {
  if (!s) return -1;
  if (!strcmp(s, "auto")) return APH_KV_AUTO;
  if (!strcmp(s, "fp8") || !strcmp(s, "fp8_e4m3")) return APH_KV_FP8_E4M3;
  if (!strcmp(s, "fp8_e5m2")) return APH_KV_FP8_E5M2;
  return -1;
}
