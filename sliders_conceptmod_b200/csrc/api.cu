// Library-wide state: version, thread-local error text, launch counter.
#include <stdlib.h>
#include "common.cuh"
#include <string.h>

namespace slider {

static thread_local char t_err[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

bool pdl_enabled() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("SLIDER_PDL");
    // on by default since the GEMM producer fetches operands ahead of the CTA start barrier (denoise forward
    // 5.39 -> 5.20 ms, layer set 2062 -> 1993 us, profiles/r01_ab_pdl_v2.txt); SLIDER_PDL=0 switches it off
    cached = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return cached == 1;
}

int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}

}  // namespace slider

extern "C" {

int slider_version(void) { return SLIDER_ABI_VERSION; }

const char* slider_last_error(void) { return slider::t_err; }

int64_t slider_launch_count(void) { return (int64_t)slider::g_launches.load(); }

}  // extern "C"
