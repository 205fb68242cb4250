// Error plumbing and version queries of the C ABI (include/b2asr.h).
#include "common.cuh"
#include <string.h>

namespace b2asr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return B2ASR_ECUDA;
}

}  // namespace b2asr

extern "C" const char* b2asr_last_error(void) { return b2asr::g_err; }
extern "C" int b2asr_version(void) { return 100; }
extern "C" int b2asr_arch(void) {
#ifdef B2ASR_ARCH
  return B2ASR_ARCH;
#else
  return 100;
#endif
}
