// Error plumbing and version of the C ABI.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

namespace hmm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("CUDA error %s (%d) in %s", cudaGetErrorString(e), (int)e, what);
  return HMM_B200_ECUDA;
}

}  // namespace hmm

extern "C" {
int hmm_b200_abi_version(void) { return 1; }
const char* hmm_b200_last_error(void) { return hmm::g_err; }
}
