// abi.cu — error bookkeeping and version probe of the C ABI.
#include <stdio.h>
#include "common.cuh"

namespace lcb {
static thread_local char g_last_error[256] = "";
void set_last_error(const char* what, cudaError_t e) {
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what, cudaGetErrorString(e));
}
}  // namespace lcb

extern "C" int lcb_abi_version(void) { return LCB_ABI_VERSION; }
extern "C" const char* lcb_last_error(void) { return lcb::g_last_error; }
