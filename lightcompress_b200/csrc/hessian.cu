// placeholder until the tcgen05 Hessian kernel lands (next commit): symbols exist, calls fail loudly.
#include "common.cuh"
extern "C" int lcb_hessian_accum(float*, const void*, int64_t, int64_t, int32_t, double, double, void*, void*) { return LCB_E_UNSUPPORTED; }
extern "C" int64_t lcb_hessian_workspace_bytes(int64_t, int64_t) { return 0; }
