#include "common.cuh"
extern "C" int lcb_awq_candidates(const void*, const void*, const void*, const void*, float*, int64_t, int64_t, int64_t, int32_t, int32_t, const lcb_quant_spec*, void*, void*) { return LCB_E_UNSUPPORTED; }
extern "C" int64_t lcb_awq_workspace_bytes(int64_t, int64_t, int64_t, int32_t) { return 0; }
