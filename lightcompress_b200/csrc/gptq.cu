#include "common.cuh"
extern "C" int lcb_gptq_weight_transform(float*, const float*, float*, float*, float*, float*, const int32_t*, int64_t, int64_t, int64_t, int32_t, int32_t, int32_t, const lcb_quant_spec*, void*, void*) { return LCB_E_UNSUPPORTED; }
extern "C" int64_t lcb_gptq_workspace_bytes(int64_t, int64_t, int32_t) { return 0; }
