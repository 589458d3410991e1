// tcgen05 / TMA / TMEM implicit-GEMM convolution (placeholder until the kernels land).
#include "rn_common.cuh"
using namespace rn;
extern "C" int rn_convtc_supported(const rn_conv_desc*, int) { return 0; }
extern "C" size_t rn_convtc_scratch_bytes(const rn_conv_desc*, int) { return 0; }
extern "C" int rn_convtc_plan_create(const rn_conv_desc*, int, const void*, const void*, void*, void*, size_t,
                                     rn_convtc_plan**) {
  return fail(RN_ERR_UNSUPPORTED, "convtc: not built");
}
extern "C" int rn_convtc_run(rn_convtc_plan*, float, void*) { return fail(RN_ERR_UNSUPPORTED, "convtc: not built"); }
extern "C" int rn_convtc_plan_destroy(rn_convtc_plan*) { return RN_OK; }
extern "C" int rn_weights_to_bf16(const float*, int, int, int, void*, void*, void*) {
  return fail(RN_ERR_UNSUPPORTED, "convtc: not built");
}
