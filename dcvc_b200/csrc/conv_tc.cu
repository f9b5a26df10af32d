// conv_tc.cu -- tcgen05 / TMEM implicit-GEMM convolution (kind::tf32, fp32 accumulate).
// Placeholder until the tensor-core kernel lands: nothing is eligible, so dcvc_conv2d always
// takes the CUDA-core kernel.
#include "common.cuh"

bool conv_tc_ok(const dcvc_conv *) { return false; }
int conv_tc_launch(const dcvc_conv *, int, int, cudaStream_t) {
  return dcvc_fail(DCVC_ERR_UNSUPPORTED, "conv_tc: not built");
}
extern "C" size_t dcvc_conv_tc_pack_bytes(int, int, int) { return 0; }
extern "C" int dcvc_conv_tc_pack(const float *, int, int, int, float *) {
  return dcvc_fail(DCVC_ERR_UNSUPPORTED, "conv_tc: not built");
}
