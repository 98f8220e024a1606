// Sparse convolution forward, tcgen05 tensor-core mode (PCC_MODE_TF32).  Placeholder until the
// UMMA kernel lands: reports "unsupported" so callers fail loudly instead of silently falling back.
#include "pcc_common.cuh"

namespace pcc {
bool conv_tf32_supported(const pcc_conv_args*) { return false; }
int conv_tf32(const pcc_conv_args*, cudaStream_t) {
  set_error("pcc_conv_fwd: PCC_MODE_TF32 kernel not built");
  return PCC_ERR_INVALID_ARG;
}
}  // namespace pcc
