// pcc_conv_fwd: argument validation and dispatch to the exact-fp32 or tcgen05 path.
#include "pcc_common.cuh"

namespace pcc {
int conv_exact(const pcc_conv_args* a, cudaStream_t st);
int conv_tf32(const pcc_conv_args* a, cudaStream_t st);
bool conv_tf32_supported(const pcc_conv_args* a);
}  // namespace pcc

extern "C" int pcc_conv_fwd(const pcc_conv_args* a, void* stream) {
  PCC_REQUIRE(a != nullptr, "null args");
  if (a->n_out == 0) return PCC_OK;
  PCC_REQUIRE(a->in && a->weight && a->nbr && a->out, "null pointer");
  PCC_REQUIRE(a->cin > 0 && a->cout > 0 && a->K > 0 && a->K <= 27, "bad shape");
  PCC_REQUIRE(a->nbr_stride >= a->n_out, "nbr_stride < n_out");
  PCC_REQUIRE(a->out_stride >= a->cout, "out_stride < cout");
  PCC_REQUIRE(a->film_c == 0 || a->film_c == 1 || a->film_c == a->cout, "film_c must be 0, 1 or cout");
  PCC_REQUIRE(a->film_c == 0 || a->film != nullptr, "film_c set without film");
  PCC_REQUIRE(a->act >= PCC_ACT_NONE && a->act <= PCC_ACT_LEAKY, "bad activation");
  cudaStream_t st = (cudaStream_t)stream;
  if (a->mode == PCC_MODE_EXACT) return pcc::conv_exact(a, st);
  PCC_REQUIRE(a->n_groups <= 1 || (a->order && a->n_groups <= a->K), "n_groups > 1 needs order and n_groups <= K");
  if (a->mode == PCC_MODE_TF32) {
    PCC_REQUIRE(a->n_groups <= 1, "PCC_MODE_TF32 takes kernel maps built with n_groups == 1");
    if (!pcc::conv_tf32_supported(a)) {
      pcc::set_error("pcc_conv_fwd: PCC_MODE_TF32 needs cin %% 32 == 0 and cout %% 32 == 0, 32 <= cout <= 256 (got %d -> %d)",
                     a->cin, a->cout);
      return PCC_ERR_INVALID_ARG;
    }
    return pcc::conv_tf32(a, st);
  }
  PCC_REQUIRE(false, "unknown mode");
}
