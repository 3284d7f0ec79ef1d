// placeholder until the tcgen05 kernel lands
#include "common.h"
bool k_gemm_limb_supported(mb_dtype, int64_t, int64_t, int64_t) { return false; }
int k_gemm_limb(mb_ctx *ctx, mb_dtype, const GemmTerm *, int, void *, int64_t, int64_t, int64_t) {
  return mb_fail(ctx, MB_ERR_UNIMPLEMENTED, "tensor-pipe GEMM not built");
}
