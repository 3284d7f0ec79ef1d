// C ABI: share-wise (local) replicated shape / reduction operators.  In the reference each of these applies the
// host kernel to the six share tensors (three distinct ones here):
//   IndexAxisOp / ExpandDimsOp / SqueezeOp / TransposeOp / ReshapeOp / BroadcastOp / SliceOp / DiagOp / ConcatOp
//                                           (moose/src/replicated/ops.rs:48-89, 267-360, 378-473, 540-650)
//   SumOp / AddNOp rep kernels              (moose/src/replicated/arith.rs:123-195)
//   CastOp::rep_reduction_kernel            (moose/src/replicated/convert.rs:499-530)
//   RingFixedpointMeanOp::rep_kernel        (moose/src/replicated/fixedpoint.rs:49-78)
// No interaction, no randomness: HBM-bound copies / reductions (shape_ops.cu, elementwise.cu).
#include <functional>

#include "common.h"

namespace {

bool valid(const mb_rep *x) { return x && x->s[0] && x->s[1] && x->s[2]; }

int sharewise(mb_ctx *ctx, const char *name, const mb_rep *x, mb_rep **out,
              const std::function<int(const mb_tensor *, mb_tensor **)> &f) {
  if (!ctx) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL context");
  if (!valid(x) || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  int rc = MB_OK;
  for (int i = 0; i < 3 && rc == MB_OK; i++) rc = f(x->s[i], &z[i]);
  if (rc == MB_OK) rc = mb_rep_from_shares(ctx, z[0], z[1], z[2], out);
  for (int i = 0; i < 3; i++)
    if (z[i]) mb_tensor_free(ctx, z[i]);
  return rc;
}

int sharewise_n(mb_ctx *ctx, const char *name, const mb_rep *const *xs, int n, mb_rep **out,
                const std::function<int(const mb_tensor *const *, int, mb_tensor **)> &f) {
  if (!ctx) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL context");
  if (!xs || n <= 0 || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": needs at least one tensor");
  for (int j = 0; j < n; j++)
    if (!valid(xs[j])) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  int rc = MB_OK;
  std::vector<const mb_tensor *> col((size_t)n);
  for (int i = 0; i < 3 && rc == MB_OK; i++) {
    for (int j = 0; j < n; j++) col[(size_t)j] = xs[j]->s[i];
    rc = f(col.data(), n, &z[i]);
  }
  if (rc == MB_OK) rc = mb_rep_from_shares(ctx, z[0], z[1], z[2], out);
  for (int i = 0; i < 3; i++)
    if (z[i]) mb_tensor_free(ctx, z[i]);
  return rc;
}

}  // namespace

extern "C" int mb_rep_index_axis(mb_ctx *ctx, int axis, int64_t index, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_index_axis", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_index_axis(ctx, axis, index, t, o); });
}
extern "C" int mb_rep_expand_dims(mb_ctx *ctx, const int64_t *axes, int naxes, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_expand_dims", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_expand_dims(ctx, axes, naxes, t, o); });
}
extern "C" int mb_rep_squeeze(mb_ctx *ctx, int axis, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_squeeze", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_squeeze(ctx, axis, t, o); });
}
extern "C" int mb_rep_transpose(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_transpose", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_transpose(ctx, t, o); });
}
extern "C" int mb_rep_reshape(mb_ctx *ctx, const mb_rep *x, int rank, const int64_t *shape, mb_rep **out) {
  return sharewise(ctx, "mb_rep_reshape", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_reshape(ctx, t, rank, shape, o); });
}
extern "C" int mb_rep_broadcast(mb_ctx *ctx, const mb_rep *x, int rank, const int64_t *shape, mb_rep **out) {
  return sharewise(ctx, "mb_rep_broadcast", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_broadcast(ctx, t, rank, shape, o); });
}
extern "C" int mb_rep_slice(mb_ctx *ctx, int n, const int64_t *start, const int64_t *end, const uint8_t *has_end, const int64_t *step,
                            const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_slice", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_slice(ctx, n, start, end, has_end, step, t, o); });
}
extern "C" int mb_rep_diag(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_diag", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_diag(ctx, t, o); });
}
extern "C" int mb_rep_sum(mb_ctx *ctx, const mb_rep *x, int axis, mb_rep **out) {
  return sharewise(ctx, "mb_rep_sum", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_sum(ctx, t, axis, o); });
}
extern "C" int mb_rep_cast_ring128_to_ring64(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_cast_ring128_to_ring64", x, out, [&](const mb_tensor *t, mb_tensor **o) { return mb_cast_ring128_to_ring64(ctx, t, o); });
}
extern "C" int mb_rep_fixedpoint_mean(mb_ctx *ctx, int axis, uint64_t scaling_base, uint32_t scaling_exp, const mb_rep *x, mb_rep **out) {
  return sharewise(ctx, "mb_rep_fixedpoint_mean", x, out,
                   [&](const mb_tensor *t, mb_tensor **o) { return mb_fixedpoint_mean(ctx, axis, scaling_base, scaling_exp, t, o); });
}
extern "C" int mb_rep_concat(mb_ctx *ctx, int axis, const mb_rep *const *xs, int n, mb_rep **out) {
  return sharewise_n(ctx, "mb_rep_concat", xs, n, out, [&](const mb_tensor *const *c, int k, mb_tensor **o) { return mb_concat(ctx, axis, c, k, o); });
}
extern "C" int mb_rep_add_n(mb_ctx *ctx, const mb_rep *const *xs, int n, mb_rep **out) {
  return sharewise_n(ctx, "mb_rep_add_n", xs, n, out, [&](const mb_tensor *const *c, int k, mb_tensor **o) { return mb_add_n(ctx, c, k, o); });
}
