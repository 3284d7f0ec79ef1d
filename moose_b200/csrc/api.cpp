// C ABI: host-placement kernel table (one entry per `[runtime]` kernel of
// moose/src/host/ops.rs / prim.rs on ring, bit, seed and key types).
#include "common.h"

#define MB_ENTER(ctx, what)                                                              \
  do {                                                                                   \
    if (!(ctx)) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, what ": NULL context"); \
    MB_CUDA((ctx), cudaSetDevice((ctx)->device));                                        \
  } while (0)

#define MB_REQUIRE(ctx, cond, code, msg) \
  do {                                   \
    if (!(cond)) return mb_fail((ctx), (code), (msg)); \
  } while (0)

// ---- primitives -------------------------------------------------------------------

extern "C" int mb_prf_key_gen(mb_ctx *ctx, uint8_t key_out[16]) {
  MB_ENTER(ctx, "mb_prf_key_gen");
  return mb_entropy16(ctx, key_out);
}

extern "C" int mb_sync_key_random(mb_ctx *ctx, uint8_t out[16]) {
  MB_ENTER(ctx, "mb_sync_key_random");
  return mb_entropy16(ctx, out);
}

extern "C" int mb_derive_seed(mb_ctx *ctx, const uint8_t sync_key[16], const uint8_t key[16], uint8_t seed_out[16]) {
  if (!ctx || !sync_key || !key || !seed_out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_derive_seed: NULL argument");
  mb_blake3_derive_seed(ctx->sid, sync_key, key, seed_out);
  return MB_OK;
}

static int sample_with_seed(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, const uint8_t seed[16],
                            int64_t max_value, mb_tensor **out, uint64_t first_block = 0) {
  MB_REQUIRE(ctx, out && (shape || rank == 0), MB_ERR_INVALID_ARGUMENT, "sample: NULL argument");
  MB_REQUIRE(ctx, dtype == MB_RING64 || dtype == MB_RING128 || dtype == MB_BIT, MB_ERR_TYPE_MISMATCH,
             "sample: dtype must be ring64, ring128 or bit");
  bool bits_as_ring = false;
  if (dtype == MB_BIT) {
    // host/ops.rs:2025-2029
    MB_REQUIRE(ctx, max_value < 0, MB_ERR_UNIMPLEMENTED,
               "SampleOp for HostBitTensor @ HostPlacement does not support max_value");
  } else if (max_value >= 0) {
    // host/ops.rs:1971-1979
    MB_REQUIRE(ctx, max_value == 1, MB_ERR_UNIMPLEMENTED,
               "SampleOp for HostRingTensor @ HostPlacement does not yet support max_value != 1");
    bits_as_ring = true;
  }
  AesKey k;
  mb_aes_expand(seed, &k);
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, dtype, rank, shape, &t));
  int rc = k_sample(ctx, k, t, bits_as_ring, first_block);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_sample_seeded(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, const uint8_t seed[16],
                                int64_t max_value, mb_tensor **out) {
  MB_ENTER(ctx, "mb_sample_seeded");
  MB_REQUIRE(ctx, seed, MB_ERR_INVALID_ARGUMENT, "mb_sample_seeded: NULL seed");
  return sample_with_seed(ctx, dtype, rank, shape, seed, max_value, out);
}

extern "C" int mb_sample_seeded_at(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, const uint8_t seed[16],
                                   uint64_t first_block, mb_tensor **out) {
  MB_ENTER(ctx, "mb_sample_seeded_at");
  MB_REQUIRE(ctx, seed, MB_ERR_INVALID_ARGUMENT, "mb_sample_seeded_at: NULL seed");
  return sample_with_seed(ctx, dtype, rank, shape, seed, -1, out, first_block);
}

extern "C" int mb_sample(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, int64_t max_value, mb_tensor **out) {
  MB_ENTER(ctx, "mb_sample");
  uint8_t seed[16];
  MB_CHECK(mb_entropy16(ctx, seed));  // AesRng::from_random_seed()
  return sample_with_seed(ctx, dtype, rank, shape, seed, max_value, out);
}

extern "C" int mb_fill(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, uint64_t lo, uint64_t hi, mb_tensor **out) {
  MB_ENTER(ctx, "mb_fill");
  MB_REQUIRE(ctx, out, MB_ERR_INVALID_ARGUMENT, "mb_fill: NULL out");
  MB_REQUIRE(ctx, dtype == MB_RING64 || dtype == MB_RING128 || dtype == MB_BIT, MB_ERR_TYPE_MISMATCH, "mb_fill: bad dtype");
  if (dtype == MB_BIT)  // kernels/constants.rs: bit fill accepts values in {0,1}
    MB_REQUIRE(ctx, lo <= 1 && hi == 0, MB_ERR_INVALID_ARGUMENT, "mb_fill: bit fill value must be 0 or 1");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, dtype, rank, shape, &t));
  int rc = k_fill(ctx, t, lo, hi);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

// ---- elementwise ------------------------------------------------------------------

static int broadcast_shape(const mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, int *rank, int64_t *shape) {
  int r = x->rank > y->rank ? x->rank : y->rank;
  for (int k = 0; k < r; k++) {
    int xk = k - (r - x->rank), yk = k - (r - y->rank);
    int64_t dx = xk >= 0 ? x->shape[xk] : 1, dy = yk >= 0 ? y->shape[yk] : 1;
    if (dx == dy) shape[k] = dx;
    else if (dx == 1) shape[k] = dy;
    else if (dy == 1) shape[k] = dx;
    else return mb_fail(ctx, MB_ERR_KERNEL, "operands could not be broadcast together");
  }
  *rank = r;
  return MB_OK;
}

static int binary_op(mb_ctx *ctx, EwOp op, const mb_tensor *x, const mb_tensor *y, mb_tensor **out, const char *name) {
  MB_REQUIRE(ctx, x && y && out, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL argument");
  MB_REQUIRE(ctx, x->dtype == y->dtype, MB_ERR_TYPE_MISMATCH, std::string(name) + ": operand dtypes differ");
  MB_REQUIRE(ctx, x->dtype != MB_FLOAT64, MB_ERR_TYPE_MISMATCH, std::string(name) + ": float tensors are not ring tensors");
  if (x->dtype == MB_BIT) {
    MB_REQUIRE(ctx, mb_same_shape(x, y), MB_ERR_KERNEL, std::string(name) + ": bit tensors must have equal shapes");
    // Z2: add = sub = xor, mul = and (kernels/boolean.rs:18-76)
    if (op == EW_ADD || op == EW_SUB) op = EW_XOR;
    if (op == EW_MUL) op = EW_AND;
  } else {
    MB_REQUIRE(ctx, op != EW_XOR && op != EW_OR, MB_ERR_TYPE_MISMATCH, std::string(name) + ": needs bit tensors");
  }
  int rank;
  int64_t shape[MB_MAX_RANK];
  MB_CHECK(broadcast_shape(ctx, x, y, &rank, shape));
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, rank, shape, &t));
  int rc = k_ew_binary(ctx, op, x, y, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_add(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_add");
  return binary_op(ctx, EW_ADD, x, y, out, "mb_add");
}
extern "C" int mb_sub(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_sub");
  return binary_op(ctx, EW_SUB, x, y, out, "mb_sub");
}
extern "C" int mb_mul(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_mul");
  return binary_op(ctx, EW_MUL, x, y, out, "mb_mul");
}
extern "C" int mb_xor(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_xor");
  MB_REQUIRE(ctx, x && x->dtype == MB_BIT, MB_ERR_TYPE_MISMATCH, "mb_xor: needs bit tensors");
  return binary_op(ctx, EW_XOR, x, y, out, "mb_xor");
}
extern "C" int mb_and(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_and");
  return binary_op(ctx, EW_AND, x, y, out, "mb_and");  // bit and ring flavours (kernels/boolean.rs)
}
extern "C" int mb_or(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_or");
  MB_REQUIRE(ctx, x && x->dtype == MB_BIT, MB_ERR_TYPE_MISMATCH, "mb_or: needs bit tensors");
  return binary_op(ctx, EW_OR, x, y, out, "mb_or");
}

extern "C" int mb_neg(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_neg");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_neg: NULL argument");
  MB_REQUIRE(ctx, x->dtype != MB_FLOAT64, MB_ERR_TYPE_MISMATCH, "mb_neg: not a ring/bit tensor");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, x->shape, &t));
  int rc = x->dtype == MB_BIT ? k_bit_not(ctx, x, t) : k_ew_neg(ctx, x, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

static int shift_op(mb_ctx *ctx, bool left, uint64_t amount, const mb_tensor *x, mb_tensor **out) {
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "shift: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "shift: needs a ring tensor");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, x->shape, &t));
  int rc = k_ew_shift(ctx, left, (unsigned)(amount & 127), x, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}
extern "C" int mb_shl(mb_ctx *ctx, uint64_t amount, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_shl");
  return shift_op(ctx, true, amount, x, out);
}
extern "C" int mb_shr(mb_ctx *ctx, uint64_t amount, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_shr");
  return shift_op(ctx, false, amount, x, out);
}

extern "C" int mb_add_n(mb_ctx *ctx, const mb_tensor *const *xs, int n, mb_tensor **out) {
  MB_ENTER(ctx, "mb_add_n");
  // host/ops.rs:1084-1103: empty input is an InvalidArgument
  MB_REQUIRE(ctx, xs && n > 0 && out, MB_ERR_INVALID_ARGUMENT, "cannot reduce on empty array of tensors");
  mb_tensor *acc = nullptr;
  for (int i = 0; i < n; i++) {
    MB_REQUIRE(ctx, xs[i], MB_ERR_INVALID_ARGUMENT, "mb_add_n: NULL operand");
    if (i == 0) {
      MB_CHECK(mb_reshape(ctx, xs[0], xs[0]->rank, xs[0]->shape, &acc));
    } else {
      mb_tensor *next = nullptr;
      int rc = binary_op(ctx, EW_ADD, acc, xs[i], &next, "mb_add_n");
      mb_tensor_free(ctx, acc);
      if (rc != MB_OK) return rc;
      acc = next;
    }
  }
  *out = acc;
  return MB_OK;
}

extern "C" int mb_sum(mb_ctx *ctx, const mb_tensor *x, int axis, mb_tensor **out) {
  MB_ENTER(ctx, "mb_sum");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_sum: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_sum: needs a ring tensor");
  MB_REQUIRE(ctx, axis < x->rank, MB_ERR_INVALID_ARGUMENT, "mb_sum: axis out of range");
  int64_t outer = 1, len = 1, inner = 1;
  int64_t oshape[MB_MAX_RANK];
  int orank = 0;
  if (axis < 0) {
    len = x->numel;
  } else {
    for (int k = 0; k < x->rank; k++) {
      if (k < axis) outer *= x->shape[k];
      else if (k == axis) len = x->shape[k];
      else inner *= x->shape[k];
      if (k != axis) oshape[orank++] = x->shape[k];
    }
  }
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, orank, oshape, &t));
  int rc = k_sum_axis(ctx, x, outer, len, inner, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_broadcast(mb_ctx *ctx, const mb_tensor *x, int rank, const int64_t *shape, mb_tensor **out) {
  MB_ENTER(ctx, "mb_broadcast");
  MB_REQUIRE(ctx, x && out && (shape || rank == 0), MB_ERR_INVALID_ARGUMENT, "mb_broadcast: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_broadcast: needs a ring tensor");
  MB_REQUIRE(ctx, rank >= x->rank && rank <= MB_MAX_RANK, MB_ERR_KERNEL, "mb_broadcast: target rank too small");
  for (int k = 0; k < x->rank; k++) {
    int64_t dx = x->shape[k], dt = shape[k + rank - x->rank];
    MB_REQUIRE(ctx, dx == dt || dx == 1, MB_ERR_KERNEL, "mb_broadcast: incompatible shapes");
  }
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, rank, shape, &t));
  int rc = k_broadcast(ctx, x, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_fixedpoint_encode(mb_ctx *ctx, mb_dtype ring, uint32_t frac, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_fixedpoint_encode");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_encode: NULL argument");
  MB_REQUIRE(ctx, x->dtype == MB_FLOAT64 && mb_is_ring(ring), MB_ERR_TYPE_MISMATCH, "mb_fixedpoint_encode: f64 -> ring only");
  MB_REQUIRE(ctx, frac < 127, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_encode: precision too large");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, ring, x->rank, x->shape, &t));
  int rc = k_fixed_encode(ctx, x, ldexp(1.0, (int)frac), t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_fixedpoint_decode(mb_ctx *ctx, uint32_t frac, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_fixedpoint_decode");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_decode: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_fixedpoint_decode: ring -> f64 only");
  MB_REQUIRE(ctx, frac < 127, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_decode: precision too large");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, MB_FLOAT64, x->rank, x->shape, &t));
  int rc = k_fixed_decode(ctx, x, ldexp(1.0, (int)frac), t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

// ---- dot ----------------------------------------------------------------------------

// Shape bookkeeping of HostRingTensor::dot (host/ops.rs:1754-1834).
int mb_dot_dims(const mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, int64_t *M, int64_t *K, int64_t *N, int *orank,
                int64_t *oshape) {
  if (x->rank != 1 && x->rank != 2)
    return mb_fail(ctx, MB_ERR_KERNEL, "Dot<HostRingTensor> not implemented for tensors of rank " + std::to_string(x->rank));
  if (y->rank != 1 && y->rank != 2)
    return mb_fail(ctx, MB_ERR_KERNEL, "Dot<HostRingTensor> cannot handle argument of rank " + std::to_string(y->rank));
  *M = x->rank == 2 ? x->shape[0] : 1;
  *K = x->shape[x->rank - 1];
  *N = y->rank == 2 ? y->shape[1] : 1;
  if (y->shape[0] != *K) return mb_fail(ctx, MB_ERR_KERNEL, "dot: inner dimensions differ");
  *orank = 0;
  if (x->rank == 2) oshape[(*orank)++] = *M;
  if (y->rank == 2) oshape[(*orank)++] = *N;
  return MB_OK;
}

extern "C" int mb_dot(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_ENTER(ctx, "mb_dot");
  MB_REQUIRE(ctx, x && y && out, MB_ERR_INVALID_ARGUMENT, "mb_dot: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype) && x->dtype == y->dtype, MB_ERR_TYPE_MISMATCH, "mb_dot: needs two ring tensors of one width");
  int64_t M, K, N, oshape[2];
  int orank;
  MB_CHECK(mb_dot_dims(ctx, x, y, &M, &K, &N, &orank, oshape));
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, orank, oshape, &t));
  int rc;
  bool limb_ok = k_gemm_limb_supported(x->dtype, M, K, N);
  if (ctx->gemm_mode == 2 && !limb_ok) {
    mb_tensor_free(ctx, t);
    return mb_fail(ctx, MB_ERR_KERNEL, "mb_dot: shape not supported by the tensor-pipe GEMM (gemm mode 2)");
  }
  if (limb_ok && (ctx->gemm_mode == 2 || (ctx->gemm_mode == 0 && k_gemm_limb_preferred(M, K, N)))) {
    GemmTerm term = {x->dptr, y->dptr, nullptr};
    rc = k_gemm_limb(ctx, x->dtype, &term, 1, t->dptr, M, K, N);
  } else {
    rc = k_gemm_imad(ctx, x->dtype, x->dptr, y->dptr, t->dptr, M, K, N, false);
  }
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

// ---- bit tensors ----------------------------------------------------------------------

extern "C" int mb_bit_decompose(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_bit_decompose");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_bit_decompose: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_bit_decompose: needs a ring tensor");
  MB_REQUIRE(ctx, x->rank < MB_MAX_RANK, MB_ERR_KERNEL, "mb_bit_decompose: rank too large");
  int64_t shape[MB_MAX_RANK];
  shape[0] = x->dtype == MB_RING128 ? 128 : 64;
  for (int k = 0; k < x->rank; k++) shape[k + 1] = x->shape[k];
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, MB_BIT, x->rank + 1, shape, &t));
  int rc = k_bit_decompose(ctx, x, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_bit_extract(mb_ctx *ctx, uint64_t bit_idx, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_bit_extract");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_bit_extract: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_bit_extract: needs a ring tensor");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, MB_BIT, x->rank, x->shape, &t));
  int rc = k_bit_extract(ctx, (unsigned)bit_idx, x, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_ring_inject(mb_ctx *ctx, mb_dtype ring, uint64_t bit_idx, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_ring_inject");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_ring_inject: NULL argument");
  MB_REQUIRE(ctx, x->dtype == MB_BIT && mb_is_ring(ring), MB_ERR_TYPE_MISMATCH, "mb_ring_inject: bit -> ring only");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, ring, x->rank, x->shape, &t));
  int rc = k_ring_inject(ctx, (unsigned)bit_idx, x, t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_shl_dim(mb_ctx *ctx, uint64_t amount, uint64_t bit_length, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_shl_dim");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_shl_dim: NULL argument");
  MB_REQUIRE(ctx, x->dtype == MB_BIT && x->rank >= 1, MB_ERR_TYPE_MISMATCH, "mb_shl_dim: needs a bit tensor of rank >= 1");
  MB_REQUIRE(ctx, amount <= bit_length && (int64_t)bit_length <= x->shape[0], MB_ERR_KERNEL, "mb_shl_dim: amount/bit_length out of range");
  int64_t height = x->shape[0] ? x->numel / x->shape[0] : 0;
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, MB_BIT, x->rank, x->shape, &t));
  // host/ops.rs:842-846: zeros(height*amount) ++ x[0 .. height*(bit_length-amount))
  int rc = k_bit_shift_copy(ctx, x, 0, height * (int64_t)amount, height * (int64_t)(bit_length - amount), t);
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_index_axis0(mb_ctx *ctx, int64_t index, const mb_tensor *x, mb_tensor **out) {
  MB_ENTER(ctx, "mb_index_axis0");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_index_axis0: NULL argument");
  MB_REQUIRE(ctx, x->rank >= 1, MB_ERR_KERNEL, "mb_index_axis0: rank 0 tensor");
  MB_REQUIRE(ctx, index >= 0 && index < x->shape[0], MB_ERR_KERNEL, "mb_index_axis0: index out of bounds");
  int64_t height = x->numel / x->shape[0];
  mb_tensor *t = nullptr;
  {
    // tensors are immutable: a 16-byte aligned slab (bit planes of n % 128 == 0 elements, ring rows) is a view
    const size_t slab = x->dtype == MB_BIT ? (size_t)height / 8 : (size_t)height * (size_t)mb_elem_bytes(x->dtype);
    const bool whole = x->dtype == MB_BIT ? height % 128 == 0 : slab % 16 == 0;
    static const bool no_views = getenv("MOOSE_B200_NO_VIEWS") != nullptr;
    if (whole && slab > 0 && !no_views) return mb_tensor_view(ctx, x, (size_t)index * slab, x->rank - 1, x->shape + 1, out);
  }
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank - 1, x->shape + 1, &t));
  int rc = MB_OK;
  if (x->dtype == MB_BIT) {
    rc = k_bit_shift_copy(ctx, x, index * height, 0, height, t);
  } else if (t->nbytes) {
    size_t eb = (size_t)mb_elem_bytes(x->dtype);
    cudaError_t e = cudaMemcpyAsync(t->dptr, (const char *)x->dptr + (size_t)(index * height) * eb, t->nbytes,
                                    cudaMemcpyDeviceToDevice, ctx->stream);
    if (e != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, cudaGetErrorString(e));
  }
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}
