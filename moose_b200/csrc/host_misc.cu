// Plaintext-side HostPlacement helpers on ring tensors: not on the MPC critical path,
// needed when a computation has host-placed tails (comparisons of revealed values, casts).
//   LessOp / GreaterOp::host_ring64/128_kernel  (moose/src/host/ops.rs:2051-2087, 2125-2161)
//   MuxOp::host_ring_kernel                     (moose/src/host/ops.rs:2227-2252)
//   SignOp::ring64/128_kernel                   (moose/src/host/ops.rs:2482-2506)
//   CastOp::ring_reduction_kernel               (moose/src/host/ops.rs:2277-2288)
//   MaximumOp::host_ring_kernel                 (moose/src/host/ops.rs:1432-1452)
#include "common.h"

namespace {
constexpr int TPB = 256;
inline unsigned blocks(int64_t n) { return (unsigned)((n + TPB - 1) / TPB); }

template <typename U> struct SignedOf;
template <> struct SignedOf<uint64_t> { typedef int64_t type; };
template <> struct SignedOf<u128> { typedef __int128 type; };

// bit = (x - y) as signed  <  0   (GREATER: > 0); one warp packs 32 results with a ballot
template <typename U, bool GREATER>
__global__ void __launch_bounds__(TPB) compare_kernel(const U *__restrict__ x, const U *__restrict__ y, uint32_t *__restrict__ out, int64_t n) {
  typedef typename SignedOf<U>::type S;
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  unsigned b = 0;
  if (e < n) {
    const S d = (S)(U)(x[e] - y[e]);
    b = GREATER ? (d > 0) : (d < 0);
  }
  const unsigned word = __ballot_sync(0xffffffffu, b);
  if ((threadIdx.x & 31) == 0 && (e & ~31ll) < n) out[e >> 5] = word;
}

template <typename U>
__global__ void __launch_bounds__(TPB) mux_kernel(const uint32_t *__restrict__ s, const U *__restrict__ x, const U *__restrict__ y,
                                                  U *__restrict__ out, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e >= n) return;
  const U sel = (U)((s[e >> 5] >> (e & 31)) & 1u);
  out[e] = sel * (U)(x[e] - y[e]) + y[e];  // s * (x - y) + y
}

template <typename U>
__global__ void __launch_bounds__(TPB) sign_kernel(const U *__restrict__ x, U *__restrict__ out, int64_t n) {
  typedef typename SignedOf<U>::type S;
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e < n) out[e] = ((S)x[e] < 0) ? (U)0 - (U)1 : (U)1;
}

__global__ void __launch_bounds__(TPB) cast128_64_kernel(const u128 *__restrict__ x, uint64_t *__restrict__ out, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e < n) out[e] = (uint64_t)x[e];  // el % 2^64
}

// unsigned comparison of Wrapping<T> (PartialOrd on the raw integer)
template <typename U>
__global__ void __launch_bounds__(TPB) max2_kernel(const U *__restrict__ a, const U *__restrict__ b, U *__restrict__ out, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e < n) out[e] = a[e] < b[e] ? b[e] : a[e];
}
}  // namespace

#define MB_REQUIRE(ctx, cond, code, msg)               \
  do {                                                 \
    if (!(cond)) return mb_fail((ctx), (code), (msg)); \
  } while (0)

static int compare_op(mb_ctx *ctx, bool greater, const mb_tensor *x, const mb_tensor *y, mb_tensor **out, const char *name) {
  MB_REQUIRE(ctx, ctx && x && y && out, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  MB_REQUIRE(ctx, mb_is_ring(x->dtype) && x->dtype == y->dtype, MB_ERR_TYPE_MISMATCH, std::string(name) + ": needs ring tensors of one width");
  // the reference subtracts with ndarray co-broadcasting; materialise the broadcast operands when shapes differ
  const mb_tensor *xs = x, *ys = y;
  mb_tensor *xb = nullptr, *yb = nullptr, *probe = nullptr;
  if (!mb_same_shape(x, y)) {
    MB_CHECK(mb_sub(ctx, x, y, &probe));
    int rc = mb_broadcast(ctx, x, probe->rank, probe->shape, &xb);
    if (rc == MB_OK) rc = mb_broadcast(ctx, y, probe->rank, probe->shape, &yb);
    if (rc != MB_OK) { mb_tensor_free(ctx, probe); mb_tensor_free(ctx, xb); return rc; }
    xs = xb; ys = yb;
  }
  mb_tensor *t = nullptr;
  int rc = mb_tensor_new(ctx, MB_BIT, xs->rank, xs->shape, &t);
  if (rc == MB_OK && t->numel) {
    cudaMemsetAsync(t->dptr, 0, t->alloc_bytes, ctx->stream);
    const int64_t n = t->numel, threads = ((n + 31) / 32) * 32;
    uint32_t *o = (uint32_t *)t->dptr;
    if (x->dtype == MB_RING128) {
      if (greater) compare_kernel<u128, true><<<blocks(threads), TPB, 0, ctx->stream>>>((const u128 *)xs->dptr, (const u128 *)ys->dptr, o, n);
      else compare_kernel<u128, false><<<blocks(threads), TPB, 0, ctx->stream>>>((const u128 *)xs->dptr, (const u128 *)ys->dptr, o, n);
    } else {
      if (greater) compare_kernel<uint64_t, true><<<blocks(threads), TPB, 0, ctx->stream>>>((const uint64_t *)xs->dptr, (const uint64_t *)ys->dptr, o, n);
      else compare_kernel<uint64_t, false><<<blocks(threads), TPB, 0, ctx->stream>>>((const uint64_t *)xs->dptr, (const uint64_t *)ys->dptr, o, n);
    }
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "host_misc");
    if (cudaGetLastError() != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, std::string(name) + ": launch failed");
  }
  mb_tensor_free(ctx, probe); mb_tensor_free(ctx, xb); mb_tensor_free(ctx, yb);
  if (rc != MB_OK) { mb_tensor_free(ctx, t); return rc; }
  *out = t;
  return MB_OK;
}

extern "C" int mb_less(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) { return compare_op(ctx, false, x, y, out, "mb_less"); }
extern "C" int mb_greater(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) { return compare_op(ctx, true, x, y, out, "mb_greater"); }

extern "C" int mb_mux(mb_ctx *ctx, const mb_tensor *s, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  MB_REQUIRE(ctx, ctx && s && x && y && out, MB_ERR_INVALID_ARGUMENT, "mb_mux: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  MB_REQUIRE(ctx, s->dtype == MB_BIT && mb_is_ring(x->dtype) && x->dtype == y->dtype, MB_ERR_TYPE_MISMATCH, "mb_mux: (bit, ring, ring)");
  MB_REQUIRE(ctx, mb_same_shape(s, x) && mb_same_shape(x, y), MB_ERR_KERNEL, "mb_mux: operands must have one shape");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, x->shape, &t));
  if (t->numel) {
    if (x->dtype == MB_RING128)
      mux_kernel<u128><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const uint32_t *)s->dptr, (const u128 *)x->dptr, (const u128 *)y->dptr, (u128 *)t->dptr, t->numel);
    else
      mux_kernel<uint64_t><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const uint32_t *)s->dptr, (const uint64_t *)x->dptr, (const uint64_t *)y->dptr, (uint64_t *)t->dptr, t->numel);
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "host_misc");
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_sign(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  MB_REQUIRE(ctx, ctx && x && out, MB_ERR_INVALID_ARGUMENT, "mb_sign: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_sign: needs a ring tensor");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, x->shape, &t));
  if (t->numel) {
    if (x->dtype == MB_RING128) sign_kernel<u128><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (u128 *)t->dptr, t->numel);
    else sign_kernel<uint64_t><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint64_t *)t->dptr, t->numel);
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "host_misc");
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_cast_ring128_to_ring64(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  MB_REQUIRE(ctx, ctx && x && out, MB_ERR_INVALID_ARGUMENT, "mb_cast_ring128_to_ring64: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  MB_REQUIRE(ctx, x->dtype == MB_RING128, MB_ERR_TYPE_MISMATCH, "mb_cast_ring128_to_ring64: needs a ring128 tensor");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, MB_RING64, x->rank, x->shape, &t));
  if (t->numel) {
    cast128_64_kernel<<<blocks(t->numel), TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (uint64_t *)t->dptr, t->numel);
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "host_misc");
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_maximum(mb_ctx *ctx, const mb_tensor *const *xs, int n, mb_tensor **out) {
  MB_REQUIRE(ctx, ctx && out, MB_ERR_INVALID_ARGUMENT, "mb_maximum: NULL argument");
  MB_REQUIRE(ctx, xs && n > 0, MB_ERR_INVALID_ARGUMENT, "cannot reduce on empty array of tensors");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int i = 0; i < n; i++)
    MB_REQUIRE(ctx, xs[i] && mb_is_ring(xs[i]->dtype) && xs[i]->dtype == xs[0]->dtype && mb_same_shape(xs[i], xs[0]),
               MB_ERR_TYPE_MISMATCH, "mb_maximum: ring tensors of one width and shape");
  mb_tensor *acc = nullptr;
  MB_CHECK(mb_reshape(ctx, xs[0], xs[0]->rank, xs[0]->shape, &acc));
  for (int i = 1; i < n && acc->numel; i++) {
    if (acc->dtype == MB_RING128)
      max2_kernel<u128><<<blocks(acc->numel), TPB, 0, ctx->stream>>>((const u128 *)acc->dptr, (const u128 *)xs[i]->dptr, (u128 *)acc->dptr, acc->numel);
    else
      max2_kernel<uint64_t><<<blocks(acc->numel), TPB, 0, ctx->stream>>>((const uint64_t *)acc->dptr, (const uint64_t *)xs[i]->dptr, (uint64_t *)acc->dptr, acc->numel);
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "host_misc");
  }
  *out = acc;
  return MB_OK;
}
