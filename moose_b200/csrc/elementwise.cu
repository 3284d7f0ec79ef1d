// Elementwise host-placement kernels on ring / bit tensors (SURVEY.md 8a rows
// a8, a9, a11, a12, a16, a20).  All are HBM-bound: 128-bit vector loads/stores,
// four independent vectors in flight per thread, no shared memory.
#include "common.h"

namespace {

constexpr int TPB = 256;
constexpr int VPT = 4;  // 16-byte vectors per thread

__device__ __forceinline__ u128 to128(ulonglong2 v) { return ((u128)v.y << 64) | (u128)v.x; }
__device__ __forceinline__ ulonglong2 from128(u128 v) { return make_ulonglong2((uint64_t)v, (uint64_t)(v >> 64)); }

template <bool W128, int OP>
__device__ __forceinline__ ulonglong2 ew_apply(ulonglong2 a, ulonglong2 b) {
  if (OP == EW_AND) return make_ulonglong2(a.x & b.x, a.y & b.y);
  if (OP == EW_XOR) return make_ulonglong2(a.x ^ b.x, a.y ^ b.y);
  if (OP == EW_OR) return make_ulonglong2(a.x | b.x, a.y | b.y);
  if (W128) {
    u128 x = to128(a), y = to128(b);
    u128 z = OP == EW_ADD ? x + y : OP == EW_SUB ? x - y : x * y;
    return from128(z);
  } else {
    if (OP == EW_ADD) return make_ulonglong2(a.x + b.x, a.y + b.y);
    if (OP == EW_SUB) return make_ulonglong2(a.x - b.x, a.y - b.y);
    return make_ulonglong2(a.x * b.x, a.y * b.y);
  }
}

// Same-shape contiguous operands.
template <bool W128, int OP>
__global__ void __launch_bounds__(TPB) ew_same_kernel(const ulonglong2 *__restrict__ x, const ulonglong2 *__restrict__ y,
                                                      ulonglong2 *__restrict__ z, int64_t nvec) {
  int64_t base = (int64_t)blockIdx.x * (TPB * VPT) + threadIdx.x;
  ulonglong2 a[VPT], b[VPT];
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i < nvec) {
      a[j] = x[i];
      b[j] = y[i];
    }
  }
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i < nvec) z[i] = ew_apply<W128, OP>(a[j], b[j]);
  }
}

// The three shares of a replicated tensor in one launch (blockIdx.y selects the share): share-wise protocol steps
// are thousands of small launches per predictor call, a third of them is enough.
struct Ptr3 {
  const ulonglong2 *x[3];
  const ulonglong2 *y[3];
  ulonglong2 *z[3];
};

template <bool W128, int OP>
__global__ void __launch_bounds__(TPB) ew_same3_kernel(Ptr3 p, int64_t nvec) {
  const ulonglong2 *__restrict__ x = p.x[blockIdx.y];
  const ulonglong2 *__restrict__ y = p.y[blockIdx.y];
  ulonglong2 *__restrict__ z = p.z[blockIdx.y];
  int64_t base = (int64_t)blockIdx.x * (TPB * VPT) + threadIdx.x;
  ulonglong2 a[VPT], b[VPT];
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i < nvec) {
      a[j] = x[i];
      b[j] = y[i];
    }
  }
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i < nvec) z[i] = ew_apply<W128, OP>(a[j], b[j]);
  }
}

// MODE 0: negate, 1: shift left, 2: shift right
template <bool W128, int MODE>
__global__ void __launch_bounds__(TPB) unary3_kernel(Ptr3 p, int64_t nvec, unsigned s) {
  const ulonglong2 *__restrict__ x = p.x[blockIdx.y];
  ulonglong2 *__restrict__ z = p.z[blockIdx.y];
  int64_t base = (int64_t)blockIdx.x * (TPB * VPT) + threadIdx.x;
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i >= nvec) continue;
    const ulonglong2 a = x[i];
    if (W128) {
      const u128 v = to128(a);
      z[i] = from128(MODE == 0 ? (u128)0 - v : MODE == 1 ? v << s : v >> s);
    } else {
      z[i] = MODE == 0   ? make_ulonglong2(0ull - a.x, 0ull - a.y)
             : MODE == 1 ? make_ulonglong2(a.x << s, a.y << s)
                         : make_ulonglong2(a.x >> s, a.y >> s);
    }
  }
}

struct BcastDesc {
  int rank;
  int64_t shape[MB_MAX_RANK];
  int64_t sx[MB_MAX_RANK];
  int64_t sy[MB_MAX_RANK];
};

// General co-broadcast (ndarray semantics, host/mod.rs:813-829): one element per thread.
template <typename T, int OP>
__global__ void __launch_bounds__(TPB) ew_bcast_kernel(const T *__restrict__ x, const T *__restrict__ y, T *__restrict__ z,
                                                       int64_t n, BcastDesc d) {
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i >= n) return;
  int64_t rem = i, ox = 0, oy = 0;
#pragma unroll
  for (int k = MB_MAX_RANK - 1; k >= 0; k--) {
    if (k < d.rank) {
      int64_t c = rem % d.shape[k];
      rem /= d.shape[k];
      ox += c * d.sx[k];
      oy += c * d.sy[k];
    }
  }
  T a = x[ox], b = y[oy];
  z[i] = OP == EW_ADD ? a + b : OP == EW_SUB ? a - b : OP == EW_MUL ? a * b : a & b;
}

template <bool W128>
__global__ void __launch_bounds__(TPB) neg_kernel(const ulonglong2 *__restrict__ x, ulonglong2 *__restrict__ z, int64_t nvec) {
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i >= nvec) return;
  ulonglong2 a = x[i];
  if (W128)
    z[i] = from128((u128)0 - to128(a));
  else
    z[i] = make_ulonglong2(0ull - a.x, 0ull - a.y);
}

template <bool W128, bool LEFT>
__global__ void __launch_bounds__(TPB) shift_kernel(const ulonglong2 *__restrict__ x, ulonglong2 *__restrict__ z, int64_t nvec,
                                                    unsigned s) {
  int64_t base = (int64_t)blockIdx.x * (TPB * VPT) + threadIdx.x;
  ulonglong2 a[VPT];
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i < nvec) a[j] = x[i];
  }
#pragma unroll
  for (int j = 0; j < VPT; j++) {
    int64_t i = base + (int64_t)j * TPB;
    if (i < nvec) {
      if (W128) {
        u128 v = to128(a[j]);
        z[i] = from128(LEFT ? v << s : v >> s);
      } else {
        z[i] = LEFT ? make_ulonglong2(a[j].x << s, a[j].y << s) : make_ulonglong2(a[j].x >> s, a[j].y >> s);
      }
    }
  }
}

__global__ void __launch_bounds__(TPB) fill_kernel(ulonglong2 *__restrict__ z, int64_t nvec, ulonglong2 v) {
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i < nvec) z[i] = v;
}

template <typename T>
__global__ void __launch_bounds__(TPB) bcast_copy_kernel(const T *__restrict__ x, T *__restrict__ z, int64_t n, BcastDesc d) {
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i >= n) return;
  int64_t rem = i, ox = 0;
#pragma unroll
  for (int k = MB_MAX_RANK - 1; k >= 0; k--) {
    if (k < d.rank) {
      int64_t c = rem % d.shape[k];
      rem /= d.shape[k];
      ox += c * d.sx[k];
    }
  }
  z[i] = x[ox];
}

// z[o, i] = sum_l x[o, l, i]; one thread per output element, coalesced over i.
template <typename T>
__global__ void __launch_bounds__(TPB) sum_axis_kernel(const T *__restrict__ x, T *__restrict__ z, int64_t outer, int64_t len,
                                                       int64_t inner) {
  int64_t idx = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (idx >= outer * inner) return;
  int64_t o = idx / inner, i = idx % inner;
  const T *p = x + o * len * inner + i;
  T acc = 0;
  for (int64_t l = 0; l < len; l++) acc += p[l * inner];
  z[idx] = acc;
}

// Rust `f64 as i128` / `as i64` (then `as uK`): truncation toward zero, saturating,
// NaN -> 0.  U is the unsigned ring type; the result is the two's complement pattern.
template <typename U>
__device__ __forceinline__ U f64_to_ring_sat(double v) {
  constexpr int BITS = sizeof(U) * 8;
  const U INT_MIN_P = (U)1 << (BITS - 1);
  const U INT_MAX_P = INT_MIN_P - 1;
  uint64_t b = (uint64_t)__double_as_longlong(v);
  bool negv = b >> 63;
  int e = (int)((b >> 52) & 0x7ff);
  uint64_t m = b & 0xfffffffffffffull;
  if (e == 0x7ff) {
    if (m) return (U)0;  // NaN
    return negv ? INT_MIN_P : INT_MAX_P;
  }
  if (e < 1023) return (U)0;  // |v| < 1 (incl. subnormals)
  m |= 1ull << 52;
  int sh = e - 1075;  // |v| = m * 2^sh
  U mag;
  if (sh >= 0) {
    if (e - 1023 >= BITS - 1) return negv ? INT_MIN_P : INT_MAX_P;  // |v| >= 2^(BITS-1)
    mag = (U)m << sh;
  } else {
    mag = (U)(m >> (-sh));
  }
  return negv ? (U)0 - mag : mag;
}

// Rust `(x as iK) as f64`: round to nearest even.  x is the two's complement pattern.
template <typename U>
__device__ __forceinline__ double ring_to_f64_rne(U x) {
  constexpr int BITS = sizeof(U) * 8;
  bool negv = (x >> (BITS - 1)) & 1;
  U mag = negv ? (U)0 - x : x;  // |INT_MIN| = 2^(BITS-1) is representable in U
  if (mag == 0) return 0.0;
  int p;  // index of the most significant set bit
  uint64_t hi64 = BITS == 128 ? (uint64_t)((u128)mag >> 64) : 0;
  if (hi64)
    p = 127 - __clzll((long long)hi64);
  else
    p = 63 - __clzll((long long)(uint64_t)mag);
  double r;
  if (p <= 52) {
    r = (double)(uint64_t)mag;
  } else {
    int sh = p - 52;
    uint64_t top = (uint64_t)(mag >> sh);  // 53 significant bits
    U rest = mag & (((U)1 << sh) - 1);
    U half = (U)1 << (sh - 1);
    if (rest > half || (rest == half && (top & 1))) top++;
    r = ldexp((double)top, sh);  // top <= 2^53 is exactly representable
  }
  return negv ? -r : r;
}

template <typename U>
__global__ void __launch_bounds__(TPB) fixed_encode_kernel(const double *__restrict__ x, U *__restrict__ z, int64_t n, double scale) {
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i < n) z[i] = f64_to_ring_sat<U>(x[i] * scale);
}
template <typename U>
__global__ void __launch_bounds__(TPB) fixed_decode_kernel(const U *__restrict__ x, double *__restrict__ z, int64_t n, double scale) {
  int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (i < n) z[i] = ring_to_f64_rne<U>(x[i]) / scale;
}

inline unsigned grid_for(int64_t n, int per_block) { return (unsigned)((n + per_block - 1) / per_block); }

void fill_strides(const mb_tensor *t, int out_rank, const int64_t *out_shape, int64_t *strides) {
  // right-aligned broadcasting; stride 0 where the operand dimension is 1 or missing
  int64_t s = 1;
  int off = out_rank - t->rank;
  for (int k = out_rank - 1; k >= 0; k--) {
    int tk = k - off;
    if (tk < 0) {
      strides[k] = 0;
    } else {
      strides[k] = (t->shape[tk] == 1 && out_shape[k] != 1) ? 0 : s;
      s *= t->shape[tk];
    }
  }
}

}  // namespace

int k_ew_binary(mb_ctx *ctx, EwOp op, const mb_tensor *x, const mb_tensor *y, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  bool same = mb_same_shape(x, out) && mb_same_shape(y, out);
  if (same || x->dtype == MB_BIT) {
    int64_t nvec = (int64_t)(out->alloc_bytes / 16);
    if (x->dtype == MB_BIT) nvec = (int64_t)((out->nbytes + 15) / 16);
    unsigned g = grid_for(nvec, TPB * VPT);
    const ulonglong2 *px = (const ulonglong2 *)x->dptr, *py = (const ulonglong2 *)y->dptr;
    ulonglong2 *pz = (ulonglong2 *)out->dptr;
    bool w = x->dtype == MB_RING128;
#define LAUNCH_SAME(OP)                                                                 \
  if (w)                                                                                \
    ew_same_kernel<true, OP><<<g, TPB, 0, ctx->stream>>>(px, py, pz, nvec);             \
  else                                                                                  \
    ew_same_kernel<false, OP><<<g, TPB, 0, ctx->stream>>>(px, py, pz, nvec);
    switch (op) {
      case EW_ADD: LAUNCH_SAME(EW_ADD) break;
      case EW_SUB: LAUNCH_SAME(EW_SUB) break;
      case EW_MUL: LAUNCH_SAME(EW_MUL) break;
      case EW_AND: LAUNCH_SAME(EW_AND) break;
      case EW_XOR: LAUNCH_SAME(EW_XOR) break;
      case EW_OR: LAUNCH_SAME(EW_OR) break;
    }
#undef LAUNCH_SAME
    MB_LAUNCH_CHECK_W(ctx, "ew_binary", 16.0 * nvec * 3, 0, 0);
    return MB_OK;
  }
  BcastDesc d;
  d.rank = out->rank;
  for (int k = 0; k < MB_MAX_RANK; k++) d.shape[k] = k < out->rank ? out->shape[k] : 1;
  fill_strides(x, out->rank, out->shape, d.sx);
  fill_strides(y, out->rank, out->shape, d.sy);
  unsigned g = grid_for(out->numel, TPB);
#define LAUNCH_BC(T, OP) \
  ew_bcast_kernel<T, OP><<<g, TPB, 0, ctx->stream>>>((const T *)x->dptr, (const T *)y->dptr, (T *)out->dptr, out->numel, d)
  if (x->dtype == MB_RING128) {
    switch (op) {
      case EW_ADD: LAUNCH_BC(u128, EW_ADD); break;
      case EW_SUB: LAUNCH_BC(u128, EW_SUB); break;
      case EW_MUL: LAUNCH_BC(u128, EW_MUL); break;
      default: LAUNCH_BC(u128, EW_AND); break;
    }
  } else {
    switch (op) {
      case EW_ADD: LAUNCH_BC(uint64_t, EW_ADD); break;
      case EW_SUB: LAUNCH_BC(uint64_t, EW_SUB); break;
      case EW_MUL: LAUNCH_BC(uint64_t, EW_MUL); break;
      default: LAUNCH_BC(uint64_t, EW_AND); break;
    }
  }
#undef LAUNCH_BC
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// x[i], y[i], out[i] (i = 0..2): same dtype and shape throughout
int k_ew_binary3(mb_ctx *ctx, EwOp op, const mb_tensor *const x[3], const mb_tensor *const y[3], mb_tensor *const out[3]) {
  if (out[0]->numel == 0) return MB_OK;
  int64_t nvec = (int64_t)(out[0]->alloc_bytes / 16);
  if (out[0]->dtype == MB_BIT) nvec = (int64_t)((out[0]->nbytes + 15) / 16);
  Ptr3 p;
  for (int i = 0; i < 3; i++) {
    p.x[i] = (const ulonglong2 *)x[i]->dptr;
    p.y[i] = (const ulonglong2 *)y[i]->dptr;
    p.z[i] = (ulonglong2 *)out[i]->dptr;
  }
  const dim3 g(grid_for(nvec, TPB * VPT), 3);
  const bool w = out[0]->dtype == MB_RING128;
#define LAUNCH_SAME3(OP)                                           \
  if (w)                                                           \
    ew_same3_kernel<true, OP><<<g, TPB, 0, ctx->stream>>>(p, nvec); \
  else                                                             \
    ew_same3_kernel<false, OP><<<g, TPB, 0, ctx->stream>>>(p, nvec);
  switch (op) {
    case EW_ADD: LAUNCH_SAME3(EW_ADD) break;
    case EW_SUB: LAUNCH_SAME3(EW_SUB) break;
    case EW_MUL: LAUNCH_SAME3(EW_MUL) break;
    case EW_AND: LAUNCH_SAME3(EW_AND) break;
    case EW_XOR: LAUNCH_SAME3(EW_XOR) break;
    case EW_OR: LAUNCH_SAME3(EW_OR) break;
  }
#undef LAUNCH_SAME3
  MB_LAUNCH_CHECK_W(ctx, "ew_binary3", 16.0 * nvec * 9, 0, 0);
  return MB_OK;
}

// mode 0: negate, 1: shl, 2: shr (ring tensors of one dtype and shape)
int k_ew_unary3(mb_ctx *ctx, int mode, unsigned amount, const mb_tensor *const x[3], mb_tensor *const out[3]) {
  if (out[0]->numel == 0) return MB_OK;
  const int64_t nvec = (int64_t)(out[0]->alloc_bytes / 16);
  Ptr3 p;
  for (int i = 0; i < 3; i++) {
    p.x[i] = (const ulonglong2 *)x[i]->dptr;
    p.y[i] = nullptr;
    p.z[i] = (ulonglong2 *)out[i]->dptr;
  }
  const dim3 g(grid_for(nvec, TPB * VPT), 3);
  const bool w = out[0]->dtype == MB_RING128;
  const unsigned s = amount & (w ? 127u : 63u);
#define LAUNCH_U3(M)                                                  \
  if (w)                                                              \
    unary3_kernel<true, M><<<g, TPB, 0, ctx->stream>>>(p, nvec, s);   \
  else                                                                \
    unary3_kernel<false, M><<<g, TPB, 0, ctx->stream>>>(p, nvec, s);
  if (mode == 0) { LAUNCH_U3(0) } else if (mode == 1) { LAUNCH_U3(1) } else { LAUNCH_U3(2) }
#undef LAUNCH_U3
  MB_LAUNCH_CHECK_W(ctx, "ew_unary3", 16.0 * nvec * 6, 0, 0);
  return MB_OK;
}

int k_ew_neg(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  int64_t nvec = (int64_t)(out->alloc_bytes / 16);
  unsigned g = grid_for(nvec, TPB);
  if (x->dtype == MB_RING128)
    neg_kernel<true><<<g, TPB, 0, ctx->stream>>>((const ulonglong2 *)x->dptr, (ulonglong2 *)out->dptr, nvec);
  else
    neg_kernel<false><<<g, TPB, 0, ctx->stream>>>((const ulonglong2 *)x->dptr, (ulonglong2 *)out->dptr, nvec);
  MB_LAUNCH_CHECK_W(ctx, "ew_neg", 16.0 * nvec * 2, 0, 0);
  return MB_OK;
}

int k_ew_shift(mb_ctx *ctx, bool left, unsigned amount, const mb_tensor *x, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  int64_t nvec = (int64_t)(out->alloc_bytes / 16);
  unsigned g = grid_for(nvec, TPB * VPT);
  const ulonglong2 *px = (const ulonglong2 *)x->dptr;
  ulonglong2 *pz = (ulonglong2 *)out->dptr;
  if (x->dtype == MB_RING128) {
    unsigned s = amount & 127;  // Wrapping<u128> shift semantics
    if (left) shift_kernel<true, true><<<g, TPB, 0, ctx->stream>>>(px, pz, nvec, s);
    else shift_kernel<true, false><<<g, TPB, 0, ctx->stream>>>(px, pz, nvec, s);
  } else {
    unsigned s = amount & 63;
    if (left) shift_kernel<false, true><<<g, TPB, 0, ctx->stream>>>(px, pz, nvec, s);
    else shift_kernel<false, false><<<g, TPB, 0, ctx->stream>>>(px, pz, nvec, s);
  }
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_fill(mb_ctx *ctx, mb_tensor *out, uint64_t lo, uint64_t hi) {
  if (out->numel == 0) return MB_OK;
  if (out->dtype == MB_BIT) {
    MB_CUDA(ctx, cudaMemsetAsync(out->dptr, (lo & 1) ? 0xff : 0x00, out->alloc_bytes, ctx->stream));
    if ((lo & 1)) {
      // restore the zero tail invariant
      size_t full = (size_t)(out->numel / 8);
      unsigned rem = (unsigned)(out->numel % 8);
      if (full < out->alloc_bytes)
        MB_CUDA(ctx, cudaMemsetAsync((char *)out->dptr + full, 0, out->alloc_bytes - full, ctx->stream));
      if (rem) MB_CUDA(ctx, cudaMemsetAsync((char *)out->dptr + full, (1u << rem) - 1, 1, ctx->stream));
    }
    return MB_OK;
  }
  int64_t nvec = (int64_t)(out->alloc_bytes / 16);
  ulonglong2 v = out->dtype == MB_RING128 ? make_ulonglong2(lo, hi) : make_ulonglong2(lo, lo);
  fill_kernel<<<grid_for(nvec, TPB), TPB, 0, ctx->stream>>>((ulonglong2 *)out->dptr, nvec, v);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_broadcast(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  BcastDesc d;
  d.rank = out->rank;
  for (int k = 0; k < MB_MAX_RANK; k++) d.shape[k] = k < out->rank ? out->shape[k] : 1;
  fill_strides(x, out->rank, out->shape, d.sx);
  for (int k = 0; k < MB_MAX_RANK; k++) d.sy[k] = 0;
  unsigned g = grid_for(out->numel, TPB);
  if (x->dtype == MB_RING128)
    bcast_copy_kernel<u128><<<g, TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (u128 *)out->dptr, out->numel, d);
  else
    bcast_copy_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint64_t *)out->dptr, out->numel, d);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_sum_axis(mb_ctx *ctx, const mb_tensor *x, int64_t outer, int64_t len, int64_t inner, mb_tensor *out) {
  int64_t n = outer * inner;
  if (n == 0) return MB_OK;
  unsigned g = grid_for(n, TPB);
  if (x->dtype == MB_RING128)
    sum_axis_kernel<u128><<<g, TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (u128 *)out->dptr, outer, len, inner);
  else
    sum_axis_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint64_t *)out->dptr, outer, len, inner);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_fixed_encode(mb_ctx *ctx, const mb_tensor *x, double scale, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  unsigned g = grid_for(out->numel, TPB);
  if (out->dtype == MB_RING128)
    fixed_encode_kernel<u128><<<g, TPB, 0, ctx->stream>>>((const double *)x->dptr, (u128 *)out->dptr, out->numel, scale);
  else
    fixed_encode_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>((const double *)x->dptr, (uint64_t *)out->dptr, out->numel, scale);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// raw-pointer form for callers that encode a tensor piece by piece (storage.cu)
int k_fixed_encode_raw(mb_ctx *ctx, const double *x_dev, mb_dtype ring, void *out_dev, int64_t n, double scale) {
  if (n == 0) return MB_OK;
  unsigned g = grid_for(n, TPB);
  if (ring == MB_RING128)
    fixed_encode_kernel<u128><<<g, TPB, 0, ctx->stream>>>(x_dev, (u128 *)out_dev, n, scale);
  else
    fixed_encode_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>(x_dev, (uint64_t *)out_dev, n, scale);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_fixed_decode(mb_ctx *ctx, const mb_tensor *x, double scale, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  unsigned g = grid_for(out->numel, TPB);
  if (x->dtype == MB_RING128)
    fixed_decode_kernel<u128><<<g, TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (double *)out->dptr, out->numel, scale);
  else
    fixed_decode_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (double *)out->dptr, out->numel, scale);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}
