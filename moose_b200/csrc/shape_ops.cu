// Shape / indexing HostPlacement kernels (SURVEY.md 8a row a19 and the remaining rows of the 8b surface):
//   ExpandDimsOp / SqueezeOp           (moose/src/host/ops.rs:1135-1209; RawShape::extend_singletons / squeeze, host/mod.rs:83-113)
//   TransposeOp                         (moose/src/host/ops.rs:1270-1299; bits: bitarray.rs:135-188)
//   ConcatOp                            (moose/src/host/ops.rs:1211-1268)
//   IndexAxisOp / SliceOp / DiagOp / SelectOp (moose/src/host/ops.rs:479-582, 604-783, 785-818; bitarray.rs:85-133)
//   DivOp::ring_kernel                  (moose/src/host/ops.rs:332-343)
//   BitDecomposeOp::host_ring64/128_kernel (ring -> ring planes, moose/src/host/ops.rs:856-894)
//   RingFixedpointMeanOp                (moose/src/host/fixedpoint.rs:41-113)
//   RingFixedpointArgmaxOp              (moose/src/host/ops.rs:2375-2437)
// All of them are copies or one-pass reductions: HBM bound, (bytes in + bytes out) per call.  Everything that is
// a strided view in ndarray is materialised here as a dense row-major tensor (the reference materialises too,
// `to_owned().into_shared()`, except transposes, which it keeps as views until the next op touches them).
#include "common.h"

namespace {
constexpr int TPB = 256;
inline unsigned blocks(int64_t n) { return (unsigned)((n + TPB - 1) / TPB); }

// Iteration domain `shape` (rank <= MB_MAX_RANK); src and dst addressed through strides in elements.
struct CopyDesc {
  int rank;
  int64_t shape[MB_MAX_RANK];
  int64_t sstride[MB_MAX_RANK];
  int64_t dstride[MB_MAX_RANK];
  int64_t soff, doff;
};

template <typename T>
__global__ void __launch_bounds__(TPB) strided_copy_kernel(const T *__restrict__ src, T *__restrict__ dst, CopyDesc d, int64_t n) {
  for (int64_t f = (int64_t)blockIdx.x * TPB + threadIdx.x; f < n; f += (int64_t)gridDim.x * TPB) {
    int64_t rem = f, s = d.soff, o = d.doff;
#pragma unroll
    for (int k = MB_MAX_RANK - 1; k >= 0; k--) {
      if (k < d.rank) {
        const int64_t i = rem % d.shape[k];
        rem /= d.shape[k];
        s += i * d.sstride[k];
        o += i * d.dstride[k];
      }
    }
    dst[o] = src[s];
  }
}

// rank-2 transpose through shared memory: both the reads and the writes are coalesced.
template <typename T>
__global__ void __launch_bounds__(256) transpose2d_kernel(const T *__restrict__ src, T *__restrict__ dst, int64_t rows, int64_t cols) {
  __shared__ T tile[32][33];
  const int64_t c0 = (int64_t)blockIdx.x * 32, r0 = (int64_t)blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int64_t r = r0 + ty + j, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + j][tx] = src[r * cols + c];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int64_t c = c0 + ty + j, r = r0 + tx;
    if (r < rows && c < cols) dst[c * rows + r] = tile[tx][ty + j];
  }
}

// Bit gather: output bit f (row-major over `shape`) = source bit soff + sum idx*sstride.  One thread per output word.
__global__ void __launch_bounds__(TPB) bit_gather_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, CopyDesc d, int64_t nbits,
                                                         int64_t nwords) {
  const int64_t w = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (w >= nwords) return;
  uint32_t word = 0;
  for (int j = 0; j < 32; j++) {
    const int64_t f = w * 32 + j;
    if (f >= nbits) break;
    int64_t rem = f, s = d.soff;
    for (int k = d.rank - 1; k >= 0; k--) {
      const int64_t i = rem % d.shape[k];
      rem /= d.shape[k];
      s += i * d.sstride[k];
    }
    word |= ((src[s >> 5] >> (s & 31)) & 1u) << j;
  }
  dst[w] = word;
}

// dst bits [dbit, dbit + nbits) |= src bits [0, nbits); the destination starts zeroed and the regions of the
// successive calls are disjoint, but they may share a word, hence the atomic.
__global__ void __launch_bounds__(TPB) bit_append_kernel(const uint32_t *__restrict__ src, int64_t src_words, uint32_t *__restrict__ dst, int64_t dbit,
                                                         int64_t nbits) {
  const int64_t w0 = dbit >> 5, w1 = (dbit + nbits - 1) >> 5;
  const int64_t w = w0 + (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (w > w1) return;
  const int64_t lo = max(w * 32, dbit), hi = min(w * 32 + 32, dbit + nbits);
  const int64_t s = lo - dbit;
  const int64_t sw = s >> 5;
  const unsigned so = (unsigned)(s & 31);
  uint64_t two = (uint64_t)src[sw];
  if (sw + 1 < src_words) two |= (uint64_t)src[sw + 1] << 32;
  const unsigned len = (unsigned)(hi - lo);
  uint32_t val = (uint32_t)(two >> so);
  if (len < 32) val &= (1u << len) - 1;
  val <<= (unsigned)(lo - w * 32);
  if (len == 32) dst[w] = val;
  else atomicOr(dst + w, val);
}

template <typename U>
__global__ void __launch_bounds__(TPB) ring_div_kernel(const U *__restrict__ x, const U *__restrict__ y, U *__restrict__ out, int64_t n, int *__restrict__ flag) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e >= n) return;
  const U d = y[e];
  if (d == 0) {
    *flag = 1;
    out[e] = 0;
  } else {
    out[e] = x[e] / d;
  }
}

// plane p, element e -> (x[e] >> p) & 1 as a ring element
template <typename U, int KBITS>
__global__ void __launch_bounds__(TPB) ring_planes_kernel(const U *__restrict__ x, U *__restrict__ out, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e >= n) return;
  const U v = x[e];
#pragma unroll 8
  for (int p = 0; p < KBITS; p++) out[(int64_t)p * n + e] = (v >> p) & (U)1;
}

template <typename U> struct SignedOf;
template <> struct SignedOf<uint64_t> { typedef int64_t type; };
template <> struct SignedOf<u128> { typedef __int128 type; };

// index of the first strict maximum along `len` (signed comparison), x viewed as [outer, len, inner]
template <typename U>
__global__ void __launch_bounds__(TPB) argmax_kernel(const U *__restrict__ x, uint64_t *__restrict__ out, int64_t outer, int64_t len, int64_t inner) {
  typedef typename SignedOf<U>::type S;
  const int64_t t = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (t >= outer * inner) return;
  const int64_t o = t / inner, i = t % inner;
  const U *p = x + o * len * inner + i;
  S best = (S)p[0];
  uint64_t arg = 0;
  for (int64_t j = 1; j < len; j++) {
    const S v = (S)p[j * inner];
    if (best < v) {
      best = v;
      arg = (uint64_t)j;
    }
  }
  out[t] = arg;
}

template <typename U>
__global__ void __launch_bounds__(TPB) scale_kernel(U *__restrict__ x, U w, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e < n) x[e] = x[e] * w;
}

void default_strides(int rank, const int64_t *shape, int64_t *st) {
  int64_t acc = 1;
  for (int k = rank - 1; k >= 0; k--) {
    st[k] = acc;
    acc *= shape[k];
  }
}

int launch_copy(mb_ctx *ctx, mb_dtype dtype, const void *src, void *dst, const CopyDesc &d) {
  int64_t n = 1;
  for (int k = 0; k < d.rank; k++) n *= d.shape[k];
  if (n == 0) return MB_OK;
  const int g = (int)std::min<int64_t>(blocks(n), (int64_t)ctx->sm_count * 16);
  if (dtype == MB_RING128)
    strided_copy_kernel<uint4><<<g, TPB, 0, ctx->stream>>>((const uint4 *)src, (uint4 *)dst, d, n);
  else
    strided_copy_kernel<uint64_t><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)src, (uint64_t *)dst, d, n);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int launch_bit_gather(mb_ctx *ctx, const mb_tensor *x, mb_tensor *t, const CopyDesc &d) {
  if (t->numel == 0) return MB_OK;
  const int64_t nwords = (int64_t)(t->alloc_bytes / 4);
  bit_gather_kernel<<<blocks(nwords), TPB, 0, ctx->stream>>>((const uint32_t *)x->dptr, (uint32_t *)t->dptr, d, t->numel, nwords);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

#define SH_REQUIRE(ctx, cond, code, msg)               \
  do {                                                 \
    if (!(cond)) return mb_fail((ctx), (code), (msg)); \
  } while (0)

#define SH_ENTER(ctx, name)                                                                   \
  do {                                                                                        \
    if (!(ctx)) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, name ": NULL context");      \
    MB_CUDA((ctx), cudaSetDevice((ctx)->device));                                             \
  } while (0)

// finish: hand `t` out on success, free it on failure
int finish(mb_ctx *ctx, int rc, mb_tensor *t, mb_tensor **out) {
  if (rc != MB_OK) {
    mb_tensor_free(ctx, t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

}  // namespace

extern "C" int mb_expand_dims(mb_ctx *ctx, const int64_t *axes, int naxes, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_expand_dims");
  SH_REQUIRE(ctx, x && out && (axes || naxes == 0) && naxes >= 0, MB_ERR_INVALID_ARGUMENT, "mb_expand_dims: NULL argument");
  SH_REQUIRE(ctx, x->rank + naxes <= MB_MAX_RANK, MB_ERR_INVALID_ARGUMENT, "mb_expand_dims: rank too large");
  // extend_singletons: axes in ascending order, each inserted at its position in the shape built so far
  std::vector<int64_t> ax(axes, axes + naxes);
  std::sort(ax.begin(), ax.end());
  std::vector<int64_t> shape(x->shape, x->shape + x->rank);
  for (int64_t a : ax) {
    SH_REQUIRE(ctx, a >= 0 && a <= (int64_t)shape.size(), MB_ERR_KERNEL, "mb_expand_dims: axis out of range");
    shape.insert(shape.begin() + a, 1);
  }
  return mb_reshape(ctx, x, (int)shape.size(), shape.data(), out);
}

extern "C" int mb_squeeze(mb_ctx *ctx, int axis, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_squeeze");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_squeeze: NULL argument");
  std::vector<int64_t> shape;
  if (axis >= 0) {
    SH_REQUIRE(ctx, axis < x->rank, MB_ERR_KERNEL, "mb_squeeze: axis out of range");
    SH_REQUIRE(ctx, x->shape[axis] == 1, MB_ERR_KERNEL,
               "mb_squeeze: cannot select an axis to squeeze out which has size not equal to one");
    for (int k = 0; k < x->rank; k++)
      if (k != axis) shape.push_back(x->shape[k]);
  } else {
    for (int k = 0; k < x->rank; k++)
      if (x->shape[k] != 1) shape.push_back(x->shape[k]);
  }
  return mb_reshape(ctx, x, (int)shape.size(), shape.data(), out);
}

extern "C" int mb_transpose(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_transpose");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_transpose: NULL argument");
  if (x->dtype == MB_BIT)
    SH_REQUIRE(ctx, x->rank <= 3, MB_ERR_KERNEL, "tranposing not implemented for 4D tensors or larger yet");
  int64_t oshape[MB_MAX_RANK], st[MB_MAX_RANK];
  for (int k = 0; k < x->rank; k++) oshape[k] = x->shape[x->rank - 1 - k];
  default_strides(x->rank, x->shape, st);
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, oshape, &t));
  CopyDesc d{};
  d.rank = x->rank;
  for (int k = 0; k < x->rank; k++) {
    d.shape[k] = oshape[k];
    d.sstride[k] = st[x->rank - 1 - k];
  }
  default_strides(x->rank, oshape, d.dstride);
  int rc = MB_OK;
  if (x->dtype == MB_BIT) {
    rc = launch_bit_gather(ctx, x, t, d);
  } else if (x->rank == 2 && x->numel) {
    dim3 grid((unsigned)((x->shape[1] + 31) / 32), (unsigned)((x->shape[0] + 31) / 32));
    if (grid.y > 65535) {
      rc = launch_copy(ctx, x->dtype, x->dptr, t->dptr, d);
    } else {
      if (x->dtype == MB_RING128)
        transpose2d_kernel<uint4><<<grid, 256, 0, ctx->stream>>>((const uint4 *)x->dptr, (uint4 *)t->dptr, x->shape[0], x->shape[1]);
      else
        transpose2d_kernel<uint64_t><<<grid, 256, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint64_t *)t->dptr, x->shape[0], x->shape[1]);
      ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "shape_ops");
      if (cudaGetLastError() != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, "mb_transpose: launch failed");
    }
  } else {
    rc = launch_copy(ctx, x->dtype, x->dptr, t->dptr, d);
  }
  return finish(ctx, rc, t, out);
}

extern "C" int mb_concat(mb_ctx *ctx, int axis, const mb_tensor *const *xs, int n, mb_tensor **out) {
  SH_ENTER(ctx, "mb_concat");
  SH_REQUIRE(ctx, xs && out && n > 0, MB_ERR_INVALID_ARGUMENT, "mb_concat: needs at least one tensor");
  const mb_tensor *x0 = xs[0];
  SH_REQUIRE(ctx, x0 && axis >= 0 && axis < x0->rank, MB_ERR_KERNEL, "mb_concat: axis out of range");
  int64_t oshape[MB_MAX_RANK];
  for (int k = 0; k < x0->rank; k++) oshape[k] = x0->shape[k];
  int64_t total = 0;
  for (int i = 0; i < n; i++) {
    SH_REQUIRE(ctx, xs[i] && xs[i]->dtype == x0->dtype && xs[i]->rank == x0->rank, MB_ERR_KERNEL, "mb_concat: incompatible tensors");
    if (x0->dtype != MB_BIT)
      for (int k = 0; k < x0->rank; k++)
        SH_REQUIRE(ctx, k == axis || xs[i]->shape[k] == x0->shape[k], MB_ERR_KERNEL, "mb_concat: incompatible shapes");
    total += xs[i]->shape[axis];
  }
  oshape[axis] = total;
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x0->dtype, x0->rank, oshape, &t));
  int rc = MB_OK;
  if (x0->dtype == MB_BIT) {
    // ConcatOp::bit_kernel appends the flat bit streams whatever the axis is and only fixes up the shape
    // (host/ops.rs:1247-1266); restated as is.
    int64_t have = 0;
    for (int i = 0; i < n; i++) have += xs[i]->numel;
    if (have != t->numel) {
      mb_tensor_free(ctx, t);
      return mb_fail(ctx, MB_ERR_KERNEL, "mb_concat: bit tensors do not fill the concatenated shape");
    }
    if (t->alloc_bytes) {
      cudaError_t e = cudaMemsetAsync(t->dptr, 0, t->alloc_bytes, ctx->stream);
      if (e != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, cudaGetErrorString(e));
    }
    int64_t pos = 0;
    for (int i = 0; i < n && rc == MB_OK; i++) {
      const int64_t nb = xs[i]->numel;
      if (nb == 0) continue;
      const int64_t words = ((pos + nb - 1) >> 5) - (pos >> 5) + 1;
      bit_append_kernel<<<blocks(words), TPB, 0, ctx->stream>>>((const uint32_t *)xs[i]->dptr, (int64_t)(xs[i]->alloc_bytes / 4), (uint32_t *)t->dptr, pos, nb);
      ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "shape_ops");
      if (cudaGetLastError() != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, "mb_concat: launch failed");
      pos += nb;
    }
    return finish(ctx, rc, t, out);
  }
  int64_t at = 0;
  for (int i = 0; i < n && rc == MB_OK; i++) {
    CopyDesc d{};
    d.rank = x0->rank;
    for (int k = 0; k < d.rank; k++) d.shape[k] = xs[i]->shape[k];
    default_strides(d.rank, xs[i]->shape, d.sstride);
    default_strides(d.rank, oshape, d.dstride);
    d.doff = at * d.dstride[axis];
    rc = launch_copy(ctx, x0->dtype, xs[i]->dptr, t->dptr, d);
    at += xs[i]->shape[axis];
  }
  return finish(ctx, rc, t, out);
}

extern "C" int mb_index_axis(mb_ctx *ctx, int axis, int64_t index, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_index_axis");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_index_axis: NULL argument");
  SH_REQUIRE(ctx, axis >= 0 && axis < x->rank, MB_ERR_INVALID_ARGUMENT, "axis too large in index axis");
  SH_REQUIRE(ctx, index >= 0 && index < x->shape[axis], MB_ERR_INVALID_ARGUMENT, "index too large in index axis");
  if (axis == 0) return mb_index_axis0(ctx, index, x, out);
  int64_t oshape[MB_MAX_RANK], st[MB_MAX_RANK];
  default_strides(x->rank, x->shape, st);
  int orank = 0;
  CopyDesc d{};
  for (int k = 0; k < x->rank; k++) {
    if (k == axis) continue;
    oshape[orank] = x->shape[k];
    d.shape[orank] = x->shape[k];
    d.sstride[orank] = st[k];
    orank++;
  }
  d.rank = orank;
  d.soff = index * st[axis];
  default_strides(orank, oshape, d.dstride);
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, orank, oshape, &t));
  int rc;
  if (x->dtype == MB_BIT) {
    // BitArrayRepr::index_axis copies new_dim.size() CONSECUTIVE bits starting at index * stride[axis]
    // (bitarray.rs:85-108), which is the selected sub-array only for axis 0 or unit leading axes; restated as is.
    rc = t->numel ? k_bit_shift_copy(ctx, x, d.soff, 0, std::min<int64_t>(t->numel, x->numel - d.soff), t) : MB_OK;
  } else {
    rc = launch_copy(ctx, x->dtype, x->dptr, t->dptr, d);
  }
  return finish(ctx, rc, t, out);
}

extern "C" int mb_slice(mb_ctx *ctx, int n, const int64_t *start, const int64_t *end, const uint8_t *has_end, const int64_t *step,
                        const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_slice");
  SH_REQUIRE(ctx, x && out && (n == 0 || (start && end && has_end && step)), MB_ERR_INVALID_ARGUMENT, "mb_slice: NULL argument");
  SH_REQUIRE(ctx, n == x->rank, MB_ERR_INVALID_ARGUMENT, "The input dimension of `info` must match the array to be sliced.");
  SH_REQUIRE(ctx, x->dtype != MB_BIT, MB_ERR_KERNEL, "slicing not implemented for BitArray yet");
  int64_t st[MB_MAX_RANK], oshape[MB_MAX_RANK];
  default_strides(x->rank, x->shape, st);
  CopyDesc d{};
  d.rank = x->rank;
  for (int k = 0; k < n; k++) {
    // ndarray Slice: negative start / end count from the back, end None = len, both must land in [0, len];
    // a negative step walks the selected range backwards from its last element
    const int64_t len = x->shape[k];
    int64_t s = start[k] < 0 ? start[k] + len : start[k];
    int64_t e = has_end[k] ? (end[k] < 0 ? end[k] + len : end[k]) : len;
    const int64_t stp = step[k];
    SH_REQUIRE(ctx, stp != 0, MB_ERR_KERNEL, "mb_slice: step must be nonzero");
    SH_REQUIRE(ctx, s >= 0 && s <= len && e >= 0 && e <= len, MB_ERR_KERNEL, "mb_slice: slice out of bounds");
    if (e < s) e = s;
    const int64_t m = e - s, a = stp < 0 ? -stp : stp;
    const int64_t cnt = (m + a - 1) / a;
    oshape[k] = cnt;
    d.shape[k] = cnt;
    if (stp > 0) {
      d.soff += s * st[k];
      d.sstride[k] = stp * st[k];
    } else {
      d.soff += (cnt ? e - 1 : s) * st[k];
      d.sstride[k] = stp * st[k];
    }
  }
  default_strides(x->rank, oshape, d.dstride);
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, oshape, &t));
  return finish(ctx, launch_copy(ctx, x->dtype, x->dptr, t->dptr, d), t, out);
}

extern "C" int mb_diag(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_diag");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_diag: NULL argument");
  int64_t len = x->rank ? x->shape[0] : 1, st[MB_MAX_RANK];
  for (int k = 1; k < x->rank; k++) len = std::min(len, x->shape[k]);
  default_strides(x->rank, x->shape, st);
  CopyDesc d{};
  d.rank = 1;
  d.shape[0] = len;
  d.dstride[0] = 1;
  for (int k = 0; k < x->rank; k++) d.sstride[0] += st[k];
  if (x->dtype == MB_BIT) {
    // BitArrayRepr::into_diag handles lengths 1 and 2 only; the second bit sits at stride_offset([1, 1]),
    // which pairs the index with the first two strides (bitarray.rs:110-128)
    SH_REQUIRE(ctx, len == 1 || len == 2, MB_ERR_UNIMPLEMENTED, "mb_diag: bit diagonals longer than 2 are not implemented in Moose");
    d.sstride[0] = 0;
    for (int k = 0; k < std::min(2, x->rank); k++) d.sstride[0] += st[k];
  }
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, 1, &len, &t));
  const int rc = x->dtype == MB_BIT ? launch_bit_gather(ctx, x, t, d) : launch_copy(ctx, x->dtype, x->dptr, t->dptr, d);
  return finish(ctx, rc, t, out);
}

extern "C" int mb_select(mb_ctx *ctx, int axis, const mb_tensor *index_bits, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_select");
  SH_REQUIRE(ctx, x && index_bits && out, MB_ERR_INVALID_ARGUMENT, "mb_select: NULL argument");
  SH_REQUIRE(ctx, index_bits->dtype == MB_BIT && x->dtype != MB_BIT, MB_ERR_TYPE_MISMATCH, "mb_select: (bit index, ring tensor)");
  SH_REQUIRE(ctx, axis >= 0 && axis < x->rank, MB_ERR_INVALID_ARGUMENT, "axis too large in index axis");
  // the output extent depends on the index bits: read them back (dim[axis] bits)
  std::vector<uint8_t> host((size_t)index_bits->nbytes);
  if (!host.empty()) {
    MB_CUDA(ctx, cudaMemcpyAsync(host.data(), index_bits->dptr, host.size(), cudaMemcpyDeviceToHost, ctx->stream));
    MB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  std::vector<int64_t> picks;
  for (int64_t i = 0; i < index_bits->numel; i++)
    if ((host[(size_t)(i >> 3)] >> (i & 7)) & 1) picks.push_back(i);
  for (int64_t p : picks) SH_REQUIRE(ctx, p < x->shape[axis], MB_ERR_KERNEL, "mb_select: index out of bounds");
  int64_t oshape[MB_MAX_RANK], st[MB_MAX_RANK];
  for (int k = 0; k < x->rank; k++) oshape[k] = x->shape[k];
  oshape[axis] = (int64_t)picks.size();
  default_strides(x->rank, x->shape, st);
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank, oshape, &t));
  // runs of consecutive picks become one strided copy each
  int rc = MB_OK;
  size_t i = 0;
  int64_t at = 0;
  while (i < picks.size() && rc == MB_OK) {
    size_t j = i + 1;
    while (j < picks.size() && picks[j] == picks[j - 1] + 1) j++;
    CopyDesc d{};
    d.rank = x->rank;
    for (int k = 0; k < d.rank; k++) d.shape[k] = x->shape[k];
    d.shape[axis] = (int64_t)(j - i);
    for (int k = 0; k < d.rank; k++) d.sstride[k] = st[k];
    default_strides(d.rank, oshape, d.dstride);
    d.soff = picks[i] * st[axis];
    d.doff = at * d.dstride[axis];
    rc = launch_copy(ctx, x->dtype, x->dptr, t->dptr, d);
    at += (int64_t)(j - i);
    i = j;
  }
  return finish(ctx, rc, t, out);
}

extern "C" int mb_div(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out) {
  SH_ENTER(ctx, "mb_div");
  SH_REQUIRE(ctx, x && y && out, MB_ERR_INVALID_ARGUMENT, "mb_div: NULL argument");
  SH_REQUIRE(ctx, mb_is_ring(x->dtype) && x->dtype == y->dtype, MB_ERR_TYPE_MISMATCH, "mb_div: needs ring tensors of one width");
  const mb_tensor *xs = x, *ys = y;
  mb_tensor *xb = nullptr, *yb = nullptr, *probe = nullptr;
  if (!mb_same_shape(x, y)) {  // ndarray co-broadcasting
    MB_CHECK(mb_sub(ctx, x, y, &probe));
    int rc = mb_broadcast(ctx, x, probe->rank, probe->shape, &xb);
    if (rc == MB_OK) rc = mb_broadcast(ctx, y, probe->rank, probe->shape, &yb);
    if (rc != MB_OK) {
      mb_tensor_free(ctx, probe);
      mb_tensor_free(ctx, xb);
      return rc;
    }
    xs = xb;
    ys = yb;
  }
  mb_tensor *t = nullptr;
  int rc = mb_tensor_new(ctx, xs->dtype, xs->rank, xs->shape, &t);
  if (rc == MB_OK && t->numel) {
    // Wrapping<T> / Wrapping(0) panics in the reference; here the call fails with MB_ERR_KERNEL
    void *flagp = nullptr;
    rc = mb_scratch(ctx, 2, 256, &flagp);
    int flag = 0;
    if (rc == MB_OK) {
      cudaMemsetAsync(flagp, 0, sizeof(int), ctx->stream);
      if (xs->dtype == MB_RING128)
        ring_div_kernel<u128><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const u128 *)xs->dptr, (const u128 *)ys->dptr, (u128 *)t->dptr, t->numel, (int *)flagp);
      else
        ring_div_kernel<uint64_t><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const uint64_t *)xs->dptr, (const uint64_t *)ys->dptr, (uint64_t *)t->dptr, t->numel, (int *)flagp);
      ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "shape_ops");
      cudaError_t e = cudaMemcpyAsync(&flag, flagp, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, cudaGetErrorString(e));
      else if (flag) rc = mb_fail(ctx, MB_ERR_KERNEL, "mb_div: attempt to divide by zero");
    }
  }
  mb_tensor_free(ctx, probe);
  mb_tensor_free(ctx, xb);
  mb_tensor_free(ctx, yb);
  return finish(ctx, rc, t, out);
}

extern "C" int mb_bit_decompose_ring(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_bit_decompose_ring");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_bit_decompose_ring: NULL argument");
  SH_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_bit_decompose_ring: needs a ring tensor");
  SH_REQUIRE(ctx, x->rank + 1 <= MB_MAX_RANK, MB_ERR_INVALID_ARGUMENT, "mb_bit_decompose_ring: rank too large");
  int64_t oshape[MB_MAX_RANK];
  oshape[0] = x->dtype == MB_RING128 ? 128 : 64;
  for (int k = 0; k < x->rank; k++) oshape[k + 1] = x->shape[k];
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, x->rank + 1, oshape, &t));
  int rc = MB_OK;
  if (x->numel) {
    if (x->dtype == MB_RING128)
      ring_planes_kernel<u128, 128><<<blocks(x->numel), TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (u128 *)t->dptr, x->numel);
    else
      ring_planes_kernel<uint64_t, 64><<<blocks(x->numel), TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint64_t *)t->dptr, x->numel);
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "shape_ops");
    if (cudaGetLastError() != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, "mb_bit_decompose_ring: launch failed");
  }
  return finish(ctx, rc, t, out);
}

extern "C" int mb_fixedpoint_mean(mb_ctx *ctx, int axis, uint64_t scaling_base, uint32_t scaling_exp, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_fixedpoint_mean");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_mean: NULL argument");
  SH_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_fixedpoint_mean: needs a ring tensor");
  SH_REQUIRE(ctx, axis < x->rank, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_mean: axis out of range");
  // weight = encode(1 / len, base^exp): ((1/len) * (scale as f64)) as i64 / i128, truncation toward zero
  const double len = axis >= 0 ? (double)x->shape[axis] : (double)x->numel;
  u128 scale = 1;
  for (uint32_t i = 0; i < scaling_exp; i++) scale *= (x->dtype == MB_RING128 ? (u128)scaling_base : (u128)(uint64_t)scaling_base);
  if (x->dtype == MB_RING64) scale = (u128)(uint64_t)scale;  // u64::pow wraps in release builds
  const double up = (1.0 / len) * (x->dtype == MB_RING128 ? (double)scale : (double)(uint64_t)scale);
  mb_tensor *t = nullptr;
  MB_CHECK(mb_sum(ctx, x, axis, &t));
  int rc = MB_OK;
  if (t->numel) {
    if (x->dtype == MB_RING128) {
      u128 w;
      if (std::isnan(up)) w = 0;
      else if (up >= 0x1p127) w = (u128)(~(u128)0 >> 1);
      else if (up <= -0x1p127) w = (u128)1 << 127;
      else w = (u128)(__int128)up;
      scale_kernel<u128><<<blocks(t->numel), TPB, 0, ctx->stream>>>((u128 *)t->dptr, w, t->numel);
    } else {
      uint64_t w;
      if (std::isnan(up)) w = 0;
      else if (up >= 0x1p63) w = (uint64_t)INT64_MAX;
      else if (up <= -0x1p63) w = (uint64_t)1 << 63;
      else w = (uint64_t)(int64_t)up;
      scale_kernel<uint64_t><<<blocks(t->numel), TPB, 0, ctx->stream>>>((uint64_t *)t->dptr, w, t->numel);
    }
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "shape_ops");
    if (cudaGetLastError() != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, "mb_fixedpoint_mean: launch failed");
  }
  return finish(ctx, rc, t, out);
}

extern "C" int mb_fixedpoint_argmax(mb_ctx *ctx, int axis, const mb_tensor *x, mb_tensor **out) {
  SH_ENTER(ctx, "mb_fixedpoint_argmax");
  SH_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_argmax: NULL argument");
  SH_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_fixedpoint_argmax: needs a ring tensor");
  SH_REQUIRE(ctx, axis >= 0 && axis < x->rank, MB_ERR_INVALID_ARGUMENT, "mb_fixedpoint_argmax: axis out of range");
  SH_REQUIRE(ctx, x->shape[axis] > 0, MB_ERR_KERNEL, "mb_fixedpoint_argmax: empty axis");
  int64_t outer = 1, inner = 1, oshape[MB_MAX_RANK];
  int orank = 0;
  for (int k = 0; k < x->rank; k++) {
    if (k < axis) outer *= x->shape[k];
    if (k > axis) inner *= x->shape[k];
    if (k != axis) oshape[orank++] = x->shape[k];
  }
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, MB_RING64, orank, oshape, &t));
  int rc = MB_OK;
  if (t->numel) {
    if (x->dtype == MB_RING128)
      argmax_kernel<u128><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (uint64_t *)t->dptr, outer, x->shape[axis], inner);
    else
      argmax_kernel<uint64_t><<<blocks(t->numel), TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint64_t *)t->dptr, outer, x->shape[axis], inner);
    ctx->launches.fetch_add(1); if (ctx->trace) mb_trace_mark(ctx, "shape_ops");
    if (cudaGetLastError() != cudaSuccess) rc = mb_fail(ctx, MB_ERR_CUDA, "mb_fixedpoint_argmax: launch failed");
  }
  return finish(ctx, rc, t, out);
}
