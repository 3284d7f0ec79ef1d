// Host-placement kernels on packed bit tensors.
// Bit tensors are flat, row-major, LSB-first packed (BitVec<u8, Lsb0>,
// moose/src/host/bitarray.rs:10-13); allocations are padded to 16 bytes and the
// bits past numel are always zero.
#include "common.h"

namespace {

constexpr int TPB = 256;

template <typename T>
__device__ __forceinline__ unsigned bit_of(T x, unsigned p) {
  return (unsigned)((x >> p) & 1);
}

// 32 x 32 bit-matrix transpose across the lanes of a warp (LSB-first): lane r holds row r on entry, lane c
// holds column c on exit.
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, int lane) {
#pragma unroll
  for (int s = 0; s < 5; s++) {
    const int j = 16 >> s;
    const uint32_t m = s == 0 ? 0x0000FFFFu : s == 1 ? 0x00FF00FFu : s == 2 ? 0x0F0F0F0Fu : s == 3 ? 0x33333333u : 0x55555555u;
    const uint32_t p = __shfl_xor_sync(0xffffffffu, x, j);
    x = (lane & j) ? ((x & ~m) | ((p >> j) & m)) : ((x & m) | ((p & m) << j));
  }
  return x;
}

// BitDecomposeOp::host_bit64/128_kernel (host/ops.rs:896-936), n % 128 == 0:
// a warp transposes 128 elements x KBITS planes (32 x 32 bit blocks, shuffle butterflies); the eight warps of a
// CTA take eight consecutive groups and exchange through shared memory so that each plane gets one full 128-byte
// line per store.
template <typename T, int KBITS>
__global__ void __launch_bounds__(TPB) bit_decompose_aligned_kernel(const T *__restrict__ x, uint32_t *__restrict__ out,
                                                                   int64_t n) {
  constexpr int WARPS = TPB / 32;
  constexpr int PL = KBITS / 32;  // planes kept per lane
  __shared__ uint4 tile[KBITS][WARPS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t ngroups = n / 128;
  const int64_t nchunks = (ngroups + WARPS - 1) / WARPS;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int64_t g0 = c * WARPS, grp = g0 + w;
    if (grp < ngroups) {
      uint32_t keep[PL][4];
#pragma unroll
      for (int it = 0; it < 4; it++) {
        T v = x[grp * 128 + it * 32 + lane];
#pragma unroll
        for (int h = 0; h < PL; h++) keep[h][it] = warp_transpose32((uint32_t)(v >> (32 * h)), lane);
      }
#pragma unroll
      for (int q = 0; q < PL; q++) tile[q * 32 + lane][w] = make_uint4(keep[q][0], keep[q][1], keep[q][2], keep[q][3]);
    }
    __syncthreads();
    uint4 *o = (uint4 *)out;
    for (int idx = threadIdx.x; idx < KBITS * WARPS; idx += TPB) {
      const int p = idx / WARPS, j = idx % WARPS;
      if (g0 + j < ngroups) o[(int64_t)p * ngroups + g0 + j] = tile[p][j];
    }
    __syncthreads();
  }
}

// Generic fallback: one thread per output word.
template <typename T, int KBITS>
__global__ void __launch_bounds__(TPB) bit_decompose_generic_kernel(const T *__restrict__ x, uint32_t *__restrict__ out,
                                                                   int64_t n, int64_t nwords) {
  int64_t w = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (w >= nwords) return;
  uint32_t word = 0;
  int64_t total = (int64_t)KBITS * n;
  for (int j = 0; j < 32; j++) {
    int64_t idx = w * 32 + j;
    if (idx < total) {
      int64_t p = idx / n, e = idx % n;
      word |= bit_of(x[e], (unsigned)p) << j;
    }
  }
  out[w] = word;
}

// BitExtractOp::kernel64/128 (host/ops.rs:1567-1595).
template <typename T>
__global__ void __launch_bounds__(TPB) bit_extract_kernel(const T *__restrict__ x, uint32_t *__restrict__ out, int64_t n,
                                                          unsigned bit_idx) {
  int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  unsigned b = e < n ? bit_of(x[e], bit_idx) : 0u;
  unsigned word = __ballot_sync(0xffffffffu, b);
  if ((threadIdx.x & 31) == 0 && (e & ~31ll) < n) out[e >> 5] = word;
}

// RingInjectOp::host_kernel (host/ops.rs:1597-1621): bit b -> b << bit_idx.
template <typename T>
__global__ void __launch_bounds__(TPB) ring_inject_kernel(const uint32_t *__restrict__ bits, T *__restrict__ out, int64_t n,
                                                          unsigned bit_idx) {
  int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (e >= n) return;
  uint32_t w = bits[e >> 5];
  out[e] = (T)((w >> (e & 31)) & 1u) << bit_idx;
}

// out bits [dst, dst+nbits) = x bits [src, src+nbits), zero elsewhere.  Used by
// ShlDimOp::host_bit_kernel (host/ops.rs:835-853) and IndexAxisOp on bit tensors
// (bitarray.rs:85-108).
__global__ void __launch_bounds__(TPB) bit_shift_copy_kernel(const uint32_t *__restrict__ x, int64_t x_words,
                                                             uint32_t *__restrict__ out, int64_t out_words, int64_t src,
                                                             int64_t dst, int64_t nbits) {
  int64_t w = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (w >= out_words) return;
  int64_t lo = max(w * 32, dst), hi = min(w * 32 + 32, dst + nbits);
  uint32_t word = 0;
  if (lo < hi) {
    int64_t s = lo - dst + src;
    int64_t sw = s >> 5;
    unsigned so = (unsigned)(s & 31);
    uint64_t two = (uint64_t)x[sw];
    if (sw + 1 < x_words) two |= (uint64_t)x[sw + 1] << 32;
    unsigned len = (unsigned)(hi - lo);
    uint32_t val = (uint32_t)(two >> so);
    if (len < 32) val &= (1u << len) - 1;
    word = val << (unsigned)(lo - w * 32);
  }
  out[w] = word;
}

__global__ void __launch_bounds__(TPB) bit_not_kernel(const uint32_t *__restrict__ x, uint32_t *__restrict__ out, int64_t nwords,
                                                      int64_t numel) {
  int64_t w = (int64_t)blockIdx.x * TPB + threadIdx.x;
  if (w >= nwords) return;
  uint32_t v = ~x[w];
  int64_t lo = w * 32;
  if (lo >= numel) v = 0;
  else if (lo + 32 > numel) v &= (1u << (unsigned)(numel - lo)) - 1;
  out[w] = v;
}

inline unsigned blocks(int64_t n) { return (unsigned)((n + TPB - 1) / TPB); }

}  // namespace

int k_bit_decompose(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out) {
  int64_t n = x->numel;
  if (n == 0) return MB_OK;
  uint32_t *o = (uint32_t *)out->dptr;
  if (n % 128 == 0) {
    int64_t warps = n / 128;
    int g = (int)std::max<int64_t>(1, std::min<int64_t>((warps * 32 + TPB - 1) / TPB, (int64_t)ctx->sm_count * 8));
    if (x->dtype == MB_RING128)
      bit_decompose_aligned_kernel<u128, 128><<<g, TPB, 0, ctx->stream>>>((const u128 *)x->dptr, o, n);
    else
      bit_decompose_aligned_kernel<uint64_t, 64><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, o, n);
  } else {
    int64_t nwords = (int64_t)(out->alloc_bytes / 4);
    if (x->dtype == MB_RING128)
      bit_decompose_generic_kernel<u128, 128><<<blocks(nwords), TPB, 0, ctx->stream>>>((const u128 *)x->dptr, o, n, nwords);
    else
      bit_decompose_generic_kernel<uint64_t, 64><<<blocks(nwords), TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, o, n, nwords);
  }
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_bit_extract(mb_ctx *ctx, unsigned bit_idx, const mb_tensor *x, mb_tensor *out) {
  int64_t n = x->numel;
  if (n == 0) return MB_OK;
  // zero the padded tail first (the kernel writes whole words only where elements exist)
  MB_CUDA(ctx, cudaMemsetAsync(out->dptr, 0, out->alloc_bytes, ctx->stream));
  int64_t threads = ((n + 31) / 32) * 32;
  if (x->dtype == MB_RING128)
    bit_extract_kernel<u128><<<blocks(threads), TPB, 0, ctx->stream>>>((const u128 *)x->dptr, (uint32_t *)out->dptr, n, bit_idx & 127);
  else
    bit_extract_kernel<uint64_t><<<blocks(threads), TPB, 0, ctx->stream>>>((const uint64_t *)x->dptr, (uint32_t *)out->dptr, n, bit_idx & 63);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_ring_inject(mb_ctx *ctx, unsigned bit_idx, const mb_tensor *bits, mb_tensor *out) {
  int64_t n = out->numel;
  if (n == 0) return MB_OK;
  if (out->dtype == MB_RING128)
    ring_inject_kernel<u128><<<blocks(n), TPB, 0, ctx->stream>>>((const uint32_t *)bits->dptr, (u128 *)out->dptr, n, bit_idx & 127);
  else
    ring_inject_kernel<uint64_t><<<blocks(n), TPB, 0, ctx->stream>>>((const uint32_t *)bits->dptr, (uint64_t *)out->dptr, n, bit_idx & 63);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_bit_shift_copy(mb_ctx *ctx, const mb_tensor *x, int64_t src_bit, int64_t dst_bit, int64_t nbits, mb_tensor *out) {
  int64_t out_words = (int64_t)(out->alloc_bytes / 4);
  int64_t x_words = (int64_t)(x->alloc_bytes / 4);
  bit_shift_copy_kernel<<<blocks(out_words), TPB, 0, ctx->stream>>>((const uint32_t *)x->dptr, x_words, (uint32_t *)out->dptr,
                                                                   out_words, src_bit, dst_bit, nbits);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_bit_not(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out) {
  int64_t nwords = (int64_t)(out->alloc_bytes / 4);
  bit_not_kernel<<<blocks(nwords), TPB, 0, ctx->stream>>>((const uint32_t *)x->dptr, (uint32_t *)out->dptr, nwords, x->numel);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}
