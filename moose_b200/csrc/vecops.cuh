// 16-byte "vector" arithmetic shared by the fused kernels.  A vector is one
// Ring128 element, two Ring64 elements or 128 packed bits -- and in every case
// exactly one AES block of each PRF stream (see prf_vec): vector index v of a
// tensor corresponds to block v of the aes-prng stream.
#pragma once
#include "aes.cuh"

enum VecKind { VK_RING64 = 0, VK_RING128 = 1, VK_BIT = 2 };

typedef ulonglong2 vec_t;

__device__ __forceinline__ u128 v128(vec_t v) { return ((u128)v.y << 64) | (u128)v.x; }
__device__ __forceinline__ vec_t vfrom(u128 v) { return make_ulonglong2((uint64_t)v, (uint64_t)(v >> 64)); }

template <int K>
struct VOps {
  static __device__ __forceinline__ vec_t zero() { return make_ulonglong2(0, 0); }
  static __device__ __forceinline__ vec_t add(vec_t a, vec_t b) {
    if (K == VK_BIT) return make_ulonglong2(a.x ^ b.x, a.y ^ b.y);
    if (K == VK_RING128) return vfrom(v128(a) + v128(b));
    return make_ulonglong2(a.x + b.x, a.y + b.y);
  }
  static __device__ __forceinline__ vec_t sub(vec_t a, vec_t b) {
    if (K == VK_BIT) return make_ulonglong2(a.x ^ b.x, a.y ^ b.y);
    if (K == VK_RING128) return vfrom(v128(a) - v128(b));
    return make_ulonglong2(a.x - b.x, a.y - b.y);
  }
  static __device__ __forceinline__ vec_t mul(vec_t a, vec_t b) {
    if (K == VK_BIT) return make_ulonglong2(a.x & b.x, a.y & b.y);
    if (K == VK_RING128) return vfrom(v128(a) * v128(b));
    return make_ulonglong2(a.x * b.x, a.y * b.y);
  }
  static __device__ __forceinline__ vec_t neg(vec_t a) { return sub(zero(), a); }
  // logical shifts with 0 <= s < width (already masked like Rust's Wrapping shifts)
  static __device__ __forceinline__ vec_t shl(vec_t a, unsigned s) {
    if (K == VK_RING128) return vfrom(v128(a) << s);
    return make_ulonglong2(a.x << s, a.y << s);
  }
  static __device__ __forceinline__ vec_t shr(vec_t a, unsigned s) {
    if (K == VK_RING128) return vfrom(v128(a) >> s);
    return make_ulonglong2(a.x >> s, a.y >> s);
  }
  static __device__ __forceinline__ vec_t splat(uint64_t lo, uint64_t hi) {
    if (K == VK_RING128) return make_ulonglong2(lo, hi);
    return make_ulonglong2(lo, lo);
  }
  // a * b where every element of b is 0 or 1
  static __device__ __forceinline__ vec_t mul01(vec_t a, vec_t b) {
    if (K == VK_RING128) return (b.x & 1) ? a : zero();
    return make_ulonglong2((b.x & 1) ? a.x : 0, (b.y & 1) ? a.y : 0);
  }
};

// PRF vector v of the stream keyed by `k`, in the element layout of kind K:
//  Ring128: element v = (u64 at bytes 0..7) << 64 + (u64 at bytes 8..15)   (host/ops.rs:1998-2000)
//  Ring64 : elements 2v, 2v+1 = the two little-endian u64 halves           (host/ops.rs:1969)
//  Bit    : bits 128v .. 128v+127 = keystream bits, LSB-first per byte     (get_bit policy 0)
template <int K>
__device__ __forceinline__ vec_t prf_vec(const AesLane &sTl, const AesKey &k, uint64_t v) {
  uint32_t o[4];
  aes_ctr_block(sTl, k, v, o);
  uint64_t first = (uint64_t)o[0] | ((uint64_t)o[1] << 32);
  uint64_t second = (uint64_t)o[2] | ((uint64_t)o[3] << 32);
  if (K == VK_RING128) return make_ulonglong2(second, first);
  return make_ulonglong2(first, second);
}

// Packed-bit invariant: bits at index >= numel are zero.
__device__ __forceinline__ vec_t bit_tail_mask(vec_t a, int64_t v, int64_t numel) {
  int64_t lo = v * 128;
  if (lo + 128 <= numel) return a;
  int64_t valid = numel - lo;  // < 128
  if (valid <= 0) return make_ulonglong2(0, 0);
  if (valid >= 64) {
    unsigned r = (unsigned)(valid - 64);
    a.y = r ? (a.y & ((1ull << r) - 1)) : 0;
  } else {
    a.y = 0;
    a.x = valid ? (a.x & ((1ull << valid) - 1)) : 0;
  }
  return a;
}

static inline VecKind vec_kind(mb_dtype d) { return d == MB_RING128 ? VK_RING128 : d == MB_BIT ? VK_BIT : VK_RING64; }
static inline int64_t vec_count(const mb_tensor *t) { return (int64_t)((t->nbytes + 15) / 16); }
