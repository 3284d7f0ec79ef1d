// Device-side AES-128 for the aes-prng compatible PRF stream:
//   block(i) = AES-128_seed(LE128(i)),   i = 0, 1, 2, ...
// (restatement pinned by moose/src/host/mod.rs:1050-1079, see oracle/).
//
// The PRF kernels are issue-bound (every SMSP issues one instruction per clock and an AES block is
// several hundred of them), so the formulation minimises INSTRUCTIONS per block, not memory traffic:
//  * T-table rounds with two tables, Te0 and Te2 = rotl16(Te0), interleaved in shared memory as
//    [256][64] words: row idx holds 32 copies of Te0[idx] then 32 copies of Te2[idx].  Lane l reads
//    column l (or 32 + l), so the data-dependent lookups never bank-conflict, and -- the row stride
//    being exactly 256 bytes and the table 64 KB aligned -- the ADDRESS of a lookup,
//    0x10000 | (byte_k(x) << 8) | lane*4 [| 128], is ONE PRMT of the state word with a per-lane constant.
//    Lookups of byte positions 1 and 3 take rotl8 of the Te0 / Te2 value (one more PRMT); positions 0 and 2
//    need no rotation.  What remains is bound by the shared-memory pipe: 148 lookups per block at one
//    32-lane LDS per clock per SM (measured 59 G blocks/s = 0.94 of that bound at 1.97 GHz).
//  * Round 1: the counter block is LE128(ctr) with ctr < 2^32 for every tensor below 64 GB, so only
//    the first state word depends on ctr; the 12 constant lookups of round 1 are folded on the host into
//    four words per key (AesKey::r1c).  Larger counters take the generic path.
//  * Last round: S[x] is byte 0 / 3 of Te2[x] and byte 1 / 2 of Te0[x]; three PRMTs assemble a column.
// Cost: the CTA's shared-memory window up to 0x20000 (one CTA per SM).
#pragma once
#include <cstdint>

#include "common.h"

#define AES_TABLE_WORDS (256 * 64)
// dynamic shared memory to request: everything up to window address 0x20000 (the window starts with the
// driver-reserved 1 KB; static shared memory of a kernel only moves the dynamic base further up)
#define AES_SMEM_BYTES (0x20000 - 1024)
// block sizes by register budget: <= 64, <= 128, <= 255 registers per thread
#define AES_TPB_LIGHT 1024
#ifndef AES_TPB_MID
#define AES_TPB_MID 512
#endif
#ifndef AES_TPB_384
#define AES_TPB_384 384  /* <= 170 registers */
#endif
#define AES_TPB_HEAVY 256

// Te0[x] = {2*S[x], S[x], S[x], 3*S[x]} as a little-endian word (one copy per translation unit,
// read once per CTA when the shared-memory table is built).
static __device__ const uint32_t c_Te0[256] = {
#include "aes_te0.inc"
};

// The table sits at the ABSOLUTE shared-memory address 0x10000 of the CTA's window: the 64 KB alignment lets the
// per-lane PRMT constant carry the base (byte 2 = 0x01), so the PRMT result IS the address of the lookup.  The
// dynamic allocation therefore reaches up to 0x20000 (one CTA per SM; block sizes 256..1024 by register need).
#define AES_TABLE_ADDR 0x10000u
__device__ __forceinline__ uint32_t aes_smem_base(const uint32_t *sT) { return (uint32_t)__cvta_generic_to_shared(sT); }
__device__ __forceinline__ void aes_load_table(uint32_t *sT) {
  uint32_t *tab = sT + (AES_TABLE_ADDR - aes_smem_base(sT)) / 4;
  for (int i = threadIdx.x; i < AES_TABLE_WORDS; i += blockDim.x) {
    uint32_t t = c_Te0[i >> 6];
    tab[i] = (i & 32) ? __byte_perm(t, t, 0x1032) : t;
  }
  __syncthreads();
}
struct AesLane {
  uint32_t l0, l2;  // (0x01 << 8) | lane*4 (Te0 half) and the same + 128 (Te2 half): bytes 0 and 2 of the address
};
__device__ __forceinline__ AesLane aes_lane(const uint32_t *) {
  AesLane a;
  a.l0 = (threadIdx.x & 31) * 4 + 0x100;
  a.l2 = a.l0 + 128;
  return a;
}
__device__ __forceinline__ uint32_t aes_lds(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
template <int K>
__device__ __forceinline__ uint32_t te0(const AesLane &a, uint32_t x) {
  return aes_lds(__byte_perm(x, a.l0, 0x6504 | (K << 4)));
}
template <int K>
__device__ __forceinline__ uint32_t te2(const AesLane &a, uint32_t x) {
  return aes_lds(__byte_perm(x, a.l2, 0x6504 | (K << 4)));
}
__device__ __forceinline__ uint32_t rotl8(uint32_t x) { return __byte_perm(x, x, 0x2103); }

// Encrypts the counter block LE128(ctr) (ctr < 2^64).  out[0..3] are the 16 keystream bytes as
// little-endian words.
__device__ __forceinline__ void aes_ctr_block(const AesLane &a, const AesKey &k, uint64_t ctr, uint32_t out[4]) {
  uint32_t s0 = (uint32_t)ctr ^ k.rk[0];
  uint32_t s1, s2, s3;
  const uint32_t hi = (uint32_t)(ctr >> 32);
  if (hi == 0) {
    // round 1 with the constant lookups pre-folded into k.r1c
    uint32_t t0 = te0<0>(a, s0) ^ k.r1c[0];
    uint32_t t1 = rotl8(te2<3>(a, s0)) ^ k.r1c[1];
    uint32_t t2 = te2<2>(a, s0) ^ k.r1c[2];
    uint32_t t3 = rotl8(te0<1>(a, s0)) ^ k.r1c[3];
    s0 = t0; s1 = t1; s2 = t2; s3 = t3;
  } else {
    s1 = hi ^ k.rk[1];
    s2 = k.rk[2];
    s3 = k.rk[3];
    uint32_t t0 = te0<0>(a, s0) ^ rotl8(te0<1>(a, s1)) ^ te2<2>(a, s2) ^ rotl8(te2<3>(a, s3)) ^ k.rk[4];
    uint32_t t1 = te0<0>(a, s1) ^ rotl8(te0<1>(a, s2)) ^ te2<2>(a, s3) ^ rotl8(te2<3>(a, s0)) ^ k.rk[5];
    uint32_t t2 = te0<0>(a, s2) ^ rotl8(te0<1>(a, s3)) ^ te2<2>(a, s0) ^ rotl8(te2<3>(a, s1)) ^ k.rk[6];
    uint32_t t3 = te0<0>(a, s3) ^ rotl8(te0<1>(a, s0)) ^ te2<2>(a, s1) ^ rotl8(te2<3>(a, s2)) ^ k.rk[7];
    s0 = t0; s1 = t1; s2 = t2; s3 = t3;
  }
#pragma unroll
  for (int r = 2; r < 10; r++) {
    uint32_t t0 = te0<0>(a, s0) ^ rotl8(te0<1>(a, s1)) ^ te2<2>(a, s2) ^ rotl8(te2<3>(a, s3)) ^ k.rk[4 * r];
    uint32_t t1 = te0<0>(a, s1) ^ rotl8(te0<1>(a, s2)) ^ te2<2>(a, s3) ^ rotl8(te2<3>(a, s0)) ^ k.rk[4 * r + 1];
    uint32_t t2 = te0<0>(a, s2) ^ rotl8(te0<1>(a, s3)) ^ te2<2>(a, s0) ^ rotl8(te2<3>(a, s1)) ^ k.rk[4 * r + 2];
    uint32_t t3 = te0<0>(a, s3) ^ rotl8(te0<1>(a, s0)) ^ te2<2>(a, s1) ^ rotl8(te2<3>(a, s2)) ^ k.rk[4 * r + 3];
    s0 = t0; s1 = t1; s2 = t2; s3 = t3;
  }
  // Last round: SubBytes + ShiftRows.  Output byte j of column c is S[byte j of s_{c+j}]:
  //   byte 0 = Te2[.].b0, byte 1 = Te0[.].b1, byte 2 = Te0[.].b2, byte 3 = Te2[.].b3.
#define AES_LAST(w0, w1, w2, w3, rkw)                                                     \
  (__byte_perm(__byte_perm(te2<0>(a, w0), te0<1>(a, w1), 0x0050),                         \
               __byte_perm(te0<2>(a, w2), te2<3>(a, w3), 0x7200), 0x7610) ^ (rkw))
  out[0] = AES_LAST(s0, s1, s2, s3, k.rk[40]);
  out[1] = AES_LAST(s1, s2, s3, s0, k.rk[41]);
  out[2] = AES_LAST(s2, s3, s0, s1, k.rk[42]);
  out[3] = AES_LAST(s3, s0, s1, s2, k.rk[43]);
#undef AES_LAST
}
