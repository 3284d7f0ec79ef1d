// Device-side AES-128 for the aes-prng compatible PRF stream:
//   block(i) = AES-128_seed(LE128(i)),   i = 0, 1, 2, ...
// (restatement pinned by moose/src/host/mod.rs:1050-1079, see oracle/).
//
// T-table formulation with ONE 256-entry table (Te0) and byte rotations.  The
// table is replicated 32x in shared memory, laid out [256][32] words, and lane l
// only ever reads column l: every lookup of a warp hits 32 distinct banks, so
// the data-dependent lookups are bank-conflict free (1 LDS wavefront each).
// Cost: 32 KB of shared memory per CTA.
#pragma once
#include <cstdint>

#include "common.h"

#define AES_TABLE_WORDS (256 * 32)

// Te0[x] = {2*S[x], S[x], S[x], 3*S[x]} as a little-endian word (one copy per
// translation unit, read once per CTA when the shared-memory table is built).
static __device__ const uint32_t c_Te0[256] = {
#include "aes_te0.inc"
};

__device__ __forceinline__ void aes_load_table(uint32_t *sT) {
  // sT[idx * 32 + lane]
  for (int i = threadIdx.x; i < AES_TABLE_WORDS; i += blockDim.x) sT[i] = c_Te0[i >> 5];
  __syncthreads();
}

__device__ __forceinline__ uint32_t rotl8(uint32_t x) { return __byte_perm(x, x, 0x2103); }
__device__ __forceinline__ uint32_t rotl16(uint32_t x) { return __byte_perm(x, x, 0x1032); }
__device__ __forceinline__ uint32_t rotl24(uint32_t x) { return __byte_perm(x, x, 0x0321); }

// Encrypts the counter block LE128(ctr) (ctr < 2^64).  `sTl` = sT + lane.
// out[0..3] are the 16 keystream bytes as little-endian words.
__device__ __forceinline__ void aes_ctr_block(const uint32_t *__restrict__ sTl, const AesKey &k, uint64_t ctr,
                                              uint32_t out[4]) {
  uint32_t s0 = (uint32_t)ctr ^ k.rk[0];
  uint32_t s1 = (uint32_t)(ctr >> 32) ^ k.rk[1];
  uint32_t s2 = k.rk[2];
  uint32_t s3 = k.rk[3];
#define TE(x) sTl[((x)&0xffu) << 5]
#pragma unroll
  for (int r = 1; r < 10; r++) {
    uint32_t t0 = TE(s0) ^ rotl8(TE(s1 >> 8)) ^ rotl16(TE(s2 >> 16)) ^ rotl24(TE(s3 >> 24)) ^ k.rk[4 * r];
    uint32_t t1 = TE(s1) ^ rotl8(TE(s2 >> 8)) ^ rotl16(TE(s3 >> 16)) ^ rotl24(TE(s0 >> 24)) ^ k.rk[4 * r + 1];
    uint32_t t2 = TE(s2) ^ rotl8(TE(s3 >> 8)) ^ rotl16(TE(s0 >> 16)) ^ rotl24(TE(s1 >> 24)) ^ k.rk[4 * r + 2];
    uint32_t t3 = TE(s3) ^ rotl8(TE(s0 >> 8)) ^ rotl16(TE(s1 >> 16)) ^ rotl24(TE(s2 >> 24)) ^ k.rk[4 * r + 3];
    s0 = t0; s1 = t1; s2 = t2; s3 = t3;
  }
  // Last round: SubBytes + ShiftRows only.  S[x] is byte 1 of Te0[x].
  // out byte j of column c = S[ byte j of s_{c+j} ].
#define SB(x, pos) ((TE(x) & 0x0000ff00u) >> 8 << (pos))
  out[0] = (SB(s0, 0) | SB(s1 >> 8, 8) | SB(s2 >> 16, 16) | SB(s3 >> 24, 24)) ^ k.rk[40];
  out[1] = (SB(s1, 0) | SB(s2 >> 8, 8) | SB(s3 >> 16, 16) | SB(s0 >> 24, 24)) ^ k.rk[41];
  out[2] = (SB(s2, 0) | SB(s3 >> 8, 8) | SB(s0 >> 16, 16) | SB(s1 >> 24, 24)) ^ k.rk[42];
  out[3] = (SB(s3, 0) | SB(s0 >> 8, 8) | SB(s1 >> 16, 16) | SB(s2 >> 24, 24)) ^ k.rk[43];
#undef SB
#undef TE
}

// PRF stream element helpers -------------------------------------------------
// Ring128 element e = (first_u64 << 64) + second_u64 of block e
// (moose/src/host/ops.rs:1998-2000): hi = bytes 0..7, lo = bytes 8..15.
__device__ __forceinline__ u128 aes_ring128(const uint32_t *sTl, const AesKey &k, uint64_t e) {
  uint32_t o[4];
  aes_ctr_block(sTl, k, e, o);
  uint64_t hi = (uint64_t)o[0] | ((uint64_t)o[1] << 32);
  uint64_t lo = (uint64_t)o[2] | ((uint64_t)o[3] << 32);
  return ((u128)hi << 64) | (u128)lo;
}
// Ring64 elements 2b and 2b+1 = the two little-endian u64 halves of block b
// (moose/src/host/ops.rs:1969).
__device__ __forceinline__ void aes_ring64_pair(const uint32_t *sTl, const AesKey &k, uint64_t b, uint64_t &e0,
                                                uint64_t &e1) {
  uint32_t o[4];
  aes_ctr_block(sTl, k, b, o);
  e0 = (uint64_t)o[0] | ((uint64_t)o[1] << 32);
  e1 = (uint64_t)o[2] | ((uint64_t)o[3] << 32);
}
