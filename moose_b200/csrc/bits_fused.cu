// Fused arithmetic -> boolean conversion:
//   BitDecomposeOp::rep_ring_kernel = PlacementSplit::split (replicated/bits.rs:84-134) +
//   BinaryAdder::binary_adder, Kogge-Stone (replicated/misc.rs:176-242), MsbOp (arith.rs:611-626) and
//   RingInjectOp::rep_kernel with gen_dabit (replicated/convert.rs:316-390, additive/dabit.rs:38-69).
//
// The reference runs the adder as ~60 host ops per call (13 replicated ANDs for k = 64, each 3 ANDs + zero
// share, plus xor / shl_dim copies) over [k, n] bit tensors.  Here every adder LEVEL is one kernel over
// 128-bit plane vectors:  g' = g ^ AND(p, g << 2^i),  p' = AND(p, p << 2^i)  with the shift resolved as a plane
// index, the xor fused, and the zero-share bits of both ANDs generated in registers at exactly the stream
// positions the reference uses (AND number, plane, element) -- so every share bit is identical.
//
// msb-only pruning: the most significant bit needs only the carry into position k-1.  Walking the recurrence
// backwards gives, per level, the set of planes whose g / p values are ever read; only those planes (183 AND
// planes instead of 832 for k = 64) are computed, stored and given PRF bits.  The msb shares are bit-identical
// to plane k-1 of the full decomposition because the skipped planes never feed it.
//
// Layout requirement of the fused path: n % 128 == 0 (every plane starts on a 16-byte vector); other shapes
// use the op-by-op composition in rep_api.cpp.
#include "vecops.cuh"

namespace {

constexpr int TPB = 256;

// one CTA per SM: the AES table takes the CTA's shared-memory window up to 0x20000 (aes.cuh)
inline int aes_grid(mb_ctx *ctx, int64_t items, int tpb) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((items + tpb - 1) / tpb, (int64_t)mb_aes_sms(ctx)));
}

#define AES_LAUNCH(ctx, kern, tpb, items, ...)                                                                  \
  do {                                                                                                          \
    MB_CUDA((ctx), cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));    \
    const mb_aes_cfg_t _cfg = mb_aes_cfg((ctx), (items), (tpb));                                                \
    kern<<<_cfg.grid_dim, _cfg.block_dim, AES_SMEM_BYTES, (ctx)->stream>>>(__VA_ARGS__);                                  \
  } while (0)

__device__ __forceinline__ vec_t vxor(vec_t a, vec_t b) { return make_ulonglong2(a.x ^ b.x, a.y ^ b.y); }
__device__ __forceinline__ vec_t vand(vec_t a, vec_t b) { return make_ulonglong2(a.x & b.x, a.y & b.y); }
__device__ __forceinline__ vec_t vzero() { return make_ulonglong2(0, 0); }

// 32 x 32 bit-matrix transpose across the lanes of a warp (LSB-first): on entry lane r holds row r, on exit
// lane c holds column c, i.e. bit r of the result in lane c is bit c of the input of lane r.  Five butterfly
// stages; each exchanges the off-diagonal j x j blocks between lanes r and r ^ j.
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

// bit planes of (x0 + x1) and of x2: [KBITS, n] packed, n % 128 == 0.  A warp transposes 128 elements x KBITS
// bits as KBITS/32 x 4 blocks of 32 x 32 bits; lane l ends up with 16 bytes of planes l, l+32, ...  The eight warps
// of a CTA take eight consecutive groups and exchange through shared memory, so that every plane receives one full
// 128-byte line per store instead of eight scattered 16-byte pieces.
template <typename T, int KBITS>
__global__ void __launch_bounds__(TPB) split_planes_kernel(const T *__restrict__ x0, const T *__restrict__ x1,
                                                          const T *__restrict__ x2, const T *__restrict__ y0,
                                                          const T *__restrict__ y1, const T *__restrict__ y2,
                                                          uint32_t *__restrict__ outB, uint32_t *__restrict__ outC, int64_t n) {
  constexpr int WARPS = TPB / 32;
  constexpr int PL = KBITS / 32;
  __shared__ uint4 tile[KBITS][WARPS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t ngroups = n / 128;
  const int64_t nchunks = (ngroups + WARPS - 1) / WARPS;  // chunks of WARPS consecutive groups, per half
  for (int64_t c = blockIdx.x; c < 2 * nchunks; c += gridDim.x) {
    const bool second = c >= nchunks;  // first half of the work list: B = bits(x0 + x1); second half: C = bits(x2)
    const int64_t g0 = (second ? c - nchunks : c) * WARPS, g = g0 + w;
    if (g < ngroups) {
      uint32_t keep[PL][4];
#pragma unroll
      for (int it = 0; it < 4; it++) {
        const int64_t e = g * 128 + it * 32 + lane;
        T v = second ? x2[e] : (T)(x0[e] + x1[e]);
        // y != NULL: the operand is the share-wise difference x - y (LessOp, compare.rs:91-104), never materialised
        if (y0) v = second ? (T)(v - y2[e]) : (T)(v - y0[e] - y1[e]);
#pragma unroll
        for (int h = 0; h < PL; h++) keep[h][it] = warp_transpose32((uint32_t)(v >> (32 * h)), lane);
      }
#pragma unroll
      for (int q = 0; q < PL; q++) tile[q * 32 + lane][w] = make_uint4(keep[q][0], keep[q][1], keep[q][2], keep[q][3]);
    }
    __syncthreads();
    uint4 *out = (uint4 *)(second ? outC : outB);
    const int64_t row_vecs = n / 128;  // 16-byte vectors per plane
    for (int idx = threadIdx.x; idx < KBITS * WARPS; idx += TPB) {
      const int p = idx / WARPS, j = idx % WARPS;
      if (g0 + j < ngroups) out[(int64_t)p * row_vecs + g0 + j] = tile[p][j];
    }
    __syncthreads();
  }
}

// split + first AND of the adder, per plane vector v (flat index over [k, n]):
//   a = share_by_player0(B) = (R, B ^ R, 0)      R = PRF_bits(seed_share)        (convert.rs:72-94)
//   b = (0, 0, C)                                                                (bits.rs:124-130)
//   g0 = AND(a, b) = (0, (B^R)&C, R&C) ^ zero-share,   p0 = a ^ b = (R, B^R, C)
struct PlanePtrs {
  vec_t *s[3];
};
__global__ void __launch_bounds__(AES_TPB_LIGHT) adder_init_kernel(AesKey k_share, AesKey a0, AesKey a1, AesKey a2,
                                                        const vec_t *__restrict__ B, const vec_t *__restrict__ C,
                                                        PlanePtrs g0, PlanePtrs p0, int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const vec_t b = B[v], c = C[v];
    const vec_t R = prf_vec<VK_BIT>(L, k_share, (uint64_t)v);
    const vec_t r0 = prf_vec<VK_BIT>(L, a0, (uint64_t)v);
    const vec_t r1 = prf_vec<VK_BIT>(L, a1, (uint64_t)v);
    const vec_t r2 = prf_vec<VK_BIT>(L, a2, (uint64_t)v);
    const vec_t bR = vxor(b, R);
    g0.s[0][v] = vxor(r0, r1);
    g0.s[1][v] = vxor(vand(bR, c), vxor(r1, r2));
    g0.s[2][v] = vxor(vand(R, c), vxor(r2, r0));
    p0.s[0][v] = R;
    p0.s[1][v] = bR;
    p0.s[2][v] = c;
  }
}

__device__ __forceinline__ void rep_and3(const vec_t x[3], const vec_t y[3], const vec_t r[3], vec_t z[3]) {
  // z_i = x_i&y_i ^ x_i&y_{i+1} ^ x_{i+1}&y_i ^ (r_i ^ r_{i+1})   (arith.rs:342-358, zero_share.rs:96-106)
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int j = (i + 1) % 3;
    vec_t v = vxor(vand(x[i], vxor(y[i], y[j])), vand(x[j], y[i]));
    z[i] = vxor(v, vxor(r[i], r[j]));
  }
}

// One Kogge-Stone level (misc.rs:216-232) on the planes listed in `d`:
//   g'[q] = g[q] ^ AND_pg(p[q], g[q - shift]),   p'[q] = AND_pp(p[q], p[q - shift])   (operands below plane
//   `shift` are the zeros shl_dim inserts).  nv = vectors per plane.
// 768 threads (<= 80 registers) with the two ANDs kept apart: measured 4.33 / 9.07 ms per MLP / LogReg step against
// 4.44 / 9.30 ms for the six-chain version at 384 threads and 162 registers
#ifndef ADDER_LEVEL_TPB
#define ADDER_LEVEL_TPB 768
#endif
__global__ void __launch_bounds__(ADDER_LEVEL_TPB) adder_level_kernel(AesKey pg0, AesKey pg1, AesKey pg2, AesKey pp0, AesKey pp1,
                                                         AesKey pp2, const __grid_constant__ LevelDesc d, PlanePtrs g_in, PlanePtrs p_in,
                                                         PlanePtrs g_out, PlanePtrs p_out, int64_t nv) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  const int64_t total = (int64_t)d.nout * nv;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(t / nv);
    const int64_t v = t - (int64_t)j * nv;
    const int q = d.plane[j];
    const uint64_t pos = (uint64_t)q * (uint64_t)nv + (uint64_t)v;  // vector index in the [k, n] bit tensor
    // Operands are loaded right before their use and the two ANDs are kept apart, so that one AND's three AES chains
    // and its operands are live at a time (the compiler otherwise interleaves all six chains: 162 registers).
    const int64_t ig = (int64_t)d.gslot[q] * nv + v, ip = (int64_t)d.pslot[q] * nv + v;
    const bool shifted = q >= d.shift;
    const bool want_p = d.pout[j] != 255;
    vec_t P[3], X[3], r[3], z[3];
#pragma unroll
    for (int i = 0; i < 3; i++) P[i] = p_in.s[i][ip];
    if (shifted) {
      const int64_t jg = (int64_t)d.gslot[q - d.shift] * nv + v;
#pragma unroll
      for (int i = 0; i < 3; i++) X[i] = g_in.s[i][jg];
    } else {
#pragma unroll
      for (int i = 0; i < 3; i++) X[i] = vzero();
    }
    r[0] = prf_vec<VK_BIT>(L, pg0, pos);
    r[1] = prf_vec<VK_BIT>(L, pg1, pos);
    r[2] = prf_vec<VK_BIT>(L, pg2, pos);
    rep_and3(P, X, r, z);
    const int64_t og = (int64_t)j * nv + v;
#pragma unroll
    for (int i = 0; i < 3; i++) g_out.s[i][og] = vxor(g_in.s[i][ig], z[i]);
    if (want_p) {
      asm volatile("" ::: "memory");
      if (shifted) {
        const int64_t jp = (int64_t)d.pslot[q - d.shift] * nv + v;
#pragma unroll
        for (int i = 0; i < 3; i++) X[i] = p_in.s[i][jp];
      }
      r[0] = prf_vec<VK_BIT>(L, pp0, pos);
      r[1] = prf_vec<VK_BIT>(L, pp1, pos);
      r[2] = prf_vec<VK_BIT>(L, pp2, pos);
      rep_and3(P, X, r, z);
      const int64_t op = (int64_t)d.pout[j] * nv + v;
#pragma unroll
      for (int i = 0; i < 3; i++) p_out.s[i][op] = z[i];
    }
  }
}

// The same level for ONE party of a party-per-GPU run: it holds components me and me+1 of g and p (arrays s[0], s[1])
// and computes component me of g' and p'; component me+1 arrives from the next party.  Zero share: r_me ^ r_{me+1}.
struct PairPtrs {
  const vec_t *s[2];
};
__global__ void __launch_bounds__(AES_TPB_MID) adder_level_party_kernel(AesKey pg_a, AesKey pg_b, AesKey pp_a, AesKey pp_b, LevelDesc d,
                                                                       PairPtrs g_in, PairPtrs p_in, vec_t *__restrict__ g_out,
                                                                       vec_t *__restrict__ p_out, vec_t *__restrict__ g_rem,
                                                                       vec_t *__restrict__ p_rem, int64_t nv) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  const int64_t total = (int64_t)d.nout * nv;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(t / nv);
    const int64_t v = t - (int64_t)j * nv;
    const int q = d.plane[j];
    const uint64_t pos = (uint64_t)q * (uint64_t)nv + (uint64_t)v;
    const int64_t ig = (int64_t)d.gslot[q] * nv + v, ip = (int64_t)d.pslot[q] * nv + v;
    const vec_t G0 = g_in.s[0][ig], P0 = p_in.s[0][ip], P1 = p_in.s[1][ip];
    const bool shifted = q >= d.shift;
    const bool want_p = d.pout[j] != 255;
    vec_t Gs0 = vzero(), Gs1 = vzero(), Ps0 = vzero(), Ps1 = vzero();
    if (shifted) {
      const int64_t jg = (int64_t)d.gslot[q - d.shift] * nv + v;
      Gs0 = g_in.s[0][jg];
      Gs1 = g_in.s[1][jg];
      if (want_p) {
        const int64_t jp = (int64_t)d.pslot[q - d.shift] * nv + v;
        Ps0 = p_in.s[0][jp];
        Ps1 = p_in.s[1][jp];
      }
    }
    vec_t z = vxor(vxor(vand(P0, vxor(Gs0, Gs1)), vand(P1, Gs0)), vxor(prf_vec<VK_BIT>(L, pg_a, pos), prf_vec<VK_BIT>(L, pg_b, pos)));
    z = vxor(G0, z);
    g_out[(int64_t)j * nv + v] = z;
    if (g_rem) g_rem[(int64_t)j * nv + v] = z;  // the reshare: straight into the previous player's ring
    if (want_p) {
      z = vxor(vxor(vand(P0, vxor(Ps0, Ps1)), vand(P1, Ps0)), vxor(prf_vec<VK_BIT>(L, pp_a, pos), prf_vec<VK_BIT>(L, pp_b, pos)));
      p_out[(int64_t)d.pout[j] * nv + v] = z;
      if (p_rem) p_rem[(int64_t)d.pout[j] * nv + v] = z;
    }
  }
}

// Bit planes [KBITS, n] of t0 +- t1 +- t2 +- t3 (bit j of `neg` set: term j is subtracted): the player-local sum a
// party decomposes in the split (bits.rs:124-130 after the share-wise subtraction of LessOp, compare.rs:91-104).
// Same tiling as split_planes_kernel.
struct SplitTerms {
  const void *t[4];
  int n, neg;
};
template <typename T, int KBITS>
__global__ void __launch_bounds__(TPB) party_split_kernel(SplitTerms terms, uint32_t *__restrict__ outB, int64_t n) {
  constexpr int WARPS = TPB / 32;
  constexpr int PL = KBITS / 32;
  __shared__ uint4 tile[KBITS][WARPS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int64_t ngroups = n / 128;
  const int64_t nchunks = (ngroups + WARPS - 1) / WARPS;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int64_t g0 = c * WARPS, g = g0 + w;
    if (g < ngroups) {
      uint32_t keep[PL][4];
#pragma unroll
      for (int it = 0; it < 4; it++) {
        const int64_t e = g * 128 + it * 32 + lane;
        T v = ((const T *)terms.t[0])[e];
        if (terms.neg & 1) v = (T)0 - v;
#pragma unroll
        for (int j = 1; j < 4; j++)
          if (j < terms.n) {
            const T u = ((const T *)terms.t[j])[e];
            v = (terms.neg >> j) & 1 ? (T)(v - u) : (T)(v + u);
          }
#pragma unroll
        for (int h = 0; h < PL; h++) keep[h][it] = warp_transpose32((uint32_t)(v >> (32 * h)), lane);
      }
#pragma unroll
      for (int q = 0; q < PL; q++) tile[q * 32 + lane][w] = make_uint4(keep[q][0], keep[q][1], keep[q][2], keep[q][3]);
    }
    __syncthreads();
    uint4 *out = (uint4 *)outB;
    const int64_t row_vecs = n / 128;
    for (int idx = threadIdx.x; idx < KBITS * WARPS; idx += TPB) {
      const int p = idx / WARPS, j = idx % WARPS;
      if (g0 + j < ngroups) out[(int64_t)p * row_vecs + g0 + j] = tile[p][j];
    }
    __syncthreads();
  }
}

// Split + first AND of the adder for ONE player (adder_init_kernel restricted to its component).  With
// a = share_by_player0(B) = (R, B ^ R, 0) and b = (0, 0, C):  p0 = a ^ b = (R, B ^ R, C), g0 = AND(a, b) =
// (0, (B ^ R) & C, R & C) ^ zero share.  `in`: B on player 0, C on players 1 and 2; `other`: B ^ R received from
// player 0 (player 1 only).  Outputs: g (this player's component); p_new: R (players 0 and 2); p_send: B ^ R
// (player 0 only, what it sends to player 1).
__global__ void __launch_bounds__(AES_TPB_LIGHT) adder_init_party_kernel(int me, AesKey k_share, AesKey ka, AesKey kb,
                                                                        const vec_t *__restrict__ in, const vec_t *__restrict__ other,
                                                                        vec_t *__restrict__ g, vec_t *__restrict__ p_new,
                                                                        vec_t *__restrict__ p_send, vec_t *__restrict__ g_rem,
                                                                        int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const vec_t x = in[v];
    const vec_t z = vxor(prf_vec<VK_BIT>(L, ka, (uint64_t)v), prf_vec<VK_BIT>(L, kb, (uint64_t)v));
    vec_t gv;
    if (me == 1) {
      gv = vxor(vand(other[v], x), z);
    } else {
      const vec_t R = prf_vec<VK_BIT>(L, k_share, (uint64_t)v);
      p_new[v] = R;
      if (me == 0) {
        p_send[v] = vxor(x, R);
        gv = z;
      } else {
        gv = vxor(vand(R, x), z);
      }
    }
    g[v] = gv;
    if (g_rem) g_rem[v] = gv;  // the reshare of g0
  }
}

// z = shl_dim(g, 1) ^ p_store (misc.rs:236-240): z[q] = g[q-1] ^ p0[q], z[0] = p0[0].  Full decomposition.
__global__ void __launch_bounds__(TPB) adder_final_full_kernel(PlanePtrs g, PlanePtrs p0, PlanePtrs z, int64_t nv, int k) {
  const int64_t total = (int64_t)k * nv;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const bool low = t < nv;
#pragma unroll
    for (int i = 0; i < 3; i++) z.s[i][t] = low ? p0.s[i][t] : vxor(g.s[i][t - nv], p0.s[i][t]);
  }
}
// msb only: out = g_final[k-2] ^ p0[k-1]
__global__ void __launch_bounds__(TPB) adder_final_msb_kernel(PlanePtrs g, int64_t gslot, PlanePtrs p0, PlanePtrs z,
                                                             int64_t nv, int k) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int i = 0; i < 3; i++) z.s[i][v] = vxor(g.s[i][gslot * nv + v], p0.s[i][(int64_t)(k - 1) * nv + v]);
  }
}

// RingInjectOp::rep_kernel after the dabit bits b have been expanded (an n-bit tensor, 1/64 .. 1/128 of the ring
// data): per ring vector
//   c      = open(x + b_bin)                 = x0 ^ x1 ^ x2 ^ b                  (convert.rs:369-371)
//   br     = ring(b), br0 = PRF(seed_br), br1 = br - br0                         (dabit.rs:50-66)
//   s      = b_ring + c - 2 b_ring c   (public c lands on share 0)               (convert.rs:374-378)
//   s    <<= bit_idx                                                             (convert.rs:379)
//   adt_to_rep: y0 = PRF(seed1), y1 = PRF(seed2), cc = (s0 - y0) + (s1 - y1)     (convert.rs:426-447)
// Output shares (y0, cc, y1).  cc depends on the additive sharing (br0, br1) of the dabit only through
// s0 + s1 = br xor c (c is a bit) = b xor (x xor b) = x, the shared bit itself: PRF(seed_br) cancels mod 2^k and so
// does the dealer's bit b.  Neither is expanded (their entropy requests are still consumed, in order): two PRF streams
// per ring vector, the same shares as the op-by-op protocol bit for bit.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) ring_inject_fused_kernel(AesKey k_y0, AesKey k_y1,
                                                               const uint32_t *__restrict__ x0, const uint32_t *__restrict__ x1,
                                                               const uint32_t *__restrict__ x2,
                                                               vec_t *__restrict__ z0, vec_t *__restrict__ z1,
                                                               vec_t *__restrict__ z2, int64_t nvec, unsigned bit_idx) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  constexpr int EPV = K == VK_RING128 ? 1 : 2;  // ring elements per vector
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = v * EPV;
    const uint32_t ws = x0[e >> 5] ^ x1[e >> 5] ^ x2[e >> 5];   // c ^ b = the shared bit itself
    const unsigned sh = (unsigned)(e & 31);
    vec_t s;
    if (K == VK_RING128) s = make_ulonglong2((ws >> sh) & 1u, 0);
    else s = make_ulonglong2((ws >> sh) & 1u, (ws >> (sh + 1)) & 1u);
    s = O::shl(s, bit_idx);
    const vec_t y0 = prf_vec<K>(L, k_y0, (uint64_t)v);
    const vec_t y1 = prf_vec<K>(L, k_y1, (uint64_t)v);
    z0[v] = y0;
    z1[v] = O::sub(O::sub(s, y0), y1);
    z2[v] = y1;
  }
}

PlanePtrs ptrs_of(mb_tensor *const t[3]) {
  PlanePtrs p;
  for (int i = 0; i < 3; i++) p.s[i] = (vec_t *)t[i]->dptr;
  return p;
}

struct Tmp3 {
  mb_ctx *ctx;
  mb_tensor *t[3];
  explicit Tmp3(mb_ctx *c) : ctx(c) { t[0] = t[1] = t[2] = nullptr; }
  ~Tmp3() { for (int i = 0; i < 3; i++) if (t[i]) mb_tensor_free(ctx, t[i]); }
  Tmp3(const Tmp3 &) = delete;
  Tmp3 &operator=(const Tmp3 &) = delete;
  int alloc(int64_t planes, int64_t n) {
    int64_t shape[2] = {planes, n};
    for (int i = 0; i < 3; i++) MB_CHECK(mb_tensor_new(ctx, MB_BIT, 2, shape, &t[i]));
    return MB_OK;
  }
  void swap(Tmp3 &o) { for (int i = 0; i < 3; i++) std::swap(t[i], o.t[i]); }
};

}  // namespace

bool k_rep_adder_supported(const mb_tensor *x) { return mb_is_ring(x->dtype) && x->numel > 0 && x->numel % 128 == 0; }

// keys: [0] split share seed, [1..3] first AND, then per level i: [4 + 6i .. 4 + 6i + 2] = AND(p, g1),
// [4 + 6i + 3 .. 4 + 6i + 5] = AND(p, p1).  out[3]: bit tensors of shape [k, shape...] (full) or [shape...] (msb).
int k_rep_adder(mb_ctx *ctx, const AesKey *keys, const mb_rep *x, bool msb_only, mb_tensor *out[3], int out_planes, const mb_rep *minus) {
  const mb_tensor *x0 = x->s[0];
  const int k = x0->dtype == MB_RING128 ? 128 : 64;
  const int levels = k == 128 ? 7 : 6;
  const int64_t n = x0->numel, nv = n / 128;
  // full decomposition of which only planes 0 .. kout-1 will be read (carries only travel upwards, so the higher
  // planes are never computed; they are left zero): top_most_index reads max_bits <= k / 2 planes
  const int kout = msb_only ? k : std::max(1, std::min(out_planes > 0 ? out_planes : k, k));

  // 1. bit planes of x0 + x1 and of x2
  mb_tensor *B = nullptr, *C = nullptr;
  int64_t pshape[2] = {k, n};
  MB_CHECK(mb_tensor_new(ctx, MB_BIT, 2, pshape, &B));
  int rc = mb_tensor_new(ctx, MB_BIT, 2, pshape, &C);
  if (rc != MB_OK) { mb_tensor_free(ctx, B); return rc; }
  struct Guard { mb_ctx *c; mb_tensor *a, *b; ~Guard() { mb_tensor_free(c, a); mb_tensor_free(c, b); } } guard{ctx, B, C};
  {
    const int64_t warps = 2 * (n / 128);
    const int g = (int)std::min<int64_t>((warps * 32 + TPB - 1) / TPB, (int64_t)ctx->sm_count * 8);
    if (k == 128)
      split_planes_kernel<u128, 128><<<g, TPB, 0, ctx->stream>>>((const u128 *)x->s[0]->dptr, (const u128 *)x->s[1]->dptr,
                                                                (const u128 *)x->s[2]->dptr, minus ? (const u128 *)minus->s[0]->dptr : nullptr,
                                                                minus ? (const u128 *)minus->s[1]->dptr : nullptr,
                                                                minus ? (const u128 *)minus->s[2]->dptr : nullptr, (uint32_t *)B->dptr,
                                                                (uint32_t *)C->dptr, n);
    else
      split_planes_kernel<uint64_t, 64><<<g, TPB, 0, ctx->stream>>>(
          (const uint64_t *)x->s[0]->dptr, (const uint64_t *)x->s[1]->dptr, (const uint64_t *)x->s[2]->dptr,
          minus ? (const uint64_t *)minus->s[0]->dptr : nullptr, minus ? (const uint64_t *)minus->s[1]->dptr : nullptr,
          minus ? (const uint64_t *)minus->s[2]->dptr : nullptr, (uint32_t *)B->dptr, (uint32_t *)C->dptr, n);
    MB_LAUNCH_CHECK_W(ctx, "split_planes", (minus ? 6.0 : 3.0) * n * (k / 8) + 2.0 * k * n / 8, 0, 0);
  }

  // 2. g0, p0 on all planes
  Tmp3 g(ctx), p(ctx), p0(ctx);
  MB_CHECK(g.alloc(k, n));
  MB_CHECK(p0.alloc(k, n));
  {
    const int64_t total = (int64_t)kout * nv;
    AES_LAUNCH(ctx, adder_init_kernel, AES_TPB_LIGHT, total, keys[0], keys[1], keys[2], keys[3], (const vec_t *)B->dptr,
               (const vec_t *)C->dptr, ptrs_of(g.t), ptrs_of(p0.t), total);
    MB_LAUNCH_CHECK_W(ctx, "adder_init", 16.0 * total * 8, 4.0 * total, 0);
  }

  // 3. plane sets needed per level (all planes for the full decomposition)
  std::vector<std::vector<int>> need_g(levels + 1), need_p(levels + 1);
  if (msb_only) {
    std::vector<char> ng(k, 0), np(k, 0);
    ng[k - 2] = 1;
    for (int i = levels; i >= 0; i--) {
      for (int q = 0; q < k; q++) {
        if (ng[q]) need_g[i].push_back(q);
        if (np[q]) need_p[i].push_back(q);
      }
      if (i == 0) break;
      const int s = 1 << (i - 1);
      std::vector<char> g2(k, 0), p2(k, 0);
      for (int q = 0; q < k; q++) {
        if (ng[q]) {  // g'[q] = g[q] ^ AND(p[q], g[q-s])
          g2[q] = 1;
          p2[q] = 1;
          if (q >= s) g2[q - s] = 1;
        }
        if (np[q]) {  // p'[q] = AND(p[q], p[q-s])
          p2[q] = 1;
          if (q >= s) p2[q - s] = 1;
        }
      }
      ng = g2;
      np = p2;
    }
  } else {
    // z[q] = g_final[q - 1] ^ p0[q] for q < kout: g and p of planes 0 .. kout-2 on every level (all planes when kout = k:
    // g of plane k-1 is computed by the reference too, and nothing distinguishes it here)
    const int top = kout == k ? k : kout - 1;
    for (int i = 0; i <= levels; i++)
      for (int q = 0; q < top; q++) {
        need_g[i].push_back(q);
        if (i < levels) need_p[i].push_back(q);  // p after the last level is never read (misc.rs:236-240)
      }
  }

  // 4. levels.  Level 0 reads the init outputs (slot = plane); p of level 0 is p0 itself.
  std::vector<int> gslot(k), pslot(k);
  for (int q = 0; q < k; q++) gslot[q] = pslot[q] = q;
  Tmp3 *p_cur = &p0;  // not owned by the loop on level 0
  for (int i = 0; i < levels; i++) {
    LevelDesc d;
    memset(&d, 255, sizeof(d));
    d.shift = 1 << i;
    d.nout = (int)need_g[i + 1].size();
    for (int q = 0; q < k; q++) {
      d.gslot[q] = (unsigned char)(gslot[q] < 0 ? 255 : gslot[q]);
      d.pslot[q] = (unsigned char)(pslot[q] < 0 ? 255 : pslot[q]);
    }
    std::vector<int> pslot_next(k, -1), gslot_next(k, -1);
    int np_out = 0;
    for (int j = 0; j < d.nout; j++) {
      const int q = need_g[i + 1][j];
      d.plane[j] = (unsigned char)q;
      gslot_next[q] = j;
      const bool want_p = std::find(need_p[i + 1].begin(), need_p[i + 1].end(), q) != need_p[i + 1].end();
      if (want_p) {
        d.pout[j] = (unsigned char)np_out;
        pslot_next[q] = np_out++;
      }
    }
    // every plane whose p' is needed also needs g' in this recurrence; assert the planner's invariant
    for (int q : need_p[i + 1])
      if (gslot_next[q] < 0) return mb_fail(ctx, MB_ERR_KERNEL, "adder plan: p needed without g");
    Tmp3 g_next(ctx), p_next(ctx);
    MB_CHECK(g_next.alloc(std::max(d.nout, 1), n));
    MB_CHECK(p_next.alloc(std::max(np_out, 1), n));
    const AesKey *kk = keys + 4 + 6 * i;
    AES_LAUNCH(ctx, adder_level_kernel, ADDER_LEVEL_TPB, (int64_t)d.nout * nv, kk[0], kk[1], kk[2], kk[3], kk[4], kk[5], d,
               ptrs_of(g.t), ptrs_of(p_cur->t), ptrs_of(g_next.t), ptrs_of(p_next.t), nv);
    {
      const double items = (double)d.nout * nv, pfrac = d.nout ? (double)np_out / d.nout : 0.0;
      MB_LAUNCH_CHECK_W(ctx, "adder_level", 16.0 * items * (12 + 3 + 3 * pfrac), items * (3 + 3 * pfrac), 0);
    }
    g.swap(g_next);
    p.swap(p_next);
    p_cur = &p;
    gslot = gslot_next;
    pslot = pslot_next;
  }

  // 5. z = shl_dim(g, 1) ^ p_store
  PlanePtrs zp = ptrs_of(out);
  if (msb_only) {
    adder_final_msb_kernel<<<(unsigned)std::max<int64_t>(1, (nv + TPB - 1) / TPB), TPB, 0, ctx->stream>>>(
        ptrs_of(g.t), (int64_t)gslot[k - 2], ptrs_of(p0.t), zp, nv, k);
  } else {
    const int64_t total = (int64_t)kout * nv;
    adder_final_full_kernel<<<(unsigned)std::min<int64_t>((total + TPB - 1) / TPB, (int64_t)ctx->sm_count * 16), TPB, 0,
                              ctx->stream>>>(ptrs_of(g.t), ptrs_of(p0.t), zp, nv, kout);
    if (kout < k)
      for (int i = 0; i < 3; i++)
        MB_CUDA(ctx, cudaMemsetAsync((char *)out[i]->dptr + (size_t)kout * nv * 16, 0, (size_t)(k - kout) * nv * 16, ctx->stream));
  }
  MB_LAUNCH_CHECK_W(ctx, "adder_final", 16.0 * 9 * nv * (msb_only ? 1 : kout), 0, 0);
  return MB_OK;
}

// x: replicated bit shares (n bits); out: ring shares (y0, cc, y1).
int k_ring_inject_fused(mb_ctx *ctx, const AesKey &k_y0, const AesKey &k_y1, unsigned bit_idx, const mb_rep *x,
                        mb_tensor *out[3]) {
  const mb_tensor *t = out[0];
  if (t->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(t);
#define ARGS k_y0, k_y1, (const uint32_t *)x->s[0]->dptr, (const uint32_t *)x->s[1]->dptr, (const uint32_t *)x->s[2]->dptr, \
             (vec_t *)out[0]->dptr, (vec_t *)out[1]->dptr, (vec_t *)out[2]->dptr, nvec
  if (t->dtype == MB_RING128)
    AES_LAUNCH(ctx, ring_inject_fused_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, ARGS, bit_idx & 127);
  else
    AES_LAUNCH(ctx, ring_inject_fused_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, ARGS, bit_idx & 63);
#undef ARGS
  MB_LAUNCH_CHECK_W(ctx, "ring_inject_fused", 16.0 * nvec * 3 + 4.0 * t->numel / 8, 2.0 * nvec, 0);
  return MB_OK;
}

// Plane sets per level of the msb-only adder (the recurrence walked backwards from plane k-2 of the final g) and the
// level descriptors, shared by the all-party kernel above and the party-per-GPU path (party.cpp).
int k_adder_plan_msb(int k, std::vector<LevelDesc> *levels_out, std::vector<int> *np_out, int *final_gslot) {
  const int levels = k == 128 ? 7 : 6;
  std::vector<std::vector<int>> need_g((size_t)levels + 1), need_p((size_t)levels + 1);
  std::vector<char> ng((size_t)k, 0), np((size_t)k, 0);
  ng[(size_t)k - 2] = 1;
  for (int i = levels; i >= 0; i--) {
    for (int q = 0; q < k; q++) {
      if (ng[(size_t)q]) need_g[(size_t)i].push_back(q);
      if (np[(size_t)q]) need_p[(size_t)i].push_back(q);
    }
    if (i == 0) break;
    const int s = 1 << (i - 1);
    std::vector<char> g2((size_t)k, 0), p2((size_t)k, 0);
    for (int q = 0; q < k; q++) {
      if (ng[(size_t)q]) {
        g2[(size_t)q] = 1;
        p2[(size_t)q] = 1;
        if (q >= s) g2[(size_t)(q - s)] = 1;
      }
      if (np[(size_t)q]) {
        p2[(size_t)q] = 1;
        if (q >= s) p2[(size_t)(q - s)] = 1;
      }
    }
    ng = g2;
    np = p2;
  }
  std::vector<int> gslot((size_t)k), pslot((size_t)k);
  for (int q = 0; q < k; q++) gslot[(size_t)q] = pslot[(size_t)q] = q;
  levels_out->clear();
  np_out->clear();
  for (int i = 0; i < levels; i++) {
    LevelDesc d;
    memset(&d, 255, sizeof(d));
    d.shift = 1 << i;
    d.nout = (int)need_g[(size_t)i + 1].size();
    for (int q = 0; q < k; q++) {
      d.gslot[q] = (unsigned char)(gslot[(size_t)q] < 0 ? 255 : gslot[(size_t)q]);
      d.pslot[q] = (unsigned char)(pslot[(size_t)q] < 0 ? 255 : pslot[(size_t)q]);
    }
    std::vector<int> pslot_next((size_t)k, -1), gslot_next((size_t)k, -1);
    int npo = 0;
    for (int j = 0; j < d.nout; j++) {
      const int q = need_g[(size_t)i + 1][(size_t)j];
      d.plane[j] = (unsigned char)q;
      gslot_next[(size_t)q] = j;
      const auto &np1 = need_p[(size_t)i + 1];
      if (std::find(np1.begin(), np1.end(), q) != np1.end()) {
        d.pout[j] = (unsigned char)npo;
        pslot_next[(size_t)q] = npo++;
      }
    }
    levels_out->push_back(d);
    np_out->push_back(npo);
    gslot = gslot_next;
    pslot = pslot_next;
  }
  *final_gslot = gslot[(size_t)k - 2];
  return MB_OK;
}

// keys: zero-share streams (me, me+1) of AND(p, g1), then of AND(p, p1).  g_in / p_in: this party's two components.
// planes [k, shape...] of the signed sum of 1..4 ring tensors (element count a multiple of 128)
int k_party_split(mb_ctx *ctx, const mb_tensor *const terms[], int nterms, int neg_mask, mb_tensor *out) {
  SplitTerms t;
  t.n = nterms;
  t.neg = neg_mask;
  for (int j = 0; j < 4; j++) t.t[j] = j < nterms ? terms[j]->dptr : nullptr;
  const int64_t n = terms[0]->numel;
  const int64_t nchunks = (n / 128 + TPB / 32 - 1) / (TPB / 32);
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(nchunks, (int64_t)ctx->sm_count * 8));
  if (terms[0]->dtype == MB_RING128)
    party_split_kernel<u128, 128><<<grid, TPB, 0, ctx->stream>>>(t, (uint32_t *)out->dptr, n);
  else
    party_split_kernel<uint64_t, 64><<<grid, TPB, 0, ctx->stream>>>(t, (uint32_t *)out->dptr, n);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// keys: [0] seed of player 0's sharing of B, [1], [2] this player's zero-share seeds of the first AND
int k_party_adder_init(mb_ctx *ctx, int me, const AesKey keys[3], const mb_tensor *in, const mb_tensor *other, mb_tensor *g,
                       mb_tensor *p_new, mb_tensor *p_send, void *g_remote) {
  const int64_t nvec = (int64_t)(in->nbytes / 16);
  AES_LAUNCH(ctx, adder_init_party_kernel, AES_TPB_LIGHT, nvec, me, keys[0], keys[1], keys[2], (const vec_t *)in->dptr,
             other ? (const vec_t *)other->dptr : nullptr, (vec_t *)g->dptr, p_new ? (vec_t *)p_new->dptr : nullptr,
             p_send ? (vec_t *)p_send->dptr : nullptr, (vec_t *)g_remote, nvec);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_party_adder_level(mb_ctx *ctx, const AesKey keys[4], const LevelDesc &d, const mb_tensor *const g_in[2],
                        const mb_tensor *const p_in[2], mb_tensor *g_out, mb_tensor *p_out, int64_t nv, void *remote) {
  PairPtrs g, p;
  for (int c = 0; c < 2; c++) {
    g.s[c] = (const vec_t *)g_in[c]->dptr;
    p.s[c] = (const vec_t *)p_in[c]->dptr;
  }
  AES_LAUNCH(ctx, adder_level_party_kernel, AES_TPB_MID, (int64_t)d.nout * nv, keys[0], keys[1], keys[2], keys[3], d, g, p,
             (vec_t *)g_out->dptr, (vec_t *)p_out->dptr, (vec_t *)remote,
             // g' and p' are two views of one buffer; the remote copy has the same layout
             remote ? (vec_t *)((uint8_t *)remote + ((uint8_t *)p_out->dptr - (uint8_t *)g_out->dptr)) : nullptr, nv);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}
