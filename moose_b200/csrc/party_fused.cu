// Party-sliced fused kernels of the party runtime (csrc/party.cpp): what ONE player computes between two messages of
// a replicated protocol, in one pass over memory with the PRF streams evaluated in registers.  The arithmetic per
// element is the reference's (mod 2^k / over Z2); the element-wise host operations it is spelled out as in the
// reference are cited per kernel.
#include "vecops.cuh"

namespace {

int aes_grid(mb_ctx *ctx, int64_t items, int tpb) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((items + tpb - 1) / tpb, (int64_t)mb_aes_sms(ctx)));
}

#define AES_LAUNCH(ctx, kern, tpb, items, ...)                                                                  \
  do {                                                                                                          \
    MB_CUDA((ctx), cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));    \
    const mb_aes_cfg_t _cfg = mb_aes_cfg((ctx), (items), (tpb));                                                \
    kern<<<_cfg.grid_dim, _cfg.block_dim, AES_SMEM_BYTES, (ctx)->stream>>>(__VA_ARGS__);                                  \
  } while (0)

#define PRF_PROLOGUE()                           \
  extern __shared__ __align__(16) uint32_t sT[]; \
  aes_load_table(sT);                            \
  const AesLane L = aes_lane(sT)

#define VEC_LOOP(v, nvec) \
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < (nvec); v += (int64_t)gridDim.x * blockDim.x)

#define CVP(t) ((const vec_t *)(t)->dptr)
#define VP(t) ((vec_t *)(t)->dptr)

#define DISPATCH_RING(dtype, CALL)                 \
  do {                                             \
    if ((dtype) == MB_RING128) { CALL(VK_RING128); } \
    else { CALL(VK_RING64); }                      \
  } while (0)

// bits 0..31 of x to the even bit positions of a 64-bit word
__device__ __forceinline__ uint64_t spread32(uint32_t x) {
  uint64_t v = x;
  v = (v | (v << 16)) & 0x0000FFFF0000FFFFull;
  v = (v | (v << 8)) & 0x00FF00FF00FF00FFull;
  v = (v | (v << 4)) & 0x0F0F0F0F0F0F0F0Full;
  v = (v | (v << 2)) & 0x3333333333333333ull;
  v = (v | (v << 1)) & 0x5555555555555555ull;
  return v;
}

template <int K>
__device__ __forceinline__ vec_t bits_at(const uint32_t *__restrict__ bits, int64_t v) {
  const int64_t e = v * (K == VK_RING128 ? 1 : 2);
  const uint32_t w = bits[e >> 5];
  const unsigned sh = (unsigned)(e & 31);
  if (K == VK_RING128) return make_ulonglong2((w >> sh) & 1u, 0);
  return make_ulonglong2((w >> sh) & 1u, (w >> (sh + 1)) & 1u);
}

// ---- bilinear steps: MulOp / AndOp rep_rep_kernel (replicated/arith.rs:317-367) for player i, with the zero share
// (zero_share.rs:96-106):  z_i = x_i (y_i + y_{i+1}) + x_{i+1} y_i + PRF(s_i) - PRF(s_{i+1})
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) party_mul_kernel(AesKey ka, AesKey kb, const vec_t *__restrict__ xa,
                                                               const vec_t *__restrict__ xb, const vec_t *__restrict__ ya,
                                                               const vec_t *__restrict__ yb, vec_t *__restrict__ z,
                                                               vec_t *__restrict__ remote, int64_t nvec, int64_t numel) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  VEC_LOOP(v, nvec) {
    const vec_t a = xa[v], b = xb[v], c = ya[v], d = yb[v];
    vec_t w = O::add(O::add(O::mul(a, O::add(c, d)), O::mul(b, c)), O::sub(prf_vec<K>(L, ka, (uint64_t)v), prf_vec<K>(L, kb, (uint64_t)v)));
    if (K == VK_BIT) w = bit_tail_mask(w, v, numel);
    z[v] = w;
    if (remote) remote[v] = w;  // the reshare: the same value lands in the previous player's ring over NVLink
  }
}

// tail of DotOp::rep_rep_kernel (arith.rs:476-484): z_i = v_i + PRF(s_i) - PRF(s_{i+1}); z may alias v
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) party_zero_add_kernel(AesKey ka, AesKey kb, const vec_t *v_in, vec_t *z,
                                                                      vec_t *__restrict__ remote, int64_t nvec, int64_t numel) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  VEC_LOOP(v, nvec) {
    vec_t w = O::add(v_in[v], O::sub(prf_vec<K>(L, ka, (uint64_t)v), prf_vec<K>(L, kb, (uint64_t)v)));
    if (K == VK_BIT) w = bit_tail_mask(w, v, numel);
    z[v] = w;
    if (remote) remote[v] = w;
  }
}

// The two products of MaximumOp / MuxOp (softmax.rs:36-50): s * y + (1 - s) * x for player i, each product with its own
// zero share, summed before the reshare.  (1 - s) touches share 0 only: one_a / one_b say whether this player's first /
// second share is share 0 (arith.rs:230-273).
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) party_mux_kernel(AesKey k1a, AesKey k1b, AesKey k2a, AesKey k2b, uint64_t one_a,
                                                               uint64_t one_b, const vec_t *__restrict__ sa, const vec_t *__restrict__ sb,
                                                               const vec_t *__restrict__ xa, const vec_t *__restrict__ xb,
                                                               const vec_t *__restrict__ ya, const vec_t *__restrict__ yb,
                                                               vec_t *__restrict__ z, vec_t *__restrict__ remote, int64_t nvec) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  VEC_LOOP(v, nvec) {
    const vec_t s0 = sa[v], s1 = sb[v];
    const vec_t n0 = O::sub(O::splat(one_a, 0), s0), n1 = O::sub(O::splat(one_b, 0), s1);
    const vec_t x0 = xa[v], x1 = xb[v], y0 = ya[v], y1 = yb[v];
    const vec_t p = O::add(O::add(O::mul(s0, O::add(y0, y1)), O::mul(s1, y0)),
                           O::sub(prf_vec<K>(L, k1a, (uint64_t)v), prf_vec<K>(L, k1b, (uint64_t)v)));
    const vec_t q = O::add(O::add(O::mul(n0, O::add(x0, x1)), O::mul(n1, x0)),
                           O::sub(prf_vec<K>(L, k2a, (uint64_t)v), prf_vec<K>(L, k2b, (uint64_t)v)));
    const vec_t w = O::add(p, q);
    z[v] = w;
    if (remote) remote[v] = w;
  }
}

// ---- TruncPr (additive/trunc.rs:22-169 through replicated/fixedpoint.rs:80-103) ----------------------------------------
// Dealer (third party): r uniform, r_top = (r << 1) >> (amount + 1), r_msb = r >> (k - 1), each additively shared
// as (PRF(seed_i), value - PRF(seed_i)) (additive/trunc.rs:63-96); o_* are the second shares.
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) trunc_dealer_kernel(AesKey kr, AesKey k0, AesKey k1, AesKey k2, unsigned amount,
                                                                  vec_t *__restrict__ o_r, vec_t *__restrict__ o_top,
                                                                  vec_t *__restrict__ o_msb, int64_t nvec) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  constexpr unsigned BITS = K == VK_RING128 ? 128 : 64;
  VEC_LOOP(v, nvec) {
    const vec_t r = prf_vec<K>(L, kr, (uint64_t)v);
    const vec_t r_top = O::shr(O::shl(r, 1), amount + 1), r_msb = O::shr(r, BITS - 1);
    o_r[v] = O::sub(r, prf_vec<K>(L, k0, (uint64_t)v));
    o_top[v] = O::sub(r_top, prf_vec<K>(L, k1, (uint64_t)v));
    o_msb[v] = O::sub(r_msb, prf_vec<K>(L, k2, (uint64_t)v));
  }
}

// First player: opens c = x + 2^(k-1) + r from its own additive share x0 + x1 and the masked share of the second
// player, then its share of the result (additive/trunc.rs:98-160):
//   y0 = c_top - r_top0 + ((r_msb0 + c_msb - 2 r_msb0 c_msb) << (k - amount)) - 2^(k-amount-1)
// followed by the first step of adt_to_rep (convert.rs:426-440): yp = PRF(seed), diff = y0 - yp.  The opened top bit
// c_msb is also written as a packed bit tensor (what travels to the second player).
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) trunc_p0_kernel(AesKey k0, AesKey k1, AesKey k2, AesKey kyp, unsigned amount,
                                                              const vec_t *__restrict__ xa, const vec_t *__restrict__ xb,
                                                              const vec_t *__restrict__ masked1, void *__restrict__ cbits,
                                                              vec_t *__restrict__ yp_out, vec_t *__restrict__ diff_out, int64_t nvec,
                                                              int64_t numel) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  constexpr unsigned BITS = K == VK_RING128 ? 128 : 64, KK = BITS - 1;
  const vec_t up = K == VK_RING128 ? make_ulonglong2(0, 1ull << 62) : make_ulonglong2(1ull << 62, 1ull << 62);  // 2^(k-1)
  const vec_t down = O::shl(O::splat(1, 0), KK - amount - 1);
  const int lane = threadIdx.x & 31;
  // warp-uniform trip count: the ballot below needs all lanes
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < nvec; base += (int64_t)gridDim.x * blockDim.x) {
    const int64_t v = base + lane;
    const bool live = v < nvec;
    vec_t c_msb = O::zero();
    if (live) {
      const vec_t cc = O::add(O::add(O::add(O::add(xa[v], xb[v]), up), prf_vec<K>(L, k0, (uint64_t)v)), masked1[v]);
      const vec_t c_top = O::shr(O::shl(cc, 1), amount + 1);
      c_msb = O::shr(cc, BITS - 1);
      const vec_t rtop0 = prf_vec<K>(L, k1, (uint64_t)v), rmsb0 = prf_vec<K>(L, k2, (uint64_t)v);
      const vec_t prod = O::mul01(rmsb0, c_msb);
      const vec_t ov = O::sub(O::sub(O::add(rmsb0, c_msb), prod), prod);
      const vec_t y = O::sub(O::add(O::sub(c_top, rtop0), O::shl(ov, KK - amount)), down);
      const vec_t yp = prf_vec<K>(L, kyp, (uint64_t)v);
      yp_out[v] = yp;
      diff_out[v] = O::sub(y, yp);
    }
    const uint32_t even = __ballot_sync(0xffffffffu, live && (c_msb.x & 1));
    if (K == VK_RING128) {
      if (lane == 0) ((uint32_t *)cbits)[base >> 5] = even;
    } else {
      const uint32_t odd = __ballot_sync(0xffffffffu, live && 2 * v + 1 < numel && (c_msb.y & 1));
      if (lane == 0) ((uint64_t *)cbits)[base >> 5] = spread32(even) | (spread32(odd) << 1);
    }
  }
}

// Second player (additive/trunc.rs:98-160): y1 = -r_top1 + ((r_msb1 - 2 r_msb1 c_msb) << (k - amount)), then
// yp = PRF(seed), diff = y1 - yp (convert.rs:426-440).  r_top1 / r_msb1 come from the dealer.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) trunc_p1_kernel(AesKey kyp, unsigned amount, const vec_t *__restrict__ r_top1,
                                                                const vec_t *__restrict__ r_msb1, const uint32_t *__restrict__ cbits, vec_t *__restrict__ yp_out,
                                                                vec_t *__restrict__ diff_out, int64_t nvec) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  constexpr unsigned KK = (K == VK_RING128 ? 128 : 64) - 1;
  VEC_LOOP(v, nvec) {
    const vec_t rtop1 = r_top1[v], rmsb1 = r_msb1[v];
    const vec_t prod = O::mul01(rmsb1, bits_at<K>(cbits, v));
    const vec_t y = O::add(O::neg(rtop1), O::shl(O::sub(O::sub(rmsb1, prod), prod), KK - amount));
    const vec_t yp = prf_vec<K>(L, kyp, (uint64_t)v);
    yp_out[v] = yp;
    diff_out[v] = O::sub(y, yp);
  }
}

// ---- ring_inject with a dabit (replicated/convert.rs:316-390, additive/dabit.rs:38-69) -----------------------------------
// Dealer: second additive share of the ring copy of its bits: br1 = inject(b) - PRF(seed_br).
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) dabit_dealer_kernel(AesKey kbr, const uint32_t *__restrict__ b, vec_t *__restrict__ br1,
                                                                    int64_t nvec) {
  PRF_PROLOGUE();
  VEC_LOOP(v, nvec) br1[v] = VOps<K>::sub(bits_at<K>(b, v), prf_vec<K>(L, kbr, (uint64_t)v));
}

// Players 0 / 1 after c = x ^ b is open: s = (b_share + [first] c - 2 b_share c) << bit_idx, where b_share is
// PRF(seed_br) on the first player and the dealt tensor on the second; then yp = PRF(seed), diff = s - yp.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) dabit_finish_kernel(AesKey kbr, AesKey kyp, int first, unsigned bit_idx,
                                                                    const vec_t *__restrict__ br1, const uint32_t *__restrict__ cbits,
                                                                    vec_t *__restrict__ yp_out, vec_t *__restrict__ diff_out, int64_t nvec) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  VEC_LOOP(v, nvec) {
    const vec_t c = bits_at<K>(cbits, v);
    const vec_t base = first ? prf_vec<K>(L, kbr, (uint64_t)v) : br1[v];
    const vec_t prod = O::mul01(base, c);
    vec_t s = O::sub(O::sub(first ? O::add(base, c) : base, prod), prod);
    s = O::shl(s, bit_idx);
    const vec_t yp = prf_vec<K>(L, kyp, (uint64_t)v);
    yp_out[v] = yp;
    diff_out[v] = O::sub(s, yp);
  }
}

}  // namespace

static int kind_of(mb_dtype d) { return d == MB_RING128 ? VK_RING128 : d == MB_BIT ? VK_BIT : VK_RING64; }

int k_party_mul(mb_ctx *ctx, const AesKey kk[2], const mb_tensor *xa, const mb_tensor *xb, const mb_tensor *ya, const mb_tensor *yb,
                mb_tensor *z, void *remote) {
  if (z->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(z);
#define ARGS kk[0], kk[1], CVP(xa), CVP(xb), CVP(ya), CVP(yb), VP(z), (vec_t *)remote, nvec, z->numel
  switch (kind_of(z->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, party_mul_kernel<VK_RING128>, AES_TPB_MID, nvec, ARGS); break;
    case VK_RING64: AES_LAUNCH(ctx, party_mul_kernel<VK_RING64>, AES_TPB_MID, nvec, ARGS); break;
    default: AES_LAUNCH(ctx, party_mul_kernel<VK_BIT>, AES_TPB_MID, nvec, ARGS); break;
  }
#undef ARGS
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_party_zero_add(mb_ctx *ctx, const AesKey kk[2], const mb_tensor *v, mb_tensor *z, void *remote) {
  if (z->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(z);
#define ARGS kk[0], kk[1], CVP(v), VP(z), (vec_t *)remote, nvec, z->numel
  switch (kind_of(z->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, party_zero_add_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, ARGS); break;
    case VK_RING64: AES_LAUNCH(ctx, party_zero_add_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, ARGS); break;
    default: AES_LAUNCH(ctx, party_zero_add_kernel<VK_BIT>, AES_TPB_LIGHT, nvec, ARGS); break;
  }
#undef ARGS
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// kk: zero-share keys of s * y (a, b) and of (1 - s) * x (a, b); me: this player's index (share 0 is the first share
// of player 0 and the second share of player 2)
int k_party_mux(mb_ctx *ctx, const AesKey kk[4], int me, const mb_tensor *const s[2], const mb_tensor *const x[2],
                const mb_tensor *const y[2], mb_tensor *z, void *remote) {
  if (z->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(z);
  const uint64_t one_a = me == 0, one_b = me == 2;
#define CALL(KIND)                                                                                                              \
  AES_LAUNCH(ctx, party_mux_kernel<KIND>, AES_TPB_MID, nvec, kk[0], kk[1], kk[2], kk[3], one_a, one_b, CVP(s[0]), CVP(s[1]), \
             CVP(x[0]), CVP(x[1]), CVP(y[0]), CVP(y[1]), VP(z), (vec_t *)remote, nvec)
  DISPATCH_RING(z->dtype, CALL);
#undef CALL
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// kk: fresh key of r, then the seeds of the first shares of r, r_top, r_msb; o: second shares
int k_party_trunc_dealer(mb_ctx *ctx, const AesKey kk[4], unsigned amount, mb_tensor *const o[3]) {
  if (o[0]->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(o[0]);
#define CALL(KIND) \
  AES_LAUNCH(ctx, trunc_dealer_kernel<KIND>, AES_TPB_MID, nvec, kk[0], kk[1], kk[2], kk[3], amount, VP(o[0]), VP(o[1]), VP(o[2]), nvec)
  DISPATCH_RING(o[0]->dtype, CALL);
#undef CALL
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// kk: seeds of the first shares of r, r_top, r_msb, then the adt_to_rep seed of this player
int k_party_trunc_p0(mb_ctx *ctx, const AesKey kk[4], unsigned amount, const mb_tensor *xa, const mb_tensor *xb, const mb_tensor *masked1,
                     mb_tensor *cbits, mb_tensor *yp, mb_tensor *diff) {
  if (yp->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(yp);
#define CALL(KIND)                                                                                                          \
  AES_LAUNCH(ctx, trunc_p0_kernel<KIND>, AES_TPB_MID, nvec, kk[0], kk[1], kk[2], kk[3], amount, CVP(xa), CVP(xb), CVP(masked1), \
             cbits->dptr, VP(yp), VP(diff), nvec, yp->numel)
  DISPATCH_RING(yp->dtype, CALL);
#undef CALL
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_party_trunc_p1(mb_ctx *ctx, const AesKey &kyp, unsigned amount, const mb_tensor *r_top1, const mb_tensor *r_msb1,
                     const mb_tensor *cbits, mb_tensor *yp, mb_tensor *diff) {
  if (yp->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(yp);
#define CALL(KIND)                                                                                                \
  AES_LAUNCH(ctx, trunc_p1_kernel<KIND>, AES_TPB_LIGHT, nvec, kyp, amount, CVP(r_top1), CVP(r_msb1),              \
             (const uint32_t *)cbits->dptr, VP(yp), VP(diff), nvec)
  DISPATCH_RING(yp->dtype, CALL);
#undef CALL
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

int k_party_dabit_dealer(mb_ctx *ctx, const AesKey &kbr, const mb_tensor *b, mb_tensor *br1) {
  if (br1->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(br1);
#define CALL(KIND) AES_LAUNCH(ctx, dabit_dealer_kernel<KIND>, AES_TPB_LIGHT, nvec, kbr, (const uint32_t *)b->dptr, VP(br1), nvec)
  DISPATCH_RING(br1->dtype, CALL);
#undef CALL
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}

// first != 0: player 0 (its share of the dabit is PRF(kbr), br1 unused); else player 1 (br1 from the dealer)
int k_party_dabit_finish(mb_ctx *ctx, const AesKey &kbr, const AesKey &kyp, int first, unsigned bit_idx, const mb_tensor *br1,
                         const mb_tensor *cbits, mb_tensor *yp, mb_tensor *diff) {
  if (yp->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(yp);
#define CALL(KIND)                                                                                                       \
  AES_LAUNCH(ctx, dabit_finish_kernel<KIND>, AES_TPB_LIGHT, nvec, kbr, kyp, first, bit_idx, br1 ? CVP(br1) : nullptr,  \
             (const uint32_t *)cbits->dptr, VP(yp), VP(diff), nvec)
  DISPATCH_RING(yp->dtype, CALL);
#undef CALL
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}
