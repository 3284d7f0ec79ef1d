// Fused replicated-protocol kernels.  Each kernel
// does, in one pass over memory and with the PRF evaluated in registers, what the
// reference executes as a few dozen host ops with materialised temporaries.  The
// arithmetic per element is exactly the reference's (mod 2^k / over Z2), so the
// resulting shares are bit-identical.
//
// Shares are stored de-duplicated: a replicated tensor is (x0, x1, x2) and party i
// holds (x_i, x_{i+1}) -- the reference's [[x00,x10],[x11,x21],[x22,x02]]
// (replicated/mod.rs:75-77) with x00==x02, x10==x11, x21==x22.
#include "vecops.cuh"

namespace {

constexpr int TPB = 256;

#define PRF_PROLOGUE()                             \
  extern __shared__ __align__(16) uint32_t sT[];   \
  aes_load_table(sT);                              \
  const AesLane sTl = aes_lane(sT)

// AES kernels need 64 KB of dynamic shared memory for the lookup table.
#define AES_LAUNCH(ctx, kern, tpb, nvec, ...)                                                                   \
  do {                                                                                                          \
    MB_CUDA((ctx), cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));    \
    const mb_aes_cfg_t _cfg = mb_aes_cfg((ctx), (nvec), (tpb));                                                 \
    kern<<<_cfg.grid_dim, _cfg.block_dim, AES_SMEM_BYTES, (ctx)->stream>>>(__VA_ARGS__);                                  \
  } while (0)

#define VEC_LOOP(v, nvec) \
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < (nvec); v += (int64_t)gridDim.x * blockDim.x)

// ShareOp::ring_kernel, owner in the placement (replicated/convert.rs:72-129):
//   x_owner = PRF(seed), x_{owner+1} = x - x_owner, x_{owner+2} = 0.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) share_kernel(AesKey key, const vec_t *__restrict__ x, vec_t *__restrict__ o_prf,
                                                    vec_t *__restrict__ o_diff, vec_t *__restrict__ o_zero, int64_t nvec,
                                                    int64_t numel) {
  PRF_PROLOGUE();
  VEC_LOOP(v, nvec) {
    vec_t r = prf_vec<K>(sTl, key, (uint64_t)v);
    if (K == VK_BIT) r = bit_tail_mask(r, v, numel);
    vec_t xv = x[v];
    o_prf[v] = r;
    o_diff[v] = VOps<K>::sub(xv, r);
    o_zero[v] = VOps<K>::zero();
  }
}

// ShareOp::ring_kernel, owner outside the placement (replicated/convert.rs:130-156):
//   x0 = PRF(seed00), x1 = PRF(seed11), x2 = x - x0 - x1.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) share_outsider_kernel(AesKey k0, AesKey k1, const vec_t *__restrict__ x,
                                                             vec_t *__restrict__ o0, vec_t *__restrict__ o1,
                                                             vec_t *__restrict__ o2, int64_t nvec, int64_t numel) {
  PRF_PROLOGUE();
  VEC_LOOP(v, nvec) {
    vec_t r0 = prf_vec<K>(sTl, k0, (uint64_t)v);
    vec_t r1 = prf_vec<K>(sTl, k1, (uint64_t)v);
    if (K == VK_BIT) {
      r0 = bit_tail_mask(r0, v, numel);
      r1 = bit_tail_mask(r1, v, numel);
    }
    vec_t xv = x[v];
    o0[v] = r0;
    o1[v] = r1;
    o2[v] = VOps<K>::sub(VOps<K>::sub(xv, r0), r1);
  }
}

// RevealOp::host_ring_kernel (replicated/convert.rs:295-311): sum of the three shares.
template <int K>
__global__ void __launch_bounds__(TPB) reveal_kernel(const vec_t *__restrict__ a, const vec_t *__restrict__ b,
                                                     const vec_t *__restrict__ c, vec_t *__restrict__ out, int64_t nvec) {
  VEC_LOOP(v, nvec) { out[v] = VOps<K>::add(VOps<K>::add(a[v], b[v]), c[v]); }
}

// MulOp::rep_rep_kernel (replicated/arith.rs:317-367) + gen_zero_share
// (replicated/zero_share.rs:82-156):
//   v_i = x_i*y_i + x_i*y_{i+1} + x_{i+1}*y_i
//   z_i = v_i + (PRF(s_i) - PRF(s_{i+1}))          s_i = DeriveSeed(k_i, sync_i)
// The reference expands six PRF streams (two per party); s10==s11 etc., so three
// distinct streams suffice.  Reads 6 vectors, writes 3, evaluates 3 AES blocks.
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) rep_mul_kernel(AesKey s0, AesKey s1, AesKey s2, const vec_t *__restrict__ x0,
                                                      const vec_t *__restrict__ x1, const vec_t *__restrict__ x2,
                                                      const vec_t *__restrict__ y0, const vec_t *__restrict__ y1,
                                                      const vec_t *__restrict__ y2, vec_t *__restrict__ z0,
                                                      vec_t *__restrict__ z1, vec_t *__restrict__ z2, int64_t nvec,
                                                      int64_t numel) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  VEC_LOOP(v, nvec) {
    vec_t a0 = x0[v], a1 = x1[v], a2 = x2[v];
    vec_t b0 = y0[v], b1 = y1[v], b2 = y2[v];
    vec_t r0 = prf_vec<K>(sTl, s0, (uint64_t)v);
    vec_t r1 = prf_vec<K>(sTl, s1, (uint64_t)v);
    vec_t r2 = prf_vec<K>(sTl, s2, (uint64_t)v);
    // x_i*(y_i + y_{i+1}) + x_{i+1}*y_i is the same value mod 2^k with one product less
    vec_t v0 = O::add(O::mul(a0, O::add(b0, b1)), O::mul(a1, b0));
    vec_t v1 = O::add(O::mul(a1, O::add(b1, b2)), O::mul(a2, b1));
    vec_t v2 = O::add(O::mul(a2, O::add(b2, b0)), O::mul(a0, b2));
    vec_t w0 = O::add(v0, O::sub(r0, r1));
    vec_t w1 = O::add(v1, O::sub(r1, r2));
    vec_t w2 = O::add(v2, O::sub(r2, r0));
    if (K == VK_BIT) {
      w0 = bit_tail_mask(w0, v, numel);
      w1 = bit_tail_mask(w1, v, numel);
      w2 = bit_tail_mask(w2, v, numel);
    }
    z0[v] = w0;
    z1[v] = w1;
    z2[v] = w2;
  }
}

// z_i = v_i + PRF(s_i) - PRF(s_{i+1}): tail of DotOp::rep_rep_kernel
// (replicated/arith.rs:476-484).  v_i may alias z_i.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) add_zero_share_kernel(AesKey s0, AesKey s1, AesKey s2, const vec_t *v0p,
                                                             const vec_t *v1p, const vec_t *v2p, vec_t *z0, vec_t *z1,
                                                             vec_t *z2, int64_t nvec, int64_t numel) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  VEC_LOOP(v, nvec) {
    vec_t r0 = prf_vec<K>(sTl, s0, (uint64_t)v);
    vec_t r1 = prf_vec<K>(sTl, s1, (uint64_t)v);
    vec_t r2 = prf_vec<K>(sTl, s2, (uint64_t)v);
    vec_t w0 = O::add(v0p[v], O::sub(r0, r1));
    vec_t w1 = O::add(v1p[v], O::sub(r1, r2));
    vec_t w2 = O::add(v2p[v], O::sub(r2, r0));
    if (K == VK_BIT) {
      w0 = bit_tail_mask(w0, v, numel);
      w1 = bit_tail_mask(w1, v, numel);
      w2 = bit_tail_mask(w2, v, numel);
    }
    z0[v] = w0;
    z1[v] = w1;
    z2[v] = w2;
  }
}

struct TruncParams {
  unsigned sh_msb;   // (L-1)
  unsigned sh_top;   // (amount+1) & (L-1)
  unsigned sh_ov;    // (k-amount) & (L-1)
  uint64_t up_lo, up_hi;      // 1 << (k-1)
  uint64_t down_lo, down_hi;  // 1 << (k-amount-1)
};

// TruncPrOp::rep_kernel (replicated/fixedpoint.rs:80-103) in one pass:
//   rep_to_adt (additive/convert.rs:28-37):   a0 = x0 + x1, a1 = x2
//   gen_trunc_mask at P2 (additive/trunc.rs:45-63): r = PRF(seed_r) [unseeded Sample],
//       r_msb = r >> (L-1), r_top = (r << 1) >> (amount+1),
//       additive shares (PRF(seed_a), r - .), (PRF(seed_b), r_top - .), (PRF(seed_c), r_msb - .)
//   trunc_pr (additive/trunc.rs:136-168): c = open(x + 2^(k-1) + r), c_top, c_msb,
//       overflow = r_msb xor c_msb = r_msb + c_msb - 2 r_msb c_msb,
//       y = c_top - r_top + (overflow << (k-amount)) - 2^(k-amount-1)
//   adt_to_rep (replicated/convert.rs:426-447,480-484): z0 = PRF(seed_1), z2 = PRF(seed_2),
//       z1 = open(y - (z0, z2)).
// With the three parties in one kernel only (z0, z1, z2) leave it, and z1 = y0 + y1 - z0 - z2 depends on the dealer's
// three additive sharings only through their sums: PRF(seed_a), PRF(seed_b), PRF(seed_c) cancel mod 2^k
// ((rm0 + c_msb - 2 rm0 c_msb) + (rm1 - 2 rm1 c_msb) = r_msb xor c_msb for a bit c_msb), so those streams are never
// expanded -- the same values as the op-by-op protocol, bit for bit (tests/test_gpu_replicated.py).
// Three PRF streams (r, z0, z2), 3 vectors read, 3 written.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) trunc_pr_kernel(AesKey kr, AesKey k1, AesKey k2, TruncParams p,
                                                       const vec_t *__restrict__ x0, const vec_t *__restrict__ x1,
                                                       const vec_t *__restrict__ x2, vec_t *__restrict__ z0,
                                                       vec_t *__restrict__ z1, vec_t *__restrict__ z2, int64_t nvec) {
  PRF_PROLOGUE();
  typedef VOps<K> O;
  const vec_t up = O::splat(p.up_lo, p.up_hi);
  const vec_t down = O::splat(p.down_lo, p.down_hi);
  VEC_LOOP(v, nvec) {
    const vec_t a = O::add(O::add(x0[v], x1[v]), x2[v]);
    const vec_t r = prf_vec<K>(sTl, kr, (uint64_t)v);
    const vec_t w0 = prf_vec<K>(sTl, k1, (uint64_t)v);
    const vec_t w2 = prf_vec<K>(sTl, k2, (uint64_t)v);
    const vec_t r_msb = O::shr(r, p.sh_msb);
    const vec_t r_top = O::shr(O::shl(r, 1), p.sh_top);
    // c = reveal(x_positive + r)
    const vec_t c = O::add(O::add(a, up), r);
    const vec_t c_top = O::shr(O::shl(c, 1), p.sh_top);
    const vec_t c_msb = O::shr(c, p.sh_msb);
    const vec_t ov = make_ulonglong2(r_msb.x ^ c_msb.x, r_msb.y ^ c_msb.y);   // both are 0 / 1 per element
    const vec_t y = O::sub(O::add(O::sub(c_top, r_top), O::shl(ov, p.sh_ov)), down);
    z0[v] = w0;
    z1[v] = O::sub(O::sub(y, w0), w2);
    z2[v] = w2;
  }
}

// one CTA per SM: the AES table takes the CTA's shared-memory window up to 0x20000 (aes.cuh)
inline int aes_grid_for(mb_ctx *ctx, int64_t nvec, int tpb) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((nvec + tpb - 1) / tpb, (int64_t)mb_aes_sms(ctx)));
}

}  // namespace

#define VP(t) ((vec_t *)(t)->dptr)
#define CVP(t) ((const vec_t *)(t)->dptr)

int k_rep_share(mb_ctx *ctx, const AesKey &key, int owner, const mb_tensor *x, mb_tensor *o0, mb_tensor *o1, mb_tensor *o2) {
  if (x->numel == 0) return MB_OK;
  mb_tensor *o[3] = {o0, o1, o2};
  mb_tensor *prf = o[owner], *diff = o[(owner + 1) % 3], *zero = o[(owner + 2) % 3];
  int64_t nvec = vec_count(x);
  switch (vec_kind(x->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, share_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, key, CVP(x), VP(prf), VP(diff), VP(zero), nvec, x->numel); break;
    case VK_RING64: AES_LAUNCH(ctx, share_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, key, CVP(x), VP(prf), VP(diff), VP(zero), nvec, x->numel); break;
    case VK_BIT: AES_LAUNCH(ctx, share_kernel<VK_BIT>, AES_TPB_LIGHT, nvec, key, CVP(x), VP(prf), VP(diff), VP(zero), nvec, x->numel); break;
  }
  MB_LAUNCH_CHECK_W(ctx, "share", 16.0 * nvec * 3, nvec, 0);
  return MB_OK;
}

int k_rep_share_outsider(mb_ctx *ctx, const AesKey &k0, const AesKey &k1, const mb_tensor *x, mb_tensor *o0, mb_tensor *o1,
                         mb_tensor *o2) {
  if (x->numel == 0) return MB_OK;
  int64_t nvec = vec_count(x);
  switch (vec_kind(x->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, share_outsider_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, k0, k1, CVP(x), VP(o0), VP(o1), VP(o2), nvec, x->numel); break;
    case VK_RING64: AES_LAUNCH(ctx, share_outsider_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, k0, k1, CVP(x), VP(o0), VP(o1), VP(o2), nvec, x->numel); break;
    case VK_BIT: AES_LAUNCH(ctx, share_outsider_kernel<VK_BIT>, AES_TPB_LIGHT, nvec, k0, k1, CVP(x), VP(o0), VP(o1), VP(o2), nvec, x->numel); break;
  }
  MB_LAUNCH_CHECK_W(ctx, "share_outsider", 16.0 * nvec * 4, 2 * nvec, 0);
  return MB_OK;
}

int k_reveal3(mb_ctx *ctx, const mb_tensor *a, const mb_tensor *b, const mb_tensor *c, mb_tensor *out) {
  if (out->numel == 0) return MB_OK;
  int64_t nvec = vec_count(out);
  int g = (int)((nvec + TPB - 1) / TPB);
  switch (vec_kind(out->dtype)) {
    case VK_RING128: reveal_kernel<VK_RING128><<<g, TPB, 0, ctx->stream>>>(CVP(a), CVP(b), CVP(c), VP(out), nvec); break;
    case VK_RING64: reveal_kernel<VK_RING64><<<g, TPB, 0, ctx->stream>>>(CVP(a), CVP(b), CVP(c), VP(out), nvec); break;
    case VK_BIT: reveal_kernel<VK_BIT><<<g, TPB, 0, ctx->stream>>>(CVP(a), CVP(b), CVP(c), VP(out), nvec); break;
  }
  MB_LAUNCH_CHECK_W(ctx, "reveal", 16.0 * nvec * 4, 0, 0);
  return MB_OK;
}

int k_rep_mul(mb_ctx *ctx, const AesKey seeds[3], const mb_rep *x, const mb_rep *y, mb_tensor *z[3]) {
  const mb_tensor *t = z[0];
  if (t->numel == 0) return MB_OK;
  int64_t nvec = vec_count(t);
#define ARGS seeds[0], seeds[1], seeds[2], CVP(x->s[0]), CVP(x->s[1]), CVP(x->s[2]), CVP(y->s[0]), CVP(y->s[1]), CVP(y->s[2]), \
             VP(z[0]), VP(z[1]), VP(z[2]), nvec, t->numel
  switch (vec_kind(t->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, rep_mul_kernel<VK_RING128>, AES_TPB_MID, nvec, ARGS); break;
    case VK_RING64: AES_LAUNCH(ctx, rep_mul_kernel<VK_RING64>, AES_TPB_MID, nvec, ARGS); break;
    case VK_BIT: AES_LAUNCH(ctx, rep_mul_kernel<VK_BIT>, AES_TPB_MID, nvec, ARGS); break;
  }
#undef ARGS
  MB_LAUNCH_CHECK_W(ctx, "rep_mul", 16.0 * nvec * 9, 3 * nvec, 0);
  return MB_OK;
}

int k_add_zero_share(mb_ctx *ctx, const AesKey seeds[3], mb_tensor *v[3], mb_tensor *z[3]) {
  const mb_tensor *t = z[0];
  if (t->numel == 0) return MB_OK;
  int64_t nvec = vec_count(t);
#define ARGS seeds[0], seeds[1], seeds[2], CVP(v[0]), CVP(v[1]), CVP(v[2]), VP(z[0]), VP(z[1]), VP(z[2]), nvec, t->numel
  switch (vec_kind(t->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, add_zero_share_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, ARGS); break;
    case VK_RING64: AES_LAUNCH(ctx, add_zero_share_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, ARGS); break;
    case VK_BIT: AES_LAUNCH(ctx, add_zero_share_kernel<VK_BIT>, AES_TPB_LIGHT, nvec, ARGS); break;
  }
#undef ARGS
  MB_LAUNCH_CHECK_W(ctx, "add_zero_share", 16.0 * nvec * 6, 3 * nvec, 0);
  return MB_OK;
}

int k_rep_trunc_pr(mb_ctx *ctx, const AesKey seeds[6], unsigned amount, const mb_rep *x, mb_tensor *z[3]) {
  const mb_tensor *t = z[0];
  if (t->numel == 0) return MB_OK;
  const unsigned L = t->dtype == MB_RING128 ? 128 : 64;
  const unsigned k = L - 1;
  TruncParams p;
  p.sh_msb = L - 1;
  p.sh_top = (amount + 1) & (L - 1);
  p.sh_ov = (k - amount) & (L - 1);
  u128 up = (u128)1 << ((k - 1) & (L - 1));
  u128 down = (u128)1 << ((k - amount - 1) & (L - 1));
  p.up_lo = (uint64_t)up; p.up_hi = (uint64_t)(up >> 64);
  p.down_lo = (uint64_t)down; p.down_hi = (uint64_t)(down >> 64);
  int64_t nvec = vec_count(t);
  // seeds[1..3] (the dealer's additive sharings of r, r_top, r_msb) cancel in the outputs: see trunc_pr_kernel
#define ARGS seeds[0], seeds[4], seeds[5], p, CVP(x->s[0]), CVP(x->s[1]), CVP(x->s[2]), VP(z[0]), VP(z[1]), VP(z[2]), nvec
  if (t->dtype == MB_RING128)
    AES_LAUNCH(ctx, trunc_pr_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, ARGS);
  else
    AES_LAUNCH(ctx, trunc_pr_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, ARGS);
#undef ARGS
  MB_LAUNCH_CHECK_W(ctx, "trunc_pr", 16.0 * nvec * 6, 3 * nvec, 0);
  return MB_OK;
}
