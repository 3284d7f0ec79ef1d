// Batched kernels for the replicated fixed-point math protocols: what the reference runs as
// hundreds of tiny replicated ops per call -- and round 1 launched one by one -- in a handful of launches.
//
//   prefix_or_level_kernel   one level of prefix_op (replicated/misc.rs:5-40) over bit planes: up to 32 independent
//                            OR(a, b) = a ^ b ^ AND(a, b), each AND a replicated multiplication over Z2 with its own
//                            zero share (replicated/arith.rs:317-367, zero_share.rs:82-156)
//   inject_sum_kernel        sum_i m_i * (ring_inject(z_i) << s_i) over up to 64 bit planes (top_most_index,
//                            replicated/division.rs:176-205; the higher bits of Pow2, replicated/exp.rs:50-70): every plane
//                            goes through RingInjectOp::rep_kernel with its own dabit and adt_to_rep seeds
//                            (replicated/convert.rs:316-390); shifts and sums are local, so only the sum is written
//   pow2_bits_kernel         Pow2FromBits (replicated/exp.rs:124-157): the first integral-precision higher bits are
//                            injected, turned into selectors and multiplied up in registers (a chain of replicated
//                            multiplications, one zero share each)
//
// Every PRF stream is the stream the op-by-op protocol uses for the same value (the host draws the keys in the
// reference's op order, csrc/rep_math.cpp), every value is computed with the same arithmetic mod 2^k, so the shares
// are identical bit for bit (tests/test_gpu_math.py).  Keys travel as kernel parameters (up to 25 KB).
#include "vecops.cuh"

namespace {

#define AES_LAUNCH(ctx, kern, tpb, items, ...)                                                                  \
  do {                                                                                                          \
    MB_CUDA((ctx), cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));    \
    const mb_aes_cfg_t _cfg = mb_aes_cfg((ctx), (items), (tpb));                                                \
    kern<<<_cfg.grid_dim, _cfg.block_dim, AES_SMEM_BYTES, (ctx)->stream>>>(__VA_ARGS__);                                  \
  } while (0)

inline int aes_grid(mb_ctx *ctx, int64_t items, int tpb) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((items + tpb - 1) / tpb, (int64_t)mb_aes_sms(ctx)));
}

__device__ __forceinline__ vec_t vx(vec_t a, vec_t b) { return make_ulonglong2(a.x ^ b.x, a.y ^ b.y); }
__device__ __forceinline__ vec_t va(vec_t a, vec_t b) { return make_ulonglong2(a.x & b.x, a.y & b.y); }

struct Planes3 {
  vec_t *s[3];
};

// W[dst] = OR(W[src], W[dst]) for the ops of one prefix level.  Ops of a level touch disjoint destinations and never
// read a plane another op of the level writes (misc.rs:14-24), so they run concurrently; a CTA walks (op, vector)
// items with the op index uniform per iteration block.
__global__ void __launch_bounds__(768) prefix_or_level_kernel(const __grid_constant__ PrefixLevel lv, Planes3 W, int64_t nv) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  const int64_t total = (int64_t)lv.nops * nv;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int op = (int)(t / nv);
    const int64_t v = t - (int64_t)op * nv;
    const int64_t ia = (int64_t)lv.src[op] * nv + v, ib = (int64_t)lv.dst[op] * nv + v;
    const vec_t a0 = W.s[0][ia], a1 = W.s[1][ia], a2 = W.s[2][ia];
    const vec_t b0 = W.s[0][ib], b1 = W.s[1][ib], b2 = W.s[2][ib];
    const vec_t r0 = prf_vec<VK_BIT>(L, lv.keys[op][0], (uint64_t)v);
    const vec_t r1 = prf_vec<VK_BIT>(L, lv.keys[op][1], (uint64_t)v);
    const vec_t r2 = prf_vec<VK_BIT>(L, lv.keys[op][2], (uint64_t)v);
    // AND: z_i = a_i b_i ^ a_i b_{i+1} ^ a_{i+1} b_i ^ r_i ^ r_{i+1};  OR = (a ^ b) ^ AND (share-wise)
    const vec_t z0 = vx(vx(va(a0, vx(b0, b1)), va(a1, b0)), vx(r0, r1));
    const vec_t z1 = vx(vx(va(a1, vx(b1, b2)), va(a2, b1)), vx(r1, r2));
    const vec_t z2 = vx(vx(va(a2, vx(b2, b0)), va(a0, b2)), vx(r2, r0));
    W.s[0][ib] = vx(vx(a0, b0), z0);
    W.s[1][ib] = vx(vx(a1, b1), z1);
    W.s[2][ib] = vx(vx(a2, b2), z2);
  }
}

// W[i] = plane idx[i] of the [k, n] bit tensors `src` (n % 128 == 0: planes are whole vectors).
struct GatherDesc {
  int nplanes;
  short idx[128];
};
__global__ void planes_gather_kernel(const __grid_constant__ GatherDesc d, Planes3 src, Planes3 W, int64_t nv) {
  const int64_t total = (int64_t)d.nplanes * nv;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int p = (int)(t / nv);
    const int64_t v = t - (int64_t)p * nv;
#pragma unroll
    for (int i = 0; i < 3; i++) W.s[i][t] = src.s[i][(int64_t)d.idx[p] * nv + v];
  }
}

struct Bits3 {
  const uint32_t *s[3];
};
struct Ring3 {
  vec_t *s[3];
};
struct CRing3 {
  const vec_t *s[3];
};

// The shared bit of element e in plane p (the XOR of the three shares: the value ring_inject opens and re-shares).
__device__ __forceinline__ uint32_t plane_word(const Bits3 &W, int p, int64_t n, int64_t e) {
  const int64_t w = ((int64_t)p * n + e) >> 5;
  return W.s[0][w] ^ W.s[1][w] ^ W.s[2][w];
}

// out = base + sign * sum_i mult_i * (inject(z_i) << shift_i),  z_i = plane pa_i ^ plane pb_i (pb_i < 0: none).
// inject(z) = (y0, z - y0 - y1, y1) with y0, y1 the adt_to_rep streams of that call (bits_fused.cu,
// ring_inject_fused_kernel).
#ifndef INJECT_SUM_TPB
#define INJECT_SUM_TPB AES_TPB_LIGHT
#endif
#ifndef INJECT_SUM_UNROLL
#define INJECT_SUM_UNROLL 1
#endif
template <int K>
__global__ void __launch_bounds__(INJECT_SUM_TPB) inject_sum_kernel(const __grid_constant__ InjectSumDesc d, Bits3 W, int64_t n,
                                                                 CRing3 base, Ring3 out, int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  constexpr int EPV = K == VK_RING128 ? 1 : 2;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = v * EPV;
    const unsigned sh = (unsigned)(e & 31);
    vec_t acc0 = O::zero(), acc1 = O::zero(), acc2 = O::zero();
    constexpr int UNROLL = INJECT_SUM_UNROLL;
#pragma unroll UNROLL
    for (int i = 0; i < d.nplanes; i++) {
      if (d.mult[i] == 0) continue;   // the value is never used (its keys were still drawn, in order)
      uint32_t w = plane_word(W, d.pa[i], n, e);
      if (d.pb[i] >= 0) w ^= plane_word(W, d.pb[i], n, e);
      vec_t s;
      if (K == VK_RING128) s = make_ulonglong2((w >> sh) & 1u, 0);
      else s = make_ulonglong2((w >> sh) & 1u, (w >> (sh + 1)) & 1u);
      const vec_t y0 = prf_vec<K>(L, d.keys[i][0], (uint64_t)v);
      const vec_t y1 = prf_vec<K>(L, d.keys[i][1], (uint64_t)v);
      const vec_t c = O::sub(O::sub(s, y0), y1);
      for (int m = 0; m < d.mult[i]; m++) {
        acc0 = O::add(acc0, O::shl(y0, d.shift[i]));
        acc1 = O::add(acc1, O::shl(c, d.shift[i]));
        acc2 = O::add(acc2, O::shl(y1, d.shift[i]));
      }
    }
    if (d.negate) {
      acc0 = O::neg(acc0);
      acc1 = O::neg(acc1);
      acc2 = O::neg(acc2);
    }
    if (base.s[0]) {
      acc0 = O::add(base.s[0][v], acc0);
      acc1 = O::add(base.s[1][v], acc1);
      acc2 = O::add(base.s[2][v], acc2);
    }
    out.s[0][v] = acc0;
    out.s[1][v] = acc1;
    out.s[2][v] = acc2;
  }
}

// Pow2 (replicated/exp.rs:50-80, 124-157) on the first `count` higher bits of |x| (planes first_plane .. of `W`):
//   hb_i  = ring_inject(plane_i)                                  (keys.inject[i])
//   part  = sum_i hb_i << (f + i)                                 (the integral part collected so far)
//   sel_i = (hb_i << (2^i mod width)) + (ones - hb_i),  ones = (1, 0, 0)
//   d     = ones * sel_0 * sel_1 * ... (replicated multiplications, zero share keys.zero[i])
// Outputs: d and rest = x_abs - part.
template <int K>
__global__ void __launch_bounds__(768) pow2_bits_kernel(const __grid_constant__ Pow2Desc d, Bits3 W, int64_t n, CRing3 x_abs,
                                                                Ring3 out_d, Ring3 out_rest, int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  constexpr int EPV = K == VK_RING128 ? 1 : 2;
  const vec_t one = O::splat(1, 0);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = v * EPV;
    const unsigned sh = (unsigned)(e & 31);
    vec_t p0 = O::zero(), p1 = O::zero(), p2 = O::zero();      // part
    vec_t a0 = one, a1 = O::zero(), a2 = O::zero();            // running product, starts as ones = (1, 0, 0)
    for (int i = 0; i < d.count; i++) {
      const uint32_t w = plane_word(W, d.first_plane + i, n, e);
      vec_t s;
      if (K == VK_RING128) s = make_ulonglong2((w >> sh) & 1u, 0);
      else s = make_ulonglong2((w >> sh) & 1u, (w >> (sh + 1)) & 1u);
      const vec_t h0 = prf_vec<K>(L, d.inject[i][0], (uint64_t)v);
      const vec_t h2 = prf_vec<K>(L, d.inject[i][1], (uint64_t)v);
      const vec_t h1 = O::sub(O::sub(s, h0), h2);
      p0 = O::add(p0, O::shl(h0, d.part_shift[i]));
      p1 = O::add(p1, O::shl(h1, d.part_shift[i]));
      p2 = O::add(p2, O::shl(h2, d.part_shift[i]));
      // selector: pos + (ones - hb)
      const vec_t b0 = O::add(O::shl(h0, d.sel_shift[i]), O::sub(one, h0));
      const vec_t b1 = O::add(O::shl(h1, d.sel_shift[i]), O::neg(h1));
      const vec_t b2 = O::add(O::shl(h2, d.sel_shift[i]), O::neg(h2));
      const vec_t r0 = prf_vec<K>(L, d.zero[i][0], (uint64_t)v);
      const vec_t r1 = prf_vec<K>(L, d.zero[i][1], (uint64_t)v);
      const vec_t r2 = prf_vec<K>(L, d.zero[i][2], (uint64_t)v);
      const vec_t z0 = O::add(O::add(O::mul(a0, O::add(b0, b1)), O::mul(a1, b0)), O::sub(r0, r1));
      const vec_t z1 = O::add(O::add(O::mul(a1, O::add(b1, b2)), O::mul(a2, b1)), O::sub(r1, r2));
      const vec_t z2 = O::add(O::add(O::mul(a2, O::add(b2, b0)), O::mul(a0, b2)), O::sub(r2, r0));
      a0 = z0; a1 = z1; a2 = z2;
    }
    out_d.s[0][v] = a0;
    out_d.s[1][v] = a1;
    out_d.s[2][v] = a2;
    out_rest.s[0][v] = O::sub(x_abs.s[0][v], p0);
    out_rest.s[1][v] = O::sub(x_abs.s[1][v], p1);
    out_rest.s[2][v] = O::sub(x_abs.s[2][v], p2);
  }
}

// ---- PolynomialEval (fixedpoint/mod.rs:95-151) ------------------------------------------------------------------------
struct TruncP {
  unsigned sh_msb, sh_top, sh_ov;
  uint64_t up_lo, up_hi, down_lo, down_hi;
};
inline TruncP make_trunc_p(mb_dtype dt, unsigned amount) {   // same constants as k_rep_trunc_pr (rep.cu)
  const unsigned Lb = dt == MB_RING128 ? 128 : 64, k = Lb - 1;
  TruncP p;
  p.sh_msb = Lb - 1;
  p.sh_top = (amount + 1) & (Lb - 1);
  p.sh_ov = (k - amount) & (Lb - 1);
  const u128 up = (u128)1 << ((k - 1) & (Lb - 1)), down = (u128)1 << ((k - amount - 1) & (Lb - 1));
  p.up_lo = (uint64_t)up; p.up_hi = (uint64_t)(up >> 64);
  p.down_lo = (uint64_t)down; p.down_hi = (uint64_t)(down >> 64);
  return p;
}
// TruncPr of the replicated value (a0, a1, a2) with the three surviving streams (r, w0, w2): see trunc_pr_kernel (rep.cu)
template <int K>
__device__ __forceinline__ void trunc3(const TruncP &p, vec_t up, vec_t down, vec_t a0, vec_t a1, vec_t a2, vec_t r, vec_t w0, vec_t w2,
                                       vec_t &z1) {
  typedef VOps<K> O;
  const vec_t a = O::add(O::add(a0, a1), a2);
  const vec_t r_msb = O::shr(r, p.sh_msb);
  const vec_t r_top = O::shr(O::shl(r, 1), p.sh_top);
  const vec_t c = O::add(O::add(a, up), r);
  const vec_t c_top = O::shr(O::shl(c, 1), p.sh_top);
  const vec_t c_msb = O::shr(c, p.sh_msb);
  const vec_t ov = make_ulonglong2(r_msb.x ^ c_msb.x, r_msb.y ^ c_msb.y);
  const vec_t y = O::sub(O::add(O::sub(c_top, r_top), O::shl(ov, p.sh_ov)), down);
  z1 = O::sub(O::sub(y, w0), w2);
}

// One batch of a prefix-mul level: P[dst] = trunc_pr(P[src] * P[dst]) for up to 16 ops (keys: zero share of the
// multiplication -- unused, see below -- then (r, y0, y1) of the truncation).  P is a [slots, n] ring buffer per share.
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) mul_trunc_level_kernel(const __grid_constant__ MulTruncLevel lv, TruncP tp, Ring3 P, int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  const vec_t up = O::splat(tp.up_lo, tp.up_hi), down = O::splat(tp.down_lo, tp.down_hi);
  const int64_t total = (int64_t)lv.nops * nvec;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int op = (int)(t / nvec);
    const int64_t v = t - (int64_t)op * nvec;
    const int64_t ia = (int64_t)lv.src[op] * nvec + v, ib = (int64_t)lv.dst[op] * nvec + v;
    const vec_t a0 = P.s[0][ia], a1 = P.s[1][ia], a2 = P.s[2][ia];
    const vec_t b0 = P.s[0][ib], b1 = P.s[1][ib], b2 = P.s[2][ib];
    // The product's zero share (keys[op][0..2]) sums to zero over the three shares and the truncation opens only that
    // sum, so its streams are not expanded (their entropy requests were drawn, in order).
    const vec_t m0 = O::add(O::mul(a0, O::add(b0, b1)), O::mul(a1, b0));
    const vec_t m1 = O::add(O::mul(a1, O::add(b1, b2)), O::mul(a2, b1));
    const vec_t m2 = O::add(O::mul(a2, O::add(b2, b0)), O::mul(a0, b2));
    const vec_t r = prf_vec<K>(L, lv.keys[op][3], (uint64_t)v);
    const vec_t w0 = prf_vec<K>(L, lv.keys[op][4], (uint64_t)v);
    const vec_t w2 = prf_vec<K>(L, lv.keys[op][5], (uint64_t)v);
    vec_t z1;
    trunc3<K>(tp, up, down, m0, m1, m2, r, w0, w2, z1);
    P.s[0][ib] = w0;
    P.s[1][ib] = z1;
    P.s[2][ib] = w2;
  }
}

// trunc_pr(x * y) on separate tensors (mb_rep_fixed_mul): the product's zero share cancels in the opened sum
template <int K>
__global__ void __launch_bounds__(AES_TPB_LIGHT) mul_trunc_kernel(AesKey kr, AesKey k1, AesKey k2, TruncP tp, CRing3 x, CRing3 y, Ring3 out,
                                                                int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  const vec_t up = O::splat(tp.up_lo, tp.up_hi), down = O::splat(tp.down_lo, tp.down_hi);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const vec_t a0 = x.s[0][v], a1 = x.s[1][v], a2 = x.s[2][v];
    const vec_t b0 = y.s[0][v], b1 = y.s[1][v], b2 = y.s[2][v];
    const vec_t m0 = O::add(O::mul(a0, O::add(b0, b1)), O::mul(a1, b0));
    const vec_t m1 = O::add(O::mul(a1, O::add(b1, b2)), O::mul(a2, b1));
    const vec_t m2 = O::add(O::mul(a2, O::add(b2, b0)), O::mul(a0, b2));
    const vec_t r = prf_vec<K>(L, kr, (uint64_t)v);
    const vec_t w0 = prf_vec<K>(L, k1, (uint64_t)v);
    const vec_t w2 = prf_vec<K>(L, k2, (uint64_t)v);
    vec_t z1;
    trunc3<K>(tp, up, down, m0, m1, m2, r, w0, w2, z1);
    out.s[0][v] = w0;
    out.s[1][v] = z1;
    out.s[2][v] = w2;
  }
}

// The select of MaximumOp (replicated/softmax.rs:40-50) after the comparison bit: with l = ring_inject(lesser, 0)
//   out = l * m2 + (ones - l) * m1          (two replicated multiplications, a public subtraction, an addition)
// in one pass: reads the bit shares and the six ring shares, writes three; 2 + 3 + 3 PRF streams.
//   l = (y0, s - y0 - y1, y1)   [s = the shared bit]      om = (1 - l_0, -l_1, -l_2)      (arith.rs:230-273)
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) maximum_select_kernel(AesKey ky0, AesKey ky1, AesKey a0k, AesKey a1k, AesKey a2k, AesKey b0k,
                                                                     AesKey b1k, AesKey b2k, Bits3 lesser, CRing3 m1, CRing3 m2, Ring3 out,
                                                                     int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  constexpr int EPV = K == VK_RING128 ? 1 : 2;
  const vec_t one = O::splat(1, 0);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = v * EPV;
    const uint32_t w = plane_word(lesser, 0, 0, e);
    const unsigned sh = (unsigned)(e & 31);
    vec_t s;
    if (K == VK_RING128) s = make_ulonglong2((w >> sh) & 1u, 0);
    else s = make_ulonglong2((w >> sh) & 1u, (w >> (sh + 1)) & 1u);
    const vec_t l0 = prf_vec<K>(L, ky0, (uint64_t)v);
    const vec_t l2 = prf_vec<K>(L, ky1, (uint64_t)v);
    const vec_t l1 = O::sub(O::sub(s, l0), l2);
    const vec_t p0 = m2.s[0][v], p1 = m2.s[1][v], p2 = m2.s[2][v];
    const vec_t q0 = m1.s[0][v], q1 = m1.s[1][v], q2 = m1.s[2][v];
    const vec_t ra0 = prf_vec<K>(L, a0k, (uint64_t)v), ra1 = prf_vec<K>(L, a1k, (uint64_t)v), ra2 = prf_vec<K>(L, a2k, (uint64_t)v);
    // a = l * m2
    const vec_t a0 = O::add(O::add(O::mul(l0, O::add(p0, p1)), O::mul(l1, p0)), O::sub(ra0, ra1));
    const vec_t a1 = O::add(O::add(O::mul(l1, O::add(p1, p2)), O::mul(l2, p1)), O::sub(ra1, ra2));
    const vec_t a2 = O::add(O::add(O::mul(l2, O::add(p2, p0)), O::mul(l0, p2)), O::sub(ra2, ra0));
    // b = (ones - l) * m1
    const vec_t n0 = O::sub(one, l0), n1 = O::neg(l1), n2 = O::neg(l2);
    const vec_t rb0 = prf_vec<K>(L, b0k, (uint64_t)v), rb1 = prf_vec<K>(L, b1k, (uint64_t)v), rb2 = prf_vec<K>(L, b2k, (uint64_t)v);
    const vec_t b0 = O::add(O::add(O::mul(n0, O::add(q0, q1)), O::mul(n1, q0)), O::sub(rb0, rb1));
    const vec_t b1 = O::add(O::add(O::mul(n1, O::add(q1, q2)), O::mul(n2, q1)), O::sub(rb1, rb2));
    const vec_t b2 = O::add(O::add(O::mul(n2, O::add(q2, q0)), O::mul(n0, q2)), O::sub(rb2, rb0));
    out.s[0][v] = O::add(a0, b0);
    out.s[1][v] = O::add(a1, b1);
    out.s[2][v] = O::add(a2, b2);
  }
}

// P[i] = x for i < slots (the `degree` copies of x that prefix_op starts from)
__global__ void replicate_kernel(CRing3 x, Ring3 P, int slots, int64_t nvec) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const vec_t a = x.s[i][v];
      for (int j = 0; j < slots; j++) P.s[i][(int64_t)j * nvec + v] = a;
    }
  }
}

// out = trunc_pr(sum_i P[i] * c_{i+1}) + c_0 (public constants; c_0 lands on share 0)
template <int K>
__global__ void __launch_bounds__(AES_TPB_MID) poly_combine_kernel(const __grid_constant__ PolyCombine pc, TruncP tp, Ring3 P, Ring3 out,
                                                                   int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  typedef VOps<K> O;
  const vec_t up = O::splat(tp.up_lo, tp.up_hi), down = O::splat(tp.down_lo, tp.down_hi);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    vec_t s0 = O::zero(), s1 = O::zero(), s2 = O::zero();
    for (int i = 0; i < pc.degree; i++) {
      const vec_t cf = O::splat(pc.coeff_lo[i + 1], pc.coeff_hi[i + 1]);
      const int64_t ix = (int64_t)i * nvec + v;
      s0 = O::add(s0, O::mul(P.s[0][ix], cf));
      s1 = O::add(s1, O::mul(P.s[1][ix], cf));
      s2 = O::add(s2, O::mul(P.s[2][ix], cf));
    }
    const vec_t r = prf_vec<K>(L, pc.keys[0], (uint64_t)v);
    const vec_t w0 = prf_vec<K>(L, pc.keys[1], (uint64_t)v);
    const vec_t w2 = prf_vec<K>(L, pc.keys[2], (uint64_t)v);
    vec_t z1;
    trunc3<K>(tp, up, down, s0, s1, s2, r, w0, w2, z1);
    out.s[0][v] = O::add(w0, O::splat(pc.coeff_lo[0], pc.coeff_hi[0]));
    out.s[1][v] = z1;
    out.s[2][v] = w2;
  }
}

}  // namespace

int k_planes_gather(mb_ctx *ctx, const mb_tensor *const src[3], const short *idx, int nplanes, mb_tensor *const W[3], int64_t n) {
  if (nplanes < 1 || nplanes > 128 || n % 128) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "planes_gather: bad plane count / ragged planes");
  GatherDesc d;
  d.nplanes = nplanes;
  for (int i = 0; i < nplanes; i++) d.idx[i] = idx[i];
  const int64_t nv = n / 128, total = nv * nplanes;
  Planes3 s, w;
  for (int i = 0; i < 3; i++) { s.s[i] = (vec_t *)src[i]->dptr; w.s[i] = (vec_t *)W[i]->dptr; }
  planes_gather_kernel<<<(unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(d, s, w, nv);
  MB_LAUNCH_CHECK_W(ctx, "planes_gather", 16.0 * total * 6, 0, 0);
  return MB_OK;
}

int k_prefix_or_level(mb_ctx *ctx, const PrefixLevel &lv, mb_tensor *const W[3], int64_t n) {
  if (lv.nops < 1 || lv.nops > MB_PREFIX_MAX_OPS || n % 128) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "prefix_or_level: bad level");
  const int64_t nv = n / 128;
  Planes3 w;
  for (int i = 0; i < 3; i++) w.s[i] = (vec_t *)W[i]->dptr;
  const int64_t total = nv * lv.nops;
  AES_LAUNCH(ctx, prefix_or_level_kernel, 768, total, lv, w, nv);
  MB_LAUNCH_CHECK_W(ctx, "prefix_or_level", 16.0 * total * 9, 3.0 * total, 0);
  return MB_OK;
}

int k_inject_sum(mb_ctx *ctx, const InjectSumDesc &d, const mb_tensor *const W[3], int64_t n, const mb_rep *base, mb_tensor *const out[3]) {
  if (d.nplanes < 1 || d.nplanes > MB_INJECT_MAX_PLANES || n % 128) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "inject_sum: bad descriptor");
  const mb_tensor *t = out[0];
  const int64_t nvec = vec_count(t);
  Bits3 w;
  Ring3 o;
  CRing3 b;
  for (int i = 0; i < 3; i++) {
    w.s[i] = (const uint32_t *)W[i]->dptr;
    o.s[i] = (vec_t *)out[i]->dptr;
    b.s[i] = base ? (const vec_t *)base->s[i]->dptr : nullptr;
  }
  int live = 0;
  for (int i = 0; i < d.nplanes; i++) live += d.mult[i] != 0;
  if (t->dtype == MB_RING128)
    AES_LAUNCH(ctx, inject_sum_kernel<VK_RING128>, INJECT_SUM_TPB, nvec, d, w, n, b, o, nvec);
  else
    AES_LAUNCH(ctx, inject_sum_kernel<VK_RING64>, INJECT_SUM_TPB, nvec, d, w, n, b, o, nvec);
  MB_LAUNCH_CHECK_W(ctx, "inject_sum", 16.0 * nvec * (3 + (base ? 3 : 0)) + 3.0 * live * n / 8, 2.0 * live * nvec, 0);
  return MB_OK;
}

int k_pow2_bits(mb_ctx *ctx, const Pow2Desc &d, const mb_tensor *const W[3], int64_t n, const mb_rep *x_abs, mb_tensor *const out_d[3],
                mb_tensor *const out_rest[3]) {
  if (d.count < 1 || d.count > MB_POW2_MAX_BITS || n % 128) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "pow2_bits: bad descriptor");
  const mb_tensor *t = out_d[0];
  const int64_t nvec = vec_count(t);
  Bits3 w;
  Ring3 od, orr;
  CRing3 xa;
  for (int i = 0; i < 3; i++) {
    w.s[i] = (const uint32_t *)W[i]->dptr;
    od.s[i] = (vec_t *)out_d[i]->dptr;
    orr.s[i] = (vec_t *)out_rest[i]->dptr;
    xa.s[i] = (const vec_t *)x_abs->s[i]->dptr;
  }
  if (t->dtype == MB_RING128)
    AES_LAUNCH(ctx, pow2_bits_kernel<VK_RING128>, 768, nvec, d, w, n, xa, od, orr, nvec);
  else
    AES_LAUNCH(ctx, pow2_bits_kernel<VK_RING64>, 768, nvec, d, w, n, xa, od, orr, nvec);
  MB_LAUNCH_CHECK_W(ctx, "pow2_bits", 16.0 * nvec * 9 + 3.0 * d.count * n / 8, 5.0 * d.count * nvec, 0);
  return MB_OK;
}

int k_replicate(mb_ctx *ctx, const mb_rep *x, mb_tensor *const P[3], int slots) {
  const int64_t nvec = vec_count(x->s[0]);
  CRing3 xs;
  Ring3 ps;
  for (int i = 0; i < 3; i++) { xs.s[i] = (const vec_t *)x->s[i]->dptr; ps.s[i] = (vec_t *)P[i]->dptr; }
  replicate_kernel<<<(unsigned)std::min<int64_t>((nvec + 255) / 256, (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(xs, ps, slots, nvec);
  MB_LAUNCH_CHECK_W(ctx, "replicate", 16.0 * nvec * 3 * (1 + slots), 0, 0);
  return MB_OK;
}

int k_mul_trunc_level(mb_ctx *ctx, const MulTruncLevel &lv, unsigned amount, mb_tensor *const P[3], int64_t nvec) {
  if (lv.nops < 1 || lv.nops > MB_MULTRUNC_MAX_OPS) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mul_trunc_level: bad level");
  const mb_dtype dt = P[0]->dtype;
  const TruncP tp = make_trunc_p(dt, amount);
  Ring3 ps;
  for (int i = 0; i < 3; i++) ps.s[i] = (vec_t *)P[i]->dptr;
  const int64_t total = nvec * lv.nops;
  if (dt == MB_RING128)
    AES_LAUNCH(ctx, mul_trunc_level_kernel<VK_RING128>, AES_TPB_LIGHT, total, lv, tp, ps, nvec);
  else
    AES_LAUNCH(ctx, mul_trunc_level_kernel<VK_RING64>, AES_TPB_LIGHT, total, lv, tp, ps, nvec);
  MB_LAUNCH_CHECK_W(ctx, "mul_trunc_level", 16.0 * total * 9, 3.0 * total, 0);
  return MB_OK;
}

int k_poly_combine(mb_ctx *ctx, const PolyCombine &pc, unsigned amount, mb_tensor *const P[3], mb_tensor *const out[3]) {
  if (pc.degree < 1 || pc.degree >= MB_POLY_MAX_COEFFS) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "poly_combine: bad degree");
  const mb_dtype dt = out[0]->dtype;
  const TruncP tp = make_trunc_p(dt, amount);
  const int64_t nvec = vec_count(out[0]);
  Ring3 ps, os;
  for (int i = 0; i < 3; i++) { ps.s[i] = (vec_t *)P[i]->dptr; os.s[i] = (vec_t *)out[i]->dptr; }
  if (dt == MB_RING128)
    AES_LAUNCH(ctx, poly_combine_kernel<VK_RING128>, AES_TPB_MID, nvec, pc, tp, ps, os, nvec);
  else
    AES_LAUNCH(ctx, poly_combine_kernel<VK_RING64>, AES_TPB_MID, nvec, pc, tp, ps, os, nvec);
  MB_LAUNCH_CHECK_W(ctx, "poly_combine", 16.0 * nvec * 3 * (pc.degree + 1), 3.0 * nvec, 0);
  return MB_OK;
}

int k_mul_trunc(mb_ctx *ctx, const AesKey tk[3], unsigned amount, const mb_rep *x, const mb_rep *y, mb_tensor *const out[3]) {
  const mb_dtype dt = out[0]->dtype;
  const TruncP tp = make_trunc_p(dt, amount);
  const int64_t nvec = vec_count(out[0]);
  CRing3 xs, ys;
  Ring3 os;
  for (int i = 0; i < 3; i++) { xs.s[i] = (const vec_t *)x->s[i]->dptr; ys.s[i] = (const vec_t *)y->s[i]->dptr; os.s[i] = (vec_t *)out[i]->dptr; }
  if (dt == MB_RING128)
    AES_LAUNCH(ctx, mul_trunc_kernel<VK_RING128>, AES_TPB_LIGHT, nvec, tk[0], tk[1], tk[2], tp, xs, ys, os, nvec);
  else
    AES_LAUNCH(ctx, mul_trunc_kernel<VK_RING64>, AES_TPB_LIGHT, nvec, tk[0], tk[1], tk[2], tp, xs, ys, os, nvec);
  MB_LAUNCH_CHECK_W(ctx, "mul_trunc", 16.0 * nvec * 9, 3.0 * nvec, 0);
  return MB_OK;
}

// keys: inject (y0, y1), zero share of l * m2 (3), zero share of (ones - l) * m1 (3)
int k_maximum_select(mb_ctx *ctx, const AesKey keys[8], const mb_rep *lesser, const mb_rep *m1, const mb_rep *m2, mb_tensor *const out[3]) {
  const mb_tensor *t = out[0];
  if (t->numel == 0) return MB_OK;
  const int64_t nvec = vec_count(t);
  Bits3 lb;
  CRing3 a, b;
  Ring3 o;
  for (int i = 0; i < 3; i++) {
    lb.s[i] = (const uint32_t *)lesser->s[i]->dptr;
    a.s[i] = (const vec_t *)m1->s[i]->dptr;
    b.s[i] = (const vec_t *)m2->s[i]->dptr;
    o.s[i] = (vec_t *)out[i]->dptr;
  }
#define ARGS keys[0], keys[1], keys[2], keys[3], keys[4], keys[5], keys[6], keys[7], lb, a, b, o, nvec
  if (t->dtype == MB_RING128)
    AES_LAUNCH(ctx, maximum_select_kernel<VK_RING128>, AES_TPB_MID, nvec, ARGS);
  else
    AES_LAUNCH(ctx, maximum_select_kernel<VK_RING64>, AES_TPB_MID, nvec, ARGS);
#undef ARGS
  MB_LAUNCH_CHECK_W(ctx, "maximum_select", 16.0 * nvec * 9 + 3.0 * t->numel / 8, 8.0 * nvec, 0);
  return MB_OK;
}
