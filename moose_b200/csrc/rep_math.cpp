// C ABI: the replicated fixed-point math protocols of the LogReg path:
//   GreaterOp (replicated/compare.rs:136-149), ExpOp / Pow2Op / pow2_from_bits / exp_from_parts
//   (replicated/exp.rs), PolynomialEval / PrefixMul (fixedpoint/mod.rs:74-151), prefix_op / prefix_or
//   (replicated/misc.rs:5-40), DivOp / approximate_reciprocal / norm / top_most(_index) / sign_from_msb
//   (replicated/division.rs), SigmoidOp (replicated/arith.rs:676-743).
// These are pure orchestration in the reference (hundreds of replicated ops per call); here they replay the
// same op sequence on the fused replicated entry points, so the entropy order -- and with it every share --
// matches the op-by-op protocol.  No arithmetic happens on the host.
#include "common.h"

#include <cstdlib>

namespace {

struct Err { int code; };
inline void ck(int rc) { if (rc != MB_OK) throw Err{rc}; }

struct R {  // owning replicated tensor
  mb_ctx *c;
  mb_rep *p;
  R(mb_ctx *c_, mb_rep *p_) : c(c_), p(p_) {}
  R(R &&o) noexcept : c(o.c), p(o.p) { o.p = nullptr; }
  R &operator=(R &&o) noexcept { if (this != &o) { if (p) mb_rep_free(c, p); c = o.c; p = o.p; o.p = nullptr; } return *this; }
  R(const R &) = delete;
  R &operator=(const R &) = delete;
  ~R() { if (p) mb_rep_free(c, p); }
  R clone() const { mb_rep *q; ck(mb_rep_from_shares(c, p->s[0], p->s[1], p->s[2], &q)); return R(c, q); }
  const mb_tensor *like() const { return p->s[0]; }
  mb_rep *release() { mb_rep *q = p; p = nullptr; return q; }
};
struct T {  // owning public tensor
  mb_ctx *c;
  mb_tensor *p;
  T(mb_ctx *c_, mb_tensor *p_) : c(c_), p(p_) {}
  T(T &&o) noexcept : c(o.c), p(o.p) { o.p = nullptr; }
  T(const T &) = delete;
  T &operator=(const T &) = delete;
  ~T() { if (p) mb_tensor_free(c, p); }
};

#define REP1(name, call) \
  R name { mb_rep *q = nullptr; ck(call); return R(c, q); }

R add(mb_ctx *c, const R &a, const R &b) { mb_rep *q = nullptr; ck(mb_rep_add(c, a.p, b.p, &q)); return R(c, q); }
R sub(mb_ctx *c, const R &a, const R &b) { mb_rep *q = nullptr; ck(mb_rep_sub(c, a.p, b.p, &q)); return R(c, q); }
R mul(mb_ctx *c, const R &a, const R &b) { mb_rep *q = nullptr; ck(mb_rep_mul(c, a.p, b.p, &q)); return R(c, q); }
R neg(mb_ctx *c, const R &a) { mb_rep *q = nullptr; ck(mb_rep_neg(c, a.p, &q)); return R(c, q); }
R shl(mb_ctx *c, uint64_t amount, const R &a) { mb_rep *q = nullptr; ck(mb_rep_shl(c, amount, a.p, &q)); return R(c, q); }
R trunc(mb_ctx *c, uint32_t amount, const R &a) { mb_rep *q = nullptr; ck(mb_rep_trunc_pr(c, amount, a.p, &q)); return R(c, q); }
R bitdec(mb_ctx *c, const R &a) { mb_rep *q = nullptr; ck(mb_rep_bit_decompose(c, a.p, &q)); return R(c, q); }
R index0(mb_ctx *c, int64_t i, const R &a) { mb_rep *q = nullptr; ck(mb_rep_index_axis0(c, i, a.p, &q)); return R(c, q); }
R inject(mb_ctx *c, mb_dtype ring, uint64_t idx, const R &a) { mb_rep *q = nullptr; ck(mb_rep_ring_inject(c, ring, idx, a.p, &q)); return R(c, q); }
R mux(mb_ctx *c, const R &s, const R &x, const R &y) { mb_rep *q = nullptr; ck(mb_rep_mux(c, s.p, x.p, y.p, &q)); return R(c, q); }
R msb_bit(mb_ctx *c, const R &a) { mb_rep *q = nullptr; ck(mb_rep_msb(c, a.p, &q)); return R(c, q); }
R add_pub(mb_ctx *c, const R &a, const T &pub) { mb_rep *q = nullptr; ck(mb_rep_add_public(c, a.p, pub.p, &q)); return R(c, q); }
R pub_sub(mb_ctx *c, const T &pub, const R &a) { mb_rep *q = nullptr; ck(mb_rep_public_sub(c, pub.p, a.p, &q)); return R(c, q); }
R mul_pub(mb_ctx *c, const R &a, const T &pub) { mb_rep *q = nullptr; ck(mb_rep_mul_public(c, a.p, pub.p, &q)); return R(c, q); }

mb_dtype ring_of(const R &x) { return x.like()->dtype; }
int bits_of(const R &x) { return ring_of(x) == MB_RING128 ? 128 : 64; }

// Constant::Fixed -> ring value (kernels/constants.rs:68-72,103-105): ((v * 2^p as f64) as iK) as uK.
u128 fixed_constant(int bits, double value, unsigned precision) {
  const double scaled = value * ldexp(1.0, (int)precision);
  u128 v;
  const double lim = ldexp(1.0, bits - 1);
  if (scaled != scaled) v = 0;
  else if (scaled >= lim) v = ((u128)1 << (bits - 1)) - 1;
  else if (scaled <= -lim) v = (u128)1 << (bits - 1);
  else v = (u128)(__int128)scaled;
  if (bits == 64) v &= (u128)0xFFFFFFFFFFFFFFFFull;
  return v;
}

// Mirrored (public) constant with the shape of `like` (ShapeFill, misc.rs:136-170).
T pub_fill(mb_ctx *c, const mb_tensor *like, u128 v) {
  mb_tensor *t = nullptr;
  ck(mb_fill(c, like->dtype, like->rank, like->shape, (uint64_t)v, (uint64_t)(v >> 64), &t));
  return T(c, t);
}
// FillOp::rep_ring64/128_kernel (replicated/ops.rs:91-160): shares (v, 0, 0).
R rep_fill(mb_ctx *c, const mb_tensor *like, u128 v) {
  mb_tensor *a = nullptr, *z1 = nullptr, *z2 = nullptr;
  ck(mb_fill(c, like->dtype, like->rank, like->shape, (uint64_t)v, (uint64_t)(v >> 64), &a));
  T ta(c, a);
  ck(mb_fill(c, like->dtype, like->rank, like->shape, 0, 0, &z1));
  T t1(c, z1);
  ck(mb_fill(c, like->dtype, like->rank, like->shape, 0, 0, &z2));
  T t2(c, z2);
  mb_rep *q = nullptr;
  ck(mb_rep_from_shares(c, a, z1, z2, &q));
  return R(c, q);
}

R greater(mb_ctx *c, const R &x, const R &y) { return msb_bit(c, sub(c, y, x)); }  // msb(y - x), compare.rs:136-149

R msb_ring(mb_ctx *c, const R &x) { return inject(c, ring_of(x), 0, msb_bit(c, x)); }  // arith.rs:628-641

// prefix_op (misc.rs:5-28) with op(res[y], res[y+k])
template <typename Op>
void prefix_op(std::vector<R> &res, Op op) {
  const int v_len = (int)res.size();
  int log_r = 0;
  while ((1 << log_r) < v_len) log_r++;
  for (int i = 0; i < log_r; i++)
    for (int j = 0; j < (1 << log_r) / (1 << (i + 1)); j++) {
      const int y = (1 << i) + j * (1 << (i + 1)) - 1;
      for (int k = 1; k < (1 << i) + 1; k++)
        if (y + k < v_len) res[y + k] = op(res[y], res[y + k]);
    }
}

// division.rs:88-104
R sign_from_msb(mb_ctx *c, const R &msb) {
  R dbl = shl(c, 1, msb);
  T ones = pub_fill(c, msb.like(), 1);
  return pub_sub(c, ones, dbl);
}

// Owning triple of freshly allocated tensors (working planes / results of the batched kernels).
struct T3 {
  mb_ctx *c;
  mb_tensor *t[3];
  explicit T3(mb_ctx *c_) : c(c_) { t[0] = t[1] = t[2] = nullptr; }
  ~T3() { for (int i = 0; i < 3; i++) if (t[i]) mb_tensor_free(c, t[i]); }
  T3(const T3 &) = delete;
  T3 &operator=(const T3 &) = delete;
  void alloc(mb_dtype d, int rank, const int64_t *shape) { for (int i = 0; i < 3; i++) ck(mb_tensor_new(c, d, rank, shape, &t[i])); }
  R to_rep() {
    mb_rep *q = nullptr;
    ck(mb_rep_from_shares(c, t[0], t[1], t[2], &q));   // retains the tensors; ours are dropped by the destructor
    return R(c, q);
  }
};

// The batched kernels (math_fused.cu) work on whole 128-bit plane vectors and on the three-launch-free layouts.
bool batched_ok(mb_ctx *c, const R &x) {
  static const bool off = getenv("MOOSE_B200_NO_BATCHED_MATH") != nullptr;
  const mb_tensor *t = x.like();
  return !off && c->fused_bits && mb_is_ring(t->dtype) && t->numel > 0 && t->numel % 128 == 0 &&
         mb_same_shape(x.p->s[1], t) && mb_same_shape(x.p->s[2], t);
}

// trunc_pr(amount, mul(a, b)) with the truncation keys given (the caller drew the multiplication's zero-share seeds and
// these keys in the reference's order): one pass, the zero share cancels in the opened sum (math_fused.cu).
bool same_shapes(const R &a, const R &b) {
  for (int i = 0; i < 3; i++)
    if (!mb_same_shape(a.p->s[i], a.like()) || !mb_same_shape(b.p->s[i], a.like())) return false;
  return a.like()->dtype == b.like()->dtype;
}
R mul_trunc_with(mb_ctx *c, const AesKey tk[3], int amount, const R &a, const R &b) {
  T3 out(c);
  out.alloc(a.like()->dtype, a.like()->rank, a.like()->shape);
  ck(k_mul_trunc(c, tk, (unsigned)amount, a.p, b.p, out.t));
  return out.to_rep();
}
// trunc(amount, mul(a, b)) as the reference issues it: mul's entropy, then trunc's
R mul_trunc(mb_ctx *c, int amount, const R &a, const R &b) {
  mb_rep *q = nullptr;
  ck(mb_rep_fixed_mul(c, (uint32_t)amount, a.p, b.p, &q));
  return R(c, q);
}

// top_most_index (division.rs:146-206) with the prefix-or as one launch per level over a plane buffer and the
// max_bits ring_injects + shifts + sum as one launch.  Keys are drawn in the reference's op order: the ANDs of prefix_or
// level by level (misc.rs:14-24), then the ring_injects for i = 0 .. max_bits - 1.
R top_most_index_batched(mb_ctx *c, int max_bits, const R &x) {
  mb_rep *xq = nullptr;
  ck(mb_rep_bit_decompose_low(c, x.p, max_bits, &xq));   // only planes 0 .. max_bits-1 are read below
  R x_bits(c, xq);
  const int64_t n = x.like()->numel;
  T3 W(c);
  const int64_t wshape[2] = {max_bits, n};
  W.alloc(MB_BIT, 2, wshape);
  short idx[128];
  for (int i = 0; i < max_bits; i++) idx[i] = (short)(max_bits - i - 1);   // x_rev
  ck(k_planes_gather(c, x_bits.p->s, idx, max_bits, W.t, n));
  int log_r = 0;
  while ((1 << log_r) < max_bits) log_r++;
  for (int i = 0; i < log_r; i++) {
    PrefixLevel lv;
    lv.nops = 0;
    for (int j = 0; j < (1 << log_r) / (1 << (i + 1)); j++) {
      const int y = (1 << i) + j * (1 << (i + 1)) - 1;
      for (int k = 1; k < (1 << i) + 1; k++)
        if (y + k < max_bits) {
          lv.src[lv.nops] = (short)y;
          lv.dst[lv.nops] = (short)(y + k);
          ck(mb_draw_zero_seeds(c, lv.keys[lv.nops]));
          lv.nops++;
        }
    }
    if (lv.nops) ck(k_prefix_or_level(c, lv, W.t, n));
  }
  // y = reverse(W); z[i] = y[i] - y[i+1] (i < max_bits - 1), z[max_bits-1] = y[max_bits-1];
  // res = (inject(z[max_bits-1]) << 0) + sum_{i >= 1} inject(z[i]) << (max_bits - i - 1)   (division.rs:196-203)
  InjectSumDesc d;
  d.nplanes = max_bits;
  d.negate = 0;
  for (int i = 0; i < max_bits; i++) {
    d.pa[i] = (short)(max_bits - 1 - i);
    d.pb[i] = (short)(i < max_bits - 1 ? max_bits - 2 - i : -1);
    d.shift[i] = (unsigned char)(max_bits - i - 1);
    d.mult[i] = (unsigned char)((i >= 1 ? 1 : 0) + (i == max_bits - 1 ? 1 : 0));
    ck(mb_draw_inject_keys(c, d.keys[i]));
  }
  T3 out(c);
  out.alloc(x.like()->dtype, x.like()->rank, x.like()->shape);
  ck(k_inject_sum(c, d, W.t, n, nullptr, out.t));
  return out.to_rep();
}

// division.rs:146-206
R top_most_index(mb_ctx *c, int max_bits, const R &x) {
  if (batched_ok(c, x) && max_bits >= 2 && max_bits <= MB_INJECT_MAX_PLANES) return top_most_index_batched(c, max_bits, x);
  R x_bits = bitdec(c, x);
  std::vector<R> y;
  for (int i = 0; i < max_bits; i++) y.push_back(index0(c, max_bits - i - 1, x_bits));
  prefix_op(y, [&](const R &a, const R &b) { return add(c, add(c, a, b), mul(c, a, b)); });  // or = xor(xor, and)
  std::reverse(y.begin(), y.end());
  std::vector<R> z;
  for (int i = 0; i + 1 < max_bits; i++) z.push_back(sub(c, y[i], y[i + 1]));
  z.push_back(y[max_bits - 1].clone());
  const mb_dtype ring = ring_of(x);
  std::vector<R> s_vec;
  for (int i = 0; i < max_bits; i++) {
    R zi = inject(c, ring, 0, z[i]);
    s_vec.push_back(shl(c, (uint64_t)(max_bits - i - 1), zi));
  }
  R res = shl(c, 0, s_vec[max_bits - 1]);
  for (int i = 1; i < max_bits; i++) res = add(c, res, s_vec[i]);
  return res;
}

// division.rs:110-139
void div_norm(mb_ctx *c, int max_bits, const R &x, R &upshifted, R &signed_topmost) {
  R msb = msb_ring(c, x);
  R sign = sign_from_msb(c, msb);
  R abs_x = mul(c, sign, x);
  R tm = top_most_index(c, max_bits, abs_x);
  upshifted = mul(c, x, tm);
  signed_topmost = mul(c, sign, tm);
}

// division.rs:208-249
R approximate_reciprocal(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  const int total = i_prec + f_prec;
  R upshifted(c, nullptr), signed_topmost(c, nullptr);
  div_norm(c, total, x, upshifted, signed_topmost);
  T alpha = pub_fill(c, x.like(), fixed_constant(bits_of(x), 2.9142, (unsigned)total));
  R d = pub_sub(c, alpha, shl(c, 1, upshifted));
  return mul_trunc(c, 2 * i_prec, d, signed_topmost);
}

// DivOp::rep_rep_kernel (division.rs:5-86)
R rep_div(mb_ctx *c, int i_prec, int f_prec, const R &x, const R &y) {
  const int bits = bits_of(x);
  const int k = i_prec + f_prec;
  if (2 * k > bits) {
    mb_fail(c, MB_ERR_INVALID_ARGUMENT, "2 * (fractional_precision + integral_precision) = " + std::to_string(2 * k) +
                                             ", BitLength = " + std::to_string(bits));
    throw Err{MB_ERR_INVALID_ARGUMENT};
  }
  const int theta = (int)std::ceil(std::log2((double)k / std::log2(17.0)));
  R w = approximate_reciprocal(c, i_prec, f_prec, y);
  T alpha = pub_fill(c, x.like(), fixed_constant(bits, 1.0, (unsigned)(2 * f_prec)));
  R init_prod = mul(c, y, w);  // broadcast to x's shape happens inside the share-wise ops below
  R a = pub_sub(c, alpha, init_prod);
  R b = mul_trunc(c, f_prec, x, w);
  for (int i = 0; i < theta; i++) {
    R a1 = add_pub(c, a, alpha);
    if (c->fused_bits && a.like()->numel > 0 && same_shapes(a, b) && same_shapes(a, a1) && 2 * f_prec + 1 <= bits - 1) {
      // reference order (division.rs:67-73): mul, mul, trunc, trunc -- draw in that order, then two fused passes
      AesKey z0[3], z1[3], t0[3], t1[3];
      ck(mb_draw_zero_seeds(c, z0));
      ck(mb_draw_zero_seeds(c, z1));
      ck(mb_draw_trunc_keys(c, t0));
      ck(mb_draw_trunc_keys(c, t1));
      R na = mul_trunc_with(c, t0, 2 * f_prec, a, a);
      R nb = mul_trunc_with(c, t1, 2 * f_prec, b, a1);
      a = std::move(na);
      b = std::move(nb);
      continue;
    }
    R next_a = mul(c, a, a);
    R next_b = mul(c, b, a1);
    a = trunc(c, (uint32_t)(2 * f_prec), next_a);
    b = trunc(c, (uint32_t)(2 * f_prec), next_b);
  }
  return mul_trunc(c, 2 * f_prec, b, add_pub(c, a, alpha));
}

const double P_1045[100] = {
#include "p1045.inc"
};

// PolynomialEval::polynomial_eval (fixedpoint/mod.rs:95-151); x has fractional precision f_prec
R polynomial_eval(mb_ctx *c, int f_prec, const double *coeff, int ncoeff, const R &x) {
  const int bits = bits_of(x);
  int degree = ncoeff - 1;
  for (int i = ncoeff - 1; i >= 0; i--) {
    if (coeff[i] < std::ldexp(1.0, -(f_prec + 1))) degree--;
    else break;
  }
  if (degree < 1) {
    mb_fail(c, MB_ERR_INVALID_ARGUMENT, "polynomial_eval: no coefficient above the precision");
    throw Err{MB_ERR_INVALID_ARGUMENT};
  }
  if (batched_ok(c, x) && degree + 1 <= MB_POLY_MAX_COEFFS) {
    // prefix_op over `degree` copies of x with op = trunc_pr(mul(a, b)): one launch per batch of <= 16 ops of a level,
    // then the weighted sum, its truncation and the constant term in one launch.  Keys per op in the reference's order
    // (multiplication, then truncation), level by level as prefix_op walks them (misc.rs:14-24).
    const int64_t nvec = x.like()->numel * (bits / 8) / 16;
    T3 P(c);
    int64_t pshape[MB_MAX_RANK + 1];
    pshape[0] = degree;
    for (int i = 0; i < x.like()->rank; i++) pshape[i + 1] = x.like()->shape[i];
    if (x.like()->rank + 1 > MB_MAX_RANK) { const int64_t flat[2] = {degree, x.like()->numel}; P.alloc(ring_of(x), 2, flat); }
    else P.alloc(ring_of(x), x.like()->rank + 1, pshape);
    ck(k_replicate(c, x.p, P.t, degree));
    int log_r = 0;
    while ((1 << log_r) < degree) log_r++;
    for (int i = 0; i < log_r; i++) {
      MulTruncLevel lv;
      lv.nops = 0;
      for (int j = 0; j < (1 << log_r) / (1 << (i + 1)); j++) {
        const int y = (1 << i) + j * (1 << (i + 1)) - 1;
        for (int k = 1; k < (1 << i) + 1; k++)
          if (y + k < degree) {
            lv.src[lv.nops] = (short)y;
            lv.dst[lv.nops] = (short)(y + k);
            ck(mb_draw_zero_seeds(c, lv.keys[lv.nops]));
            ck(mb_draw_trunc_keys(c, lv.keys[lv.nops] + 3));
            if (++lv.nops == MB_MULTRUNC_MAX_OPS) {
              ck(k_mul_trunc_level(c, lv, (unsigned)f_prec, P.t, nvec));
              lv.nops = 0;
            }
          }
      }
      if (lv.nops) ck(k_mul_trunc_level(c, lv, (unsigned)f_prec, P.t, nvec));
    }
    PolyCombine pc;
    pc.degree = degree;
    for (int i = 0; i <= degree; i++) {
      const u128 v = fixed_constant(bits, coeff[i], (unsigned)f_prec);
      pc.coeff_lo[i] = (uint64_t)v;
      pc.coeff_hi[i] = (uint64_t)(v >> 64);
    }
    ck(mb_draw_trunc_keys(c, pc.keys));
    T3 out(c);
    out.alloc(ring_of(x), x.like()->rank, x.like()->shape);
    ck(k_poly_combine(c, pc, (unsigned)f_prec, P.t, out.t));
    return out.to_rep();
  }
  std::vector<T> coeffs;
  for (int i = 0; i <= degree; i++) coeffs.push_back(pub_fill(c, x.like(), fixed_constant(bits, coeff[i], (unsigned)f_prec)));
  std::vector<R> pre;
  for (int i = 0; i < degree; i++) pre.push_back(x.clone());
  prefix_op(pre, [&](const R &a, const R &b) { return trunc(c, (uint32_t)f_prec, mul(c, a, b)); });
  R acc = mul_pub(c, pre[0], coeffs[1]);
  for (int i = 1; i < (int)pre.size(); i++) acc = add(c, acc, mul_pub(c, pre[i], coeffs[i + 1]));
  acc = trunc(c, (uint32_t)f_prec, acc);
  return add_pub(c, acc, coeffs[0]);
}
R polynomial_eval(mb_ctx *c, int f_prec, const R &x) { return polynomial_eval(c, f_prec, P_1045, 100, x); }

// Pow2FromBits (exp.rs:124-157)
R pow2_from_bits(mb_ctx *c, const std::vector<R> &xs, int count) {
  R ones = rep_fill(c, xs[0].like(), 1);
  std::vector<R> selectors;
  for (int i = 0; i < count; i++) {
    R pos = shl(c, (uint64_t)1 << i, xs[i]);
    R ng = sub(c, ones, xs[i]);
    selectors.push_back(add(c, pos, ng));
  }
  R acc = ones.clone();
  for (auto &y : selectors) acc = mul(c, acc, y);
  return acc;
}

// ExpFromParts (exp.rs:176-217)
R exp_from_parts(mb_ctx *c, const R &e_int, const R &e_frac, int f, int k) {
  const int amount = k - 2 - f;
  R x = shl(c, (uint64_t)amount, e_frac);
  R e_approx = polynomial_eval(c, k - 2, x);
  return mul_trunc(c, amount, e_int, e_approx);
}

// Pow2Op::rep_rep_kernel (exp.rs:11-111)
R rep_pow2(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  const int bits = bits_of(x);
  const mb_dtype ring = ring_of(x);
  // only the sign plane of this decomposition is read: the msb-pruned adder computes exactly that plane (the same
  // entropy requests, the same shares as plane bits-1 of the full decomposition)
  R msb = inject(c, ring, 0, batched_ok(c, x) ? msb_bit(c, x) : index0(c, bits - 1, bitdec(c, x)));
  R abs_x = mux(c, msb, neg(c, x), x);
  R absolute_bits = bitdec(c, abs_x);
  if (batched_ok(c, x) && i_prec >= 1 && i_prec <= MB_POW2_MAX_BITS && i_prec <= bits - f_prec) {
    // The higher bits (ring_inject of planes f .. bits-1), their weighted sum, the fractional part and Pow2FromBits
    // in 1 + ceil((bits - f - i) / 64) launches.  Keys in the reference's op order: every ring_inject first
    // (exp.rs:50-58), then the i_prec multiplications of pow2_from_bits (exp.rs:150-153).
    const int nhigh = bits - f_prec;
    const int64_t n = x.like()->numel;
    std::vector<AesKey> ik((size_t)nhigh * 2);
    for (int i = 0; i < nhigh; i++) ck(mb_draw_inject_keys(c, &ik[2 * (size_t)i]));
    Pow2Desc pd;
    pd.count = i_prec;
    pd.first_plane = f_prec;
    for (int i = 0; i < i_prec; i++) {
      pd.part_shift[i] = (unsigned char)((f_prec + i) & (bits - 1));
      pd.sel_shift[i] = (unsigned char)(((uint64_t)1 << i) & (uint64_t)(bits - 1));   // Rust masks the shift amount
      pd.inject[i][0] = ik[2 * (size_t)i];
      pd.inject[i][1] = ik[2 * (size_t)i + 1];
      ck(mb_draw_zero_seeds(c, pd.zero[i]));
    }
    T3 d3(c), rest(c);
    d3.alloc(ring, x.like()->rank, x.like()->shape);
    rest.alloc(ring, x.like()->rank, x.like()->shape);
    ck(k_pow2_bits(c, pd, absolute_bits.p->s, n, abs_x.p, d3.t, rest.t));
    R fractional_part = rest.to_rep();
    for (int lo = i_prec; lo < nhigh; lo += MB_INJECT_MAX_PLANES) {
      InjectSumDesc sd;
      sd.nplanes = std::min(MB_INJECT_MAX_PLANES, nhigh - lo);
      sd.negate = 1;
      for (int j = 0; j < sd.nplanes; j++) {
        sd.pa[j] = (short)(f_prec + lo + j);
        sd.pb[j] = -1;
        sd.shift[j] = (unsigned char)((f_prec + lo + j) & (bits - 1));
        sd.mult[j] = 1;
        sd.keys[j][0] = ik[2 * (size_t)(lo + j)];
        sd.keys[j][1] = ik[2 * (size_t)(lo + j) + 1];
      }
      T3 next(c);
      next.alloc(ring, x.like()->rank, x.like()->shape);
      ck(k_inject_sum(c, sd, absolute_bits.p->s, n, fractional_part.p, next.t));
      fractional_part = next.to_rep();
    }
    R d = d3.to_rep();
    R g = exp_from_parts(c, d, fractional_part, f_prec, f_prec + i_prec);
    R one_fixed = shl(c, (uint64_t)f_prec, rep_fill(c, x.like(), 1));
    R inverse = rep_div(c, i_prec, f_prec, one_fixed, g);
    return mux(c, msb, inverse, g);
  }
  std::vector<R> higher_bits;
  for (int i = f_prec; i < bits; i++) higher_bits.push_back(inject(c, ring, 0, index0(c, i, absolute_bits)));
  R acc = rep_fill(c, x.like(), 0);
  for (int i = 0; i < (int)higher_bits.size(); i++) acc = add(c, acc, shl(c, (uint64_t)(f_prec + i), higher_bits[i]));
  R fractional_part = sub(c, abs_x, acc);
  R d = pow2_from_bits(c, higher_bits, i_prec);
  R g = exp_from_parts(c, d, fractional_part, f_prec, f_prec + i_prec);
  R one_fixed = shl(c, (uint64_t)f_prec, rep_fill(c, x.like(), 1));
  R inverse = rep_div(c, i_prec, f_prec, one_fixed, g);
  return mux(c, msb, inverse, g);
}

// ExpOp::rep_rep_kernel (exp.rs:219-244)
R rep_exp(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  T log2e = pub_fill(c, x.like(), fixed_constant(bits_of(x), std::log2(std::exp(1.0)), (unsigned)f_prec));
  R exponent = trunc(c, (uint32_t)f_prec, mul_pub(c, x, log2e));
  return rep_pow2(c, i_prec, f_prec, exponent);
}

// SigmoidOp::rep_rep_kernel (arith.rs:676-743)
R rep_sigmoid(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  const int bits = bits_of(x);
  const mb_dtype ring = ring_of(x);
  R ones_rep = rep_fill(c, x.like(), fixed_constant(bits, 1.0, (unsigned)f_prec));
  R zeros_rep = rep_fill(c, x.like(), 0);
  R denominator = add(c, ones_rep, rep_exp(c, i_prec, f_prec, neg(c, x)));
  R output = rep_div(c, i_prec, f_prec, ones_rep, denominator);
  const double max_val = std::log(std::pow(2.0, (double)(i_prec - 1)));
  R max_val_rep = rep_fill(c, x.like(), fixed_constant(bits, max_val, (unsigned)f_prec));
  R upper_ring = inject(c, ring, 0, greater(c, x, max_val_rep));
  R upper_wall = mux(c, upper_ring, ones_rep, output);
  R lower_ring = inject(c, ring, 0, greater(c, neg(c, max_val_rep), x));
  return mux(c, lower_ring, zeros_rep, upper_wall);
}

// xor(ones, x) on shared bits: a public value is added to share 0 only (arith.rs:39-79)
R not_bits(mb_ctx *c, const R &x) {
  mb_tensor *n = nullptr;
  ck(mb_neg(c, x.p->s[0], &n));
  T tn(c, n);
  mb_rep *q = nullptr;
  ck(mb_rep_from_shares(c, n, x.p->s[1], x.p->s[2], &q));
  return R(c, q);
}

// EqualZeroOp::bitdec_bit_kernel (compare.rs:53-72): tree_reduce (misc.rs:54-74) of AND over the negated bits
R and_tree(mb_ctx *c, std::vector<R> &v, size_t lo, size_t hi) {
  if (hi - lo == 1) return v[lo].clone();
  const size_t mid = lo + (hi - lo) / 2;
  R a = and_tree(c, v, lo, mid);
  R b = and_tree(c, v, mid, hi);
  return mul(c, a, b);
}
R equal_zero(mb_ctx *c, const R &x_bits, int nbits) {
  std::vector<R> v_not;
  for (int i = 0; i < nbits; i++) v_not.push_back(not_bits(c, index0(c, i, x_bits)));
  return and_tree(c, v_not, 0, v_not.size());
}

// EqualOp::rep_kernel (compare.rs:24-40)
R rep_equal(mb_ctx *c, const R &x, const R &y) { return equal_zero(c, bitdec(c, sub(c, x, y)), bits_of(x)); }

R sub_pub(mb_ctx *c, const R &a, const T &pub) { mb_rep *q = nullptr; ck(mb_rep_sub_public(c, a.p, pub.p, &q)); return R(c, q); }

// Int2FL::int2fl (log.rs:166-219): v, p with (1 - 2s)(1 - z) v 2^p = x
void int2fl(mb_ctx *c, const R &x, int max_bit_len, int f_prec, R &v_out, R &p_out) {
  const int bits = bits_of(x);
  const mb_dtype ring = ring_of(x);
  const int lam = max_bit_len - 1;
  R x_bits = bitdec(c, x);
  R sign = inject(c, ring, 0, index0(c, bits - 1, x_bits));
  R is_zero = inject(c, ring, 0, equal_zero(c, x_bits, bits));
  R x_pos = mux(c, sign, neg(c, x), x);
  R x_pos_bits = bitdec(c, x_pos);
  std::vector<R> b;
  for (int i = lam - 1; i >= 0; i--) b.push_back(index0(c, i, x_pos_bits));
  prefix_op(b, [&](const R &p, const R &q) { return add(c, add(c, p, q), mul(c, p, q)); });  // prefix_or (misc.rs:30-40)
  std::vector<R> b_ring;
  for (int i = 0; i < lam; i++) b_ring.push_back(inject(c, ring, 0, b[i]));
  T ones = pub_fill(c, x.like(), 1);
  R neg_b_sum = rep_fill(c, x.like(), 0);
  for (int i = 0; i < lam; i++) neg_b_sum = add(c, neg_b_sum, shl(c, (uint64_t)i, pub_sub(c, ones, b_ring[i])));
  R x_upshifted = mul(c, x_pos, add_pub(c, neg_b_sum, ones));
  v_out = trunc(c, (uint32_t)(max_bit_len - 1 - f_prec), x_upshifted);
  R bit_count = rep_fill(c, x.like(), 0);
  for (int i = 0; i < lam; i++) bit_count = add(c, bit_count, b_ring[i]);
  T ften = pub_fill(c, x.like(), (u128)(uint8_t)f_prec);
  p_out = mul(c, sub_pub(c, bit_count, ften), pub_sub(c, ones, is_zero));
}

const double P_2524[4] = {-2.05466671951, -8.8626599391, 6.10585199015, 4.81147460989};  // log.rs:10
const double Q_2524[4] = {0.353553425277, 4.54517087629, 6.42784209029, 1.0};             // log.rs:11

// Log2Op::rep_rep_kernel (log.rs:13-82)
R rep_log2(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  R v(c, nullptr), p(c, nullptr);
  int2fl(c, x, i_prec + f_prec, f_prec, v, p);
  R p2524 = polynomial_eval(c, f_prec, P_2524, 4, v);
  R q2524 = polynomial_eval(c, f_prec, Q_2524, 4, v);
  R quotient = rep_div(c, i_prec, f_prec, p2524, q2524);
  return add(c, shl(c, (uint64_t)f_prec, p), quotient);
}

// LogOp::rep_rep_kernel (log.rs:84-111)
R rep_log(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  T ln2 = pub_fill(c, x.like(), fixed_constant(bits_of(x), std::log(2.0), (unsigned)f_prec));
  return trunc(c, (uint32_t)f_prec, mul_pub(c, rep_log2(c, i_prec, f_prec, x), ln2));
}

// SqrtOp::rep_rep_kernel (sqrt.rs:7-34): 2^(0.5 * log2 x)
R rep_sqrt(mb_ctx *c, int i_prec, int f_prec, const R &x) {
  R log2_x = rep_log2(c, i_prec, f_prec, x);
  T half = pub_fill(c, x.like(), fixed_constant(bits_of(x), 0.5, (unsigned)f_prec));
  R exponent = trunc(c, (uint32_t)f_prec, mul_pub(c, log2_x, half));
  return rep_pow2(c, i_prec, f_prec, exponent);
}

// share-wise helpers (rep_shape.cpp)
R index_axis(mb_ctx *c, int axis, int64_t i, const R &a) { mb_rep *q = nullptr; ck(mb_rep_index_axis(c, axis, i, a.p, &q)); return R(c, q); }
R expand_dim(mb_ctx *c, int axis, const R &a) { const int64_t ax = axis; mb_rep *q = nullptr; ck(mb_rep_expand_dims(c, &ax, 1, a.p, &q)); return R(c, q); }
R sum_axis(mb_ctx *c, int axis, const R &a) { mb_rep *q = nullptr; ck(mb_rep_sum(c, a.p, axis, &q)); return R(c, q); }
R less_bit(mb_ctx *c, const R &x, const R &y) { return msb_bit(c, sub(c, x, y)); }  // compare.rs:91-104

// AbsOp::rep_ring_kernel (arith.rs:655-674)
R rep_abs(mb_ctx *c, const R &x) {
  R msb = msb_ring(c, x);
  return mul(c, sign_from_msb(c, msb), x);
}

// SoftmaxOp::rep_fixed_kernel (softmax.rs:55-127)
R rep_softmax(mb_ctx *c, int axis, int upmost_index, int i_prec, int f_prec, const R &x) {
  const int bits = bits_of(x);
  std::vector<mb_rep *> raw;
  std::vector<R> xs;
  for (int i = 0; i < upmost_index; i++) xs.push_back(index_axis(c, axis, i, x));
  for (auto &t : xs) raw.push_back(t.p);
  mb_rep *mq = nullptr;
  ck(mb_rep_maximum(c, raw.data(), (int)raw.size(), &mq));
  R xmax = expand_dim(c, axis, R(c, mq));
  R max_diff = sub(c, x, xmax);
  R e_x = rep_exp(c, i_prec, f_prec, max_diff);
  R zeros_rep = rep_fill(c, e_x.like(), 0);
  const double min_val = -1.0 * std::log(std::pow(2.0, (double)(i_prec - 1)));
  R lower_bound = rep_fill(c, max_diff.like(), fixed_constant(bits, min_val, (unsigned)f_prec));
  R threshold_ring = inject(c, ring_of(x), 0, greater(c, lower_bound, max_diff));
  R normalized = mux(c, threshold_ring, zeros_rep, e_x);
  R e_x_sum = expand_dim(c, axis, sum_axis(c, axis, normalized));
  return rep_div(c, i_prec, f_prec, normalized, e_x_sum);
}

// RingFixedpointArgmaxOp::rep_ring_kernel with TreeReduceArgmax (argmax.rs:10-92); the result is Ring64
struct IdxVal { R idx, val; };
IdxVal argmax_tree(mb_ctx *c, std::vector<IdxVal> &v, size_t lo, size_t hi) {
  if (hi - lo == 1) return IdxVal{v[lo].idx.clone(), v[lo].val.clone()};
  const size_t mid = lo + (hi - lo) / 2;
  IdxVal a = argmax_tree(c, v, lo, mid);
  IdxVal b = argmax_tree(c, v, mid, hi);
  R comp_ring = inject(c, ring_of(a.val), 0, less_bit(c, a.val, b.val));
  R idx = mux(c, comp_ring, b.idx, a.idx);
  R val = mux(c, comp_ring, b.val, a.val);
  return IdxVal{std::move(idx), std::move(val)};
}
R rep_argmax(mb_ctx *c, int axis, int upmost_index, const R &x) {
  std::vector<IdxVal> pairs;
  for (int i = 0; i < upmost_index; i++) {
    R item = index_axis(c, axis, i, x);
    R idx = rep_fill(c, item.like(), (u128)(uint8_t)i);  // (i as u8).into()
    pairs.push_back(IdxVal{std::move(idx), std::move(item)});
  }
  IdxVal top = argmax_tree(c, pairs, 0, pairs.size());
  if (ring_of(x) == MB_RING64) return std::move(top.idx);
  mb_rep *q = nullptr;
  ck(mb_rep_cast_ring128_to_ring64(c, top.idx.p, &q));
  return R(c, q);
}

bool valid(const mb_rep *x) { return x && x->s[0] && x->s[1] && x->s[2]; }

template <typename F>
int guarded(mb_ctx *ctx, const char *name, F f) {
  if (!ctx) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL context");
  try {
    return f();
  } catch (const Err &e) {
    return e.code;  // message already recorded by the failing call
  } catch (const std::exception &e) {
    return mb_fail(ctx, MB_ERR_KERNEL, std::string(name) + ": " + e.what());
  }
}

int check_fixed(mb_ctx *ctx, const mb_rep *x, uint32_t i_prec, uint32_t f_prec, const char *name) {
  if (!valid(x)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL argument");
  if (!mb_is_ring(x->s[0]->dtype)) return mb_fail(ctx, MB_ERR_TYPE_MISMATCH, std::string(name) + ": needs ring shares");
  const int bits = x->s[0]->dtype == MB_RING128 ? 128 : 64;
  if (i_prec == 0 || f_prec == 0 || (int)(i_prec + f_prec) >= bits)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": bad fixed-point precision");
  return MB_OK;
}

}  // namespace

extern "C" int mb_rep_fill(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, uint64_t lo, uint64_t hi, mb_rep **out) {
  return guarded(ctx, "mb_rep_fill", [&]() {
    if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_fill: NULL out");
    mb_tensor *like = nullptr;
    ck(mb_fill(ctx, dtype, rank, shape, lo, hi, &like));
    T tl(ctx, like);
    mb_tensor *z1 = nullptr, *z2 = nullptr;
    ck(mb_fill(ctx, dtype, rank, shape, 0, 0, &z1));
    T t1(ctx, z1);
    ck(mb_fill(ctx, dtype, rank, shape, 0, 0, &z2));
    T t2(ctx, z2);
    return mb_rep_from_shares(ctx, like, z1, z2, out);
  });
}

extern "C" int mb_rep_greater(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  return guarded(ctx, "mb_rep_greater", [&]() { return mb_rep_less(ctx, y, x, out); });  // msb(y - x)
}

#define BORROW(x) R(ctx, nullptr)
static R borrow(mb_ctx *ctx, const mb_rep *x) {
  mb_rep *q = nullptr;
  ck(mb_rep_from_shares(ctx, x->s[0], x->s[1], x->s[2], &q));
  return R(ctx, q);
}

extern "C" int mb_rep_div(mb_ctx *ctx, uint32_t i_prec, uint32_t f_prec, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  return guarded(ctx, "mb_rep_div", [&]() {
    MB_CHECK(check_fixed(ctx, x, i_prec, f_prec, "mb_rep_div"));
    MB_CHECK(check_fixed(ctx, y, i_prec, f_prec, "mb_rep_div"));
    if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_div: NULL out");
    R r = rep_div(ctx, (int)i_prec, (int)f_prec, borrow(ctx, x), borrow(ctx, y));
    *out = r.release();
    return (int)MB_OK;
  });
}
extern "C" int mb_rep_pow2(mb_ctx *ctx, uint32_t i_prec, uint32_t f_prec, const mb_rep *x, mb_rep **out) {
  return guarded(ctx, "mb_rep_pow2", [&]() {
    MB_CHECK(check_fixed(ctx, x, i_prec, f_prec, "mb_rep_pow2"));
    if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_pow2: NULL out");
    R r = rep_pow2(ctx, (int)i_prec, (int)f_prec, borrow(ctx, x));
    *out = r.release();
    return (int)MB_OK;
  });
}
extern "C" int mb_rep_exp(mb_ctx *ctx, uint32_t i_prec, uint32_t f_prec, const mb_rep *x, mb_rep **out) {
  return guarded(ctx, "mb_rep_exp", [&]() {
    MB_CHECK(check_fixed(ctx, x, i_prec, f_prec, "mb_rep_exp"));
    if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_exp: NULL out");
    R r = rep_exp(ctx, (int)i_prec, (int)f_prec, borrow(ctx, x));
    *out = r.release();
    return (int)MB_OK;
  });
}
extern "C" int mb_rep_sigmoid(mb_ctx *ctx, uint32_t i_prec, uint32_t f_prec, const mb_rep *x, mb_rep **out) {
  return guarded(ctx, "mb_rep_sigmoid", [&]() {
    MB_CHECK(check_fixed(ctx, x, i_prec, f_prec, "mb_rep_sigmoid"));
    if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_sigmoid: NULL out");
    R r = rep_sigmoid(ctx, (int)i_prec, (int)f_prec, borrow(ctx, x));
    *out = r.release();
    return (int)MB_OK;
  });
}

extern "C" int mb_rep_abs(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  return guarded(ctx, "mb_rep_abs", [&]() {
    if (!valid(x) || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_abs: NULL argument");
    if (!mb_is_ring(x->s[0]->dtype)) return mb_fail(ctx, MB_ERR_TYPE_MISMATCH, "mb_rep_abs: needs ring shares");
    R r = rep_abs(ctx, borrow(ctx, x));
    *out = r.release();
    return (int)MB_OK;
  });
}
extern "C" int mb_rep_softmax(mb_ctx *ctx, int axis, int upmost_index, uint32_t i_prec, uint32_t f_prec, const mb_rep *x, mb_rep **out) {
  return guarded(ctx, "mb_rep_softmax", [&]() {
    MB_CHECK(check_fixed(ctx, x, i_prec, f_prec, "mb_rep_softmax"));
    if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_softmax: NULL out");
    if (axis < 0 || axis >= x->s[0]->rank || upmost_index < 1 || upmost_index > x->s[0]->shape[axis])
      return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_softmax: axis / upmost_index out of range");
    R r = rep_softmax(ctx, axis, upmost_index, (int)i_prec, (int)f_prec, borrow(ctx, x));
    *out = r.release();
    return (int)MB_OK;
  });
}
extern "C" int mb_rep_argmax(mb_ctx *ctx, int axis, int upmost_index, const mb_rep *x, mb_rep **out) {
  return guarded(ctx, "mb_rep_argmax", [&]() {
    if (!valid(x) || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_argmax: NULL argument");
    if (!mb_is_ring(x->s[0]->dtype)) return mb_fail(ctx, MB_ERR_TYPE_MISMATCH, "mb_rep_argmax: needs ring shares");
    if (axis < 0 || axis >= x->s[0]->rank || upmost_index < 1 || upmost_index > x->s[0]->shape[axis] || upmost_index > 256)
      return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_argmax: axis / upmost_index out of range");
    R r = rep_argmax(ctx, axis, upmost_index, borrow(ctx, x));
    *out = r.release();
    return (int)MB_OK;
  });
}

extern "C" int mb_rep_equal(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out_bits) {
  return guarded(ctx, "mb_rep_equal", [&]() {
    if (!valid(x) || !valid(y) || !out_bits) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_rep_equal: NULL argument");
    if (!mb_is_ring(x->s[0]->dtype) || x->s[0]->dtype != y->s[0]->dtype) return mb_fail(ctx, MB_ERR_TYPE_MISMATCH, "mb_rep_equal: needs ring shares of one width");
    R r = rep_equal(ctx, borrow(ctx, x), borrow(ctx, y));
    *out_bits = r.release();
    return (int)MB_OK;
  });
}
#define MB_FIXED_UNARY(NAME, FN)                                                                                             \
  extern "C" int NAME(mb_ctx *ctx, uint32_t i_prec, uint32_t f_prec, const mb_rep *x, mb_rep **out) {                      \
    return guarded(ctx, #NAME, [&]() {                                                                                      \
      MB_CHECK(check_fixed(ctx, x, i_prec, f_prec, #NAME));                                                                 \
      if (!out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, #NAME ": NULL out");                                           \
      R r = FN(ctx, (int)i_prec, (int)f_prec, borrow(ctx, x));                                                              \
      *out = r.release();                                                                                                   \
      return (int)MB_OK;                                                                                                    \
    });                                                                                                                     \
  }
MB_FIXED_UNARY(mb_rep_log2, rep_log2)
MB_FIXED_UNARY(mb_rep_log, rep_log)
MB_FIXED_UNARY(mb_rep_sqrt, rep_sqrt)
