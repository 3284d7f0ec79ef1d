// C ABI: replicated-placement protocols (moose/src/replicated/*.rs,
// moose/src/additive/*.rs).  The host side only derives keys / seeds (16-byte
// values, in the reference's entropy order) and launches fused kernels.
#include "common.h"

int mb_dot_dims(const mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, int64_t *M, int64_t *K, int64_t *N, int *orank,
                int64_t *oshape);

#define MB_ENTER(ctx, what)                                                              \
  do {                                                                                   \
    if (!(ctx)) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, what ": NULL context"); \
    MB_CUDA((ctx), cudaSetDevice((ctx)->device));                                        \
  } while (0)
#define MB_REQUIRE(ctx, cond, code, msg)               \
  do {                                                 \
    if (!(cond)) return mb_fail((ctx), (code), (msg)); \
  } while (0)

namespace {

// Owning tensor handle for temporaries.
struct Tmp {
  mb_ctx *ctx;
  mb_tensor *p;
  explicit Tmp(mb_ctx *c) : ctx(c), p(nullptr) {}
  ~Tmp() { if (p) mb_tensor_free(ctx, p); }
  Tmp(const Tmp &) = delete;
  Tmp &operator=(const Tmp &) = delete;
  mb_tensor *release() { mb_tensor *t = p; p = nullptr; return t; }
  int alloc_like(const mb_tensor *x) { return mb_tensor_new(ctx, x->dtype, x->rank, x->shape, &p); }
  int alloc(mb_dtype d, int rank, const int64_t *shape) { return mb_tensor_new(ctx, d, rank, shape, &p); }
};

struct RepTmp {
  mb_ctx *ctx;
  mb_rep *p;
  explicit RepTmp(mb_ctx *c) : ctx(c), p(nullptr) {}
  ~RepTmp() { if (p) mb_rep_free(ctx, p); }
  RepTmp(const RepTmp &) = delete;
  RepTmp &operator=(const RepTmp &) = delete;
  mb_rep *release() { mb_rep *t = p; p = nullptr; return t; }
};

mb_rep *rep_make(mb_tensor *a, mb_tensor *b, mb_tensor *c) {
  mb_rep *r = new mb_rep();
  r->s[0] = a; r->s[1] = b; r->s[2] = c;
  r->deferred = nullptr;
  return r;
}

int rep_alloc_like(mb_ctx *ctx, const mb_tensor *like, mb_rep **out) {
  Tmp a(ctx), b(ctx), c(ctx);
  MB_CHECK(a.alloc_like(like));
  MB_CHECK(b.alloc_like(like));
  MB_CHECK(c.alloc_like(like));
  *out = rep_make(a.release(), b.release(), c.release());
  return MB_OK;
}

int rep_alloc(mb_ctx *ctx, mb_dtype d, int rank, const int64_t *shape, mb_rep **out) {
  Tmp a(ctx), b(ctx), c(ctx);
  MB_CHECK(a.alloc(d, rank, shape));
  MB_CHECK(b.alloc(d, rank, shape));
  MB_CHECK(c.alloc(d, rank, shape));
  *out = rep_make(a.release(), b.release(), c.release());
  return MB_OK;
}

bool rep_valid(const mb_rep *x) { return x && x->s[0] && x->s[1] && x->s[2]; }
int alloc3(mb_ctx *ctx, const mb_tensor *like, mb_tensor *z[3]);

int ensure_setup(mb_ctx *ctx) {
  // RepSetup: k00, k11, k22 generated in this order (replicated/setup.rs:46-48);
  // k10 = k11, k21 = k22, k02 = k00 are relabelled copies (:50-52).
  // one thread generates the keys; concurrent callers (the reference runs every op as its own task,
  // execution/asynchronous.rs:453-531) wait for it and then all use the same setup
  std::lock_guard<std::mutex> setup_lock(ctx->setup_mu);
  if (!ctx->have_setup) {
    uint8_t k[3][16];
    for (int i = 0; i < 3; i++) MB_CHECK(mb_entropy16(ctx, k[i]));
    std::lock_guard<std::mutex> lock(ctx->mu);
    memcpy(ctx->setup_keys, k, sizeof(k));
    ctx->have_setup = true;
  }
  return MB_OK;
}

void seed_to_aes(mb_ctx *ctx, const uint8_t sync_key[16], const uint8_t key[16], AesKey *out) {
  uint8_t seed[16];
  mb_blake3_derive_seed(ctx->sid, sync_key, key, seed);
  mb_aes_expand(seed, out);
}

// SeedsGen::gen_seeds (replicated/zero_share.rs:129-156): s_i = DeriveSeed(k_i, sync_i);
// s10 == s11, s21 == s22, s02 == s00 because the keys are copies.
int gen_zero_seeds(mb_ctx *ctx, AesKey seeds[3]) {
  MB_CHECK(ensure_setup(ctx));
  uint8_t sync[3][16];
  for (int i = 0; i < 3; i++) MB_CHECK(mb_entropy16(ctx, sync[i]));
  for (int i = 0; i < 3; i++) seed_to_aes(ctx, sync[i], ctx->setup_keys[i], &seeds[i]);
  return MB_OK;
}

}  // namespace

int mb_draw_zero_seeds(mb_ctx *ctx, AesKey seeds[3]) { return gen_zero_seeds(ctx, seeds); }

// RingInjectOp::rep_kernel: gen_dabit (dabit.rs:49-55): Sample seed, PrfKeyGen, 2 x SyncKey; adt_to_rep
// (convert.rs:426-431): 2 x SyncKey, PrfKeyGen.  Only the adt_to_rep streams survive in the result.
int mb_draw_inject_keys(mb_ctx *ctx, AesKey keys[2]) {
  uint8_t seed_bits[16], key[16], sync_b[16], sync_br[16], sync0[16], sync1[16], key2[16];
  MB_CHECK(mb_entropy16(ctx, seed_bits));
  MB_CHECK(mb_entropy16(ctx, key));
  MB_CHECK(mb_entropy16(ctx, sync_b));
  MB_CHECK(mb_entropy16(ctx, sync_br));
  MB_CHECK(mb_entropy16(ctx, sync0));
  MB_CHECK(mb_entropy16(ctx, sync1));
  MB_CHECK(mb_entropy16(ctx, key2));
  seed_to_aes(ctx, sync0, key2, &keys[0]);
  seed_to_aes(ctx, sync1, key2, &keys[1]);
  return MB_OK;
}

// TruncPrOp::rep_kernel: gen_trunc_mask (trunc.rs:45-63): Sample seed, PrfKeyGen, 3 x SyncKey::random();
// adt_to_rep (convert.rs:426-431): 2 x SyncKey::random(), PrfKeyGen.  (r, y0, y1) survive in the result.
int mb_draw_trunc_keys(mb_ctx *ctx, AesKey keys[3]) {
  uint8_t seed_r[16], key[16], sync[3][16], sync0[16], sync1[16], key2[16];
  MB_CHECK(mb_entropy16(ctx, seed_r));
  MB_CHECK(mb_entropy16(ctx, key));
  for (int i = 0; i < 3; i++) MB_CHECK(mb_entropy16(ctx, sync[i]));
  MB_CHECK(mb_entropy16(ctx, sync0));
  MB_CHECK(mb_entropy16(ctx, sync1));
  MB_CHECK(mb_entropy16(ctx, key2));
  mb_aes_expand(seed_r, &keys[0]);
  seed_to_aes(ctx, sync0, key2, &keys[1]);
  seed_to_aes(ctx, sync1, key2, &keys[2]);
  return MB_OK;
}

namespace {

// three fresh tensors shaped like `like`
int alloc3(mb_ctx *ctx, const mb_tensor *like, mb_tensor *z[3]) {
  z[0] = z[1] = z[2] = nullptr;
  for (int i = 0; i < 3; i++) {
    int rc = mb_tensor_new(ctx, like->dtype, like->rank, like->shape, &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  return MB_OK;
}

bool uniform_shapes(const mb_rep *x, const mb_rep *y) {
  static const bool no_batch = getenv("MOOSE_B200_NO_SHARE_BATCH") != nullptr;
  if (no_batch) return false;
  for (int i = 0; i < 3; i++)
    if (!mb_same_shape(x->s[i], x->s[0]) || (y && !mb_same_shape(y->s[i], x->s[0]))) return false;
  return true;
}

int sharewise_binary(mb_ctx *ctx, EwOp op, const mb_rep *x, const mb_rep *y, mb_rep **out, const char *name) {
  MB_REQUIRE(ctx, rep_valid(x) && rep_valid(y) && out, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL argument");
  MB_REQUIRE(ctx, x->s[0]->dtype == y->s[0]->dtype, MB_ERR_TYPE_MISMATCH, std::string(name) + ": dtypes differ");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  if (uniform_shapes(x, y) && x->s[0]->dtype != MB_FLOAT64) {
    // all six operands of one shape: the three share-wise ops go out as ONE launch
    EwOp o = op;
    if (x->s[0]->dtype == MB_BIT) o = op == EW_MUL ? EW_AND : EW_XOR;  // Z2: add = sub = xor, mul = and
    MB_CHECK(alloc3(ctx, x->s[0], z));
    int rc = k_ew_binary3(ctx, o, x->s, y->s, z);
    if (rc != MB_OK) {
      for (int j = 0; j < 3; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
    *out = rep_make(z[0], z[1], z[2]);
    return MB_OK;
  }
  for (int i = 0; i < 3; i++) {
    int rc = op == EW_ADD ? mb_add(ctx, x->s[i], y->s[i], &z[i]) : op == EW_SUB ? mb_sub(ctx, x->s[i], y->s[i], &z[i])
                                                                                 : mb_mul(ctx, x->s[i], y->s[i], &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

int sharewise_unary(mb_ctx *ctx, int mode, uint64_t amount, const mb_rep *x, mb_rep **out) {
  mb_tensor *z[3];
  MB_CHECK(alloc3(ctx, x->s[0], z));
  int rc = k_ew_unary3(ctx, mode, (unsigned)(amount & 127), x->s, z);
  if (rc != MB_OK) {
    for (int j = 0; j < 3; j++) mb_tensor_free(ctx, z[j]);
    return rc;
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

}  // namespace

// ---- handles ------------------------------------------------------------------------

extern "C" int mb_rep_from_shares(mb_ctx *ctx, const mb_tensor *x0, const mb_tensor *x1, const mb_tensor *x2, mb_rep **out) {
  MB_REQUIRE(ctx, x0 && x1 && x2 && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_from_shares: NULL argument");
  MB_REQUIRE(ctx, x0->dtype == x1->dtype && x1->dtype == x2->dtype && mb_same_shape(x0, x1) && mb_same_shape(x1, x2),
             MB_ERR_TYPE_MISMATCH, "mb_rep_from_shares: shares differ in dtype or shape");
  *out = rep_make(mb_tensor_retain((mb_tensor *)x0), mb_tensor_retain((mb_tensor *)x1), mb_tensor_retain((mb_tensor *)x2));
  return MB_OK;
}
extern "C" const mb_tensor *mb_rep_share_tensor(const mb_rep *x, int i) { return (x && i >= 0 && i < 3) ? x->s[i] : nullptr; }
extern "C" int mb_rep_free(mb_ctx *ctx, mb_rep *x) {
  if (!x) return MB_OK;
  for (int i = 0; i < 3; i++) mb_tensor_free(ctx, x->s[i]);
  if (x->deferred) {
    mb_tensor_free(ctx, x->deferred->src);
    delete x->deferred;
  }
  delete x;
  return MB_OK;
}

extern "C" int mb_rep_setup(mb_ctx *ctx) {
  MB_ENTER(ctx, "mb_rep_setup");
  return ensure_setup(ctx);
}

// ---- share / reveal -------------------------------------------------------------------

extern "C" int mb_rep_share(mb_ctx *ctx, const mb_tensor *x, int owner, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_share");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_share: NULL argument");
  MB_REQUIRE(ctx, x->dtype != MB_FLOAT64, MB_ERR_TYPE_MISMATCH, "mb_rep_share: float tensors must be fixed-point encoded first");
  MB_REQUIRE(ctx, owner >= 0 && owner <= 3, MB_ERR_INVALID_ARGUMENT, "mb_rep_share: owner must be 0..3");
  MB_CHECK(ensure_setup(ctx));  // convert.rs:67: sess.setup(plc) precedes SyncKey::random()
  RepTmp r(ctx);
  MB_CHECK(rep_alloc_like(ctx, x, &r.p));
  if (owner < 3) {
    uint8_t sync[16];
    MB_CHECK(mb_entropy16(ctx, sync));
    AesKey k;
    seed_to_aes(ctx, sync, ctx->setup_keys[owner], &k);
    MB_CHECK(k_rep_share(ctx, k, owner, x, r.p->s[0], r.p->s[1], r.p->s[2]));
  } else {
    uint8_t sync0[16], sync1[16];
    MB_CHECK(mb_entropy16(ctx, sync0));
    MB_CHECK(mb_entropy16(ctx, sync1));
    AesKey k0, k1;
    seed_to_aes(ctx, sync0, ctx->setup_keys[0], &k0);  // seed00 (= seed02)
    seed_to_aes(ctx, sync1, ctx->setup_keys[1], &k1);  // seed11 (= seed10)
    MB_CHECK(k_rep_share_outsider(ctx, k0, k1, x, r.p->s[0], r.p->s[1], r.p->s[2]));
  }
  *out = r.release();
  return MB_OK;
}

// ShareOp by a placement member with the PRF expansion postponed: the sync key is drawn NOW (the reference's entropy
// order), the shares x_owner = PRF(seed), x_{owner+1} = x - x_owner, x_{owner+2} = 0 are produced by the consumer --
// mb_rep_dot_concat / mb_rep_fixed_dot_concat expand them inside the skinny dot kernel -- or by mb_rep_materialize.
// Any other entry point rejects a deferred handle as a NULL argument (it has no share tensors yet).
extern "C" int mb_rep_share_deferred(mb_ctx *ctx, const mb_tensor *x, int owner, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_share_deferred");
  MB_REQUIRE(ctx, x && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_share_deferred: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->dtype), MB_ERR_TYPE_MISMATCH, "mb_rep_share_deferred: needs a ring tensor");
  MB_REQUIRE(ctx, owner >= 0 && owner <= 2, MB_ERR_INVALID_ARGUMENT, "mb_rep_share_deferred: owner must be a placement member (0..2)");
  MB_CHECK(ensure_setup(ctx));
  uint8_t sync[16];
  MB_CHECK(mb_entropy16(ctx, sync));
  mb_rep *r = rep_make(nullptr, nullptr, nullptr);
  r->deferred = new mb_rep_deferred();
  r->deferred->src = mb_tensor_retain(const_cast<mb_tensor *>(x));
  r->deferred->owner = owner;
  seed_to_aes(ctx, sync, ctx->setup_keys[owner], &r->deferred->key);
  *out = r;
  return MB_OK;
}
extern "C" int mb_rep_materialize(mb_ctx *ctx, mb_rep *x) {
  MB_ENTER(ctx, "mb_rep_materialize");
  MB_REQUIRE(ctx, x, MB_ERR_INVALID_ARGUMENT, "mb_rep_materialize: NULL argument");
  if (!x->deferred) return MB_OK;
  mb_rep_deferred *d = x->deferred;
  mb_tensor *z[3];
  MB_CHECK(alloc3(ctx, d->src, z));
  int rc = k_rep_share(ctx, d->key, d->owner, d->src, z[0], z[1], z[2]);
  if (rc != MB_OK) {
    for (int i = 0; i < 3; i++) mb_tensor_free(ctx, z[i]);
    return rc;
  }
  for (int i = 0; i < 3; i++) x->s[i] = z[i];
  mb_tensor_free(ctx, d->src);
  delete d;
  x->deferred = nullptr;
  return MB_OK;
}

extern "C" int mb_rep_reveal(mb_ctx *ctx, const mb_rep *x, int receiver, mb_tensor **out) {
  MB_ENTER(ctx, "mb_rep_reveal");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_reveal: NULL argument");
  (void)receiver;  // every receiver adds the same three distinct shares (convert.rs:295-311)
  Tmp t(ctx);
  MB_CHECK(t.alloc_like(x->s[0]));
  MB_CHECK(k_reveal3(ctx, x->s[0], x->s[1], x->s[2], t.p));
  *out = t.release();
  return MB_OK;
}

// ---- linear ops -----------------------------------------------------------------------

extern "C" int mb_rep_add(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_add");
  return sharewise_binary(ctx, EW_ADD, x, y, out, "mb_rep_add");
}
extern "C" int mb_rep_sub(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_sub");
  return sharewise_binary(ctx, EW_SUB, x, y, out, "mb_rep_sub");
}

extern "C" int mb_rep_neg(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_neg");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_neg: NULL argument");
  if (mb_is_ring(x->s[0]->dtype) && uniform_shapes(x, nullptr)) return sharewise_unary(ctx, 0, 0, x, out);
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  for (int i = 0; i < 3; i++) {
    int rc = MB_OK;
    // NegOp::rep_bit_kernel flips only x00 / x02 = share 0 (arith.rs:525-558); the other shares are shared as is
    if (x->s[0]->dtype == MB_BIT && i != 0) z[i] = mb_tensor_retain(x->s[i]);
    else rc = mb_neg(ctx, x->s[i], &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

extern "C" int mb_rep_shl(mb_ctx *ctx, uint64_t amount, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_shl");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_shl: NULL argument");
  if (mb_is_ring(x->s[0]->dtype) && uniform_shapes(x, nullptr)) return sharewise_unary(ctx, 1, amount, x, out);
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  for (int i = 0; i < 3; i++) {
    int rc = mb_shl(ctx, amount, x->s[i], &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

// Mirrored operand: only share 0 (= x00 and x02) sees the public value, the other shares are
// broadcast to the result shape (arith.rs:81-121, 275-315).
static int rep_public_linear(mb_ctx *ctx, bool subtract, const mb_rep *x, const mb_tensor *pub, mb_rep **out) {
  MB_REQUIRE(ctx, rep_valid(x) && pub && out, MB_ERR_INVALID_ARGUMENT, "rep op public: NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  int rc = subtract ? mb_sub(ctx, x->s[0], pub, &z[0]) : mb_add(ctx, x->s[0], pub, &z[0]);
  if (rc != MB_OK) return rc;
  for (int i = 1; i < 3; i++) {
    // tensors are immutable, so a share whose shape already is the result shape is shared, not copied
    if (mb_same_shape(x->s[i], z[0])) {
      z[i] = mb_tensor_retain(x->s[i]);
      continue;
    }
    rc = mb_broadcast(ctx, x->s[i], z[0]->rank, z[0]->shape, &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}
extern "C" int mb_rep_add_public(mb_ctx *ctx, const mb_rep *x, const mb_tensor *pub, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_add_public");
  return rep_public_linear(ctx, false, x, pub, out);
}
extern "C" int mb_rep_sub_public(mb_ctx *ctx, const mb_rep *x, const mb_tensor *pub, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_sub_public");
  return rep_public_linear(ctx, true, x, pub, out);
}
extern "C" int mb_rep_public_sub(mb_ctx *ctx, const mb_tensor *pub, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_public_sub");
  // SubOp::mir_rep_kernel (arith.rs:230-273): z0 = pub - x0, z1 = -x1, z2 = -x2 (broadcast)
  MB_REQUIRE(ctx, rep_valid(x) && pub && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_public_sub: NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  int rc = mb_sub(ctx, pub, x->s[0], &z[0]);
  if (rc != MB_OK) return rc;
  for (int i = 1; i < 3; i++) {
    Tmp n(ctx);
    rc = mb_neg(ctx, x->s[i], &n.p);
    if (rc == MB_OK) {
      if (mb_same_shape(n.p, z[0])) z[i] = n.release();
      else rc = mb_broadcast(ctx, n.p, z[0]->rank, z[0]->shape, &z[i]);
    }
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}
extern "C" int mb_rep_mul_public(mb_ctx *ctx, const mb_rep *x, const mb_tensor *pub, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_mul_public");
  MB_REQUIRE(ctx, rep_valid(x) && pub && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_mul_public: NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  for (int i = 0; i < 3; i++) {
    int rc = mb_mul(ctx, x->s[i], pub, &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

// ---- multiplication -------------------------------------------------------------------

extern "C" int mb_rep_mul(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_mul");
  MB_REQUIRE(ctx, rep_valid(x) && rep_valid(y) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_mul: NULL argument");
  MB_REQUIRE(ctx, x->s[0]->dtype == y->s[0]->dtype, MB_ERR_TYPE_MISMATCH, "mb_rep_mul: dtypes differ");
  const mb_rep *xx = x, *yy = y;
  RepTmp xb(ctx), yb(ctx);
  if (!mb_same_shape(x->s[0], y->s[0])) {
    // ndarray co-broadcast of the share products: materialise broadcast operands
    MB_REQUIRE(ctx, mb_is_ring(x->s[0]->dtype), MB_ERR_KERNEL, "mb_rep_mul: bit operands must have equal shapes");
    Tmp probe(ctx);
    MB_CHECK(mb_add(ctx, x->s[0], y->s[0], &probe.p));  // result shape (and validity) of the broadcast
    mb_tensor *a[3] = {nullptr, nullptr, nullptr}, *b[3] = {nullptr, nullptr, nullptr};
    int rc = MB_OK;
    for (int i = 0; i < 3 && rc == MB_OK; i++) rc = mb_broadcast(ctx, x->s[i], probe.p->rank, probe.p->shape, &a[i]);
    for (int i = 0; i < 3 && rc == MB_OK; i++) rc = mb_broadcast(ctx, y->s[i], probe.p->rank, probe.p->shape, &b[i]);
    if (rc != MB_OK) {
      for (int i = 0; i < 3; i++) { mb_tensor_free(ctx, a[i]); mb_tensor_free(ctx, b[i]); }
      return rc;
    }
    xb.p = rep_make(a[0], a[1], a[2]);
    yb.p = rep_make(b[0], b[1], b[2]);
    xx = xb.p; yy = yb.p;
  }
  RepTmp z(ctx);
  MB_CHECK(rep_alloc_like(ctx, xx->s[0], &z.p));
  AesKey seeds[3];
  MB_CHECK(gen_zero_seeds(ctx, seeds));
  MB_CHECK(k_rep_mul(ctx, seeds, xx, yy, z.p->s));
  *out = z.release();
  return MB_OK;
}

// v_i = x_i.(y_i + y_{i+1}) + x_{i+1}.y_i for the three parties in one pass over x (gemm_imad.cu)
static int rep_dot_skinny(mb_ctx *ctx, mb_dtype dt, const RepDotParts &parts, const mb_rep *y, mb_rep *z, int64_t M, int64_t N) {
  Tmp ys0(ctx), ys1(ctx), ys2(ctx);
  MB_CHECK(mb_add(ctx, y->s[0], y->s[1], &ys0.p));
  MB_CHECK(mb_add(ctx, y->s[1], y->s[2], &ys1.p));
  MB_CHECK(mb_add(ctx, y->s[2], y->s[0], &ys2.p));
  const void *ysum[3] = {ys0.p->dptr, ys1.p->dptr, ys2.p->dptr};
  const void *yy[3] = {y->s[0]->dptr, y->s[1]->dptr, y->s[2]->dptr};
  void *zz[3] = {z->s[0]->dptr, z->s[1]->dptr, z->s[2]->dptr};
  return k_rep_dot_skinny(ctx, dt, parts, ysum, yy, zz, M, N);
}

static int rep_dot_skinny_deferred(mb_ctx *ctx, mb_dtype dt, const RepDotParts &parts, const RepDotDeferred &lazy, const mb_rep *y, mb_rep *z,
                                   int64_t M, int64_t N) {
  Tmp ys0(ctx), ys1(ctx), ys2(ctx);
  MB_CHECK(mb_add(ctx, y->s[0], y->s[1], &ys0.p));
  MB_CHECK(mb_add(ctx, y->s[1], y->s[2], &ys1.p));
  MB_CHECK(mb_add(ctx, y->s[2], y->s[0], &ys2.p));
  const void *ysum[3] = {ys0.p->dptr, ys1.p->dptr, ys2.p->dptr};
  const void *yy[3] = {y->s[0]->dptr, y->s[1]->dptr, y->s[2]->dptr};
  void *zz[3] = {z->s[0]->dptr, z->s[1]->dptr, z->s[2]->dptr};
  return k_rep_dot_skinny_deferred(ctx, dt, parts, lazy, ysum, yy, zz, M, N);
}

// dot(concatenate(xs, axis = 1), y) without materialising the concatenation when the skinny kernel applies
// (LinearPredictor's bias trick, pymoose linear_predictor.py:33-58: the [B, 1 + features] operand is 8 GB per share
// at BASELINE config 3).  Same values, same zero share, hence the same shares as ConcatOp + DotOp.
static int rep_dot_core(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, bool zero_share, mb_rep **out);
static int rep_dot_concat_core(mb_ctx *ctx, const mb_rep *const *xs, int n, const mb_rep *y, bool zero_share, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_dot_concat");
  MB_REQUIRE(ctx, xs && n > 0 && rep_valid(y) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_dot_concat: NULL argument");
  const mb_dtype dt = y->s[0]->dtype;
  bool fused = n <= 4 && mb_is_ring(dt) && y->s[0]->rank == 2 && ctx->gemm_mode != 2;
  int64_t M = 0, Ktot = 0;
  bool any_deferred = false;
  for (int p = 0; p < n; p++) {
    MB_REQUIRE(ctx, xs[p] && (rep_valid(xs[p]) || xs[p]->deferred), MB_ERR_INVALID_ARGUMENT, "mb_rep_dot_concat: NULL argument");
    const mb_tensor *t = xs[p]->deferred ? xs[p]->deferred->src : xs[p]->s[0];
    any_deferred = any_deferred || xs[p]->deferred;
    if (t->dtype != dt || t->rank != 2 || (p && t->shape[0] != M)) fused = false;
    if (p == 0) M = t->rank ? t->shape[0] : 0;
    Ktot += t->rank == 2 ? t->shape[1] : 0;
  }
  const int64_t N = y->s[0]->rank == 2 ? y->s[0]->shape[1] : 0;
  fused = fused && Ktot == y->s[0]->shape[0] && k_rep_dot_skinny_supported(M, N) &&
          !(ctx->gemm_mode == 0 && k_gemm_limb_supported(dt, M, 2 * Ktot, N) && k_gemm_limb_preferred(M, 2 * Ktot, N));
  RepDotParts parts;
  RepDotDeferred lazy;
  if (fused) {
    parts.n = n;
    int64_t row = 0;
    for (int p = 0; p < n; p++) {
      const mb_rep_deferred *d = xs[p]->deferred;
      lazy.owner[p] = d ? d->owner : -1;
      lazy.plain[p] = d ? d->src->dptr : nullptr;
      if (d) lazy.key[p] = d->key;
      for (int i = 0; i < 3; i++) parts.p[p].x[i] = d ? nullptr : xs[p]->s[i]->dptr;
      parts.p[p].K = d ? d->src->shape[1] : xs[p]->s[0]->shape[1];
      parts.p[p].krow0 = row;
      row += parts.p[p].K;
    }
  }
  const bool in_kernel = fused && any_deferred && ctx->fused_bits && k_rep_dot_skinny_deferred_supported(dt, parts, N);
  if (any_deferred && !in_kernel) {
    // no kernel expands the sharing for this shape: write the shares out (what mb_rep_share would have done) and go on
    for (int p = 0; p < n; p++) MB_CHECK(mb_rep_materialize(ctx, const_cast<mb_rep *>(xs[p])));
    return rep_dot_concat_core(ctx, xs, n, y, zero_share, out);
  }
  if (!fused) {
    RepTmp cat(ctx);
    MB_CHECK(mb_rep_concat(ctx, 1, xs, n, &cat.p));
    return rep_dot_core(ctx, cat.p, y, zero_share, out);
  }
  RepTmp z(ctx);
  const int64_t oshape[2] = {M, N};
  MB_CHECK(rep_alloc(ctx, dt, 2, oshape, &z.p));
  if (in_kernel) MB_CHECK(rep_dot_skinny_deferred(ctx, dt, parts, lazy, y, z.p, M, N));
  else MB_CHECK(rep_dot_skinny(ctx, dt, parts, y, z.p, M, N));
  AesKey seeds[3];
  MB_CHECK(gen_zero_seeds(ctx, seeds));
  if (zero_share) MB_CHECK(k_add_zero_share(ctx, seeds, z.p->s, z.p->s));
  *out = z.release();
  return MB_OK;
}
extern "C" int mb_rep_dot_concat(mb_ctx *ctx, const mb_rep *const *xs, int n, const mb_rep *y, mb_rep **out) {
  return rep_dot_concat_core(ctx, xs, n, y, true, out);
}

// zero_share = false: the products v_i without alpha_i (the zero share's entropy requests are still drawn, in order) --
// for callers that only open the sum v_0 + v_1 + v_2, where the alphas cancel (the fused fixed-point dot below)
static int rep_dot_core(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, bool zero_share, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_dot");
  MB_REQUIRE(ctx, rep_valid(x) && rep_valid(y) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_dot: NULL argument");
  mb_dtype dt = x->s[0]->dtype;
  MB_REQUIRE(ctx, mb_is_ring(dt) && dt == y->s[0]->dtype, MB_ERR_TYPE_MISMATCH, "mb_rep_dot: needs ring tensors of one width");
  int64_t M, K, N, oshape[2];
  int orank;
  MB_CHECK(mb_dot_dims(ctx, x->s[0], y->s[0], &M, &K, &N, &orank, oshape));
  RepTmp z(ctx);
  MB_CHECK(rep_alloc(ctx, dt, orank, oshape, &z.p));
  bool limb_ok = k_gemm_limb_supported(dt, M, 2 * K, N);
  MB_REQUIRE(ctx, !(ctx->gemm_mode == 2 && !limb_ok), MB_ERR_KERNEL,
             "mb_rep_dot: shape not supported by the tensor-pipe GEMM (gemm mode 2)");
  const bool use_limb = limb_ok && (ctx->gemm_mode == 2 || (ctx->gemm_mode == 0 && k_gemm_limb_preferred(M, 2 * K, N)));
  // v_i = x_i.y_i + x_i.y_{i+1} + x_{i+1}.y_i = x_i.(y_i + y_{i+1}) + x_{i+1}.y_i   (arith.rs:461-469;
  // identical mod 2^k, one GEMM less per party)
  if (use_limb) {
    // one batched launch for the three parties; x_0,x_1,x_2 are limb-split once each
    LimbOperandA As[3];
    LimbOperandB Bs[6];
    int a_idx[3][2], b_idx[3][2];
    void *Cs[3];
    for (int i = 0; i < 3; i++) {
      int j = (i + 1) % 3;
      As[i].ptr = x->s[i]->dptr;
      Bs[i].b0 = y->s[i]->dptr;      Bs[i].b1 = y->s[j]->dptr;   // y_i + y_{i+1}
      Bs[3 + i].b0 = y->s[i]->dptr;  Bs[3 + i].b1 = nullptr;     // y_i
      a_idx[i][0] = i;  b_idx[i][0] = i;
      a_idx[i][1] = j;  b_idx[i][1] = 3 + i;
      Cs[i] = z.p->s[i]->dptr;
    }
    MB_CHECK(k_gemm_limb_batched(ctx, dt, As, 3, Bs, 6, a_idx, b_idx, 3, 2, Cs, M, K, N));
  } else if (k_rep_dot_skinny_supported(M, N) && x->s[0]->rank == 2 && y->s[0]->rank == 2) {
    RepDotParts parts;
    parts.n = 1;
    for (int i = 0; i < 3; i++) parts.p[0].x[i] = x->s[i]->dptr;
    parts.p[0].K = K;
    parts.p[0].krow0 = 0;
    MB_CHECK(rep_dot_skinny(ctx, dt, parts, y, z.p, M, N));
  } else {
    for (int i = 0; i < 3; i++) {
      int j = (i + 1) % 3;
      Tmp ysum(ctx);
      MB_CHECK(mb_add(ctx, y->s[i], y->s[j], &ysum.p));
      MB_CHECK(k_gemm_imad(ctx, dt, x->s[i]->dptr, ysum.p->dptr, z.p->s[i]->dptr, M, K, N, false));
      MB_CHECK(k_gemm_imad(ctx, dt, x->s[j]->dptr, y->s[i]->dptr, z.p->s[i]->dptr, M, K, N, true));
    }
  }
  AesKey seeds[3];
  MB_CHECK(gen_zero_seeds(ctx, seeds));
  if (zero_share) MB_CHECK(k_add_zero_share(ctx, seeds, z.p->s, z.p->s));
  *out = z.release();
  return MB_OK;
}
extern "C" int mb_rep_dot(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out) { return rep_dot_core(ctx, x, y, true, out); }

// Fixed-point dot / mul of the logical layer: DotOp / MulOp followed by TruncPrOp (fixedpoint/ops.rs:761-928, 548-672;
// logical/ops.rs:776-821, 553-600).  With the three parties in one process only the truncated shares leave the call, and
// TruncPr opens x_0 + x_1 + x_2, in which the product's zero share cancels: its streams are never expanded and the
// separate pass that adds it disappears.  Entropy is drawn exactly as by mb_rep_dot + mb_rep_trunc_pr, so the result
// shares are those of the two-call sequence bit for bit (tests/test_gpu_replicated.py).
extern "C" int mb_rep_fixed_dot(mb_ctx *ctx, uint32_t amount, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_fixed_dot");
  RepTmp v(ctx);
  MB_CHECK(rep_dot_core(ctx, x, y, !ctx->fused_bits, &v.p));   // fused kernels off (A/B tests): the two-call sequence
  return mb_rep_trunc_pr(ctx, amount, v.p, out);
}
extern "C" int mb_rep_fixed_dot_concat(mb_ctx *ctx, uint32_t amount, const mb_rep *const *xs, int n, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_fixed_dot_concat");
  RepTmp v(ctx);
  MB_CHECK(rep_dot_concat_core(ctx, xs, n, y, !ctx->fused_bits, &v.p));
  return mb_rep_trunc_pr(ctx, amount, v.p, out);
}
extern "C" int mb_rep_fixed_mul(mb_ctx *ctx, uint32_t amount, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_fixed_mul");
  MB_REQUIRE(ctx, rep_valid(x) && rep_valid(y) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_fixed_mul: NULL argument");
  const mb_dtype dt = x->s[0]->dtype;
  const bool same = mb_is_ring(dt) && dt == y->s[0]->dtype && mb_same_shape(x->s[0], y->s[0]) && uniform_shapes(x, y) && x->s[0]->numel > 0;
  const unsigned k = (dt == MB_RING128 ? 128 : 64) - 1;
  if (!same || amount + 1 > k || !ctx->fused_bits) {   // broadcasting operands / argument errors: the two-call sequence handles and reports them
    RepTmp m(ctx);
    MB_CHECK(mb_rep_mul(ctx, x, y, &m.p));
    return mb_rep_trunc_pr(ctx, amount, m.p, out);
  }
  AesKey zero[3], tk[3];
  MB_CHECK(gen_zero_seeds(ctx, zero));          // drawn, not expanded
  MB_CHECK(mb_draw_trunc_keys(ctx, tk));
  RepTmp z(ctx);
  MB_CHECK(rep_alloc_like(ctx, x->s[0], &z.p));
  MB_CHECK(k_mul_trunc(ctx, tk, amount, x, y, z.p->s));
  *out = z.release();
  return MB_OK;
}

// ---- truncation -----------------------------------------------------------------------

extern "C" int mb_rep_trunc_pr(mb_ctx *ctx, uint32_t amount, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_trunc_pr");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_trunc_pr: NULL argument");
  mb_dtype dt = x->s[0]->dtype;
  MB_REQUIRE(ctx, mb_is_ring(dt), MB_ERR_TYPE_MISMATCH, "mb_rep_trunc_pr: needs ring shares");
  const unsigned k = (dt == MB_RING128 ? 128 : 64) - 1;
  // trunc.rs:139-140: `k - amount - 1` on usize panics when it underflows
  MB_REQUIRE(ctx, amount + 1 <= k, MB_ERR_INVALID_ARGUMENT, "mb_rep_trunc_pr: amount too large for the ring");
  // Entropy order: gen_trunc_mask (trunc.rs:45-63): Sample seed, PrfKeyGen, 3 x SyncKey::random();
  // adt_to_rep (convert.rs:426-431): 2 x SyncKey::random(), PrfKeyGen.
  uint8_t seed_r[16], key[16], sync[3][16], sync0[16], sync1[16], key2[16];
  MB_CHECK(mb_entropy16(ctx, seed_r));
  MB_CHECK(mb_entropy16(ctx, key));
  for (int i = 0; i < 3; i++) MB_CHECK(mb_entropy16(ctx, sync[i]));
  MB_CHECK(mb_entropy16(ctx, sync0));
  MB_CHECK(mb_entropy16(ctx, sync1));
  MB_CHECK(mb_entropy16(ctx, key2));
  AesKey seeds[6];
  mb_aes_expand(seed_r, &seeds[0]);
  for (int i = 0; i < 3; i++) seed_to_aes(ctx, sync[i], key, &seeds[1 + i]);  // r, r_top, r_msb shares
  seed_to_aes(ctx, sync0, key2, &seeds[4]);
  seed_to_aes(ctx, sync1, key2, &seeds[5]);
  RepTmp z(ctx);
  MB_CHECK(rep_alloc_like(ctx, x->s[0], &z.p));
  MB_CHECK(k_rep_trunc_pr(ctx, seeds, amount, x, z.p->s));
  *out = z.release();
  return MB_OK;
}

// ---- bits -----------------------------------------------------------------------------

extern "C" int mb_rep_index_axis0(mb_ctx *ctx, int64_t index, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_index_axis0");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_index_axis0: NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  for (int i = 0; i < 3; i++) {
    int rc = mb_index_axis0(ctx, index, x->s[i], &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

extern "C" int mb_rep_shl_dim(mb_ctx *ctx, uint64_t amount, uint64_t bit_length, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_shl_dim");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_shl_dim: NULL argument");
  mb_tensor *z[3] = {nullptr, nullptr, nullptr};
  for (int i = 0; i < 3; i++) {
    int rc = mb_shl_dim(ctx, amount, bit_length, x->s[i], &z[i]);
    if (rc != MB_OK) {
      for (int j = 0; j < i; j++) mb_tensor_free(ctx, z[j]);
      return rc;
    }
  }
  *out = rep_make(z[0], z[1], z[2]);
  return MB_OK;
}

// Fused split + Kogge-Stone adder (bits_fused.cu).  Seeds are derived in the order the op-by-op protocol draws
// them: share of bits(x0+x1) by player0 (setup, 1 sync key), first AND (3), then per level AND(p,g1) (3), AND(p,p1) (3).
static int rep_adder_fused(mb_ctx *ctx, const mb_rep *x, bool msb_only, mb_rep **out, int out_planes = 0, const mb_rep *minus = nullptr) {
  const mb_tensor *x0 = x->s[0];
  const int k = x0->dtype == MB_RING128 ? 128 : 64;
  const int levels = k == 128 ? 7 : 6;
  std::vector<AesKey> keys(4 + 6 * levels);
  MB_CHECK(ensure_setup(ctx));
  uint8_t sync[16];
  MB_CHECK(mb_entropy16(ctx, sync));
  seed_to_aes(ctx, sync, ctx->setup_keys[0], &keys[0]);
  MB_CHECK(gen_zero_seeds(ctx, &keys[1]));
  for (int i = 0; i < levels; i++) {
    MB_CHECK(gen_zero_seeds(ctx, &keys[4 + 6 * i]));
    MB_CHECK(gen_zero_seeds(ctx, &keys[4 + 6 * i + 3]));
  }
  int64_t shape[MB_MAX_RANK];
  int rank = 0;
  if (!msb_only) shape[rank++] = k;
  for (int d = 0; d < x0->rank; d++) shape[rank++] = x0->shape[d];
  RepTmp z(ctx);
  MB_CHECK(rep_alloc(ctx, MB_BIT, rank, shape, &z.p));
  MB_CHECK(k_rep_adder(ctx, keys.data(), x, msb_only, z.p->s, out_planes, minus));
  *out = z.release();
  return MB_OK;
}

// BitDecomposeOp::rep_ring_kernel = split (bits.rs:84-134) + Kogge-Stone binary_adder
// (misc.rs:176-242): fused per-level kernels when every plane is vector aligned, else composed from
// the fused share / and / xor kernels.
extern "C" int mb_rep_bit_decompose(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_bit_decompose");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_bit_decompose: NULL argument");
  mb_dtype dt = x->s[0]->dtype;
  MB_REQUIRE(ctx, mb_is_ring(dt), MB_ERR_TYPE_MISMATCH, "mb_rep_bit_decompose: needs ring shares");
  MB_REQUIRE(ctx, x->s[0]->rank < MB_MAX_RANK, MB_ERR_KERNEL, "mb_rep_bit_decompose: rank too large");
  if (ctx->fused_bits && k_rep_adder_supported(x->s[0])) return rep_adder_fused(ctx, x, false, out);
  const unsigned L = dt == MB_RING128 ? 128 : 64;
  // split: bsl = bits(x0 + x1) shared by player0; bsr = (0, 0, bits(x2))
  Tmp left(ctx), bsl(ctx), x2b(ctx);
  MB_CHECK(mb_add(ctx, x->s[0], x->s[1], &left.p));
  MB_CHECK(mb_bit_decompose(ctx, left.p, &bsl.p));
  MB_CHECK(mb_bit_decompose(ctx, x->s[2], &x2b.p));
  RepTmp a(ctx), b(ctx);
  MB_CHECK(mb_rep_share(ctx, bsl.p, 0, &a.p));
  {
    Tmp z0(ctx), z1(ctx);
    MB_CHECK(mb_fill(ctx, MB_BIT, x2b.p->rank, x2b.p->shape, 0, 0, &z0.p));
    MB_CHECK(mb_fill(ctx, MB_BIT, x2b.p->rank, x2b.p->shape, 0, 0, &z1.p));
    b.p = rep_make(z0.release(), z1.release(), x2b.release());
  }
  // binary adder
  RepTmp g(ctx), p_store(ctx), p(ctx);
  MB_CHECK(mb_rep_mul(ctx, a.p, b.p, &g.p));
  MB_CHECK(mb_rep_add(ctx, a.p, b.p, &p_store.p));
  MB_CHECK(mb_rep_from_shares(ctx, p_store.p->s[0], p_store.p->s[1], p_store.p->s[2], &p.p));
  unsigned log_r = L == 128 ? 7 : 6;
  for (unsigned i = 0; i < log_r; i++) {
    RepTmp p1(ctx), g1(ctx), pg(ctx), gn(ctx), pn(ctx);
    MB_CHECK(mb_rep_shl_dim(ctx, 1ull << i, L, p.p, &p1.p));
    MB_CHECK(mb_rep_shl_dim(ctx, 1ull << i, L, g.p, &g1.p));
    MB_CHECK(mb_rep_mul(ctx, p.p, g1.p, &pg.p));
    MB_CHECK(mb_rep_add(ctx, g.p, pg.p, &gn.p));
    MB_CHECK(mb_rep_mul(ctx, p.p, p1.p, &pn.p));
    mb_rep_free(ctx, g.p); g.p = gn.release();
    mb_rep_free(ctx, p.p); p.p = pn.release();
  }
  RepTmp c(ctx);
  MB_CHECK(mb_rep_shl_dim(ctx, 1, L, g.p, &c.p));
  return mb_rep_add(ctx, c.p, p_store.p, out);
}

int mb_rep_bit_decompose_low(mb_ctx *ctx, const mb_rep *x, int planes, mb_rep **out) {
  if (ctx && ctx->fused_bits && rep_valid(x) && mb_is_ring(x->s[0]->dtype) && x->s[0]->rank < MB_MAX_RANK && k_rep_adder_supported(x->s[0]))
    return rep_adder_fused(ctx, x, false, out, planes);
  return mb_rep_bit_decompose(ctx, x, out);
}

extern "C" int mb_rep_msb(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_msb");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_msb: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(x->s[0]->dtype), MB_ERR_TYPE_MISMATCH, "mb_rep_msb: needs ring shares");
  // msb-only pruned adder: bit-identical to plane k-1 of the full decomposition
  if (ctx->fused_bits && k_rep_adder_supported(x->s[0])) return rep_adder_fused(ctx, x, true, out);
  RepTmp bits(ctx);
  MB_CHECK(mb_rep_bit_decompose(ctx, x, &bits.p));
  int L = x->s[0]->dtype == MB_RING128 ? 128 : 64;
  return mb_rep_index_axis0(ctx, L - 1, bits.p, out);
}

// RingInjectOp::rep_kernel (convert.rs:316-390) with gen_dabit (dabit.rs:38-69).
extern "C" int mb_rep_ring_inject(mb_ctx *ctx, mb_dtype ring, uint64_t bit_idx, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_ring_inject");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_ring_inject: NULL argument");
  MB_REQUIRE(ctx, x->s[0]->dtype == MB_BIT && mb_is_ring(ring), MB_ERR_TYPE_MISMATCH, "mb_rep_ring_inject: bit shares -> ring");
  const mb_tensor *like = x->s[0];
  if (ctx->fused_bits) {
    // Entropy order: gen_dabit (dabit.rs:49-55): Sample seed, PrfKeyGen, 2 x SyncKey; adt_to_rep
    // (convert.rs:426-431): 2 x SyncKey, PrfKeyGen.  The dealer's bit b and both of its additive sharings cancel in
    // the result (the opened value is x ^ b and b_ring + c - 2 b_ring c = b ^ c = x), so none of their streams is
    // expanded; their entropy requests are consumed so that everything after this call sees the reference's order.
    uint8_t seed_bits[16], key[16], sync_b[16], sync_br[16], sync0[16], sync1[16], key2[16];
    MB_CHECK(mb_entropy16(ctx, seed_bits));
    MB_CHECK(mb_entropy16(ctx, key));
    MB_CHECK(mb_entropy16(ctx, sync_b));
    MB_CHECK(mb_entropy16(ctx, sync_br));
    MB_CHECK(mb_entropy16(ctx, sync0));
    MB_CHECK(mb_entropy16(ctx, sync1));
    MB_CHECK(mb_entropy16(ctx, key2));
    AesKey k_y0, k_y1;
    seed_to_aes(ctx, sync0, key2, &k_y0);
    seed_to_aes(ctx, sync1, key2, &k_y1);
    RepTmp z(ctx);
    MB_CHECK(rep_alloc(ctx, ring, like->rank, like->shape, &z.p));
    MB_CHECK(k_ring_inject_fused(ctx, k_y0, k_y1, (unsigned)bit_idx, x, z.p->s));
    *out = z.release();
    return MB_OK;
  }
  // gen_dabit: b = Sample (bits), br = ring_inject(b); key; seed_b, seed_br
  Tmp b(ctx), br(ctx), b0(ctx), b1(ctx), br0(ctx), br1(ctx);
  MB_CHECK(mb_sample(ctx, MB_BIT, like->rank, like->shape, -1, &b.p));
  MB_CHECK(mb_ring_inject(ctx, ring, 0, b.p, &br.p));
  uint8_t key[16], sync_b[16], sync_br[16], seed_b[16], seed_br[16];
  MB_CHECK(mb_entropy16(ctx, key));
  MB_CHECK(mb_entropy16(ctx, sync_b));
  MB_CHECK(mb_entropy16(ctx, sync_br));
  mb_blake3_derive_seed(ctx->sid, sync_b, key, seed_b);
  mb_blake3_derive_seed(ctx->sid, sync_br, key, seed_br);
  MB_CHECK(mb_sample_seeded(ctx, MB_BIT, like->rank, like->shape, seed_b, -1, &b0.p));
  MB_CHECK(mb_sub(ctx, b.p, b0.p, &b1.p));
  MB_CHECK(mb_sample_seeded(ctx, ring, like->rank, like->shape, seed_br, -1, &br0.p));
  MB_CHECK(mb_sub(ctx, br.p, br0.p, &br1.p));
  // x_adt = rep_to_adt(x) = [x0 + x1, x2]; c = open(x_adt + b_bin)
  Tmp xa0(ctx), c0(ctx), c1(ctx), c_open(ctx), c_ring(ctx);
  MB_CHECK(mb_add(ctx, x->s[0], x->s[1], &xa0.p));
  MB_CHECK(mb_add(ctx, xa0.p, b0.p, &c0.p));
  MB_CHECK(mb_add(ctx, x->s[2], b1.p, &c1.p));
  MB_CHECK(mb_add(ctx, c0.p, c1.p, &c_open.p));
  MB_CHECK(mb_ring_inject(ctx, ring, 0, c_open.p, &c_ring.p));
  // x_adt_ring = b_ring + c_ring - 2 b_ring c_ring  (public c_ring lands on share 0)
  Tmp m0(ctx), m1(ctx), t0(ctx), t0b(ctx), s0(ctx), t1(ctx), s1(ctx), sh0(ctx), sh1(ctx);
  MB_CHECK(mb_mul(ctx, br0.p, c_ring.p, &m0.p));
  MB_CHECK(mb_mul(ctx, br1.p, c_ring.p, &m1.p));
  MB_CHECK(mb_add(ctx, br0.p, c_ring.p, &t0.p));
  MB_CHECK(mb_sub(ctx, t0.p, m0.p, &t0b.p));
  MB_CHECK(mb_sub(ctx, t0b.p, m0.p, &s0.p));
  MB_CHECK(mb_sub(ctx, br1.p, m1.p, &t1.p));
  MB_CHECK(mb_sub(ctx, t1.p, m1.p, &s1.p));
  MB_CHECK(mb_shl(ctx, bit_idx, s0.p, &sh0.p));
  MB_CHECK(mb_shl(ctx, bit_idx, s1.p, &sh1.p));
  // adt_to_rep (convert.rs:426-447,480-484)
  uint8_t sync0[16], sync1[16], key2[16], seed1[16], seed2[16];
  MB_CHECK(mb_entropy16(ctx, sync0));
  MB_CHECK(mb_entropy16(ctx, sync1));
  MB_CHECK(mb_entropy16(ctx, key2));
  mb_blake3_derive_seed(ctx->sid, sync0, key2, seed1);
  mb_blake3_derive_seed(ctx->sid, sync1, key2, seed2);
  Tmp y0(ctx), y1(ctx), d0(ctx), d1(ctx), cc(ctx);
  MB_CHECK(mb_sample_seeded(ctx, ring, like->rank, like->shape, seed1, -1, &y0.p));
  MB_CHECK(mb_sample_seeded(ctx, ring, like->rank, like->shape, seed2, -1, &y1.p));
  MB_CHECK(mb_sub(ctx, sh0.p, y0.p, &d0.p));
  MB_CHECK(mb_sub(ctx, sh1.p, y1.p, &d1.p));
  MB_CHECK(mb_add(ctx, d0.p, d1.p, &cc.p));
  *out = rep_make(y0.release(), cc.release(), y1.release());
  return MB_OK;
}

extern "C" int mb_rep_less(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_less");
  // msb(x - y): the share-wise subtraction is folded into the adder's split when both operands are plain ring
  // tensors of one shape (no entropy is involved in the subtraction, so the order of requests is untouched)
  if (ctx->fused_bits && rep_valid(x) && rep_valid(y) && out && mb_is_ring(x->s[0]->dtype) && x->s[0]->dtype == y->s[0]->dtype &&
      uniform_shapes(x, y) && k_rep_adder_supported(x->s[0]))
    return rep_adder_fused(ctx, x, true, out, 0, y);
  RepTmp d(ctx);
  MB_CHECK(mb_rep_sub(ctx, x, y, &d.p));
  return mb_rep_msb(ctx, d.p, out);
}

extern "C" int mb_rep_mux(mb_ctx *ctx, const mb_rep *s, const mb_rep *x, const mb_rep *y, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_mux");
  MB_REQUIRE(ctx, rep_valid(s) && rep_valid(x) && rep_valid(y) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_mux: NULL argument");
  RepTmp sr(ctx), d(ctx), m(ctx);
  if (s->s[0]->dtype == MB_BIT && mb_is_ring(x->s[0]->dtype)) {
    // rep_bit_selector_kernel (control_flow.rs:27-45): the selector is injected into the ring first
    MB_CHECK(mb_rep_ring_inject(ctx, x->s[0]->dtype, 0, s, &sr.p));
    s = sr.p;
  }
  MB_CHECK(mb_rep_sub(ctx, x, y, &d.p));
  MB_CHECK(mb_rep_mul(ctx, s, d.p, &m.p));
  return mb_rep_add(ctx, m.p, y, out);
}

// BitComposeOp::rep_kernel (replicated/bits.rs:23-46): fold over the k planes, acc = acc + ring_inject(i, plane_i),
// starting from a replicated zero; one dabit per plane, drawn in plane order like the reference's fold.
extern "C" int mb_rep_bit_compose(mb_ctx *ctx, mb_dtype ring, const mb_rep *x_bits, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_bit_compose");
  MB_REQUIRE(ctx, rep_valid(x_bits) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_bit_compose: NULL argument");
  MB_REQUIRE(ctx, mb_is_ring(ring), MB_ERR_TYPE_MISMATCH, "mb_rep_bit_compose: the result type must be a ring");
  const mb_tensor *p = x_bits->s[0];
  const int64_t k = ring == MB_RING64 ? 64 : 128;
  MB_REQUIRE(ctx, p->dtype == MB_BIT, MB_ERR_TYPE_MISMATCH, "mb_rep_bit_compose: needs replicated bit planes");
  MB_REQUIRE(ctx, p->rank >= 1 && p->shape[0] == k, MB_ERR_INVALID_ARGUMENT, "mb_rep_bit_compose: axis 0 must hold one plane per ring bit");
  RepTmp acc(ctx);
  MB_CHECK(mb_rep_fill(ctx, ring, p->rank - 1, p->shape + 1, 0, 0, &acc.p));
  for (int64_t i = 0; i < k; i++) {
    RepTmp plane(ctx), inj(ctx), next(ctx);
    MB_CHECK(mb_rep_index_axis0(ctx, i, x_bits, &plane.p));
    MB_CHECK(mb_rep_ring_inject(ctx, ring, (uint64_t)i, plane.p, &inj.p));
    MB_CHECK(mb_rep_add(ctx, acc.p, inj.p, &next.p));
    std::swap(acc.p, next.p);
  }
  *out = acc.release();
  return MB_OK;
}

extern "C" int mb_rep_maximum(mb_ctx *ctx, const mb_rep *const *xs, int n, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_maximum");
  MB_REQUIRE(ctx, xs && out && n > 0, MB_ERR_INVALID_ARGUMENT, "maximum op needs a non-empty array of tensors");
  if (n == 1) {
    MB_REQUIRE(ctx, rep_valid(xs[0]), MB_ERR_INVALID_ARGUMENT, "mb_rep_maximum: NULL operand");
    return mb_rep_from_shares(ctx, xs[0]->s[0], xs[0]->s[1], xs[0]->s[2], out);
  }
  RepTmp m1(ctx), m2(ctx);
  MB_CHECK(mb_rep_maximum(ctx, xs, n / 2, &m1.p));
  MB_CHECK(mb_rep_maximum(ctx, xs + n / 2, n - n / 2, &m2.p));
  mb_dtype dt = m1.p->s[0]->dtype;
  RepTmp lesser(ctx), lr(ctx), a(ctx), om(ctx), b(ctx);
  MB_CHECK(mb_rep_less(ctx, m1.p, m2.p, &lesser.p));
  if (ctx->fused_bits && mb_is_ring(dt) && uniform_shapes(m1.p, m2.p) && m1.p->s[0]->numel > 0 && mb_same_shape(lesser.p->s[0], m1.p->s[0]) &&
      lesser.p->s[0]->dtype == MB_BIT) {
    // ring_inject, the two multiplications, the public subtraction and the addition in one pass; keys in the op order
    // below (ring_inject, mul, mul)
    AesKey keys[8];
    MB_CHECK(mb_draw_inject_keys(ctx, keys));
    MB_CHECK(gen_zero_seeds(ctx, keys + 2));
    MB_CHECK(gen_zero_seeds(ctx, keys + 5));
    RepTmp z(ctx);
    MB_CHECK(rep_alloc_like(ctx, m1.p->s[0], &z.p));
    MB_CHECK(k_maximum_select(ctx, keys, lesser.p, m1.p, m2.p, z.p->s));
    *out = z.release();
    return MB_OK;
  }
  MB_CHECK(mb_rep_ring_inject(ctx, dt, 0, lesser.p, &lr.p));
  Tmp ones(ctx);
  MB_CHECK(mb_fill(ctx, dt, lr.p->s[0]->rank, lr.p->s[0]->shape, 1, 0, &ones.p));
  // lesser_ring * max2 + (ones - lesser_ring) * max1, evaluated left to right (softmax.rs:45-49)
  MB_CHECK(mb_rep_mul(ctx, lr.p, m2.p, &a.p));
  MB_CHECK(mb_rep_public_sub(ctx, ones.p, lr.p, &om.p));
  MB_CHECK(mb_rep_mul(ctx, om.p, m1.p, &b.p));
  return mb_rep_add(ctx, a.p, b.p, out);
}

extern "C" int mb_rep_relu(mb_ctx *ctx, const mb_rep *x, mb_rep **out) {
  MB_ENTER(ctx, "mb_rep_relu");
  MB_REQUIRE(ctx, rep_valid(x) && out, MB_ERR_INVALID_ARGUMENT, "mb_rep_relu: NULL argument");
  mb_dtype dt = x->s[0]->dtype;
  RepTmp msb(ctx), sign(ctx), zeros(ctx);
  MB_CHECK(mb_rep_msb(ctx, x, &msb.p));
  MB_CHECK(mb_rep_ring_inject(ctx, dt, 0, msb.p, &sign.p));
  MB_CHECK(rep_alloc_like(ctx, x->s[0], &zeros.p));
  for (int i = 0; i < 3; i++) MB_CHECK(k_fill(ctx, zeros.p->s[i], 0, 0));
  return mb_rep_mux(ctx, sign.p, zeros.p, x, out);
}
