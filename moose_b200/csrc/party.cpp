// C ABI: ONE party of the three-party replicated protocols per process / GPU (BASELINE config 5, SURVEY.md 8e
// "party-per-GPU").  Party i holds the shares (x_i, x_{i+1}) of every replicated tensor and computes what player i
// computes in the reference; everything the reference moves between hosts (reshare after mul / dot / and, the
// dealer's messages of TruncPr and of the dabit, the openings) goes through the transport callback, which the
// host binds to NCCL send/recv over NVLink.  Protocol sources restated here, per player:
//   setup / zero shares    moose/src/replicated/setup.rs:39-58, zero_share.rs:82-156
//   share / reveal         moose/src/replicated/convert.rs:49-160, 280-313
//   mul / dot / and        moose/src/replicated/arith.rs:317-367, 436-508
//   TruncPr                moose/src/replicated/fixedpoint.rs:80-103; additive/convert.rs:12-79; additive/trunc.rs:22-169;
//                          replicated/convert.rs:392-497 (adt_to_rep)
//   msb                    moose/src/replicated/bits.rs:6-134, misc.rs:176-242 (split + Kogge-Stone), arith.rs:611-626
//   ring_inject + dabit    moose/src/replicated/convert.rs:316-390; additive/dabit.rs:38-69
//   less / maximum         moose/src/replicated/compare.rs:91-104, softmax.rs:10-53
// The heavy steps use the fused kernels of the single-GPU path restricted to this party's component: the msb-pruned
// Kogge-Stone levels (one exchange of the surviving planes per level instead of a [k, n] bit tensor per AND) and the
// two-term tensor-pipe GEMM.  Values a seed determines travel as the 16-byte seed, not as the expanded tensor.
//
// Randomness: PRF keys and dealer masks come from this party's own entropy; sync keys are public and must agree on
// the three parties, so they are drawn from a counter-mode stream keyed by the session id (all parties execute the
// same protocol sequence, the reference fixes them at compile time, zero_share.rs:82-85).
#include <chrono>
#include <deque>
#include <functional>

#include "common.h"

int mb_dot_dims(const mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, int64_t *M, int64_t *K, int64_t *N, int *orank,
                int64_t *oshape);

struct mb_party {
  mb_ctx *ctx;
  int me;
  mb_transport_fn tx;
  void *user;
  std::string names[3];
  bool have_setup;
  bool stream_ordered;  // transport capability bit 0: device payloads are ordered on the context's stream
  bool can_reserve;     // capability bit 1: a kernel may write its payload into the receiver's memory (modes 6 / 7)
  uint8_t keys[2][16];  // k_me (own) and k_{me+1} (copy received from the next party)
  AesKey sync_cipher;
  uint64_t sync_counter;
  std::deque<mb_tensor *> in_flight;
  uint64_t sends_reported, bytes_sent, transfers;
  double transport_seconds, sync_seconds;
};

struct mb_pshare {
  mb_tensor *t[2];  // shares me and me+1
};

namespace {

struct PErr { int code; };

struct T {  // owning tensor handle
  mb_ctx *c;
  mb_tensor *p;
  T(mb_ctx *c_, mb_tensor *p_) : c(c_), p(p_) {}
  T(T &&o) noexcept : c(o.c), p(o.p) { o.p = nullptr; }
  T &operator=(T &&o) noexcept {
    if (this != &o) {
      if (p) mb_tensor_free(c, p);
      c = o.c;
      p = o.p;
      o.p = nullptr;
    }
    return *this;
  }
  T(const T &) = delete;
  T &operator=(const T &) = delete;
  ~T() { if (p) mb_tensor_free(c, p); }
  T clone() const { return T(c, p ? mb_tensor_retain(p) : nullptr); }
  mb_tensor *release() { mb_tensor *q = p; p = nullptr; return q; }
};

struct Pair {  // this party's two shares
  T a, b;
};

struct P {
  mb_party *pt;
  mb_ctx *c;
  int me, nx, pv;
  explicit P(mb_party *p) : pt(p), c(p->ctx), me(p->me), nx((p->me + 1) % 3), pv((p->me + 2) % 3) {}

  void ck(int rc) { if (rc != MB_OK) throw PErr{rc}; }
  [[noreturn]] void fail(int code, const std::string &m) { mb_fail(c, code, m); throw PErr{code}; }

  // ---- transport --------------------------------------------------------------------------------------
  static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
  void raw(int mode, int peer, void *buf, size_t n, int dev) {
    const double t0 = now_s();
    const int rc = pt->tx(pt->user, mode, peer >= 0 ? pt->names[peer].c_str() : "", buf, n, dev);
    pt->transport_seconds += now_s() - t0;
    if (rc != 0) fail(MB_ERR_KERNEL, "party transport failed");
  }
  void sync_stream() {
    if (pt->stream_ordered) return;
    const double t0 = now_s();
    ck(mb_ctx_sync(c));
    pt->sync_seconds += now_s() - t0;
  }
  // receive buffer: off the main stream's allocation order unless the transport writes it in stream order
  mb_tensor *recv_buf(mb_dtype d, int rank, const int64_t *shape) {
    mb_tensor *t = nullptr;
    ck(pt->stream_ordered ? mb_tensor_new(c, d, rank, shape, &t) : mb_tensor_new_detached(c, d, rank, shape, &t));
    return t;
  }
  void reap(bool all) {
    if (pt->in_flight.empty() && !all) return;
    int64_t done = 0;
    raw(all ? 3 : 2, -1, &done, sizeof(done), 0);
    while (pt->sends_reported < (uint64_t)done && !pt->in_flight.empty()) {
      mb_tensor_free(c, pt->in_flight.front());
      pt->in_flight.pop_front();
      pt->sends_reported++;
    }
  }
  void send_t(int peer, const mb_tensor *t) {
    int64_t hdr[10] = {0};
    hdr[0] = 0;
    hdr[1] = t->dtype;
    hdr[2] = t->rank;
    for (int k = 0; k < t->rank; k++) hdr[3 + k] = t->shape[k];
    hdr[9] = (int64_t)t->nbytes;
    sync_stream();  // the payload is read by another stream / library
    pt->transfers++;
    raw(1, peer, hdr, sizeof(hdr), 0);
    if (t->nbytes) {
      raw(1, peer, t->dptr, t->nbytes, 1);
      pt->in_flight.push_back(mb_tensor_retain(const_cast<mb_tensor *>(t)));
    }
    pt->bytes_sent += t->nbytes;
    reap(false);
  }
  T recv_t(int peer) {
    int64_t hdr[10] = {0};
    raw(0, peer, hdr, sizeof(hdr), 0);
    if (hdr[0] != 0 || hdr[2] < 0 || hdr[2] > MB_MAX_RANK) fail(MB_ERR_KERNEL, "party transport: expected a tensor");
    if (hdr[1] != MB_RING64 && hdr[1] != MB_RING128 && hdr[1] != MB_BIT && hdr[1] != MB_FLOAT64)
      fail(MB_ERR_KERNEL, "party transport: unknown dtype in a tensor header");
    mb_tensor *t = recv_buf((mb_dtype)hdr[1], (int)hdr[2], hdr + 3);
    T out(c, t);
    if ((int64_t)t->nbytes != hdr[9]) fail(MB_ERR_KERNEL, "party transport: payload size mismatch");
    if (t->nbytes) raw(0, peer, t->dptr, t->nbytes, 1);
    return out;
  }
  void send_b(int peer, const uint8_t *b, size_t n) {
    int64_t hdr[10] = {0};
    hdr[0] = 1;
    hdr[9] = (int64_t)n;
    raw(1, peer, hdr, sizeof(hdr), 0);
    raw(1, peer, const_cast<uint8_t *>(b), n, 0);
  }
  void recv_b(int peer, uint8_t *b, size_t n) {
    int64_t hdr[10] = {0};
    raw(0, peer, hdr, sizeof(hdr), 0);
    if (hdr[0] != 1 || hdr[9] != (int64_t)n) fail(MB_ERR_KERNEL, "party transport: expected a byte string");
    raw(0, peer, b, n, 0);
  }
  // One directed transfer in the global order every party follows: tensor valid at `from`, result valid at `to`.
  T move(int from, int to, const mb_tensor *t) {
    if (me == from) send_t(to, t);
    if (me == to) return recv_t(from);
    return T(c, nullptr);
  }
  // The same when the receiver knows dtype and shape already (SPMD: it does almost everywhere): payload only.
  T move_like(int from, int to, const mb_tensor *t, mb_dtype d, const mb_tensor *shape_of) {
    if (me == from) {
      sync_stream();
      pt->transfers++;
      if (t->nbytes) {
        raw(1, to, t->dptr, t->nbytes, 1);
        pt->in_flight.push_back(mb_tensor_retain(const_cast<mb_tensor *>(t)));
      }
      pt->bytes_sent += t->nbytes;
      reap(false);
    }
    if (me == to) {
      mb_tensor *r = recv_buf(d, shape_of->rank, shape_of->shape);
      T out(c, r);
      if (r->nbytes) raw(0, from, r->dptr, r->nbytes, 1);
      return out;
    }
    return T(c, nullptr);
  }
  void move_b(int from, int to, uint8_t *b, size_t n) {
    if (me == from) send_b(to, b, n);
    if (me == to) recv_b(from, b, n);
  }
  // Every party hands its component to the previous party and gets the next party's, which has the same dtype and
  // shape: one grouped send + receive (transport mode 4), the three transfers run concurrently.
  T ring(const mb_tensor *mine) {
    mb_tensor *t = recv_buf(mine->dtype, mine->rank, mine->shape);
    T got(c, t);
    if (mine->nbytes == 0) return got;
    sync_stream();  // the payload is read by another stream / library
    mb_sendrecv d;
    d.send_peer = pt->names[pv].c_str();
    d.send_buf = mine->dptr;
    d.send_bytes = mine->nbytes;
    d.recv_peer = pt->names[nx].c_str();
    d.recv_buf = t->dptr;
    d.recv_bytes = t->nbytes;
    pt->transfers++;
    raw(4, -1, &d, sizeof(d), 1);
    pt->bytes_sent += mine->nbytes;
    return got;
  }

  // Reshare fused into the producing kernel: memory of the previous party for the next payload of `nbytes` (NULL:
  // the transport cannot, or the payload is too large for its ring -- the kernel then writes locally only and
  // ring_after() sends the ordinary way).
  void *reserve_prev(size_t nbytes) {
    if (!pt->can_reserve || nbytes == 0) return nullptr;
    mb_reserve r{pt->names[pv].c_str(), nbytes, nullptr};
    raw(6, -1, &r, sizeof(r), 1);
    return r.ptr;
  }
  // Second half of ring(): the kernel enqueued since reserve_prev() has stored `mine` into `remote` as well.
  T ring_after(const mb_tensor *mine, void *remote) {
    if (!remote) return ring(mine);
    mb_reserve r{pt->names[pv].c_str(), mine->nbytes, remote};
    raw(7, -1, &r, sizeof(r), 1);
    pt->transfers++;
    pt->bytes_sent += mine->nbytes;
    T got(c, recv_buf(mine->dtype, mine->rank, mine->shape));
    raw(0, nx, got.p->dptr, got.p->nbytes, 1);
    return got;
  }

  // ---- keys ---------------------------------------------------------------------------------------------
  void sync_key(uint8_t out[16]) {
    uint8_t ctr[16] = {0};
    const uint64_t v = pt->sync_counter++;
    memcpy(ctr, &v, 8);
    mb_aes_encrypt_host(&pt->sync_cipher, ctr, out);
  }
  void derive(const uint8_t key[16], const uint8_t sync[16], uint8_t seed[16]) { mb_blake3_derive_seed(c->sid, sync, key, seed); }
  void setup() {
    if (pt->have_setup) return;
    ck(mb_entropy16(c, pt->keys[0]));
    // k_i is generated by player i and copied to player i-1 (setup.rs:39-58)
    for (int from = 0; from < 3; from++) {
      const int to = (from + 2) % 3;
      uint8_t buf[16];
      memcpy(buf, pt->keys[0], 16);
      move_b(from, to, buf, 16);
      if (me == to) memcpy(pt->keys[1], buf, 16);
    }
    pt->have_setup = true;
  }
  // seeds of this party's zero share: s_me = DeriveSeed(k_me, sync_me), s_{me+1} = DeriveSeed(k_{me+1}, sync_{me+1})
  void zero_seeds(uint8_t sa[16], uint8_t sb[16]) {
    setup();
    uint8_t sk[3][16];
    for (int i = 0; i < 3; i++) sync_key(sk[i]);
    derive(pt->keys[0], sk[me], sa);
    derive(pt->keys[1], sk[nx], sb);
  }

  // ---- host-op shorthands ---------------------------------------------------------------------------------
  T add(const mb_tensor *x, const mb_tensor *y) { mb_tensor *t = nullptr; ck(mb_add(c, x, y, &t)); return T(c, t); }
  T sub(const mb_tensor *x, const mb_tensor *y) { mb_tensor *t = nullptr; ck(mb_sub(c, x, y, &t)); return T(c, t); }
  T mul(const mb_tensor *x, const mb_tensor *y) { mb_tensor *t = nullptr; ck(mb_mul(c, x, y, &t)); return T(c, t); }
  T neg(const mb_tensor *x) { mb_tensor *t = nullptr; ck(mb_neg(c, x, &t)); return T(c, t); }
  T shl(uint64_t a, const mb_tensor *x) { mb_tensor *t = nullptr; ck(mb_shl(c, a, x, &t)); return T(c, t); }
  T shr(uint64_t a, const mb_tensor *x) { mb_tensor *t = nullptr; ck(mb_shr(c, a, x, &t)); return T(c, t); }
  T xor_(const mb_tensor *x, const mb_tensor *y) { mb_tensor *t = nullptr; ck(mb_xor(c, x, y, &t)); return T(c, t); }
  T and_(const mb_tensor *x, const mb_tensor *y) { mb_tensor *t = nullptr; ck(mb_and(c, x, y, &t)); return T(c, t); }
  T fill(mb_dtype d, const mb_tensor *like, u128 v) {
    mb_tensor *t = nullptr;
    ck(mb_fill(c, d, like->rank, like->shape, (uint64_t)v, (uint64_t)(v >> 64), &t));
    return T(c, t);
  }
  T prf(mb_dtype d, const mb_tensor *like, const uint8_t seed[16]) {
    mb_tensor *t = nullptr;
    ck(mb_sample_seeded(c, d, like->rank, like->shape, seed, -1, &t));
    return T(c, t);
  }
  T fresh(mb_dtype d, const mb_tensor *like) {
    mb_tensor *t = nullptr;
    ck(mb_sample(c, d, like->rank, like->shape, -1, &t));
    return T(c, t);
  }
  T plane(int64_t i, const mb_tensor *x) { mb_tensor *t = nullptr; ck(mb_index_axis0(c, i, x, &t)); return T(c, t); }

  // ---- share / reveal ---------------------------------------------------------------------------------------
  // `like` carries dtype and shape on every party; x is the plaintext on the owner only.
  Pair share(int owner, const mb_tensor *x, const mb_tensor *like) {
    setup();
    uint8_t sk[16], seed[16];
    sync_key(sk);
    const mb_dtype d = like->dtype;
    const int onx = (owner + 1) % 3;
    if (me == owner) {
      derive(pt->keys[0], sk, seed);
      T xo = prf(d, like, seed);
      T xn = sub(x, xo.p);
      move(owner, onx, xn.p);
      return Pair{std::move(xo), std::move(xn)};
    }
    if (me == onx) {
      T xn = move(owner, onx, nullptr);
      return Pair{std::move(xn), fill(d, like, 0)};
    }
    derive(pt->keys[1], sk, seed);  // k_owner, held as the copy of the next party's key
    return Pair{fill(d, like, 0), prf(d, like, seed)};
  }
  // the receiver adds its two shares and the third one, sent by the next party (convert.rs:295-311)
  T reveal(int receiver, const Pair &x) {
    const int sender = (receiver + 1) % 3;
    T third = move_like(sender, receiver, x.b.p, x.a.p->dtype, x.a.p);
    if (me != receiver) return T(c, nullptr);
    return add(add(x.a.p, x.b.p).p, third.p);
  }

  // ---- share-wise ---------------------------------------------------------------------------------------------
  Pair padd(const Pair &x, const Pair &y) { return Pair{add(x.a.p, y.a.p), add(x.b.p, y.b.p)}; }
  Pair psub(const Pair &x, const Pair &y) { return Pair{sub(x.a.p, y.a.p), sub(x.b.p, y.b.p)}; }
  Pair pxor(const Pair &x, const Pair &y) { return Pair{xor_(x.a.p, y.a.p), xor_(x.b.p, y.b.p)}; }
  // a public value touches share 0 only: first share of party 0, second share of party 2 (arith.rs:81-121)
  Pair add_public(const Pair &x, const mb_tensor *pub) {
    if (me == 0) return Pair{add(x.a.p, pub), x.b.clone()};
    if (me == 2) return Pair{x.a.clone(), add(x.b.p, pub)};
    return Pair{x.a.clone(), x.b.clone()};
  }
  Pair public_sub(const mb_tensor *pub, const Pair &x) {  // pub - x (arith.rs:230-273)
    T a = me == 0 ? sub(pub, x.a.p) : neg(x.a.p);
    T b = me == 2 ? sub(pub, x.b.p) : neg(x.b.p);
    return Pair{std::move(a), std::move(b)};
  }

  // ---- bilinear: z_me = v_me + PRF(s_me) - PRF(s_{me+1}), then the reshare ring ---------------------------------
  void zero_keys(AesKey kk[2]) {
    uint8_t sa[16], sb[16];
    zero_seeds(sa, sb);
    mb_aes_expand(sa, &kk[0]);
    mb_aes_expand(sb, &kk[1]);
  }
  T like_new(mb_dtype d, const mb_tensor *like) {
    mb_tensor *t = nullptr;
    ck(mb_tensor_new(c, d, like->rank, like->shape, &t));
    return T(c, t);
  }
  Pair finish_bilinear(T v) {  // v is this party's own temporary: the zero share is added in place
    AesKey kk[2];
    zero_keys(kk);
    void *remote = reserve_prev(v.p->nbytes);
    ck(k_party_zero_add(c, kk, v.p, v.p, remote));
    T zn = ring_after(v.p, remote);
    return Pair{std::move(v), std::move(zn)};
  }
  static bool same_shape(const mb_tensor *a, const mb_tensor *b) {
    if (a->rank != b->rank || a->dtype != b->dtype) return false;
    for (int i = 0; i < a->rank; i++)
      if (a->shape[i] != b->shape[i]) return false;
    return true;
  }
  Pair pmul(const Pair &x, const Pair &y) {  // v = x_me (y_me + y_{me+1}) + x_{me+1} y_me
    if (same_shape(x.a.p, y.a.p)) {  // one pass: products, zero share
      AesKey kk[2];
      zero_keys(kk);
      T z = like_new(x.a.p->dtype, x.a.p);
      void *remote = reserve_prev(z.p->nbytes);
      ck(k_party_mul(c, kk, x.a.p, x.b.p, y.a.p, y.b.p, z.p, remote));
      T zn = ring_after(z.p, remote);
      return Pair{std::move(z), std::move(zn)};
    }
    T v = add(mul(x.a.p, add(y.a.p, y.b.p).p).p, mul(x.b.p, y.a.p).p);  // broadcasting operands
    return finish_bilinear(std::move(v));
  }
  Pair pdot(const Pair &x, const Pair &y) {
    int64_t M, K, N, oshape[2];
    int orank;
    ck(mb_dot_dims(c, x.a.p, y.a.p, &M, &K, &N, &orank, oshape));
    const mb_dtype d = x.a.p->dtype;
    mb_tensor *vt = nullptr;
    ck(mb_tensor_new(c, d, orank, oshape, &vt));
    T v(c, vt);
    if (c->gemm_mode != 1 && k_gemm_limb_supported(d, M, 2 * K, N) && (c->gemm_mode == 2 || k_gemm_limb_preferred(M, 2 * K, N))) {
      GemmTerm terms[2] = {{x.a.p->dptr, y.a.p->dptr, y.b.p->dptr}, {x.b.p->dptr, y.a.p->dptr, nullptr}};
      ck(k_gemm_limb(c, d, terms, 2, v.p->dptr, M, K, N));
    } else {
      T ys = add(y.a.p, y.b.p);
      ck(k_gemm_imad(c, d, x.a.p->dptr, ys.p->dptr, v.p->dptr, M, K, N, false));
      ck(k_gemm_imad(c, d, x.b.p->dptr, y.a.p->dptr, v.p->dptr, M, K, N, true));
    }
    return finish_bilinear(std::move(v));
  }

  // ---- adt_to_rep (convert.rs:392-497): additive (y0 at party 0, y1 at party 1), provider party 2 ------------------
  // The provider's seeds: seed1 to party 0, seed2 to party 1 (it keeps both).
  void adt_seeds(uint8_t seed1[16], uint8_t seed2[16]) {
    uint8_t sk0[16], sk1[16], key[16];
    sync_key(sk0);
    sync_key(sk1);
    if (me == 2) {
      ck(mb_entropy16(c, key));
      derive(key, sk0, seed1);
      derive(key, sk1, seed2);
    }
    move_b(2, 0, seed1, 16);
    move_b(2, 1, seed2, 16);
  }
  // parties 0 / 1 hold yp = PRF(seed) and diff = y - yp already (fused into the step that produced y)
  Pair adt_finish(T yp, T diff, mb_dtype d, const mb_tensor *like, const uint8_t seed1[16], const uint8_t seed2[16]) {
    if (me == 2) return Pair{prf(d, like, seed2), prf(d, like, seed1)};  // (y1', y0')
    T d1 = move_like(1, 0, diff.p, d, like);       // party 0 opens c = (y0 - y0') + (y1 - y1')
    T cc(c, nullptr);
    if (me == 0) cc = add(diff.p, d1.p);
    T c1 = move_like(0, 1, cc.p, d, like);
    if (me == 0) return Pair{std::move(yp), std::move(cc)};
    return Pair{std::move(c1), std::move(yp)};
  }

  // ---- TruncPr ---------------------------------------------------------------------------------------------------
  Pair trunc_pr(unsigned amount, const Pair &x) {
    const mb_tensor *like = x.a.p;
    const mb_dtype d = like->dtype;
    const unsigned bits = d == MB_RING128 ? 128 : 64, k = bits - 1;
    if (amount + 1 > k) fail(MB_ERR_INVALID_ARGUMENT, "trunc_pr: amount too large for the ring");
    // dealer (party 2): r, r_top, r_msb, additively shared: share 0 = PRF(seed) (the seed goes to party 0), share 1 to party 1
    uint8_t key[16], seeds[3][16], sk[3][16], seed1[16], seed2[16];
    for (int i = 0; i < 3; i++) sync_key(sk[i]);
    T m1[3] = {T(c, nullptr), T(c, nullptr), T(c, nullptr)};
    if (me == 2) {
      AesKey kk[4];
      ck(mb_entropy16(c, key));
      mb_aes_expand(key, &kk[0]);  // r itself
      ck(mb_entropy16(c, key));
      for (int i = 0; i < 3; i++) {
        derive(key, sk[i], seeds[i]);
        mb_aes_expand(seeds[i], &kk[1 + i]);
        m1[i] = like_new(d, like);
      }
      mb_tensor *const o[3] = {m1[0].p, m1[1].p, m1[2].p};
      ck(k_party_trunc_dealer(c, kk, amount, o));
    }
    for (int i = 0; i < 3; i++) move_b(2, 0, seeds[i], 16);
    adt_seeds(seed1, seed2);
    // message order: r to the second player first (all it needs to answer the first player), its masked share to the
    // first player, then r_top and r_msb behind it
    T r1[3] = {T(c, nullptr), T(c, nullptr), T(c, nullptr)};
    r1[0] = move_like(2, 1, m1[0].p, d, like);
    T masked1(c, nullptr);
    if (me == 1) masked1 = add(x.b.p, r1[0].p);                          // a1 = x2, masked with r1
    masked1 = move_like(1, 0, masked1.p, d, like);
    for (int i = 1; i < 3; i++) r1[i] = move_like(2, 1, m1[i].p, d, like);
    T yp(c, nullptr), diff(c, nullptr);
    if (me == 0) {
      AesKey kk[4];
      for (int i = 0; i < 3; i++) mb_aes_expand(seeds[i], &kk[i]);
      mb_aes_expand(seed1, &kk[3]);
      T cbits = like_new(MB_BIT, like);  // the opened top bit travels as a bit tensor (1/64 .. 1/128 of the ring tensor)
      yp = like_new(d, like);
      diff = like_new(d, like);
      ck(k_party_trunc_p0(c, kk, amount, x.a.p, x.b.p, masked1.p, cbits.p, yp.p, diff.p));
      move_like(0, 1, cbits.p, MB_BIT, like);
    } else if (me == 1) {
      T cbits = move_like(0, 1, nullptr, MB_BIT, like);
      AesKey kyp;
      mb_aes_expand(seed2, &kyp);
      yp = like_new(d, like);
      diff = like_new(d, like);
      ck(k_party_trunc_p1(c, kyp, amount, r1[1].p, r1[2].p, cbits.p, yp.p, diff.p));
    }
    return adt_finish(std::move(yp), std::move(diff), d, like, seed1, seed2);
  }

  // ---- msb: split + msb-pruned Kogge-Stone on this party's components -----------------------------------------------
  // msb(x - y) (y may be null): the split decomposes this player's local sum of the difference directly
  Pair msb_diff(const Pair &x, const Pair *y) {
    setup();
    const mb_tensor *like = x.a.p;
    const int k = like->dtype == MB_RING128 ? 128 : 64;
    const int64_t n = like->numel;
    if (n == 0 || n % 128 || like->rank + 1 > MB_MAX_RANK)
      fail(MB_ERR_UNIMPLEMENTED, "party msb: element count must be a positive multiple of 128 (rank <= 5)");
    const int64_t nv = n / 128;
    int64_t psh_all[MB_MAX_RANK];
    psh_all[0] = k;
    for (int r = 0; r < like->rank; r++) psh_all[r + 1] = like->shape[r];
    auto planes = [&]() {
      mb_tensor *t = nullptr;
      ck(mb_tensor_new(c, MB_BIT, like->rank + 1, psh_all, &t));
      return T(c, t);
    };
    // operands of the adder: a = bits(x0 + x1) shared by party 0 as (R, B ^ R, 0); b = (0, 0, bits(x2));
    // p0 = a ^ b = (R, B ^ R, C), g0 = a & b = (0, (B ^ R) & C, R & C) ^ zero share
    T own = planes();  // B on party 0, C = bits(x2) on parties 1 and 2
    {
      const mb_tensor *terms[4];
      int nt = 0, neg = 0;
      if (me == 0) {
        terms[nt++] = x.a.p;
        terms[nt++] = x.b.p;
        if (y) { neg = 0xC; terms[nt++] = y->a.p; terms[nt++] = y->b.p; }
      } else {
        terms[nt++] = me == 1 ? x.b.p : x.a.p;
        if (y) { neg = 0x2; terms[nt++] = me == 1 ? y->b.p : y->a.p; }
      }
      ck(k_party_split(c, terms, nt, neg, own.p));
    }
    uint8_t sk[16], seed[16] = {0};
    sync_key(sk);
    AesKey keys[3];
    if (me == 0) derive(pt->keys[0], sk, seed);
    if (me == 2) derive(pt->keys[1], sk, seed);
    mb_aes_expand(seed, &keys[0]);
    zero_keys(keys + 1);
    T g_me = planes();
    Pair p0{T(c, nullptr), T(c, nullptr)};
    void *g_remote = nullptr;
    if (me == 0) {
      T R = planes(), bR = planes();
      g_remote = reserve_prev(g_me.p->nbytes);
      ck(k_party_adder_init(c, 0, keys, own.p, nullptr, g_me.p, R.p, bR.p, g_remote));
      move_like(0, 1, bR.p, MB_BIT, own.p);
      p0 = Pair{std::move(R), std::move(bR)};
    } else if (me == 1) {
      T bR = move_like(0, 1, nullptr, MB_BIT, own.p);
      g_remote = reserve_prev(g_me.p->nbytes);
      ck(k_party_adder_init(c, 1, keys, own.p, bR.p, g_me.p, nullptr, nullptr, g_remote));
      p0 = Pair{std::move(bR), std::move(own)};
    } else {
      T R = planes();
      g_remote = reserve_prev(g_me.p->nbytes);
      ck(k_party_adder_init(c, 2, keys, own.p, nullptr, g_me.p, R.p, nullptr, g_remote));
      p0 = Pair{std::move(own), std::move(R)};
    }
    T g_nx = ring_after(g_me.p, g_remote);
    Pair g{std::move(g_me), std::move(g_nx)};
    // pruned levels
    std::vector<LevelDesc> levels;
    std::vector<int> npo;
    int final_slot = 0;
    ck(k_adder_plan_msb(k, &levels, &npo, &final_slot));
    Pair p{p0.a.clone(), p0.b.clone()};
    for (size_t i = 0; i < levels.size(); i++) {
      uint8_t s_pg[2][16], s_pp[2][16];
      zero_seeds(s_pg[0], s_pg[1]);
      zero_seeds(s_pp[0], s_pp[1]);
      AesKey kk[4];
      mb_aes_expand(s_pg[0], &kk[0]);
      mb_aes_expand(s_pg[1], &kk[1]);
      mb_aes_expand(s_pp[0], &kk[2]);
      mb_aes_expand(s_pp[1], &kk[3]);
      // one buffer [nout + npo, shape...] for this party's g' and p' components: one exchange per level
      int64_t sh[MB_MAX_RANK];
      const int64_t ng = std::max(levels[i].nout, 1), np_ = npo[i];
      sh[0] = ng + np_;
      for (int r = 0; r < like->rank; r++) sh[r + 1] = like->shape[r];
      mb_tensor *buf = nullptr;
      ck(mb_tensor_new(c, MB_BIT, like->rank + 1, sh, &buf));
      T mine(c, buf);
      int64_t gsh[MB_MAX_RANK], psh[MB_MAX_RANK];
      memcpy(gsh, sh, sizeof(sh));
      memcpy(psh, sh, sizeof(sh));
      gsh[0] = ng;
      psh[0] = std::max<int64_t>(np_, 1);
      const size_t plane_bytes = (size_t)n / 8;
      auto views = [&](const mb_tensor *whole, T &gv, T &pv_) {
        mb_tensor *a = nullptr, *b = nullptr;
        ck(mb_tensor_view(c, whole, 0, like->rank + 1, gsh, &a));
        gv = T(c, a);
        if (np_ > 0) {
          ck(mb_tensor_view(c, whole, (size_t)ng * plane_bytes, like->rank + 1, psh, &b));
          pv_ = T(c, b);
        } else {
          pv_ = gv.clone();  // never read
        }
      };
      T g_me(c, nullptr), p_me(c, nullptr), g_nx(c, nullptr), p_nx(c, nullptr);
      views(mine.p, g_me, p_me);
      const mb_tensor *gin[2] = {g.a.p, g.b.p}, *pin[2] = {p.a.p, p.b.p};
      void *remote = reserve_prev(mine.p->nbytes);
      ck(k_party_adder_level(c, kk, levels[i], gin, pin, g_me.p, p_me.p, nv, remote));
      T theirs = ring_after(mine.p, remote);
      views(theirs.p, g_nx, p_nx);
      g = Pair{std::move(g_me), std::move(g_nx)};
      p = Pair{std::move(p_me), std::move(p_nx)};
    }
    // msb = g_final[k-2] ^ p0[k-1]
    return Pair{xor_(plane(final_slot, g.a.p).p, plane(k - 1, p0.a.p).p), xor_(plane(final_slot, g.b.p).p, plane(k - 1, p0.b.p).p)};
  }

  // ---- Z2 -> ring with a dabit from the dealer (party 2) --------------------------------------------------------------
  Pair ring_inject(mb_dtype ring, unsigned bit_idx, const Pair &x) {
    const mb_tensor *like = x.a.p;  // shape carrier (a bit tensor)
    uint8_t key[16], sk_b[16], sk_br[16], seed_b[16], seed_br[16], seed1[16], seed2[16];
    sync_key(sk_b);
    sync_key(sk_br);
    T b1(c, nullptr), br1(c, nullptr);
    if (me == 2) {
      T b = fresh(MB_BIT, like);
      ck(mb_entropy16(c, key));
      derive(key, sk_b, seed_b);
      derive(key, sk_br, seed_br);
      b1 = xor_(b.p, prf(MB_BIT, like, seed_b).p);
      AesKey kbr;
      mb_aes_expand(seed_br, &kbr);
      br1 = like_new(ring, like);
      ck(k_party_dabit_dealer(c, kbr, b.p, br1.p));  // inject(b) - PRF(seed_br)
    }
    move_b(2, 0, seed_b, 16);
    move_b(2, 0, seed_br, 16);
    adt_seeds(seed1, seed2);
    // message order: the dealer's bit share, the second player's masked bit, the dealer's ring share (in flight while
    // the first player opens c), the opened bit
    T b1r = move_like(2, 1, b1.p, MB_BIT, like);
    T c0(c, nullptr), c1(c, nullptr);
    if (me == 0) c0 = xor_(xor_(x.a.p, x.b.p).p, prf(MB_BIT, like, seed_b).p);  // rep_to_adt over Z2: x0 ^ x1, masked
    if (me == 1) c1 = xor_(x.b.p, b1r.p);
    c1 = move_like(1, 0, c1.p, MB_BIT, like);
    T br1r = move_like(2, 1, br1.p, ring, like);
    T yp(c, nullptr), diff(c, nullptr);
    if (me != 2) {
      T cbits(c, nullptr);
      if (me == 0) cbits = xor_(c0.p, c1.p);                               // the opened bit c = x ^ b
      T got = move_like(0, 1, cbits.p, MB_BIT, like);
      if (me == 1) cbits = std::move(got);
      // s = (b_share + [party 0] c - 2 b_share c) << bit_idx, then yp = PRF(seed), diff = s - yp
      AesKey kbr, kyp;
      mb_aes_expand(seed_br, &kbr);
      mb_aes_expand(me == 0 ? seed1 : seed2, &kyp);
      yp = like_new(ring, like);
      diff = like_new(ring, like);
      ck(k_party_dabit_finish(c, kbr, kyp, me == 0, bit_idx, me == 0 ? nullptr : br1r.p, cbits.p, yp.p, diff.p));
    }
    return adt_finish(std::move(yp), std::move(diff), ring, like, seed1, seed2);
  }

  Pair msb(const Pair &x) { return msb_diff(x, nullptr); }
  Pair less(const Pair &x, const Pair &y) { return msb_diff(x, &y); }

  // MaximumOp::kernel (softmax.rs:10-53) as a tree
  Pair maximum(std::vector<Pair> &xs, size_t lo, size_t hi) {
    if (hi - lo == 1) return Pair{xs[lo].a.clone(), xs[lo].b.clone()};
    const size_t mid = lo + (hi - lo) / 2;
    Pair m1 = maximum(xs, lo, mid), m2 = maximum(xs, mid, hi);
    const mb_dtype ring = m1.a.p->dtype;
    Pair lesser = ring_inject(ring, 0, less(m1, m2));
    // lesser * m2 + (1 - lesser) * m1: both products and their zero shares in one pass, the sum reshared once
    AesKey kk[4];
    zero_keys(kk);
    zero_keys(kk + 2);
    T z = like_new(m1.a.p->dtype, m1.a.p);
    const mb_tensor *sv[2] = {lesser.a.p, lesser.b.p}, *xv[2] = {m1.a.p, m1.b.p}, *yv[2] = {m2.a.p, m2.b.p};
    void *remote = reserve_prev(z.p->nbytes);
    ck(k_party_mux(c, kk, me, sv, xv, yv, z.p, remote));
    T zn = ring_after(z.p, remote);
    return Pair{std::move(z), std::move(zn)};
  }
};

template <typename F>
int guarded(mb_party *p, const char *name, F f) {
  if (!p) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, std::string(name) + ": NULL party");
  try {
    cudaSetDevice(p->ctx->device);
    return f();
  } catch (const PErr &e) {
    return e.code;
  } catch (const std::exception &e) {
    return mb_fail(p->ctx, MB_ERR_KERNEL, std::string(name) + ": " + e.what());
  }
}

mb_pshare *make_share(Pair &&pr) {
  mb_pshare *s = new mb_pshare();
  s->t[0] = pr.a.release();
  s->t[1] = pr.b.release();
  return s;
}
Pair borrow(mb_party *p, const mb_pshare *s) {
  return Pair{T(p->ctx, mb_tensor_retain(s->t[0])), T(p->ctx, mb_tensor_retain(s->t[1]))};
}
bool valid(const mb_pshare *s) { return s && s->t[0] && s->t[1]; }

}  // namespace

extern "C" int mb_party_create(mb_ctx *ctx, int me, mb_transport_fn fn, void *user, const char *const role_names[3], mb_party **out) {
  if (!ctx || !fn || !role_names || !out || me < 0 || me > 2) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_create: bad argument");
  mb_party *p = new mb_party();
  p->ctx = ctx;
  p->me = me;
  p->tx = fn;
  p->user = user;
  for (int i = 0; i < 3; i++) p->names[i] = role_names[i] ? role_names[i] : "";
  p->have_setup = false;
  int64_t caps = 0;
  if (fn(user, 5, "", &caps, sizeof(caps), 0) != 0) caps = 0;  // optional mode: an error means "none"
  p->stream_ordered = (caps & 1) != 0;
  p->can_reserve = (caps & 2) != 0 && !getenv("MB_PARTY_NO_REMOTE_STORES");
  mb_aes_expand(ctx->sid, &p->sync_cipher);  // public sync-key stream, the same on the three parties
  p->sync_counter = 0;
  p->sends_reported = p->bytes_sent = p->transfers = 0;
  p->transport_seconds = p->sync_seconds = 0.0;
  *out = p;
  return MB_OK;
}

extern "C" int mb_party_free(mb_party *p) {
  if (!p) return MB_OK;
  for (auto *t : p->in_flight) mb_tensor_free(p->ctx, t);
  delete p;
  return MB_OK;
}

extern "C" uint64_t mb_party_bytes_sent(const mb_party *p) { return p ? p->bytes_sent : 0; }
// which: 0 host seconds inside the transport, 1 host seconds waiting for the stream before a send, 2 tensor sends
extern "C" double mb_party_stat(const mb_party *p, int which) {
  if (!p) return 0.0;
  return which == 0 ? p->transport_seconds : which == 1 ? p->sync_seconds : (double)p->transfers;
}

// Waits for every posted send; call before the tensors involved are reused outside the party runtime.
extern "C" int mb_party_flush(mb_party *p) {
  return guarded(p, "mb_party_flush", [&]() {
    P(p).reap(true);
    return (int)MB_OK;
  });
}

extern "C" int mb_pshare_free(mb_party *p, mb_pshare *s) {
  if (!s) return MB_OK;
  for (int i = 0; i < 2; i++) mb_tensor_free(p ? p->ctx : nullptr, s->t[i]);
  delete s;
  return MB_OK;
}
extern "C" const mb_tensor *mb_pshare_tensor(const mb_pshare *s, int i) { return (s && i >= 0 && i < 2) ? s->t[i] : nullptr; }

extern "C" int mb_party_share(mb_party *p, int owner, const mb_tensor *x, mb_dtype dtype, int rank, const int64_t *shape, mb_pshare **out) {
  return guarded(p, "mb_party_share", [&]() {
    if (!out || owner < 0 || owner > 2 || (p->me == owner && !x)) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_share: bad argument");
    mb_tensor *like = nullptr;
    P rt(p);
    rt.ck(mb_tensor_new(p->ctx, dtype, rank, shape, &like));
    T lk(p->ctx, like);
    if (p->me == owner && (x->dtype != dtype || !mb_same_shape(x, like))) return mb_fail(p->ctx, MB_ERR_TYPE_MISMATCH, "mb_party_share: x does not match dtype / shape");
    *out = make_share(rt.share(owner, x, like));
    return (int)MB_OK;
  });
}

extern "C" int mb_party_reveal(mb_party *p, int receiver, const mb_pshare *x, mb_tensor **out) {
  return guarded(p, "mb_party_reveal", [&]() {
    if (!valid(x) || !out || receiver < 0 || receiver > 2) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_reveal: bad argument");
    P rt(p);
    T r = rt.reveal(receiver, borrow(p, x));
    rt.reap(true);
    *out = r.release();  // NULL on the parties that are not the receiver
    return (int)MB_OK;
  });
}

#define PARTY_BINARY(NAME, CALL)                                                                                  \
  extern "C" int NAME(mb_party *p, const mb_pshare *x, const mb_pshare *y, mb_pshare **out) {                     \
    return guarded(p, #NAME, [&]() {                                                                              \
      if (!valid(x) || !valid(y) || !out) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, #NAME ": NULL argument"); \
      P rt(p);                                                                                                    \
      Pair a = borrow(p, x), b = borrow(p, y);                                                                    \
      *out = make_share(rt.CALL(a, b));                                                                           \
      return (int)MB_OK;                                                                                          \
    });                                                                                                           \
  }
PARTY_BINARY(mb_party_add, padd)
PARTY_BINARY(mb_party_sub, psub)
PARTY_BINARY(mb_party_mul, pmul)
PARTY_BINARY(mb_party_dot, pdot)
PARTY_BINARY(mb_party_less, less)

extern "C" int mb_party_add_public(mb_party *p, const mb_pshare *x, const mb_tensor *pub, mb_pshare **out) {
  return guarded(p, "mb_party_add_public", [&]() {
    if (!valid(x) || !pub || !out) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_add_public: NULL argument");
    P rt(p);
    *out = make_share(rt.add_public(borrow(p, x), pub));
    return (int)MB_OK;
  });
}

extern "C" int mb_party_trunc_pr(mb_party *p, uint32_t amount, const mb_pshare *x, mb_pshare **out) {
  return guarded(p, "mb_party_trunc_pr", [&]() {
    if (!valid(x) || !out || !mb_is_ring(x->t[0]->dtype)) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_trunc_pr: needs ring shares");
    P rt(p);
    *out = make_share(rt.trunc_pr(amount, borrow(p, x)));
    return (int)MB_OK;
  });
}

extern "C" int mb_party_msb(mb_party *p, const mb_pshare *x, mb_pshare **out_bits) {
  return guarded(p, "mb_party_msb", [&]() {
    if (!valid(x) || !out_bits || !mb_is_ring(x->t[0]->dtype)) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_msb: needs ring shares");
    P rt(p);
    *out_bits = make_share(rt.msb(borrow(p, x)));
    return (int)MB_OK;
  });
}

extern "C" int mb_party_ring_inject(mb_party *p, mb_dtype ring, uint64_t bit_idx, const mb_pshare *x_bits, mb_pshare **out) {
  return guarded(p, "mb_party_ring_inject", [&]() {
    if (!valid(x_bits) || !out || x_bits->t[0]->dtype != MB_BIT || !mb_is_ring(ring))
      return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_ring_inject: bit shares -> ring");
    P rt(p);
    *out = make_share(rt.ring_inject(ring, (unsigned)bit_idx, borrow(p, x_bits)));
    return (int)MB_OK;
  });
}

extern "C" int mb_party_maximum(mb_party *p, const mb_pshare *const *xs, int n, mb_pshare **out) {
  return guarded(p, "mb_party_maximum", [&]() {
    if (!xs || n <= 0 || !out) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "maximum op needs a non-empty array of tensors");
    std::vector<Pair> v;
    for (int i = 0; i < n; i++) {
      if (!valid(xs[i])) return mb_fail(p->ctx, MB_ERR_INVALID_ARGUMENT, "mb_party_maximum: NULL argument");
      v.push_back(borrow(p, xs[i]));
    }
    P rt(p);
    *out = make_share(rt.maximum(v, 0, v.size()));
    return (int)MB_OK;
  });
}
