// Internal declarations shared by the translation units of libmoose_b200.so.
#pragma once

#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/moose_b200.h"

typedef unsigned __int128 u128;
constexpr int64_t MB_MAX_NUMEL = (int64_t)1 << 40;   // 16 TB of Ring128: beyond any device; rejects wrapped / hostile shapes

struct mb_tensor {
  mb_dtype dtype;
  int rank;
  int64_t shape[MB_MAX_RANK];
  int64_t numel;
  size_t nbytes;        // logical bytes (bits: (numel+7)/8)
  size_t alloc_bytes;   // allocation size (multiple of 16)
  void *dptr;
  std::atomic<int> refs;
  int device;
  mb_tensor *parent;    // non-null: dptr points into parent's allocation (a retained, immutable view)
};

struct AesKey;
struct mb_rep_deferred;   // a sharing whose PRF expansion is postponed into its consumer (rep_api.cpp)
struct mb_rep {
  mb_tensor *s[3];
  mb_rep_deferred *deferred;  // non-null: s[] are NULL until mb_rep_materialize
};

// AES-128 expanded key as 44 little-endian column words, plus the round-1 constants of the counter-mode
// stream (the part of round 1 that does not depend on a counter < 2^32, see aes.cuh).
struct AesKey {
  uint32_t rk[44];
  uint32_t r1c[4];
};

struct mb_ctx {
  int device;
  cudaStream_t stream;
  cudaStream_t aux_stream;    // allocations that must not be ordered after `stream` (receive buffers)
  cudaEvent_t ev0, ev1;
  int sm_count;
  std::mutex mu;             // guards entropy, setup, last-error bookkeeping
  std::mutex setup_mu;       // serialises the one-time generation of the replicated setup
  std::mutex gemm_mu;        // limb GEMM: scratch lookup + split + GEMM + recombine are enqueued as one unit
  int reserve_sms;           // SMs the PRF kernels (one CTA per SM, whole register file) leave free for resident waits
  bool gemm_registered;      // counted in the per-device tally of contexts that run limb GEMMs
  uint8_t sid[16];
  bool deterministic;
  uint8_t entropy_key[16];
  uint64_t entropy_counter;
  bool have_setup;
  uint8_t setup_keys[3][16]; // k0,k1,k2 (k10=k11 etc., replicated/setup.rs:39-58)
  int gemm_mode;
  bool gemm_round_barrier;    // 2-CTA limb GEMM: keep clusters in step per scheduling round (L2 sharing)
  bool fused_bits;            // fused adder / ring_inject kernels (false = op-by-op composition, for A/B tests)
  int gemm_pair_mode;         // 0 = CTA pairs when possible, 1 = single-CTA tiles only (A/B comparison)
  std::atomic<uint64_t> launches;
  std::mutex prof_mu;         // guards the event pool and the profile / trace lists
  bool profile;
  bool trace;                 // per-kernel trace: one event behind every launch (mb_ctx_trace_*)
  struct TraceRec { const char *name; cudaEvent_t ev; double bytes, aes_blocks, tensor_ops; };
  std::vector<TraceRec> trace_events;  // event after each launch; name NULL = start marker
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;  // pending (start, stop) pairs
  std::vector<cudaEvent_t> prof_pool;
  // scratch for the limb GEMM (grown on demand, stream ordered)
  void *scratch[4];
  size_t scratch_bytes[4];
};

// ---- error handling ---------------------------------------------------------
void mb_set_error(const mb_ctx *ctx, const std::string &msg);
int mb_fail(const mb_ctx *ctx, int code, const std::string &msg);

#define MB_CUDA(ctx, expr)                                                           \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess)                                                           \
      return mb_fail((ctx), MB_ERR_CUDA,                                             \
                     std::string(#expr) + ": " + cudaGetErrorString(_e));            \
  } while (0)

#define MB_CHECK(expr)            \
  do {                            \
    int _rc = (expr);             \
    if (_rc != MB_OK) return _rc; \
  } while (0)

// After every kernel launch: count it, and -- when the per-kernel trace is on (mb_ctx_trace_enable) -- record a CUDA
// event behind it on the context's stream, tagged with the launcher's name.
// MB_LAUNCH_CHECK_W also states the launch's ALGORITHMIC work: HBM bytes (operands read + results
// written, each once), AES blocks of PRF stream, int8 tensor operations -- what the roofline fractions are computed from.
#define MB_LAUNCH_CHECK_W(ctx, name, bytes, aes_blocks, tensor_ops)                                                    \
  do {                                                                                                                 \
    (ctx)->launches.fetch_add(1);                                                                                      \
    MB_CUDA((ctx), cudaGetLastError());                                                                                \
    if ((ctx)->trace) MB_CHECK(mb_trace_mark((ctx), (name), (double)(bytes), (double)(aes_blocks), (double)(tensor_ops))); \
  } while (0)
#define MB_LAUNCH_CHECK_AS(ctx, name) MB_LAUNCH_CHECK_W(ctx, name, 0, 0, 0)
#define MB_LAUNCH_CHECK(ctx) MB_LAUNCH_CHECK_AS(ctx, __func__)

// ---- tensors (tensor.cpp) ---------------------------------------------------
static inline int mb_elem_bytes(mb_dtype d) { return d == MB_RING128 ? 16 : 8; }
static inline bool mb_is_ring(mb_dtype d) { return d == MB_RING64 || d == MB_RING128; }
int mb_tensor_new(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out);
int mb_tensor_new_detached(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out);
mb_tensor *mb_tensor_retain(mb_tensor *t);
// Zero-copy view of `nbytes`-aligned contents of `parent` starting at byte_offset (must be a multiple of 16).
int mb_tensor_view(mb_ctx *ctx, const mb_tensor *parent, size_t byte_offset, int rank, const int64_t *shape, mb_tensor **out);
bool mb_same_shape(const mb_tensor *a, const mb_tensor *b);
// CTAs of a PRF kernel (one per SM): a 1024-thread CTA at 64 registers takes an SM's whole register file, so on a GPU
// where other streams keep small kernels resident (the spin waits of the peer-memory transport, several party lanes) a
// grid of sm_count CTAs runs in two waves; such contexts reserve a few SMs (mb_ctx_reserve_sms).
static inline int mb_aes_sms(const mb_ctx *ctx) { return ctx->sm_count - ctx->reserve_sms > 0 ? ctx->sm_count - ctx->reserve_sms : 1; }
// Launch shape of a PRF kernel over `items` grid-stride items: one CTA per SM, and a block size (<= tpb_max, a multiple
// of 32) chosen so that every thread gets the same number of items.  With full blocks a tensor of 1.65 waves
// (250 000 vectors on 148 x 1024 threads: a LogReg slice at 8 GPUs) leaves 35 % of the threads idle in the last wave.
struct mb_aes_cfg_t { int grid_dim, block_dim; };
static inline mb_aes_cfg_t mb_aes_cfg(const mb_ctx *ctx, int64_t items, int tpb_max) {
  const int64_t sms = mb_aes_sms(ctx);
  if (items <= 0) return mb_aes_cfg_t{1, 32};
  static const bool full_blocks = getenv("MOOSE_B200_FULL_BLOCKS") != nullptr;   // A/B: the fixed block size of round 1
  const int64_t per_wave = sms * tpb_max;
  const int64_t iters = (items + per_wave - 1) / per_wave;
  // Only where it matters: beyond a few waves the idle tail is a few per cent, and odd block sizes were measured to
  // cost more than that on long loops (inject_sum over 2 M vectors: 22.1 ms at 992 threads against 15.8 ms at 1024).
  if (full_blocks || iters > 3) {
    const int64_t g = (items + tpb_max - 1) / tpb_max;
    return mb_aes_cfg_t{(int)(g < sms ? g : sms), tpb_max};
  }
  const int64_t threads = (items + iters - 1) / iters;
  if (threads < sms * 128) {   // tiny: few CTAs of 128 threads
    const int64_t g = (threads + 127) / 128;
    return mb_aes_cfg_t{(int)(g < 1 ? 1 : g), 128};
  }
  int64_t tpb = ((threads + sms - 1) / sms + 31) / 32 * 32;
  if (tpb > tpb_max) tpb = tpb_max;
  return mb_aes_cfg_t{(int)sms, (int)tpb};
}
int mb_scratch(mb_ctx *ctx, int slot, size_t bytes, void **out);
// contexts on `device` that have launched the limb GEMM so far (the round barrier is used only by a sole one)
int mb_gemm_ctx_register(mb_ctx *ctx);
int mb_prof_begin(mb_ctx *ctx, cudaEvent_t *start);  // no-ops unless profiling is enabled
int mb_prof_end(mb_ctx *ctx, cudaEvent_t start);
int mb_trace_mark(mb_ctx *ctx, const char *name, double bytes = 0, double aes_blocks = 0, double tensor_ops = 0);

// ---- host-side crypto (blake3.cpp, aes_host.cpp) ----------------------------
void mb_blake3_session_id(const char *s, size_t len, uint8_t out[16]);
void mb_blake3_derive_seed(const uint8_t sid[16], const uint8_t sync_key[16], const uint8_t key[16],
                           uint8_t out[16]);
void mb_aes_expand(const uint8_t key[16], AesKey *out);
void mb_aes_encrypt_host(const AesKey *k, const uint8_t in[16], uint8_t out[16]);
int mb_entropy16(mb_ctx *ctx, uint8_t out[16]);  // takes ctx->mu

// ---- kernel launchers (elementwise.cu, sample.cu, bits.cu, gemm_*.cu, rep.cu)
enum EwOp { EW_ADD = 0, EW_SUB = 1, EW_MUL = 2, EW_AND = 3, EW_XOR = 4, EW_OR = 5 };

// Binary op with ndarray co-broadcasting: strides in elements, 0 for broadcast axes.
int k_ew_binary(mb_ctx *ctx, EwOp op, const mb_tensor *x, const mb_tensor *y, mb_tensor *out);
int k_ew_neg(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out);
int k_ew_binary3(mb_ctx *ctx, EwOp op, const mb_tensor *const x[3], const mb_tensor *const y[3], mb_tensor *const out[3]);
int k_ew_unary3(mb_ctx *ctx, int mode, unsigned amount, const mb_tensor *const x[3], mb_tensor *const out[3]);
int k_ew_shift(mb_ctx *ctx, bool left, unsigned amount, const mb_tensor *x, mb_tensor *out);
int k_fill(mb_ctx *ctx, mb_tensor *out, uint64_t lo, uint64_t hi);
int k_broadcast(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out);
int k_sum_axis(mb_ctx *ctx, const mb_tensor *x, int64_t outer, int64_t len, int64_t inner, mb_tensor *out);
int k_fixed_encode(mb_ctx *ctx, const mb_tensor *x, double scale, mb_tensor *out);
int k_fixed_decode(mb_ctx *ctx, const mb_tensor *x, double scale, mb_tensor *out);
int k_fixed_encode_raw(mb_ctx *ctx, const double *x_dev, mb_dtype ring, void *out_dev, int64_t n, double scale);

int k_sample(mb_ctx *ctx, const AesKey &key, mb_tensor *out, bool bits_as_ring, uint64_t first_block = 0);

int k_bit_decompose(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out);
int k_bit_extract(mb_ctx *ctx, unsigned bit_idx, const mb_tensor *x, mb_tensor *out);
int k_ring_inject(mb_ctx *ctx, unsigned bit_idx, const mb_tensor *bits, mb_tensor *out);
int k_bit_shift_copy(mb_ctx *ctx, const mb_tensor *x, int64_t src_bit, int64_t dst_bit, int64_t nbits,
                     mb_tensor *out);  // out must be zero-filled where not covered
int k_bit_not(mb_ctx *ctx, const mb_tensor *x, mb_tensor *out);

// Ring GEMM C[M,N] (+)= A[M,K] * B[K,N] mod 2^k on raw device pointers.
int k_gemm_imad(mb_ctx *ctx, mb_dtype dtype, const void *A, const void *B, void *C, int64_t M, int64_t K,
                int64_t N, bool accumulate);
// Skinny replicated dot, three parties in one pass; the left operand is the concatenation along K of `n` parts.
struct RepDotPart { const void *x[3]; int64_t K; int64_t krow0; };
struct RepDotParts { int n; RepDotPart p[4]; };
// Parts whose sharing is deferred (ShareOp by a placement member, replicated/convert.rs:72-129): x_owner = PRF(key),
// x_{owner+1} = plain - PRF(key), x_{owner+2} = 0, expanded inside the dot instead of written out and read back.
struct RepDotDeferred { int owner[4]; const void *plain[4]; AesKey key[4]; };  // owner < 0: the part is materialised
bool k_rep_dot_skinny_supported(int64_t M, int64_t N);
int k_rep_dot_skinny(mb_ctx *ctx, mb_dtype dtype, const RepDotParts &parts, const void *const ysum[3], const void *const y[3],
                     void *const z[3], int64_t M, int64_t N);
bool k_rep_dot_skinny_deferred_supported(mb_dtype dtype, const RepDotParts &parts, int64_t N);
int k_rep_dot_skinny_deferred(mb_ctx *ctx, mb_dtype dtype, const RepDotParts &parts, const RepDotDeferred &lazy, const void *const ysum[3],
                              const void *const y[3], void *const z[3], int64_t M, int64_t N);
bool k_gemm_limb_supported(mb_dtype dtype, int64_t M, int64_t K, int64_t N);
bool k_gemm_limb_preferred(int64_t M, int64_t K, int64_t N);
// C = sum_t A_t * B_t for nterms (1 or 2) operand pairs, optional fused epilogue.
struct GemmTerm { const void *A; const void *B0; const void *B1; };  // B = B0 (+ B1 if non-null)
int k_gemm_limb(mb_ctx *ctx, mb_dtype dtype, const GemmTerm *terms, int nterms, void *C, int64_t M,
                int64_t K, int64_t N);
// Batched form with de-duplicated operands: C[b] = sum_t A[a_idx[b][t]] * B[b_idx[b][t]].
struct LimbOperandA { const void *ptr; };
struct LimbOperandB { const void *b0; const void *b1; };  // B = b0 (+ b1 if non-null)
int k_gemm_limb_batched(mb_ctx *ctx, mb_dtype dtype, const LimbOperandA *As, int nA, const LimbOperandB *Bs, int nB,
                        const int (*a_idx)[2], const int (*b_idx)[2], int nbatch, int nterms, void *const *C,
                        int64_t M, int64_t K, int64_t N);

// One Kogge-Stone level on a pruned set of bit planes (bits_fused.cu)
struct LevelDesc {
  int nout;                 // number of output planes
  int shift;                // 2^i
  unsigned char plane[128]; // output plane numbers
  unsigned char gslot[128]; // input slot of plane q in g_in (255 = not stored)
  unsigned char pslot[128]; // input slot of plane q in p_in
  unsigned char pout[128];  // output slot in p_out for output index j (255 = p' not needed)
};
int k_adder_plan_msb(int k, std::vector<LevelDesc> *levels_out, std::vector<int> *np_out, int *final_gslot);
int k_party_split(mb_ctx *ctx, const mb_tensor *const terms[], int nterms, int neg_mask, mb_tensor *out);
int k_party_adder_init(mb_ctx *ctx, int me, const AesKey keys[3], const mb_tensor *in, const mb_tensor *other, mb_tensor *g,
                       mb_tensor *p_new, mb_tensor *p_send, void *g_remote);
// party-sliced fused steps (party_fused.cu)
// `remote` (may be NULL): device memory of the previous player (reserved through the transport) that receives the
// same values as z -- the reshare is done by the kernel that computes the share
int k_party_mul(mb_ctx *ctx, const AesKey kk[2], const mb_tensor *xa, const mb_tensor *xb, const mb_tensor *ya, const mb_tensor *yb,
                mb_tensor *z, void *remote);
int k_party_zero_add(mb_ctx *ctx, const AesKey kk[2], const mb_tensor *v, mb_tensor *z, void *remote);
int k_party_mux(mb_ctx *ctx, const AesKey kk[4], int me, const mb_tensor *const s[2], const mb_tensor *const x[2],
                const mb_tensor *const y[2], mb_tensor *z, void *remote);
int k_party_trunc_dealer(mb_ctx *ctx, const AesKey kk[4], unsigned amount, mb_tensor *const o[3]);
int k_party_trunc_p0(mb_ctx *ctx, const AesKey kk[4], unsigned amount, const mb_tensor *xa, const mb_tensor *xb, const mb_tensor *masked1,
                     mb_tensor *cbits, mb_tensor *yp, mb_tensor *diff);
int k_party_trunc_p1(mb_ctx *ctx, const AesKey &kyp, unsigned amount, const mb_tensor *r_top1, const mb_tensor *r_msb1,
                     const mb_tensor *cbits, mb_tensor *yp, mb_tensor *diff);
int k_party_dabit_dealer(mb_ctx *ctx, const AesKey &kbr, const mb_tensor *b, mb_tensor *br1);
int k_party_dabit_finish(mb_ctx *ctx, const AesKey &kbr, const AesKey &kyp, int first, unsigned bit_idx, const mb_tensor *br1,
                         const mb_tensor *cbits, mb_tensor *yp, mb_tensor *diff);
int k_party_adder_level(mb_ctx *ctx, const AesKey keys[4], const LevelDesc &d, const mb_tensor *const g_in[2],
                        const mb_tensor *const p_in[2], mb_tensor *g_out, mb_tensor *p_out, int64_t nv, void *remote);

// Fused arithmetic->boolean conversion (bits_fused.cu)
bool k_rep_adder_supported(const mb_tensor *x);
// out_planes (full decomposition only): planes 0 .. out_planes-1 are computed, the rest of the [k, n] result is zero (0 = all)
// minus (may be NULL): the adder decomposes x - minus, the subtraction folded into the split
int k_rep_adder(mb_ctx *ctx, const AesKey *keys, const mb_rep *x, bool msb_only, mb_tensor *out[3], int out_planes = 0, const mb_rep *minus = nullptr);
// BitDecomposeOp::rep_ring_kernel of which the caller reads only planes 0 .. planes-1 (same entropy requests, same values there)
int mb_rep_bit_decompose_low(mb_ctx *ctx, const mb_rep *x, int planes, mb_rep **out);
int k_ring_inject_fused(mb_ctx *ctx, const AesKey &k_y0, const AesKey &k_y1, unsigned bit_idx, const mb_rep *x,
                        mb_tensor *out[3]);

// Batched kernels of the fixed-point math protocols (math_fused.cu); descriptors travel as kernel parameters
constexpr int MB_PREFIX_MAX_OPS = 32;
struct PrefixLevel {             // one level of prefix_op on bit planes: W[dst] = OR(W[src], W[dst])
  int nops;
  short src[MB_PREFIX_MAX_OPS], dst[MB_PREFIX_MAX_OPS];
  AesKey keys[MB_PREFIX_MAX_OPS][3];   // zero-share seeds of the AND inside each OR
};
constexpr int MB_INJECT_MAX_PLANES = 64;
struct InjectSumDesc {           // out = base + (negate ? -1 : 1) * sum_i mult_i * (ring_inject(plane pa_i ^ plane pb_i) << shift_i)
  int nplanes, negate;
  short pa[MB_INJECT_MAX_PLANES], pb[MB_INJECT_MAX_PLANES];   // pb < 0: no second plane
  unsigned char shift[MB_INJECT_MAX_PLANES], mult[MB_INJECT_MAX_PLANES];
  AesKey keys[MB_INJECT_MAX_PLANES][2];  // adt_to_rep seeds (y0, y1) of each ring_inject
};
constexpr int MB_POW2_MAX_BITS = 32;
struct Pow2Desc {                // Pow2FromBits over planes first_plane .. first_plane + count - 1
  int count, first_plane;
  unsigned char part_shift[MB_POW2_MAX_BITS], sel_shift[MB_POW2_MAX_BITS];
  AesKey inject[MB_POW2_MAX_BITS][2];
  AesKey zero[MB_POW2_MAX_BITS][3];
};
constexpr int MB_MULTRUNC_MAX_OPS = 16;
struct MulTruncLevel {           // P[dst] = trunc_pr(P[src] * P[dst]) for a batch of ops of one prefix-mul level
  int nops;
  short src[MB_MULTRUNC_MAX_OPS], dst[MB_MULTRUNC_MAX_OPS];
  AesKey keys[MB_MULTRUNC_MAX_OPS][6];  // zero share (3), then trunc (r, y0, y1)
};
constexpr int MB_POLY_MAX_COEFFS = 104;
struct PolyCombine {             // out = trunc_pr(sum_{i < degree} P[i] * coeff[i + 1]) + coeff[0]
  int degree;
  uint64_t coeff_lo[MB_POLY_MAX_COEFFS], coeff_hi[MB_POLY_MAX_COEFFS];
  AesKey keys[3];
};
int k_mul_trunc(mb_ctx *ctx, const AesKey tk[3], unsigned amount, const mb_rep *x, const mb_rep *y, mb_tensor *const out[3]);
int k_maximum_select(mb_ctx *ctx, const AesKey keys[8], const mb_rep *lesser, const mb_rep *m1, const mb_rep *m2, mb_tensor *const out[3]);
int k_replicate(mb_ctx *ctx, const mb_rep *x, mb_tensor *const P[3], int slots);
int k_mul_trunc_level(mb_ctx *ctx, const MulTruncLevel &lv, unsigned amount, mb_tensor *const P[3], int64_t nvec);
int k_poly_combine(mb_ctx *ctx, const PolyCombine &pc, unsigned amount, mb_tensor *const P[3], mb_tensor *const out[3]);
int k_planes_gather(mb_ctx *ctx, const mb_tensor *const src[3], const short *idx, int nplanes, mb_tensor *const W[3], int64_t n);
int k_prefix_or_level(mb_ctx *ctx, const PrefixLevel &lv, mb_tensor *const W[3], int64_t n);
int k_inject_sum(mb_ctx *ctx, const InjectSumDesc &d, const mb_tensor *const W[3], int64_t n, const mb_rep *base, mb_tensor *const out[3]);
int k_pow2_bits(mb_ctx *ctx, const Pow2Desc &d, const mb_tensor *const W[3], int64_t n, const mb_rep *x_abs, mb_tensor *const out_d[3],
                mb_tensor *const out_rest[3]);
// key derivation in the reference's entropy order, shared by the per-op entry points and the batched paths (rep_api.cpp)
int mb_draw_zero_seeds(mb_ctx *ctx, AesKey seeds[3]);      // gen_zero_share: 3 x SyncKey (after the one-time setup)
int mb_draw_inject_keys(mb_ctx *ctx, AesKey keys[2]);      // ring_inject: dabit (4 requests) + adt_to_rep (3): returns (y0, y1)
int mb_draw_trunc_keys(mb_ctx *ctx, AesKey keys[3]);       // trunc_pr: mask (5 requests) + adt_to_rep (3): returns (r, y0, y1)

struct mb_rep_deferred {   // see mb_rep_share_deferred (rep_api.cpp)
  mb_tensor *src;          // the plain tensor (retained)
  AesKey key;              // PRF stream of the owner's share
  int owner;
};

// Fused replicated kernels (rep.cu)
int k_rep_share(mb_ctx *ctx, const AesKey &key, int owner, const mb_tensor *x, mb_tensor *o0, mb_tensor *o1,
                mb_tensor *o2);
int k_rep_share_outsider(mb_ctx *ctx, const AesKey &k0, const AesKey &k1, const mb_tensor *x, mb_tensor *o0,
                         mb_tensor *o1, mb_tensor *o2);
int k_reveal3(mb_ctx *ctx, const mb_tensor *a, const mb_tensor *b, const mb_tensor *c, mb_tensor *out);
// z_i = x_i*y_i + x_i*y_{i+1} + x_{i+1}*y_i + PRF(s_i) - PRF(s_{i+1})
int k_rep_mul(mb_ctx *ctx, const AesKey seeds[3], const mb_rep *x, const mb_rep *y, mb_tensor *z[3]);
// z_i = v_i + PRF(s_i) - PRF(s_{i+1})  (v_i may alias z_i)
int k_add_zero_share(mb_ctx *ctx, const AesKey seeds[3], mb_tensor *v[3], mb_tensor *z[3]);
// seeds: r, r0, rtop0, rmsb0, y0, y1
int k_rep_trunc_pr(mb_ctx *ctx, const AesKey seeds[6], unsigned amount, const mb_rep *x, mb_tensor *z[3]);
