/*
 * moose_b200.h -- C ABI of libmoose_b200.so: B200 (sm_100a) kernels for Moose's
 * replicated-secret-sharing hot path.
 *
 * Drop-in boundary.  After Moose's lowering pass only `[runtime]` HostPlacement
 * kernels survive in a computation (moose/src/compilation/lowering.rs:4-6); they
 * are plain Rust fns `fn(&S, &HostPlacement, attrs.., T0[, T1]) -> Result<U>`
 * registered by modelled_kernel! (moose/src/kernels/*.rs) and implemented in
 * moose/src/host/ops.rs and moose/src/host/prim.rs.  Section "host kernels"
 * below has one entry point per such kernel on ring / bit / seed / key types;
 * each comment cites the Rust kernel it replaces.  Section "replicated
 * protocols" adds fused entry points that execute a whole replicated protocol
 * (moose/src/replicated/*.rs, moose/src/additive/*.rs) in a handful of kernel
 * launches and produce bit-identical shares to the op-by-op path.
 *
 * Conventions (mirroring moose/src/kernels/mod.rs:34-106 and error.rs:8-59):
 *  - every call returns MB_OK (0) or a negative mb_status; mb_last_error()
 *    gives the message a Rust shim maps to Error::KernelError(String);
 *  - operands are immutable, results are freshly allocated device tensors
 *    owned by the caller (free with mb_tensor_free / mb_rep_free);
 *  - no torch / C++ types cross the boundary: opaque handles, plain pointers,
 *    sizes;
 *  - all device work of a context is issued on that context's CUDA stream and
 *    is asynchronous; mb_tensor_to_host / mb_ctx_sync synchronise;
 *  - thread safe: concurrent calls on one context (Moose runs every op as its own
 *    tokio task, moose/src/execution/asynchronous.rs:453-531) serialise on its stream;
 *    a call's launches reach the stream as one unit where they share per-context
 *    scratch (the tensor-pipe GEMM), the entropy source, the one-time replicated
 *    setup and the event pools are guarded.  A deferred sharing handle
 *    (mb_rep_share_deferred) must not be consumed by two threads at once;
 *  - there is NO CPU fallback: creating a context without a CUDA device fails.
 *
 * Layouts: MB_RING64 = uint64_t, MB_RING128 = little-endian 128-bit integer
 * (identical to Rust u128 / unsigned __int128 on x86-64), row-major "standard
 * layout" (moose/src/host/mod.rs:388).  MB_BIT = packed bits, flat row-major
 * bit j in byte j/8, bit j%8 (BitVec<u8, Lsb0>, moose/src/host/bitarray.rs:10-13).
 *
 * Determinism contract: PrfKeyGen, unseeded Sample and
 * SyncKey::random() draw 16 bytes each from the context's entropy source in
 * program order.  By default that is the OS (getrandom); with
 * mb_ctx_set_entropy_key() it is AES-128-CTR(key) block i for the i-th request,
 * which makes every share reproducible (used by the parity tests).
 */
#ifndef MOOSE_B200_H
#define MOOSE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mb_ctx mb_ctx;       /* one per (device, session) */
typedef struct mb_tensor mb_tensor; /* device tensor: HostRingTensor<u64|u128> / HostBitTensor */
typedef struct mb_rep mb_rep;       /* RepTensor: three distinct shares x0,x1,x2 (party i holds x_i, x_{i+1}) */

typedef enum { MB_RING64 = 0, MB_RING128 = 1, MB_BIT = 2, MB_FLOAT64 = 3 } mb_dtype;

typedef enum {
  MB_OK = 0,
  MB_ERR_INVALID_ARGUMENT = -1, /* Error::InvalidArgument */
  MB_ERR_KERNEL = -2,           /* Error::KernelError */
  MB_ERR_UNIMPLEMENTED = -3,    /* Error::UnimplementedOperator */
  MB_ERR_TYPE_MISMATCH = -4,    /* Error::TypeMismatch */
  MB_ERR_CUDA = -5,
  MB_ERR_NO_DEVICE = -6
} mb_status;

#define MB_MAX_RANK 6

/* ---------------------------------------------------------------- context */

/* Creates a context bound to CUDA device `device`. Fails with MB_ERR_NO_DEVICE
 * when no CUDA device is present (there is no CPU path). */
int mb_ctx_create(int device, mb_ctx **out);
int mb_ctx_destroy(mb_ctx *ctx);
/* Message of the last failing call on this thread for `ctx` (ctx may be NULL
 * for mb_ctx_create failures). */
const char *mb_last_error(const mb_ctx *ctx);
/* SessionId::try_from(&str): secure id = blake3(logical)[0..16]
 * (moose/src/computation.rs:108-124). */
int mb_ctx_set_session(mb_ctx *ctx, const char *logical_id);
int mb_ctx_set_session_raw(mb_ctx *ctx, const uint8_t sid[16]);
int mb_ctx_get_session_raw(const mb_ctx *ctx, uint8_t sid[16]);
/* Deterministic entropy source (see header comment); resets the request counter
 * and forgets the cached replicated setup. */
int mb_ctx_set_entropy_key(mb_ctx *ctx, const uint8_t key[16]);
int mb_ctx_sync(mb_ctx *ctx);
/* cudaStream_t of the context, for callers that interoperate (e.g. torch). */
void *mb_ctx_stream(mb_ctx *ctx);
/* CUDA-event timer on the context stream. */
int mb_ctx_timer_start(mb_ctx *ctx);
int mb_ctx_timer_stop(mb_ctx *ctx, float *elapsed_ms);
/* Number of kernels this library launched on `ctx` since creation. */
uint64_t mb_ctx_launch_count(const mb_ctx *ctx);
/* Per-kernel device timing for roofline reports: while enabled, launches of the
 * ring-GEMM main kernel are bracketed by CUDA events on the context stream.
 * mb_ctx_profile_read synchronises, returns the summed duration and launch count
 * since the last read and resets the counters.  kernel: 0 = tensor-pipe limb GEMM. */
/* The PRF-consuming kernels run one CTA per SM that fills the SM's register file.  A context whose GPU also hosts
 * long-lived small kernels of other streams (the spin waits of the peer-memory transport with several party lanes)
 * leaves `count` SMs free so that its grids still run in one wave.  No counterpart in the reference. */
int mb_ctx_reserve_sms(mb_ctx *ctx, int count);
int mb_ctx_profile_enable(mb_ctx *ctx, int enable);
int mb_ctx_profile_read(mb_ctx *ctx, int kernel, float *total_ms, uint64_t *launches);
/* Per-kernel trace (measurement only; no counterpart in the reference, whose telemetry is
 * moose/src/execution/asynchronous.rs tracing spans): while enabled, a CUDA event is recorded behind EVERY
 * kernel this library launches on `ctx`.  mb_ctx_trace_read synchronises and writes one line per launcher,
 * "name<TAB>launches<TAB>summed_ms<TAB>hbm_bytes<TAB>aes_blocks<TAB>tensor_ops<NL>" (the last three: the
 * algorithmic work the launcher states; 0 where it states none), NUL-terminated, into buf (if cap suffices; *needed = bytes
 * required), then starts a new interval.  A launch's time is the distance to the preceding event on the
 * stream, i.e. its run time plus any memset / copy enqueued between the two launches. */
int mb_ctx_trace_enable(mb_ctx *ctx, int enable);
int mb_ctx_trace_read(mb_ctx *ctx, char *buf, size_t cap, size_t *needed);
/* Page-locked host buffers for callers that stage tensors (H2D / D2H at full
 * PCIe rate); plain malloc'ed memory is accepted everywhere too. */
int mb_host_alloc_pinned(mb_ctx *ctx, size_t nbytes, void **out);
int mb_host_free_pinned(mb_ctx *ctx, void *ptr);

/* ---------------------------------------------------------------- tensors */

int mb_tensor_from_host(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape,
                        const void *host_data, mb_tensor **out);
int mb_tensor_alloc(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out);
/* Copies the tensor to host memory (nbytes must equal mb_tensor_nbytes) and
 * synchronises the stream. */
int mb_tensor_to_host(mb_ctx *ctx, const mb_tensor *t, void *host_out, size_t nbytes);
int mb_tensor_free(mb_ctx *ctx, mb_tensor *t);
/* Asynchronous transfers for callers that overlap H2D / compute / D2H with several contexts on one device (one
 * stream each): the host buffers must be page-locked (mb_host_alloc_pinned) and stay untouched until the stream
 * has passed the copy.  mb_ctx_wait orders the contexts: work enqueued on `ctx` afterwards starts only when
 * everything enqueued on `other` so far has finished.  A tensor must be freed through a context whose stream is
 * ordered after its last use. */
int mb_tensor_from_host_async(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, const void *pinned_host,
                              mb_tensor **out);
int mb_tensor_to_host_async(mb_ctx *ctx, const mb_tensor *t, void *pinned_host_out, size_t nbytes);
int mb_ctx_wait(mb_ctx *ctx, mb_ctx *other);
/* Device-to-device copies between a tensor and a caller-owned device buffer of mb_tensor_nbytes bytes,
 * asynchronous on the context stream (for callers that move tensors with their own transport, e.g.
 * NCCL send/recv in party-per-GPU mode: the Send/Receive ops of moose/src/networking/mod.rs:16-54). */
int mb_tensor_copy_to_device(mb_ctx *ctx, const mb_tensor *t, void *dst_dev);
int mb_tensor_copy_from_device(mb_ctx *ctx, mb_tensor *t, const void *src_dev);
mb_dtype mb_tensor_dtype(const mb_tensor *t);
int mb_tensor_rank(const mb_tensor *t);
const int64_t *mb_tensor_shape(const mb_tensor *t);
int64_t mb_tensor_numel(const mb_tensor *t);
size_t mb_tensor_nbytes(const mb_tensor *t); /* bits: (numel+7)/8 */
void *mb_tensor_device_ptr(const mb_tensor *t);
/* ReshapeOp (host/ops.rs:1465-1505): new handle sharing no storage (copy). */
int mb_reshape(mb_ctx *ctx, const mb_tensor *x, int rank, const int64_t *shape, mb_tensor **out);

/* ------------------------------------------------- host kernels: primitives */

/* PrfKeyGenOp::kernel (moose/src/host/prim.rs:113-121). */
int mb_prf_key_gen(mb_ctx *ctx, uint8_t key_out[16]);
/* SyncKey::random() (moose/src/host/prim.rs:73-79). */
int mb_sync_key_random(mb_ctx *ctx, uint8_t sync_key_out[16]);
/* DeriveSeedOp::kernel (moose/src/host/prim.rs:123-149):
 * blake3_keyed(derive_key("Derive Seed", key))(sid || sync_key)[0..16]. */
int mb_derive_seed(mb_ctx *ctx, const uint8_t sync_key[16], const uint8_t key[16], uint8_t seed_out[16]);
/* SampleSeededOp::{ring64,ring128,bit}_kernel (moose/src/host/ops.rs:1959-2036).
 * max_value < 0: None (uniform); max_value == 1: ring element = get_bit();
 * any other value: MB_ERR_UNIMPLEMENTED like the reference. For MB_BIT a
 * max_value >= 0 is MB_ERR_UNIMPLEMENTED (ops.rs:2025-2029). */
int mb_sample_seeded(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape,
                     const uint8_t seed[16], int64_t max_value, mb_tensor **out);
/* The same uniform stream (max_value = None) entered at AES block `first_block`: the tensor's 16-byte vector v is
 * block first_block + v of AES-128-CTR(seed), i.e. what elements [first_block*16/w, ...) of a longer
 * SampleSeededOp tensor hold (moose/src/host/ops.rs:1959-2017; aes-prng keeps incrementing one 128-bit counter).
 * For callers that cut one logical tensor into row slices and for streams past 2^32 blocks. */
int mb_sample_seeded_at(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape,
                        const uint8_t seed[16], uint64_t first_block, mb_tensor **out);
/* SampleOp::* (moose/src/host/ops.rs:1883-1957): seed drawn from the entropy source. */
int mb_sample(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, int64_t max_value,
              mb_tensor **out);
/* FillOp::host_{ring64,ring128,bit}_kernel (moose/src/host/ops.rs:1623-1655). */
int mb_fill(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, uint64_t value_lo,
            uint64_t value_hi, mb_tensor **out);

/* ------------------------------------------------ host kernels: ring arithmetic */

/* AddOp/SubOp/MulOp::ring_kernel (moose/src/host/ops.rs:1709-1737,299-310) with
 * ndarray co-broadcasting; on MB_BIT operands add = sub = xor, mul = and
 * (moose/src/kernels/boolean.rs:18-76). */
int mb_add(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
int mb_sub(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
int mb_mul(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
/* NegOp::ring_kernel / NegOp::bit_kernel (moose/src/host/ops.rs:1739-1752,1555-1565). */
int mb_neg(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
/* ShlOp/ShrOp::ring_kernel (moose/src/host/ops.rs:1853-1881); the amount is
 * masked mod the ring width like Rust's Wrapping<T> shifts. */
int mb_shl(mb_ctx *ctx, uint64_t amount, const mb_tensor *x, mb_tensor **out);
int mb_shr(mb_ctx *ctx, uint64_t amount, const mb_tensor *x, mb_tensor **out);
/* DotOp::ring_kernel / HostRingTensor::dot (moose/src/host/ops.rs:1754-1851):
 * rank (1,1), (1,2), (2,1), (2,2) contraction mod 2^k. */
int mb_dot(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
/* No counterpart in the reference: a ring matrix product C[M,N] = A[M,K] * B[K,N] mod 2^64 / 2^128 on
 * the tensor pipe that never allocates.  It is meant for hosts that cannot let this library call
 * cudaMalloc: services whose device memory is handed out up front to another allocator (so a late
 * cudaMalloc would fail or eat someone else's budget), and request paths that cannot afford the
 * device-wide stall an allocation or free may cause.  mb_dot keeps its byte-limb planes and partial
 * sums in buffers it grows on demand; this call keeps them in a fixed 8 MiB array inside the library
 * image (present once on every GPU that loads the library, for the life of the process).
 *  - a_dev, b_dev, c_dev are DEVICE pointers supplied by the caller (row-major, same element layout as
 *    MB_RING64 / MB_RING128 tensors); nothing is allocated or freed, on the device or by any library.
 *  - Limits: M >= 128, K >= 128, N >= 64, N even, and the limb planes plus partial sums,
 *    L*Kp*(M+N) + 8*passes*M*N bytes with L = 8 or 16 limbs and Kp = K rounded up to 128, must fit in
 *    8 MiB (Ring128 256 x 512 x 256 fills it exactly; Ring64 512 x 512 x 256 needs 7 MiB).  Larger shapes are refused with
 *    MB_ERR_INVALID_ARGUMENT and a message that states both sizes; use mb_dot for them.
 *  - The first call on each device looks up the array's device address; make that first call before
 *    starting any stream capture on the context's stream.
 *  - All calls on a device share the one array and it keeps the bytes of the previous call; the call clears
 *    the part that has to start at zero itself.  Two calls that overlap in time on the same GPU (different
 *    contexts / streams) would corrupt each other: serialise them. */
int mb_dot_noalloc(mb_ctx *ctx, mb_dtype dtype, const void *a_dev, const void *b_dev, void *c_dev,
                   int64_t M, int64_t K, int64_t N);
/* AddNOp::host_kernel (moose/src/host/ops.rs:1084-1103). */
int mb_add_n(mb_ctx *ctx, const mb_tensor *const *xs, int n, mb_tensor **out);
/* SumOp::host_ring_kernel (moose/src/host/ops.rs:1067-1081); axis < 0 = all. */
int mb_sum(mb_ctx *ctx, const mb_tensor *x, int axis, mb_tensor **out);
/* BroadcastOp::host_ring_kernel (moose/src/host/ops.rs:1668-1681). */
int mb_broadcast(mb_ctx *ctx, const mb_tensor *x, int rank, const int64_t *shape, mb_tensor **out);
/* RingFixedpointEncodeOp / DecodeOp (moose/src/host/ops.rs:1362-1418,
 * moose/src/host/fixedpoint.rs:10-38): (x * 2^f) as iK as uK, truncation toward
 * zero, saturating; decode = (x as iK) as f64 / 2^f.  MB_FLOAT64 tensors hold
 * IEEE doubles. */
int mb_fixedpoint_encode(mb_ctx *ctx, mb_dtype ring, uint32_t fractional_precision,
                         const mb_tensor *x_f64, mb_tensor **out);
int mb_fixedpoint_decode(mb_ctx *ctx, uint32_t fractional_precision, const mb_tensor *x,
                         mb_tensor **out_f64);

/* Plaintext-side helpers on ring tensors (host-placed tails of a computation):
 * LessOp / GreaterOp::host_ring64/128_kernel (moose/src/host/ops.rs:2051-2087,2125-2161): bit = (x - y) as signed
 * < 0 / > 0; MuxOp::host_ring_kernel (:2227-2252): s * (x - y) + y with a bit selector; SignOp (:2482-2506): -1 / 1;
 * CastOp::ring_reduction_kernel (:2277-2288): Ring128 -> Ring64 = low 64 bits; MaximumOp::host_ring_kernel
 * (:1432-1452): element-wise maximum under the UNSIGNED order of Wrapping<T>. */
int mb_less(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out_bits);
int mb_greater(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out_bits);
int mb_mux(mb_ctx *ctx, const mb_tensor *s_bits, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
int mb_sign(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
int mb_cast_ring128_to_ring64(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
int mb_maximum(mb_ctx *ctx, const mb_tensor *const *xs, int n, mb_tensor **out);

/* ------------------------------------------------ host kernels: bit tensors */

/* XorOp/AndOp/OrOp::host_kernel (moose/src/host/ops.rs:1507-1553). */
int mb_xor(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
int mb_and(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
int mb_or(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
/* BitDecomposeOp::host_bit64/128_kernel (moose/src/host/ops.rs:896-936):
 * result shape [k, shape...], plane i = (x >> i) & 1. */
int mb_bit_decompose(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
/* BitExtractOp::kernel64/128 (moose/src/host/ops.rs:1567-1595). */
int mb_bit_extract(mb_ctx *ctx, uint64_t bit_idx, const mb_tensor *x, mb_tensor **out);
/* RingInjectOp::host_kernel (moose/src/host/ops.rs:1597-1621). */
int mb_ring_inject(mb_ctx *ctx, mb_dtype ring, uint64_t bit_idx, const mb_tensor *x_bits,
                   mb_tensor **out);
/* ShlDimOp::host_bit_kernel (moose/src/host/ops.rs:835-853). */
int mb_shl_dim(mb_ctx *ctx, uint64_t amount, uint64_t bit_length, const mb_tensor *x,
               mb_tensor **out);
/* IndexAxisOp (moose/src/host/ops.rs:740-783, bitarray.rs:85-108), axis 0 only. */
int mb_index_axis0(mb_ctx *ctx, int64_t index, const mb_tensor *x, mb_tensor **out);
/* IndexAxisOp on any axis (moose/src/host/ops.rs:740-783). Rings: the selected sub-array. Bits: like
 * BitArrayRepr::index_axis (bitarray.rs:85-108) the result is the run of CONSECUTIVE bits that starts at
 * index * stride[axis], which equals the sub-array only for axis 0 / unit leading axes. */
int mb_index_axis(mb_ctx *ctx, int axis, int64_t index, const mb_tensor *x, mb_tensor **out);
/* ExpandDimsOp (moose/src/host/ops.rs:1135-1165; RawShape::extend_singletons, host/mod.rs:83-92):
 * axes are positions in the resulting shape, any order. */
int mb_expand_dims(mb_ctx *ctx, const int64_t *axes, int naxes, const mb_tensor *x, mb_tensor **out);
/* SqueezeOp (moose/src/host/ops.rs:1167-1209; RawShape::squeeze, host/mod.rs:99-113): axis < 0 = None
 * (drop every unit axis); a named axis of extent != 1 is MB_ERR_KERNEL (the reference panics). */
int mb_squeeze(mb_ctx *ctx, int axis, const mb_tensor *x, mb_tensor **out);
/* TransposeOp = reversed axes (moose/src/host/ops.rs:1270-1299); bit tensors up to rank 3
 * (bitarray.rs:135-188). The result is materialised row-major. */
int mb_transpose(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
/* ConcatOp (moose/src/host/ops.rs:1211-1268). Bit tensors: the flat bit streams are appended in order and
 * only shape[axis] is summed, exactly like ConcatOp::bit_kernel (correct for axis 0). */
int mb_concat(mb_ctx *ctx, int axis, const mb_tensor *const *xs, int n, mb_tensor **out);
/* SliceOp::host_ring_kernel (moose/src/host/ops.rs:523-535, 686-704; SliceInfo -> ndarray Slice,
 * host/mod.rs:153-165): per axis start, optional end (has_end[k] == 0: to the end), step (1 when the
 * reference has None); negative start/end count from the back, a negative step walks the selected range
 * backwards. Bit tensors: MB_ERR_KERNEL "slicing not implemented for BitArray yet" (bitarray.rs:130-133). */
int mb_slice(mb_ctx *ctx, int n, const int64_t *start, const int64_t *end, const uint8_t *has_end,
             const int64_t *step, const mb_tensor *x, mb_tensor **out);
/* DiagOp::ring_kernel / bit_kernel (moose/src/host/ops.rs:549-582; bitarray.rs:110-128: bit diagonals of
 * length 1 and 2 only, longer ones are MB_ERR_UNIMPLEMENTED where the reference hits todo!()). */
int mb_diag(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
/* SelectOp::host_ring_kernel (moose/src/host/ops.rs:803-818, 662-684): keep the positions along `axis`
 * whose index bit is set. Reads the index bits back to the host (the output extent depends on them). */
int mb_select(mb_ctx *ctx, int axis, const mb_tensor *index_bits, const mb_tensor *x, mb_tensor **out);
/* DivOp::ring_kernel (moose/src/host/ops.rs:332-343): elementwise unsigned integer quotient. A zero
 * divisor is MB_ERR_KERNEL (the reference panics); the call synchronises the stream to find out. */
int mb_div(mb_ctx *ctx, const mb_tensor *x, const mb_tensor *y, mb_tensor **out);
/* BitDecomposeOp::host_ring64/128_kernel (moose/src/host/ops.rs:856-894): [k, shape...] ring tensor,
 * plane i = (x >> i) & 1. */
int mb_bit_decompose_ring(mb_ctx *ctx, const mb_tensor *x, mb_tensor **out);
/* RingFixedpointMeanOp::ring64/128_kernel (moose/src/host/fixedpoint.rs:41-113): sum over `axis`
 * (axis < 0 = all) times encode(1 / len, scaling_base^scaling_exp). */
int mb_fixedpoint_mean(mb_ctx *ctx, int axis, uint64_t scaling_base, uint32_t scaling_exp,
                       const mb_tensor *x, mb_tensor **out);
/* RingFixedpointArgmaxOp::host_ring64/128_kernel (moose/src/host/ops.rs:2375-2437): index (Ring64) of the
 * first maximum along `axis` under the signed interpretation. */
int mb_fixedpoint_argmax(mb_ctx *ctx, int axis, const mb_tensor *x, mb_tensor **out);

/* ----------------------------------------------------- replicated protocols */

/* Party indices 0,1,2 = the replicated placement's owners; 3 = a host outside it. */

/* Takes ownership of nothing: the three share tensors are copied by reference
 * count.  Shares must have identical dtype and shape. */
int mb_rep_from_shares(mb_ctx *ctx, const mb_tensor *x0, const mb_tensor *x1, const mb_tensor *x2,
                       mb_rep **out);
/* Borrowed handle to share i (valid until mb_rep_free). */
const mb_tensor *mb_rep_share_tensor(const mb_rep *x, int i);
int mb_rep_free(mb_ctx *ctx, mb_rep *x);

/* RepSetup generation (moose/src/replicated/setup.rs:39-58 via
 * SyncSession::setup, execution/synchronous.rs:297-307): lazily draws the three
 * PRF keys; exposed so callers can force the entropy order. */
int mb_rep_setup(mb_ctx *ctx);
/* ShareOp::ring_kernel (moose/src/replicated/convert.rs:49-160). */
int mb_rep_share(mb_ctx *ctx, const mb_tensor *x, int owner, mb_rep **out);
/* ShareOp::ring_kernel by a placement member (moose/src/replicated/convert.rs:72-129) with the expansion of
 * the owner's PRF share postponed into the consumer: the sync key is drawn at this call (the reference's entropy
 * order); mb_rep_dot_concat / mb_rep_fixed_dot_concat with a skinny right operand compute x_owner = PRF(seed),
 * x_{owner+1} = x - x_owner, x_{owner+2} = 0 inside the dot kernel (same values, never written to memory);
 * mb_rep_materialize writes the three share tensors like mb_rep_share.  Other entry points reject a handle that is
 * still deferred (MB_ERR_INVALID_ARGUMENT). */
int mb_rep_share_deferred(mb_ctx *ctx, const mb_tensor *x, int owner, mb_rep **out);
int mb_rep_materialize(mb_ctx *ctx, mb_rep *x);
/* RevealOp::host_ring_kernel (moose/src/replicated/convert.rs:280-313). */
int mb_rep_reveal(mb_ctx *ctx, const mb_rep *x, int receiver, mb_tensor **out);
/* AddOp/SubOp::rep_rep_kernel (moose/src/replicated/arith.rs:7-37,197-228);
 * XorOp::rep_kernel on bit shares (:510-523). */
int mb_rep_add(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out);
int mb_rep_sub(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out);
/* NegOp::rep_rep_kernel / rep_bit_kernel (moose/src/replicated/arith.rs:525-583). */
int mb_rep_neg(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
/* ShlOp::rep_kernel (moose/src/replicated/arith.rs:585-609). */
int mb_rep_shl(mb_ctx *ctx, uint64_t amount, const mb_rep *x, mb_rep **out);
/* AddOp::rep_mir_kernel / SubOp::rep_mir_kernel / SubOp::mir_rep_kernel /
 * MulOp::rep_mir_kernel (moose/src/replicated/arith.rs:81-121,275-315,230-273,
 * 402-434) with the mirrored value given once as a public tensor. */
int mb_rep_add_public(mb_ctx *ctx, const mb_rep *x, const mb_tensor *pub, mb_rep **out);
int mb_rep_sub_public(mb_ctx *ctx, const mb_rep *x, const mb_tensor *pub, mb_rep **out);
int mb_rep_public_sub(mb_ctx *ctx, const mb_tensor *pub, const mb_rep *x, mb_rep **out);
int mb_rep_mul_public(mb_ctx *ctx, const mb_rep *x, const mb_tensor *pub, mb_rep **out);
/* MulOp::rep_rep_kernel (moose/src/replicated/arith.rs:317-367) incl. the
 * zero-share (replicated/zero_share.rs:82-156); AndOp::rep_kernel on bits. */
int mb_rep_mul(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out);
/* DotOp::rep_rep_kernel (moose/src/replicated/arith.rs:436-493). */
int mb_rep_dot(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out);
/* dot(concatenate(xs, axis = 1), y) -- ConcatOp::rep_kernel followed by DotOp::rep_rep_kernel (moose/src/replicated/
 * ops.rs:48-89, arith.rs:436-493), the bias trick of pymoose's linear predictors (linear_predictor.py:33-58) -- without
 * writing the concatenated operand when y is skinny (N <= 32): identical values and shares. */
int mb_rep_dot_concat(mb_ctx *ctx, const mb_rep *const *xs, int n, const mb_rep *y, mb_rep **out);
/* TruncPrOp::rep_kernel (moose/src/replicated/fixedpoint.rs:80-103) =
 * rep_to_adt (additive/convert.rs:12-79) + TruncPrProvider::trunc_pr
 * (additive/trunc.rs:22-65,114-169) + AdtToRepOp (replicated/convert.rs:392-497). */
int mb_rep_trunc_pr(mb_ctx *ctx, uint32_t amount, const mb_rep *x, mb_rep **out);
/* Fixed-point dot / mul as the logical layer issues them: DotOp / MulOp::rep_rep_kernel followed by
 * TruncPrOp::rep_kernel (moose/src/fixedpoint/ops.rs:761-928, 548-672; moose/src/logical/ops.rs:776-821, 553-600).
 * Same entropy requests and the same result shares as mb_rep_dot / mb_rep_mul + mb_rep_trunc_pr; the product's
 * zero share cancels in the value TruncPr opens and is not expanded.  mb_rep_fixed_dot_concat takes the left operand
 * as the concatenation along axis 1 of `n` tensors (mb_rep_dot_concat). */
int mb_rep_fixed_dot(mb_ctx *ctx, uint32_t amount, const mb_rep *x, const mb_rep *y, mb_rep **out);
int mb_rep_fixed_dot_concat(mb_ctx *ctx, uint32_t amount, const mb_rep *const *xs, int n, const mb_rep *y, mb_rep **out);
int mb_rep_fixed_mul(mb_ctx *ctx, uint32_t amount, const mb_rep *x, const mb_rep *y, mb_rep **out);
/* BitDecomposeOp::rep_ring_kernel (moose/src/replicated/bits.rs:6-20,84-134;
 * misc.rs:176-242): replicated bit shares of shape [k, shape...]. */
int mb_rep_bit_decompose(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
/* MsbOp::rep_bit_kernel (moose/src/replicated/arith.rs:611-626). */
int mb_rep_msb(mb_ctx *ctx, const mb_rep *x, mb_rep **out_bits);
/* RingInjectOp::rep_kernel (moose/src/replicated/convert.rs:316-390) with
 * gen_dabit (additive/dabit.rs:38-69). */
int mb_rep_ring_inject(mb_ctx *ctx, mb_dtype ring, uint64_t bit_idx, const mb_rep *x_bits,
                       mb_rep **out);
/* MuxOp::rep_kernel (moose/src/replicated/control_flow.rs:9-25): s*(x-y)+y with a
 * ring selector; a replicated BIT selector takes rep_bit_selector_kernel
 * (control_flow.rs:27-45): ring_inject(s, 0) first. */
int mb_rep_mux(mb_ctx *ctx, const mb_rep *s, const mb_rep *x, const mb_rep *y, mb_rep **out);
/* BitComposeOp::rep_kernel (moose/src/replicated/bits.rs:23-46): replicated bit
 * planes [k, shape...] -> ring value, sum_i ring_inject(i, plane_i). */
int mb_rep_bit_compose(mb_ctx *ctx, mb_dtype ring, const mb_rep *x_bits, mb_rep **out);
/* LessOp::rep_kernel (moose/src/replicated/compare.rs:91-104). */
int mb_rep_less(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out_bits);
/* MaximumOp::kernel (moose/src/replicated/softmax.rs:10-53). */
int mb_rep_maximum(mb_ctx *ctx, const mb_rep *const *xs, int n, mb_rep **out);
/* ReluOp::rep_ring_kernel (moose/src/fixedpoint/ops.rs:314-330). */
int mb_rep_relu(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
/* FillOp::rep_ring64/128_kernel (moose/src/replicated/ops.rs:91-160): public constant as a replicated tensor,
 * shares (v, 0, 0). */
int mb_rep_fill(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, uint64_t value_lo, uint64_t value_hi,
                mb_rep **out);
/* GreaterOp::rep_kernel (moose/src/replicated/compare.rs:136-149): msb(y - x). */
int mb_rep_greater(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out_bits);
/* Fixed-point math on replicated ring shares with RepFixedTensor's (integral, fractional) precision passed
 * explicitly.  DivOp::rep_rep_kernel (moose/src/replicated/division.rs:5-86, Goldschmidt; requires
 * 2*(i+f) <= ring bits like the reference), Pow2Op / ExpOp (moose/src/replicated/exp.rs:11-111,219-244,
 * polynomial P_1045 via fixedpoint/mod.rs:95-151), SigmoidOp (moose/src/replicated/arith.rs:676-743). */
int mb_rep_div(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x,
               const mb_rep *y, mb_rep **out);
int mb_rep_pow2(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x,
                mb_rep **out);
int mb_rep_exp(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x,
               mb_rep **out);
int mb_rep_sigmoid(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x,
                   mb_rep **out);
/* IndexAxisOp::rep_kernel axis 0 / ShlDimOp::rep_bit_kernel
 * (moose/src/replicated/ops.rs:331-376,475-504). */
int mb_rep_index_axis0(mb_ctx *ctx, int64_t index, const mb_rep *x, mb_rep **out);
int mb_rep_shl_dim(mb_ctx *ctx, uint64_t amount, uint64_t bit_length, const mb_rep *x, mb_rep **out);
/* Share-wise (local) replicated shape / reduction operators: the host kernel of the same name applied to each
 * share (moose/src/replicated/ops.rs:48-89 Concat, :267-295 ExpandDims, :297-329 Reshape, :331-360 IndexAxis,
 * :378-405 Diag, :407-473 Slice, :540-576 Broadcast, :578-620 Squeeze, :622-650 Transpose;
 * moose/src/replicated/arith.rs:123-165 AddN, :167-195 Sum; moose/src/replicated/convert.rs:499-530 Cast
 * (Ring128 -> Ring64 reduction of every share); moose/src/replicated/fixedpoint.rs:49-78 mean). */
int mb_rep_index_axis(mb_ctx *ctx, int axis, int64_t index, const mb_rep *x, mb_rep **out);
int mb_rep_expand_dims(mb_ctx *ctx, const int64_t *axes, int naxes, const mb_rep *x, mb_rep **out);
int mb_rep_squeeze(mb_ctx *ctx, int axis, const mb_rep *x, mb_rep **out);
int mb_rep_transpose(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
int mb_rep_reshape(mb_ctx *ctx, const mb_rep *x, int rank, const int64_t *shape, mb_rep **out);
int mb_rep_broadcast(mb_ctx *ctx, const mb_rep *x, int rank, const int64_t *shape, mb_rep **out);
int mb_rep_slice(mb_ctx *ctx, int n, const int64_t *start, const int64_t *end, const uint8_t *has_end,
                 const int64_t *step, const mb_rep *x, mb_rep **out);
int mb_rep_diag(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
int mb_rep_concat(mb_ctx *ctx, int axis, const mb_rep *const *xs, int n, mb_rep **out);
int mb_rep_add_n(mb_ctx *ctx, const mb_rep *const *xs, int n, mb_rep **out);
int mb_rep_sum(mb_ctx *ctx, const mb_rep *x, int axis, mb_rep **out);
int mb_rep_cast_ring128_to_ring64(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
int mb_rep_fixedpoint_mean(mb_ctx *ctx, int axis, uint64_t scaling_base, uint32_t scaling_exp,
                           const mb_rep *x, mb_rep **out);
/* AbsOp::rep_ring_kernel (moose/src/replicated/arith.rs:655-674): (1 - 2 msb(x)) * x. */
int mb_rep_abs(mb_ctx *ctx, const mb_rep *x, mb_rep **out);
/* SoftmaxOp::rep_fixed_kernel (moose/src/replicated/softmax.rs:55-127): max over the first `upmost_index`
 * entries of `axis`, exp of the difference, entries below -ln 2^(i-1) forced to 0, division by the sum. */
int mb_rep_softmax(mb_ctx *ctx, int axis, int upmost_index, uint32_t integral_precision,
                   uint32_t fractional_precision, const mb_rep *x, mb_rep **out);
/* EqualOp::rep_kernel (moose/src/replicated/compare.rs:24-40): shared bit [x == y] = AND-tree over the negated
 * bits of x - y (EqualZeroOp::bitdec_bit_kernel, compare.rs:53-72). */
int mb_rep_equal(mb_ctx *ctx, const mb_rep *x, const mb_rep *y, mb_rep **out_bits);
/* Log2Op / LogOp::rep_rep_kernel (moose/src/replicated/log.rs:13-111; Int2FL :166-219, P_2524 / Q_2524 rational
 * approximation) and SqrtOp::rep_rep_kernel (moose/src/replicated/sqrt.rs:7-34: 2^(0.5 log2 x)); x > 0. */
int mb_rep_log2(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x, mb_rep **out);
int mb_rep_log(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x, mb_rep **out);
int mb_rep_sqrt(mb_ctx *ctx, uint32_t integral_precision, uint32_t fractional_precision, const mb_rep *x, mb_rep **out);
/* RingFixedpointArgmaxOp::rep_ring_kernel + TreeReduceArgmax (moose/src/replicated/argmax.rs:10-92): shared
 * Ring64 index of the maximum over the first `upmost_index` (<= 256) entries of `axis`. */
int mb_rep_argmax(mb_ctx *ctx, int axis, int upmost_index, const mb_rep *x, mb_rep **out);

/* ------------------------------------------------------------------ storage */

/* .npy / .csv files straight into device tensors (LoadOp / SaveOp of moose/src/host/ops.rs:193-225 over
 * moose/src/storage/filesystem/numpy.rs:8-140 and csv.rs:10-140).  Files are streamed through page-locked
 * staging buffers; little-endian C-order <f8 <f4 <i8 <i4 <u8 <u4 are read (the types read_numpy accepts) and
 * widened to float64 on the device.  mb_load_npy_fixed additionally applies RingFixedpointEncode (2^frac) chunk
 * by chunk, so an 8 GB input never exists as a second host copy or as a full float64 image in HBM. */
int mb_load_npy(mb_ctx *ctx, const char *path, mb_tensor **out);
int mb_load_npy_fixed(mb_ctx *ctx, const char *path, mb_dtype ring, uint32_t frac, mb_tensor **out);
/* float64 tensors are written as <f8, ring64 tensors as <u8 (HostUint64Tensor in the reference). */
int mb_save_npy(mb_ctx *ctx, const mb_tensor *t, const char *path);
/* read_csv: header row, records parsed to f64, `columns` selects by header name (ncolumns = 0: all). */
int mb_load_csv(mb_ctx *ctx, const char *path, const char *const *columns, int ncolumns, mb_tensor **out);
/* write_csv: header col_0 .. col_{n-1}, one record per row. */
int mb_save_csv(mb_ctx *ctx, const mb_tensor *t, const char *path);

/* ------------------------------------------------------------ graph replay */

/* Reader for Moose's textual computation format, host dialect (grammar of moose/src/textual/parsing.rs), and a
 * replay executor that walks a lowered computation in topological order and dispatches every operation to the
 * entry point above that replaces its Rust kernel -- the role SyncSession / AsyncSession play in the reference
 * (moose/src/execution/synchronous.rs:347-361, asynchronous.rs:453-531; dispatch table moose/src/kernels/mod.rs:34-106).
 * Only @Host operations are executable (what remains after moose/src/compilation/lowering.rs). */
typedef struct mb_graph mb_graph;
typedef struct mb_exec mb_exec;
int mb_graph_parse(mb_ctx *ctx, const char *text, mb_graph **out);
int mb_graph_free(mb_graph *g);
int64_t mb_graph_num_ops(const mb_graph *g);
/* Name of the i-th operation in execution order (every Receive directly after its Send); NULL past the end. */
const char *mb_graph_op_name(const mb_graph *g, int64_t i);
/* role == NULL: all parties in this process, Send/Receive are a mailbox hand-off (networking/local.rs:10-53).
 * role != NULL: run only the operations placed on `role`; Send/Receive go through the transport. */
int mb_exec_create(mb_ctx *ctx, const mb_graph *g, const char *role, mb_exec **out);
int mb_exec_free(mb_exec *e);
/* Transport for Send / Receive (moose/src/networking/mod.rs:16-54) in party-per-process mode.  `mode`:
 *   1 send / 0 receive `nbytes` at `buf` to / from `peer_role`.  The executor calls it once for a 10-word int64
 *     header in host memory (kind, dtype, rank, shape[6], payload bytes) and once more for the payload; tensor
 *     payloads are DEVICE memory (on_device = 1), keys / seeds / strings host memory.  A receive returns when the
 *     data is there; a send of device memory may return while the transfer is in flight -- the executor keeps the
 *     tensor referenced until the transport reports it complete:
 *   2 poll: write to *(int64_t *)buf the number of device sends completed so far (in posting order);
 *   3 flush: wait for all posted sends, then report like 2;
 *   4 grouped exchange (party runtime only): buf points to an mb_sendrecv; send and receive the two device
 *     buffers concurrently and return when both are complete;
 *   5 capabilities (optional): write a bit mask to *(int64_t *)buf.  Bit 0: device payloads are
 *     ordered on the context's stream -- the caller need not synchronise the stream before a send, and a received
 *     buffer is valid in stream order rather than at return.  Bit 1: modes 6 / 7 exist.  A transport that does not
 *     know the mode leaves 0;
 *   6 reserve (only with capability bit 1): buf points to an mb_reserve; the transport sets `ptr` to device memory
 *     of the RECEIVER (peer memory over NVLink) where the next payload of `nbytes` for `peer` goes, valid for work
 *     enqueued on the context's stream after the call, or to NULL (the caller then sends the ordinary way);
 *   7 publish: the payload reserved last for `peer` has been written by work enqueued on the context's stream; the
 *     receiver's matching receive (mode 0 / 4 with the same size) may complete.  This is how a kernel that computes
 *     a share also delivers it: its stores go to the local tensor and to the reserved peer memory.
 * Return 0 on success.  The caller binds it to NCCL send/recv between the parties' GPUs. */
typedef struct {
  const char *send_peer;
  void *send_buf;
  size_t send_bytes;
  const char *recv_peer;
  void *recv_buf;
  size_t recv_bytes;
} mb_sendrecv;
typedef struct {
  const char *peer;
  size_t nbytes;
  void *ptr;
} mb_reserve;
typedef int (*mb_transport_fn)(void *user, int mode, const char *peer_role, void *buf, size_t nbytes, int on_device);
int mb_exec_set_transport(mb_exec *e, mb_transport_fn fn, void *user);
/* Binds a tensor to an Input operation (by arg_name) or replaces the result of the named operation (e.g. the
 * Constant inputs of a compiled graph by larger tensors: lowered graphs are shape-agnostic). */
int mb_exec_bind_tensor(mb_exec *e, const char *name, const mb_tensor *t);
/* keep = 1 (default): every operation's result stays readable through mb_exec_value after the run.
 * keep = 0: a value is released once its last consumer in this process has run (outputs and saved values stay). */
int mb_exec_set_keep_values(mb_exec *e, int keep);
int mb_exec_run(mb_exec *e);
uint64_t mb_exec_executed(const mb_exec *e);
uint64_t mb_exec_bytes_sent(const mb_exec *e);
/* Host wall-clock seconds of the last runs spent inside the transport (which = 0) and waiting for the
 * context's stream before handing a buffer to it (which = 1). */
double mb_exec_seconds(const mb_exec *e, int which);
/* Result of the named operation after a run.  kind: 0 tensor (a new reference in *tensor, free it with
 * mb_tensor_free), 1 key / seed bytes, 2 shape, 3 unit, 4 string, -1 not executed. */
int mb_exec_value(mb_exec *e, const char *op_name, int *kind, mb_tensor **tensor, uint8_t *bytes, size_t bytes_cap,
                  size_t *nbytes, int64_t *shape, int *rank);
/* Tensor recorded by an Output operation under its tag (saved = 0) or by a Save operation under its key
 * (saved = 1); a new reference. */
int mb_exec_output(mb_exec *e, int saved, const char *key, mb_tensor **tensor);
const char *mb_exec_output_key(const mb_exec *e, int saved, int64_t i);

/* Native NCCL binding of the transport: ncclSend / ncclRecv (grouped for mode 4) on a dedicated stream; libnccl is
 * opened at run time.  One rank creates the unique id, the host distributes it (torch.distributed, a file, MPI);
 * role_names[r] is the peer name of rank r.  Pass mb_nccl_transport_fn() and the handle as (fn, user) to
 * mb_exec_set_transport / mb_party_create.  No counterpart in the reference (its transport is gRPC / TCP). */
typedef struct mb_nccl_transport mb_nccl_transport;
int mb_nccl_unique_id(uint8_t out[128]);
int mb_nccl_transport_create(mb_ctx *ctx, const uint8_t unique_id[128], int rank, int nranks, const char *const *role_names,
                             mb_nccl_transport **out);
mb_transport_fn mb_nccl_transport_fn(void);
int mb_nccl_transport_free(mb_nccl_transport *t);

/* Peer-memory binding of the transport for parties on the GPUs of one NVLink / NVSwitch node.  Every party owns an
 * inbox in its HBM (per peer: a byte ring of `ring_bytes`, 0 = 1 GiB, and a ring of 4 KiB message slots); the peers
 * map it (CUDA IPC between processes, directly between threads of one process) and a transfer is a copy into the
 * receiver's ring over NVLink plus a release store of a cumulative byte counter.  Device payloads are enqueued on the
 * context's stream on both sides (wait for space, copy, signal / wait for the counter, copy out, acknowledge): the
 * callback returns at once, a received buffer is valid in stream order, and the caller need not synchronise before a
 * send (mode 5 of the callback reports this as capability bit 0).  Start-up: create on every rank, exchange the
 * 128-byte handles through the host (torch.distributed, a file, MPI), connect with all of them in rank order.  All
 * ranks must have stopped communicating (a host barrier) before any of them frees its transport.
 * No counterpart in the reference (its transport is gRPC / TCP). */
typedef struct mb_peer_transport mb_peer_transport;
int mb_peer_transport_create(mb_ctx *ctx, int rank, int nranks, const char *const *role_names, size_t ring_bytes,
                             mb_peer_transport **out);
int mb_peer_transport_handle(mb_peer_transport *t, uint8_t out[128]);
int mb_peer_transport_connect(mb_peer_transport *t, const uint8_t *handles);
mb_transport_fn mb_peer_transport_fn(void);
int mb_peer_transport_free(mb_peer_transport *t);

/* ------------------------------------------------------- party per process */

/* ONE party of the replicated protocols per process / GPU (BASELINE config 5).  Party `me` holds the shares
 * (x_me, x_{me+1}) of every replicated tensor (mb_pshare) and computes what player `me` computes in the reference
 * (moose/src/replicated/{setup,zero_share,convert,arith,fixedpoint,bits,misc,compare,softmax}.rs,
 * moose/src/additive/{convert,trunc,dabit}.rs); what the reference moves between hosts -- reshares, the dealer's
 * messages, openings -- goes through the transport (same callback contract as mb_exec_set_transport; role_names[i]
 * is the peer name handed to it for party i).  The three parties must issue the same sequence of calls.  The
 * heavy steps are the fused kernels of the single-GPU path restricted to this party's component: msb-pruned
 * Kogge-Stone levels with one exchange per level, the two-term tensor-pipe GEMM. */
typedef struct mb_party mb_party;
typedef struct mb_pshare mb_pshare;
int mb_party_create(mb_ctx *ctx, int me, mb_transport_fn fn, void *user, const char *const role_names[3], mb_party **out);
int mb_party_free(mb_party *p);
int mb_party_flush(mb_party *p); /* waits for every posted send */
uint64_t mb_party_bytes_sent(const mb_party *p);
/* which: 0 host seconds inside the transport, 1 host seconds waiting for the stream before a send, 2 tensor sends */
double mb_party_stat(const mb_party *p, int which);
int mb_pshare_free(mb_party *p, mb_pshare *s);
const mb_tensor *mb_pshare_tensor(const mb_pshare *s, int i); /* i = 0: share me, 1: share me+1 (borrowed) */
/* ShareOp (convert.rs:49-160): x is the plaintext on the owner (NULL elsewhere); dtype / shape on every party. */
int mb_party_share(mb_party *p, int owner, const mb_tensor *x, mb_dtype dtype, int rank, const int64_t *shape, mb_pshare **out);
/* RevealOp (convert.rs:280-313): *out is the value on the receiver, NULL on the other parties. */
int mb_party_reveal(mb_party *p, int receiver, const mb_pshare *x, mb_tensor **out);
int mb_party_add(mb_party *p, const mb_pshare *x, const mb_pshare *y, mb_pshare **out);
int mb_party_sub(mb_party *p, const mb_pshare *x, const mb_pshare *y, mb_pshare **out);
int mb_party_add_public(mb_party *p, const mb_pshare *x, const mb_tensor *pub, mb_pshare **out);
int mb_party_mul(mb_party *p, const mb_pshare *x, const mb_pshare *y, mb_pshare **out);
int mb_party_dot(mb_party *p, const mb_pshare *x, const mb_pshare *y, mb_pshare **out);
int mb_party_trunc_pr(mb_party *p, uint32_t amount, const mb_pshare *x, mb_pshare **out);
int mb_party_msb(mb_party *p, const mb_pshare *x, mb_pshare **out_bits);
int mb_party_ring_inject(mb_party *p, mb_dtype ring, uint64_t bit_idx, const mb_pshare *x_bits, mb_pshare **out);
int mb_party_less(mb_party *p, const mb_pshare *x, const mb_pshare *y, mb_pshare **out_bits);
int mb_party_maximum(mb_party *p, const mb_pshare *const *xs, int n, mb_pshare **out);

/* Selects the ring-GEMM implementation used by mb_dot / mb_rep_dot:
 * 0 = automatic (int8 tensor pipe for large shapes, 64-bit IMAD CUDA-core
 * kernel for skinny / small ones), 1 = force IMAD, 2 = force tensor pipe
 * (fails for shapes it cannot tile), 3 = like 2 but single-CTA tiles only (no
 * cta_group::2 pairs).  For ncu comparisons and tests. */
int mb_ctx_set_gemm_mode(mb_ctx *ctx, int mode);
/* 1 (default): mb_rep_bit_decompose / mb_rep_msb / mb_rep_ring_inject use the fused per-level adder and
 * dabit kernels; 0: the op-by-op composition of the same protocols (identical shares; for tests and ncu). */
int mb_ctx_set_fused_bits(mb_ctx *ctx, int enable);

#ifdef __cplusplus
}
#endif
#endif /* MOOSE_B200_H */
