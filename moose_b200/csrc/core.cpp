// Context, tensor handles, error reporting.
#include "common.h"

#include <cstdlib>

static thread_local std::string g_last_error;
static std::atomic<int> g_gemm_ctxs[64];  // per device: contexts that have run the limb GEMM

void mb_set_error(const mb_ctx *, const std::string &msg) { g_last_error = msg; }
int mb_fail(const mb_ctx *ctx, int code, const std::string &msg) {
  mb_set_error(ctx, msg);
  return code;
}

extern "C" const char *mb_last_error(const mb_ctx *) { return g_last_error.c_str(); }

extern "C" int mb_ctx_create(int device, mb_ctx **out) {
  if (!out) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, "mb_ctx_create: out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return mb_fail(nullptr, MB_ERR_NO_DEVICE,
                   std::string("mb_ctx_create: no CUDA device (libmoose_b200 has no CPU path): ") +
                       cudaGetErrorString(e));
  if (device < 0 || device >= n) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, "mb_ctx_create: bad device index");
  mb_ctx *ctx = new mb_ctx();
  ctx->device = device;
  ctx->deterministic = false;
  ctx->entropy_counter = 0;
  ctx->have_setup = false;
  ctx->gemm_mode = 0;
  ctx->gemm_pair_mode = 0;
  ctx->fused_bits = true;
  ctx->gemm_round_barrier = getenv("MOOSE_B200_NO_ROUND_BARRIER") == nullptr;
  ctx->launches = 0;
  ctx->profile = false;
  ctx->trace = false;
  ctx->gemm_registered = false;
  ctx->reserve_sms = 0;
  memset(ctx->scratch, 0, sizeof(ctx->scratch));
  memset(ctx->scratch_bytes, 0, sizeof(ctx->scratch_bytes));
  MB_CUDA(ctx, cudaSetDevice(device));
  MB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  MB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking));
  MB_CUDA(ctx, cudaEventCreate(&ctx->ev0));
  MB_CUDA(ctx, cudaEventCreate(&ctx->ev1));
  MB_CUDA(ctx, cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device));
  // keep freed blocks in the stream-ordered pool (tensors are allocated per op, like the reference)
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    uint64_t thresh = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh);
    // an allocation on the auxiliary stream must never wait for work of the main stream (mb_tensor_new_detached)
    int no = 0;
    cudaMemPoolSetAttribute(pool, cudaMemPoolReuseAllowInternalDependencies, &no);
  }
  // default session id: SessionId::try_from("b200") -- callers normally set their own
  mb_blake3_session_id("b200", 4, ctx->sid);
  *out = ctx;
  return MB_OK;
}

extern "C" int mb_ctx_destroy(mb_ctx *ctx) {
  if (!ctx) return MB_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->gemm_registered) g_gemm_ctxs[ctx->device & 63].fetch_sub(1);
  for (int i = 0; i < 4; i++)
    if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
  for (auto &pr : ctx->prof_events) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  for (auto ev : ctx->prof_pool) cudaEventDestroy(ev);
  for (auto &te : ctx->trace_events) cudaEventDestroy(te.ev);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaStreamDestroy(ctx->aux_stream);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
  return MB_OK;
}

extern "C" int mb_ctx_set_session(mb_ctx *ctx, const char *logical_id) {
  if (!ctx || !logical_id) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_set_session: NULL argument");
  size_t len = strlen(logical_id);
  if (len > 1024) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_set_session: id longer than 1024 bytes");
  std::lock_guard<std::mutex> lock(ctx->mu);
  mb_blake3_session_id(logical_id, len, ctx->sid);
  return MB_OK;
}
extern "C" int mb_ctx_set_session_raw(mb_ctx *ctx, const uint8_t sid[16]) {
  if (!ctx || !sid) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_set_session_raw: NULL argument");
  std::lock_guard<std::mutex> lock(ctx->mu);
  memcpy(ctx->sid, sid, 16);
  return MB_OK;
}
extern "C" int mb_ctx_get_session_raw(const mb_ctx *ctx, uint8_t sid[16]) {
  if (!ctx || !sid) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_get_session_raw: NULL argument");
  memcpy(sid, ctx->sid, 16);
  return MB_OK;
}
extern "C" int mb_ctx_set_entropy_key(mb_ctx *ctx, const uint8_t key[16]) {
  if (!ctx || !key) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_set_entropy_key: NULL argument");
  std::lock_guard<std::mutex> lock(ctx->mu);
  memcpy(ctx->entropy_key, key, 16);
  ctx->deterministic = true;
  ctx->entropy_counter = 0;
  ctx->have_setup = false;
  return MB_OK;
}
extern "C" int mb_ctx_sync(mb_ctx *ctx) {
  if (!ctx) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_sync: NULL ctx");
  MB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MB_OK;
}
extern "C" void *mb_ctx_stream(mb_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
extern "C" int mb_ctx_timer_start(mb_ctx *ctx) {
  MB_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  return MB_OK;
}
extern "C" int mb_ctx_timer_stop(mb_ctx *ctx, float *ms) {
  MB_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  MB_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  MB_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
  return MB_OK;
}
extern "C" uint64_t mb_ctx_launch_count(const mb_ctx *ctx) { return ctx ? ctx->launches.load() : 0; }
extern "C" int mb_ctx_set_fused_bits(mb_ctx *ctx, int enable) {
  if (!ctx) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_set_fused_bits: NULL ctx");
  ctx->fused_bits = enable != 0;
  return MB_OK;
}
extern "C" int mb_ctx_reserve_sms(mb_ctx *ctx, int count) {
  if (!ctx || count < 0 || count > 64) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_reserve_sms: bad count");
  ctx->reserve_sms = count;
  return MB_OK;
}
extern "C" int mb_ctx_set_gemm_mode(mb_ctx *ctx, int mode) {
  // modes 0..2 select the implementation; 3 = tensor pipe restricted to single-CTA tiles (no cta_group::2)
  if (!ctx || mode < 0 || mode > 3) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_set_gemm_mode: bad mode");
  ctx->gemm_pair_mode = mode == 3 ? 1 : 0;
  ctx->gemm_mode = mode == 3 ? 2 : mode;
  return MB_OK;
}

static int prof_event(mb_ctx *ctx, cudaEvent_t *ev) {
  if (!ctx->prof_pool.empty()) {
    *ev = ctx->prof_pool.back();
    ctx->prof_pool.pop_back();
    return MB_OK;
  }
  MB_CUDA(ctx, cudaEventCreate(ev));
  return MB_OK;
}
int mb_prof_begin(mb_ctx *ctx, cudaEvent_t *start) {
  *start = nullptr;
  if (!ctx->profile) return MB_OK;
  std::lock_guard<std::mutex> lock(ctx->prof_mu);
  MB_CHECK(prof_event(ctx, start));
  MB_CUDA(ctx, cudaEventRecord(*start, ctx->stream));
  return MB_OK;
}
int mb_prof_end(mb_ctx *ctx, cudaEvent_t start) {
  if (!start) return MB_OK;
  cudaEvent_t stop;
  std::lock_guard<std::mutex> lock(ctx->prof_mu);
  MB_CHECK(prof_event(ctx, &stop));
  MB_CUDA(ctx, cudaEventRecord(stop, ctx->stream));
  ctx->prof_events.emplace_back(start, stop);
  return MB_OK;
}
extern "C" int mb_ctx_profile_enable(mb_ctx *ctx, int enable) {
  if (!ctx) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_profile_enable: NULL ctx");
  ctx->profile = enable != 0;
  return MB_OK;
}
extern "C" int mb_ctx_profile_read(mb_ctx *ctx, int kernel, float *total_ms, uint64_t *launches) {
  if (!ctx || !total_ms || !launches || kernel != 0)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_profile_read: bad argument");
  MB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  std::lock_guard<std::mutex> lock(ctx->prof_mu);
  float total = 0.f;
  for (auto &pr : ctx->prof_events) {
    float ms = 0.f;
    MB_CUDA(ctx, cudaEventElapsedTime(&ms, pr.first, pr.second));
    total += ms;
    ctx->prof_pool.push_back(pr.first);
    ctx->prof_pool.push_back(pr.second);
  }
  *total_ms = total;
  *launches = ctx->prof_events.size();
  ctx->prof_events.clear();
  return MB_OK;
}
// ---- per-kernel trace ---------------------------------------------------------------------------------------------
// One event behind every launch while the trace is on.  A launch's time is the distance to the event before it on the
// same stream (the previous launch's event or the start marker): on a busy stream that is the kernel's run time plus
// whatever memset / copy / allocation was enqueued between the two launches.
int mb_trace_mark(mb_ctx *ctx, const char *name, double bytes, double aes_blocks, double tensor_ops) {
  cudaEvent_t ev;
  std::lock_guard<std::mutex> lock(ctx->prof_mu);
  MB_CHECK(prof_event(ctx, &ev));
  MB_CUDA(ctx, cudaEventRecord(ev, ctx->stream));
  ctx->trace_events.push_back(mb_ctx::TraceRec{name, ev, bytes, aes_blocks, tensor_ops});
  return MB_OK;
}
extern "C" int mb_ctx_trace_enable(mb_ctx *ctx, int enable) {
  if (!ctx) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_trace_enable: NULL ctx");
  {
    std::lock_guard<std::mutex> lock(ctx->prof_mu);
    for (auto &te : ctx->trace_events) ctx->prof_pool.push_back(te.ev);
    ctx->trace_events.clear();
  }
  ctx->trace = enable != 0;
  if (ctx->trace) MB_CHECK(mb_trace_mark(ctx, nullptr));  // start marker
  return MB_OK;
}
extern "C" int mb_ctx_trace_read(mb_ctx *ctx, char *buf, size_t cap, size_t *needed) {
  if (!ctx || !needed) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_trace_read: NULL argument");
  MB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  std::lock_guard<std::mutex> lock(ctx->prof_mu);
  struct Acc { std::string name; uint64_t n; double ms, bytes, blocks, ops; };
  std::vector<Acc> acc;
  for (size_t i = 1; i < ctx->trace_events.size(); i++) {
    const auto &te = ctx->trace_events[i];
    if (!te.name) continue;
    float ms = 0.f;
    MB_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->trace_events[i - 1].ev, te.ev));
    size_t j = 0;
    for (; j < acc.size(); j++)
      if (acc[j].name == te.name) break;
    if (j == acc.size()) acc.push_back(Acc{te.name, 0, 0.0, 0.0, 0.0, 0.0});
    acc[j].n++;
    acc[j].ms += ms;
    acc[j].bytes += te.bytes;
    acc[j].blocks += te.aes_blocks;
    acc[j].ops += te.tensor_ops;
  }
  std::string text;
  char line[320];
  for (auto &a : acc) {
    snprintf(line, sizeof(line), "%s\t%llu\t%.6f\t%.6e\t%.6e\t%.6e\n", a.name.c_str(), (unsigned long long)a.n, a.ms, a.bytes,
             a.blocks, a.ops);
    text += line;
  }
  *needed = text.size() + 1;
  if (buf && cap >= text.size() + 1) memcpy(buf, text.c_str(), text.size() + 1);
  // keep the last event as the new start marker so consecutive reads partition the stream
  if (!ctx->trace_events.empty()) {
    for (size_t i = 0; i + 1 < ctx->trace_events.size(); i++) ctx->prof_pool.push_back(ctx->trace_events[i].ev);
    auto last = ctx->trace_events.back();
    ctx->trace_events.clear();
    ctx->trace_events.push_back(mb_ctx::TraceRec{nullptr, last.ev, 0, 0, 0});
  }
  return MB_OK;
}

extern "C" int mb_host_alloc_pinned(mb_ctx *ctx, size_t nbytes, void **out) {
  if (!ctx || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_host_alloc_pinned: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  MB_CUDA(ctx, cudaHostAlloc(out, nbytes ? nbytes : 1, cudaHostAllocDefault));
  return MB_OK;
}
extern "C" int mb_host_free_pinned(mb_ctx *ctx, void *ptr) {
  if (ptr) MB_CUDA(ctx, cudaFreeHost(ptr));
  return MB_OK;
}

int mb_gemm_ctx_register(mb_ctx *ctx) {
  if (!ctx->gemm_registered) {
    ctx->gemm_registered = true;
    g_gemm_ctxs[ctx->device & 63].fetch_add(1);
  }
  return g_gemm_ctxs[ctx->device & 63].load();
}

int mb_scratch(mb_ctx *ctx, int slot, size_t bytes, void **out) {
  // grown in stream order (no device-wide synchronisation: a peer may be waiting for work this thread has yet to
  // enqueue, csrc/peer_transport.cu)
  if (ctx->scratch_bytes[slot] < bytes) {
    if (ctx->scratch[slot]) {
      MB_CUDA(ctx, cudaFreeAsync(ctx->scratch[slot], ctx->stream));
      ctx->scratch[slot] = nullptr;
      ctx->scratch_bytes[slot] = 0;
    }
    MB_CUDA(ctx, cudaMallocAsync(&ctx->scratch[slot], bytes, ctx->stream));
    ctx->scratch_bytes[slot] = bytes;
  }
  *out = ctx->scratch[slot];
  return MB_OK;
}

// ---- tensors ------------------------------------------------------------------

static int tensor_new_on(mb_ctx *ctx, cudaStream_t stream, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out);

int mb_tensor_new(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out) {
  return tensor_new_on(ctx, ctx ? ctx->stream : nullptr, dtype, rank, shape, out);
}

// A tensor whose memory is NOT ordered after the context's stream: allocated on the auxiliary stream, which is
// drained before returning, so another stream or library (a transport receiving into it) may write it at once
// while earlier work of the context is still running.  The writer must be complete before the context reads it.
int mb_tensor_new_detached(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out) {
  if (!ctx) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "tensor: NULL argument");
  MB_CHECK(tensor_new_on(ctx, ctx->aux_stream, dtype, rank, shape, out));
  MB_CUDA(ctx, cudaStreamSynchronize(ctx->aux_stream));
  return MB_OK;
}

static int tensor_new_on(mb_ctx *ctx, cudaStream_t stream, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out) {
  if (!ctx || !out) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "tensor: NULL argument");
  if (rank < 0 || rank > MB_MAX_RANK) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "tensor: rank out of range");
  if (dtype != MB_RING64 && dtype != MB_RING128 && dtype != MB_BIT && dtype != MB_FLOAT64)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "tensor: bad dtype");
  // shapes can come straight off the wire (a peer's 10-word header): the element count must not wrap
  int64_t numel = 1;
  for (int i = 0; i < rank; i++) {
    if (shape[i] < 0) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "tensor: negative dimension");
    if (__builtin_mul_overflow(numel, shape[i], &numel) || numel > MB_MAX_NUMEL)
      return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "tensor: element count overflows / exceeds 2^40");
  }
  mb_tensor *t = new mb_tensor();
  t->dtype = dtype;
  t->rank = rank;
  for (int i = 0; i < MB_MAX_RANK; i++) t->shape[i] = i < rank ? shape[i] : 1;
  t->numel = numel;
  t->nbytes = dtype == MB_BIT ? (size_t)(numel + 7) / 8 : (size_t)numel * (dtype == MB_RING128 ? 16 : 8);
  t->alloc_bytes = ((t->nbytes + 15) / 16) * 16;
  if (t->alloc_bytes == 0) t->alloc_bytes = 16;
  t->refs = 1;
  t->device = ctx->device;
  t->dptr = nullptr;
  t->parent = nullptr;
  cudaError_t e = cudaMallocAsync(&t->dptr, t->alloc_bytes, stream);
  if (e != cudaSuccess) {
    delete t;
    return mb_fail(ctx, MB_ERR_CUDA, std::string("cudaMallocAsync: ") + cudaGetErrorString(e));
  }
  if (dtype == MB_BIT) {
    // packed-bit invariant: bits beyond numel (and padding bytes) are zero
    e = cudaMemsetAsync((char *)t->dptr + (t->alloc_bytes - 16), 0, 16, stream);
    if (e != cudaSuccess) {
      cudaFreeAsync(t->dptr, stream);
      delete t;
      return mb_fail(ctx, MB_ERR_CUDA, std::string("cudaMemsetAsync: ") + cudaGetErrorString(e));
    }
  }
  *out = t;
  return MB_OK;
}

mb_tensor *mb_tensor_retain(mb_tensor *t) {
  t->refs.fetch_add(1);
  return t;
}

bool mb_same_shape(const mb_tensor *a, const mb_tensor *b) {
  if (a->rank != b->rank) return false;
  for (int i = 0; i < a->rank; i++)
    if (a->shape[i] != b->shape[i]) return false;
  return true;
}

extern "C" int mb_tensor_alloc(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, mb_tensor **out) {
  return mb_tensor_new(ctx, dtype, rank, shape, out);
}

extern "C" int mb_tensor_from_host(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, const void *host_data,
                                   mb_tensor **out) {
  if (!host_data && rank >= 0) {
    int64_t n = 1;
    for (int i = 0; i < rank; i++) n *= shape[i];
    if (n > 0) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_from_host: NULL data");
  }
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, dtype, rank, shape, &t));
  if (t->nbytes) {
    cudaError_t e = cudaMemcpyAsync(t->dptr, host_data, t->nbytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
      mb_tensor_free(ctx, t);
      return mb_fail(ctx, MB_ERR_CUDA, std::string("cudaMemcpyAsync H2D: ") + cudaGetErrorString(e));
    }
    // pageable host memory: the copy has been staged when the call returns only after a sync
    MB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  *out = t;
  return MB_OK;
}

// Asynchronous forms for callers that pipeline transfers against compute with several contexts on one device
// (loader / compute / drainer streams): the host buffers must be page-locked and stay valid until the stream has
// passed the copy; ordering between contexts is the caller's (mb_ctx_wait).
extern "C" int mb_tensor_from_host_async(mb_ctx *ctx, mb_dtype dtype, int rank, const int64_t *shape, const void *pinned_host,
                                         mb_tensor **out) {
  if (!ctx || !out || !pinned_host) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_from_host_async: NULL argument");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, dtype, rank, shape, &t));
  if (t->nbytes) {
    cudaError_t e = cudaMemcpyAsync(t->dptr, pinned_host, t->nbytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
      mb_tensor_free(ctx, t);
      return mb_fail(ctx, MB_ERR_CUDA, std::string("cudaMemcpyAsync H2D: ") + cudaGetErrorString(e));
    }
  }
  *out = t;
  return MB_OK;
}

extern "C" int mb_tensor_to_host_async(mb_ctx *ctx, const mb_tensor *t, void *pinned_host_out, size_t nbytes) {
  if (!ctx || !t || (!pinned_host_out && nbytes)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_to_host_async: NULL");
  if (nbytes != t->nbytes) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_to_host_async: size mismatch");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (nbytes) MB_CUDA(ctx, cudaMemcpyAsync(pinned_host_out, t->dptr, nbytes, cudaMemcpyDeviceToHost, ctx->stream));
  return MB_OK;
}

// Work enqueued on `ctx` after this call starts only when everything enqueued on `other` so far has finished.
extern "C" int mb_ctx_wait(mb_ctx *ctx, mb_ctx *other) {
  if (!ctx || !other) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_wait: NULL context");
  if (ctx->device != other->device) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_ctx_wait: contexts on different devices");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  std::lock_guard<std::mutex> lock(other->mu);
  MB_CUDA(ctx, cudaEventRecord(other->ev1, other->stream));
  MB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, other->ev1, 0));
  return MB_OK;
}

extern "C" int mb_tensor_to_host(mb_ctx *ctx, const mb_tensor *t, void *host_out, size_t nbytes) {
  if (!ctx || !t || (!host_out && nbytes)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_to_host: NULL");
  if (nbytes != t->nbytes) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_to_host: size mismatch");
  if (nbytes) MB_CUDA(ctx, cudaMemcpyAsync(host_out, t->dptr, nbytes, cudaMemcpyDeviceToHost, ctx->stream));
  MB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return MB_OK;
}

extern "C" int mb_tensor_free(mb_ctx *ctx, mb_tensor *t) {
  if (!t) return MB_OK;
  if (t->refs.fetch_sub(1) == 1) {
    mb_tensor *parent = t->parent;
    if (!parent && t->dptr) cudaFreeAsync(t->dptr, ctx ? ctx->stream : 0);
    delete t;
    if (parent) return mb_tensor_free(ctx, parent);
  }
  return MB_OK;
}

int mb_tensor_view(mb_ctx *ctx, const mb_tensor *parent, size_t byte_offset, int rank, const int64_t *shape, mb_tensor **out) {
  if (!ctx || !parent || !out || rank < 0 || rank > MB_MAX_RANK) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "view: bad argument");
  int64_t numel = 1;
  for (int i = 0; i < rank; i++)
    if (shape[i] < 0 || __builtin_mul_overflow(numel, shape[i], &numel) || numel > MB_MAX_NUMEL)
      return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "view: bad shape");
  mb_tensor *t = new mb_tensor();
  t->dtype = parent->dtype;
  t->rank = rank;
  for (int i = 0; i < MB_MAX_RANK; i++) t->shape[i] = i < rank ? shape[i] : 1;
  t->numel = numel;
  t->nbytes = t->dtype == MB_BIT ? (size_t)(numel + 7) / 8 : (size_t)numel * (t->dtype == MB_RING128 ? 16 : 8);
  t->alloc_bytes = ((t->nbytes + 15) / 16) * 16;
  if (t->alloc_bytes == 0) t->alloc_bytes = 16;
  if (byte_offset % 16 || t->nbytes % 16 || byte_offset + t->nbytes > parent->alloc_bytes) {
    delete t;
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "view: not a 16-byte aligned sub-range");
  }
  t->refs = 1;
  t->device = parent->device;
  t->dptr = (char *)parent->dptr + byte_offset;
  // views of views hang off the root allocation
  mb_tensor *root = const_cast<mb_tensor *>(parent);
  while (root->parent) root = root->parent;
  t->parent = mb_tensor_retain(root);
  *out = t;
  return MB_OK;
}

extern "C" int mb_tensor_copy_to_device(mb_ctx *ctx, const mb_tensor *t, void *dst) {
  if (!ctx || !t || (!dst && t->nbytes)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_copy_to_device: NULL");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (t->nbytes) MB_CUDA(ctx, cudaMemcpyAsync(dst, t->dptr, t->nbytes, cudaMemcpyDeviceToDevice, ctx->stream));
  return MB_OK;
}
extern "C" int mb_tensor_copy_from_device(mb_ctx *ctx, mb_tensor *t, const void *src) {
  if (!ctx || !t || (!src && t->nbytes)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_tensor_copy_from_device: NULL");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (t->nbytes) MB_CUDA(ctx, cudaMemcpyAsync(t->dptr, src, t->nbytes, cudaMemcpyDeviceToDevice, ctx->stream));
  return MB_OK;
}
extern "C" mb_dtype mb_tensor_dtype(const mb_tensor *t) { return t->dtype; }
extern "C" int mb_tensor_rank(const mb_tensor *t) { return t->rank; }
extern "C" const int64_t *mb_tensor_shape(const mb_tensor *t) { return t->shape; }
extern "C" int64_t mb_tensor_numel(const mb_tensor *t) { return t->numel; }
extern "C" size_t mb_tensor_nbytes(const mb_tensor *t) { return t->nbytes; }
extern "C" void *mb_tensor_device_ptr(const mb_tensor *t) { return t->dptr; }

extern "C" int mb_reshape(mb_ctx *ctx, const mb_tensor *x, int rank, const int64_t *shape, mb_tensor **out) {
  if (!x) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_reshape: NULL");
  int64_t n = 1;
  for (int i = 0; i < rank; i++) n *= shape[i];
  if (n != x->numel) return mb_fail(ctx, MB_ERR_KERNEL, "mb_reshape: element count differs");
  mb_tensor *t = nullptr;
  MB_CHECK(mb_tensor_new(ctx, x->dtype, rank, shape, &t));
  if (x->nbytes)
    MB_CUDA(ctx, cudaMemcpyAsync(t->dptr, x->dptr, x->nbytes, cudaMemcpyDeviceToDevice, ctx->stream));
  *out = t;
  return MB_OK;
}
