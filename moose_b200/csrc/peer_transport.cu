// Peer-memory binding of the Send / Receive transport (mb_transport_fn) for party-per-GPU runs on one NVSwitch node:
// every party owns an inbox in its HBM (one byte ring + one small message ring per peer), the peers map it through
// CUDA IPC, and a transfer is a copy INTO the receiver's ring over NVLink followed by a release store of a
// cumulative byte counter.  Everything is enqueued on the context's stream -- wait for ring space, copy, signal on
// the sender; wait for the counter, copy out, acknowledge on the receiver -- so a party never synchronises its
// stream with the host for a tensor transfer and its kernels, copies and waits run back to back (capability bit 0 of
// mode 5).  Host payloads (keys, seeds, headers) never touch the GPU: they go through message rings in POSIX shared
// memory between the processes (release / acquire counters), so a seed arrives while the GPUs are busy.
//
// No counterpart in the reference: its transport is gRPC / TCP between hosts (moose/src/networking/mod.rs:16-54).
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <map>

#include "common.h"

#include <cstdlib>

namespace {

constexpr size_t CTRL_SLOT = 4096;  // host payloads: headers, keys, seeds, short strings
constexpr int CTRL_SLOTS = 64;
constexpr size_t WORDS_BYTES = 128;  // per peer, device: data_flag (written by the sender), data_ack (by the receiver)
constexpr size_t ALIGN = 256;
constexpr unsigned long long WAIT_NS = 60ull * 1000000000ull;

enum { W_DATA_FLAG = 0, W_DATA_ACK = 2 };

struct Blob {  // what the parties exchange at start-up (128 bytes)
  cudaIpcMemHandle_t ipc;
  int64_t pid;
  uint64_t raw;
  int32_t device, rank;
  uint64_t ring_bytes;
  char shm_name[128 - sizeof(cudaIpcMemHandle_t) - 32];  // this party's host message segment
};

// Host message segment of one party (POSIX shared memory): per sending peer a ring of CTRL_SLOTS slots.
struct HostRing {
  alignas(64) std::atomic<uint64_t> written;   // by the sender
  alignas(64) std::atomic<uint64_t> consumed;  // by the owner
  alignas(64) uint8_t slots[CTRL_SLOTS][CTRL_SLOT];
};
static_assert(sizeof(Blob) == 128, "handle blob is 128 bytes");

__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// one thread: spin until *word >= target; gives up after `wait_ns` and raises *status (checked by flush).  The
// failure is sticky: once *status is set, later waits return at once and later signals publish nothing, so no
// acknowledgement or data flag goes out for bytes that never arrived and the peers time out in turn instead of
// computing on stale ring contents.
__global__ void k_peer_wait(const uint64_t *word, uint64_t target, int *status, uint64_t wait_ns) {
  if (*(volatile int *)status) return;
  const uint64_t t0 = globaltimer_ns();
  while (ld_acquire_sys(word) < target) {
    __nanosleep(200);
    if (globaltimer_ns() - t0 > wait_ns) {
      *status = 1;
      return;
    }
  }
}

// one thread: everything before this kernel in the stream (the copy) is complete; publish the counter
__global__ void k_peer_signal(uint64_t *word, uint64_t value, const int *status) {
  if (*(volatile const int *)status) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(word), "l"(value) : "memory");
}

}  // namespace

struct mb_peer_transport {
  mb_ctx *ctx;
  int rank, nranks;
  size_t ring_bytes;
  std::map<std::string, int> rank_of;
  uint8_t *inbox;                  // this party's allocation
  std::vector<uint8_t *> peer_in;  // peer_in[q]: q's inbox mapped here (nullptr for q == rank)
  std::vector<bool> ipc_opened;
  std::vector<uint64_t> sent_data, recv_data;
  std::string shm_name;
  HostRing *host_in;                 // this party's message segment: host_in[slot_of(rank, p)] is written by p
  std::vector<HostRing *> peer_host;  // peer_host[q]: q's segment mapped here
  std::vector<bool> shm_mapped;
  int *status;  // device word: a wait timed out
  uint64_t wait_ns;  // give-up time of a device-side wait (MOOSE_B200_PEER_TIMEOUT_S, default 60 s)
  int64_t posted;
  bool connected;
};

namespace {

int pfail(mb_peer_transport *t, const std::string &m) {
  mb_fail(t ? t->ctx : nullptr, MB_ERR_KERNEL, "peer transport: " + m);
  return 1;
}

#define PCU(t, call)                                                                               \
  do {                                                                                             \
    cudaError_t _e = (call);                                                                       \
    if (_e != cudaSuccess) return pfail((t), std::string(#call) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// Wait on the stream until *word >= target: one spinning thread.  (Stream memory operations -- cuStreamWaitValue64 --
// were measured instead: 48.1 ms against 43.3 ms per MLP batch with processes per GPU, and three contexts of one
// process on one GPU block each other for good; not used.)
int enqueue_wait(mb_peer_transport *t, cudaStream_t s, uint64_t *word, uint64_t target) {
  k_peer_wait<<<1, 1, 0, s>>>(word, target, t->status, t->wait_ns);
  if (t->ctx->trace && s == t->ctx->stream) mb_trace_mark(t->ctx, "peer_wait");   // time spent waiting for the peer
  return 0;
}

inline int slot_of(int owner, int peer) { return peer - (peer > owner ? 1 : 0); }
inline size_t inbox_bytes(int nranks, size_t ring) { return (size_t)(nranks - 1) * (WORDS_BYTES + ring); }
// layout of an inbox: [words of every peer][byte rings of every peer]
inline uint64_t *words(uint8_t *inbox, int slot) { return (uint64_t *)(inbox + (size_t)slot * WORDS_BYTES); }
inline uint8_t *data_ring(uint8_t *inbox, int nranks, size_t ring, int slot) {
  return inbox + (size_t)(nranks - 1) * WORDS_BYTES + (size_t)slot * ring;
}
inline size_t host_segment_bytes(int nranks) { return (size_t)(nranks - 1) * sizeof(HostRing); }
inline size_t round_up(size_t n) { return (n + ALIGN - 1) / ALIGN * ALIGN; }

int peer_rank_of(mb_peer_transport *t, const char *name, int *out) {
  auto it = t->rank_of.find(name ? name : "");
  if (it == t->rank_of.end() || it->second == t->rank) return pfail(t, std::string("unknown peer ") + (name ? name : "(null)"));
  *out = it->second;
  return 0;
}

// A message (<= ring / 2, padded to 256 bytes) occupies contiguous ring space: when it does not fit before the end of
// the ring both sides skip to the start (the skipped bytes count as sent / consumed).  Returns the position and
// advances the cumulative counter past the message.
inline size_t place(uint64_t &counter, size_t C, size_t padded, uint64_t *before) {
  size_t pos = counter % C;
  if (pos + padded > C) {
    counter += C - pos;
    pos = 0;
  }
  *before = counter;
  counter += padded;
  return pos;
}

// Ring space in q for a message of n bytes, on the context's stream: waits until the receiver has consumed enough,
// returns where the message goes.  Whoever fills it (a copy, or the kernel that computes the payload) is followed
// by publish().
int reserve(mb_peer_transport *t, int q, size_t n, uint8_t **where) {
  const size_t C = t->ring_bytes, padded = round_up(n);
  uint64_t before = 0;
  const size_t pos = place(t->sent_data[q], C, padded, &before);
  uint64_t *mine = words(t->inbox, slot_of(t->rank, q));
  if (t->sent_data[q] > C && enqueue_wait(t, t->ctx->stream, mine + W_DATA_ACK, t->sent_data[q] - C)) return 1;
  *where = data_ring(t->peer_in[q], t->nranks, C, slot_of(q, t->rank)) + pos;
  return 0;
}
int publish(mb_peer_transport *t, int q) {
  uint64_t *theirs = words(t->peer_in[q], slot_of(q, t->rank));
  k_peer_signal<<<1, 1, 0, t->ctx->stream>>>(theirs + W_DATA_FLAG, t->sent_data[q], t->status);
  PCU(t, cudaGetLastError());
  if (t->ctx->trace) mb_trace_mark(t->ctx, "peer_publish");
  return 0;
}

// One chunk (<= ring / 2) of a device payload into q's ring, on the context's stream.
int send_chunk(mb_peer_transport *t, int q, const uint8_t *src, size_t n) {
  uint8_t *where = nullptr;
  if (reserve(t, q, n, &where)) return 1;
  PCU(t, cudaMemcpyAsync(where, src, n, cudaMemcpyDefault, t->ctx->stream));
  if (t->ctx->trace) mb_trace_mark(t->ctx, "peer_copy_into_ring", (double)n);
  return publish(t, q);
}

// One chunk out of this party's ring for peer p, on the context's stream.
int recv_chunk(mb_peer_transport *t, int p, uint8_t *dst, size_t n) {
  cudaStream_t s = t->ctx->stream;
  const size_t C = t->ring_bytes, padded = round_up(n);
  uint64_t before = 0;
  const size_t pos = place(t->recv_data[p], C, padded, &before);
  uint64_t *mine = words(t->inbox, slot_of(t->rank, p)), *theirs = words(t->peer_in[p], slot_of(p, t->rank));
  if (enqueue_wait(t, s, mine + W_DATA_FLAG, t->recv_data[p])) return 1;
  const uint8_t *ring = data_ring(t->inbox, t->nranks, C, slot_of(t->rank, p));
  PCU(t, cudaMemcpyAsync(dst, ring + pos, n, cudaMemcpyDeviceToDevice, s));
  if (t->ctx->trace && s == t->ctx->stream) mb_trace_mark(t->ctx, "peer_copy_out_of_ring", 2.0 * (double)n);
  k_peer_signal<<<1, 1, 0, s>>>(theirs + W_DATA_ACK, t->recv_data[p], t->status);
  PCU(t, cudaGetLastError());
  if (t->ctx->trace && s == t->ctx->stream) mb_trace_mark(t->ctx, "peer_ack");
  return 0;
}

// Send and / or receive device payloads; the chunks of the two directions alternate so that an exchange larger
// than the rings cannot block (every party sends chunk k, then receives chunk k).
int exchange(mb_peer_transport *t, int q, const uint8_t *src, size_t ns, int p, uint8_t *dst, size_t nr) {
  const size_t chunk = t->ring_bytes / 2;
  size_t so = 0, ro = 0;
  while (so < ns || ro < nr) {
    if (so < ns) {
      const size_t n = ns - so < chunk ? ns - so : chunk;
      if (send_chunk(t, q, src + so, n)) return 1;
      so += n;
    }
    if (ro < nr) {
      const size_t n = nr - ro < chunk ? nr - ro : chunk;
      if (recv_chunk(t, p, dst + ro, n)) return 1;
      ro += n;
    }
  }
  return 0;
}

template <class Pred>
bool spin_until(Pred ready) {
  const auto t0 = std::chrono::steady_clock::now();
  for (unsigned i = 0; !ready(); i++) {
    if ((i & 1023) == 1023) {
      sched_yield();
      if (std::chrono::steady_clock::now() - t0 > std::chrono::nanoseconds(WAIT_NS)) return false;
    }
  }
  return true;
}

// Host payloads: shared-memory message rings, no GPU work (a seed must not queue behind a long kernel).
int host_send(mb_peer_transport *t, int q, const void *buf, size_t n) {
  if (n > CTRL_SLOT) return pfail(t, "host payload larger than a message slot");
  HostRing *r = t->peer_host[q] + slot_of(q, t->rank);
  const uint64_t w = r->written.load(std::memory_order_relaxed);
  if (!spin_until([&] { return w - r->consumed.load(std::memory_order_acquire) < (uint64_t)CTRL_SLOTS; }))
    return pfail(t, "host send: the peer does not drain its message ring");
  memcpy(r->slots[w % CTRL_SLOTS], buf, n);
  r->written.store(w + 1, std::memory_order_release);
  return 0;
}

int host_recv(mb_peer_transport *t, int p, void *buf, size_t n) {
  if (n > CTRL_SLOT) return pfail(t, "host payload larger than a message slot");
  HostRing *r = t->host_in + slot_of(t->rank, p);
  const uint64_t c = r->consumed.load(std::memory_order_relaxed);
  if (!spin_until([&] { return r->written.load(std::memory_order_acquire) > c; })) return pfail(t, "host receive timed out");
  memcpy(buf, r->slots[c % CTRL_SLOTS], n);
  r->consumed.store(c + 1, std::memory_order_release);
  return 0;
}

int check_status(mb_peer_transport *t) {
  int st = 0;
  PCU(t, cudaMemcpy(&st, t->status, sizeof(int), cudaMemcpyDeviceToHost));
  if (st) return pfail(t, "a wait for a peer timed out");
  return 0;
}

// mb_transport_fn (include/moose_b200.h)
int peer_transport_cb(void *user, int mode, const char *peer, void *buf, size_t nbytes, int on_device) {
  mb_peer_transport *t = (mb_peer_transport *)user;
  if (!t) return 1;
  if (!t->connected) return pfail(t, "not connected");
  PCU(t, cudaSetDevice(t->ctx->device));
  if (mode == 5) {  // capabilities: device payloads are ordered on the context's stream; reserve / publish
    *(int64_t *)buf = 3;
    return 0;
  }
  if (mode == 6 || mode == 7) {  // the payload is written into the receiver's ring by the caller's own kernel
    mb_reserve *r = (mb_reserve *)buf;
    int to = 0;
    if (peer_rank_of(t, r->peer, &to)) return 1;
    if (mode == 7) {
      t->posted++;
      return publish(t, to);
    }
    r->ptr = nullptr;
    if (r->nbytes == 0 || r->nbytes > t->ring_bytes / 2) return 0;  // the caller falls back to an ordinary send
    uint8_t *where = nullptr;
    if (reserve(t, to, r->nbytes, &where)) return 1;
    r->ptr = where;
    return 0;
  }
  if (mode == 2 || mode == 3) {
    if (mode == 3) {
      PCU(t, cudaStreamSynchronize(t->ctx->stream));
      if (check_status(t)) return 1;
    }
    *(int64_t *)buf = t->posted;  // the source of a send is read in stream order: nothing to keep alive
    return 0;
  }
  if (mode == 4) {
    const mb_sendrecv *d = (const mb_sendrecv *)buf;
    int to = 0, from = 0;
    if (peer_rank_of(t, d->send_peer, &to) || peer_rank_of(t, d->recv_peer, &from)) return 1;
    return exchange(t, to, (const uint8_t *)d->send_buf, d->send_bytes, from, (uint8_t *)d->recv_buf, d->recv_bytes);
  }
  if (mode != 0 && mode != 1) return pfail(t, "unknown mode");
  int other = 0;
  if (peer_rank_of(t, peer, &other)) return 1;
  if (!on_device) return mode == 1 ? host_send(t, other, buf, nbytes) : host_recv(t, other, buf, nbytes);
  if (mode == 1) {
    t->posted++;
    return exchange(t, other, (const uint8_t *)buf, nbytes, -1, nullptr, 0);
  }
  return exchange(t, -1, nullptr, 0, other, (uint8_t *)buf, nbytes);
}

}  // namespace

static int peer_init(mb_peer_transport *t);

extern "C" int mb_peer_transport_create(mb_ctx *ctx, int rank, int nranks, const char *const *role_names, size_t ring_bytes,
                                        mb_peer_transport **out) {
  if (!ctx || !role_names || !out || nranks < 2 || rank < 0 || rank >= nranks)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_peer_transport_create: bad argument");
  if (ring_bytes == 0) ring_bytes = (size_t)1 << 30;
  if (ring_bytes % (2 * ALIGN)) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_peer_transport_create: ring_bytes must be a multiple of 512");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  mb_peer_transport *t = new mb_peer_transport();
  t->ctx = ctx;
  t->rank = rank;
  t->nranks = nranks;
  t->ring_bytes = ring_bytes;
  t->posted = 0;
  t->connected = false;
  for (int i = 0; i < nranks; i++) t->rank_of[role_names[i] ? role_names[i] : ""] = i;
  t->peer_in.assign(nranks, nullptr);
  t->ipc_opened.assign(nranks, false);
  t->sent_data.assign(nranks, 0);
  t->recv_data.assign(nranks, 0);
  t->peer_host.assign(nranks, nullptr);
  t->shm_mapped.assign(nranks, false);
  t->inbox = nullptr;
  t->status = nullptr;
  t->wait_ns = WAIT_NS;
  if (const char *e = getenv("MOOSE_B200_PEER_TIMEOUT_S")) {
    const double sec = atof(e);
    if (sec > 0) t->wait_ns = (uint64_t)(sec * 1e9);
  }
  t->host_in = nullptr;
  const int rc = peer_init(t);
  if (rc != MB_OK) {
    mb_peer_transport_free(t);
    return rc;
  }
  *out = t;
  return MB_OK;
}

// allocations of mb_peer_transport_create; the caller frees the half-built transport when this fails
static int peer_init(mb_peer_transport *t) {
  mb_ctx *ctx = t->ctx;
  const int nranks = t->nranks, rank = t->rank;
  const size_t ring_bytes = t->ring_bytes;
  // cudaMalloc, not the context's pool: IPC handles exist for plain allocations only
  MB_CUDA(ctx, cudaMalloc((void **)&t->inbox, inbox_bytes(nranks, ring_bytes)));
  MB_CUDA(ctx, cudaMemset(t->inbox, 0, (size_t)(nranks - 1) * WORDS_BYTES));
  MB_CUDA(ctx, cudaMalloc((void **)&t->status, sizeof(int)));
  MB_CUDA(ctx, cudaMemset(t->status, 0, sizeof(int)));
  // host message segment: POSIX shared memory, zero-filled by ftruncate
  static std::atomic<unsigned> serial{0};
  char name[32];
  snprintf(name, sizeof(name), "/mbp_%x_%x_%x", (unsigned)getpid(), serial.fetch_add(1), (unsigned)rank);
  t->shm_name = name;
  const int fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
  if (fd < 0 || ftruncate(fd, (off_t)host_segment_bytes(nranks)) != 0) {
    if (fd >= 0) close(fd);
    return mb_fail(ctx, MB_ERR_KERNEL, std::string("mb_peer_transport_create: shm_open ") + name + " failed");
  }
  void *seg = mmap(nullptr, host_segment_bytes(nranks), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (seg == MAP_FAILED) return mb_fail(ctx, MB_ERR_KERNEL, "mb_peer_transport_create: mmap failed");
  t->host_in = (HostRing *)seg;
  MB_CUDA(ctx, cudaDeviceSynchronize());
  return MB_OK;
}

extern "C" int mb_peer_transport_handle(mb_peer_transport *t, uint8_t out[128]) {
  if (!t || !out) return mb_fail(t ? t->ctx : nullptr, MB_ERR_INVALID_ARGUMENT, "mb_peer_transport_handle: NULL argument");
  Blob b;
  memset(&b, 0, sizeof(b));
  MB_CUDA(t->ctx, cudaSetDevice(t->ctx->device));
  MB_CUDA(t->ctx, cudaIpcGetMemHandle(&b.ipc, t->inbox));
  b.pid = (int64_t)getpid();
  b.raw = (uint64_t)(uintptr_t)t->inbox;
  b.device = t->ctx->device;
  b.rank = t->rank;
  b.ring_bytes = t->ring_bytes;
  snprintf(b.shm_name, sizeof(b.shm_name), "%s", t->shm_name.c_str());
  memcpy(out, &b, sizeof(b));
  return MB_OK;
}

// handles: nranks x 128 bytes, handles[r] from rank r.  Inboxes of the same process are used directly (threads, one
// context per party), the others are mapped through CUDA IPC with peer access enabled on first use.
extern "C" int mb_peer_transport_connect(mb_peer_transport *t, const uint8_t *handles) {
  if (!t || !handles) return mb_fail(t ? t->ctx : nullptr, MB_ERR_INVALID_ARGUMENT, "mb_peer_transport_connect: NULL argument");
  mb_ctx *ctx = t->ctx;
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int q = 0; q < t->nranks; q++) {
    if (q == t->rank) continue;
    Blob b;
    memcpy(&b, handles + (size_t)q * 128, sizeof(b));
    if (b.rank != q || b.ring_bytes != t->ring_bytes) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_peer_transport_connect: handle does not match");
    if (b.pid == (int64_t)getpid()) {
      if (b.device != ctx->device) {
        cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) return mb_fail(ctx, MB_ERR_KERNEL, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
      }
      t->peer_in[q] = (uint8_t *)(uintptr_t)b.raw;
    } else {
      void *p = nullptr;
      MB_CUDA(ctx, cudaIpcOpenMemHandle(&p, b.ipc, cudaIpcMemLazyEnablePeerAccess));
      t->peer_in[q] = (uint8_t *)p;
      t->ipc_opened[q] = true;
    }
    b.shm_name[sizeof(b.shm_name) - 1] = 0;
    const int fd = shm_open(b.shm_name, O_RDWR, 0600);
    if (fd < 0) return mb_fail(ctx, MB_ERR_KERNEL, std::string("mb_peer_transport_connect: shm_open ") + b.shm_name + " failed");
    void *seg = mmap(nullptr, host_segment_bytes(t->nranks), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (seg == MAP_FAILED) return mb_fail(ctx, MB_ERR_KERNEL, "mb_peer_transport_connect: mmap failed");
    t->peer_host[q] = (HostRing *)seg;
    t->shm_mapped[q] = true;
  }
  t->connected = true;
  return MB_OK;
}

extern "C" mb_transport_fn mb_peer_transport_fn(void) { return peer_transport_cb; }

extern "C" int mb_peer_transport_free(mb_peer_transport *t) {
  if (!t) return MB_OK;
  cudaSetDevice(t->ctx->device);
  cudaStreamSynchronize(t->ctx->stream);
  for (int q = 0; q < t->nranks; q++) {
    if (t->ipc_opened[q]) cudaIpcCloseMemHandle(t->peer_in[q]);
    if (t->shm_mapped[q]) munmap(t->peer_host[q], host_segment_bytes(t->nranks));
  }
  if (t->host_in) munmap(t->host_in, host_segment_bytes(t->nranks));
  if (!t->shm_name.empty()) shm_unlink(t->shm_name.c_str());
  if (t->status) cudaFree(t->status);
  if (t->inbox) cudaFree(t->inbox);
  delete t;
  return MB_OK;
}
