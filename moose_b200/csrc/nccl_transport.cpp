// Native NCCL binding of the Send / Receive transport (mb_transport_fn) for party-per-GPU runs: ncclSend / ncclRecv
// over NVLink on a dedicated stream, grouped send + receive for the ring exchange, no interpreter in the loop.
// libnccl is opened at run time (dlopen), so libmoose_b200.so itself has no NCCL dependency; a process that already
// loaded NCCL (torch) shares that copy.  The unique id is created on one rank and distributed by the host
// (torch.distributed, a file, MPI ...): that bootstrap is the only thing left outside.
#include <dlfcn.h>

#include <deque>
#include <map>
#include <mutex>

#include "common.h"

namespace {

typedef struct { char internal[128]; } NcclUniqueId;
typedef void *NcclComm;
typedef int NcclResult;

struct NcclApi {
  void *lib = nullptr;
  NcclResult (*GetUniqueId)(NcclUniqueId *) = nullptr;
  NcclResult (*CommInitRank)(NcclComm *, int, NcclUniqueId, int) = nullptr;
  NcclResult (*CommDestroy)(NcclComm) = nullptr;
  NcclResult (*Send)(const void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  NcclResult (*Recv)(void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  NcclResult (*GroupStart)() = nullptr;
  NcclResult (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(NcclResult) = nullptr;
};

NcclApi *nccl_api(std::string *err) {
  static NcclApi api;
  static std::once_flag once;
  static std::string why;
  std::call_once(once, [] {
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
      api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (!api.lib) {
      why = std::string("cannot open libnccl: ") + dlerror();
    } else {
#define LOAD(field, sym)                                         \
  api.field = (decltype(api.field))dlsym(api.lib, sym);          \
  if (!api.field) why = std::string("libnccl lacks ") + sym;
      LOAD(GetUniqueId, "ncclGetUniqueId")
      LOAD(CommInitRank, "ncclCommInitRank")
      LOAD(CommDestroy, "ncclCommDestroy")
      LOAD(Send, "ncclSend")
      LOAD(Recv, "ncclRecv")
      LOAD(GroupStart, "ncclGroupStart")
      LOAD(GroupEnd, "ncclGroupEnd")
      LOAD(GetErrorString, "ncclGetErrorString")
#undef LOAD
    }
  });
  if (!why.empty()) {
    if (err) *err = why;
    return nullptr;
  }
  return &api;
}

constexpr int NCCL_UINT8 = 1;     // ncclUint8
constexpr size_t SLOT_BYTES = 4096;  // host payloads (headers, keys, seeds, short strings)
constexpr int SLOTS = 64;

}  // namespace

struct mb_nccl_transport {
  mb_ctx *ctx;
  NcclApi *api;
  NcclComm comm;
  cudaStream_t stream;
  int rank, nranks;
  std::map<std::string, int> rank_of;
  uint8_t *host_slots;  // pinned, SLOTS x SLOT_BYTES
  uint8_t *dev_slots;
  int next_slot;
  cudaEvent_t slot_done[64];  // recorded behind the last send that used the slot
  bool slot_busy[64];
  std::deque<cudaEvent_t> posted;  // device sends in flight, oldest first
  std::vector<cudaEvent_t> pool;
  int64_t completed;
};

namespace {

int tfail(mb_nccl_transport *t, const std::string &m) { return mb_fail(t ? t->ctx : nullptr, MB_ERR_KERNEL, "nccl transport: " + m) ? 1 : 1; }

#define NC(t, call)                                                                        \
  do {                                                                                     \
    NcclResult _r = (call);                                                                \
    if (_r != 0) return tfail((t), std::string(#call) + ": " + (t)->api->GetErrorString(_r)); \
  } while (0)
#define CU(t, call)                                                              \
  do {                                                                           \
    cudaError_t _e = (call);                                                     \
    if (_e != cudaSuccess) return tfail((t), std::string(#call) + ": " + cudaGetErrorString(_e)); \
  } while (0)

int peer_rank(mb_nccl_transport *t, const char *name, int *out) {
  auto it = t->rank_of.find(name ? name : "");
  if (it == t->rank_of.end()) return tfail(t, std::string("unknown peer ") + (name ? name : "(null)"));
  *out = it->second;
  return 0;
}

void reap_events(mb_nccl_transport *t, bool wait_all) {
  while (!t->posted.empty()) {
    cudaEvent_t ev = t->posted.front();
    if (wait_all) cudaEventSynchronize(ev);
    else if (cudaEventQuery(ev) != cudaSuccess) break;
    t->posted.pop_front();
    t->pool.push_back(ev);
    t->completed++;
  }
}

// mb_transport_fn (include/moose_b200.h)
int nccl_transport_cb(void *user, int mode, const char *peer, void *buf, size_t nbytes, int on_device) {
  mb_nccl_transport *t = (mb_nccl_transport *)user;
  if (!t) return 1;
  CU(t, cudaSetDevice(t->ctx->device));
  if (mode == 5) {  // capabilities: transfers run on this transport's own stream
    *(int64_t *)buf = 0;
    return 0;
  }
  if (mode == 2 || mode == 3) {
    reap_events(t, mode == 3);
    if (mode == 3) CU(t, cudaStreamSynchronize(t->stream));
    *(int64_t *)buf = t->completed;
    return 0;
  }
  if (mode == 4) {
    const mb_sendrecv *d = (const mb_sendrecv *)buf;
    int to = 0, from = 0;
    if (peer_rank(t, d->send_peer, &to) || peer_rank(t, d->recv_peer, &from)) return 1;
    NC(t, t->api->GroupStart());
    NC(t, t->api->Send(d->send_buf, d->send_bytes, NCCL_UINT8, to, t->comm, t->stream));
    NC(t, t->api->Recv(d->recv_buf, d->recv_bytes, NCCL_UINT8, from, t->comm, t->stream));
    NC(t, t->api->GroupEnd());
    CU(t, cudaStreamSynchronize(t->stream));
    return 0;
  }
  int other = 0;
  if (peer_rank(t, peer, &other)) return 1;
  if (on_device) {
    if (mode == 1) {
      NC(t, t->api->Send(buf, nbytes, NCCL_UINT8, other, t->comm, t->stream));
      cudaEvent_t ev;
      if (t->pool.empty()) CU(t, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
      else { ev = t->pool.back(); t->pool.pop_back(); }
      CU(t, cudaEventRecord(ev, t->stream));
      t->posted.push_back(ev);
    } else {
      NC(t, t->api->Recv(buf, nbytes, NCCL_UINT8, other, t->comm, t->stream));
      CU(t, cudaStreamSynchronize(t->stream));
    }
    return 0;
  }
  // host payloads go through a ring of page-locked / device staging slots
  if (nbytes > SLOT_BYTES) return tfail(t, "host payload larger than a staging slot");
  const int slot = t->next_slot;
  t->next_slot = (t->next_slot + 1) % SLOTS;
  uint8_t *h = t->host_slots + (size_t)slot * SLOT_BYTES, *dv = t->dev_slots + (size_t)slot * SLOT_BYTES;
  // A send returns before its H2D copy and ncclSend have run (the Send blocks the stream until the peer posts its
  // Recv): the slot is reused only after the event recorded behind its last use has completed.
  if (t->slot_busy[slot]) {
    CU(t, cudaEventSynchronize(t->slot_done[slot]));
    t->slot_busy[slot] = false;
  }
  if (mode == 1) {
    memcpy(h, buf, nbytes);
    CU(t, cudaMemcpyAsync(dv, h, nbytes, cudaMemcpyHostToDevice, t->stream));
    NC(t, t->api->Send(dv, nbytes, NCCL_UINT8, other, t->comm, t->stream));
    CU(t, cudaEventRecord(t->slot_done[slot], t->stream));
    t->slot_busy[slot] = true;
  } else {
    NC(t, t->api->Recv(dv, nbytes, NCCL_UINT8, other, t->comm, t->stream));
    CU(t, cudaMemcpyAsync(h, dv, nbytes, cudaMemcpyDeviceToHost, t->stream));
    CU(t, cudaStreamSynchronize(t->stream));
    memcpy(buf, h, nbytes);
  }
  return 0;
}

}  // namespace

extern "C" int mb_nccl_unique_id(uint8_t out[128]) {
  std::string err;
  NcclApi *api = nccl_api(&err);
  if (!api || !out) return mb_fail(nullptr, MB_ERR_KERNEL, "mb_nccl_unique_id: " + (api ? std::string("NULL argument") : err));
  NcclUniqueId id;
  NcclResult r = api->GetUniqueId(&id);
  if (r != 0) return mb_fail(nullptr, MB_ERR_KERNEL, std::string("ncclGetUniqueId: ") + api->GetErrorString(r));
  memcpy(out, id.internal, 128);
  return MB_OK;
}

extern "C" int mb_nccl_transport_create(mb_ctx *ctx, const uint8_t unique_id[128], int rank, int nranks, const char *const *role_names,
                                        mb_nccl_transport **out) {
  if (!ctx || !unique_id || !role_names || !out || rank < 0 || rank >= nranks) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_nccl_transport_create: bad argument");
  std::string err;
  NcclApi *api = nccl_api(&err);
  if (!api) return mb_fail(ctx, MB_ERR_KERNEL, "mb_nccl_transport_create: " + err);
  MB_CUDA(ctx, cudaSetDevice(ctx->device));
  mb_nccl_transport *t = new mb_nccl_transport();
  t->ctx = ctx;
  t->api = api;
  t->rank = rank;
  t->nranks = nranks;
  t->next_slot = 0;
  t->completed = 0;
  t->comm = nullptr;
  t->stream = nullptr;
  t->host_slots = t->dev_slots = nullptr;
  for (int i = 0; i < SLOTS; i++) { t->slot_done[i] = nullptr; t->slot_busy[i] = false; }
  for (int i = 0; i < nranks; i++) t->rank_of[role_names[i] ? role_names[i] : ""] = i;
  NcclUniqueId id;
  memcpy(id.internal, unique_id, 128);
  NcclResult r = api->CommInitRank(&t->comm, nranks, id, rank);
  if (r != 0) {
    delete t;
    return mb_fail(ctx, MB_ERR_KERNEL, std::string("ncclCommInitRank: ") + api->GetErrorString(r));
  }
  cudaError_t e = cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMallocHost((void **)&t->host_slots, SLOTS * SLOT_BYTES);
  if (e == cudaSuccess) e = cudaMalloc((void **)&t->dev_slots, SLOTS * SLOT_BYTES);
  for (int i = 0; i < SLOTS && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&t->slot_done[i], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    mb_nccl_transport_free(t);
    return mb_fail(ctx, MB_ERR_CUDA, std::string("mb_nccl_transport_create: ") + cudaGetErrorString(e));
  }
  *out = t;
  return MB_OK;
}

extern "C" mb_transport_fn mb_nccl_transport_fn(void) { return nccl_transport_cb; }

extern "C" int mb_nccl_transport_free(mb_nccl_transport *t) {
  if (!t) return MB_OK;
  cudaSetDevice(t->ctx->device);
  if (t->stream) cudaStreamSynchronize(t->stream);
  for (auto ev : t->posted) cudaEventDestroy(ev);
  for (auto ev : t->pool) cudaEventDestroy(ev);
  for (int i = 0; i < SLOTS; i++)
    if (t->slot_done[i]) cudaEventDestroy(t->slot_done[i]);
  if (t->comm) t->api->CommDestroy(t->comm);
  if (t->host_slots) cudaFreeHost(t->host_slots);
  if (t->dev_slots) cudaFree(t->dev_slots);
  if (t->stream) cudaStreamDestroy(t->stream);
  delete t;
  return MB_OK;
}
