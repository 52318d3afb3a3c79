// Native stage-boundary transport: NCCL point-to-point on a dedicated side stream per
// communicator, ordered against the compute stream with CUDA events only (the host never
// blocks in the hot loop).
//
// The reference moves the stage-boundary activation with a blocking TensorPipe RPC
// (/root/reference/simple_distributed.py:47-49: `.to_here()` on an RRef, payload pickled and
// staged through host memory, ref:80 `.cpu()`), and the gradient through dist-autograd's own
// RPCs (ref:112).  Here one `P2PChannel` = one NCCL communicator of two ranks for ONE direction
// of ONE stage boundary, so a receive that is posted early can never sit in front of a send the
// neighbour is waiting for, and activations never queue behind gradients.
//
//   send:  side stream waits for an event recorded on the producer (compute) stream
//          -> ncclSend -> completion event (ticket)
//   recv:  ncclRecv on the side stream -> completion event (ticket); the consumer stream waits
//          for the ticket with cudaStreamWaitEvent right before the kernel that reads the buffer
//
// NCCL is resolved at run time from the libnccl already loaded into the process (PyTorch's),
// falling back to the system library; nothing links against it at build time.
//
// Failure handling (the reference disables every timeout, ref:36 / ref:167, and hangs forever
// on a dead peer): `sdml_p2p_wait_host` polls the ticket and the communicator's asynchronous
// error state with a deadline and reports which it was.
#include "sdml_host.h"

#include <dlfcn.h>
#include <nccl.h>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>

namespace sdml {
namespace {

struct NcclApi {
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclCommAbort) CommAbort = nullptr;
  decltype(&ncclSend) Send = nullptr;
  decltype(&ncclRecv) Recv = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclCommGetAsyncError) CommGetAsyncError = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
  bool ok = false;
};

NcclApi& nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    // prefer the copy that is already mapped (torch's bundled NCCL): two NCCL instances in one
    // process would each bring their own proxy threads and buffers
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
#define SDML_NCCL_SYM(name)                                                        \
  api.name = reinterpret_cast<decltype(api.name)>(dlsym(h, "nccl" #name));          \
  if (!api.name) return;
    SDML_NCCL_SYM(GetUniqueId)
    SDML_NCCL_SYM(CommInitRank)
    SDML_NCCL_SYM(CommDestroy)
    SDML_NCCL_SYM(CommAbort)
    SDML_NCCL_SYM(Send)
    SDML_NCCL_SYM(Recv)
    SDML_NCCL_SYM(GroupStart)
    SDML_NCCL_SYM(GroupEnd)
    SDML_NCCL_SYM(GetErrorString)
    SDML_NCCL_SYM(CommGetAsyncError)
    SDML_NCCL_SYM(GetVersion)
#undef SDML_NCCL_SYM
    api.ok = true;
  });
  return api;
}

int nccl_error(ncclResult_t r, const char* what) {
  char buf[256];
  snprintf(buf, sizeof(buf), "%s: NCCL error %d (%s)", what, static_cast<int>(r),
           nccl().GetErrorString ? nccl().GetErrorString(r) : "?");
  return set_error(buf);
}

#define SDML_NCCL_OK(expr)                                  \
  do {                                                      \
    ncclResult_t _r = (expr);                               \
    if (_r != ncclSuccess) return nccl_error(_r, #expr);    \
  } while (0)

constexpr int kTickets = 256;  // completion events in flight per channel (ring)

struct P2PChannel {
  ncclComm_t comm = nullptr;
  int nranks = 0, rank = 0, device = 0;
  cudaStream_t side = nullptr;
  cudaEvent_t ready = nullptr;       // producer -> side stream
  cudaEvent_t done[kTickets];        // side stream -> consumer, indexed by ticket % kTickets
  long long next_ticket = 0;
};

}  // namespace
}  // namespace sdml

using namespace sdml;

struct SdmlP2POp {
  int is_send;            // 1 = send, 0 = receive
  int peer;               // rank inside this channel's communicator
  void* ptr;
  unsigned long long bytes;
};

// NCCL version the engine is bound to (e.g. 22809), or -1 when no libnccl could be loaded.
extern "C" int sdml_p2p_nccl_version() {
  if (!nccl().ok) return -1;
  int v = 0;
  if (nccl().GetVersion(&v) != ncclSuccess) return -1;
  return v;
}

// 128-byte rendezvous token; create it on one rank and hand it to the others out of band.
extern "C" int sdml_p2p_unique_id(void* out128) {
  if (!nccl().ok) return set_error("sdml_p2p: libnccl.so.2 not found");
  ncclUniqueId id;
  SDML_NCCL_OK(nccl().GetUniqueId(&id));
  static_assert(sizeof(id) == 128, "ncclUniqueId size");
  memcpy(out128, &id, sizeof(id));
  return 0;
}

// Collective over the `nranks` ranks that share `uid128`: builds the communicator on the CURRENT
// device plus the channel's side stream and events.
extern "C" int sdml_p2p_create(const void* uid128, int nranks, int rank, void** handle) {
  if (!nccl().ok) return set_error("sdml_p2p: libnccl.so.2 not found");
  if (!uid128 || !handle || nranks < 1 || rank < 0 || rank >= nranks)
    return set_error("sdml_p2p_create: bad arguments");
  auto* ch = new P2PChannel();
  ch->nranks = nranks;
  ch->rank = rank;
  if (cudaGetDevice(&ch->device) != cudaSuccess) {
    delete ch;
    return set_error("sdml_p2p_create: cudaGetDevice failed");
  }
  ncclUniqueId id;
  memcpy(&id, uid128, sizeof(id));
  ncclResult_t r = nccl().CommInitRank(&ch->comm, nranks, id, rank);
  if (r != ncclSuccess) {
    delete ch;
    return nccl_error(r, "ncclCommInitRank");
  }
  // transfers are on the critical path of the neighbour stage: highest priority
  int lo = 0, hi = 0;
  cudaError_t e = cudaDeviceGetStreamPriorityRange(&lo, &hi);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&ch->side, cudaStreamNonBlocking, hi);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ch->ready, cudaEventDisableTiming);
  int made = 0;
  for (; e == cudaSuccess && made < kTickets; ++made)
    e = cudaEventCreateWithFlags(&ch->done[made], cudaEventDisableTiming);
  if (e != cudaSuccess) {  // give back what was created: a half-built channel is never returned
    for (int i = 0; i + 1 < made; ++i) cudaEventDestroy(ch->done[i]);
    if (ch->ready) cudaEventDestroy(ch->ready);
    if (ch->side) cudaStreamDestroy(ch->side);
    nccl().CommAbort(ch->comm);
    delete ch;
    return set_cuda_error(e, "sdml_p2p_create: stream / event set-up", __FILE__, __LINE__);
  }
  *handle = ch;
  return 0;
}

extern "C" int sdml_p2p_destroy(void* handle, int abort) {
  auto* ch = static_cast<P2PChannel*>(handle);
  if (!ch) return 0;
  cudaSetDevice(ch->device);
  if (abort) {
    nccl().CommAbort(ch->comm);   // peer is gone: do not wait for outstanding operations
  } else {
    cudaStreamSynchronize(ch->side);
    nccl().CommDestroy(ch->comm);
  }
  cudaEventDestroy(ch->ready);
  for (int i = 0; i < kTickets; ++i) cudaEventDestroy(ch->done[i]);
  cudaStreamDestroy(ch->side);
  delete ch;
  return 0;
}

// Enqueue `n_ops` transfers as ONE NCCL group on the channel's side stream.  With `order_after`
// != 0 the group starts only after everything queued on `after_stream` so far (the kernels that
// produce the send buffers); a separate flag because the legacy default stream's handle IS 0.
// Returns a ticket through `ticket_out`; the transfer is complete -- receive buffers readable,
// send buffers reusable -- once the ticket's event fires.
extern "C" int sdml_p2p_enqueue(void* handle, const SdmlP2POp* ops, int n_ops, void* after_stream,
                                int order_after, long long* ticket_out) {
  auto* ch = static_cast<P2PChannel*>(handle);
  if (!ch || !ops || n_ops < 1) return set_error("sdml_p2p_enqueue: bad arguments");
  for (int i = 0; i < n_ops; ++i)
    if (ops[i].peer < 0 || ops[i].peer >= ch->nranks || !ops[i].ptr || ops[i].bytes == 0)
      return set_error("sdml_p2p_enqueue: bad op (peer out of range, null buffer or zero bytes)");
  if (order_after) {
    SDML_CUDA_OK(cudaEventRecord(ch->ready, static_cast<cudaStream_t>(after_stream)));
    SDML_CUDA_OK(cudaStreamWaitEvent(ch->side, ch->ready, 0));
  }
  const bool grouped = n_ops > 1;
  if (grouped) SDML_NCCL_OK(nccl().GroupStart());
  for (int i = 0; i < n_ops; ++i) {
    if (ops[i].is_send)
      SDML_NCCL_OK(nccl().Send(ops[i].ptr, ops[i].bytes, ncclInt8, ops[i].peer, ch->comm, ch->side));
    else
      SDML_NCCL_OK(nccl().Recv(ops[i].ptr, ops[i].bytes, ncclInt8, ops[i].peer, ch->comm, ch->side));
  }
  if (grouped) SDML_NCCL_OK(nccl().GroupEnd());
  const long long t = ch->next_ticket++;
  SDML_CUDA_OK(cudaEventRecord(ch->done[t % kTickets], ch->side));
  if (ticket_out) *ticket_out = t;
  return 0;
}

static int ticket_event(P2PChannel* ch, long long ticket, cudaEvent_t* ev) {
  if (!ch || ticket < 0 || ticket >= ch->next_ticket) return set_error("sdml_p2p: unknown ticket");
  if (ch->next_ticket - ticket > kTickets)
    return set_error("sdml_p2p: ticket too old (more than 256 transfers were enqueued since)");
  *ev = ch->done[ticket % kTickets];
  return 0;
}

// Make `stream` wait (on the device) for the transfer behind `ticket`.  The host does not block.
extern "C" int sdml_p2p_wait_stream(void* handle, long long ticket, void* stream) {
  auto* ch = static_cast<P2PChannel*>(handle);
  cudaEvent_t ev;
  if (int rc = ticket_event(ch, ticket, &ev)) return rc;
  SDML_CUDA_OK(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), ev, 0));
  return 0;
}

// Host-side bounded wait: 0 = complete, 1 = still pending after timeout_ms, -1 = NCCL / CUDA error
// (message through sdml_last_error()).  timeout_ms == 0 polls once.
extern "C" int sdml_p2p_wait_host(void* handle, long long ticket, int timeout_ms) {
  auto* ch = static_cast<P2PChannel*>(handle);
  cudaEvent_t ev;
  if (int rc = ticket_event(ch, ticket, &ev)) return rc;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  for (;;) {
    const cudaError_t q = cudaEventQuery(ev);
    if (q == cudaSuccess) return 0;
    if (q != cudaErrorNotReady) return set_cuda_error(q, "cudaEventQuery", __FILE__, __LINE__);
    ncclResult_t async = ncclSuccess;
    const ncclResult_t r = nccl().CommGetAsyncError(ch->comm, &async);
    if (r != ncclSuccess) return nccl_error(r, "ncclCommGetAsyncError");
    if (async != ncclSuccess && async != ncclInProgress) return nccl_error(async, "asynchronous communicator error");
    if (std::chrono::steady_clock::now() >= deadline) return 1;
    std::this_thread::sleep_for(std::chrono::microseconds(50));
  }
}
