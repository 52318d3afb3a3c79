// Native op-list executor: walks one rank's schedule (F / B / W / BW ops from
// csrc/schedule.cpp) for one training step without returning to Python.
//
// The reference's hot loop is four Python statements per batch
// (/root/reference/simple_distributed.py:108-113: forward -> rpc_sync -> dist_autograd.backward
// -> dist_optim.step), each of which blocks on an RPC.  Here a step is a list of ops per rank;
// this file is the loop that runs it:
//
//   * `run_step` is the control flow, written once against a small table of hooks
//     (compute / post_recv / wait_recv / send / wait_send / end_step).  It is the same
//     discipline as the Python executor (`parallel/pipeline.py`): a receive is posted as soon
//     as the op before its consumer has been LAUNCHED (`lookahead` = 0 on GPUs, see the
//     dead-lock note there), sends never block the compute stream, all sends are waited for at
//     the end of the step because their buffers are reused by the next one.
//   * `sdml_exec_run_hooks` exposes it with caller-supplied hooks (used by the CPU tests to
//     check the order of actions against the Python executor).
//   * `ExecPlan` is the GPU binding: every op is a pre-captured CUDA graph (a handful of kernel
//     launches each; `cudaGraphLaunch` replaces the Python-side launch code), transfers go
//     through the native NCCL engine (`csrc/p2p_engine.cu`), ordering is CUDA events only.
//     One call = one step; the host never blocks in it.
#include "sdml_host.h"

#include <cstdio>
#include <vector>

extern "C" {
struct SdmlP2POp {
  int is_send;
  int peer;
  void* ptr;
  unsigned long long bytes;
};
int sdml_p2p_enqueue(void* handle, const SdmlP2POp* ops, int n_ops, void* after_stream,
                     int order_after, long long* ticket_out);
int sdml_p2p_wait_stream(void* handle, long long ticket, void* stream);
}

namespace sdml {
namespace {

enum { OP_F = 0, OP_B = 1, OP_W = 2, OP_BW = 3 };
enum { DIR_ACT = 0, DIR_GRAD = 1 };

struct Hooks {
  void* user;
  int (*compute)(void* user, int kind, int mb);
  int (*post_recv)(void* user, int dir, int mb, long long* ticket);
  int (*wait_recv)(void* user, int dir, int mb, long long ticket);
  int (*send)(void* user, int dir, int mb, long long* ticket);
  int (*wait_send)(void* user, int dir, int mb, long long ticket);
  int (*end_step)(void* user);
};

#define SDML_HOOK_OK(expr)     \
  do {                         \
    const int _rc = (expr);    \
    if (_rc != 0) return _rc;  \
  } while (0)

int run_step(const int* ops, int n_ops, bool is_first, bool is_last, int lookahead, const Hooks& h) {
  if (!ops || n_ops <= 0) return set_error("sdml_exec: empty op list");
  struct Pending {
    int op_index, dir, mb;
    long long ticket;
  };
  std::vector<int> needs_recv;  // indices of ops that consume a transfer
  for (int i = 0; i < n_ops; ++i) {
    const int kind = ops[2 * i];
    if (kind < OP_F || kind > OP_BW) return set_error("sdml_exec: unknown op kind");
    if (kind == OP_F && !is_first) needs_recv.push_back(i);
    if ((kind == OP_B || kind == OP_BW) && !is_last) needs_recv.push_back(i);
  }
  std::vector<Pending> posted;  // receives in flight, in posting order (= consumption order)
  std::vector<Pending> sends;
  size_t next_recv = 0, next_wait = 0;

  for (int i = 0; i < n_ops; ++i) {
    // post every receive whose consumer is at most `lookahead` ops away
    while (next_recv < needs_recv.size() && needs_recv[next_recv] <= i + lookahead) {
      const int j = needs_recv[next_recv++];
      Pending p{j, ops[2 * j] == OP_F ? DIR_ACT : DIR_GRAD, ops[2 * j + 1], -1};
      SDML_HOOK_OK(h.post_recv(h.user, p.dir, p.mb, &p.ticket));
      posted.push_back(p);
    }
    const int kind = ops[2 * i], mb = ops[2 * i + 1];
    const bool consumes = (kind == OP_F && !is_first) || ((kind == OP_B || kind == OP_BW) && !is_last);
    if (consumes) {
      if (next_wait >= posted.size() || posted[next_wait].op_index != i)
        return set_error("sdml_exec: internal error, receive of this op was not posted");
      const Pending& p = posted[next_wait++];
      SDML_HOOK_OK(h.wait_recv(h.user, p.dir, p.mb, p.ticket));
    }
    if (kind == OP_F) {
      SDML_HOOK_OK(h.compute(h.user, OP_F, mb));
      if (!is_last) {
        Pending s{i, DIR_ACT, mb, -1};
        SDML_HOOK_OK(h.send(h.user, DIR_ACT, mb, &s.ticket));
        sends.push_back(s);
      }
    } else if (kind == OP_B || kind == OP_BW) {
      SDML_HOOK_OK(h.compute(h.user, OP_B, mb));
      if (!is_first) {
        Pending s{i, DIR_GRAD, mb, -1};
        SDML_HOOK_OK(h.send(h.user, DIR_GRAD, mb, &s.ticket));
        sends.push_back(s);
      }
      if (kind == OP_BW) SDML_HOOK_OK(h.compute(h.user, OP_W, mb));
    } else {
      SDML_HOOK_OK(h.compute(h.user, OP_W, mb));
    }
  }
  SDML_HOOK_OK(h.end_step(h.user));
  for (const Pending& s : sends) SDML_HOOK_OK(h.wait_send(h.user, s.dir, s.mb, s.ticket));
  return 0;
}

// ---------------------------------------------------------------------------
// GPU binding: ops are CUDA graphs, transfers go through the p2p engine
// ---------------------------------------------------------------------------
struct ExecPlan {
  int n_mb = 0;
  bool is_first = false, is_last = false;
  int lookahead = 0;
  std::vector<int> ops;                       // (kind, mb) pairs
  std::vector<cudaGraphExec_t> graph[3];      // [F / B / W][mb]; nullptr = nothing to launch
  cudaGraphExec_t end_graph = nullptr;        // flush of pending wgrads + optimizer
  cudaStream_t stream = nullptr;              // compute stream (0 = the default stream)
  void* chan_recv[2] = {nullptr, nullptr};    // [DIR_ACT] from the previous stage, [DIR_GRAD] from the next
  void* chan_send[2] = {nullptr, nullptr};    // [DIR_ACT] to the next stage, [DIR_GRAD] to the previous
  std::vector<void*> recv_ptr[2], send_ptr[2];  // per micro-batch buffers
  std::vector<unsigned long long> recv_bytes[2], send_bytes[2];
  long long steps = 0;
};

int plan_compute(void* u, int kind, int mb) {
  auto* p = static_cast<ExecPlan*>(u);
  cudaGraphExec_t g = p->graph[kind][mb];
  if (g == nullptr) return 0;
  SDML_CUDA_OK(cudaGraphLaunch(g, p->stream));
  count_launch();
  return 0;
}
int plan_post_recv(void* u, int dir, int mb, long long* ticket) {
  auto* p = static_cast<ExecPlan*>(u);
  if (!p->chan_recv[dir] || !p->recv_ptr[dir][mb]) return set_error("sdml_exec: receive channel / buffer not set");
  // sender is rank 0 of its channel; ordered after the compute stream (the slot may still be read)
  const SdmlP2POp op{0, 0, p->recv_ptr[dir][mb], p->recv_bytes[dir][mb]};
  return sdml_p2p_enqueue(p->chan_recv[dir], &op, 1, p->stream, 1, ticket);
}
int plan_wait_recv(void* u, int dir, int, long long ticket) {
  auto* p = static_cast<ExecPlan*>(u);
  return sdml_p2p_wait_stream(p->chan_recv[dir], ticket, p->stream);
}
int plan_send(void* u, int dir, int mb, long long* ticket) {
  auto* p = static_cast<ExecPlan*>(u);
  if (!p->chan_send[dir] || !p->send_ptr[dir][mb]) return set_error("sdml_exec: send channel / buffer not set");
  const SdmlP2POp op{1, 1, p->send_ptr[dir][mb], p->send_bytes[dir][mb]};
  return sdml_p2p_enqueue(p->chan_send[dir], &op, 1, p->stream, 1, ticket);
}
int plan_wait_send(void* u, int dir, int, long long ticket) {
  auto* p = static_cast<ExecPlan*>(u);
  return sdml_p2p_wait_stream(p->chan_send[dir], ticket, p->stream);
}
int plan_end_step(void* u) {
  auto* p = static_cast<ExecPlan*>(u);
  if (p->end_graph) {
    SDML_CUDA_OK(cudaGraphLaunch(p->end_graph, p->stream));
    count_launch();
  }
  return 0;
}

}  // namespace
}  // namespace sdml

using namespace sdml;

struct SdmlExecHooks {
  void* user;
  int (*compute)(void* user, int kind, int mb);                        /* kind: 0 F, 1 B, 2 W */
  int (*post_recv)(void* user, int dir, int mb, long long* ticket);    /* dir: 0 activation, 1 gradient */
  int (*wait_recv)(void* user, int dir, int mb, long long ticket);
  int (*send)(void* user, int dir, int mb, long long* ticket);
  int (*wait_send)(void* user, int dir, int mb, long long ticket);
  int (*end_step)(void* user);
};

// Run one step of `ops` ((kind, mb) pairs) against caller-supplied hooks.  A hook returning
// non-zero aborts the step with that code.
extern "C" int sdml_exec_run_hooks(const int* ops, int n_ops, int is_first, int is_last, int lookahead,
                                   const SdmlExecHooks* hooks) {
  if (!hooks || !hooks->compute || !hooks->post_recv || !hooks->wait_recv || !hooks->send ||
      !hooks->wait_send || !hooks->end_step)
    return set_error("sdml_exec_run_hooks: every hook must be set");
  Hooks h{hooks->user, hooks->compute, hooks->post_recv, hooks->wait_recv,
          hooks->send, hooks->wait_send, hooks->end_step};
  return run_step(ops, n_ops, is_first != 0, is_last != 0, lookahead, h);
}

extern "C" int sdml_exec_plan_create(const int* ops, int n_ops, int n_mb, int is_first, int is_last,
                                     void* compute_stream, void** plan_out) {
  if (!ops || n_ops <= 0 || n_mb <= 0 || !plan_out) return set_error("sdml_exec_plan_create: bad arguments");
  for (int i = 0; i < n_ops; ++i)
    if (ops[2 * i] < OP_F || ops[2 * i] > OP_BW || ops[2 * i + 1] < 0 || ops[2 * i + 1] >= n_mb)
      return set_error("sdml_exec_plan_create: op out of range");
  auto* p = new ExecPlan();
  p->n_mb = n_mb;
  p->is_first = is_first != 0;
  p->is_last = is_last != 0;
  p->ops.assign(ops, ops + 2 * n_ops);
  for (auto& g : p->graph) g.assign(n_mb, nullptr);
  for (int d = 0; d < 2; ++d) {
    p->recv_ptr[d].assign(n_mb, nullptr);
    p->send_ptr[d].assign(n_mb, nullptr);
    p->recv_bytes[d].assign(n_mb, 0);
    p->send_bytes[d].assign(n_mb, 0);
  }
  p->stream = static_cast<cudaStream_t>(compute_stream);
  *plan_out = p;
  return 0;
}

// kind 0 / 1 / 2 = F / B / W of micro-batch `mb`; kind 3 = the end-of-step graph (mb ignored).
// `graph_exec` is a cudaGraphExec_t that stays owned by the caller (torch.cuda.CUDAGraph).
extern "C" int sdml_exec_plan_set_graph(void* plan, int kind, int mb, void* graph_exec) {
  auto* p = static_cast<ExecPlan*>(plan);
  if (!p) return set_error("sdml_exec_plan_set_graph: null plan");
  if (kind == 3) {
    p->end_graph = static_cast<cudaGraphExec_t>(graph_exec);
    return 0;
  }
  if (kind < 0 || kind > 2 || mb < 0 || mb >= p->n_mb) return set_error("sdml_exec_plan_set_graph: out of range");
  p->graph[kind][mb] = static_cast<cudaGraphExec_t>(graph_exec);
  return 0;
}

// dir 0 = activations, 1 = gradients; is_send selects the outgoing / incoming side.
extern "C" int sdml_exec_plan_set_channel(void* plan, int dir, int is_send, void* p2p_handle) {
  auto* p = static_cast<ExecPlan*>(plan);
  if (!p || dir < 0 || dir > 1) return set_error("sdml_exec_plan_set_channel: bad arguments");
  (is_send ? p->chan_send : p->chan_recv)[dir] = p2p_handle;
  return 0;
}

extern "C" int sdml_exec_plan_set_buffer(void* plan, int dir, int is_send, int mb, void* ptr,
                                         unsigned long long bytes) {
  auto* p = static_cast<ExecPlan*>(plan);
  if (!p || dir < 0 || dir > 1 || mb < 0 || mb >= p->n_mb || !ptr || bytes == 0)
    return set_error("sdml_exec_plan_set_buffer: bad arguments");
  (is_send ? p->send_ptr : p->recv_ptr)[dir][mb] = ptr;
  (is_send ? p->send_bytes : p->recv_bytes)[dir][mb] = bytes;
  return 0;
}

// One training step: every graph launch, receive, wait and send of this rank's op list.
extern "C" int sdml_exec_plan_step(void* plan) {
  auto* p = static_cast<ExecPlan*>(plan);
  if (!p) return set_error("sdml_exec_plan_step: null plan");
  Hooks h{p, plan_compute, plan_post_recv, plan_wait_recv, plan_send, plan_wait_send, plan_end_step};
  const int rc = run_step(p->ops.data(), static_cast<int>(p->ops.size() / 2), p->is_first, p->is_last,
                          p->lookahead, h);
  if (rc == 0) ++p->steps;
  return rc;
}

extern "C" int sdml_exec_plan_destroy(void* plan) {
  delete static_cast<ExecPlan*>(plan);
  return 0;
}
