// Native launch sequences: a whole step's host-side CUDA calls issued by ONE call from Python.
//
// The single-GPU step of the reference model is one CUDA graph of ~50 us; around it the public
// per-step API (`Trainer.train_batch`) has to copy the batch host->device, order the copy
// against the previous replay, launch the graph and queue the loss read-back -- a dozen runtime
// calls.  Issued one by one from Python they cost about as much host time as the graph costs
// device time, so any scheduling hiccup on the host idles the GPU (and, with data-parallel
// replicas that meet in an in-kernel all-reduce every step, a hiccup on ONE rank idles ALL of
// them).  A sequence is built once (array of primitive ops below); per step the caller patches the
// host source pointers and makes one call.
//
// The reference has nothing comparable: its loop blocks in an RPC per statement
// (/root/reference/simple_distributed.py:108-113).
#include "sdml_host.h"

extern "C" {

enum SdmlLaunchKind {
  SDML_LAUNCH_WAIT_EVENT = 0,    // stream waits for event            (a = stream, b = event)
  SDML_LAUNCH_RECORD_EVENT = 1,  // record event on stream            (a = stream, b = event)
  SDML_LAUNCH_MEMCPY = 2,        // cudaMemcpyAsync(dst, src, bytes)  (a = stream, b = dst, c = src, bytes)
  SDML_LAUNCH_GRAPH = 3,         // cudaGraphLaunch(exec, stream)     (a = stream, b = graph exec)
};

struct SdmlLaunchOp {
  int kind;
  int flags;            // bit 0: skip this op (lets a caller disable e.g. the first WAIT without rebuilding)
  void* a;
  void* b;
  void* c;
  unsigned long long bytes;
};

// Events for sequences (timing disabled: they only order streams).  Created / destroyed here so
// that Python never has to hold a torch.cuda.Event per step.
int sdml_event_create(void** event_out) {
  if (!event_out) return sdml::set_error("sdml_event_create: null output");
  cudaEvent_t ev;
  SDML_CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  *event_out = ev;
  return 0;
}

int sdml_event_destroy(void* event) {
  if (event) SDML_CUDA_OK(cudaEventDestroy(static_cast<cudaEvent_t>(event)));
  return 0;
}

// 0 = complete, 1 = not yet, negative = error.
int sdml_event_query(void* event) {
  const cudaError_t q = cudaEventQuery(static_cast<cudaEvent_t>(event));
  if (q == cudaSuccess) return 0;
  if (q == cudaErrorNotReady) return 1;
  return sdml::set_cuda_error(q, "cudaEventQuery", __FILE__, __LINE__);
}

int sdml_event_synchronize(void* event) {
  SDML_CUDA_OK(cudaEventSynchronize(static_cast<cudaEvent_t>(event)));
  return 0;
}

// Issue `n_ops` primitive ops in order.  Stops at the first failure.
int sdml_launch_seq(const SdmlLaunchOp* ops, int n_ops) {
  if (!ops || n_ops < 0) return sdml::set_error("sdml_launch_seq: bad arguments");
  for (int i = 0; i < n_ops; ++i) {
    const SdmlLaunchOp& op = ops[i];
    if (op.flags & 1) continue;
    cudaStream_t stream = static_cast<cudaStream_t>(op.a);
    switch (op.kind) {
      case SDML_LAUNCH_WAIT_EVENT:
        SDML_CUDA_OK(cudaStreamWaitEvent(stream, static_cast<cudaEvent_t>(op.b), 0));
        break;
      case SDML_LAUNCH_RECORD_EVENT:
        SDML_CUDA_OK(cudaEventRecord(static_cast<cudaEvent_t>(op.b), stream));
        break;
      case SDML_LAUNCH_MEMCPY:
        SDML_CUDA_OK(cudaMemcpyAsync(op.b, op.c, op.bytes, cudaMemcpyDefault, stream));
        break;
      case SDML_LAUNCH_GRAPH:
        SDML_CUDA_OK(cudaGraphLaunch(static_cast<cudaGraphExec_t>(op.b), stream));
        break;
      default:
        return sdml::set_error("sdml_launch_seq: unknown op kind");
    }
  }
  return 0;
}

}  // extern "C"
