// Stage-boundary transport over symmetric (peer-mapped) memory: push + flag, no NCCL.
//
// The reference moves an activation from rank 0 to rank 1 and the gradient back with one blocking
// TensorPipe RPC each (/root/reference/simple_distributed.py:47-49 `rpc_sync`, :71/:80 the
// `.cpu()` round trip, :112 the backward RPC).  Here a boundary tensor lands in a slot that
// lives in the RECEIVER's memory and is mapped into the sender's address space
// (csrc/symm.cu): the producer writes the slot over NVLink -- either the producing GEMM's TMA
// store epilogue targets the slot directly, or one of the copy kernels below moves it -- and a
// 4-byte flag published with a system-scope release tells the consumer that the slot is
// complete.  No receive is ever posted, no NCCL kernel runs, and every piece is an ordinary
// stream-ordered kernel, so a whole training step (including its transfers) can be captured
// into ONE CUDA graph per rank.
//
// Protocol of one direction (sender S -> receiver R), counted in "rounds" (a training step or an
// evaluation batch).  Both sides keep a device-resident round counter (`cnt`, uint32, number of
// completed rounds); a captured graph therefore replays correctly.
//
//   S: [first write of a round]  wait   ack_S        >= cnt_S        (R has finished round cnt_S-1:
//                                                                     its slots may be overwritten)
//      write slot i (peer memory), then  signal flag_R[i] = cnt_S + 1
//      [end of round]            cnt_S += 1
//   R: before consuming slot i:  wait   flag_R[i]    >= cnt_R + 1
//      [end of round, after the last reader was enqueued]
//                                signal ack_S = cnt_R + 1 ;  cnt_R += 1
//
// Comparisons are on the signed 32-bit difference, so the counters may wrap.  Every wait is
// bounded: after `timeout_ns` it writes an error record into a host-visible word and returns
// (later waits return at once), and the host raises at its next synchronisation point -- the
// reference runs with every timeout disabled (ref:36, ref:167) and hangs forever if the peer dies.
#include "sdml_host.h"
#include "sdml_ptx.cuh"

namespace sdml {

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Error record of a timed-out wait: {code, flag value seen, value wanted, 0}, 4 x uint32.
// `dev` lives in DEVICE memory and is read at the start of every wait, so a dead peer costs one
// timeout, not one per wait.  `host` (optional) is pinned host memory that receives the same
// record, so the host can poll it without touching the device.
struct BndErr {
  uint32_t* dev;
  uint32_t* host;
};

__device__ __forceinline__ bool bnd_spin(const uint32_t* flag, uint32_t want,
                                         unsigned long long timeout_ns, BndErr e, uint32_t code) {
  uint32_t* err = e.dev;
  if (err != nullptr && *reinterpret_cast<volatile uint32_t*>(err) != 0u) return false;  // already gave up
  if (static_cast<int32_t>(ld_acquire_sys(flag) - want) >= 0) return true;
  const unsigned long long t0 = global_timer_ns();
  uint32_t spins = 0;
  for (;;) {
    const uint32_t v = ld_acquire_sys(flag);
    if (static_cast<int32_t>(v - want) >= 0) return true;
    if (((++spins) & 63u) == 0u) {
      if (timeout_ns != 0ull && global_timer_ns() - t0 > timeout_ns) {
        if (err != nullptr) {
          err[1] = v;
          err[2] = want;
          err[0] = code;
        }
        if (e.host != nullptr) {
          e.host[1] = v;
          e.host[2] = want;
          __threadfence_system();
          e.host[0] = code;
          __threadfence_system();
        }
        return false;
      }
      __nanosleep(64);
    }
  }
}

__global__ void bnd_wait_kernel(const uint32_t* flag, const uint32_t* cnt, uint32_t add,
                                unsigned long long timeout_ns, BndErr err, uint32_t code) {
  bnd_spin(flag, *cnt + add, timeout_ns, err, code);
}

__global__ void bnd_signal_kernel(uint32_t* flag_peer, uint32_t* cnt, uint32_t add, uint32_t bump) {
  const uint32_t c = *cnt;
  if (flag_peer != nullptr) {
    __threadfence_system();
    st_release_sys(flag_peer, c + add);
  }
  if (bump) *cnt = c + bump;
}

// Small messages (<= 128 KiB): ONE CTA copies the tensor into the peer's slot with 16-byte stores
// and publishes the flag -- one launch per transfer, nothing else on the wire.  Optionally waits
// first for the receiver's acknowledgement of the previous round (slot reuse).
__global__ void __launch_bounds__(1024)
bnd_push_small_kernel(uint4* __restrict__ dst_peer, const uint4* __restrict__ src, int n16,
                      uint32_t* flag_peer, const uint32_t* cnt, uint32_t add, const uint32_t* ack,
                      unsigned long long timeout_ns, BndErr err, uint32_t code) {
  if (ack != nullptr) {
    __shared__ int ok;
    if (threadIdx.x == 0) ok = bnd_spin(ack, *cnt, timeout_ns, err, code) ? 1 : 0;
    __syncthreads();
    if (!ok) return;
  }
  for (int i = threadIdx.x; i < n16; i += blockDim.x) dst_peer[i] = src[i];
  // The barrier orders every thread's stores before thread 0's fence, and a fence is cumulative:
  // one system-scope fence + release store publishes the whole slot (a fence in EVERY thread cost
  // ~8 us per push under ncu, most of this kernel).
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    st_release_sys(flag_peer, *cnt + add);
  }
}

}  // namespace sdml

using namespace sdml;

// Stream-ordered wait: the stream does not proceed until *flag >= *cnt + add (signed 32-bit
// difference), or until timeout_ns has passed (0 = unbounded), in which case {code, seen, wanted}
// is written to err_dev (4 x uint32, device memory, must start zeroed; a non-zero first word makes
// every later wait return at once) and to err_host (4 x uint32 pinned host memory, optional), and
// the stream continues.
extern "C" int sdml_bnd_wait(const void* flag, const void* cnt, unsigned int add,
                             unsigned long long timeout_ns, void* err_dev, void* err_host,
                             unsigned int code, cudaStream_t stream) {
  const BndErr err{static_cast<uint32_t*>(err_dev), static_cast<uint32_t*>(err_host)};
  if (!flag || !cnt) return set_error("sdml_bnd_wait: null flag / counter");
  bnd_wait_kernel<<<1, 1, 0, stream>>>(static_cast<const uint32_t*>(flag), static_cast<const uint32_t*>(cnt),
                                       add, timeout_ns, err, code);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// Stream-ordered signal: everything enqueued on `stream` before this call (including writes to
// peer memory) is visible to a peer that observes *flag_peer == *cnt + add.  flag_peer may be
// NULL (only bump the counter); bump is added to *cnt afterwards.
extern "C" int sdml_bnd_signal(void* flag_peer, void* cnt, unsigned int add, unsigned int bump,
                               cudaStream_t stream) {
  if (!cnt) return set_error("sdml_bnd_signal: null counter");
  bnd_signal_kernel<<<1, 1, 0, stream>>>(static_cast<uint32_t*>(flag_peer), static_cast<uint32_t*>(cnt), add, bump);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// Copy `bytes` (multiple of 16, both pointers 16-byte aligned) from local memory into the peer's
// slot and publish flag_peer = *cnt + add.  ack != NULL: first wait until *ack >= *cnt.
// Up to 128 KiB the copy and the flag are one single-CTA kernel; larger transfers use the copy
// engine (cudaMemcpyAsync device -> peer) followed by the signal kernel, so no SM is spent on them.
extern "C" int sdml_bnd_push(void* dst_peer, const void* src, long long bytes, void* flag_peer,
                             const void* cnt, unsigned int add, const void* ack,
                             unsigned long long timeout_ns, void* err_dev, void* err_host,
                             unsigned int code, cudaStream_t stream) {
  const BndErr err{static_cast<uint32_t*>(err_dev), static_cast<uint32_t*>(err_host)};
  if (!dst_peer || !src || bytes <= 0 || !flag_peer || !cnt) return set_error("sdml_bnd_push: bad arguments");
  if ((bytes & 15) || (reinterpret_cast<uintptr_t>(dst_peer) & 15) || (reinterpret_cast<uintptr_t>(src) & 15))
    return set_error("sdml_bnd_push: size and pointers must be multiples of 16 bytes");
  if (bytes <= (128 << 10)) {
    const int n16 = static_cast<int>(bytes / 16);
    int threads = ((n16 + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    if (threads < 32) threads = 32;
    bnd_push_small_kernel<<<1, threads, 0, stream>>>(
        static_cast<uint4*>(dst_peer), static_cast<const uint4*>(src), n16, static_cast<uint32_t*>(flag_peer),
        static_cast<const uint32_t*>(cnt), add, static_cast<const uint32_t*>(ack), timeout_ns, err, code);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
    return 0;
  }
  if (ack != nullptr) {
    bnd_wait_kernel<<<1, 1, 0, stream>>>(static_cast<const uint32_t*>(ack), static_cast<const uint32_t*>(cnt), 0u,
                                         timeout_ns, err, code);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
  }
  SDML_CUDA_OK(cudaMemcpyAsync(dst_peer, src, static_cast<size_t>(bytes), cudaMemcpyDeviceToDevice, stream));
  bnd_signal_kernel<<<1, 1, 0, stream>>>(static_cast<uint32_t*>(flag_peer),
                                         const_cast<uint32_t*>(static_cast<const uint32_t*>(cnt)), add, 0u);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}
