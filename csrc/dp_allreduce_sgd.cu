// Data-parallel gradient all-reduce FUSED with the optimizer step, for stages whose
// parameters are small enough that latency, not bandwidth, is the cost (the reference
// model: 5 280 parameters on stage 0, 16 560 on stage 1).  One kernel per step and rank:
//
//   phase 1  copy (and pre-scale) this rank's gradients into its symmetric staging
//            buffer; the last CTA to finish publishes "epoch e ready" into every peer's
//            flag array with a system-scope release store over NVLink;
//   phase 2  spin (acquire, system scope) until every peer has published epoch e, then
//            every rank sums all replicas' staged gradients straight out of the peers'
//            memory (NVSwitch P2P loads) and applies SGD-with-momentum to its own copy of
//            the parameters, zeroing the gradient accumulators for the next step.
//
// All replicas compute the same sum in the same order, so their weights stay bit-identical.
// No NCCL call, no separate optimizer launch.  Replaces, for the DP x PP hybrid, the
// reference's per-owner `DistributedOptimizer.step` RPC fan-out
// (/root/reference/simple_distributed.py:113) plus the all-reduce the reference never had.
//
// Staging is double-buffered by epoch parity: a rank can only be two epochs ahead of a
// peer that still reads its buffer after that peer has itself published the epoch in
// between, which it does only after finishing its reads.
#include "sdml_host.h"
#include "sdml_ptx.cuh"

#include <mutex>

namespace sdml {

constexpr int kDpMaxTensors = 32;
constexpr int kDpMaxRanks = 8;
constexpr int kDpThreads = 256;

struct DpTable {
  float* p[kDpMaxTensors];
  float* g[kDpMaxTensors];
  float* m[kDpMaxTensors];
  long long n[kDpMaxTensors];
  long long off[kDpMaxTensors];  // offset of the tensor in the flat staging buffer (floats)
  int count;
};

struct DpPeers {
  float* stage[kDpMaxRanks];      // staging buffer of rank r for THIS epoch parity (peer-mapped)
  float* stage_alt[kDpMaxRanks];  // device-epoch mode: buffers of the other parity (odd epochs use `stage`)
  const uint32_t* epoch_dev;      // device-epoch mode: *epoch_dev = steps completed so far (else null)
  uint32_t* flags[kDpMaxRanks];   // flag array of rank r: flags[r][src] = last epoch published by src
  uint32_t* local_counter;        // CTA completion counter (local memory)
  WaitGuard guard;                // deadline of the wait for the replicas' flags
  int rank, world;
};

__device__ __forceinline__ float ld_volatile_f32(const float* p) {
  float v;
  asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kDpThreads)
dp_allreduce_sgd_kernel(const __grid_constant__ DpTable t, const __grid_constant__ DpPeers pr,
                        uint32_t epoch, float lr, float momentum, float weight_decay,
                        float grad_scale) {
  const long long tid = static_cast<long long>(blockIdx.x) * kDpThreads + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * kDpThreads;
  // device-epoch mode (CUDA-graph replays): the step counter lives in device memory and is
  // advanced by a separate stream-ordered op after this kernel
  if (pr.epoch_dev) epoch = *pr.epoch_dev + 1u;
  float* const* stages = (pr.epoch_dev && !(epoch & 1u)) ? pr.stage_alt : pr.stage;
  float* my_stage = stages[pr.rank];

  // ---- phase 1: stage local gradients ----
  for (int i = 0; i < t.count; ++i) {
    const float* __restrict__ g = t.g[i];
    float* __restrict__ s = my_stage + t.off[i];
    for (long long k = tid; k < t.n[i]; k += stride) s[k] = g[k] * grad_scale;
  }
  // device-scope release is enough here: the CTA that turns out to be last acquires the counter
  // at device scope and then releases the flags at system scope (causality is transitive)
  __threadfence();
  __syncthreads();
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(pr.local_counter, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    // every CTA of this rank has staged and fenced: publish to all peers (and self)
    if (threadIdx.x < pr.world) {
      __threadfence_system();
      st_release_sys(pr.flags[threadIdx.x] + pr.rank, epoch);
    }
    if (threadIdx.x == 0) *pr.local_counter = 0;  // ready for the next launch
  }

  // ---- phase 2: wait for every replica, reduce, update ----
  if (threadIdx.x < pr.world) {
    const uint32_t* f = pr.flags[pr.rank] + threadIdx.x;
    spin_until_ge_sys(f, epoch, pr.guard, 15u);
  }
  __syncthreads();
  for (int i = 0; i < t.count; ++i) {
    float* __restrict__ p = t.p[i];
    float* __restrict__ g = t.g[i];
    float* __restrict__ m = t.m[i];
    const long long off = t.off[i];
    for (long long k = tid; k < t.n[i]; k += stride) {
      // issue the loads from every replica back to back (each is a ~2 us NVLink round trip);
      // only then consume them -- the latencies overlap instead of adding up
      float v[kDpMaxRanks];
#pragma unroll
      for (int r = 0; r < kDpMaxRanks; ++r)
        v[r] = (r < pr.world) ? ld_volatile_f32(stages[r] + off + k) : 0.f;
      float sum = 0.f;
#pragma unroll
      for (int r = 0; r < kDpMaxRanks; ++r) sum += v[r];   // fixed order: replicas stay bit-identical
      float pv = p[k];
      float d = sum + weight_decay * pv;
      if (momentum != 0.f) {
        const float mv = momentum * m[k] + d;
        m[k] = mv;
        d = mv;
      }
      p[k] = pv - lr * d;
      g[k] = 0.f;
    }
  }
}

}  // namespace sdml

using namespace sdml;

// params/grads/moms: HOST arrays of n_tensors device pointers (fp32); sizes: element counts.
// peer_stage / peer_flags: HOST arrays of `world` device pointers (entry `rank` is the local
// buffer, the others are peer-mapped): staging buffers for this epoch's parity and flag arrays
// (world uint32 each).  local_counter: one zero-initialised uint32 in local memory.
// grad_scale is applied before the sum (pass 1/world for the mean over replicas).
extern "C" int sdml_dp_allreduce_sgd(int n_tensors, void* const* params, void* const* grads,
                                     void* const* moms, const long long* sizes, int rank,
                                     int world, void* const* peer_stage, void* const* peer_flags,
                                     void* local_counter, unsigned int epoch, float lr,
                                     float momentum, float weight_decay, float grad_scale,
                                     cudaStream_t stream) {
  if (n_tensors <= 0 || n_tensors > kDpMaxTensors)
    return set_error("sdml_dp_allreduce_sgd: 1..32 tensors per call");
  if (world < 1 || world > kDpMaxRanks || rank < 0 || rank >= world)
    return set_error("sdml_dp_allreduce_sgd: world must be 1..8");
  DpTable t;
  long long total = 0;
  t.count = n_tensors;
  for (int i = 0; i < n_tensors; ++i) {
    t.p[i] = static_cast<float*>(params[i]);
    t.g[i] = static_cast<float*>(grads[i]);
    t.m[i] = static_cast<float*>(moms[i]);
    t.n[i] = sizes[i];
    t.off[i] = total;
    total += sizes[i];
  }
  DpPeers pr;
  for (int r = 0; r < world; ++r) {
    pr.stage[r] = static_cast<float*>(peer_stage[r]);
    pr.flags[r] = static_cast<uint32_t*>(peer_flags[r]);
  }
  pr.local_counter = static_cast<uint32_t*>(local_counter);
  pr.guard = WaitGuard{wait_guard().timeout_ns, wait_guard().err};
  pr.rank = rank;
  pr.world = world;
  pr.epoch_dev = nullptr;
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  // every CTA spins on peers: the whole grid must be co-resident (<= one CTA per SM)
  long long blocks = (total + kDpThreads * 4 - 1) / (kDpThreads * 4);
  if (blocks > num_sms(dev)) blocks = num_sms(dev);
  if (blocks < 1) blocks = 1;
  dp_allreduce_sgd_kernel<<<(int)blocks, kDpThreads, 0, stream>>>(t, pr, epoch, lr, momentum,
                                                                 weight_decay, grad_scale);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// CUDA-graph friendly form: the epoch is read from device memory (*epoch_dev = number of
// completed calls; the caller advances it with a stream-ordered increment AFTER this launch),
// so a captured launch can be replayed.  peer_stage_odd / peer_stage_even: HOST arrays of the
// `world` staging pointers used by odd / even epochs.
extern "C" int sdml_dp_allreduce_sgd_dev_epoch(int n_tensors, void* const* params,
                                               void* const* grads, void* const* moms,
                                               const long long* sizes, int rank, int world,
                                               void* const* peer_stage_odd,
                                               void* const* peer_stage_even,
                                               void* const* peer_flags, void* local_counter,
                                               const void* epoch_dev, float lr, float momentum,
                                               float weight_decay, float grad_scale,
                                               cudaStream_t stream) {
  if (n_tensors <= 0 || n_tensors > kDpMaxTensors)
    return set_error("sdml_dp_allreduce_sgd_dev_epoch: 1..32 tensors per call");
  if (world < 1 || world > kDpMaxRanks || rank < 0 || rank >= world || !epoch_dev)
    return set_error("sdml_dp_allreduce_sgd_dev_epoch: bad world / rank / epoch pointer");
  DpTable t;
  long long total = 0;
  t.count = n_tensors;
  for (int i = 0; i < n_tensors; ++i) {
    t.p[i] = static_cast<float*>(params[i]);
    t.g[i] = static_cast<float*>(grads[i]);
    t.m[i] = static_cast<float*>(moms[i]);
    t.n[i] = sizes[i];
    t.off[i] = total;
    total += sizes[i];
  }
  DpPeers pr;
  for (int r = 0; r < world; ++r) {
    pr.stage[r] = static_cast<float*>(peer_stage_odd[r]);
    pr.stage_alt[r] = static_cast<float*>(peer_stage_even[r]);
    pr.flags[r] = static_cast<uint32_t*>(peer_flags[r]);
  }
  pr.local_counter = static_cast<uint32_t*>(local_counter);
  pr.guard = WaitGuard{wait_guard().timeout_ns, wait_guard().err};
  pr.rank = rank;
  pr.world = world;
  pr.epoch_dev = static_cast<const uint32_t*>(epoch_dev);
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  long long blocks = (total + kDpThreads - 1) / kDpThreads;   // ~1 element per thread: latency-bound
  if (blocks > num_sms(dev)) blocks = num_sms(dev);
  if (blocks < 1) blocks = 1;
  dp_allreduce_sgd_kernel<<<(int)blocks, kDpThreads, 0, stream>>>(t, pr, 0u, lr, momentum,
                                                                 weight_decay, grad_scale);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// ============================================================================
// Low-latency variant: data and flag travel in the same 8-byte word
// ============================================================================
// The staged protocol above costs two NVLink round trips plus two system fences per step
// (measured 27 us per call at 2 GPUs, almost all of it latency).  Here every replica PUSHES its
// scaled gradients straight into a slot it owns in every peer's buffer as (value, epoch) pairs
// -- one 64-bit store each, so a pair can never be seen half-written -- and every replica then
// polls its OWN memory until all pairs of the current epoch have arrived.  One one-way NVLink
// trip, no fences, no flags, no counters.  Buffers alternate by epoch parity (a replica can be
// at most one step ahead of a peer: it needs that peer's pairs of the step in between).
// Layout per replica: [parity 2][source rank `world`][total] x {float value, uint32 epoch}.
namespace sdml {

struct DpLLPeers {
  unsigned long long* buf[kDpMaxRanks];  // base of rank r's landing buffer (peer-mapped)
  uint32_t* epoch_dev;                   // *epoch_dev = completed calls; advanced by the last CTA to finish
  uint32_t* exit_counter;                // CTAs that have finished this call (local memory, starts at 0)
  unsigned long long* stamps;            // optional (may be null): globaltimer probes, see below
  unsigned long long timeout_ns;         // 0 = spin without bound
  uint32_t* err;                         // optional: {code, peer, epoch, 0} left by a timed-out poll
  long long total;                       // elements per (parity, source) slot
  int rank, world;
};

__device__ __forceinline__ void st_pair(unsigned long long* p, float v, uint32_t e) {
  const unsigned long long w = (static_cast<unsigned long long>(e) << 32) | __float_as_uint(v);
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long ld_pair(const unsigned long long* p) {
  unsigned long long w;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  return w;
}
__device__ __forceinline__ unsigned long long dp_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// The parameter set is ONE flat index space (element k of tensor i sits at off[i] + k): every
// thread owns whole elements of it, first PUSHES all of them to every peer, and only then polls.
// (Walking the tensors one after the other, push + poll each, costs one NVLink round trip per
// TENSOR for the threads that own an element of each -- 8 serial round trips for the reference
// model, measured 18 us per call at 2 GPUs.)
__global__ void __launch_bounds__(kDpThreads)
dp_ll_allreduce_sgd_kernel(const __grid_constant__ DpTable t, const __grid_constant__ DpLLPeers pr,
                           float lr, float momentum, float weight_decay, float grad_scale) {
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(pr.epoch_dev) + 1u;
  const long long slot = static_cast<long long>(epoch & 1u) * pr.world * pr.total;
  const long long tid = static_cast<long long>(blockIdx.x) * kDpThreads + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * kDpThreads;
  unsigned long long t_enter = 0;
  if (pr.stamps != nullptr && threadIdx.x == 0) t_enter = dp_timer_ns();

  // ---- phase 1: push every element this thread owns into every peer's slot for source = me ----
  for (long long k = tid; k < pr.total; k += stride) {
    int i = 0;
    while (i + 1 < t.count && k >= t.off[i + 1]) ++i;
    const float mine = t.g[i][k - t.off[i]] * grad_scale;
    const long long at = slot + pr.rank * pr.total + k;
#pragma unroll
    for (int r = 0; r < kDpMaxRanks; ++r)
      if (r < pr.world && r != pr.rank) st_pair(pr.buf[r] + at, mine, epoch);
  }
  unsigned long long t_pushed = 0;
  if (pr.stamps != nullptr && threadIdx.x == 0) t_pushed = dp_timer_ns();

  // ---- phase 2: poll the local slots until this epoch's pairs are there, sum, update ----
  bool gave_up = false;
  for (long long k = tid; k < pr.total; k += stride) {
    int i = 0;
    while (i + 1 < t.count && k >= t.off[i + 1]) ++i;
    const long long j = k - t.off[i];
    float* __restrict__ g = t.g[i];
    const float mine = g[j] * grad_scale;
    const unsigned long long* local = pr.buf[pr.rank] + slot + k;
    float v[kDpMaxRanks];
#pragma unroll
    for (int r = 0; r < kDpMaxRanks; ++r) {
      v[r] = 0.f;
      if (r < pr.world) {
        if (r == pr.rank) {
          v[r] = mine;
        } else {
          unsigned long long w = ld_pair(local + r * pr.total);
          if (static_cast<uint32_t>(w >> 32) != epoch && !gave_up) {
            const unsigned long long t0 = dp_timer_ns();
            uint32_t spins = 0;
            do {
              w = ld_pair(local + r * pr.total);
              if (((++spins) & 1023u) == 0u && pr.timeout_ns != 0ull && dp_timer_ns() - t0 > pr.timeout_ns) {
                gave_up = true;      // dead or stalled replica: leave a record, finish the kernel
                if (pr.err != nullptr) {
                  pr.err[1] = static_cast<uint32_t>(r);
                  pr.err[2] = epoch;
                  __threadfence_system();
                  pr.err[0] = 7u;
                }
                break;
              }
            } while (static_cast<uint32_t>(w >> 32) != epoch);
          }
          v[r] = __uint_as_float(static_cast<uint32_t>(w));
        }
      }
    }
    float sum = 0.f;
#pragma unroll
    for (int r = 0; r < kDpMaxRanks; ++r) sum += v[r];  // fixed order: replicas stay bit-identical
    float* __restrict__ p = t.p[i];
    float* __restrict__ m = t.m[i];
    const float pv = p[j];
    float d = sum + weight_decay * pv;
    if (momentum != 0.f) {
      const float mv = momentum * m[j] + d;
      m[j] = mv;
      d = mv;
    }
    p[j] = pv - lr * d;
    g[j] = 0.f;
  }

  // ---- completion: the last CTA to finish advances the device-resident epoch ----
  // (every CTA has read the epoch before it can arrive here, and only the last arrival writes it)
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t_done = 0;
    if (pr.stamps != nullptr) {
      t_done = dp_timer_ns();
      unsigned long long* st = pr.stamps + 4 * (epoch & 63u);
      atomicMin(st + 0, t_enter);
      atomicMax(st + 1, t_pushed);
      atomicMax(st + 2, t_done);
    }
    __threadfence();
    const uint32_t prev = atomicAdd(pr.exit_counter, 1u);
    if (prev == gridDim.x - 1) {
      *pr.exit_counter = 0u;
      *pr.epoch_dev = epoch;
      if (pr.stamps != nullptr) {      // prepare the ring entry of the next call
        unsigned long long* nx = pr.stamps + 4 * ((epoch + 1u) & 63u);
        nx[0] = ~0ull;
        nx[1] = 0ull;
        nx[2] = 0ull;
      }
    }
  }
}

}  // namespace sdml

// peer_buf: HOST array of `world` device pointers (entry `rank` local, the rest peer-mapped) to
// zero-initialised landing buffers of 2 * world * total * 8 bytes, `total` = sum of sizes.
// epoch_dev: device uint32 holding the number of completed calls -- the kernel itself advances it
// (last CTA to finish), so a captured launch replays; exit_counter: device uint32, zero before the
// first call.  All replicas must call in lock-step.  timeout_ns / err (optional, 4 x uint32
// host-visible): a poll that waits longer leaves {7, peer, epoch} in err and the kernel finishes.
// stamps (optional, 64 x 4 uint64): per-call globaltimer probes {first CTA in, last push issued, last
// CTA done} at index 4 * (epoch % 64), for profiling.
extern "C" int sdml_dp_ll_allreduce_sgd(int n_tensors, void* const* params, void* const* grads,
                                        void* const* moms, const long long* sizes, int rank,
                                        int world, void* const* peer_buf, void* epoch_dev,
                                        void* exit_counter, float lr, float momentum,
                                        float weight_decay, float grad_scale,
                                        unsigned long long timeout_ns, void* err, void* stamps,
                                        cudaStream_t stream) {
  if (n_tensors <= 0 || n_tensors > kDpMaxTensors)
    return set_error("sdml_dp_ll_allreduce_sgd: 1..32 tensors per call");
  if (world < 1 || world > kDpMaxRanks || rank < 0 || rank >= world || !epoch_dev || !exit_counter)
    return set_error("sdml_dp_ll_allreduce_sgd: bad world / rank / epoch / counter pointer");
  DpTable t;
  long long total = 0;
  t.count = n_tensors;
  for (int i = 0; i < n_tensors; ++i) {
    t.p[i] = static_cast<float*>(params[i]);
    t.g[i] = static_cast<float*>(grads[i]);
    t.m[i] = static_cast<float*>(moms[i]);
    t.n[i] = sizes[i];
    t.off[i] = total;
    total += sizes[i];
  }
  DpLLPeers pr;
  for (int r = 0; r < world; ++r) pr.buf[r] = static_cast<unsigned long long*>(peer_buf[r]);
  for (int r = world; r < kDpMaxRanks; ++r) pr.buf[r] = pr.buf[0];
  pr.epoch_dev = static_cast<uint32_t*>(epoch_dev);
  pr.exit_counter = static_cast<uint32_t*>(exit_counter);
  pr.stamps = static_cast<unsigned long long*>(stamps);
  pr.timeout_ns = timeout_ns;
  pr.err = static_cast<uint32_t*>(err);
  pr.total = total;
  pr.rank = rank;
  pr.world = world;
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  // threads spin on data pushed by peers whose kernels must be running too: stay co-resident
  long long blocks = (total + kDpThreads - 1) / kDpThreads;
  if (blocks > num_sms(dev)) blocks = num_sms(dev);
  if (blocks < 1) blocks = 1;
  dp_ll_allreduce_sgd_kernel<<<(int)blocks, kDpThreads, 0, stream>>>(t, pr, lr, momentum,
                                                                    weight_decay, grad_scale);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// ============================================================================
// Allocation-free variant (single replica)
// ============================================================================
// sdml_dp_allreduce_sgd (above) leaves it to the caller to provide the staging buffers,
// the flag array and the CTA counter, which in practice means somebody calls cudaMalloc.
// Some hosts of this library cannot do that once they are up and running:
//   * a runtime that grabs (nearly) all device memory for its own arena at start-up, so
//     a later cudaMalloc fails or eats into a budget it does not own;
//   * a latency-critical training / serving loop, where cudaMalloc and cudaFree
//     synchronise the whole device and stall every stream;
//   * code that runs with allocation forbidden by policy (fixed memory plans, MPS
//     partitions with a hard limit).
// For them the same kernel is offered with its scratch in a fixed, statically sized
// __device__ array that lives in this library's image on every GPU that loads it and
// stays there until the process exits.  The price: a hard cap on the problem size, the
// sharing rules written on the entry point below, and ONE replica only -- other GPUs can
// neither IPC-map nor peer-address a module-global array (a peer load of it faults, tried
// on 2 x B200), so the cross-GPU form keeps needing caller-provided symmetric buffers.
namespace sdml {

constexpr long long kDpStaticMaxFloats = 65536;  // floats per staging buffer (256 KiB)
constexpr size_t kDpStaticStageBytes = static_cast<size_t>(kDpStaticMaxFloats) * sizeof(float);
constexpr size_t kDpStaticFlagsOff = 2 * kDpStaticStageBytes;      // two parities
constexpr size_t kDpStaticCounterOff = kDpStaticFlagsOff + 256;    // kDpMaxRanks x u32, padded
constexpr size_t kDpStaticBytes = kDpStaticCounterOff + 256;       // = 512 KiB + 512 B per GPU

// One instance per device that loads the module.  In HOST code this name is only a
// handle for cudaGetSymbolAddress -- it is not a device address.
static __device__ __align__(256) unsigned char g_dp_static_scratch[kDpStaticBytes];

struct DpStaticDevice {
  unsigned char* base = nullptr;  // device address of g_dp_static_scratch on this GPU
  unsigned int epoch = 0;         // calls made on this GPU so far
};
static DpStaticDevice g_dp_static_dev[64];
static std::mutex g_dp_static_mu;

}  // namespace sdml

// Same operation and same kernel as sdml_dp_allreduce_sgd with world == 1, but the entry
// point itself owns the scratch: it calls no cudaMalloc / cudaFree and no allocating
// library routine.
//
//   params / grads / moms / sizes : as for sdml_dp_allreduce_sgd (HOST arrays of device
//                                   pointers to fp32 tensors on the CURRENT device);
//                                   p -= lr * (m = momentum * m + grad_scale * g + wd * p),
//                                   all in place, gradients end up zeroed.
//
// Limits and rules
//   * the tensors of one call may hold at most 65 536 fp32 elements in total (the reference
//     model has 21 840); a larger call is rejected with an error and launches nothing;
//   * the scratch is 512.5 KiB per GPU, present from the moment the library is loaded on
//     that GPU until the process ends;
//   * the first call on a device looks up that device's scratch address
//     (cudaGetSymbolAddress with that device current, which also forces the module to load
//     there).  Make that first call outside of any stream capture;
//   * every call on a device uses the same array and finds it as the previous call left it.
//     The CTA counter inside it is cleared on `stream` before each launch; staging data is
//     overwritten and the epoch flag only ever advances (the epoch is counted inside the
//     library, per device).  Two calls whose kernels overlap on one GPU -- different
//     streams -- corrupt each other: serialise them.
extern "C" int sdml_dp_allreduce_sgd_static(int n_tensors, void* const* params,
                                            void* const* grads, void* const* moms,
                                            const long long* sizes, float lr, float momentum,
                                            float weight_decay, float grad_scale,
                                            cudaStream_t stream) {
  if (n_tensors <= 0 || n_tensors > kDpMaxTensors)
    return set_error("sdml_dp_allreduce_sgd_static: 1..32 tensors per call");
  long long total = 0;
  for (int i = 0; i < n_tensors; ++i) {
    if (sizes[i] < 0) return set_error("sdml_dp_allreduce_sgd_static: negative tensor size");
    total += sizes[i];
  }
  if (total > kDpStaticMaxFloats)
    return set_error(
        "sdml_dp_allreduce_sgd_static: more than 65536 fp32 elements in one call; the built-in "
        "scratch is fixed -- use sdml_dp_allreduce_sgd with caller-provided buffers");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return set_error("sdml_dp_allreduce_sgd_static: device ordinal out of range");

  DpTable t;
  t.count = n_tensors;
  long long off = 0;
  for (int i = 0; i < n_tensors; ++i) {
    t.p[i] = static_cast<float*>(params[i]);
    t.g[i] = static_cast<float*>(grads[i]);
    t.m[i] = static_cast<float*>(moms[i]);
    t.n[i] = sizes[i];
    t.off[i] = off;
    off += sizes[i];
  }

  DpPeers pr;
  unsigned int epoch;
  {
    std::lock_guard<std::mutex> lock(g_dp_static_mu);
    DpStaticDevice& d = g_dp_static_dev[dev];
    if (!d.base) {
      // set-up, once per device, with that device current: the array's name is a host-side
      // stand-in, the kernel needs the address it has on THIS GPU
      unsigned char* ptr = nullptr;
      SDML_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_dp_static_scratch));
      d.base = ptr;
    }
    epoch = ++d.epoch;
    pr.stage[0] = reinterpret_cast<float*>(d.base + (epoch & 1u) * kDpStaticStageBytes);
    pr.flags[0] = reinterpret_cast<uint32_t*>(d.base + kDpStaticFlagsOff);
    pr.local_counter = reinterpret_cast<uint32_t*>(d.base + kDpStaticCounterOff);
  }
  pr.guard = WaitGuard{0ull, nullptr};   // one replica: the flag it waits for is its own
  pr.rank = 0;
  pr.world = 1;
  pr.epoch_dev = nullptr;

  // the only part of the shared array that has to start from zero
  SDML_CUDA_OK(cudaMemsetAsync(pr.local_counter, 0, sizeof(uint32_t), stream));
  long long blocks = (total + kDpThreads * 4 - 1) / (kDpThreads * 4);
  if (blocks > num_sms(dev)) blocks = num_sms(dev);
  if (blocks < 1) blocks = 1;
  dp_allreduce_sgd_kernel<<<(int)blocks, kDpThreads, 0, stream>>>(t, pr, epoch, lr, momentum,
                                                                 weight_decay, grad_scale);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}
