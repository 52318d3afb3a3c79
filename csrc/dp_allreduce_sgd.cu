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
  uint32_t* flags[kDpMaxRanks];   // flag array of rank r: flags[r][src] = last epoch published by src
  uint32_t* local_counter;        // CTA completion counter (local memory)
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
  float* my_stage = pr.stage[pr.rank];

  // ---- phase 1: stage local gradients ----
  for (int i = 0; i < t.count; ++i) {
    const float* __restrict__ g = t.g[i];
    float* __restrict__ s = my_stage + t.off[i];
    for (long long k = tid; k < t.n[i]; k += stride) s[k] = g[k] * grad_scale;
  }
  __threadfence_system();
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
    while (static_cast<int32_t>(ld_acquire_sys(f) - epoch) < 0) {
    }
  }
  __syncthreads();
  for (int i = 0; i < t.count; ++i) {
    float* __restrict__ p = t.p[i];
    float* __restrict__ g = t.g[i];
    float* __restrict__ m = t.m[i];
    const long long off = t.off[i];
    for (long long k = tid; k < t.n[i]; k += stride) {
      float sum = 0.f;
      for (int r = 0; r < pr.world; ++r) sum += ld_volatile_f32(pr.stage[r] + off + k);
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
  pr.rank = rank;
  pr.world = world;
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
