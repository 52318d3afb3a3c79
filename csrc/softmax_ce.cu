// Fused softmax cross-entropy: one pass computes, per row, the loss
// (logsumexp - logit[target]), the gradient (softmax - onehot) * grad_scale and
// the arg-max hit, i.e. the reference's log_softmax (simple_distributed.py:79)
// + nll_loss (:111 train / :126 eval sum) + backward + the eval accuracy
// reduction (:127-128), in a single kernel on the last pipeline stage.
//
// One CTA per row; the row is read twice (second read hits L1/L2).  bf16 or fp32
// logits, fp32 math.  loss_sum / correct are accumulated with one atomic per row.
#include "sdml_host.h"
#include "sdml_ptx.cuh"

namespace sdml {

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

constexpr int kCeThreads = 256;

struct MaxSumIdx {
  float m;   // running max
  float s;   // running sum of exp(x - m)
  int idx;   // arg max
};

__device__ __forceinline__ MaxSumIdx combine(MaxSumIdx a, MaxSumIdx b) {
  if (b.s == 0.f) return a;  // empty partial (also avoids -inf - -inf)
  if (a.s == 0.f) return b;
  MaxSumIdx r;
  if (b.m > a.m || (b.m == a.m && b.idx < a.idx)) {
    r.m = b.m; r.idx = b.idx;
    r.s = b.s + a.s * __expf(a.m - b.m);
  } else {
    r.m = a.m; r.idx = a.idx;
    r.s = a.s + b.s * __expf(b.m - a.m);
  }
  return r;
}

template <typename T>
__global__ void __launch_bounds__(kCeThreads)
softmax_ce_kernel(const T* __restrict__ logits, long long ld, const long long* __restrict__ targets,
                  int rows, int classes, float* __restrict__ loss_sum, T* __restrict__ dlogits,
                  long long ld_d, float grad_scale, int* __restrict__ correct) {
  const int row = blockIdx.x;
  if (row >= rows) return;
  const T* x = logits + static_cast<long long>(row) * ld;
  MaxSumIdx st{-INFINITY, 0.f, 0x7fffffff};
  for (int c = threadIdx.x; c < classes; c += kCeThreads) {
    const float v = to_f32(x[c]);
    MaxSumIdx e{v, 1.f, c};
    st = combine(st, e);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSumIdx other;
    other.m = __shfl_xor_sync(0xffffffffu, st.m, o);
    other.s = __shfl_xor_sync(0xffffffffu, st.s, o);
    other.idx = __shfl_xor_sync(0xffffffffu, st.idx, o);
    st = combine(st, other);
  }
  __shared__ MaxSumIdx warp_res[kCeThreads / 32];
  __shared__ MaxSumIdx final_res;
  if ((threadIdx.x & 31) == 0) warp_res[threadIdx.x >> 5] = st;
  __syncthreads();
  if (threadIdx.x < 32) {
    MaxSumIdx v = threadIdx.x < kCeThreads / 32 ? warp_res[threadIdx.x]
                                                : MaxSumIdx{-INFINITY, 0.f, 0x7fffffff};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      MaxSumIdx other;
      other.m = __shfl_xor_sync(0xffffffffu, v.m, o);
      other.s = __shfl_xor_sync(0xffffffffu, v.s, o);
      other.idx = __shfl_xor_sync(0xffffffffu, v.idx, o);
      v = combine(v, other);
    }
    if (threadIdx.x == 0) final_res = v;
  }
  __syncthreads();
  const float mx = final_res.m;
  const float lse = mx + __logf(final_res.s);
  const int tgt = static_cast<int>(targets[row]);
  if (threadIdx.x == 0) {
    if (loss_sum) atomicAdd(loss_sum, lse - to_f32(x[tgt]));
    if (correct && final_res.idx == tgt) atomicAdd(correct, 1);
  }
  if (dlogits) {
    T* d = dlogits + static_cast<long long>(row) * ld_d;
    for (int c = threadIdx.x; c < classes; c += kCeThreads) {
      float p = __expf(to_f32(x[c]) - lse);
      if (c == tgt) p -= 1.f;
      d[c] = from_f32<T>(p * grad_scale);
    }
  }
}

}  // namespace sdml

using namespace sdml;

// dtype: 0 = fp32 logits/grad, 1 = bf16 logits/grad.  targets are int64.
// loss_sum (fp32 scalar) and correct (int32 scalar) are ACCUMULATED; either may be null.
extern "C" int sdml_softmax_ce(const void* logits, long long ld, int dtype, const void* targets,
                               int rows, int classes, float* loss_sum, void* dlogits,
                               long long ld_d, float grad_scale, int* correct,
                               cudaStream_t stream) {
  if (rows <= 0 || classes <= 0) return set_error("sdml_softmax_ce: empty problem");
  if (dtype == 0) {
    softmax_ce_kernel<float><<<rows, kCeThreads, 0, stream>>>(
        static_cast<const float*>(logits), ld, static_cast<const long long*>(targets), rows,
        classes, loss_sum, static_cast<float*>(dlogits), ld_d, grad_scale, correct);
  } else if (dtype == 1) {
    softmax_ce_kernel<__nv_bfloat16><<<rows, kCeThreads, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(logits), ld, static_cast<const long long*>(targets),
        rows, classes, loss_sum, static_cast<__nv_bfloat16*>(dlogits), ld_d, grad_scale, correct);
  } else {
    return set_error("sdml_softmax_ce: dtype must be 0 (fp32) or 1 (bf16)");
  }
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}
