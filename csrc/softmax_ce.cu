// Fused softmax cross-entropy: one pass computes, per row, the loss
// (logsumexp - logit[target]), the gradient (softmax - onehot) * grad_scale and
// the arg-max hit, i.e. the reference's log_softmax (simple_distributed.py:79)
// + nll_loss (:111 train / :126 eval sum) + backward + the eval accuracy
// reduction (:127-128), in a single kernel on the last pipeline stage.
//
// Fast path (row length a multiple of 8 bf16 / 4 fp32 values, up to 8192 bf16 / 4096 fp32 classes,
// 16-byte aligned rows): a persistent grid of CTAs walks the rows; a row is read ONCE with 16-byte
// loads and lives in registers (max + arg-max reduction, one exp per element, sum reduction, the
// gradient is written with 16-byte stores from the same registers); loss and hit count are
// accumulated per CTA and leave it as ONE atomic each.  Other shapes: one CTA per row, scalar
// loads, the row read twice.  A target outside [0, classes) never indexes memory: its row
// contributes NaN to the loss (visible at once) and a gradient without the -1 term.
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
  const long long tgt64 = targets[row];
  const bool tgt_ok = tgt64 >= 0 && tgt64 < classes;
  const int tgt = tgt_ok ? static_cast<int>(tgt64) : -1;
  if (threadIdx.x == 0) {
    if (loss_sum) atomicAdd(loss_sum, tgt_ok ? lse - to_f32(x[tgt]) : __int_as_float(0x7fc00000));
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

// ---- fast path: row in registers, 16-byte accesses, one atomic per CTA ----
template <typename T>
struct Vec16;
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int W = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = bf16_lo(w[i]);
      v[2 * i + 1] = bf16_hi(w[i]);
    }
  }
  __device__ static void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 u;
    u.x = pack_bf16x2(v[0], v[1]);
    u.y = pack_bf16x2(v[2], v[3]);
    u.z = pack_bf16x2(v[4], v[5]);
    u.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};
template <>
struct Vec16<float> {
  static constexpr int W = 4;
  __device__ static void load(const float* p, float (&v)[4]) {
    const float4 u = *reinterpret_cast<const float4*>(p);
    v[0] = u.x; v[1] = u.y; v[2] = u.z; v[3] = u.w;
  }
  __device__ static void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};

// VPT = 16-byte vectors per thread (the row has at most VPT * 256 of them)
template <typename T, int VPT>
__global__ void __launch_bounds__(kCeThreads)
softmax_ce_rows_kernel(const T* __restrict__ logits, long long ld, const long long* __restrict__ targets,
                       int rows, int classes, float* __restrict__ loss_sum, T* __restrict__ dlogits,
                       long long ld_d, float grad_scale, int* __restrict__ correct) {
  constexpr int W = Vec16<T>::W;
  const int nvec = classes / W;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __shared__ float s_max[kCeThreads / 32];
  __shared__ int s_idx[kCeThreads / 32];
  __shared__ float s_sum[kCeThreads / 32];
  float cta_loss = 0.f;   // thread 0 only
  int cta_hits = 0;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* x = logits + static_cast<long long>(row) * ld;
    float v[VPT][W];
    float mx = -INFINITY;
    int arg = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int vi = threadIdx.x + j * kCeThreads;
      if (vi < nvec) {
        Vec16<T>::load(x + vi * W, v[j]);
#pragma unroll
        for (int e = 0; e < W; ++e)
          if (v[j][e] > mx) { mx = v[j][e]; arg = vi * W + e; }   // ascending index: first maximum wins
      } else {
#pragma unroll
        for (int e = 0; e < W; ++e) v[j][e] = -INFINITY;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oi = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oi < arg)) { mx = om; arg = oi; }
    }
    if (lane == 0) { s_max[warp] = mx; s_idx[warp] = arg; }
    __syncthreads();
#pragma unroll
    for (int w = 0; w < kCeThreads / 32; ++w) {
      const float om = s_max[w];
      const int oi = s_idx[w];
      if (om > mx || (om == mx && oi < arg)) { mx = om; arg = oi; }
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < VPT; ++j)
#pragma unroll
      for (int e = 0; e < W; ++e) {
        v[j][e] = __expf(v[j][e] - mx);     // exp(-inf) = 0 for the padding
        sum += v[j][e];
      }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) s_sum[warp] = sum;
    __syncthreads();
    sum = 0.f;
#pragma unroll
    for (int w = 0; w < kCeThreads / 32; ++w) sum += s_sum[w];   // same order in every thread
    const long long tgt64 = targets[row];
    const bool tgt_ok = tgt64 >= 0 && tgt64 < classes;
    const int tgt = tgt_ok ? static_cast<int>(tgt64) : -1;
    if (threadIdx.x == 0) {
      // logit[target] is read back once (L1/L2 hit); lse - logit
      cta_loss += tgt_ok ? (mx + __logf(sum)) - to_f32(x[tgt]) : __int_as_float(0x7fc00000);
      cta_hits += (arg == tgt) ? 1 : 0;
    }
    if (dlogits) {
      T* d = dlogits + static_cast<long long>(row) * ld_d;
      const float inv = grad_scale / sum;
#pragma unroll
      for (int j = 0; j < VPT; ++j) {
        const int vi = threadIdx.x + j * kCeThreads;
        if (vi < nvec) {
          float g[W];
#pragma unroll
          for (int e = 0; e < W; ++e) g[e] = v[j][e] * inv - ((vi * W + e == tgt) ? grad_scale : 0.f);
          Vec16<T>::store(d + vi * W, g);
        }
      }
    }
    __syncthreads();   // s_max / s_sum are reused by the next row
  }
  if (threadIdx.x == 0) {
    if (loss_sum) atomicAdd(loss_sum, cta_loss);
    if (correct && cta_hits) atomicAdd(correct, cta_hits);
  }
}

template <typename T>
static bool launch_rows(const void* logits, long long ld, const void* targets, int rows, int classes,
                        float* loss_sum, void* dlogits, long long ld_d, float grad_scale, int* correct,
                        cudaStream_t stream) {
  constexpr int W = Vec16<T>::W;
  const size_t esz = sizeof(T);
  if (classes % W || classes > 4 * kCeThreads * W) return false;
  if ((reinterpret_cast<uintptr_t>(logits) & 15) || ((ld * esz) & 15)) return false;
  if (dlogits && ((reinterpret_cast<uintptr_t>(dlogits) & 15) || ((ld_d * esz) & 15))) return false;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  const int nvec = classes / W;
  int grid = 4 * num_sms(dev);      // 4 x 256 threads per SM; each CTA walks rows with this stride
  if (grid > rows) grid = rows;
  const T* x = static_cast<const T*>(logits);
  const long long* t = static_cast<const long long*>(targets);
  T* d = static_cast<T*>(dlogits);
  if (nvec <= kCeThreads)
    softmax_ce_rows_kernel<T, 1><<<grid, kCeThreads, 0, stream>>>(x, ld, t, rows, classes, loss_sum, d, ld_d, grad_scale, correct);
  else if (nvec <= 2 * kCeThreads)
    softmax_ce_rows_kernel<T, 2><<<grid, kCeThreads, 0, stream>>>(x, ld, t, rows, classes, loss_sum, d, ld_d, grad_scale, correct);
  else
    softmax_ce_rows_kernel<T, 4><<<grid, kCeThreads, 0, stream>>>(x, ld, t, rows, classes, loss_sum, d, ld_d, grad_scale, correct);
  return true;
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
  if (dtype != 0 && dtype != 1) return set_error("sdml_softmax_ce: dtype must be 0 (fp32) or 1 (bf16)");
  const bool fast = dtype == 0
      ? launch_rows<float>(logits, ld, targets, rows, classes, loss_sum, dlogits, ld_d, grad_scale, correct, stream)
      : launch_rows<__nv_bfloat16>(logits, ld, targets, rows, classes, loss_sum, dlogits, ld_d, grad_scale, correct, stream);
  if (fast) {
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
    return 0;
  }
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
