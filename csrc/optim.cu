// Fused multi-tensor optimizer updates (SGD with momentum, Adam) for one pipeline
// stage's parameters.  One launch updates up to kMaxTensors tensors: tensor table is
// passed in kernel-parameter space (no device-side metadata buffers), blockIdx.y
// selects the tensor, blockIdx.x grid-strides over it with 128-bit accesses.
//
// Fused into the update: gradient un-scaling (grad_scale, e.g. 1/micro-batches or
// 1/dp), weight decay, the bf16 cast of the updated fp32 master weight into the
// model's compute copy, and zeroing of the fp32 gradient accumulator for the next
// step.  Replaces the reference's DistributedOptimizer(SGD, lr=0.1, momentum=0.5)
// (/root/reference/simple_distributed.py:100-104, :113 -> torch functional_sgd).
#include "sdml_host.h"
#include "sdml_ptx.cuh"

namespace sdml {

constexpr int kMaxTensors = 48;
constexpr int kOptThreads = 256;

struct TensorTable {
  float* p[kMaxTensors];
  float* g[kMaxTensors];
  float* s0[kMaxTensors];          // momentum buffer / Adam m
  float* s1[kMaxTensors];          // Adam v
  __nv_bfloat16* lp[kMaxTensors];  // optional low-precision copy
  long long n[kMaxTensors];
};

struct SgdHyper {
  float lr, momentum, weight_decay, grad_scale;
  int first_step, zero_grad;
};

__device__ __forceinline__ float sgd_one(float p, float g, float& m, const SgdHyper& h) {
  g = g * h.grad_scale + h.weight_decay * p;
  if (h.momentum != 0.f) {
    m = h.first_step ? g : h.momentum * m + g;
    g = m;
  }
  return p - h.lr * g;
}

__global__ void __launch_bounds__(kOptThreads)
sgd_multi_kernel(const __grid_constant__ TensorTable t, const SgdHyper h) {
  const int ti = blockIdx.y;
  float* __restrict__ p = t.p[ti];
  float* __restrict__ g = t.g[ti];
  float* __restrict__ m = t.s0[ti];
  __nv_bfloat16* __restrict__ lp = t.lp[ti];
  const long long n = t.n[ti];
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                     reinterpret_cast<uintptr_t>(m)) & 15) == 0 &&
                   (lp == nullptr || (reinterpret_cast<uintptr_t>(lp) & 7) == 0);
  const long long tid = static_cast<long long>(blockIdx.x) * kOptThreads + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * kOptThreads;
  const long long n4 = vec ? n / 4 : 0;
  for (long long i = tid; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 gv = reinterpret_cast<float4*>(g)[i];
    float4 mv = h.momentum != 0.f && !h.first_step ? reinterpret_cast<float4*>(m)[i]
                                                   : make_float4(0.f, 0.f, 0.f, 0.f);
    pv.x = sgd_one(pv.x, gv.x, mv.x, h);
    pv.y = sgd_one(pv.y, gv.y, mv.y, h);
    pv.z = sgd_one(pv.z, gv.z, mv.z, h);
    pv.w = sgd_one(pv.w, gv.w, mv.w, h);
    reinterpret_cast<float4*>(p)[i] = pv;
    if (h.momentum != 0.f) reinterpret_cast<float4*>(m)[i] = mv;
    if (h.zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (lp) {
      uint2 o;
      o.x = pack_bf16x2(pv.x, pv.y);
      o.y = pack_bf16x2(pv.z, pv.w);
      reinterpret_cast<uint2*>(lp)[i] = o;
    }
  }
  for (long long i = n4 * 4 + tid; i < n; i += stride) {
    float mv = h.momentum != 0.f && !h.first_step ? m[i] : 0.f;
    const float pv = sgd_one(p[i], g[i], mv, h);
    p[i] = pv;
    if (h.momentum != 0.f) m[i] = mv;
    if (h.zero_grad) g[i] = 0.f;
    if (lp) lp[i] = __float2bfloat16(pv);
  }
}

struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay, grad_scale;
  float bias1, bias2_sqrt;  // 1 - beta1^t, sqrt(1 - beta2^t)
  int zero_grad;
};

__device__ __forceinline__ float adam_one(float p, float g, float& m, float& v, const AdamHyper& h) {
  g = g * h.grad_scale + h.weight_decay * p;
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  const float denom = sqrtf(v) / h.bias2_sqrt + h.eps;
  return p - (h.lr / h.bias1) * (m / denom);
}

__global__ void __launch_bounds__(kOptThreads)
adam_multi_kernel(const __grid_constant__ TensorTable t, const AdamHyper h) {
  const int ti = blockIdx.y;
  float* __restrict__ p = t.p[ti];
  float* __restrict__ g = t.g[ti];
  float* __restrict__ m = t.s0[ti];
  float* __restrict__ v = t.s1[ti];
  __nv_bfloat16* __restrict__ lp = t.lp[ti];
  const long long n = t.n[ti];
  const long long tid = static_cast<long long>(blockIdx.x) * kOptThreads + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * kOptThreads;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                     reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0 &&
                   (lp == nullptr || (reinterpret_cast<uintptr_t>(lp) & 7) == 0);
  const long long n4 = vec ? n / 4 : 0;
  for (long long i = tid; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    pv.x = adam_one(pv.x, gv.x, mv.x, vv.x, h);
    pv.y = adam_one(pv.y, gv.y, mv.y, vv.y, h);
    pv.z = adam_one(pv.z, gv.z, mv.z, vv.z, h);
    pv.w = adam_one(pv.w, gv.w, mv.w, vv.w, h);
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (h.zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (lp) {
      uint2 o;
      o.x = pack_bf16x2(pv.x, pv.y);
      o.y = pack_bf16x2(pv.z, pv.w);
      reinterpret_cast<uint2*>(lp)[i] = o;
    }
  }
  for (long long i = n4 * 4 + tid; i < n; i += stride) {
    float mv = m[i], vv = v[i];
    const float pv = adam_one(p[i], g[i], mv, vv, h);
    p[i] = pv; m[i] = mv; v[i] = vv;
    if (h.zero_grad) g[i] = 0.f;
    if (lp) lp[i] = __float2bfloat16(pv);
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                     long long n) {
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15) | (reinterpret_cast<uintptr_t>(dst) & 7)) == 0;
  const long long n4 = vec ? n / 4 : 0;
  for (long long i = tid; i < n4; i += stride) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    uint2 o;
    o.x = pack_bf16x2(v.x, v.y);
    o.y = pack_bf16x2(v.z, v.w);
    reinterpret_cast<uint2*>(dst)[i] = o;
  }
  for (long long i = n4 * 4 + tid; i < n; i += stride) dst[i] = __float2bfloat16(src[i]);
}

// column sums of a bf16 [rows, cols] matrix into fp32 (bias gradient).
// grid = (col tiles of 256, row chunks); each thread owns 2 adjacent columns.
constexpr int kColsumRows = 256;
__global__ void __launch_bounds__(128)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, long long ld, int rows, int cols,
                   float* __restrict__ out) {
  const int c = (blockIdx.x * 128 + threadIdx.x) * 2;
  if (c >= cols) return;
  const int r0 = blockIdx.y * kColsumRows;
  const int r1 = min(rows, r0 + kColsumRows);
  float s0 = 0.f, s1 = 0.f;
  if (c + 1 < cols && (ld & 1) == 0 && (reinterpret_cast<uintptr_t>(x) & 3) == 0) {
    for (int r = r0; r < r1; ++r) {
      const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(x + static_cast<long long>(r) * ld + c));
      s0 += bf16_lo(v);
      s1 += bf16_hi(v);
    }
  } else {
    for (int r = r0; r < r1; ++r) {
      s0 += __bfloat162float(x[static_cast<long long>(r) * ld + c]);
      if (c + 1 < cols) s1 += __bfloat162float(x[static_cast<long long>(r) * ld + c + 1]);
    }
  }
  atomicAdd(out + c, s0);
  if (c + 1 < cols) atomicAdd(out + c + 1, s1);
}

static int fill_table(TensorTable& t, int base, int cnt, void* const* params, void* const* grads,
                      void* const* s0, void* const* s1, void* const* lowp, const long long* sizes,
                      long long& max_n) {
  max_n = 0;
  for (int i = 0; i < cnt; ++i) {
    t.p[i] = static_cast<float*>(params[base + i]);
    t.g[i] = static_cast<float*>(grads[base + i]);
    t.s0[i] = s0 ? static_cast<float*>(s0[base + i]) : nullptr;
    t.s1[i] = s1 ? static_cast<float*>(s1[base + i]) : nullptr;
    t.lp[i] = lowp ? static_cast<__nv_bfloat16*>(lowp[base + i]) : nullptr;
    t.n[i] = sizes[base + i];
    if (sizes[base + i] > max_n) max_n = sizes[base + i];
  }
  return 0;
}

static int opt_grid_x(long long max_n, int dev) {
  long long want = (max_n + kOptThreads * 16 - 1) / (kOptThreads * 16);
  const long long cap = 4LL * num_sms(dev);
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return static_cast<int>(want);
}

}  // namespace sdml

using namespace sdml;

// params/grads/moms/lowp are HOST arrays of n_tensors device pointers (lowp may be null
// or contain nulls); sizes is a host array of element counts.  All tensors fp32.
extern "C" int sdml_sgd_multi(int n_tensors, void* const* params, void* const* grads,
                              void* const* moms, void* const* lowp, const long long* sizes,
                              float lr, float momentum, float weight_decay, float grad_scale,
                              int first_step, int zero_grad, cudaStream_t stream) {
  if (n_tensors <= 0) return 0;
  if (momentum != 0.f && !moms) return set_error("sdml_sgd_multi: momentum buffers required");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  SgdHyper h{lr, momentum, weight_decay, grad_scale, first_step, zero_grad};
  for (int base = 0; base < n_tensors; base += kMaxTensors) {
    const int cnt = n_tensors - base < kMaxTensors ? n_tensors - base : kMaxTensors;
    TensorTable t;
    long long max_n;
    fill_table(t, base, cnt, params, grads, moms, nullptr, lowp, sizes, max_n);
    dim3 grid(opt_grid_x(max_n, dev), cnt);
    sgd_multi_kernel<<<grid, kOptThreads, 0, stream>>>(t, h);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
  }
  return 0;
}

extern "C" int sdml_adam_multi(int n_tensors, void* const* params, void* const* grads,
                               void* const* m, void* const* v, void* const* lowp,
                               const long long* sizes, float lr, float beta1, float beta2,
                               float eps, float weight_decay, float grad_scale, int step,
                               int zero_grad, cudaStream_t stream) {
  if (n_tensors <= 0) return 0;
  if (!m || !v) return set_error("sdml_adam_multi: moment buffers required");
  if (step < 1) return set_error("sdml_adam_multi: step counts from 1");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  AdamHyper h;
  h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps;
  h.weight_decay = weight_decay; h.grad_scale = grad_scale;
  h.bias1 = 1.f - powf(beta1, (float)step);
  h.bias2_sqrt = sqrtf(1.f - powf(beta2, (float)step));
  h.zero_grad = zero_grad;
  for (int base = 0; base < n_tensors; base += kMaxTensors) {
    const int cnt = n_tensors - base < kMaxTensors ? n_tensors - base : kMaxTensors;
    TensorTable t;
    long long max_n;
    fill_table(t, base, cnt, params, grads, m, v, lowp, sizes, max_n);
    dim3 grid(opt_grid_x(max_n, dev), cnt);
    adam_multi_kernel<<<grid, kOptThreads, 0, stream>>>(t, h);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
  }
  return 0;
}

extern "C" int sdml_cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t stream) {
  if (n <= 0) return 0;
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  long long blocks = (n + 256 * 8 - 1) / (256 * 8);
  const long long cap = 8LL * num_sms(dev);
  if (blocks > cap) blocks = cap;
  cast_f32_bf16_kernel<<<(int)blocks, 256, 0, stream>>>(src, static_cast<__nv_bfloat16*>(dst), n);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// out[c] (+)= sum_r x[r, c];  accumulate == 0 zeroes `out` first (async memset on `stream`).
extern "C" int sdml_colsum_bf16(const void* x, long long ld, int rows, int cols, float* out,
                                int accumulate, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return set_error("sdml_colsum_bf16: empty problem");
  if (!accumulate) SDML_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(float) * cols, stream));
  dim3 grid((cols + 255) / 256, (rows + kColsumRows - 1) / kColsumRows);
  colsum_bf16_kernel<<<grid, 128, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), ld, rows,
                                               cols, out);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}
