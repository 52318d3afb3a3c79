// Small-shape fp32 Linear (+bias, +ReLU, +dropout) forward and backward on CUDA
// cores, for the reference model's stage-1 layers fc1 320->50 and fc2 50->10 at
// batch 60 (/root/reference/simple_distributed.py:63-64, :75-77).  These shapes are
// far below one tensor-core tile (and 50 is not a TMA-legal stride), and the
// reference computes them in fp32, so they get an exact-fp32 SIMT path; the large
// bf16 layers go through gemm_tcgen05.cu.
//
// Dropout uses a counter-based hash RNG keyed by (seed, offset + element index),
// so the backward pass regenerates the mask instead of storing it.  The offset is an
// immediate plus an optional device-resident int64 (offset_dev): a step captured in a
// CUDA graph advances that counter on the device and gets a fresh mask on every replay.
#include "sdml_host.h"
#include "sdml_ptx.cuh"

namespace sdml {

constexpr int kT = 16;

__device__ __forceinline__ float uniform01(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ULL * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z = z ^ (z >> 31);
  return static_cast<float>(z >> 40) * (1.0f / 16777216.0f);
}

// y = dropout(relu(x . w^T + b))
__global__ void __launch_bounds__(kT* kT)
linear_f32_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                      const float* __restrict__ b, float* __restrict__ y, int M, int N, int K,
                      int relu, float drop_p, unsigned long long seed, unsigned long long offset,
                      const long long* __restrict__ offset_dev) {
  __shared__ float xs[kT][kT + 1];
  __shared__ float ws[kT][kT + 1];
  if (offset_dev) offset += static_cast<unsigned long long>(*offset_dev);
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int row = blockIdx.y * kT + ty;
  const int col = blockIdx.x * kT + tx;
  const int wrow = blockIdx.x * kT + ty;
  float acc = 0.f;
  for (int k0 = 0; k0 < K; k0 += kT) {
    xs[ty][tx] = (row < M && k0 + tx < K) ? x[static_cast<long long>(row) * K + k0 + tx] : 0.f;
    ws[ty][tx] = (wrow < N && k0 + tx < K) ? w[static_cast<long long>(wrow) * K + k0 + tx] : 0.f;
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kT; ++kk) acc += xs[ty][kk] * ws[tx][kk];
    __syncthreads();
  }
  if (row < M && col < N) {
    float v = acc + (b ? b[col] : 0.f);
    if (relu) v = fmaxf(v, 0.f);
    if (drop_p > 0.f) {
      const unsigned long long idx = offset + static_cast<unsigned long long>(row) * N + col;
      v = uniform01(seed, idx) >= drop_p ? v * (1.f / (1.f - drop_p)) : 0.f;
    }
    y[static_cast<long long>(row) * N + col] = v;
  }
}

// gradient w.r.t. the pre-activation: undo dropout (regenerated mask) and ReLU (y > 0)
__device__ __forceinline__ float dz_at(const float* __restrict__ dy, const float* __restrict__ y,
                                       int m, int n, int N, int relu, float drop_p,
                                       unsigned long long seed, unsigned long long offset) {
  float g = dy[static_cast<long long>(m) * N + n];
  if (drop_p > 0.f) {
    const unsigned long long idx = offset + static_cast<unsigned long long>(m) * N + n;
    g = uniform01(seed, idx) >= drop_p ? g * (1.f / (1.f - drop_p)) : 0.f;
  }
  if (relu && !(y[static_cast<long long>(m) * N + n] > 0.f)) g = 0.f;
  return g;
}

// dx[M,K] = dz[M,N] . w[N,K]
__global__ void __launch_bounds__(kT* kT)
linear_f32_dx_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                     const float* __restrict__ w, float* __restrict__ dx, int M, int N, int K,
                     int relu, float drop_p, unsigned long long seed, unsigned long long offset,
                     const long long* __restrict__ offset_dev) {
  __shared__ float zs[kT][kT + 1];
  __shared__ float ws[kT][kT + 1];
  if (offset_dev) offset += static_cast<unsigned long long>(*offset_dev);
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int row = blockIdx.y * kT + ty;  // m
  const int col = blockIdx.x * kT + tx;  // k
  float acc = 0.f;
  for (int n0 = 0; n0 < N; n0 += kT) {
    zs[ty][tx] = (row < M && n0 + tx < N) ? dz_at(dy, y, row, n0 + tx, N, relu, drop_p, seed, offset) : 0.f;
    ws[ty][tx] = (n0 + ty < N && col < K) ? w[static_cast<long long>(n0 + ty) * K + col] : 0.f;
    __syncthreads();
#pragma unroll
    for (int nn = 0; nn < kT; ++nn) acc += zs[ty][nn] * ws[nn][tx];
    __syncthreads();
  }
  if (row < M && col < K) dx[static_cast<long long>(row) * K + col] = acc;
}

// dw[N,K] (+)= dz^T[N,M] . x[M,K];  blocks with blockIdx.x == 0 also produce db[N] (+)= sum_m dz
__global__ void __launch_bounds__(kT* kT)
linear_f32_dw_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                     const float* __restrict__ x, float* __restrict__ dw, float* __restrict__ db,
                     int M, int N, int K, int relu, float drop_p, unsigned long long seed,
                     unsigned long long offset, const long long* __restrict__ offset_dev,
                     int accumulate) {
  __shared__ float zs[kT][kT + 1];  // [m][n]
  __shared__ float xs[kT][kT + 1];  // [m][k]
  if (offset_dev) offset += static_cast<unsigned long long>(*offset_dev);
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int n = blockIdx.y * kT + ty;
  const int k = blockIdx.x * kT + tx;
  float acc = 0.f, bacc = 0.f;
  for (int m0 = 0; m0 < M; m0 += kT) {
    const int nn = blockIdx.y * kT + tx;
    zs[ty][tx] = (m0 + ty < M && nn < N) ? dz_at(dy, y, m0 + ty, nn, N, relu, drop_p, seed, offset) : 0.f;
    xs[ty][tx] = (m0 + ty < M && k < K) ? x[static_cast<long long>(m0 + ty) * K + k] : 0.f;
    __syncthreads();
#pragma unroll
    for (int mm = 0; mm < kT; ++mm) {
      acc += zs[mm][ty] * xs[mm][tx];
      bacc += zs[mm][ty];
    }
    __syncthreads();
  }
  if (n < N && k < K) {
    float* o = dw + static_cast<long long>(n) * K + k;
    *o = accumulate ? *o + acc : acc;
  }
  if (db && blockIdx.x == 0 && tx == 0 && n < N) db[n] = accumulate ? db[n] + bacc : bacc;
}

}  // namespace sdml

using namespace sdml;

extern "C" int sdml_linear_f32_fwd(const float* x, const float* w, const float* b, float* y, int M,
                                   int N, int K, int relu, float drop_p, long long seed,
                                   long long offset, const long long* offset_dev,
                                   cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return set_error("sdml_linear_f32_fwd: empty problem");
  if (drop_p < 0.f || drop_p >= 1.f) return set_error("sdml_linear_f32_fwd: drop_p must be in [0,1)");
  dim3 block(kT, kT), grid((N + kT - 1) / kT, (M + kT - 1) / kT);
  linear_f32_fwd_kernel<<<grid, block, 0, stream>>>(x, w, b, y, M, N, K, relu, drop_p,
                                                    (unsigned long long)seed,
                                                    (unsigned long long)offset, offset_dev);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// dx and/or dw/db may be null to skip that gradient.
extern "C" int sdml_linear_f32_bwd(const float* x, const float* w, const float* y, const float* dy,
                                   float* dx, float* dw, float* db, int M, int N, int K, int relu,
                                   float drop_p, long long seed, long long offset,
                                   const long long* offset_dev, int accumulate,
                                   cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return set_error("sdml_linear_f32_bwd: empty problem");
  dim3 block(kT, kT);
  if (dx) {
    dim3 grid((K + kT - 1) / kT, (M + kT - 1) / kT);
    linear_f32_dx_kernel<<<grid, block, 0, stream>>>(dy, y, w, dx, M, N, K, relu, drop_p,
                                                     (unsigned long long)seed,
                                                     (unsigned long long)offset, offset_dev);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
  }
  if (dw) {
    dim3 grid((K + kT - 1) / kT, (N + kT - 1) / kT);
    linear_f32_dw_kernel<<<grid, block, 0, stream>>>(dy, y, x, dw, db, M, N, K, relu, drop_p,
                                                     (unsigned long long)seed,
                                                     (unsigned long long)offset, offset_dev,
                                                     accumulate);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
  }
  return 0;
}
