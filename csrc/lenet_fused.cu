// Fused kernels for the reference model (LeNet on 28x28 images, batch ~60).
//
// The reference runs this model as ~50 ATen kernels per step
// (/root/reference/simple_distributed.py:39-50, 69-80, 109-113).  At batch 60 that is pure
// launch / tiny-kernel overhead, so here the whole step is FOUR kernels plus the fused SGD:
//
//   lenet_s0_fwd      one CTA per sample: conv1 -> maxpool -> ReLU -> conv2 -> Dropout2d ->
//                     maxpool -> ReLU -> flatten, everything in shared memory
//                     (ref:42-46).  Saves z1, the pooling arg-maxes and the channel mask.
//   lenet_s1_fwd_bwd  one CTA per sample: fc1 -> ReLU -> dropout -> fc2 -> log-softmax ->
//                     NLL (ref:75-79, :111), then straight back: dlogits, dW2/db2/db1
//                     (atomics), dz4, and d(z3) for stage 0.  Also the eval path
//                     (loss sum + correct count, ref:126-128).
//   lenet_fc1_wgrad   dW1 = dz4^T . z3 over the batch.
//   lenet_s0_bwd      one CTA per sample: un-pool / ReLU / dropout backward, conv2 wgrad and
//                     dgrad exploiting that only one position per pooling window carries
//                     gradient, conv1 wgrad; per-sample weight gradients reduced with atomics.
//
// fp32 throughout (the reference is fp32).  Dropout masks come from a counter-based hash of
// (seed, offset + index) where the offset is read from device memory, so the backward
// kernels regenerate them and CUDA-graph replays see fresh masks.
#include "sdml_host.h"
#include <atomic>
#include "sdml_ptx.cuh"

namespace sdml {

__device__ __forceinline__ float ln_uniform01(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ULL * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z = z ^ (z >> 31);
  return static_cast<float>(z >> 40) * (1.0f / 16777216.0f);
}

constexpr int kS0Threads = 512;

// ---------------------------------------------------------------------------------------------
// stage 0 forward
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kS0Threads)
lenet_s0_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w1,
                    const float* __restrict__ b1, const float* __restrict__ w2,
                    const float* __restrict__ b2, float* __restrict__ z3,
                    float* __restrict__ z1_save, unsigned char* __restrict__ idx1,
                    unsigned char* __restrict__ idx2, float* __restrict__ mask2, float drop_p,
                    unsigned long long seed, const long long* __restrict__ offset_dev,
                    long long offset_imm, float* __restrict__ zero_this,
                    long long* __restrict__ bump_counter, long long bump_by) {
  __shared__ float s_img[784];
  __shared__ float s_w1[250];
  __shared__ float s_b1[10];
  __shared__ float s_z1[1440];
  __shared__ float s_w2[5000];
  __shared__ float s_b2[20];
  __shared__ float s_c2[1280];
  __shared__ float s_mask[20];
  const int b = blockIdx.x;
  const int t = threadIdx.x;
  if (b == 0 && t == 0) {
    // housekeeping folded into this launch: clear the loss accumulator the NEXT kernel adds to,
    // and advance the RNG position that kernel will read (never the one this kernel reads)
    if (zero_this) *zero_this = 0.f;
    if (bump_counter) *bump_counter += bump_by;
  }
  {
    // all global loads of the prologue are issued before the first shared-memory store, so the
    // CTA pays one L2 round trip instead of a dozen dependent ones
    float rw2[10], rimg[2], rw1 = 0.f;
#pragma unroll
    for (int j = 0; j < 10; ++j) rw2[j] = (t + j * kS0Threads < 5000) ? w2[t + j * kS0Threads] : 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j)
      rimg[j] = (t + j * kS0Threads < 784) ? x[static_cast<long long>(b) * 784 + t + j * kS0Threads] : 0.f;
    if (t < 250) rw1 = w1[t];
#pragma unroll
    for (int j = 0; j < 10; ++j)
      if (t + j * kS0Threads < 5000) s_w2[t + j * kS0Threads] = rw2[j];
#pragma unroll
    for (int j = 0; j < 2; ++j)
      if (t + j * kS0Threads < 784) s_img[t + j * kS0Threads] = rimg[j];
    if (t < 250) s_w1[t] = rw1;
  }
  if (t < 10) s_b1[t] = b1[t];
  if (t < 20) {
    s_b2[t] = b2[t];
    float m = 1.f;
    if (drop_p > 0.f) {
      const unsigned long long off =
          static_cast<unsigned long long>(offset_imm) + (offset_dev ? static_cast<unsigned long long>(*offset_dev) : 0ULL);
      m = ln_uniform01(seed, off + static_cast<unsigned long long>(b) * 20 + t) >= drop_p
              ? 1.f / (1.f - drop_p) : 0.f;
    }
    s_mask[t] = m;
    mask2[b * 20 + t] = m;
  }
  __syncthreads();
  // conv1 (1->10, k5) + maxpool 2 + ReLU : 10 x 12 x 12 pooled outputs
  for (int o = t; o < 1440; o += kS0Threads) {
    const int c = o / 144, r = o - c * 144, py = r / 12, px = r - py * 12;
    float patch[36];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j) patch[i * 6 + j] = s_img[(2 * py + i) * 28 + 2 * px + j];
    float a00 = s_b1[c], a01 = a00, a10 = a00, a11 = a00;
#pragma unroll
    for (int ky = 0; ky < 5; ++ky)
#pragma unroll
      for (int kx = 0; kx < 5; ++kx) {
        const float w = s_w1[c * 25 + ky * 5 + kx];
        a00 += w * patch[ky * 6 + kx];
        a01 += w * patch[ky * 6 + kx + 1];
        a10 += w * patch[(ky + 1) * 6 + kx];
        a11 += w * patch[(ky + 1) * 6 + kx + 1];
      }
    float best = a00;
    int bi = 0;
    if (a01 > best) { best = a01; bi = 1; }
    if (a10 > best) { best = a10; bi = 2; }
    if (a11 > best) { best = a11; bi = 3; }
    const float v = fmaxf(best, 0.f);
    s_z1[o] = v;
    z1_save[static_cast<long long>(b) * 1440 + o] = v;
    idx1[static_cast<long long>(b) * 1440 + o] = static_cast<unsigned char>(bi);
  }
  __syncthreads();
  // conv2 (10->20, k5): item = (oc, y, half) computes 4 adjacent outputs of one row
  for (int it = t; it < 320; it += kS0Threads) {
    const int oc = it >> 4, y = (it >> 1) & 7, x0 = (it & 1) * 4;
    float acc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = s_b2[oc];
    for (int ic = 0; ic < 10; ++ic) {
#pragma unroll
      for (int ky = 0; ky < 5; ++ky) {
        float row[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) row[i] = s_z1[ic * 144 + (y + ky) * 12 + x0 + i];
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) {
          const float w = s_w2[((oc * 10 + ic) * 5 + ky) * 5 + kx];
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[i] += w * row[i + kx];
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) s_c2[oc * 64 + y * 8 + x0 + i] = acc[i];
  }
  __syncthreads();
  // Dropout2d (per channel) + maxpool 2 + ReLU -> z3[b, oc*16 + py*4 + px]
  for (int o = t; o < 320; o += kS0Threads) {
    const int oc = o >> 4, r = o & 15, py = r >> 2, px = r & 3;
    const float* c = s_c2 + oc * 64 + (2 * py) * 8 + 2 * px;
    float best = c[0];
    int bi = 0;
    if (c[1] > best) { best = c[1]; bi = 1; }
    if (c[8] > best) { best = c[8]; bi = 2; }
    if (c[9] > best) { best = c[9]; bi = 3; }
    z3[static_cast<long long>(b) * 320 + o] = fmaxf(best * s_mask[oc], 0.f);
    idx2[static_cast<long long>(b) * 320 + o] = static_cast<unsigned char>(bi);
  }
}

// ---------------------------------------------------------------------------------------------
// stage 1: forward + loss + backward (train) or forward + metrics (eval); one CTA per sample
// ---------------------------------------------------------------------------------------------
constexpr int kS1Threads = 256;

__global__ void __launch_bounds__(kS1Threads)
lenet_s1_kernel(const float* __restrict__ z3, const long long* __restrict__ targets,
                const float* __restrict__ w1, const float* __restrict__ b1,
                const float* __restrict__ w2, const float* __restrict__ b2, int train,
                float grad_scale, float drop_p, unsigned long long seed,
                const long long* __restrict__ offset_dev, long long offset_imm,
                float* __restrict__ loss_sum, float loss_scale, int* __restrict__ correct,
                float* __restrict__ dz3, float* __restrict__ dz4_save, float* __restrict__ dw2,
                float* __restrict__ db2, float* __restrict__ db1,
                long long* __restrict__ bump_counter, long long bump_by) {
  __shared__ float s_x[320];
  __shared__ float s_z4[50];
  __shared__ float s_keep[50];
  __shared__ float s_logit[10];
  __shared__ float s_dl[10];
  __shared__ float s_dz4[50];
  const int b = blockIdx.x, t = threadIdx.x, warp = t >> 5, lane = t & 31;
  for (int i = t; i < 320; i += kS1Threads) s_x[i] = z3[static_cast<long long>(b) * 320 + i];
  __syncthreads();
  // fc1 + ReLU + dropout: warp per output, lanes over k (coalesced weight rows)
  const unsigned long long off =
      static_cast<unsigned long long>(offset_imm) + (offset_dev ? static_cast<unsigned long long>(*offset_dev) : 0ULL);
  {
    // 8 warps x 7 outputs: all 70 weight loads of a lane are in flight together (one L2 round
    // trip for the whole layer instead of one per output)
    constexpr int kPer = 7;
    float acc[kPer];
    float xv[10];
#pragma unroll
    for (int j = 0; j < 10; ++j) xv[j] = s_x[lane + 32 * j];
#pragma unroll
    for (int q = 0; q < kPer; ++q) {
      const int n = warp * kPer + q;
      acc[q] = 0.f;
      if (n < 50) {
#pragma unroll
        for (int j = 0; j < 10; ++j) acc[q] += xv[j] * __ldg(w1 + n * 320 + lane + 32 * j);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int q = 0; q < kPer; ++q) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], o);
    if (lane < kPer) {
      float a = acc[0];
#pragma unroll
      for (int q = 1; q < kPer; ++q)
        if (lane == q) a = acc[q];
      const int n = warp * kPer + lane;
      if (n < 50) {
        const float v = fmaxf(a + b1[n], 0.f);
        float keep = 1.f;
        if (drop_p > 0.f)
          keep = ln_uniform01(seed, off + (1ULL << 20) + static_cast<unsigned long long>(b) * 50 + n) >= drop_p
                     ? 1.f / (1.f - drop_p) : 0.f;
        s_keep[n] = keep;
        s_z4[n] = v * keep;
      }
    }
  }
  __syncthreads();
  // fc2: 10 logits
  if (t < 10) {
    float acc = b2[t];
#pragma unroll
    for (int j = 0; j < 50; ++j) acc += s_z4[j] * __ldg(w2 + t * 50 + j);
    s_logit[t] = acc;
  }
  __syncthreads();
  if (t == 0) {
    float mx = s_logit[0];
    int am = 0;
    for (int c = 1; c < 10; ++c)
      if (s_logit[c] > mx) { mx = s_logit[c]; am = c; }
    float sum = 0.f;
    for (int c = 0; c < 10; ++c) sum += __expf(s_logit[c] - mx);
    const float lse = mx + __logf(sum);
    const int tgt = static_cast<int>(targets[b]);
    if (loss_sum) atomicAdd(loss_sum, (lse - s_logit[tgt]) * loss_scale);
    if (b == 0 && bump_counter) *bump_counter += bump_by;   // RNG position of the NEXT stage-0 forward
    if (correct && am == tgt) atomicAdd(correct, 1);
    for (int c = 0; c < 10; ++c) s_dl[c] = (__expf(s_logit[c] - lse) - (c == tgt ? 1.f : 0.f)) * grad_scale;
  }
  if (!train) return;
  __syncthreads();
  // fc2 backward: dW2 += dl (x) z4, db2 += dl, dz4 = (dl . W2) * relu' * keep
  for (int i = t; i < 500; i += kS1Threads) {
    const int c = i / 50, j = i - c * 50;
    atomicAdd(dw2 + i, s_dl[c] * s_z4[j]);
  }
  if (t < 10) atomicAdd(db2 + t, s_dl[t]);
  if (t < 50) {
    float g = 0.f;
#pragma unroll
    for (int c = 0; c < 10; ++c) g += s_dl[c] * __ldg(w2 + c * 50 + t);
    // z4 = relu(.) * keep : zero where the unit was dropped or inactive
    g = (s_z4[t] > 0.f) ? g * s_keep[t] : 0.f;
    s_dz4[t] = g;
    dz4_save[b * 50 + t] = g;
    atomicAdd(db1 + t, g);
  }
  __syncthreads();
  // d(z3) = dz4 . W1 : thread per k (coalesced over k), all 50 weight loads in flight at once;
  // threads 0..63 also take k + 256
  {
    float g0 = 0.f, g1 = 0.f;
    const bool two = t < 64;
#pragma unroll
    for (int j = 0; j < 50; ++j) {
      const float d = s_dz4[j];
      g0 += d * __ldg(w1 + j * 320 + t);
      if (two) g1 += d * __ldg(w1 + j * 320 + 256 + t);
    }
    dz3[static_cast<long long>(b) * 320 + t] = g0;
    if (two) dz3[static_cast<long long>(b) * 320 + 256 + t] = g1;
  }
}

// dW1[50,320] += dz4^T[50,B] . z3[B,320]
__global__ void __launch_bounds__(320)
lenet_fc1_wgrad_kernel(const float* __restrict__ dz4, const float* __restrict__ z3, int B,
                       float* __restrict__ dw1) {
  const int n = blockIdx.x;   // 0..49
  const int k = threadIdx.x;  // 0..319
  float acc = 0.f;
  for (int b = 0; b < B; ++b) acc += __ldg(dz4 + b * 50 + n) * __ldg(z3 + static_cast<long long>(b) * 320 + k);
  dw1[n * 320 + k] += acc;
}

// ---------------------------------------------------------------------------------------------
// stage 0 backward
// ---------------------------------------------------------------------------------------------
// One CTA of 512 threads per sample.  The kernel is latency-bound (short dependent chains of
// shared-memory loads), so the three independent pieces of work run side by side on
// different warps instead of one after the other:
//   threads   0..239  conv2 input gradient: thread (half, ic, y) computes row y of dz1[ic] for
//                     output channels [10*half, 10*half+10) as a dense "full" convolution of
//                     the 8x8 gradient map, 12 outputs in registers (the sparse scatter was
//                     3x slower: divergent, 4-way bank-conflicted read-modify-writes);
//   threads 256..511  conv2 weight gradient (5000 elements, 16 live positions each) + bias;
// then all threads push the gradient through ReLU / max-pool of conv1, and 500 threads (two
// per element, even / odd positions) do the conv1 weight gradient.
constexpr int kS0BwdThreads = 512;
constexpr int kS0BwdSmemBytes =
    (784 + 1440 + 5000 + 1440 + 1440 + 320 + 1440 + 1280) * 4 + 320 + 1440 * 2 + 64;

__global__ void __launch_bounds__(kS0BwdThreads)
lenet_s0_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w2,
                    const float* __restrict__ z3, const float* __restrict__ dz3,
                    const float* __restrict__ z1_save, const unsigned char* __restrict__ idx1,
                    const unsigned char* __restrict__ idx2, const float* __restrict__ mask2,
                    float* __restrict__ dw1, float* __restrict__ db1, float* __restrict__ dw2,
                    float* __restrict__ db2) {
  extern __shared__ __align__(16) float s_all[];
  float* s_img = s_all;             // 784
  float* s_z1 = s_img + 784;        // 1440
  float* s_w2 = s_z1 + 1440;        // 5000
  float* s_dz1a = s_w2 + 5000;      // 1440
  float* s_dz1b = s_dz1a + 1440;    // 1440
  float* s_g2 = s_dz1b + 1440;      // 320: gradient of the one live position per conv2 pooling window
  float* s_g1 = s_g2 + 320;         // 1440
  float* s_dc2 = s_g1 + 1440;       // 1280: dense gradient w.r.t. the conv2 output (20 x 8 x 8)
  unsigned short* s_p1 = reinterpret_cast<unsigned short*>(s_dc2 + 1280);  // live pos in the 24x24 map
  unsigned char* s_p2 = reinterpret_cast<unsigned char*>(s_p1 + 1440);    // live pos in the 8x8 map
  const int b = blockIdx.x, t = threadIdx.x;
  {
    float rw2[10], rimg[2], rz1[3];
#pragma unroll
    for (int j = 0; j < 10; ++j) rw2[j] = (t + j * kS0BwdThreads < 5000) ? w2[t + j * kS0BwdThreads] : 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j)
      rimg[j] = (t + j * kS0BwdThreads < 784) ? x[static_cast<long long>(b) * 784 + t + j * kS0BwdThreads] : 0.f;
#pragma unroll
    for (int j = 0; j < 3; ++j)
      rz1[j] = (t + j * kS0BwdThreads < 1440) ? z1_save[static_cast<long long>(b) * 1440 + t + j * kS0BwdThreads] : 0.f;
#pragma unroll
    for (int j = 0; j < 10; ++j)
      if (t + j * kS0BwdThreads < 5000) s_w2[t + j * kS0BwdThreads] = rw2[j];
#pragma unroll
    for (int j = 0; j < 2; ++j)
      if (t + j * kS0BwdThreads < 784) s_img[t + j * kS0BwdThreads] = rimg[j];
#pragma unroll
    for (int j = 0; j < 3; ++j)
      if (t + j * kS0BwdThreads < 1440) s_z1[t + j * kS0BwdThreads] = rz1[j];
  }
  for (int i = t; i < 1280; i += kS0BwdThreads) s_dc2[i] = 0.f;
  float my_g = 0.f;
  int my_pos = 0;
  if (t < 320) {
    const int oc = t >> 4, r = t & 15, py = r >> 2, px = r & 3;
    const long long gi = static_cast<long long>(b) * 320 + t;
    const float m = mask2[b * 20 + oc];
    // z3 = relu(pool(c2) * m): live iff z3 > 0; d/dc2 = m at the arg-max position
    my_g = (z3[gi] > 0.f) ? dz3[gi] * m : 0.f;
    const int bi = idx2[gi];
    my_pos = (2 * py + (bi >> 1)) * 8 + 2 * px + (bi & 1);
    s_g2[t] = my_g;
    s_p2[t] = static_cast<unsigned char>(my_pos);
  }
  __syncthreads();
  if (t < 320) s_dc2[(t >> 4) * 64 + my_pos] = my_g;   // one live position per pooling window
  __syncthreads();
  if (t < 240) {
    const int half = t / 120, u = t - half * 120, ic = u / 12, y = u - ic * 12;
    float acc[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = 0.f;
    for (int oc = half * 10; oc < half * 10 + 10; ++oc) {
#pragma unroll
      for (int ky = 0; ky < 5; ++ky) {
        const int yy = y - ky;
        if (yy >= 0 && yy < 8) {
          const float* g = s_dc2 + oc * 64 + yy * 8;
          const float* w = s_w2 + (oc * 10 + ic) * 25 + ky * 5;
          float gr[8], wr[5];
#pragma unroll
          for (int i = 0; i < 8; ++i) gr[i] = g[i];
#pragma unroll
          for (int i = 0; i < 5; ++i) wr[i] = w[i];
#pragma unroll
          for (int xx = 0; xx < 8; ++xx)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) acc[xx + kx] += gr[xx] * wr[kx];
        }
      }
    }
    float* row = (half ? s_dz1b : s_dz1a) + ic * 144 + y * 12;
#pragma unroll
    for (int i = 0; i < 12; ++i) row[i] = acc[i];
  } else if (t >= 256) {
    const int u = t - 256;
    if (u < 20) {
      float sum = 0.f;
#pragma unroll
      for (int p = 0; p < 16; ++p) sum += s_g2[u * 16 + p];
      atomicAdd(db2 + u, sum);
    }
    // dW2[oc][ic][ky][kx] = sum_p g2[oc][p] * z1[ic][y_p+ky][x_p+kx]
    for (int i = u; i < 5000; i += 256) {
      const int oc = i / 250, r = i - oc * 250, ic = r / 25, kk = r - ic * 25, ky = kk / 5, kx = kk - ky * 5;
      float acc = 0.f;
#pragma unroll
      for (int p = 0; p < 16; ++p) {
        const int pos = s_p2[oc * 16 + p];
        acc += s_g2[oc * 16 + p] * s_z1[ic * 144 + ((pos >> 3) + ky) * 12 + (pos & 7) + kx];
      }
      atomicAdd(dw2 + i, acc);
    }
  }
  __syncthreads();
  // through ReLU + maxpool of conv1: one live position per pooled output
  for (int o = t; o < 1440; o += kS0BwdThreads) {
    const int r = o % 144, py = r / 12, px = r - py * 12;
    s_g1[o] = (s_z1[o] > 0.f) ? s_dz1a[o] + s_dz1b[o] : 0.f;
    const int bi = idx1[static_cast<long long>(b) * 1440 + o];
    s_p1[o] = static_cast<unsigned short>((2 * py + (bi >> 1)) * 24 + 2 * px + (bi & 1));
  }
  __syncthreads();
  if (t < 500) {
    // dW1[oc][ky][kx] = sum_p g1[oc][p] * img[y_p+ky][x_p+kx]; two threads per element
    const int o = t >> 1, h = t & 1;
    const int oc = o / 25, kk = o - oc * 25, ky = kk / 5, kx = kk - ky * 5;
    float acc = 0.f;
#pragma unroll 8
    for (int p = h; p < 144; p += 2) {
      const int pos = s_p1[oc * 144 + p];
      acc += s_g1[oc * 144 + p] * s_img[(pos / 24 + ky) * 28 + pos % 24 + kx];
    }
    atomicAdd(dw1 + o, acc);
  } else if (t < 510) {
    const int oc = t - 500;
    float sum = 0.f;
#pragma unroll 8
    for (int p = 0; p < 144; ++p) sum += s_g1[oc * 144 + p];
    atomicAdd(db1 + oc, sum);
  }
}

}  // namespace sdml

using namespace sdml;

// x [B,1,28,28] -> z3 [B,320]; saves z1 [B,1440], idx1 [B,1440] (u8), idx2 [B,320] (u8),
// mask2 [B,20] for the backward kernel.  drop_p = 0 disables Dropout2d (eval).
extern "C" int sdml_lenet_s0_fwd(const float* x, const float* w1, const float* b1, const float* w2,
                                 const float* b2, int B, float* z3, float* z1_save, void* idx1,
                                 void* idx2, float* mask2, float drop_p, long long seed,
                                 const long long* offset_dev, long long offset_imm,
                                 float* zero_this, long long* bump_counter, long long bump_by,
                                 cudaStream_t stream) {
  if (B <= 0) return set_error("sdml_lenet_s0_fwd: empty batch");
  lenet_s0_fwd_kernel<<<B, kS0Threads, 0, stream>>>(
      x, w1, b1, w2, b2, z3, z1_save, static_cast<unsigned char*>(idx1),
      static_cast<unsigned char*>(idx2), mask2, drop_p, (unsigned long long)seed, offset_dev,
      offset_imm, zero_this, bump_counter, bump_by);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// loss_sum += loss_scale * sum CE.  bump_counter / bump_by: optional device counter advanced by
// this launch (the RNG position the next stage-0 forward reads).  wgrad_stream: optional stream
// for the fc1 weight-gradient kernel (NULL = `stream`).
// train != 0: gradients dW1/db1/dW2/db2 (accumulated) and dz3 [B,320]
// (gradient for stage 0), grad_scale multiplies dlogits (1 / global batch for the mean loss).
// dz4_scratch: [B,50] workspace handed from the per-sample kernel to the fc1 wgrad kernel.
// train == 0: loss_sum / correct only.
extern "C" int sdml_lenet_s1(const float* z3, const void* targets, const float* w1,
                             const float* b1, const float* w2, const float* b2, int B, int train,
                             float grad_scale, float drop_p, long long seed,
                             const long long* offset_dev, long long offset_imm, float* loss_sum,
                             float loss_scale, int* correct, float* dz3, float* dz4_scratch,
                             float* dw1, float* db1, float* dw2, float* db2,
                             long long* bump_counter, long long bump_by, cudaStream_t stream,
                             cudaStream_t wgrad_stream) {
  if (B <= 0) return set_error("sdml_lenet_s1: empty batch");
  lenet_s1_kernel<<<B, kS1Threads, 0, stream>>>(
      z3, static_cast<const long long*>(targets), w1, b1, w2, b2, train, grad_scale, drop_p,
      (unsigned long long)seed, offset_dev, offset_imm, loss_sum, loss_scale, correct, dz3,
      dz4_scratch, dw2, db2, db1, bump_counter, bump_by);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  if (train) {
    // dW1 only depends on this kernel's dz4: on a side stream (ordered after `stream` by an
    // event here; the caller joins it back before the optimizer) it overlaps the stage-0 backward
    if (wgrad_stream && wgrad_stream != stream) {
      static cudaEvent_t ev[64] = {};
      int dev = 0;
      SDML_CUDA_OK(cudaGetDevice(&dev));
      if (!ev[dev & 63]) SDML_CUDA_OK(cudaEventCreateWithFlags(&ev[dev & 63], cudaEventDisableTiming));
      SDML_CUDA_OK(cudaEventRecord(ev[dev & 63], stream));
      SDML_CUDA_OK(cudaStreamWaitEvent(wgrad_stream, ev[dev & 63], 0));
    }
    lenet_fc1_wgrad_kernel<<<50, 320, 0, wgrad_stream ? wgrad_stream : stream>>>(dz4_scratch, z3, B, dw1);
    SDML_CUDA_OK(cudaGetLastError());
    count_launch();
  }
  return 0;
}

extern "C" int sdml_lenet_s0_bwd(const float* x, const float* w2, const float* z3, const float* dz3,
                                 const float* z1_save, const void* idx1, const void* idx2,
                                 const float* mask2, int B, float* dw1, float* db1, float* dw2,
                                 float* db2, cudaStream_t stream) {
  if (B <= 0) return set_error("sdml_lenet_s0_bwd: empty batch");
  static std::atomic<bool> configured[64];
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    SDML_CUDA_OK(cudaFuncSetAttribute(lenet_s0_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      kS0BwdSmemBytes));
    configured[dev & 63] = true;
  }
  lenet_s0_bwd_kernel<<<B, kS0BwdThreads, kS0BwdSmemBytes, stream>>>(
      x, w2, z3, dz3, z1_save, static_cast<const unsigned char*>(idx1),
      static_cast<const unsigned char*>(idx2), mask2, dw1, db1, dw2, db2);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}
