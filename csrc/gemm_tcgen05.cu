// bf16 GEMM on the 5th-gen tensor cores (tcgen05 + TMEM), operands staged by TMA.
//
//   C[M,N] = epilogue( alpha * op(A)[M,K] . op(B)[N,K]^T )
//
// One persistent CTA per SM, warp-specialised:
//   warp 0      : TMA producer (one lane) -> STAGES-deep smem ring (128B swizzle)
//   warp 1      : TMEM allocator + single-lane tcgen05.mma issuer
//   warps 2..5  : epilogue, tcgen05.ld the fp32 accumulator out of TMEM, apply the
//                 fused epilogue and store; TMEM is double-buffered so the epilogue
//                 of tile i overlaps the main loop of tile i+1.
//
// Both operands may be K-major or MN-major (instruction-descriptor transpose bits), so
// the same kernel serves the three GEMMs of a Linear layer without any transposed copy:
//   forward  Y  = X  . W^T   (A = X   K-major, B = W K-major)
//   dgrad    dX = dZ . W     (A = dZ  K-major, B = W MN-major)
//   wgrad    dW = dZ^T . X   (A = dZ MN-major, B = X MN-major)
//
// Fused epilogues: + bias[n], ReLU, ReLU-backward mask (aux > 0), fp32 accumulate
// (C += ...), bf16 or fp32 output.  This replaces K7/K9 of the reference
// (/root/reference/simple_distributed.py:63-64,75-77 run through ATen addmm + relu).
#include "sdml_ptx.cuh"
#include "sdml_host.h"

namespace sdml {

struct GemmParams {
  int M, N, K;
  void* C;
  long long ldc;
  const float* bias;          // [N] or nullptr
  const __nv_bfloat16* aux;   // [M, ld_aux] or nullptr : multiply by (aux > 0)
  long long ld_aux;
  float alpha;
  int relu;
  int accumulate;  // C (fp32) += result
  int out_fp32;
};

constexpr int kBM = 128;
constexpr int kBK = 64;   // 64 bf16 = one 128-byte swizzle row
constexpr int kUK = 16;   // K per tcgen05.mma for 16-bit inputs
constexpr int kGroupM = 16;
constexpr int kGemmThreads = 192;

template <int GROUP_M = kGroupM>
__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk,
                                            int& n_blk) {
  // grouped ordering: a wave of CTAs covers a roughly square block of tiles so that
  // operand k-slices are shared through the 126 MB L2.
  const int per_group = GROUP_M * num_n;
  const int group = tile / per_group;
  const int first_m = group * GROUP_M;
  const int gsize = min(num_m - first_m, GROUP_M);
  const int in_group = tile - group * per_group;
  m_blk = first_m + in_group % gsize;
  n_blk = in_group / gsize;
}

// Epilogue for one 32-column chunk held in registers (one accumulator row per thread).
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t (&r)[32], int row,
                                               int col0, bool vec_ok_bf16, bool vec_ok_f32,
                                               bool aux_vec_ok) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
  const bool full = (col0 + 32 <= p.N);
  if (p.bias) {
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i));
        v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (col0 + i < p.N) v[i] += __ldg(p.bias + col0 + i);
    }
  }
  if (p.relu) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
  }
  if (p.aux) {
    const __nv_bfloat16* ap = p.aux + static_cast<long long>(row) * p.ld_aux + col0;
    if (full && aux_vec_ok) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(ap + i));
        const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (!(bf16_lo(w[j]) > 0.f)) v[i + 2 * j] = 0.f;
          if (!(bf16_hi(w[j]) > 0.f)) v[i + 2 * j + 1] = 0.f;
        }
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (col0 + i < p.N && !(__bfloat162float(ap[i]) > 0.f)) v[i] = 0.f;
    }
  }
  if (p.out_fp32) {
    float* cp = reinterpret_cast<float*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
    if (full && vec_ok_f32) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 o = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        if (p.accumulate) {
          const float4 old = *reinterpret_cast<const float4*>(cp + i);
          o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
        }
        *reinterpret_cast<float4*>(cp + i) = o;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (col0 + i < p.N) cp[i] = p.accumulate ? cp[i] + v[i] : v[i];
    }
  } else {
    __nv_bfloat16* cp =
        reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
    if (full && vec_ok_bf16) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 o;
        o.x = pack_bf16x2(v[i], v[i + 1]);
        o.y = pack_bf16x2(v[i + 2], v[i + 3]);
        o.z = pack_bf16x2(v[i + 4], v[i + 5]);
        o.w = pack_bf16x2(v[i + 6], v[i + 7]);
        *reinterpret_cast<uint4*>(cp + i) = o;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (col0 + i < p.N) cp[i] = __float2bfloat16(v[i]);
    }
  }
}

// alpha / bias / ReLU / ReLU-backward mask on one 32-column chunk (registers only).
__device__ __forceinline__ void epilogue_math(const GemmParams& p, const uint32_t (&r)[32], int row,
                                              bool row_ok, int col0, bool aux_vec_ok, float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
  const bool full = (col0 + 32 <= p.N);
  if (p.bias) {
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i));
        v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (col0 + i < p.N) v[i] += __ldg(p.bias + col0 + i);
    }
  }
  if (p.relu) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
  }
  if (p.aux && row_ok) {
    const __nv_bfloat16* ap = p.aux + static_cast<long long>(row) * p.ld_aux + col0;
    if (full && aux_vec_ok) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(ap + i));
        const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (!(bf16_lo(w[j]) > 0.f)) v[i + 2 * j] = 0.f;
          if (!(bf16_hi(w[j]) > 0.f)) v[i + 2 * j + 1] = 0.f;
        }
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (col0 + i < p.N && !(__bfloat162float(ap[i]) > 0.f)) v[i] = 0.f;
    }
  }
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmParams p) {
  constexpr uint32_t A_BYTES = kBM * kBK * 2;
  constexpr uint32_t B_BYTES = BN * kBK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  static_assert(TMEM_COLS <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "TMEM columns");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t smem_base = (raw_addr + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw_addr));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int num_m = (p.M + kBM - 1) / kBM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = (p.K + kBK - 1) / kBK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_slot, TMEM_COLS);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, m_blk, n_blk);
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t fb = full_bar(stage);
          const uint32_t a_dst = smem_base + stage * STAGE_BYTES;
          const uint32_t b_dst = a_dst + A_BYTES;
          if (elect_one_sync()) {
            mbar_arrive_expect_tx(fb, STAGE_BYTES);
            if constexpr (!A_MN) {
              tma_load_2d(a_dst, &tmA, fb, kb * kBK, m_blk * kBM);
            } else {
#pragma unroll
              for (int j = 0; j < kBM / 64; ++j)
                tma_load_2d(a_dst + j * (kBK * 128), &tmA, fb, m_blk * kBM + j * 64, kb * kBK);
            }
            if constexpr (!B_MN) {
              tma_load_2d(b_dst, &tmB, fb, kb * kBK, n_blk * BN);
            } else {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                tma_load_2d(b_dst + j * (kBK * 128), &tmB, fb, n_blk * BN + j * 64, kb * kBK);
            }
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc_f16(1u, A_MN ? 1u : 0u, B_MN ? 1u : 0u, kBM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int tile_iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tile_iter) {
      const int acc = tile_iter & 1;
      const uint32_t acc_phase = (tile_iter >> 1) & 1;
      mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_k; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tcgen05_fence_after();
        const uint32_t a_base = smem_base + stage * STAGE_BYTES;
        const uint32_t b_base = a_base + A_BYTES;
        if (elect_one_sync()) {
#pragma unroll
          for (int k = 0; k < kBK / kUK; ++k) {
            // K-major : rows of 128 B, 8-row atoms 1024 B apart (SBO); step 32 B per UMMA_K.
            // MN-major: [k][64 mn] boxes, 8-k atoms 1024 B apart (SBO), next 64 mn (LBO)
            //           one box (kBK*128 B) further; step 16 k-rows = 2048 B per UMMA_K.
            const uint64_t da = A_MN ? make_smem_desc(a_base + k * 2048, kBK * 128, 1024, kSwizzle128B)
                                     : make_smem_desc(a_base + k * 32, 16, 1024, kSwizzle128B);
            const uint64_t db = B_MN ? make_smem_desc(b_base + k * 2048, kBK * 128, 1024, kSwizzle128B)
                                     : make_smem_desc(b_base + k * 32, 16, 1024, kSwizzle128B);
            umma_f16<1>(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));            // smem slot free once these MMAs retire
          if (kb == num_k - 1) umma_commit(tfull_bar(acc));  // accumulator complete
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int row_in_tile = quad * 32 + lane;
    int tile_iter = 0;
    const bool vec_ok_bf16 = ((p.ldc & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
    const bool vec_ok_f32 = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
    const bool aux_vec_ok =
        p.aux && ((p.ld_aux & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.aux) & 15) == 0);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tile_iter) {
      int m_blk, n_blk;
      tile_coords(tile, num_m, num_n, m_blk, n_blk);
      const int acc = tile_iter & 1;
      const uint32_t acc_phase = (tile_iter >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tcgen05_fence_after();
      const int row = m_blk * kBM + row_in_tile;
      const bool row_ok = row < p.M;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n_blk * BN + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN + c * 32, r);
        tmem_ld_wait();
        if (row_ok) epilogue_chunk(p, r, row, col0, vec_ok_bf16, vec_ok_f32, aux_vec_ok);
      }
      // all tcgen05.ld of this warp are complete (wait::ld above): hand the buffer back
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, TMEM_COLS);
  }
}

// ----------------------------------------------------------------------------
// 2-CTA variant: a cluster of two CTAs (one TPC) computes a 256 x BN tile with
// tcgen05.mma.cta_group::2.  Each CTA stages its own 128 rows of A and HALF of the B tile
// (BN/2 rows); the pair's tensor cores read both halves, so per-SM shared-memory and L2
// operand traffic for B is halved and the smem ring is 6 stages deep instead of 4.  Only
// the leader CTA (cluster rank 0) issues MMAs; its tcgen05.commit multicasts to the
// barriers of both CTAs.  Each CTA drains its own 128 accumulator rows from its own TMEM.
// ----------------------------------------------------------------------------
template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tmA,
                      const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  constexpr int BNH = BN / 2;  // B rows staged per CTA
  constexpr uint32_t A_BYTES = kBM * kBK * 2;
  constexpr uint32_t B_BYTES = BNH * kBK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  static_assert(TMEM_COLS <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "TMEM columns");
  static_assert(BNH % 64 == 0, "half B tile must be whole 64-wide boxes");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t smem_base = (raw_addr + 1023u) & ~1023u;
  // epilogue staging: 4 warps x 2 buffers x (32 rows x 128 B), 128B-swizzled for TMA stores
  constexpr uint32_t EPI_BUF_BYTES = 32 * 128;
  constexpr uint32_t EPI_BYTES = 4 * 2 * EPI_BUF_BYTES;
  const uint32_t epi_base = smem_base + STAGES * STAGE_BYTES;
  const uint32_t bar_base = epi_base + EPI_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw_addr));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  constexpr int BM2 = 2 * kBM;
  const int num_m = (p.M + BM2 - 1) / BM2;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = (p.K + kBK - 1) / kBK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);   // leader: its own arrive.expect_tx (bytes of BOTH CTAs)
      mbar_init(empty_bar(s), 1);  // one multicast commit per use
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 8);  // 4 epilogue warps x 2 CTAs (only the leader's is waited on)
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<2>(tmem_slot, TMEM_COLS);
  }
  tcgen05_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / TMA signal
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int m_blk, n_blk;
        tile_coords<8>(tile, num_m, num_n, m_blk, n_blk);
        const int m0 = m_blk * BM2 + static_cast<int>(cta_rank) * kBM;
        const int n0 = n_blk * BN + static_cast<int>(cta_rank) * BNH;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          // transaction bytes of both CTAs are counted on the LEADER's full barrier
          const uint32_t fb = mapa_shared(full_bar(stage), 0);
          const uint32_t a_dst = smem_base + stage * STAGE_BYTES;
          const uint32_t b_dst = a_dst + A_BYTES;
          if (elect_one_sync()) {
            if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * STAGE_BYTES);
            if constexpr (!A_MN) {
              tma_load_2d_2sm(a_dst, &tmA, fb, kb * kBK, m0);
            } else {
#pragma unroll
              for (int j = 0; j < kBM / 64; ++j)
                tma_load_2d_2sm(a_dst + j * (kBK * 128), &tmA, fb, m0 + j * 64, kb * kBK);
            }
            if constexpr (!B_MN) {
              tma_load_2d_2sm(b_dst, &tmB, fb, kb * kBK, n0);
            } else {
#pragma unroll
              for (int j = 0; j < BNH / 64; ++j)
                tma_load_2d_2sm(b_dst + j * (kBK * 128), &tmB, fb, n0 + j * 64, kb * kBK);
            }
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader) {
      constexpr uint32_t idesc = make_idesc_f16(1u, A_MN ? 1u : 0u, B_MN ? 1u : 0u, BM2, BN);
      int stage = 0;
      uint32_t phase = 0;
      int tile_iter = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++tile_iter) {
        const int acc = tile_iter & 1;
        const uint32_t acc_phase = (tile_iter >> 1) & 1;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t a_base = smem_base + stage * STAGE_BYTES;
          const uint32_t b_base = a_base + A_BYTES;
          if (elect_one_sync()) {
#pragma unroll
            for (int k = 0; k < kBK / kUK; ++k) {
              const uint64_t da = A_MN ? make_smem_desc(a_base + k * 2048, kBK * 128, 1024, kSwizzle128B)
                                       : make_smem_desc(a_base + k * 32, 16, 1024, kSwizzle128B);
              const uint64_t db = B_MN ? make_smem_desc(b_base + k * 2048, kBK * 128, 1024, kSwizzle128B)
                                       : make_smem_desc(b_base + k * 32, 16, 1024, kSwizzle128B);
              umma_f16<2>(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit_2sm(empty_bar(stage), 0x3);  // frees the stage in both CTAs
            if (kb == num_k - 1) umma_commit_2sm(tfull_bar(acc), 0x3);
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5, both CTAs) =====================
    // TMEM -> registers -> (alpha, bias, ReLU, mask) -> 128B-swizzled smem -> TMA store, or
    // TMA reduce-add for fp32 accumulation (wgrad over micro-batches): the tile leaves the SM
    // as whole 128-byte lines and C is never read back by the SM.
    const int quad = warp & 3;
    const uint32_t my_epi = epi_base + static_cast<uint32_t>(warp - 2) * 2 * EPI_BUF_BYTES;
    const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;
    const uint32_t swz = static_cast<uint32_t>(lane & 7);
    const bool aux_vec_ok =
        p.aux && ((p.ld_aux & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.aux) & 15) == 0);
    int tile_iter = 0;
    uint32_t buf_it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++tile_iter) {
      int m_blk, n_blk;
      tile_coords<8>(tile, num_m, num_n, m_blk, n_blk);
      const int acc = tile_iter & 1;
      const uint32_t acc_phase = (tile_iter >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tcgen05_fence_after();
      const int row0 = m_blk * BM2 + static_cast<int>(cta_rank) * kBM + quad * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
      const uint32_t tm = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN;
      if (p.out_fp32) {
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          const int col0 = n_blk * BN + c;
          if (col0 >= p.N || row0 >= p.M) break;  // warp-uniform
          uint32_t r[32];
          tmem_ld_32x32b_x32(tm + c, r);
          const uint32_t buf = my_epi + (buf_it & 1u) * EPI_BUF_BYTES;
          if (lane == 0) tma_store_wait_read<1>();  // the store that last used this buffer has read it
          __syncwarp();
          tmem_ld_wait();
          float v[32];
          epilogue_math(p, r, row, row_ok, col0, aux_vec_ok, v);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(buf + row_off + ((static_cast<uint32_t>(j) ^ swz) << 4),
                         __float_as_uint(v[4 * j]), __float_as_uint(v[4 * j + 1]),
                         __float_as_uint(v[4 * j + 2]), __float_as_uint(v[4 * j + 3]));
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (p.accumulate) tma_reduce_add_2d(&tmC, buf, col0, row0);
            else tma_store_2d(&tmC, buf, col0, row0);
            tma_store_commit();
          }
          ++buf_it;
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN; c += 64) {
          const int col0 = n_blk * BN + c;
          if (col0 >= p.N || row0 >= p.M) break;
          uint32_t r0[32], r1[32];
          tmem_ld_32x32b_x32(tm + c, r0);
          tmem_ld_32x32b_x32(tm + c + 32, r1);
          const uint32_t buf = my_epi + (buf_it & 1u) * EPI_BUF_BYTES;
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          tmem_ld_wait();
          float v0[32], v1[32];
          epilogue_math(p, r0, row, row_ok, col0, aux_vec_ok, v0);
          epilogue_math(p, r1, row, row_ok, col0 + 32, aux_vec_ok, v1);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            st_shared_v4(buf + row_off + ((static_cast<uint32_t>(j) ^ swz) << 4),
                         pack_bf16x2(v0[8 * j], v0[8 * j + 1]), pack_bf16x2(v0[8 * j + 2], v0[8 * j + 3]),
                         pack_bf16x2(v0[8 * j + 4], v0[8 * j + 5]), pack_bf16x2(v0[8 * j + 6], v0[8 * j + 7]));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            st_shared_v4(buf + row_off + ((static_cast<uint32_t>(j + 4) ^ swz) << 4),
                         pack_bf16x2(v1[8 * j], v1[8 * j + 1]), pack_bf16x2(v1[8 * j + 2], v1[8 * j + 3]),
                         pack_bf16x2(v1[8 * j + 4], v1[8 * j + 5]), pack_bf16x2(v1[8 * j + 6], v1[8 * j + 7]));
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmC, buf, col0, row0);
            tma_store_commit();
          }
          ++buf_it;
        }
      }
      // every tcgen05.ld of this warp has completed (wait::ld): hand the accumulator back
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tempty_bar(acc), 0);  // leader's barrier, both CTAs arrive
    }
    if (lane == 0) tma_store_wait<0>();  // smem must stay valid until the last store has read it
  }

  tcgen05_fence_before();
  cluster_sync_all();  // the peer's smem / barriers stay valid until every MMA and commit retired
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<2>(tmem_base, TMEM_COLS);
  }
}

// ----------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------
template <int BN, int STAGES, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                       cudaStream_t stream) {
  constexpr size_t smem = size_t(STAGES) * (kBM * kBK * 2 + BN * kBK * 2) + 1024 + 256;
  auto kern = gemm_bf16_kernel<BN, STAGES, A_MN, B_MN>;
  static bool configured[64] = {};
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    SDML_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev & 63] = true;
  }
  const int num_tiles = ((p.M + kBM - 1) / kBM) * ((p.N + BN - 1) / BN);
  const int grid = num_tiles < num_sms(dev) ? num_tiles : num_sms(dev);
  kern<<<grid, kGemmThreads, smem, stream>>>(tmA, tmB, p);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
static int launch_gemm_2cta(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                            const GemmParams& p, cudaStream_t stream) {
  constexpr size_t smem =
      size_t(STAGES) * (kBM * kBK * 2 + (BN / 2) * kBK * 2) + 4 * 2 * 32 * 128 + 1024 + 256;
  auto kern = gemm_bf16_2cta_kernel<BN, STAGES, A_MN, B_MN>;
  static bool configured[64] = {};
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    SDML_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev & 63] = true;
  }
  const int num_tiles = ((p.M + 2 * kBM - 1) / (2 * kBM)) * ((p.N + BN - 1) / BN);
  const int max_clusters = num_sms(dev) / 2;
  const int clusters = num_tiles < max_clusters ? num_tiles : max_clusters;
  kern<<<2 * clusters, kGemmThreads, smem, stream>>>(tmA, tmB, tmC, p);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

template <int BN, int STAGES>
static int dispatch_major_2cta(int a_mn, int b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB,
                               const CUtensorMap& tmC, const GemmParams& p, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm_2cta<BN, STAGES, false, false>(tmA, tmB, tmC, p, stream);
  if (!a_mn && b_mn) return launch_gemm_2cta<BN, STAGES, false, true>(tmA, tmB, tmC, p, stream);
  if (a_mn && !b_mn) return launch_gemm_2cta<BN, STAGES, true, false>(tmA, tmB, tmC, p, stream);
  return launch_gemm_2cta<BN, STAGES, true, true>(tmA, tmB, tmC, p, stream);
}

template <int BN, int STAGES>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB,
                          const GemmParams& p, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm<BN, STAGES, false, false>(tmA, tmB, p, stream);
  if (!a_mn && b_mn) return launch_gemm<BN, STAGES, false, true>(tmA, tmB, p, stream);
  if (a_mn && !b_mn) return launch_gemm<BN, STAGES, true, false>(tmA, tmB, p, stream);
  return launch_gemm<BN, STAGES, true, true>(tmA, tmB, p, stream);
}

}  // namespace sdml

using namespace sdml;

extern "C" int sdml_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                              long long lda, long long ldb, long long ldc, int a_mn, int b_mn,
                              const float* bias, const void* aux, long long ld_aux, int relu,
                              int accumulate, int out_fp32, float alpha, int bn,
                              cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return set_error("sdml_gemm_bf16: empty problem");
  if ((lda & 7) || (ldb & 7)) return set_error("sdml_gemm_bf16: lda/ldb must be multiples of 8 (TMA 16 B stride)");
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return set_error("sdml_gemm_bf16: A/B must be 16-byte aligned");
  if (accumulate && !out_fp32) return set_error("sdml_gemm_bf16: accumulate needs fp32 output");
  // bn: 64/128/256 = single-CTA tiles of 128 x bn; 1256 / 1128 = CTA-pair tiles of 256 x 256 / 256 x 128
  // (tcgen05 cta_group::2); 0 = choose.
  if (bn == 0) {
    if (N >= 256 && M >= 512) bn = 1256;
    else bn = (N >= 256) ? 256 : (N > 64 ? 128 : 64);
  }
  // the CTA-pair kernels write C through TMA (16-byte aligned rows); otherwise use 128 x bn tiles
  const long long c_row_bytes = ldc * (out_fp32 ? 4 : 2);
  const bool c_tma_ok = (c_row_bytes % 16 == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
  if (bn > 1000 && !c_tma_ok) bn = (bn - 1000 >= 256 && N >= 256) ? 256 : (N > 64 ? 128 : 64);
  const bool two_cta = bn > 1000;
  const int bn_tile = two_cta ? bn - 1000 : bn;

  CUtensorMap tmA, tmB;
  // K-major operand: global [rows, K] (K contiguous), box {64 k, rows_per_tile}
  // MN-major operand: global [K, rows] (rows contiguous), box {64 rows, 64 k}
  int rc;
  if (!a_mn) rc = make_tmap_2d(&tmA, A, TMAP_BF16, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kBK, kBM, TMAP_SWIZZLE_128B);
  else       rc = make_tmap_2d(&tmA, A, TMAP_BF16, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, kBK, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d(&tmB, B, TMAP_BF16, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, kBK,
                               (uint32_t)(two_cta ? bn_tile / 2 : bn_tile), TMAP_SWIZZLE_128B);
  else       rc = make_tmap_2d(&tmB, B, TMAP_BF16, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, kBK, TMAP_SWIZZLE_128B);
  if (rc) return rc;

  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.C = C; p.ldc = ldc;
  p.bias = bias;
  p.aux = reinterpret_cast<const __nv_bfloat16*>(aux);
  p.ld_aux = ld_aux;
  p.alpha = alpha;
  p.relu = relu; p.accumulate = accumulate; p.out_fp32 = out_fp32;

  CUtensorMap tmC;
  if (two_cta) {
    rc = out_fp32 ? make_tmap_2d(&tmC, C, TMAP_F32, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 32, 32, TMAP_SWIZZLE_128B)
                  : make_tmap_2d(&tmC, C, TMAP_BF16, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 64, 32, TMAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  switch (bn) {
    case 1256: return dispatch_major_2cta<256, 6>(a_mn, b_mn, tmA, tmB, tmC, p, stream);
    case 1128: return dispatch_major_2cta<128, 8>(a_mn, b_mn, tmA, tmB, tmC, p, stream);
    case 256: return dispatch_major<256, 4>(a_mn, b_mn, tmA, tmB, p, stream);
    case 128: return dispatch_major<128, 6>(a_mn, b_mn, tmA, tmB, p, stream);
    case 64:  return dispatch_major<64, 8>(a_mn, b_mn, tmA, tmB, p, stream);
    default:  return set_error("sdml_gemm_bf16: bn must be 0, 64, 128, 256, 1128 or 1256");
  }
}
