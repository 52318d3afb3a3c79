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
#include <atomic>

#include <string.h>

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

// ----------------------------------------------------------------------------
// Communication hooks of the CTA-pair kernel (tensor-parallel layers)
// ----------------------------------------------------------------------------
// all-gather -> GEMM : A is row-sharded over `ag_world` GPUs.  Every rank keeps the full
//   [M, K] operand buffer in symmetric memory with only its own rows valid.  Two extra
//   warps per CTA pull the peers' rows over NVLink into the local buffer chunk by chunk and
//   count completed chunks; the TMA producer starts with the M-tiles of the local shard and
//   only waits for a chunk's counter right before loading tiles that need it, so the
//   transfer hides behind the math of earlier tiles.
// GEMM -> reduce-scatter : C is a partial sum; every 32-row store box is TMA-stored (bf16)
//   straight out of the epilogue into the landing slot this rank owns in the memory of the
//   GPU that owns those rows -- across NVSwitch for remote owners, no atomics -- and a
//   per-warp completion count tells each owner when all contributions have landed;
//   sdml_rs_finalize then sums the slots in fp32 and applies bias / ReLU / mask.
//   (A first version TMA-reduce-added fp32 tiles into one buffer: correct, but the L2 /
//   NVLink atomic rate capped it at 51 % of the roofline on 2 x B200.)
constexpr int kCommMaxRanks = 8;

struct CommParams {
  // ---- all-gather of A rows ----
  int ag_world, ag_rank;
  int ag_rows_per_rank;                          // M / ag_world (multiple of 256)
  int ag_chunk_rows;                             // rows per completion chunk (multiple of 256)
  long long ag_ld;                               // leading dimension of the operand buffer (elements)
  int ag_k;                                      // columns to copy (elements, multiple of 8)
  const __nv_bfloat16* ag_src[kCommMaxRanks];    // rank r's operand buffer (peer-mapped)
  __nv_bfloat16* ag_dst;                         // local operand buffer (what TMA reads)
  uint32_t* ag_flags[kCommMaxRanks];             // rank r's flag array; we publish flags[r][ag_rank]
  uint32_t* ag_count;                            // local chunk counters [ag_world][chunks]
  int ag_max_chunks;
  uint32_t epoch;
  // ---- reduce-scatter of C rows ----
  int rs_world, rs_rank;
  int rs_rows_per_rank;                          // M / rs_world (multiple of 32)
  uint32_t* rs_done[kCommMaxRanks];              // rank r's completion counter (peer-mapped)
  WaitGuard guard;                               // deadline of the cross-GPU flag waits
};

struct TmapArray {
  CUtensorMap m[kCommMaxRanks];
};

struct PtrArray {
  void* p[kCommMaxRanks];
};

__device__ __forceinline__ int comm_rotate_m(const CommParams* cp, int m_blk, int num_m) {
  if (cp->ag_world <= 1) return m_blk;
  const int per_rank = num_m / cp->ag_world;       // M-blocks per shard
  int m = m_blk + cp->ag_rank * per_rank;          // own shard first, then rank+1, ...
  return m >= num_m ? m - num_m : m;
}

// Publish "my shard is complete" (it was produced by earlier work on this stream) to all peers.
__device__ __forceinline__ void comm_announce(const CommParams* cp) {
  if (cp->ag_world > 1 && blockIdx.x == 0 && threadIdx.x < static_cast<unsigned>(cp->ag_world) &&
      static_cast<int>(threadIdx.x) != cp->ag_rank) {
    __threadfence_system();
    st_release_sys(cp->ag_flags[threadIdx.x] + cp->ag_rank, cp->epoch);
  }
}

__device__ __forceinline__ uint4 ld_peer_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

// Executed by 64 threads of every CTA.  `t` = thread index within the CTA's gather group.
__device__ __forceinline__ void comm_gather_rows(const CommParams* cp, int t, int group) {
  if (cp->ag_world <= 1) return;
  const long long gtid = static_cast<long long>(blockIdx.x) * group + t;
  const long long gthreads = static_cast<long long>(gridDim.x) * group;
  const int vec_per_row = cp->ag_k / 8;
  const int chunks = cp->ag_rows_per_rank / cp->ag_chunk_rows;
  const long long vec_per_chunk = static_cast<long long>(cp->ag_chunk_rows) * vec_per_row;
  for (int step = 1; step < cp->ag_world; ++step) {
    int r = cp->ag_rank + step;
    if (r >= cp->ag_world) r -= cp->ag_world;
    // peer r has published its shard for this epoch?
    const uint32_t* flag = cp->ag_flags[cp->ag_rank] + r;
    spin_until_ge_sys(flag, cp->epoch, cp->guard, 11u);   // (on expiry: copy what is there, host raises)
    for (int c = 0; c < chunks; ++c) {
      const long long row_base = static_cast<long long>(r) * cp->ag_rows_per_rank +
                                 static_cast<long long>(c) * cp->ag_chunk_rows;
      const __nv_bfloat16* src = cp->ag_src[r] + row_base * cp->ag_ld;
      __nv_bfloat16* dst = cp->ag_dst + row_base * cp->ag_ld;
      long long i = gtid;
      // 4 independent 16-byte loads in flight per thread
      for (; i + 3 * gthreads < vec_per_chunk; i += 4 * gthreads) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long j = i + u * gthreads;
          const long long row = j / vec_per_row, col = j - row * vec_per_row;
          v[u] = ld_peer_v4(reinterpret_cast<const uint4*>(src + row * cp->ag_ld) + col);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long j = i + u * gthreads;
          const long long row = j / vec_per_row, col = j - row * vec_per_row;
          reinterpret_cast<uint4*>(dst + row * cp->ag_ld)[col] = v[u];
        }
      }
      for (; i < vec_per_chunk; i += gthreads) {
        const long long row = i / vec_per_row, col = i - row * vec_per_row;
        reinterpret_cast<uint4*>(dst + row * cp->ag_ld)[col] =
            ld_peer_v4(reinterpret_cast<const uint4*>(src + row * cp->ag_ld) + col);
      }
      __threadfence();
      __syncwarp();
      if ((t & 31) == 0) atomicAdd(cp->ag_count + r * cp->ag_max_chunks + c, 1u);
    }
  }
}

// TMA producer: block until the rows [row0, row0 + rows) of the operand buffer are local.
__device__ __forceinline__ void comm_wait_rows(const CommParams* cp, int row0, int rows) {
  if (cp->ag_world <= 1) return;
  const int r = row0 / cp->ag_rows_per_rank;
  if (r == cp->ag_rank) return;
  const int in_shard = row0 - r * cp->ag_rows_per_rank;
  const int c0 = in_shard / cp->ag_chunk_rows;
  const int c1 = (in_shard + rows - 1) / cp->ag_chunk_rows;
  const uint32_t want = cp->epoch * (gridDim.x * 2u);  // 2 gather warps per CTA, counters never reset
  for (int c = c0; c <= c1; ++c) {
    const uint32_t* cnt = cp->ag_count + r * cp->ag_max_chunks + c;
    uint32_t v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
    } while (static_cast<int32_t>(v - want) < 0);
  }
  // generic-proxy writes of the gather warps -> visible to the async proxy (TMA loads)
  asm volatile("fence.proxy.async.global;" ::: "memory");
}

template <int BN, int STAGES, bool A_MN, bool B_MN, bool COMM>
__device__ __forceinline__ void gemm_2cta_body(const CUtensorMap* tmA, const CUtensorMap* tmB,
                                               const CUtensorMap* tmC, const GemmParams& p,
                                               const CommParams* cp, const CUtensorMap* tmRS) {
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
    tma_prefetch_desc(tmA);
    tma_prefetch_desc(tmB);
    tma_prefetch_desc(tmC);
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
  if constexpr (COMM) comm_announce(cp);

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int m_blk, n_blk;
        tile_coords<8>(tile, num_m, num_n, m_blk, n_blk);
        if constexpr (COMM) m_blk = comm_rotate_m(cp, m_blk, num_m);
        const int m0 = m_blk * BM2 + static_cast<int>(cta_rank) * kBM;
        const int n0 = n_blk * BN + static_cast<int>(cta_rank) * BNH;
        if constexpr (COMM) comm_wait_rows(cp, m_blk * BM2, BM2);  // gathered rows have landed
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          // transaction bytes of both CTAs are counted on the LEADER's full barrier
          const uint32_t fb = mapa_shared(full_bar(stage), 0);
          const uint32_t a_dst = smem_base + stage * STAGE_BYTES;
          const uint32_t b_dst = a_dst + A_BYTES;
          if (elect_one_sync()) {
            if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * STAGE_BYTES);
            if constexpr (!A_MN) {
              tma_load_2d_2sm(a_dst, tmA, fb, kb * kBK, m0);
            } else {
#pragma unroll
              for (int j = 0; j < kBM / 64; ++j)
                tma_load_2d_2sm(a_dst + j * (kBK * 128), tmA, fb, m0 + j * 64, kb * kBK);
            }
            if constexpr (!B_MN) {
              tma_load_2d_2sm(b_dst, tmB, fb, kb * kBK, n0);
            } else {
#pragma unroll
              for (int j = 0; j < BNH / 64; ++j)
                tma_load_2d_2sm(b_dst + j * (kBK * 128), tmB, fb, n0 + j * 64, kb * kBK);
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
  } else if (COMM && warp >= 6) {
    // ===================== gather warps (6..7): all-gather of A rows over NVLink ==========
    comm_gather_rows(cp, (warp - 6) * 32 + lane, 64);
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
      if constexpr (COMM) m_blk = comm_rotate_m(cp, m_blk, num_m);
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
            if (p.accumulate) tma_reduce_add_2d(tmC, buf, col0, row0);
            else tma_store_2d(tmC, buf, col0, row0);
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
            if (COMM && cp->rs_world > 1) {
              // reduce-scatter: this 32-row box of the partial result goes to the landing slot
              // reserved for this rank in the memory of the GPU that owns the rows (a plain
              // TMA store across NVSwitch for remote owners); the owner sums the slots.
              const int owner = row0 / cp->rs_rows_per_rank;
              tma_store_2d(tmRS + owner, buf, col0, row0 - owner * cp->rs_rows_per_rank);
            } else {
              tma_store_2d(tmC, buf, col0, row0);
            }
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
    if (lane == 0) {
      tma_store_wait<0>();  // smem must stay valid until the last store has read it
      if (COMM && cp->rs_world > 1) {
        // all reduce-adds issued by this warp have been performed: tell every row owner
        __threadfence_system();
        for (int r = 0; r < cp->rs_world; ++r) red_add_release_sys(cp->rs_done[r], 1u);
      }
    }
  }

  tcgen05_fence_before();
  cluster_sync_all();  // the peer's smem / barriers stay valid until every MMA and commit retired
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<2>(tmem_base, TMEM_COLS);
  }
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tmA,
                      const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  gemm_2cta_body<BN, STAGES, A_MN, B_MN, false>(&tmA, &tmB, &tmC, p, nullptr, nullptr);
}

constexpr int kGemmCommThreads = 256;  // + 2 gather warps

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmCommThreads, 1)
gemm_bf16_2cta_comm_kernel(const __grid_constant__ CUtensorMap tmA,
                           const __grid_constant__ CUtensorMap tmB,
                           const __grid_constant__ CUtensorMap tmC, const GemmParams p,
                           const __grid_constant__ CommParams cp,
                           const __grid_constant__ TmapArray tmRS) {
  gemm_2cta_body<BN, STAGES, A_MN, B_MN, true>(&tmA, &tmB, &tmC, p, &cp, tmRS.m);
}

// ----------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------
template <int BN, int STAGES, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                       cudaStream_t stream) {
  constexpr size_t smem = size_t(STAGES) * (kBM * kBK * 2 + BN * kBK * 2) + 1024 + 256;
  auto kern = gemm_bf16_kernel<BN, STAGES, A_MN, B_MN>;
  static std::atomic<bool> configured[64];
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
  static std::atomic<bool> configured[64];
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

// ----------------------------------------------------------------------------
// GEMM fused with tensor-parallel collectives (CTA-pair kernel + communication hooks)
// ----------------------------------------------------------------------------
namespace sdml {

template <int BN, int STAGES, bool A_MN, bool B_MN>
static int launch_gemm_comm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                            const GemmParams& p, const CommParams& cp, const TmapArray& tmRS,
                            cudaStream_t stream) {
  constexpr size_t smem =
      size_t(STAGES) * (kBM * kBK * 2 + (BN / 2) * kBK * 2) + 4 * 2 * 32 * 128 + 1024 + 256;
  auto kern = gemm_bf16_2cta_comm_kernel<BN, STAGES, A_MN, B_MN>;
  static std::atomic<bool> configured[64];
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    SDML_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev & 63] = true;
  }
  // the gather warps of EVERY CTA take part in the chunk counters: always launch the full grid
  const int clusters = num_sms(dev) / 2;
  kern<<<2 * clusters, kGemmCommThreads, smem, stream>>>(tmA, tmB, tmC, p, cp, tmRS);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

__global__ void __launch_bounds__(256)
rs_finalize_kernel(const __nv_bfloat16* __restrict__ landing, int slots, long long slot_stride,
                   long long ld_land, int rows, int cols, const float* __restrict__ bias,
                   const __nv_bfloat16* __restrict__ aux, long long ld_aux, int relu,
                   __nv_bfloat16* __restrict__ out, long long ld_out, const uint32_t* done,
                   uint32_t expected, WaitGuard guard) {
  if (threadIdx.x == 0) spin_until_ge_sys(done, expected, guard, 12u);
  __syncthreads();
  const int c8 = cols / 8;
  const long long n8 = static_cast<long long>(rows) * c8;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / c8;
    const int c = static_cast<int>(i - r * c8) * 8;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < slots; ++s) {
      // written by remote TMA stores during the GEMM kernels: bypass L1
      const uint4 a = ld_peer_v4(reinterpret_cast<const uint4*>(landing + s * slot_stride + r * ld_land + c));
      const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[2 * j] += bf16_lo(w[j]);
        v[2 * j + 1] += bf16_hi(w[j]);
      }
    }
    if (bias) {
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + c));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + c + 4));
      v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
      v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    if (aux) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(aux + r * ld_aux + c));
      const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (!(bf16_lo(w[j]) > 0.f)) v[2 * j] = 0.f;
        if (!(bf16_hi(w[j]) > 0.f)) v[2 * j + 1] = 0.f;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(out + r * ld_out + c) = o;
  }
}

}  // namespace sdml

// C (or the reduce-scatter destinations) = epilogue(alpha * A_full . op(B)^T) with
//   * ag_world > 1: A_full [M, K] is the local row-stacked operand buffer in symmetric memory;
//     rows of rank r are pulled from ag_src[r] (HOST array of peer-mapped pointers to the peers'
//     operand buffers, same layout) while the GEMM runs.  ag_flags: HOST array, entry r = rank
//     r's flag array (>= ag_world uint32, zero-initialised); ag_count: local counters
//     [ag_world * ag_max_chunks] (zero-initialised, never reset); epoch = 1, 2, 3, ... per
//     (buffer, flags, counters) set, identical on all ranks.
//   * rs_world > 1: the bf16 partial result rows owned by rank r are stored into rs_dst[r]
//     (HOST array: for every owner r the peer-mapped address of the landing slot reserved for
//     THIS rank in r's memory, bf16 [M / rs_world, rs_ld]); rs_done: HOST array of the owners'
//     completion counters.  bias / relu / aux are applied by sdml_rs_finalize on the sum.
// A is K-major; B may be K-major (b_mn = 0) or MN-major (b_mn = 1).  M % (256 * world) == 0.
extern "C" int sdml_gemm_bf16_comm(const void* A, const void* B, void* C, int M, int N, int K,
                                   long long lda, long long ldb, long long ldc, int b_mn,
                                   const float* bias, const void* aux, long long ld_aux, int relu,
                                   int out_fp32, float alpha,
                                   int ag_world, int ag_rank, void* const* ag_src,
                                   void* const* ag_flags, void* ag_count, int ag_max_chunks,
                                   int ag_chunk_rows, unsigned int epoch,
                                   int rs_world, int rs_rank, void* const* rs_dst, long long rs_ld,
                                   void* const* rs_done, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return set_error("sdml_gemm_bf16_comm: empty problem");
  if ((lda & 7) || (ldb & 7) || (K & 7)) return set_error("sdml_gemm_bf16_comm: lda/ldb/K must be multiples of 8");
  if (ag_world < 1) ag_world = 1;
  if (rs_world < 1) rs_world = 1;
  if (ag_world > kCommMaxRanks || rs_world > kCommMaxRanks)
    return set_error("sdml_gemm_bf16_comm: at most 8 ranks");
  if (ag_world > 1 && (M % (256 * ag_world) || ag_chunk_rows % 256 || (M / ag_world) % ag_chunk_rows))
    return set_error("sdml_gemm_bf16_comm: all-gather needs M % (256*world) == 0 and chunk rows % 256 == 0");
  if (rs_world > 1 && (M % (32 * rs_world) || (rs_ld & 7)))
    return set_error("sdml_gemm_bf16_comm: reduce-scatter needs M % (32*world) == 0 and rs_ld % 8 == 0");
  const bool rs = rs_world > 1;
  if (rs) out_fp32 = 0;

  CUtensorMap tmA, tmB, tmC;
  int rc = make_tmap_2d(&tmA, A, TMAP_BF16, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kBK, kBM, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d(&tmB, B, TMAP_BF16, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, kBK, 128, TMAP_SWIZZLE_128B);
  else       rc = make_tmap_2d(&tmB, B, TMAP_BF16, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, kBK, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  TmapArray tmRS;
  memset(&tmRS, 0, sizeof(tmRS));
  if (rs) {
    for (int r = 0; r < rs_world; ++r) {
      rc = make_tmap_2d(&tmRS.m[r], rs_dst[r], TMAP_BF16, (uint64_t)N, (uint64_t)(M / rs_world),
                        (uint64_t)rs_ld, 64, 32, TMAP_SWIZZLE_128B);
      if (rc) return rc;
    }
    tmC = tmRS.m[rs_rank];
  } else {
    rc = out_fp32 ? make_tmap_2d(&tmC, C, TMAP_F32, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 32, 32, TMAP_SWIZZLE_128B)
                  : make_tmap_2d(&tmC, C, TMAP_BF16, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 64, 32, TMAP_SWIZZLE_128B);
    if (rc) return rc;
  }

  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.C = C; p.ldc = ldc;
  p.bias = rs ? nullptr : bias;
  p.aux = rs ? nullptr : reinterpret_cast<const __nv_bfloat16*>(aux);
  p.ld_aux = ld_aux;
  p.alpha = alpha;
  p.relu = rs ? 0 : relu;
  p.accumulate = 0;
  p.out_fp32 = out_fp32;

  CommParams cp;
  memset(&cp, 0, sizeof(cp));
  cp.guard = WaitGuard{wait_guard().timeout_ns, wait_guard().err};
  cp.ag_world = ag_world; cp.ag_rank = ag_rank;
  cp.ag_rows_per_rank = M / ag_world;
  cp.ag_chunk_rows = ag_chunk_rows > 0 ? ag_chunk_rows : 256;
  cp.ag_ld = lda; cp.ag_k = K;
  cp.ag_dst = reinterpret_cast<__nv_bfloat16*>(const_cast<void*>(A));
  cp.ag_count = static_cast<uint32_t*>(ag_count);
  cp.ag_max_chunks = ag_max_chunks;
  cp.epoch = epoch;
  for (int r = 0; r < ag_world && ag_world > 1; ++r) {
    cp.ag_src[r] = static_cast<const __nv_bfloat16*>(ag_src[r]);
    cp.ag_flags[r] = static_cast<uint32_t*>(ag_flags[r]);
  }
  if (ag_world > 1 && (M / ag_world) / cp.ag_chunk_rows > ag_max_chunks)
    return set_error("sdml_gemm_bf16_comm: ag_max_chunks too small");
  cp.rs_world = rs_world; cp.rs_rank = rs_rank;
  cp.rs_rows_per_rank = M / rs_world;
  for (int r = 0; r < rs_world && rs; ++r) cp.rs_done[r] = static_cast<uint32_t*>(rs_done[r]);

  return b_mn ? launch_gemm_comm<256, 6, false, true>(tmA, tmB, tmC, p, cp, tmRS, stream)
              : launch_gemm_comm<256, 6, false, false>(tmA, tmB, tmC, p, cp, tmRS, stream);
}

// Owner side of GEMM -> reduce-scatter: wait until `done` reaches `expected` (every epilogue
// warp of every rank's kernel adds 1 per call: expected = calls * world * 4 * grid), then
// out = mask(relu(sum over `slots` landing slots + bias)) as bf16.  landing: local bf16 buffer,
// slot s at landing + s * slot_stride elements, rows x cols with leading dimension ld_land.
extern "C" int sdml_rs_finalize(const void* landing, int slots, long long slot_stride,
                                long long ld_land, int rows, int cols, const float* bias,
                                const void* aux, long long ld_aux, int relu, void* out,
                                long long ld_out, const void* done, unsigned int expected,
                                cudaStream_t stream) {
  if (rows <= 0 || cols <= 0 || (cols & 7)) return set_error("sdml_rs_finalize: cols must be a multiple of 8");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  long long blocks = (static_cast<long long>(rows) * (cols / 8) + 256 * 2 - 1) / (256 * 2);
  if (blocks > 2LL * num_sms(dev)) blocks = 2LL * num_sms(dev);   // all CTAs co-resident (they spin)
  rs_finalize_kernel<<<(int)blocks, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(landing), slots, slot_stride, ld_land, rows, cols, bias,
      static_cast<const __nv_bfloat16*>(aux), ld_aux, relu, static_cast<__nv_bfloat16*>(out), ld_out,
      static_cast<const uint32_t*>(done), expected, WaitGuard{wait_guard().timeout_ns, wait_guard().err});
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// ----------------------------------------------------------------------------
// Weight-gradient GEMM fused with the data-parallel reduction and the optimizer
// ----------------------------------------------------------------------------
// dW = dz^T . x of one replica never touches local HBM as a gradient: the epilogue of
// sdml_wgrad_rs scales it, casts it to bf16 and TMA-stores every 32-row box into the landing
// slot this replica owns on the GPU that owns those rows of the weight (reduce-scatter over
// NVSwitch, overlapped tile by tile with the MMA main loop).  sdml_rs_finalize_sgd, on the
// owner, sums the replicas' slots in fp32, applies SGD-with-momentum to ITS rows of the fp32
// master weight (optimizer state is sharded over the replicas) and writes the refreshed bf16
// weights straight into every replica's compute copy (the all-gather half, as P2P stores),
// then publishes a counter the replicas wait on before the next forward.
namespace sdml {

__global__ void __launch_bounds__(256)
rs_finalize_sgd_kernel(const __nv_bfloat16* __restrict__ landing, int slots, long long slot_stride,
                       long long ld_land, int rows, int cols, float* __restrict__ master,
                       float* __restrict__ mom, long long ld_w, float lr, float momentum,
                       float weight_decay, const __grid_constant__ PtrArray wout, long long ld_wout,
                       int world, const uint32_t* done, uint32_t expected,
                       const __grid_constant__ PtrArray ready, uint32_t* exit_counter,
                       WaitGuard guard) {
  if (threadIdx.x == 0) spin_until_ge_sys(done, expected, guard, 13u);
  __syncthreads();
  const int c8 = cols / 8;
  const long long n8 = static_cast<long long>(rows) * c8;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / c8;
    const int c = static_cast<int>(i - r * c8) * 8;
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < slots; ++s) {
      const uint4 a = ld_peer_v4(reinterpret_cast<const uint4*>(landing + s * slot_stride + r * ld_land + c));
      const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[2 * j] += bf16_lo(w[j]);
        g[2 * j + 1] += bf16_hi(w[j]);
      }
    }
    float* pw = master + r * ld_w + c;
    float* pm = mom + r * ld_w + c;
    float4 w0 = *reinterpret_cast<float4*>(pw), w1 = *reinterpret_cast<float4*>(pw + 4);
    float4 m0 = *reinterpret_cast<float4*>(pm), m1 = *reinterpret_cast<float4*>(pm + 4);
    float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
    float mv[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float d = g[j] + weight_decay * wv[j];
      if (momentum != 0.f) {
        mv[j] = momentum * mv[j] + d;
        d = mv[j];
      }
      wv[j] -= lr * d;
    }
    *reinterpret_cast<float4*>(pw) = make_float4(wv[0], wv[1], wv[2], wv[3]);
    *reinterpret_cast<float4*>(pw + 4) = make_float4(wv[4], wv[5], wv[6], wv[7]);
    if (momentum != 0.f) {
      *reinterpret_cast<float4*>(pm) = make_float4(mv[0], mv[1], mv[2], mv[3]);
      *reinterpret_cast<float4*>(pm + 4) = make_float4(mv[4], mv[5], mv[6], mv[7]);
    }
    uint4 o;
    o.x = pack_bf16x2(wv[0], wv[1]);
    o.y = pack_bf16x2(wv[2], wv[3]);
    o.z = pack_bf16x2(wv[4], wv[5]);
    o.w = pack_bf16x2(wv[6], wv[7]);
    for (int q = 0; q < world; ++q)   // refreshed compute copy for every replica (P2P stores)
      *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(wout.p[q]) + r * ld_wout + c) = o;
  }
  // last CTA out: all stores of this kernel are fenced -> tell every replica this shard is fresh
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(exit_counter, 1u);
    if (prev == gridDim.x - 1) {
      *exit_counter = 0;
      __threadfence_system();
      for (int q = 0; q < world; ++q) red_add_release_sys(static_cast<uint32_t*>(ready.p[q]), 1u);
    }
  }
}

__global__ void wait_counter_kernel(const uint32_t* counter, uint32_t expected, WaitGuard guard) {
  spin_until_ge_sys(counter, expected, guard, 14u);
}

}  // namespace sdml

// dW partial of this replica -> reduce-scatter into the row owners' landing slots.
// dz: bf16 [K rows, M = out features] (ld_dz), x: bf16 [K rows, N = in features] (ld_x);
// alpha scales the gradient (e.g. 1 / dp).  rs_dst / rs_done as for sdml_gemm_bf16_comm.
extern "C" int sdml_wgrad_rs(const void* dz, const void* x, int M, int N, int K, long long ld_dz,
                             long long ld_x, float alpha, int rs_world, int rs_rank,
                             void* const* rs_dst, long long rs_ld, void* const* rs_done,
                             cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return set_error("sdml_wgrad_rs: empty problem");
  if ((ld_dz & 7) || (ld_x & 7)) return set_error("sdml_wgrad_rs: leading dimensions must be multiples of 8");
  if (rs_world < 2 || rs_world > kCommMaxRanks) return set_error("sdml_wgrad_rs: 2..8 replicas");
  if (M % (32 * rs_world) || (rs_ld & 7)) return set_error("sdml_wgrad_rs: M % (32*world) == 0 and rs_ld % 8 == 0 required");
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, dz, TMAP_BF16, (uint64_t)M, (uint64_t)K, (uint64_t)ld_dz, 64, kBK, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_2d(&tmB, x, TMAP_BF16, (uint64_t)N, (uint64_t)K, (uint64_t)ld_x, 64, kBK, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  TmapArray tmRS;
  memset(&tmRS, 0, sizeof(tmRS));
  for (int r = 0; r < rs_world; ++r) {
    rc = make_tmap_2d(&tmRS.m[r], rs_dst[r], TMAP_BF16, (uint64_t)N, (uint64_t)(M / rs_world),
                      (uint64_t)rs_ld, 64, 32, TMAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K;
  p.alpha = alpha;
  CommParams cp;
  memset(&cp, 0, sizeof(cp));
  cp.guard = WaitGuard{wait_guard().timeout_ns, wait_guard().err};
  cp.ag_world = 1;
  cp.ag_rows_per_rank = M; cp.ag_chunk_rows = 256;
  cp.rs_world = rs_world; cp.rs_rank = rs_rank;
  cp.rs_rows_per_rank = M / rs_world;
  for (int r = 0; r < rs_world; ++r) cp.rs_done[r] = static_cast<uint32_t*>(rs_done[r]);
  return launch_gemm_comm<256, 6, true, true>(tmA, tmB, tmRS.m[rs_rank], p, cp, tmRS, stream);
}

// Owner side: wait for `expected` on `done`, then for the owned rows: g = sum of landing slots,
// SGD-with-momentum on master/mom (fp32, row-major, ld_w), refreshed bf16 rows stored to
// wout[q] (HOST array of `world` peer-mapped pointers to the same rows of every replica's bf16
// weight, ld_wout), and +1 (release, system scope) on ready[q] for every replica once done.
// exit_counter: one zeroed uint32 in local memory.
extern "C" int sdml_rs_finalize_sgd(const void* landing, int slots, long long slot_stride,
                                    long long ld_land, int rows, int cols, float* master,
                                    float* mom, long long ld_w, float lr, float momentum,
                                    float weight_decay, void* const* wout, long long ld_wout,
                                    int world, const void* done, unsigned int expected,
                                    void* const* ready, void* exit_counter, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0 || (cols & 7) || (ld_w & 3) || (ld_wout & 7))
    return set_error("sdml_rs_finalize_sgd: cols % 8, ld_w % 4, ld_wout % 8 must be 0");
  if (world < 1 || world > kCommMaxRanks) return set_error("sdml_rs_finalize_sgd: 1..8 replicas");
  PtrArray w, rdy;
  memset(&w, 0, sizeof(w));
  memset(&rdy, 0, sizeof(rdy));
  for (int q = 0; q < world; ++q) {
    w.p[q] = wout[q];
    rdy.p[q] = ready[q];
  }
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  long long blocks = (static_cast<long long>(rows) * (cols / 8) + 256 * 2 - 1) / (256 * 2);
  if (blocks > 2LL * num_sms(dev)) blocks = 2LL * num_sms(dev);
  rs_finalize_sgd_kernel<<<(int)blocks, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(landing), slots, slot_stride, ld_land, rows, cols, master,
      mom, ld_w, lr, momentum, weight_decay, w, ld_wout, world, static_cast<const uint32_t*>(done),
      expected, rdy, static_cast<uint32_t*>(exit_counter),
      WaitGuard{wait_guard().timeout_ns, wait_guard().err});
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// Block `stream` until *counter >= expected (one spinning thread; pairs with the `ready`
// counters of sdml_rs_finalize_sgd).
extern "C" int sdml_wait_counter(const void* counter, unsigned int expected, cudaStream_t stream) {
  const HostWaitGuard hg = wait_guard();
  wait_counter_kernel<<<1, 1, 0, stream>>>(static_cast<const uint32_t*>(counter), expected,
                                           WaitGuard{hg.timeout_ns, hg.err});
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

extern "C" int sdml_gemm_comm_grid(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  return 2 * (sdml::num_sms(dev) / 2);
}
