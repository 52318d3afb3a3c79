// MX-fp8 (E4M3 values, one UE8M0 scale per 32 K-elements) quantisation and block-scaled
// GEMM on tcgen05: `tcgen05.mma.kind::mxf8f6f4.block_scale`, scale factors staged
// global -> smem (bulk copy) -> TMEM (`tcgen05.cp`), fp32 accumulation in TMEM.
//
//   C[M,N] = epilogue( sum_k (A[m,k] * sfa[m, k/32]) * (B[n,k] * sfb[n, k/32]) )
//
// Both operands are K-major fp8 with 128-byte (= 128 element) swizzled rows, so one smem
// stage holds BLOCK_K = 128 elements = 4 MMA k-steps of 32; the 4 scale bytes of a row for
// those 4 k-steps form one 32-bit TMEM column entry and the instruction descriptor's
// a_sf_id / b_sf_id select the byte.
//
// Scale-factor memory layout (written by sdml_quant_mxfp8, read verbatim by the GEMM): for
// every (128-row group g, 128-column block kb) one 512-byte atom
//     atom[(r % 32) * 16 + ((r % 128) / 32) * 4 + j] = UE8M0 scale of row r, columns kb*128 + 32*j ..
// i.e. exactly the 32 x 128-bit shape `tcgen05.cp.32x128b.warpx4` copies into 4 TMEM columns
// (lane = r % 32 replicated to the four lane quadrants, column = (r % 128) / 32).  Atoms are
// ordered kb-major: offset = (kb * groups + g) * 512.
//
// This is the fp8 form of the reference's fc1 Linear+bias+ReLU
// (/root/reference/simple_distributed.py:63, :75) for BASELINE.json config 5.
#include "sdml_host.h"
#include <atomic>
#include "sdml_ptx.cuh"

#include <cuda_fp8.h>
#include <stdlib.h>
#include <string.h>

namespace sdml {

// ----------------------------------------------------------------------------
// quantisation: bf16 [rows, cols] -> fp8 e4m3 [rows, cols] + UE8M0 scales (1 x 32 blocks)
// ----------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
quant_mxfp8_kernel(const __nv_bfloat16* __restrict__ src, long long ld_src, int rows, int cols,
                   uint8_t* __restrict__ dst, long long ld_dst, uint8_t* __restrict__ sf,
                   int row_groups) {
  const int blocks_per_row = cols / 32;
  const long long total = static_cast<long long>(rows) * blocks_per_row;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / blocks_per_row);
    const int b = static_cast<int>(i - static_cast<long long>(r) * blocks_per_row);
    const uint4* sp = reinterpret_cast<const uint4*>(src + static_cast<long long>(r) * ld_src + b * 32);
    uint4 in[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) in[j] = __ldg(sp + j);
    float v[32];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t w[4] = {in[j].x, in[j].y, in[j].z, in[j].w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        v[8 * j + 2 * q] = bf16_lo(w[q]);
        v[8 * j + 2 * q + 1] = bf16_hi(w[q]);
      }
    }
    float amax = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) amax = fmaxf(amax, fabsf(v[j]));
    // smallest power of two s with amax / s <= 448 (E4M3 max): no saturation, <= 1 bit lost
    int e = 0;
    if (amax > 0.f) {
      e = static_cast<int>(ceilf(log2f(amax * (1.0f / 448.0f))));
      e = max(-127, min(127, e));
    } else {
      e = -127;
    }
    const float inv = exp2f(static_cast<float>(-e));
    uint32_t packed[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const __nv_fp8x2_storage_t lo = __nv_cvt_float2_to_fp8x2(
          make_float2(v[4 * j] * inv, v[4 * j + 1] * inv), __NV_SATFINITE, __NV_E4M3);
      const __nv_fp8x2_storage_t hi = __nv_cvt_float2_to_fp8x2(
          make_float2(v[4 * j + 2] * inv, v[4 * j + 3] * inv), __NV_SATFINITE, __NV_E4M3);
      packed[j] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
    }
    uint4* dp = reinterpret_cast<uint4*>(dst + static_cast<long long>(r) * ld_dst + b * 32);
    dp[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    dp[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    const int kb = b >> 2, j4 = b & 3, g = r >> 7;
    sf[(static_cast<long long>(kb) * row_groups + g) * 512 + (r & 31) * 16 + ((r & 127) >> 5) * 4 + j4] =
        static_cast<uint8_t>(e + 127);
  }
}

// Power-of-two block scale: smallest e with amax / 2^e <= 448 (E4M3 max); -127 for an all-zero block.
__device__ __forceinline__ int mx_exponent(float amax) {
  int e = -127;
  if (amax > 0.f) {
    e = static_cast<int>(ceilf(log2f(amax * (1.0f / 448.0f))));
    e = max(-127, min(127, e));
  }
  return e;
}

// One MX block: 32 fp32 values -> 32 E4M3 bytes (8 packed words) + the UE8M0 exponent byte.
__device__ __forceinline__ uint8_t quant_block32(const float (&v)[32], uint32_t (&packed)[8]) {
  float amax = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) amax = fmaxf(amax, fabsf(v[j]));
  const int e = mx_exponent(amax);
  const float inv = exp2f(static_cast<float>(-e));
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const __nv_fp8x2_storage_t lo = __nv_cvt_float2_to_fp8x2(
        make_float2(v[4 * j] * inv, v[4 * j + 1] * inv), __NV_SATFINITE, __NV_E4M3);
    const __nv_fp8x2_storage_t hi = __nv_cvt_float2_to_fp8x2(
        make_float2(v[4 * j + 2] * inv, v[4 * j + 3] * inv), __NV_SATFINITE, __NV_E4M3);
    packed[j] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
  }
  return static_cast<uint8_t>(e + 127);
}

// ----------------------------------------------------------------------------
// dual quantisation: ONE read of a bf16 [rows, cols] tensor produces
//   * dst_row [rows, cols]  fp8, scaled along the columns (1 x 32 blocks)  -> the K-major operand
//     of a GEMM that contracts over `cols` (forward: x, W; dgrad: dz), and / or
//   * dst_col [cols, rows]  fp8, the TRANSPOSE, scaled along the original rows (32 x 1 blocks)
//     -> the K-major operand of a GEMM that contracts over `rows` (wgrad: x^T, dz^T; dgrad: W^T).
// A training step needs both copies of every activation, gradient and weight (MX scales must
// run along the contraction dimension), so reading the source once halves the traffic.
// Tiles of 128 rows x 64 columns staged in shared memory (rows padded to 72 elements so both
// the row-wise 16-byte reads and the column-wise 2-byte reads are bank-conflict free).
// ----------------------------------------------------------------------------
constexpr int kQR = 128, kQC = 64, kQLD = kQC + 8;

__global__ void __launch_bounds__(256)
quant_mxfp8_dual_kernel(const __nv_bfloat16* __restrict__ src, long long ld_src, int rows, int cols,
                        uint8_t* __restrict__ dst_row, long long ld_row, uint8_t* __restrict__ sf_row,
                        int groups_row, uint8_t* __restrict__ dst_col, long long ld_col,
                        uint8_t* __restrict__ sf_col, int groups_col) {
  __shared__ __align__(16) __nv_bfloat16 tile[kQR * kQLD];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int tiles_c = cols / kQC;
  const int n_tiles = (rows / kQR) * tiles_c;
  for (int tl = blockIdx.x; tl < n_tiles; tl += gridDim.x) {
    const int tr = tl / tiles_c, tc = tl - tr * tiles_c;
    const int row0 = tr * kQR, col0 = tc * kQC;
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {
      const int r = pass * 32 + (t >> 3), q = t & 7;
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + static_cast<long long>(row0 + r) * ld_src + col0) + q);
      *reinterpret_cast<uint4*>(tile + r * kQLD + q * 8) = v;
    }
    __syncthreads();
    if (dst_row != nullptr) {
      const int r = t >> 1, cb = t & 1;   // two lanes per row: 64 contiguous output bytes
      float v[32];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 a = *reinterpret_cast<const uint4*>(tile + r * kQLD + cb * 32 + j * 8);
        const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          v[8 * j + 2 * q] = bf16_lo(w[q]);
          v[8 * j + 2 * q + 1] = bf16_hi(w[q]);
        }
      }
      uint32_t pk[8];
      const uint8_t e = quant_block32(v, pk);
      const int gr = row0 + r;
      uint4* dp = reinterpret_cast<uint4*>(dst_row + static_cast<long long>(gr) * ld_row + col0 + cb * 32);
      dp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      dp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
      const int b = (col0 >> 5) + cb;
      sf_row[(static_cast<long long>(b >> 2) * groups_row + (gr >> 7)) * 512 + (gr & 31) * 16 +
             ((gr & 127) >> 5) * 4 + (b & 3)] = e;
    }
    if (dst_col != nullptr) {
      const int c = (warp >> 2) * 32 + lane, rb = warp & 3;   // a warp: 32 columns x one 32-row block
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __bfloat162float(tile[(rb * 32 + j) * kQLD + c]);
      uint32_t pk[8];
      const uint8_t e = quant_block32(v, pk);
      const int gc = col0 + c;            // row of the transposed operand
      uint4* dp = reinterpret_cast<uint4*>(dst_col + static_cast<long long>(gc) * ld_col + row0 + rb * 32);
      dp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      dp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
      // contraction index of the transposed operand = original row: k-block tr, sub-block rb
      sf_col[(static_cast<long long>(tr) * groups_col + (gc >> 7)) * 512 + (gc & 31) * 16 +
             ((gc & 127) >> 5) * 4 + rb] = e;
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------
// block-scaled GEMM
// ----------------------------------------------------------------------------
struct Fp8Params {
  int M, N, K;
  void* C;
  long long ldc;
  const float* bias;
  const __nv_bfloat16* aux;
  long long ld_aux;
  float alpha;
  int relu, out_fp32, accumulate;
  const uint8_t* sfa;  // atoms [K/128][ceil(M/128)][512]
  const uint8_t* sfb;  // atoms [K/128][ceil(N/128)][512]
  int groups_a, groups_b;
  // Optional fused quantisation of the OUTPUT (CTA-pair kernel, bf16 output, M % 128 == 0,
  // N % 32 == 0): the epilogue also emits the MX-fp8 copies the next GEMMs need, so the
  // stand-alone dual quantiser pass over C disappears.
  uint8_t* q_row;        // C as a row-wise operand [M, N] (scales along N), or nullptr
  long long ld_qrow;
  uint8_t* sf_qrow;      // atoms [N/128][M/128][512]
  uint8_t* q_col;        // C^T as an operand [N, M] scaled along M (already offset to this call's
  long long ld_qcol;     //   row range inside a step-wide buffer), or nullptr
  uint8_t* sf_qcol;      // atoms of that row range: [M/128][groups_qcol][512]
  int groups_qcol;       // ceil(rows of the transposed operand / 128) = ceil(N / 128)
  // CTA-pair kernel, bf16 output with 16-byte aligned rows: the epilogue stages 32 x 64 boxes in
  // shared memory and TMA-stores them (whole 128-byte lines; the form that also travels well over
  // NVLink when C is a pipeline neighbour's receive slot) instead of per-thread 16-byte stores
  int c_tma;
};

// GEMM -> reduce-scatter hook of the CTA-pair kernel (tensor-parallel row-parallel layer): the rows
// of the bf16 partial result owned by rank r leave as 32 x 64 TMA boxes into the landing slot this
// rank owns in r's memory (map[r]; remote slots over NVLink), and every epilogue warp adds 1 to every
// owner's completion counter when its stores are done -- the same contract as the bf16 kernel's
// (csrc/gemm_tcgen05.cu), so sdml_rs_finalize sums the slots of either.
constexpr int kFp8MaxRanks = 8;
struct Fp8Rs {
  CUtensorMap map[kFp8MaxRanks];
  uint32_t* done[kFp8MaxRanks];
  int world;            // <= 1: no reduce-scatter
  int rows_per_rank;    // M / world (multiple of 32)
};

constexpr int kFBM = 128, kFBN = 128, kFBK = 128, kFStages = 6;
constexpr int kFThreads = 192;

__device__ __forceinline__ void bulk_load_1d(uint32_t smem_dst, const void* gsrc, uint32_t bytes,
                                             uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_dst),
      "l"(gsrc), "r"(bytes), "r"(bar)
      : "memory");
}

// direct (per-thread) epilogue store of one 32-column chunk, bf16 or fp32 output
__device__ __forceinline__ void fp8_store_chunk(const Fp8Params& p, const float (&v)[32], int row,
                                                int col0) {
  const bool full = col0 + 32 <= p.N;
  if (p.out_fp32) {
    float* cp = reinterpret_cast<float*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
    if (full && (p.ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 o = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        if (p.accumulate) {   // weight gradients summed over micro-batch groups
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
    __nv_bfloat16* cp = reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
    if (full && (p.ldc & 7) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0) {
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

__global__ void __launch_bounds__(kFThreads, 1)
gemm_mxfp8_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const Fp8Params p) {
  constexpr uint32_t A_BYTES = kFBM * kFBK;  // 1 byte per element
  constexpr uint32_t B_BYTES = kFBN * kFBK;
  constexpr uint32_t SF_BYTES = 512;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES + 2 * SF_BYTES;  // 33 792, 1024-aligned
  constexpr uint32_t TX_BYTES = A_BYTES + B_BYTES + 2 * SF_BYTES;
  constexpr uint32_t TMEM_COLS = 512;        // 2 x 128 accumulator + 8 scale columns, power of two
  constexpr uint32_t SFA_COL = 256, SFB_COL = 260;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t smem_base = (raw_addr + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + kFStages * STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kFStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kFStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kFStages + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kFStages + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw_addr));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_m = (p.M + kFBM - 1) / kFBM;
  const int num_n = (p.N + kFBN - 1) / kFBN;
  const int num_tiles = num_m * num_n;
  const int num_k = p.K / kFBK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kFStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_slot, TMEM_COLS);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  auto tile_mn = [&](int tile, int& m_blk, int& n_blk) {
    const int per_group = 16 * num_n;
    const int group = tile / per_group;
    const int first_m = group * 16;
    const int gsize = min(num_m - first_m, 16);
    const int in_group = tile - group * per_group;
    m_blk = first_m + in_group % gsize;
    n_blk = in_group / gsize;
  };

  if (warp == 0) {
    // ===================== producer: operand tiles (TMA) + scale atoms (bulk copies) =========
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_mn(tile, m_blk, n_blk);
      for (int kb = 0; kb < num_k; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t fb = full_bar(stage);
        const uint32_t a_dst = smem_base + stage * STAGE_BYTES;
        const uint32_t b_dst = a_dst + A_BYTES;
        const uint32_t sfa_dst = b_dst + B_BYTES;
        const uint32_t sfb_dst = sfa_dst + SF_BYTES;
        if (elect_one_sync()) {
          mbar_arrive_expect_tx(fb, TX_BYTES);
          tma_load_2d(a_dst, &tmA, fb, kb * kFBK, m_blk * kFBM);
          tma_load_2d(b_dst, &tmB, fb, kb * kFBK, n_blk * kFBN);
          bulk_load_1d(sfa_dst, p.sfa + (static_cast<long long>(kb) * p.groups_a + m_blk) * 512, SF_BYTES, fb);
          bulk_load_1d(sfb_dst, p.sfb + (static_cast<long long>(kb) * p.groups_b + n_blk) * 512, SF_BYTES, fb);
        }
        __syncwarp();
        if (++stage == kFStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc0 = make_idesc_mxf8(0u, 0u, kFBM, kFBN, 0u, 0u);  // E4M3 x E4M3, UE8M0 scales
    int stage = 0;
    uint32_t phase = 0;
    int tile_iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tile_iter) {
      const int acc = tile_iter & 1;
      const uint32_t acc_phase = (tile_iter >> 1) & 1;
      mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + acc * kFBN;
      for (int kb = 0; kb < num_k; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tcgen05_fence_after();
        const uint32_t a_base = smem_base + stage * STAGE_BYTES;
        const uint32_t b_base = a_base + A_BYTES;
        const uint32_t sfa_s = b_base + B_BYTES;
        const uint32_t sfb_s = sfa_s + SF_BYTES;
        if (elect_one_sync()) {
          // scale atoms smem -> TMEM (32 lanes x 128 bit, replicated to the 4 lane quadrants)
          tcgen05_cp_32x128b_warpx4(tmem_base + SFA_COL, make_smem_desc(sfa_s, 0, 128, kSwizzleNone));
          tcgen05_cp_32x128b_warpx4(tmem_base + SFB_COL, make_smem_desc(sfb_s, 0, 128, kSwizzleNone));
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t da = make_smem_desc(a_base + k * 32, 16, 1024, kSwizzle128B);
            const uint64_t db = make_smem_desc(b_base + k * 32, 16, 1024, kSwizzle128B);
            const uint32_t idesc = idesc0 | (static_cast<uint32_t>(k) << 29) | (static_cast<uint32_t>(k) << 4);
            umma_mxf8(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u, tmem_base + SFA_COL,
                      tmem_base + SFB_COL);
          }
          umma_commit(empty_bar(stage));
          if (kb == num_k - 1) umma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == kFStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else {
    // ===================== epilogue =====================
    const int quad = warp & 3;
    int tile_iter = 0;
    const bool aux_vec_ok =
        p.aux && ((p.ld_aux & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.aux) & 15) == 0);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tile_iter) {
      int m_blk, n_blk;
      tile_mn(tile, m_blk, n_blk);
      const int acc = tile_iter & 1;
      const uint32_t acc_phase = (tile_iter >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tcgen05_fence_after();
      const int row = m_blk * kFBM + quad * 32 + lane;
      const bool row_ok = row < p.M;
#pragma unroll 1
      for (int c = 0; c < kFBN / 32; ++c) {
        const int col0 = n_blk * kFBN + c * 32;
        if (col0 >= p.N) break;
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * kFBN + c * 32, r);
        tmem_ld_wait();
        if (!row_ok) continue;
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
        const bool full = col0 + 32 <= p.N;
        if (p.bias) {
          for (int i = 0; i < 32; ++i)
            if (full || col0 + i < p.N) v[i] += __ldg(p.bias + col0 + i);
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
        fp8_store_chunk(p, v, row, col0);
      }
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
// CTA-pair (cta_group::2) block-scaled GEMM: 256 x BN output tile per cluster of two CTAs.
// The 128 x 128 single-CTA kernel above moves 32 KB of operands through L2 -> smem per
// 4.2 MFLOP, which at fp8 tensor rates is the L2 feed limit, not the tensor pipe.  Pairing two
// SMs on one 256 x 192 tile (each CTA stages its 128 A rows and HALF of the B rows; the MMA
// reads both halves) brings that to 28 KB per 6.3 MFLOP.  BN = 192 rather than 256 because
// TMEM has 512 columns: two 192-column accumulators (epilogue of tile i overlaps the MMAs of
// tile i+1) + 4 columns of A scales + 8 columns of B scales = 396.
//
// Scale factors: every CTA needs the scales of its own 128 A rows (one 512-byte atom) and of
// ALL BN columns of B (each CTA's tensor core produces all BN columns of its 128 rows): the
// 1 or 2 consecutive atoms that cover [n0, n0 + BN).  A tile that starts in the middle of an
// atom (odd tiles at BN = 192 start at row 64) addresses the B scales 2 TMEM columns further
// in.  Scale atoms travel by TMA tensor loads (the scale array viewed as [atoms, 128] fp32)
// because, unlike the 1-D bulk copy, those can signal the LEADER CTA's mbarrier.
// ----------------------------------------------------------------------------
constexpr int kF2Threads = 192;

template <int BN>
struct F2Cfg {
  static constexpr int BNH = BN / 2;
  static constexpr int SFB_ATOMS = BN > 128 ? 2 : 1;
  static constexpr uint32_t A_BYTES = 128 * 128;
  static constexpr uint32_t B_BYTES = BNH * 128;
  static constexpr uint32_t SFA_BYTES = 512;
  static constexpr uint32_t SFB_BYTES = 512 * SFB_ATOMS;
  static constexpr uint32_t TX_BYTES = A_BYTES + B_BYTES + SFA_BYTES + SFB_BYTES;
  static constexpr uint32_t STAGE_BYTES = (TX_BYTES + 1023u) & ~1023u;
  // epilogue staging for the TMA-store path: 4 warps x 2 buffers x (32 rows x 128 B, 128B-swizzled)
  static constexpr uint32_t EPI_BUF_BYTES = 32 * 128;
  static constexpr uint32_t EPI_BYTES = 4 * 2 * EPI_BUF_BYTES;
  static constexpr uint32_t FIXED_BYTES = 1024 + 256 + 4 * 32 * 36 + EPI_BYTES;   // align slack, barriers, QOUT tiles, staging
  static constexpr int STAGES = (227 * 1024 - FIXED_BYTES) / STAGE_BYTES;         // 6 at BN = 192, 7 at BN = 128
  static constexpr uint32_t SFA_COL = 2 * BN;
  static constexpr uint32_t SFB_COL = 2 * BN + 4;
  static constexpr size_t SMEM = size_t(STAGES) * STAGE_BYTES + FIXED_BYTES;
  static_assert(2 * BN + 4 + 4 * SFB_ATOMS <= 512, "TMEM columns");
  static_assert(BNH % 8 == 0 && BN % 32 == 0, "tile shape");
};

// alpha, bias, ReLU and the ReLU-backward mask on one 32-column chunk of row `row` (r = raw TMEM words)
__device__ __forceinline__ void fp8_epilogue_math(const Fp8Params& p, const uint32_t (&r)[32], int row,
                                                  int col0, bool aux_vec_ok, float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
  const bool full = col0 + 32 <= p.N;
  if (p.bias) {
    for (int i = 0; i < 32; ++i)
      if (full || col0 + i < p.N) v[i] += __ldg(p.bias + col0 + i);
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
}

// Fused output quantisation for one 32 x 32 chunk held by a warp (lane = row, v[i] = column i).
// The values are first rounded to bf16 -- exactly what is stored in C -- so both copies are
// bit-identical to what the stand-alone quantisers would produce from C.
//   row-wise copy : the thread's 32 columns ARE one MX block of its row.
//   transposed    : a column's block is spread over the 32 lanes: warp-max for the scale, one
//                   byte per lane, a 32 x 32 byte transpose through shared memory, then lane i
//                   owns column i and writes its 32 bytes (rows row0..row0+31) contiguously.
__device__ __forceinline__ void fused_quant_chunk(const Fp8Params& p, const float (&v)[32], int row,
                                                  int row0, int col0, int lane, uint8_t* tile) {
  float vb[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) vb[i] = __bfloat162float(__float2bfloat16(v[i]));
  if (p.q_row != nullptr) {
    uint32_t pk[8];
    const uint8_t e = quant_block32(vb, pk);
    uint4* dp = reinterpret_cast<uint4*>(p.q_row + static_cast<long long>(row) * p.ld_qrow + col0);
    dp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    dp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    const int b = col0 >> 5;
    const int groups = p.M >> 7;
    p.sf_qrow[(static_cast<long long>(b >> 2) * groups + (row >> 7)) * 512 + (row & 31) * 16 +
              ((row & 127) >> 5) * 4 + (b & 3)] = e;
  }
  if (p.q_col != nullptr) {
    constexpr int LD = 36;  // padded row of the byte tile: conflict-free word writes / byte reads
    uint32_t words[8];
    uint8_t my_e = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      uint32_t packed = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int i = 4 * w + q;
        float a = fabsf(vb[i]);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, off));
        const int e = mx_exponent(a);
        if (lane == i) my_e = static_cast<uint8_t>(e + 127);
        const __nv_fp8_storage_t byte =
            __nv_cvt_float_to_fp8(vb[i] * exp2f(static_cast<float>(-e)), __NV_SATFINITE, __NV_E4M3);
        packed |= static_cast<uint32_t>(byte) << (8 * q);
      }
      words[w] = packed;
    }
    uint32_t* trow = reinterpret_cast<uint32_t*>(tile + lane * LD);
#pragma unroll
    for (int w = 0; w < 8; ++w) trow[w] = words[w];
    __syncwarp();
    uint32_t out[8];
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      uint32_t o = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) o |= static_cast<uint32_t>(tile[(4 * w + q) * LD + lane]) << (8 * q);
      out[w] = o;
    }
    __syncwarp();
    const int col = col0 + lane;  // row of the transposed operand owned by this lane
    uint4* dp = reinterpret_cast<uint4*>(p.q_col + static_cast<long long>(col) * p.ld_qcol + row0);
    dp[0] = make_uint4(out[0], out[1], out[2], out[3]);
    dp[1] = make_uint4(out[4], out[5], out[6], out[7]);
    p.sf_qcol[(static_cast<long long>(row0 >> 7) * p.groups_qcol + (col >> 7)) * 512 + (col & 31) * 16 +
              ((col & 127) >> 5) * 4 + ((row0 & 127) >> 5)] = my_e;
  }
}

template <int BN, bool QOUT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kF2Threads, 1)
gemm_mxfp8_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmSFA,
                       const __grid_constant__ CUtensorMap tmSFB,
                       const __grid_constant__ CUtensorMap tmC, const Fp8Params p,
                       const __grid_constant__ Fp8Rs rs) {
  using C = F2Cfg<BN>;
  constexpr int STAGES = C::STAGES;
  constexpr uint32_t STAGE_BYTES = C::STAGE_BYTES;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t smem_base = (raw_addr + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + STAGES * STAGE_BYTES;      // 1024-byte aligned (swizzled TMA source)
  const uint32_t bar_base = epi_base + C::EPI_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw_addr));
  // QOUT: 4 epilogue warps x (32 x 36 B) byte tiles for the transposed output copy
  uint8_t* qtiles = smem_raw + (tmem_slot + 16u - raw_addr);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_m = (p.M + 255) / 256;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = p.K / kFBK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmSFA);
    tma_prefetch_desc(&tmSFB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);   // the leader's arrive.expect_tx covers the bytes of BOTH CTAs
      mbar_init(empty_bar(s), 1);  // one multicast commit per use
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 8);  // 4 epilogue warps x 2 CTAs arrive on the leader's barrier
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<2>(tmem_slot, 512);
  tcgen05_fence_before();
  cluster_sync_all();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  auto tile_mn = [&](int tile, int& m_blk, int& n_blk) {
    const int per_group = 8 * num_n;
    const int group = tile / per_group;
    const int first_m = group * 8;
    const int gsize = min(num_m - first_m, 8);
    const int in_group = tile - group * per_group;
    m_blk = first_m + in_group % gsize;
    n_blk = in_group / gsize;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int m_blk, n_blk;
      tile_mn(tile, m_blk, n_blk);
      const int m0 = m_blk * 256 + static_cast<int>(cta_rank) * 128;
      const int n0 = n_blk * BN;
      const int sfb_atom0 = n0 >> 7;
      for (int kb = 0; kb < num_k; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t fb = mapa_shared(full_bar(stage), 0);
        const uint32_t a_dst = smem_base + stage * STAGE_BYTES;
        const uint32_t b_dst = a_dst + C::A_BYTES;
        const uint32_t sfa_dst = b_dst + C::B_BYTES;
        const uint32_t sfb_dst = sfa_dst + C::SFA_BYTES;
        if (elect_one_sync()) {
          if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * C::TX_BYTES);
          tma_load_2d_2sm(a_dst, &tmA, fb, kb * kFBK, m0);
          tma_load_2d_2sm(b_dst, &tmB, fb, kb * kFBK, n0 + static_cast<int>(cta_rank) * C::BNH);
          tma_load_2d_2sm(sfa_dst, &tmSFA, fb, 0, kb * p.groups_a + (m0 >> 7));
          tma_load_2d_2sm(sfb_dst, &tmSFB, fb, 0, kb * p.groups_b + sfb_atom0);
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader) {
      const uint32_t idesc0 = make_idesc_mxf8(0u, 0u, 256, BN, 0u, 0u);
      int stage = 0;
      uint32_t phase = 0;
      int tile_iter = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++tile_iter) {
        int m_blk, n_blk;
        tile_mn(tile, m_blk, n_blk);
        const uint32_t sfb_shift = static_cast<uint32_t>(((n_blk * BN) & 127) >> 5);
        const int acc = tile_iter & 1;
        const uint32_t acc_phase = (tile_iter >> 1) & 1;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t a_base = smem_base + stage * STAGE_BYTES;
          const uint32_t b_base = a_base + C::A_BYTES;
          const uint32_t sfa_s = b_base + C::B_BYTES;
          const uint32_t sfb_s = sfa_s + C::SFA_BYTES;
          if (elect_one_sync()) {
            tcgen05_cp_32x128b_warpx4_2sm(tmem_base + C::SFA_COL, make_smem_desc(sfa_s, 0, 128, kSwizzleNone));
#pragma unroll
            for (int a = 0; a < C::SFB_ATOMS; ++a)
              tcgen05_cp_32x128b_warpx4_2sm(tmem_base + C::SFB_COL + 4 * a,
                                            make_smem_desc(sfb_s + 512 * a, 0, 128, kSwizzleNone));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t da = make_smem_desc(a_base + k * 32, 16, 1024, kSwizzle128B);
              const uint64_t db = make_smem_desc(b_base + k * 32, 16, 1024, kSwizzle128B);
              const uint32_t idesc = idesc0 | (static_cast<uint32_t>(k) << 29) | (static_cast<uint32_t>(k) << 4);
              umma_mxf8_2sm(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u, tmem_base + C::SFA_COL,
                            tmem_base + C::SFB_COL + sfb_shift);
            }
            umma_commit_2sm(empty_bar(stage), 0x3);
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
    const int quad = warp & 3;
    int tile_iter = 0;
    uint32_t epi_it = 0;
    const bool aux_vec_ok =
        p.aux && ((p.ld_aux & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.aux) & 15) == 0);
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++tile_iter) {
      int m_blk, n_blk;
      tile_mn(tile, m_blk, n_blk);
      const int acc = tile_iter & 1;
      const uint32_t acc_phase = (tile_iter >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tcgen05_fence_after();
      const int row0 = m_blk * 256 + static_cast<int>(cta_rank) * 128 + quad * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
      const uint32_t tm = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN;
      if (!QOUT && p.c_tma == 2) {
        // fp32 output (weight gradients): 32 x 32 fp32 boxes leave through TMA, and an ACCUMULATING
        // call (one wgrad GEMM per micro-batch group of a pipeline step) becomes a TMA reduce-add --
        // dW is never read back by the SM (a float4 read-modify-write per thread made those GEMMs
        // memory-latency bound: 536 MB of dependent traffic per 8192^2 accumulate).
        const uint32_t my_epi = epi_base + static_cast<uint32_t>(warp - 2) * 2 * C::EPI_BUF_BYTES;
        const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;
        const uint32_t swz = static_cast<uint32_t>(lane & 7);
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          const int col0 = n_blk * BN + c;
          if (col0 >= p.N || row0 >= p.M) break;   // warp-uniform
          uint32_t r[32];
          tmem_ld_32x32b_x32(tm + c, r);
          const uint32_t buf = my_epi + (epi_it & 1u) * C::EPI_BUF_BYTES;
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          tmem_ld_wait();
          float v[32];
          if (row_ok) {
            fp8_epilogue_math(p, r, row, col0, aux_vec_ok, v);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
          }
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
          ++epi_it;
        }
      } else if (!QOUT && p.c_tma == 1) {
        // TMEM -> registers -> epilogue math -> 128B-swizzled smem box (32 rows x 64 bf16) -> TMA store
        const uint32_t my_epi = epi_base + static_cast<uint32_t>(warp - 2) * 2 * C::EPI_BUF_BYTES;
        const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;
        const uint32_t swz = static_cast<uint32_t>(lane & 7);
#pragma unroll 1
        for (int c = 0; c < BN; c += 64) {
          const int col0 = n_blk * BN + c;
          if (col0 >= p.N || row0 >= p.M) break;   // warp-uniform
          uint32_t r0[32], r1[32];
          tmem_ld_32x32b_x32(tm + c, r0);
          tmem_ld_32x32b_x32(tm + c + 32, r1);
          const uint32_t buf = my_epi + (epi_it & 1u) * C::EPI_BUF_BYTES;
          if (lane == 0) tma_store_wait_read<1>();   // the store that last used this buffer has read it
          __syncwarp();
          tmem_ld_wait();
          float v0[32], v1[32];
          if (row_ok) {
            fp8_epilogue_math(p, r0, row, col0, aux_vec_ok, v0);
            if (col0 + 32 < p.N) fp8_epilogue_math(p, r1, row, col0 + 32, aux_vec_ok, v1);
          }
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
            if (rs.world > 1) {
              const int owner = row0 / rs.rows_per_rank;
              tma_store_2d(&rs.map[owner], buf, col0, row0 - owner * rs.rows_per_rank);
            } else {
              tma_store_2d(&tmC, buf, col0, row0);   // rows >= M / columns >= N are clipped by the TMA unit
            }
            tma_store_commit();
          }
          ++epi_it;
        }
      } else {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n_blk * BN + c * 32;
        if (col0 >= p.N) break;
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm + c * 32, r);
        tmem_ld_wait();
        if (!row_ok) continue;
        float v[32];
        fp8_epilogue_math(p, r, row, col0, aux_vec_ok, v);
        fp8_store_chunk(p, v, row, col0);
        if constexpr (QOUT) {
          // host guarantees M % 128 == 0 and N % 32 == 0: whole warps, whole chunks
          fused_quant_chunk(p, v, row, row - lane, col0, lane, qtiles + (warp - 2) * (32 * 36));
        }
      }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tempty_bar(acc), 0);
    }
    if (lane == 0) {
      tma_store_wait<0>();   // smem must stay valid until the last store has read it; all stores performed
      if (rs.world > 1) {
        __threadfence_system();
        for (int r = 0; r < rs.world; ++r) red_add_release_sys(rs.done[r], 1u);
      }
    }
  }

  tcgen05_fence_before();
  cluster_sync_all();  // the peer's smem / barriers stay valid until every MMA and commit retired
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<2>(tmem_base, 512);
  }
}

}  // namespace sdml

using namespace sdml;

// Bytes needed for the scale-factor array of a [rows, cols] operand.
extern "C" long long sdml_mxfp8_sf_bytes(int rows, int cols) {
  return static_cast<long long>((cols + 127) / 128) * ((rows + 127) / 128) * 512;
}

// src bf16 [rows, cols] (ld_src) -> dst fp8 e4m3 [rows, cols] (ld_dst, bytes) + sf (layout above).
// cols % 128 == 0; src rows 16-byte aligned.  sf must hold sdml_mxfp8_sf_bytes(rows, cols) bytes;
// entries of rows >= `rows` inside the last 128-row group are left untouched (zero them once).
extern "C" int sdml_quant_mxfp8(const void* src, long long ld_src, int rows, int cols, void* dst,
                                long long ld_dst, void* sf, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0 || (cols & 127)) return set_error("sdml_quant_mxfp8: cols must be a multiple of 128");
  if ((ld_src & 7) || (ld_dst & 15)) return set_error("sdml_quant_mxfp8: ld_src % 8 and ld_dst % 16 must be 0");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  const long long total = static_cast<long long>(rows) * (cols / 32);
  long long blocks = (total + 255) / 256;
  const long long cap = 16LL * num_sms(dev);
  if (blocks > cap) blocks = cap;
  quant_mxfp8_kernel<<<(int)blocks, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(src), ld_src, rows, cols, static_cast<uint8_t*>(dst), ld_dst,
      static_cast<uint8_t*>(sf), (rows + 127) / 128);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// src bf16 [rows, cols] -> (optional) dst_row fp8 [rows, cols] scaled along cols + (optional)
// dst_col fp8 [cols, rows] (the transpose) scaled along the original rows, from one read of src.
// rows % 128 == 0, cols % 128 == 0.  ld_row / ld_col in bytes (multiples of 16).  sf_row holds
// sdml_mxfp8_sf_bytes(rows, cols) bytes, sf_col sdml_mxfp8_sf_bytes(cols, rows).  To fill a row
// range [r0, r0+rows) of a larger transposed operand (a micro-batch of a step-wide buffer) pass
// dst_col + r0, sf_col + (r0 / 128) * groups_col * 512 and groups_col = ceil(total_cols / 128).
extern "C" int sdml_quant_mxfp8_dual(const void* src, long long ld_src, int rows, int cols,
                                     void* dst_row, long long ld_row, void* sf_row,
                                     void* dst_col, long long ld_col, void* sf_col, int groups_col,
                                     cudaStream_t stream) {
  if (rows <= 0 || cols <= 0 || (rows & 127) || (cols & 127))
    return set_error("sdml_quant_mxfp8_dual: rows and cols must be positive multiples of 128");
  if (ld_src & 7) return set_error("sdml_quant_mxfp8_dual: ld_src % 8 must be 0");
  if (dst_row == nullptr && dst_col == nullptr) return set_error("sdml_quant_mxfp8_dual: no output requested");
  if (dst_row && ((ld_row & 15) || sf_row == nullptr)) return set_error("sdml_quant_mxfp8_dual: bad row-wise output");
  if (dst_col && ((ld_col & 15) || sf_col == nullptr || groups_col < (cols + 127) / 128))
    return set_error("sdml_quant_mxfp8_dual: bad transposed output");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  long long blocks = static_cast<long long>(rows / kQR) * (cols / kQC);
  const long long cap = 8LL * num_sms(dev);
  if (blocks > cap) blocks = cap;
  quant_mxfp8_dual_kernel<<<(int)blocks, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(src), ld_src, rows, cols, static_cast<uint8_t*>(dst_row), ld_row,
      static_cast<uint8_t*>(sf_row), (rows + 127) / 128, static_cast<uint8_t*>(dst_col), ld_col,
      static_cast<uint8_t*>(sf_col), groups_col);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// Same dual quantiser, writing the ROW-WISE copy into a row range of a larger operand: dst_row points
// at the first of the `rows` destination rows, sf_row at the scale atom of that row group inside an
// operand of groups_row = ceil(total_rows / 128) row groups (rows and the offset are multiples of 128).
// Lets every tensor-parallel rank quantise just its shard of a gathered operand in place.
extern "C" int sdml_quant_mxfp8_dual_at(const void* src, long long ld_src, int rows, int cols,
                                        void* dst_row, long long ld_row, void* sf_row, int groups_row,
                                        void* dst_col, long long ld_col, void* sf_col, int groups_col,
                                        cudaStream_t stream) {
  if (rows <= 0 || cols <= 0 || (rows & 127) || (cols & 127))
    return set_error("sdml_quant_mxfp8_dual_at: rows and cols must be positive multiples of 128");
  if (ld_src & 7) return set_error("sdml_quant_mxfp8_dual_at: ld_src % 8 must be 0");
  if (dst_row == nullptr && dst_col == nullptr) return set_error("sdml_quant_mxfp8_dual_at: no output requested");
  if (dst_row && ((ld_row & 15) || sf_row == nullptr || groups_row < rows / 128))
    return set_error("sdml_quant_mxfp8_dual_at: bad row-wise output");
  if (dst_col && ((ld_col & 15) || sf_col == nullptr || groups_col < (cols + 127) / 128))
    return set_error("sdml_quant_mxfp8_dual_at: bad transposed output");
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  long long blocks = static_cast<long long>(rows / kQR) * (cols / kQC);
  const long long cap = 8LL * num_sms(dev);
  if (blocks > cap) blocks = cap;
  quant_mxfp8_dual_kernel<<<(int)blocks, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(src), ld_src, rows, cols, static_cast<uint8_t*>(dst_row), ld_row,
      static_cast<uint8_t*>(sf_row), groups_row, static_cast<uint8_t*>(dst_col), ld_col,
      static_cast<uint8_t*>(sf_col), groups_col);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// C[M,N] = epilogue(alpha * (A . sfa)[M,K] . (B . sfb)[N,K]^T); A, B fp8 e4m3 K-major
// (lda / ldb in bytes, multiples of 16), K % 128 == 0, scale arrays from sdml_quant_mxfp8(_dual).
// accumulate != 0 (fp32 output only): C += ... instead of C = ....
extern "C" int sdml_gemm_mxfp8(const void* A, const void* sfa, const void* B, const void* sfb,
                               void* C, int M, int N, int K, long long lda, long long ldb,
                               long long ldc, const float* bias, const void* aux, long long ld_aux,
                               int relu, int out_fp32, int accumulate, float alpha, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 127)) return set_error("sdml_gemm_mxfp8: K must be a positive multiple of 128");
  if (accumulate && !out_fp32) return set_error("sdml_gemm_mxfp8: accumulate needs an fp32 output");
  if ((lda & 15) || (ldb & 15)) return set_error("sdml_gemm_mxfp8: lda / ldb must be multiples of 16 bytes");
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, TMAP_U8, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kFBK, kFBM, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_2d(&tmB, B, TMAP_U8, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, kFBK, kFBN, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  Fp8Params p{};
  p.M = M; p.N = N; p.K = K;
  p.C = C; p.ldc = ldc;
  p.bias = bias;
  p.aux = static_cast<const __nv_bfloat16*>(aux);
  p.ld_aux = ld_aux;
  p.alpha = alpha; p.relu = relu; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.sfa = static_cast<const uint8_t*>(sfa);
  p.sfb = static_cast<const uint8_t*>(sfb);
  p.groups_a = (M + 127) / 128;
  p.groups_b = (N + 127) / 128;
  constexpr size_t smem = size_t(kFStages) * (kFBM * kFBK + kFBN * kFBK + 1024) + 1024 + 256;
  static std::atomic<bool> configured[64];
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    SDML_CUDA_OK(cudaFuncSetAttribute(gemm_mxfp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev & 63] = true;
  }
  const int num_tiles = ((M + kFBM - 1) / kFBM) * ((N + kFBN - 1) / kFBN);
  const int grid = num_tiles < num_sms(dev) ? num_tiles : num_sms(dev);
  gemm_mxfp8_kernel<<<grid, kFThreads, smem, stream>>>(tmA, tmB, p);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// The TMA-store epilogue pays for its shared-memory staging pass when C is local (8-layer MLP step
// 27.7 vs 26.2 ms on one B200) and wins by a wide margin when C is a pipeline neighbour's receive
// slot: per-thread 64-byte stores crawl over NVLink, whole 128-byte lines do not.  So: peer memory
// -> TMA stores, local memory -> direct stores.  SDML_FP8_TMA_STORE=1 / 0 forces one form (tests).
static bool want_tma_store(const void* C, int dev) {
  static const int forced = [] {
    const char* e = getenv("SDML_FP8_TMA_STORE");
    return e == nullptr ? -1 : (e[0] == '1' ? 1 : 0);
  }();
  if (forced >= 0) return forced == 1;
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, C) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return attr.type == cudaMemoryTypeDevice && attr.device != dev;
}

// fp32 accumulation through TMA reduce-add (default) or the per-thread read-modify-write
// (SDML_FP8_TMA_REDUCE=0; kept for A/B measurements and the equivalence test)
static bool want_tma_reduce() {
  static const bool on = [] {
    const char* e = getenv("SDML_FP8_TMA_REDUCE");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

template <int BN, bool QOUT>
static int launch_mxfp8_2cta(const CUtensorMap& tmA, const void* B, long long ldb, const Fp8Params& p,
                             cudaStream_t stream, const Fp8Rs* rs_in = nullptr) {
  using C = F2Cfg<BN>;
  CUtensorMap tmB, tmSFA, tmSFB;
  int rc = make_tmap_2d(&tmB, B, TMAP_U8, (uint64_t)p.K, (uint64_t)p.N, (uint64_t)ldb, kFBK, C::BNH, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  const uint64_t kblocks = (uint64_t)(p.K / kFBK);
  // scale arrays as [atoms, 128] fp32: one box row = one 512-byte atom
  rc = make_tmap_2d(&tmSFA, p.sfa, TMAP_F32, 128, kblocks * p.groups_a, 128, 128, 1, TMAP_SWIZZLE_NONE);
  if (rc) return rc;
  rc = make_tmap_2d(&tmSFB, p.sfb, TMAP_F32, 128, kblocks * p.groups_b, 128, 128, C::SFB_ATOMS, TMAP_SWIZZLE_NONE);
  if (rc) return rc;
  static std::atomic<bool> configured[64];
  int dev = 0;
  SDML_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    SDML_CUDA_OK(cudaFuncSetAttribute(gemm_mxfp8_2cta_kernel<BN, QOUT>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
    configured[dev & 63] = true;
  }
  const int num_tiles = ((p.M + 255) / 256) * ((p.N + BN - 1) / BN);
  int clusters = num_sms(dev) / 2;
  if (clusters > num_tiles) clusters = num_tiles;
  // bf16 output with 16-byte aligned rows leaves through TMA stores of 32 x 64 boxes
  Fp8Params pp = p;
  CUtensorMap tmC = tmA;
  pp.c_tma = 0;
  if (!QOUT && !p.out_fp32 && ((p.ldc * 2) % 16 == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0) &&
      want_tma_store(p.C, dev)) {
    rc = make_tmap_2d(&tmC, p.C, TMAP_BF16, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ldc, 64, 32, TMAP_SWIZZLE_128B);
    if (rc) return rc;
    pp.c_tma = 1;
  } else if (!QOUT && p.out_fp32 && p.accumulate && ((p.ldc * 4) % 16 == 0) &&
             ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0) && want_tma_reduce()) {
    rc = make_tmap_2d(&tmC, p.C, TMAP_F32, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ldc, 32, 32, TMAP_SWIZZLE_128B);
    if (rc) return rc;
    pp.c_tma = 2;
  }
  static Fp8Rs no_rs = [] {
    Fp8Rs z;
    memset(&z, 0, sizeof(z));
    return z;
  }();
  const Fp8Rs& rs = rs_in ? *rs_in : no_rs;
  if (rs.world > 1) {
    // every CTA's epilogue warps take part in the completion counters: always the full grid,
    // and the partial result leaves through the per-owner maps
    clusters = num_sms(dev) / 2;
    pp.c_tma = 1;
  }
  gemm_mxfp8_2cta_kernel<BN, QOUT><<<2 * clusters, kF2Threads, C::SMEM, stream>>>(tmA, tmB, tmSFA, tmSFB, tmC, pp, rs);
  SDML_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

// Same contract as sdml_gemm_mxfp8 on the CTA-pair kernel; bn = 192 (default tile) or 128.
// Additionally needs both scale arrays 16-byte aligned.
extern "C" int sdml_gemm_mxfp8_2cta(const void* A, const void* sfa, const void* B, const void* sfb,
                                    void* C, int M, int N, int K, long long lda, long long ldb,
                                    long long ldc, const float* bias, const void* aux, long long ld_aux,
                                    int relu, int out_fp32, int accumulate, float alpha, int bn,
                                    cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 127)) return set_error("sdml_gemm_mxfp8_2cta: K must be a positive multiple of 128");
  if ((lda & 15) || (ldb & 15)) return set_error("sdml_gemm_mxfp8_2cta: lda / ldb must be multiples of 16 bytes");
  if (accumulate && !out_fp32) return set_error("sdml_gemm_mxfp8_2cta: accumulate needs an fp32 output");
  if ((reinterpret_cast<uintptr_t>(sfa) & 15) || (reinterpret_cast<uintptr_t>(sfb) & 15))
    return set_error("sdml_gemm_mxfp8_2cta: scale arrays must be 16-byte aligned");
  CUtensorMap tmA;
  int rc = make_tmap_2d(&tmA, A, TMAP_U8, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kFBK, 128, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  Fp8Params p{};
  p.M = M; p.N = N; p.K = K;
  p.C = C; p.ldc = ldc;
  p.bias = bias;
  p.aux = static_cast<const __nv_bfloat16*>(aux);
  p.ld_aux = ld_aux;
  p.alpha = alpha; p.relu = relu; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.sfa = static_cast<const uint8_t*>(sfa);
  p.sfb = static_cast<const uint8_t*>(sfb);
  p.groups_a = (M + 127) / 128;
  p.groups_b = (N + 127) / 128;
  p.q_row = nullptr; p.ld_qrow = 0; p.sf_qrow = nullptr;
  p.q_col = nullptr; p.ld_qcol = 0; p.sf_qcol = nullptr; p.groups_qcol = 0;
  if (bn == 128) return launch_mxfp8_2cta<128, false>(tmA, B, ldb, p, stream);
  if (bn == 192 || bn == 0) return launch_mxfp8_2cta<192, false>(tmA, B, ldb, p, stream);
  return set_error("sdml_gemm_mxfp8_2cta: bn must be 128 or 192");
}

// MX-fp8 GEMM -> reduce-scatter (tensor-parallel row-parallel layer, fp8 operands): the bf16 partial
// result A[M,K] . B[N,K]^T * alpha never exists as a whole -- its rows go straight into the landing
// slots of their owners (rs_dst[r]: HOST array, for every owner r the peer-mapped address of the
// slot reserved for THIS rank in r's memory, bf16 [M / rs_world, rs_ld]); rs_done: HOST array of the
// owners' completion counters.  Finish with sdml_rs_finalize on every owner (expected count per
// call: rs_world * 4 * sdml_gemm_comm_grid()).  M % (32 * rs_world) == 0, K % 128 == 0.
// Counterpart of ref:75-77 (fc1 / fc2) sharded over GPUs; the reference never shards a layer.
extern "C" int sdml_gemm_mxfp8_2cta_rs(const void* A, const void* sfa, const void* B, const void* sfb,
                                       int M, int N, int K, long long lda, long long ldb, float alpha,
                                       int rs_world, int rs_rank, void* const* rs_dst, long long rs_ld,
                                       void* const* rs_done, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 127)) return set_error("sdml_gemm_mxfp8_2cta_rs: K must be a positive multiple of 128");
  if ((lda & 15) || (ldb & 15)) return set_error("sdml_gemm_mxfp8_2cta_rs: lda / ldb must be multiples of 16 bytes");
  if ((reinterpret_cast<uintptr_t>(sfa) & 15) || (reinterpret_cast<uintptr_t>(sfb) & 15))
    return set_error("sdml_gemm_mxfp8_2cta_rs: scale arrays must be 16-byte aligned");
  if (rs_world < 2 || rs_world > kFp8MaxRanks || rs_rank < 0 || rs_rank >= rs_world || !rs_dst || !rs_done)
    return set_error("sdml_gemm_mxfp8_2cta_rs: 2..8 ranks and their slot / counter arrays");
  if (M % (32 * rs_world) || (rs_ld & 7)) return set_error("sdml_gemm_mxfp8_2cta_rs: M % (32*world) == 0 and rs_ld % 8 == 0 required");
  CUtensorMap tmA;
  int rc = make_tmap_2d(&tmA, A, TMAP_U8, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kFBK, 128, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  Fp8Rs rs;
  memset(&rs, 0, sizeof(rs));
  rs.world = rs_world;
  rs.rows_per_rank = M / rs_world;
  for (int r = 0; r < rs_world; ++r) {
    if ((reinterpret_cast<uintptr_t>(rs_dst[r]) & 15)) return set_error("sdml_gemm_mxfp8_2cta_rs: landing slots must be 16-byte aligned");
    rc = make_tmap_2d(&rs.map[r], rs_dst[r], TMAP_BF16, (uint64_t)N, (uint64_t)(M / rs_world), (uint64_t)rs_ld,
                      64, 32, TMAP_SWIZZLE_128B);
    if (rc) return rc;
    rs.done[r] = static_cast<uint32_t*>(rs_done[r]);
  }
  Fp8Params p{};
  p.M = M; p.N = N; p.K = K;
  p.C = rs_dst[rs_rank]; p.ldc = rs_ld;      // (unused by the reduce-scatter epilogue)
  p.alpha = alpha; p.relu = 0; p.out_fp32 = 0; p.accumulate = 0;
  p.sfa = static_cast<const uint8_t*>(sfa);
  p.sfb = static_cast<const uint8_t*>(sfb);
  p.groups_a = (M + 127) / 128;
  p.groups_b = (N + 127) / 128;
  return launch_mxfp8_2cta<192, false>(tmA, B, ldb, p, stream, &rs);
}

// CTA-pair GEMM (256 x 192 tiles) whose epilogue ALSO writes MX-fp8 copies of the bf16 output:
//   q_row / sf_qrow : C [M, N] quantised along N (operand of a GEMM that contracts over N), optional
//   q_col / sf_qcol : C^T [N, M] quantised along M (operand of a GEMM that contracts over M),
//                     optional; to fill rows [r0, r0 + M) of a step-wide transposed buffer pass
//                     q_col + r0 and sf_qcol + (r0 / 128) * groups_qcol * 512
// Both copies are bit-identical to sdml_quant_mxfp8_dual applied to C.  Needs a bf16 output,
// M % 128 == 0, N % 128 == 0, 16-byte aligned pointers / leading dimensions.
extern "C" int sdml_gemm_mxfp8_2cta_qout(const void* A, const void* sfa, const void* B, const void* sfb,
                                         void* C, int M, int N, int K, long long lda, long long ldb,
                                         long long ldc, const float* bias, const void* aux,
                                         long long ld_aux, int relu, float alpha, void* q_row,
                                         long long ld_qrow, void* sf_qrow, void* q_col,
                                         long long ld_qcol, void* sf_qcol, int groups_qcol,
                                         cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 127) || (M & 127) || (N & 127))
    return set_error("sdml_gemm_mxfp8_2cta_qout: M, N, K must be positive multiples of 128");
  if ((lda & 15) || (ldb & 15)) return set_error("sdml_gemm_mxfp8_2cta_qout: lda / ldb must be multiples of 16 bytes");
  if ((reinterpret_cast<uintptr_t>(sfa) & 15) || (reinterpret_cast<uintptr_t>(sfb) & 15))
    return set_error("sdml_gemm_mxfp8_2cta_qout: scale arrays must be 16-byte aligned");
  if (q_row == nullptr && q_col == nullptr) return set_error("sdml_gemm_mxfp8_2cta_qout: no quantised output requested");
  if (q_row && ((ld_qrow & 15) || (reinterpret_cast<uintptr_t>(q_row) & 15) || sf_qrow == nullptr))
    return set_error("sdml_gemm_mxfp8_2cta_qout: bad row-wise output");
  if (q_col && ((ld_qcol & 15) || (reinterpret_cast<uintptr_t>(q_col) & 15) || sf_qcol == nullptr ||
                groups_qcol < N / 128))
    return set_error("sdml_gemm_mxfp8_2cta_qout: bad transposed output");
  CUtensorMap tmA;
  int rc = make_tmap_2d(&tmA, A, TMAP_U8, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kFBK, 128, TMAP_SWIZZLE_128B);
  if (rc) return rc;
  Fp8Params p{};
  p.M = M; p.N = N; p.K = K;
  p.C = C; p.ldc = ldc;
  p.bias = bias;
  p.aux = static_cast<const __nv_bfloat16*>(aux);
  p.ld_aux = ld_aux;
  p.alpha = alpha; p.relu = relu; p.out_fp32 = 0; p.accumulate = 0;
  p.sfa = static_cast<const uint8_t*>(sfa);
  p.sfb = static_cast<const uint8_t*>(sfb);
  p.groups_a = M / 128;
  p.groups_b = N / 128;
  p.q_row = static_cast<uint8_t*>(q_row); p.ld_qrow = ld_qrow; p.sf_qrow = static_cast<uint8_t*>(sf_qrow);
  p.q_col = static_cast<uint8_t*>(q_col); p.ld_qcol = ld_qcol; p.sf_qcol = static_cast<uint8_t*>(sf_qcol);
  p.groups_qcol = groups_qcol;
  return launch_mxfp8_2cta<192, true>(tmA, B, ldb, p, stream);
}
