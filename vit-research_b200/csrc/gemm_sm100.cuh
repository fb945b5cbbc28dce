// gemm_sm100.cuh — persistent warp-specialised tcgen05 GEMM  C[M,N] = A[M,K] · W[N,K]^T (+ epilogue)
//
// Replaces, for the hot path, every dense contraction the reference runs through cuBLAS/cuDNN:
//   patch embedding  (HF modeling_vit.py:131-167 — Conv2d k16 s16 == GEMM over im2col'd patches)
//   Q/K/V projection (modeling_vit.py:216-230, fused into one N=3D GEMM)
//   attention output (modeling_vit.py:254-268) + residual (modeling_vit.py:337)
//   MLP fc1 + GELU   (modeling_vit.py:287-299), fc2 + residual (modeling_vit.py:302-312)
//   similarity GEMM  (nba_proj/retrieval/ratt_chunk_retriever.py:123  sims = Q @ C.T)
//
// Shape of the kernel (one CTA per SM, 384 threads):
//   warp 0      TMA producer: A tile 128x64 and W tile BNx64 (bf16, 128B swizzle) per k-block
//   warp 1      MMA issuer: one thread issues tcgen05.mma (M=128, N=BN, K=16) x4 per k-block,
//               accumulating in TMEM; tcgen05.commit releases smem stages / publishes the tile
//   warp 2      TMEM allocator (2 accumulator stages x BN columns)
//   warps 4-11  epilogue: tcgen05.ld TMEM->registers, bias/GELU/residual/table, swizzled
//               st.shared into a per-warp staging tile, TMA store to global (coalesced, async)
// The TMEM accumulator is double buffered so the epilogue of tile i overlaps the MMAs of tile i+1.
#pragma once
#include "vitk_common.cuh"

namespace vitk {

enum GemmEpilogue : int {
  EPI_BIAS_BF16 = 0,       // out bf16 = acc + bias[n]
  EPI_BIAS_GELU_BF16 = 1,  // out bf16 = gelu_erf(acc + bias[n])
  EPI_BIAS_GELUT_BF16 = 2, // out bf16 = gelu_tanh(acc + bias[n])
  EPI_BIAS_RESID_F32 = 3,  // out f32  = acc + bias[n] + resid[m,n]   (resid may alias out)
  EPI_TABLE_F32 = 4,       // out f32  = acc + table[m % table_rows, n]  (patch embed: pos+bias/CLS)
  EPI_F32 = 5,             // out f32  = acc                              (similarity scores)
};

struct GemmParams {
  int M, N, K;
  const float* bias;   // [N]
  const float* resid;  // [M, ldr]
  int ldr;
  const float* table;  // [table_rows, N]
  int table_rows;
};

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 384;
constexpr int GEMM_EPI_WARPS = 8;
constexpr int GEMM_STAGING_BYTES = 4096;  // 32 rows x 128 B per buffer

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = GEMM_BM * GEMM_BK * 2;  // 16 KB
  static constexpr int kBBytes = BN * GEMM_BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagingBytes = GEMM_EPI_WARPS * GEMM_STAGING_BYTES;  // 32 KB
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kBarBytes + 1024;
  static constexpr uint32_t kTmemCols = 2 * BN;
};

template <int EPI>
struct EpiTraits {
  static constexpr bool kOutBf16 = (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_GELU_BF16 ||
                                    EPI == EPI_BIAS_GELUT_BF16);
  static constexpr int kChunkCols = kOutBf16 ? 64 : 32;  // 128 B of output per row per chunk
};

template <int BN, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  constexpr int CH = EpiTraits<EPI>::kChunkCols;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * Cfg::kABytes;
  uint8_t* sStage = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStage + Cfg::kStagingBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tfull_bar = bars + 2 * kStages;
  uint64_t* tempty_bar = bars + 2 * kStages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (p.M + GEMM_BM - 1) / GEMM_BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = p.K / GEMM_BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], GEMM_EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          tma_load_2d(sA + stage * Cfg::kABytes, &tmA, &full_bar[stage], kb * GEMM_BK,
                      m_blk * GEMM_BM);
          tma_load_2d(sB + stage * Cfg::kBBytes, &tmB, &full_bar[stage], kb * GEMM_BK, n_blk * BN);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(GEMM_BM, BN, false, false);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint64_t da = make_smem_desc_sw128(smem_u32(sA + stage * Cfg::kABytes), 1024, 0);
          const uint64_t db = make_smem_desc_sw128(smem_u32(sB + stage * Cfg::kBBytes), 1024, 0);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            // advance 16 bf16 = 32 B along K inside the 128 B swizzle atom: +2 in the >>4 field
            umma_bf16_ss(d_tmem, da + static_cast<uint64_t>(2 * k), db + static_cast<uint64_t>(2 * k),
                         idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // smem stage reusable once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&tfull_bar[acc]);  // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue ------------------------------
    const int ew = warp - 4;            // 0..7
    const int q = warp & 3;             // TMEM lane quarter this warp may access
    const int half = ew >> 2;           // which half of the BN columns
    constexpr int kColsPerWarp = BN / 2;
    constexpr int kChunks = kColsPerWarp / CH;
    uint8_t* sbuf = sStage + ew * GEMM_STAGING_BYTES;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int row0 = m_blk * GEMM_BM + q * 32;
      const int row = row0 + lane;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after_sync();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                             static_cast<uint32_t>(acc * BN + half * kColsPerWarp);
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c) {
        const int col0 = n_blk * BN + half * kColsPerWarp + c * CH;
        if constexpr (CH == 32) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(t_row + c * CH, r);
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          if constexpr (EPI == EPI_BIAS_RESID_F32) {
            if (row < p.M) {
              const float4* rp = reinterpret_cast<const float4*>(p.resid + (size_t)row * p.ldr + col0);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (col0 + 4 * j < p.N) {
                  float4 x = rp[j];
                  v[4 * j + 0] += x.x; v[4 * j + 1] += x.y; v[4 * j + 2] += x.z; v[4 * j + 3] += x.w;
                }
              }
            }
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < p.N) v[j] += __ldg(p.bias + col0 + j);
          } else if constexpr (EPI == EPI_TABLE_F32) {
            if (row < p.M) {
              const float4* tp = reinterpret_cast<const float4*>(
                  p.table + (size_t)(row % p.table_rows) * p.N + col0);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (col0 + 4 * j < p.N) {
                  float4 x = __ldg(tp + j);
                  v[4 * j + 0] += x.x; v[4 * j + 1] += x.y; v[4 * j + 2] += x.z; v[4 * j + 3] += x.w;
                }
              }
            }
          }
          // the TMA store that last read the staging tile must have finished reading it
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 o = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            *reinterpret_cast<float4*>(sbuf + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
          }
        } else {
          uint32_t r0[32], r1[32];
          tmem_ld_32x32b_x32(t_row + c * CH, r0);
          tmem_ld_32x32b_x32(t_row + c * CH + 32, r1);
          tmem_ld_wait();
          float v[64];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] = __uint_as_float(r0[j]);
            v[32 + j] = __uint_as_float(r1[j]);
          }
#pragma unroll
          for (int j = 0; j < 64; ++j) {
            float b = (col0 + j < p.N) ? __ldg(p.bias + col0 + j) : 0.f;
            float x = v[j] + b;
            if constexpr (EPI == EPI_BIAS_GELU_BF16) x = gelu_erf(x);
            if constexpr (EPI == EPI_BIAS_GELUT_BF16) x = gelu_tanh(x);
            v[j] = x;
          }
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 o;
            o.x = pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
            o.y = pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
            o.z = pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
            o.w = pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
            *reinterpret_cast<uint4*>(sbuf + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
          }
        }
        if (c == kChunks - 1) {
          // all TMEM reads of this accumulator stage are done: hand it back to the MMA warp
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (col0 < p.N && row0 < p.M) tma_store_2d(&tmC, sbuf, col0, row0);
          tma_store_commit();
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

}  // namespace vitk
