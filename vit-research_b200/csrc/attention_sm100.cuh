// attention_sm100.cuh — fused softmax(Q K^T / sqrt(d)) V for short sequences (S <= 256, d = 64)
//
// Replaces HF ViTSelfAttention / SDPA (modeling_vit.py:171-196, 220-252) for the 197-token
// ViT-B/16 shape.  One CTA (128 threads) per (frame, head, 128-query block); the whole K and V
// of the head live in shared memory (S*64*2 B = 25 KB each), so there is no online-softmax
// tiling.  Scores and probabilities never leave the SM:
//   S = Q_blk K^T     tcgen05.mma M=128, N=SK (S rounded up to 16), K=64   -> TMEM cols [0,SK) f32
//   softmax           one thread per query row (= TMEM lane): tcgen05.ld, max, exp2, sum
//   P (bf16)          tcgen05.st back into TMEM cols [0,SK/2), two bf16 per 32-bit cell,
//                     overwriting the S columns this thread has already consumed
//   O = P V           tcgen05.mma with A = P read FROM TMEM, B = V in smem consumed MN-major as
//                     TMA delivered it, K = SK; accumulator in TMEM cols [SK/2, SK/2+64)
//   O / rowsum -> bf16 -> staged in Q's shared memory -> TMA store (3-D map clips tokens >= S)
// v1 kept P in shared memory (190 KB/CTA, one CTA/SM, every phase serialised: 232 us/layer,
// 8.6 % tensor pipe, profiles/r01_v1_ncu_attention.txt).  With P in TMEM a CTA needs 70 KB of
// smem and 256 TMEM columns, so two CTAs are resident per SM and one's loads / MMAs overlap
// the other's softmax.
#pragma once
#include "vitk_common.cuh"

namespace vitk {

constexpr int ATT_THREADS = 128;
constexpr int ATT_D = 64;
constexpr uint32_t ATT_TMEM_COLS = 256;
// Two CTAs per SM by construction (2 x 256 TMEM columns = all 512): ask for at least this much
// dynamic smem so a third CTA is never co-scheduled and left spinning inside tcgen05.alloc.
constexpr int ATT_MIN_SMEM = 78 * 1024;

struct AttnSmemLayout {
  int sk;  // keys rounded up to 16 (MMA N / K extent)
  int off_q, off_k, off_v, off_bar, total;
};

__host__ __device__ inline AttnSmemLayout attn_smem_layout(int S) {
  AttnSmemLayout L;
  L.sk = (S + 15) & ~15;
  int kv_bytes = ((L.sk * 128) + 1023) & ~1023;
  L.off_q = 0;
  L.off_k = 16384;
  L.off_v = L.off_k + kv_bytes;
  L.off_bar = L.off_v + kv_bytes;
  L.total = L.off_bar + 64 + 1024;  // + alignment slack
  if (L.total < ATT_MIN_SMEM) L.total = ATT_MIN_SMEM;
  return L;
}

struct AttnParams {
  int S;        // tokens per frame
  int heads;
  int frames;
  int q_rows;   // queries per frame that are needed (== S, or fewer: last-layer CLS shortcut)
  int nblk;     // query blocks per (frame, head): 1 if q_rows <= 128 else 2
  float scale_log2e;  // (1/sqrt(d)) * log2(e)
};

// tmQ : 3-D map over qkv [frames, S, 3*D] bf16, box {64, 128, 1}   (Q block)
// tmKV: same tensor,                          box {64, sk, 1}      (K and V)
// tmO : 3-D map over out [frames, S, D] bf16, box {64, 128, 1}
__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                 const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const AttnSmemLayout L = attn_smem_layout(p.S);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* bar_qk = bars + 0;
  uint64_t* bar_v = bars + 1;
  uint64_t* bar_s = bars + 2;
  uint64_t* bar_o = bars + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int blk = blockIdx.x % p.nblk;              // query block (fastest: K/V re-read hits L2)
  const int fh = blockIdx.x / p.nblk;
  const int frame = fh / p.heads;
  const int head = fh % p.heads;
  const int D = p.heads * ATT_D;
  const int S = p.S, SK = L.sk;
  const int rows_valid = min(128, p.q_rows - blk * 128);  // query rows of this block needed
  const bool warp_active = warp * 32 < rows_valid;  // warps whose 32 rows are all padding idle

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmO);
    mbar_init(bar_qk, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_mbar_init();
    // ---- loads: Q block + K on one barrier, V on another (issued before the TMEM alloc) ----
    const uint32_t kv_bytes = SK * 128;
    mbar_arrive_expect_tx(bar_qk, 128 * 128 + kv_bytes);
    tma_load_3d(smem + L.off_q, &tmQ, bar_qk, head * ATT_D, blk * 128, frame);
    tma_load_3d(smem + L.off_k, &tmKV, bar_qk, D + head * ATT_D, 0, frame);
    mbar_arrive_expect_tx(bar_v, kv_bytes);
    tma_load_3d(smem + L.off_v, &tmKV, bar_v, 2 * D + head * ATT_D, 0, frame);
  }
  if (warp == 1) tmem_alloc<ATT_TMEM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_p = tmem_base;                                 // bf16 pairs, cols [0,SK/2)
  const uint32_t tmem_o = tmem_base + static_cast<uint32_t>(SK / 2); // f32, 64 cols

  if (tid == 0) {
    // ---- S = Q K^T ----
    mbar_wait(bar_qk, 0);
    tc_fence_after_sync();
    const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
    const uint64_t dk = make_smem_desc_sw128(smem_u32(smem + L.off_k), 1024, 0);
    const uint64_t dq = make_smem_desc_sw128(smem_u32(smem + L.off_q), 1024, 0);
#pragma unroll
    for (int k = 0; k < ATT_D / 16; ++k)
      umma_bf16_ss(tmem_base, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
    umma_commit(bar_s);
  }

  const int r = warp * 32 + lane;  // row within the block == TMEM lane
  const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
  float sum = 0.f;
  if (warp_active) {
    mbar_wait(bar_s, 0);
    tc_fence_after_sync();
    const uint32_t t_s = tmem_base + lane_off;
    // pass 1: row max over the S valid key columns (32-column loads; SK is a multiple of 16)
    float mx = -INFINITY;
    int c0 = 0;
    for (; c0 + 32 <= SK; c0 += 32) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(t_s + c0, v);
      tmem_ld_wait();
      if (c0 + 32 <= S) {
#pragma unroll
        for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(v[j]));
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (c0 + j < S) mx = fmaxf(mx, __uint_as_float(v[j]));
      }
    }
    if (c0 < SK) {
      uint32_t v[16];
      tmem_ld_32x32b_x16(t_s + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < S) mx = fmaxf(mx, __uint_as_float(v[j]));
    }
    const float mxs = mx * p.scale_log2e;
    // pass 2: p = exp2(s*scale*log2e - max*scale*log2e) -> bf16 pairs back into TMEM.  The P
    // cells [c0/2, c0/2+16) written in iteration c0 only cover S columns < c0+32 that this
    // thread has already loaded (tcgen05.wait::ld above), so the in-place overwrite is safe.
    const uint32_t t_p = tmem_p + lane_off;
    for (c0 = 0; c0 + 32 <= SK; c0 += 32) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(t_s + c0, v);
      tmem_ld_wait();
      uint32_t pk[16];
      const bool full = c0 + 32 <= S;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float e0 = ex2_approx(fmaf(__uint_as_float(v[2 * j]), p.scale_log2e, -mxs));
        float e1 = ex2_approx(fmaf(__uint_as_float(v[2 * j + 1]), p.scale_log2e, -mxs));
        if (!full) {
          e0 = (c0 + 2 * j < S) ? e0 : 0.f;
          e1 = (c0 + 2 * j + 1 < S) ? e1 : 0.f;
        }
        sum += e0 + e1;
        pk[j] = pack_bf16x2(e0, e1);
      }
      tmem_st_32x32b_x16(t_p + (c0 >> 1), pk);
    }
    if (c0 < SK) {
      uint32_t v[16];
      tmem_ld_32x32b_x16(t_s + c0, v);
      tmem_ld_wait();
      uint32_t pk[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float e0 = ex2_approx(fmaf(__uint_as_float(v[2 * j]), p.scale_log2e, -mxs));
        float e1 = ex2_approx(fmaf(__uint_as_float(v[2 * j + 1]), p.scale_log2e, -mxs));
        e0 = (c0 + 2 * j < S) ? e0 : 0.f;
        e1 = (c0 + 2 * j + 1 < S) ? e1 : 0.f;
        sum += e0 + e1;
        pk[j] = pack_bf16x2(e0, e1);
      }
      tmem_st_32x32b_x8(t_p + (c0 >> 1), pk);
    }
    tmem_st_wait();
  }
  // P complete in TMEM, S fully consumed by every row of the block
  tc_fence_before_sync();
  __syncthreads();

  if (tid == 0) {
    // ---- O = P V : A = P from TMEM (8 cells per K=16 step), V MN-major (tokens x 64 dims) ----
    tc_fence_after_sync();
    mbar_wait(bar_v, 0);
    tc_fence_after_sync();
    const uint32_t idesc_o = make_idesc_bf16_f32(128, ATT_D, false, true);
    for (int kk = 0; kk < SK / 16; ++kk) {
      // 16 K rows (tokens) of V = 2 groups of 8 rows x 128 B
      const uint64_t dv =
          make_smem_desc_sw128(smem_u32(smem + L.off_v) + kk * 16 * 128, 1024, 1024);
      umma_bf16_ts(tmem_o, tmem_p + kk * 8, dv, idesc_o, kk != 0 ? 1u : 0u);
    }
    umma_commit(bar_o);
  }

  // ---- epilogue: O / rowsum -> bf16 -> Q's smem tile (dead by now) -> TMA store ----
  uint8_t* sO = smem + L.off_q;
  if (warp_active) {
    mbar_wait(bar_o, 0);
    tc_fence_after_sync();
    const float inv = 1.0f / sum;
    const uint32_t t_o = tmem_o + lane_off;
#pragma unroll
    for (int c0 = 0; c0 < ATT_D; c0 += 32) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(t_o + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(v[8 * j + 0]) * inv, __uint_as_float(v[8 * j + 1]) * inv);
        o.y = pack_bf16x2(__uint_as_float(v[8 * j + 2]) * inv, __uint_as_float(v[8 * j + 3]) * inv);
        o.z = pack_bf16x2(__uint_as_float(v[8 * j + 4]) * inv, __uint_as_float(v[8 * j + 5]) * inv);
        o.w = pack_bf16x2(__uint_as_float(v[8 * j + 6]) * inv, __uint_as_float(v[8 * j + 7]) * inv);
        const int ci = (c0 >> 3) + j;
        *reinterpret_cast<uint4*>(sO + r * 128 + ((ci ^ (r & 7)) << 4)) = o;
      }
    }
    fence_proxy_async_smem();
  }
  tc_fence_before_sync();
  __syncthreads();
  if (tid == 0) {
    // rows of idle warps hold stale Q bytes: they are tokens >= S, which the 3-D map clips
    tma_store_3d(&tmO, sO, head * ATT_D, blk * 128, frame);
    tma_store_commit();
    tma_store_wait_read<0>();
  }
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<ATT_TMEM_COLS>(tmem_base);
  }
}

}  // namespace vitk
