// attention_sm100.cuh — fused softmax(Q K^T / sqrt(d)) V for short sequences (S <= 256, d = 64)
//
// Replaces HF ViTSelfAttention / SDPA (modeling_vit.py:171-196, 220-252) for the 197-token
// ViT-B/16 shape.  One CTA per (frame, head); the whole K and V of the head live in shared
// memory (S*64*2 B = 25 KB each), so there is no online-softmax tiling:
//   S_blk = Q_blk K^T   tcgen05.mma M=128, N=SK (S rounded up to 16), K=64 -> TMEM (f32)
//   softmax             one thread per query row (TMEM lane): tcgen05.ld, max, exp2, sum
//   P_blk (bf16)        written to shared memory in the K-major 128B-swizzled UMMA layout
//   O_blk = P_blk V     tcgen05.mma M=128, N=64, K=SK, V consumed MN-major as TMA delivered it
//   O / rowsum -> bf16 -> staged in Q's shared memory -> TMA store (3-D map clips tokens >= S)
// Query rows are processed as two M=128 blocks (rows 0..127 by warps 0-3, 128..255 by
// warps 4-7).  S and O accumulators stay in TMEM, P in shared memory; scores never touch HBM.
#pragma once
#include "vitk_common.cuh"

namespace vitk {

constexpr int ATT_THREADS = 256;
constexpr int ATT_D = 64;

struct AttnSmemLayout {
  int sk;         // keys rounded up to 16 (MMA N / K extent)
  int katoms;     // ceil(sk / 64) 64-column swizzle atoms of P
  int rows1;      // rows of M-block 1 that are backed by smem (rounded to 8)
  int off_q0, off_q1, off_k, off_v, off_p1, off_p0, off_bar, total;
};

__host__ __device__ inline AttnSmemLayout attn_smem_layout(int S) {
  AttnSmemLayout L;
  L.sk = (S + 15) & ~15;
  L.katoms = (L.sk + 63) / 64;
  int r1 = S > 128 ? S - 128 : 0;
  L.rows1 = (r1 + 7) & ~7;
  int kv_bytes = ((L.sk * 128) + 1023) & ~1023;
  L.off_q0 = 0;
  L.off_q1 = 16384;
  L.off_k = 32768;
  L.off_v = L.off_k + kv_bytes;
  // P of block 1 sits BEFORE P of block 0: the M=128 MMA on block 1 reads 128 rows per atom
  // although only rows1 are backed; the over-read lands in P0 (valid smem, results discarded).
  L.off_p1 = L.off_v + kv_bytes;
  L.off_p0 = L.off_p1 + L.katoms * L.rows1 * 128;
  L.off_bar = L.off_p0 + L.katoms * 16384;
  L.total = L.off_bar + 128 + 1024;  // + alignment slack
  return L;
}

struct AttnParams {
  int S;        // tokens per frame
  int heads;
  int frames;
  float scale_log2e;  // (1/sqrt(d)) * log2(e)
};

// tmQKV: 3-D map over qkv [frames, S, 3*D] bf16, box {64, 128, 1}   (Q blocks)
// tmKV : same tensor,                          box {64, sk, 1}      (K and V)
// tmO  : 3-D map over out [frames, S, D] bf16, box {64, 128, 1}
__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                 const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const AttnSmemLayout L = attn_smem_layout(p.S);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* bar_qk = bars + 0;
  uint64_t* bar_v = bars + 1;
  uint64_t* bar_s = bars + 2;  // [2]
  uint64_t* bar_o = bars + 4;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 6);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int frame = blockIdx.x / p.heads;
  const int head = blockIdx.x % p.heads;
  const int D = p.heads * ATT_D;
  const int nblk = p.S > 128 ? 2 : 1;
  const int S = p.S, SK = L.sk;

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmO);
    mbar_init(bar_qk, 1);
    mbar_init(bar_v, 1);
    mbar_init(&bar_s[0], 1);
    mbar_init(&bar_s[1], 1);
    mbar_init(&bar_o[0], 1);
    mbar_init(&bar_o[1], 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (tid == 0) {
    // ---- loads: Q (both blocks) + K on one barrier, V on another ----
    const uint32_t q_bytes = 128 * 128 * nblk;
    const uint32_t kv_bytes = SK * 128;
    mbar_arrive_expect_tx(bar_qk, q_bytes + kv_bytes);
    tma_load_3d(smem + L.off_q0, &tmQ, bar_qk, head * ATT_D, 0, frame);
    if (nblk == 2) tma_load_3d(smem + L.off_q1, &tmQ, bar_qk, head * ATT_D, 128, frame);
    tma_load_3d(smem + L.off_k, &tmKV, bar_qk, D + head * ATT_D, 0, frame);
    mbar_arrive_expect_tx(bar_v, kv_bytes);
    tma_load_3d(smem + L.off_v, &tmKV, bar_v, 2 * D + head * ATT_D, 0, frame);

    // ---- S_blk = Q_blk K^T ----
    mbar_wait(bar_qk, 0);
    tc_fence_after_sync();
    const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
    const uint64_t dk = make_smem_desc_sw128(smem_u32(smem + L.off_k), 1024, 0);
    for (int b = 0; b < nblk; ++b) {
      const uint64_t dq =
          make_smem_desc_sw128(smem_u32(smem + (b ? L.off_q1 : L.off_q0)), 1024, 0);
#pragma unroll
      for (int k = 0; k < ATT_D / 16; ++k)
        umma_bf16_ss(tmem_base + b * 256, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
      umma_commit(&bar_s[b]);
    }
  }

  const int blk = warp >> 2;  // which M block this warp's rows belong to
  if (blk < nblk) {
    const int q = warp & 3;
    const int r = q * 32 + lane;       // row within the block == TMEM lane
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + blk * 256;
    uint8_t* sP = smem + (blk ? L.off_p1 : L.off_p0);
    const int atom_bytes = blk ? L.rows1 * 128 : 16384;
    const bool row_backed = blk == 0 || r < L.rows1;

    mbar_wait(&bar_s[blk], 0);
    tc_fence_after_sync();

    // pass 1: row max over the S valid key columns
    float mx = -INFINITY;
    for (int c0 = 0; c0 < SK; c0 += 16) {
      uint32_t v[16];
      tmem_ld_32x32b_x16(t_row + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < S) mx = fmaxf(mx, __uint_as_float(v[j]));
    }
    const float mxs = mx * p.scale_log2e;
    // pass 2: p = exp2(s*scale*log2e - max*scale*log2e); row sum; bf16 P -> swizzled smem
    float sum = 0.f;
    for (int c0 = 0; c0 < SK; c0 += 16) {
      uint32_t v[16];
      tmem_ld_32x32b_x16(t_row + c0, v);
      tmem_ld_wait();
      float e[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float x = exp2f(fmaf(__uint_as_float(v[j]), p.scale_log2e, -mxs));
        x = (c0 + j < S) ? x : 0.f;
        // the row sum is taken over the bf16-rounded probabilities the P.V MMA will consume
        x = __bfloat162float(__float2bfloat16_rn(x));
        e[j] = x;
        sum += x;
      }
      if (row_backed) {
        const int atom = c0 >> 6;
        const int ci = (c0 & 63) >> 3;  // 16-byte chunk index inside the 128-byte row
        uint8_t* rowp = sP + atom * atom_bytes + r * 128;
        uint4 o0, o1;
        o0.x = pack_bf16x2(e[0], e[1]);   o0.y = pack_bf16x2(e[2], e[3]);
        o0.z = pack_bf16x2(e[4], e[5]);   o0.w = pack_bf16x2(e[6], e[7]);
        o1.x = pack_bf16x2(e[8], e[9]);   o1.y = pack_bf16x2(e[10], e[11]);
        o1.z = pack_bf16x2(e[12], e[13]); o1.w = pack_bf16x2(e[14], e[15]);
        *reinterpret_cast<uint4*>(rowp + (((ci + 0) ^ (r & 7)) << 4)) = o0;
        *reinterpret_cast<uint4*>(rowp + (((ci + 1) ^ (r & 7)) << 4)) = o1;
      }
    }
    // P visible to the tensor core (async proxy); S fully consumed by every row of the block
    fence_proxy_async_smem();
    tc_fence_before_sync();
    named_bar_sync(1 + blk, 128);

    if (q == 0 && lane == 0) {
      // ---- O_blk = P_blk V : K = SK in steps of 16, V is MN-major (tokens x 64 dims) ----
      tc_fence_after_sync();
      mbar_wait(bar_v, 0);
      tc_fence_after_sync();
      const uint32_t idesc_o = make_idesc_bf16_f32(128, ATT_D, false, true);
      for (int kk = 0; kk < SK / 16; ++kk) {
        const uint64_t dp = make_smem_desc_sw128(
            smem_u32(sP + (kk >> 2) * atom_bytes) + (kk & 3) * 32, 1024, 0);
        // 16 K rows (tokens) of V = 2 groups of 8 rows x 128 B
        const uint64_t dv =
            make_smem_desc_sw128(smem_u32(smem + L.off_v) + kk * 16 * 128, 1024, 1024);
        umma_bf16_ss(tmem_base + blk * 256, dp, dv, idesc_o, kk != 0 ? 1u : 0u);
      }
      umma_commit(&bar_o[blk]);
    }

    // ---- epilogue: O / rowsum -> bf16 -> Q's smem tile (dead by now) -> TMA store ----
    mbar_wait(&bar_o[blk], 0);
    tc_fence_after_sync();
    const float inv = 1.0f / sum;
    uint8_t* sO = smem + (blk ? L.off_q1 : L.off_q0);
#pragma unroll
    for (int c0 = 0; c0 < ATT_D; c0 += 32) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(t_row + c0, v);
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
    tc_fence_before_sync();
    named_bar_sync(1 + blk, 128);
    if (q == 0 && lane == 0) {
      tma_store_3d(&tmO, sO, head * ATT_D, blk * 128, frame);
      tma_store_commit();
      tma_store_wait_all<0>();
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace vitk
