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
  int sk;      // keys rounded up to 16 (MMA N / K extent)
  int kv_box;  // rows per K / V TMA box: sk, or for sk > 256 (box and MMA-N limit) the first of
               // two halves, a multiple of 16
  int off_q, off_k, off_v, off_bar, total;
};
constexpr int ATT_MAX_S = 448;  // wide form: both halves of the keys within the 256 box / MMA-N limit

__host__ __device__ inline AttnSmemLayout attn_smem_layout(int S) {
  AttnSmemLayout L;
  L.sk = (S + 15) & ~15;
  L.kv_box = L.sk <= 256 ? L.sk : ((L.sk / 2 + 15) & ~15);
  // rows the K / V loads deliver: sk, or both halves of the wide form (2*kv_box >= sk)
  const int kv_rows = L.sk <= 256 ? L.sk : 2 * L.kv_box;
  int kv_bytes = ((kv_rows * 128) + 1023) & ~1023;
  L.off_q = 0;
  L.off_k = 16384;
  L.off_v = L.off_k + kv_bytes;
  L.off_bar = L.off_v + kv_bytes;
  L.total = L.off_bar + 64 + 1024;  // + alignment slack
  if (L.total < ATT_MIN_SMEM) L.total = ATT_MIN_SMEM;
  // wide form: the CTA owns all 512 TMEM columns, so at most ONE CTA may be resident per SM
  if (L.sk > 256 && L.total < 116 * 1024) L.total = 116 * 1024;
  return L;
}

struct AttnParams {
  int S;        // tokens per frame
  int heads;
  int frames;
  int q_rows;   // queries per frame that are needed (== S, or fewer: last-layer CLS shortcut)
  int nblk;     // query blocks per (frame, head): 1 if q_rows <= 128 else 2
  int reverse;  // v3: walk the (frame, head) units from the last to the first ("snake" order)
  int one_issuer; // v3: 1 = both TMEM slots' MMAs issued by one polling thread (round-1 form)
  int early_pv; // v3, S padded to 208 keys: P.V of the keys 96-191 issued in the middle of the softmax pass
  int handover; // v3: 32-column chunk of the softmax pass after which the exp2 turn passes to the
                // other warpgroup (>= number of chunks: at the end of the pass)
  float scale_log2e;  // (1/sqrt(d)) * log2(e)
  long long* dbg;     // developer timeline (tools/dev_attn_timeline.py): null in production
};

// tmQ : 3-D map over qkv [frames, S, 3*D] bf16, box {64, 128, 1}   (Q block)
// tmKV: same tensor,                          box {64, sk, 1}      (K and V)
// tmO : 3-D map over out [frames, S, D] bf16, box {64, 128, 1}
// WIDE = false: S <= 256, 256 TMEM columns, two CTAs per SM.  WIDE = true: 256 < S <= 448 (the
// reference's TF-path geometry: 432x768 frames, patch 32 -> 313 tokens, generate_clips_hmm.py:48-56):
// 512 TMEM columns, one CTA per SM, K / V loaded as two TMA boxes and S = Q K^T issued as two
// N-halves (box and MMA-N extents stop at 256).
template <bool WIDE>
__global__ void __launch_bounds__(ATT_THREADS, WIDE ? 1 : 2)
attention_v2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                 const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  // aligned by POINTER arithmetic on the __shared__ array: an integer round trip makes every access
  // through the result a generic LD / ST (with R2UR descriptor traffic) instead of LDS / STS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const AttnSmemLayout L = attn_smem_layout(p.S);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* bar_qk = bars + 0;
  uint64_t* bar_v = bars + 1;
  uint64_t* bar_s = bars + 2;
  uint64_t* bar_o = bars + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx(), lane = tid & 31;
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
    const uint32_t kv_bytes = (WIDE ? 2 * L.kv_box : SK) * 128;
    mbar_arrive_expect_tx(bar_qk, 128 * 128 + kv_bytes);
    tma_load_3d(smem + L.off_q, &tmQ, bar_qk, head * ATT_D, blk * 128, frame);
    tma_load_3d(smem + L.off_k, &tmKV, bar_qk, D + head * ATT_D, 0, frame);
    if constexpr (WIDE)
      tma_load_3d(smem + L.off_k + L.kv_box * 128, &tmKV, bar_qk, D + head * ATT_D, L.kv_box, frame);
    mbar_arrive_expect_tx(bar_v, kv_bytes);
    tma_load_3d(smem + L.off_v, &tmKV, bar_v, 2 * D + head * ATT_D, 0, frame);
    if constexpr (WIDE)
      tma_load_3d(smem + L.off_v + L.kv_box * 128, &tmKV, bar_v, 2 * D + head * ATT_D, L.kv_box, frame);
  }
  if (warp == 1) tmem_alloc<WIDE ? 512u : ATT_TMEM_COLS>(tmem_slot);
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
    const uint64_t dq = make_smem_desc_sw128(smem_u32(smem + L.off_q), 1024, 0);
    const int n1 = WIDE ? L.kv_box : SK;  // keys [0, n1) and, wide only, [n1, SK)
    const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)n1, false, false);
    const uint64_t dk = make_smem_desc_sw128(smem_u32(smem + L.off_k), 1024, 0);
#pragma unroll
    for (int k = 0; k < ATT_D / 16; ++k)
      umma_bf16_ss(tmem_base, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
    if constexpr (WIDE) {
      const uint32_t idesc_s2 = make_idesc_bf16_f32(128, (uint32_t)(SK - n1), false, false);
      const uint64_t dk2 = make_smem_desc_sw128(smem_u32(smem + L.off_k + n1 * 128), 1024, 0);
#pragma unroll
      for (int k = 0; k < ATT_D / 16; ++k)
        umma_bf16_ss(tmem_base + (uint32_t)n1, dq + 2 * k, dk2 + 2 * k, idesc_s2, k != 0 ? 1u : 0u);
    }
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
    tmem_dealloc<WIDE ? 512u : ATT_TMEM_COLS>(tmem_base);
  }
}

// =============================================================================================
// v3: persistent, warp-specialised.  One CTA per SM loops over (frame, head) units; both query
// blocks of a unit share ONE K/V load.  Roles (384 threads):
//   warp 0        TMA loader: Q0 | Q1 | K on kv_full[stage], V on v_full[stage]; 2 stages of 84 KB
//   warp 1        MMA issuer: two independent per-TMEM-slot state machines, advanced by polling
//                 (S = Q K^T when the slot is free, O = P V when that slot's P is ready)
//   warp 2        TMEM allocator (2 slots x 256 columns: S f32 | P bf16 aliasing S | O f32)
//   warps 4-7     softmax + epilogue of slot 0      warps 8-11  softmax + epilogue of slot 1
// Query block b of unit j uses slot b ^ (j & 1), so the 128-row and the 69-row block alternate
// between the two softmax warpgroups; while one group is in its exp2 loop the other group's
// MMAs, loads and stores proceed.  v2 ran 2 independent CTAs/SM with every phase serialised
// inside a CTA and K/V loaded twice (127 us/layer at 256 frames).
// =============================================================================================
// ---- softmax building blocks of the v3 fast path (one thread == one query row) ----
// The compiler must not move arithmetic on registers filled by an in-flight tcgen05.ld above the
// tcgen05.wait::ld that completes it: thread the registers through an empty asm after the wait.
__device__ __forceinline__ void reg_fence16(uint32_t* r) {
  asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]),
                    "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]),
                    "+r"(r[13]), "+r"(r[14]), "+r"(r[15]));
}
// exp2 on the FMA / ALU pipes (Cody-Waite split + degree-3 minimax of 2^f on [-0.5, 0.5], max relative
// error 7.5e-5 -- 1/50 of the bf16 rounding of P): x = n + f, 2^x = poly(f) with n added to the exponent
// field.  Inputs are clamped to [-126, 127.9]: an overflowing row still reaches > 1e30 and takes the
// exact redo.  The softmax pass is MUFU-bound (one ex2 per score, 8 cycles per warp instruction); the
// pairs selected by VITK_ATT_POLY_MASK (bit i%16 of the pair index inside a chunk) take this route
// instead, so that the MUFU and the FMA pipe share the exponentials (cf. FlashAttention-4's software exp2).
// 1: clock64 stamps for tools/dev_attn*_timeline.py are compiled in (developer builds: VITK_NVCC_EXTRA=-DVITK_ATT_TIMELINE=1)
#ifndef VITK_ATT_TIMELINE
#define VITK_ATT_TIMELINE 0
#endif
// Default 0x8420: pairs 5, 10 and 15 of every 16 (19 % of the scores).  Measured on the final kernel
// (147 registers): none 70.2 us, 0x8000 69.3, 0x8080 67.5, 0x8420 67.1, 0x8888 68.6, 0x4924 70.8, 0xaaaa 77.0
// (profiles/r02_attn_handover.txt); on the round-1 / mid-round kernels, which sat at the register limit,
// every mask was slower.  -DVITK_ATT_POLY_MASK=0u restores the all-MUFU form.
#ifndef VITK_ATT_POLY_MASK
#define VITK_ATT_POLY_MASK 0x8420u
#endif
__device__ __forceinline__ float2 exp2_poly2(float2 x) {
  const float2 xc = make_float2(fminf(fmaxf(x.x, -126.0f), 127.9f), fminf(fmaxf(x.y, -126.0f), 127.9f));
  const float2 magic = make_float2(12582912.0f, 12582912.0f);   // 1.5 * 2^23
  const float2 t = __fadd2_rn(xc, magic);
  const float2 n = __fadd2_rn(t, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = __ffma2_rn(n, make_float2(-1.0f, -1.0f), xc);
  float2 q = make_float2(5.517124757e-02f, 5.517124757e-02f);
  q = __ffma2_rn(q, f, make_float2(2.426100969e-01f, 2.426100969e-01f));
  q = __ffma2_rn(q, f, make_float2(6.932609677e-01f, 6.932609677e-01f));
  q = __ffma2_rn(q, f, make_float2(9.999281168e-01f, 9.999281168e-01f));
  return make_float2(__int_as_float(__float_as_int(q.x) + (__float_as_int(t.x) << 23)),
                     __int_as_float(__float_as_int(q.y) + (__float_as_int(t.y) << 23)));
}
__device__ __forceinline__ float2 att_exp2_pair(float2 x, int i) {
  if ((VITK_ATT_POLY_MASK >> (i & 15)) & 1u) return exp2_poly2(x);
  return make_float2(ex2_approx(x.x), ex2_approx(x.y));
}

// exp2(s*scale - mxs) of NC = 32 or 16 scores starting at column c0 -> bf16 pairs stored over S
// cells [c0/2, c0/2 + NC/2) (already consumed); accumulates the row sum.  Packed fp32 (FFMA2 /
// FADD2): 2.5 issue slots per element next to its one MUFU op — a single softmax warp per
// scheduler is issue-limited, not MUFU-limited.  No exponent clamp: a row whose maximum lies far
// above the shift overflows to +inf, the row sum becomes inf and the block is redone exactly.
// FULL = 1: the caller knows that every column of the chunk is a valid key (no mask, no branch).
// p_col >= 0 overrides the destination column of the packed probabilities.
template <int NC, int FULL = 0>
__device__ __forceinline__ void att_exp_chunk(const uint32_t* v, int c0, int S, float scale, float mxs,
                                              float2 (&s2)[2], uint32_t t_slot, int p_col = -1) {
  uint32_t pk[NC / 2];
  const bool full = FULL == 1 ? true : (c0 + NC <= S);
  const float2 sc2 = make_float2(scale, scale), nm2 = make_float2(-mxs, -mxs);
#pragma unroll
  for (int i = 0; i < NC / 2; ++i) {
    const float2 x = __ffma2_rn(make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), sc2, nm2);
    float2 e = att_exp2_pair(x, i);
    if (!full) {
      e.x = (c0 + 2 * i < S) ? e.x : 0.f;
      e.y = (c0 + 2 * i + 1 < S) ? e.y : 0.f;
    }
    s2[i & 1] = __fadd2_rn(s2[i & 1], e);
    pk[i] = pack_bf16x2(e.x, e.y);
  }
  const uint32_t dst = t_slot + static_cast<uint32_t>(p_col >= 0 ? p_col : (c0 >> 1));
  if constexpr (NC == 32) tmem_st_32x32b_x16(dst, *reinterpret_cast<const uint32_t(*)[16]>(pk));
  else tmem_st_32x32b_x8(dst, *reinterpret_cast<const uint32_t(*)[8]>(pk));
}

template <int V> struct AttConst { static constexpr int value = V; };
constexpr int ATT3_THREADS = 384;
constexpr int ATT3_STAGES = 2;

struct Attn3SmemLayout {
  int sk, kv_bytes, stage_bytes, off_o, off_bar, total;
};
__host__ __device__ inline Attn3SmemLayout attn3_smem_layout(int S) {
  Attn3SmemLayout L;
  L.sk = (S + 15) & ~15;
  L.kv_bytes = ((L.sk * 128) + 1023) & ~1023;
  L.stage_bytes = 2 * 16384 + 2 * L.kv_bytes;  // Q0 | Q1 | K | V
  L.off_o = ATT3_STAGES * L.stage_bytes;  // 2 x 16 KB output staging, one per softmax warpgroup
  L.off_bar = L.off_o + 2 * 16384;
  L.total = L.off_bar + 256 + 1024;
  return L;
}

// EARLY = true: the instantiation for 193-208 keys (the ViT case) whose softmax pass takes the chunks
// in the order 3, 4, 5 | 0, 1, 2, tail and lets the MMA warp start O = P V in the middle of it; it
// contains ONLY that pass (both passes in one kernel push ptxas over the register budget and the
// spills cost more than the early start wins).  EARLY = false: any S <= 256.
template <bool EARLY>
__global__ void __launch_bounds__(ATT3_THREADS, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                 const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  // aligned by POINTER arithmetic on the __shared__ array: an integer round trip makes every access
  // through the result a generic LD / ST (with R2UR descriptor traffic) instead of LDS / STS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const Attn3SmemLayout L = attn3_smem_layout(p.S);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* kv_full = bars + 0;      // [2] stage
  uint64_t* v_full = bars + 2;       // [2] stage
  uint64_t* stage_empty = bars + 4;  // [2] stage, count 2 (one per softmax warpgroup)
  uint64_t* s_full = bars + 6;       // [2] slot
  uint64_t* p_full = bars + 8;       // [2] slot, count 4 (one per warp)
  uint64_t* o_full = bars + 10;      // [2] slot
  uint64_t* slot_free = bars + 12;   // [2] slot, count 4
  uint64_t* redo_bar = bars + 14;    // [2] slot: exact-path recomputation of S
  uint64_t* turn_bar = bars + 16;    // [2] group: "your turn on the exp2 units", count 4
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);
  uint64_t* p_half = bars + 22;      // [2] slot, count 4: P of keys 96-191 written, S columns 96-191 read (early P.V)
  uint64_t* epv_done = bars + 24;    // [2] slot: the early P.V instructions have been issued
  volatile int* redo_flag = reinterpret_cast<volatile int*>(bars + 26);  // [2] slot: this unit took the exact path
  uint64_t* qk_empty = bars + 20;    // [2] stage, count 2: Q / K of the stage are dead (both blocks' passes done)

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx(), lane = tid & 31;
  const int D = p.heads * ATT_D;
  const int S = p.S, SK = L.sk;
  const int units = p.frames * p.heads;
  const int nblk = p.nblk;

  griddep_launch_dependents();
  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmO);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&stage_empty[i], 2);
      mbar_init(&qk_empty[i], 2);
      mbar_init(&p_half[i], 4);
      mbar_init(&epv_done[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1);
      mbar_init(&slot_free[i], 4);
      mbar_init(&redo_bar[i], 1);
      mbar_init(&turn_bar[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();  // the QKV GEMM's output is read (and `out` written) only from here on

  if (warp == 0) {
    // ------------------------------ loader ------------------------------
    if (lane == 0) {
      int j = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const int uu = p.reverse ? units - 1 - u : u;
        const int frame = uu / p.heads, head = uu % p.heads;
        uint8_t* sb = smem + st * L.stage_bytes;
        // Q and K of a stage die when both blocks' softmax passes are over (the exact-path redo, the
        // last reader of Q / K, is decided there), V only when both P.V chains have retired ~800
        // cycles later: the next-but-one unit's Q / K load -- 58 KB that take 2 200-3 300 cycles to
        // land and sit on the critical path of the next S = Q K^T -- starts that much earlier
        mbar_wait(&qk_empty[st], ph ^ 1u);
        if ((VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && j < 16) p.dbg[(2 * 16 + j) * 8 + 0] = clock64();
        mbar_arrive_expect_tx(&kv_full[st], 16384 * nblk + SK * 128);
        tma_load_3d(sb, &tmQ, &kv_full[st], head * ATT_D, 0, frame);
        if (nblk == 2) tma_load_3d(sb + 16384, &tmQ, &kv_full[st], head * ATT_D, 128, frame);
        tma_load_3d(sb + 32768, &tmKV, &kv_full[st], D + head * ATT_D, 0, frame);
        mbar_wait(&stage_empty[st], ph ^ 1u);
        mbar_arrive_expect_tx(&v_full[st], SK * 128);
        tma_load_3d(sb + 32768 + L.kv_bytes, &tmKV, &v_full[st], 2 * D + head * ATT_D, 0, frame);
      }
    }
  } else if (warp == 1 || warp == 3) {
    // ------------------------------ MMA issuers ------------------------------
    // One issuing thread PER TMEM SLOT (warp 1: slot 0, warp 3: slot 1), each walking its own
    // chain  S = Q K^T -> (softmax) -> O = P V  with blocking waits.  Round 1 ran both chains as
    // polled state machines in ONE thread; the timeline (tools/dev_attn_timeline.py) showed the
    // other slot's 13 TS-form P.V instructions (~1 075 cycles of issue) sitting between a slot
    // becoming free and its next S being issued (wait_S ~800 cycles, P.V issued ~300 cycles late).
    // p.one_issuer != 0 keeps the single-thread form for A/B runs.
    const bool both = p.one_issuer != 0;
    if (!both || (lane == 0 && warp == 1)) {
      const int my_units = (units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
      const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
      const uint32_t idesc_o = make_idesc_bf16_f32(128, ATT_D, false, true);
      if (!both) {
        // the whole warp walks the chain (warp-uniform control flow) and ONE elected lane issues:
        // the tcgen05.mma operands then sit in uniform registers, one instruction each, where
        // `if (lane == 0)` code wraps every one of them in an ELECT / R2UR loop (see gemm_sm100.cuh)
        const int s = warp == 1 ? 0 : 1;
        const uint32_t t_slot = tmem_base + s * 256;
        for (int j = 0; j < my_units; ++j) {
          const int st = j & 1;
          const uint32_t sph = (j >> 1) & 1, uph = j & 1;
          const int b = s ^ (j & 1);  // query block this slot holds in unit j
          uint8_t* sb = smem + st * L.stage_bytes;
          const bool ldbg = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && j < 16 && lane == 0;
          mbar_wait(&kv_full[st], sph);
          if (ldbg) p.dbg[(2 * 16 + j) * 8 + 1 + s] = clock64();
          mbar_wait(&slot_free[s], uph ^ 1u);
          if (ldbg) p.dbg[(2 * 16 + j) * 8 + 3 + s] = clock64();
          tc_fence_after_sync();
          if (elect_one_sync()) {
            if (b < nblk) {
              const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
              const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
              for (int k = 0; k < ATT_D / 16; ++k)
                umma_bf16_ss(t_slot, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
            }
            umma_commit(&s_full[s]);
          }
          __syncwarp();
          constexpr bool early = EARLY;
          if (early) {
            // keys 96-191 (k-steps 6-11): their probabilities sit in the slot's spare columns
            // 208-255 and the S columns under the O accumulator (104-167) have been read by all
            // four warps, so these six instructions run under the rest of the softmax pass
            mbar_wait(&p_half[s], uph);
            mbar_wait(&v_full[st], sph);
            tc_fence_after_sync();
            if (elect_one_sync()) {
              if (b < nblk) {
                const uint64_t dv0 = make_smem_desc_sw128(smem_u32(sb + 32768 + L.kv_bytes), 1024, 1024);
#pragma unroll
                for (int kk = 6; kk < 12; ++kk)
                  umma_bf16_ts(t_slot + 104, t_slot + 208 + (kk - 6) * 8, dv0 + (uint64_t)(kk * 128), idesc_o,
                               kk != 6 ? 1u : 0u);
              }
              mbar_arrive(&epv_done[s]);
            }
            __syncwarp();
          }
          mbar_wait(&p_full[s], uph);
          mbar_wait(&v_full[st], sph);
          tc_fence_after_sync();
          const bool mdbg = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && j < 16 && lane == 0;
          if (mdbg) p.dbg[(s * 16 + j) * 8 + 6] = clock64();
          if (early && redo_flag[s] == 0) {
            if (elect_one_sync()) {
              if (b < nblk) {
                const uint64_t dv0 = make_smem_desc_sw128(smem_u32(sb + 32768 + L.kv_bytes), 1024, 1024);
#pragma unroll
                for (int kk = 0; kk < 6; ++kk)
                  umma_bf16_ts(t_slot + 104, t_slot + kk * 8, dv0 + (uint64_t)(kk * 128), idesc_o, 1u);
                umma_bf16_ts(t_slot + 104, t_slot + 96, dv0 + (uint64_t)(12 * 128), idesc_o, 1u);
              }
              umma_commit(&o_full[s]);
            }
          } else
          if (elect_one_sync()) {
            if (b < nblk) {
              const uint32_t sv = smem_u32(sb + 32768 + L.kv_bytes);
              const uint64_t dv0 = make_smem_desc_sw128(sv, 1024, 1024);
              const uint32_t t_o_acc = t_slot + SK / 2;
              umma_bf16_ts(t_o_acc, t_slot, dv0, idesc_o, 0u);
#pragma unroll 4
              for (int kk = 1; kk < SK / 16; ++kk)
                umma_bf16_ts(t_o_acc, t_slot + kk * 8, dv0 + (uint64_t)(kk * 128), idesc_o, 1u);
            }
            umma_commit(&o_full[s]);
          }
          __syncwarp();
          if (mdbg) p.dbg[(s * 16 + j) * 8 + 7] = clock64();
        }
      } else {
      int uj[2] = {0, 0};   // unit each slot's state machine is working on
      int phs[2] = {0, 0};  // 0: S = Q K^T is next, 1: O = P V is next
      long long t0 = clock64();
      while (uj[0] < my_units || uj[1] < my_units) {
        bool progressed = false;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const int j = uj[s];
          if (j >= my_units) continue;
          const int st = j & 1;
          const uint32_t sph = (j >> 1) & 1, uph = j & 1;
          const int b = s ^ (j & 1);  // query block this slot holds in unit j
          uint8_t* sb = smem + st * L.stage_bytes;
          const uint32_t t_slot = tmem_base + s * 256;
          if (phs[s] == 0) {
            if (!mbar_test_wait(&kv_full[st], sph) || !mbar_test_wait(&slot_free[s], uph ^ 1u)) continue;
            tc_fence_after_sync();
            if (b < nblk) {
              const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
              const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
              for (int k = 0; k < ATT_D / 16; ++k)
                umma_bf16_ss(t_slot, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
            }
            umma_commit(&s_full[s]);
            phs[s] = 1;
          } else {
            if (!mbar_test_wait(&p_full[s], uph) || !mbar_test_wait(&v_full[st], sph)) continue;
            tc_fence_after_sync();
            const bool mdbg = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && j < 16;
            if (mdbg) p.dbg[(s * 16 + j) * 8 + 6] = clock64();
            if (b < nblk) {
              const uint32_t sv = smem_u32(sb + 32768 + L.kv_bytes);
              // descriptor of K-step kk = first one + kk * (16 rows * 128 B >> 4) in the address field
              const uint64_t dv0 = make_smem_desc_sw128(sv, 1024, 1024);
              const uint32_t t_o_acc = t_slot + SK / 2;
              umma_bf16_ts(t_o_acc, t_slot, dv0, idesc_o, 0u);
#pragma unroll 4
              for (int kk = 1; kk < SK / 16; ++kk)
                umma_bf16_ts(t_o_acc, t_slot + kk * 8, dv0 + (uint64_t)(kk * 128), idesc_o, 1u);
            }
            umma_commit(&o_full[s]);
            if (mdbg) p.dbg[(s * 16 + j) * 8 + 7] = clock64();
            phs[s] = 0;
            uj[s] = j + 1;
          }
          progressed = true;
        }
        if (progressed) t0 = clock64();
        else if (clock64() - t0 > 4000000000LL) {
          printf("vitk: attention MMA poll timeout block %d\n", (int)blockIdx.x);
          __trap();
        }
      }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ softmax + epilogue ------------------------------
    const int g = (warp - 4) >> 2;  // warpgroup == TMEM slot
    const int wq = warp & 3;        // TMEM lane quarter
    const int r = wq * 32 + lane;   // row within the query block == TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t t_slot = tmem_base + g * 256 + lane_off;
    const uint32_t t_o = t_slot + static_cast<uint32_t>(SK / 2);
    int j = 0;
    uint32_t redo_phase = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
      const int st = j & 1;
      const uint32_t uph = j & 1;
      const int b = g ^ (j & 1);
      const int uu = p.reverse ? units - 1 - u : u;
        const int frame = uu / p.heads, head = uu % p.heads;
      const bool blk_active = b < nblk;
      const int rows_valid = blk_active ? min(128, p.q_rows - b * 128) : 0;
      const bool warp_active = wq * 32 < rows_valid;
      uint8_t* sO = smem + L.off_o + g * 16384;  // this warpgroup's private output staging

      const bool dbg_on = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && wq == 0 && lane == 0 && j < 16;
      long long* dbg = p.dbg + (g * 16 + j) * 8;
      if (dbg_on) dbg[0] = clock64();
      mbar_wait(&s_full[g], uph);
      if (dbg_on) p.dbg[(2 * 16 + j) * 8 + 5 + g] = clock64();
      // The two warpgroups take turns in the softmax: g0(unit j) -> g1(unit j) -> g0(unit j+1) ...
      // Left alone they run in lock-step — both in the exp2 loop at once (the MUFU pipe is the
      // loop's bound, so the pass takes twice as long), then both waiting while the single MMA
      // thread issues their two P.V chains back to back (~1100 cycles each: 13 instructions at
      // ~85 cycles).  Alternating puts one group's MMAs / epilogue / next S under the other
      // group's exp2 pass (tools/dev_attn_timeline.py: 8300 -> ~5500 cycles per unit).
      if (g == 0) { if (j > 0) mbar_wait(&turn_bar[0], (uint32_t)(j - 1) & 1u); }
      else mbar_wait(&turn_bar[1], uph);
      tc_fence_after_sync();
      if (dbg_on) dbg[1] = clock64();
      float sum = 0.f;
      bool handed = false;  // this warp has passed the exp2 turn on for this unit
      // exp2 of one 32- or 16-column chunk against the shift mxs -> bf16 pairs over the consumed
      // S cells; accumulates the row sum.  kClamp bounds the exponent (fast path only).
      auto exp_pass = [&](float mxs, bool clamp) {
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
        int c0 = 0;
        for (; c0 + 32 <= SK; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_slot + c0, v);
          tmem_ld_wait();
          uint32_t pk[16];
          const bool full = c0 + 32 <= S;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x0 = fmaf(__uint_as_float(v[2 * i]), p.scale_log2e, -mxs);
            float x1 = fmaf(__uint_as_float(v[2 * i + 1]), p.scale_log2e, -mxs);
            if (clamp) { x0 = fminf(x0, 100.f); x1 = fminf(x1, 100.f); }
            float e0 = ex2_approx(x0), e1 = ex2_approx(x1);
            if (!full) {
              e0 = (c0 + 2 * i < S) ? e0 : 0.f;
              e1 = (c0 + 2 * i + 1 < S) ? e1 : 0.f;
            }
            s4[i & 3] += e0 + e1;
            pk[i] = pack_bf16x2(e0, e1);
          }
          tmem_st_32x32b_x16(t_slot + (c0 >> 1), pk);
        }
        if (c0 < SK) {
          uint32_t v[16];
          tmem_ld_32x32b_x16(t_slot + c0, v);
          tmem_ld_wait();
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float x0 = fmaf(__uint_as_float(v[2 * i]), p.scale_log2e, -mxs);
            float x1 = fmaf(__uint_as_float(v[2 * i + 1]), p.scale_log2e, -mxs);
            if (clamp) { x0 = fminf(x0, 100.f); x1 = fminf(x1, 100.f); }
            float e0 = ex2_approx(x0), e1 = ex2_approx(x1);
            e0 = (c0 + 2 * i < S) ? e0 : 0.f;
            e1 = (c0 + 2 * i + 1 < S) ? e1 : 0.f;
            s4[i & 3] += e0 + e1;
            pk[i] = pack_bf16x2(e0, e1);
          }
          tmem_st_32x32b_x8(t_slot + (c0 >> 1), pk);
        }
        tmem_st_wait();
        return (s4[0] + s4[1]) + (s4[2] + s4[3]);
      };
      // row max over the first `upto` columns (<= S valid ones), 4 independent chains
      auto max_pass = [&](int upto) {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        int c0 = 0;
        for (; c0 + 32 <= SK && c0 < upto; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_slot + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c0 + i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
        }
        if (c0 < SK && c0 < upto) {
          uint32_t v[16];
          tmem_ld_32x32b_x16(t_slot + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (c0 + i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
        }
        return fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      };
      // FAST PATH — one pass.  Softmax is shift invariant and bf16 / fp32 keep their relative
      // precision at any magnitude, so the shift need not be the exact row maximum: any m0 with
      // (max - m0)*scale inside the fp32 exponent range yields the same normalised
      // probabilities.  m0 = max of the row's first <= 32 scores.  A row whose true maximum lies
      // ~100 binades or more above m0 (never observed on real activations) overflows towards
      // +inf and shows up as a huge / non-finite sum ...
      bool redo = false;
      constexpr bool early = EARLY;
      if constexpr (EARLY) {
        // 197 keys padded to 208, the ViT case.  The 32-key chunks are taken in the order
        // 3, 4, 5 | 0, 1, 2, tail: chunks 3-5 are the S columns that lie under the O accumulator
        // (104-167), and their probabilities go to the slot's spare columns 208-255 instead of
        // over S.  After them the MMA warp starts O = P V with those six k-steps while this warp
        // is still in the exp2 loop of the other chunks.
        if (warp_active) {
          uint32_t va[32], vb[32];
          auto ld32 = [&](uint32_t (&dst)[32], int col) { tmem_ld_32x32b_x32(t_slot + col, dst); };
          auto landed = [&](uint32_t (&dst)[32]) { tmem_ld_wait(); reg_fence16(dst); reg_fence16(dst + 16); };
          ld32(va, 96);
          landed(va);
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
          for (int i = 0; i < 32; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(va[i]));
          const float mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2e;
          float2 s4[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
          // the exp2 turn passes on after the fifth chunk (p.handover <= 4) or at the end of the pass
          auto maybe_hand_over = [&]() {};
          auto hand_over_now = [&]() {
            if (p.handover <= 4 && !handed) {
              handed = true;
              __syncwarp();
              if (lane == 0) mbar_arrive(&turn_bar[g ^ 1]);
            }
          };
          ld32(vb, 128);
          att_exp_chunk<32, 1>(va, 96, S, p.scale_log2e, mxs, s4, t_slot, 208);
          maybe_hand_over();
          landed(vb);
          ld32(va, 160);
          att_exp_chunk<32, 1>(vb, 128, S, p.scale_log2e, mxs, s4, t_slot, 224);
          maybe_hand_over();
          landed(va);
          ld32(vb, 0);
          att_exp_chunk<32, 1>(va, 160, S, p.scale_log2e, mxs, s4, t_slot, 240);
          maybe_hand_over();
          // columns 96-191 of S are consumed and their P is on its way to TMEM
          tmem_st_wait();
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_half[g]);
          landed(vb);
          ld32(va, 32);
          att_exp_chunk<32, 1>(vb, 0, S, p.scale_log2e, mxs, s4, t_slot);
          maybe_hand_over();
          landed(va);
          ld32(vb, 64);
          att_exp_chunk<32, 1>(va, 32, S, p.scale_log2e, mxs, s4, t_slot);
          hand_over_now();
          landed(vb);
          tmem_ld_32x32b_x16(t_slot + 192, *reinterpret_cast<uint32_t(*)[16]>(va));
          att_exp_chunk<32, 1>(vb, 64, S, p.scale_log2e, mxs, s4, t_slot);
          maybe_hand_over();
          tmem_ld_wait();
          reg_fence16(va);
          att_exp_chunk<16>(va, 192, S, p.scale_log2e, mxs, s4, t_slot);
          tmem_st_wait();
          sum = (s4[0].x + s4[0].y) + (s4[1].x + s4[1].y);
          redo = !(sum < 1.0e30f);
        } else {
          tc_fence_before_sync();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_half[g]);
        }
      } else
      if (!EARLY && warp_active) {
        // Software-pipelined: the tcgen05.ld of chunk c+1 is in flight while chunk c goes
        // through the exp2 loop (two register buffers), and the first chunk serves both the
        // shift estimate and the first exp2 pass — with two softmax warps per scheduler the
        // TMEM load latency was otherwise exposed once per chunk.
        uint32_t va[32], vb[32];
        const int n32 = SK >> 5;
        const bool tail = (SK & 16) != 0;
        if (n32 > 0) tmem_ld_32x32b_x32(t_slot, va);
        else tmem_ld_32x32b_x16(t_slot, *reinterpret_cast<uint32_t(*)[16]>(va));
        tmem_ld_wait();
        reg_fence16(va);
        if (n32 > 0) reg_fence16(va + 16);
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        const int first = n32 > 0 ? 32 : 16;
        if (SK == 208 && S > 192) {  // the ViT case: all 32 scores of the first chunk are valid keys
#pragma unroll
          for (int i = 0; i < 32; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(va[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (i < first && i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(va[i]));
        }
        const float mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2e;
        float2 s4[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        int c = 0;
        // two 32-column chunks: va holds chunk c_, vb receives c_ + 1, va is refilled with c_ + 2.
        // kFull = 1: every column of both chunks is a valid key -- no mask and no branch in the exp2 code
        auto pair = [&](int c_, auto full_tag) {
          constexpr int kFull = decltype(full_tag)::value;
          tmem_ld_32x32b_x32(t_slot + (c_ + 1) * 32, vb);
          att_exp_chunk<32, kFull>(va, c_ * 32, S, p.scale_log2e, mxs, s4, t_slot);
          if (c_ >= p.handover && !handed) {  // early hand-over: the other group's TMEM loads can run
            handed = true;                    // under the rest of this pass's exp2 work
            __syncwarp();
            if (lane == 0) mbar_arrive(&turn_bar[g ^ 1]);
          }
          tmem_ld_wait();
          reg_fence16(vb);
          reg_fence16(vb + 16);
          if (c_ + 2 < n32) tmem_ld_32x32b_x32(t_slot + (c_ + 2) * 32, va);
          else if (tail) tmem_ld_32x32b_x16(t_slot + (c_ + 2) * 32, *reinterpret_cast<uint32_t(*)[16]>(va));
          att_exp_chunk<32, kFull>(vb, (c_ + 1) * 32, S, p.scale_log2e, mxs, s4, t_slot);
          tmem_ld_wait();
          reg_fence16(va);
          reg_fence16(va + 16);
        };
        if (SK == 208 && S > 192) {  // the ViT case, unrolled: chunks 0-5 hold valid keys only
          pair(0, AttConst<1>{});
          pair(2, AttConst<1>{});
          pair(4, AttConst<1>{});
          c = 6;
        } else {
          for (; c + 2 <= n32; c += 2) pair(c, AttConst<0>{});
        }
        if (c < n32) {  // odd number of 32-column chunks: va holds the last one
          if (tail) tmem_ld_32x32b_x16(t_slot + (c + 1) * 32, *reinterpret_cast<uint32_t(*)[16]>(vb));
          att_exp_chunk<32>(va, c * 32, S, p.scale_log2e, mxs, s4, t_slot);
          tmem_ld_wait();
          reg_fence16(vb);
          if (tail) att_exp_chunk<16>(vb, (c + 1) * 32, S, p.scale_log2e, mxs, s4, t_slot);
        } else if (tail) {  // va holds the 16-column tail
          att_exp_chunk<16>(va, n32 * 32, S, p.scale_log2e, mxs, s4, t_slot);
        }
        tmem_st_wait();
        sum = (s4[0].x + s4[0].y) + (s4[1].x + s4[1].y);
        redo = !(sum < 1.0e30f);
      }
      if (dbg_on) dbg[2] = clock64();
      // ... in which case the whole query block takes the EXACT PATH: S = Q K^T is recomputed
      // into the slot (P has overwritten it) and the classic max pass + exp pass run.
      tc_fence_before_sync();
      const bool took_redo = named_bar_or(2 + g, 128, redo);
      if (early && wq == 0 && lane == 0) redo_flag[g] = took_redo ? 1 : 0;  // read by the MMA warp after p_full
      if (took_redo) {
        if (wq == 0 && lane == 0 && blk_active) {
          // the early P.V instructions write O over S columns 104-167: they must be in the tensor
          // pipe BEFORE S is recomputed (they then run first and the recomputed S overwrites them)
          if (early) mbar_wait(&epv_done[g], uph);
          tc_fence_after_sync();
          uint8_t* sb = smem + st * L.stage_bytes;
          const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
          const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
          const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
          for (int k = 0; k < ATT_D / 16; ++k)
            umma_bf16_ss(tmem_base + g * 256, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
          umma_commit(&redo_bar[g]);
        }
        if (blk_active) {
          mbar_wait(&redo_bar[g], redo_phase);
          redo_phase ^= 1u;
          tc_fence_after_sync();
        }
        if (warp_active) {
          const float mxs = max_pass(SK) * p.scale_log2e;
          sum = exp_pass(mxs, false);
        }
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&p_full[g]);
        if (!handed) mbar_arrive(&turn_bar[g ^ 1]);  // hand the exp2 units to the other warpgroup
        if (wq == 0) mbar_arrive(&qk_empty[st]);     // past the redo decision: this block is done with Q / K
      }
      if (dbg_on) dbg[3] = clock64();

      mbar_wait(&o_full[g], uph);
      tc_fence_after_sync();
      if (dbg_on) dbg[4] = clock64();
      // O = P V has retired: this warpgroup no longer needs the stage's Q / K / V
      if (wq == 0 && lane == 0) {
        mbar_arrive(&stage_empty[st]);
        tma_store_wait_read<0>();  // previous unit's store has long drained the staging tile
      }
      named_bar_sync(2 + g, 128);
      if (warp_active) {
        const float inv = 1.0f / sum;
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(t_o, v0);
        tmem_ld_32x32b_x32(t_o + 32, v1);
        tmem_ld_wait();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[g]);  // O is in registers: the slot can be refilled
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t* src = i < 4 ? v0 + 8 * i : v1 + 8 * (i - 4);
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(src[0]) * inv, __uint_as_float(src[1]) * inv);
          o.y = pack_bf16x2(__uint_as_float(src[2]) * inv, __uint_as_float(src[3]) * inv);
          o.z = pack_bf16x2(__uint_as_float(src[4]) * inv, __uint_as_float(src[5]) * inv);
          o.w = pack_bf16x2(__uint_as_float(src[6]) * inv, __uint_as_float(src[7]) * inv);
          *reinterpret_cast<uint4*>(sO + r * 128 + ((i ^ (r & 7)) << 4)) = o;
        }
        fence_proxy_async_smem();
      } else {
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[g]);
      }
      named_bar_sync(2 + g, 128);
      if (wq == 0 && lane == 0) {
        if (blk_active) {
          // rows of idle warps hold stale bytes: tokens >= S are clipped by the 3-D map, and
          // with q_rows < S (last-layer CLS shortcut) the extra rows are never read
          tma_store_3d(&tmO, sO, head * ATT_D, b * 128, frame);
          tma_store_commit();
        }
      }
      if (dbg_on) dbg[5] = clock64();
    }
  }

  if (warp >= 4 && (warp & 3) == 0 && lane == 0) tma_store_wait_read<0>();  // before smem goes away
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

// =============================================================================================
// v5: the v3 pipeline with the softmax split over TWO warp roles per TMEM lane quarter.
//
// What bounds v3 (tools/dev_attn_timeline.py): per query block a softmax warp spends
// ~1300 cycles in tcgen05.ld (the S block leaves TMEM at ~80 B/clk per SM, and the instruction holds
// the issuing warp until its data is back) and ~1850 cycles in the exp2 loop (MUFU-bound), one
// after the other, and the two warpgroups take turns.  Prefetching, an FMA-pipe exp2 and a column
// split between two warps of a quarter all failed to overlap the two because each warp still did
// both jobs.  Here the jobs belong to different warps of the same scheduler:
//   warps 4-7   "loaders": tcgen05.ld the S chunks and hand them over through a shared-memory
//               ring (2 x 4 KB per lane quarter, 128B-swizzled, mbarrier full / empty pairs);
//               afterwards the epilogue of the PREVIOUS block (O from TMEM, 1/rowsum, bf16, TMA store)
//   warps 8-11  "math": first chunk straight from TMEM (it yields the shift), the others from the
//               ring with ld.shared (which does not hold the warp), exp2, row sum, P back into TMEM
// so one quarter's TMEM reads run under the same quarter's MUFU work, block after block, and no
// turn-taking is needed.  Arithmetic (shift, exp2 inputs, summation order, rounding of P and O) is
// v3's own code, so the output is bit-identical to v3.  K/V/Q staging, the MMA issuer and the
// overflow redo are v3's.  Shared memory: 2 stages x (Q0|Q1|K|V) + one 16 KB output tile + the
// 32 KB ring: fits for S <= 208 (ViT @224: S = 197); longer sequences keep v3.
// =============================================================================================
constexpr int ATT5_MAX_SK = 208;

struct Attn5SmemLayout {
  int sk, kv_bytes, stage_bytes, off_o, off_ring, off_sum, off_bar, total;
};
__host__ __device__ inline Attn5SmemLayout attn5_smem_layout(int S) {
  Attn5SmemLayout L;
  L.sk = (S + 15) & ~15;
  L.kv_bytes = ((L.sk * 128) + 1023) & ~1023;
  L.stage_bytes = 2 * 16384 + 2 * L.kv_bytes;   // Q0 | Q1 | K | V
  L.off_o = ATT3_STAGES * L.stage_bytes;        // one 16 KB output tile (only the loader group stores)
  L.off_ring = L.off_o + 16384;                 // [quarter][buffer] 4 KB chunks of S
  L.off_sum = L.off_ring + 4 * 2 * 4096;        // [slot][row] row sums, math -> loaders
  L.off_bar = L.off_sum + 2 * 128 * 4;
  L.total = L.off_bar + 512 + 1024;
  return L;
}

__global__ void __launch_bounds__(ATT3_THREADS, 1)
attention_split_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                       const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  // aligned by POINTER arithmetic on the __shared__ array: an integer round trip makes every access
  // through the result a generic LD / ST (with R2UR descriptor traffic) instead of LDS / STS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const Attn5SmemLayout L = attn5_smem_layout(p.S);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* kv_full = bars + 0;      // [2] stage
  uint64_t* v_full = bars + 2;       // [2] stage
  uint64_t* stage_empty = bars + 4;  // [2] stage, count 2 (one per query block of the unit)
  uint64_t* s_full = bars + 6;       // [2] slot
  uint64_t* p_full = bars + 8;       // [2] slot, count 4 (math warps)
  uint64_t* o_full = bars + 10;      // [2] slot
  uint64_t* slot_free = bars + 12;   // [2] slot, count 4 (loader warps)
  uint64_t* redo_bar = bars + 14;    // [2] slot
  uint64_t* sum_ready = bars + 16;   // [2] slot, count 4 (math warps): row sums are in shared memory
  uint64_t* ring_full = bars + 18;   // [4 quarters][2 buffers]
  uint64_t* ring_empty = bars + 26;  // [4 quarters][2 buffers]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 34);
  float* s_sum = reinterpret_cast<float*>(smem + L.off_sum);

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx(), lane = tid & 31;
  const int D = p.heads * ATT_D;
  const int S = p.S, SK = L.sk;
  const int units = p.frames * p.heads;
  const int nblk = p.nblk;
  const int n32 = SK >> 5;
  const bool tail = (SK & 16) != 0;
  const int nchunks = n32 + (tail ? 1 : 0);

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmO);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&stage_empty[i], 2);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_full[i], 1);
      mbar_init(&slot_free[i], 4);
      mbar_init(&redo_bar[i], 1);
      mbar_init(&sum_ready[i], 4);
    }
    for (int i = 0; i < 8; ++i) {
      mbar_init(&ring_full[i], 1);
      mbar_init(&ring_empty[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ loader of Q / K / V (as v3) ------------------------------
    if (lane == 0) {
      int j = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const int uu = p.reverse ? units - 1 - u : u;
        const int frame = uu / p.heads, head = uu % p.heads;
        uint8_t* sb = smem + st * L.stage_bytes;
        mbar_wait(&stage_empty[st], ph ^ 1u);
        mbar_arrive_expect_tx(&kv_full[st], 16384 * nblk + SK * 128);
        tma_load_3d(sb, &tmQ, &kv_full[st], head * ATT_D, 0, frame);
        if (nblk == 2) tma_load_3d(sb + 16384, &tmQ, &kv_full[st], head * ATT_D, 128, frame);
        tma_load_3d(sb + 32768, &tmKV, &kv_full[st], D + head * ATT_D, 0, frame);
        mbar_arrive_expect_tx(&v_full[st], SK * 128);
        tma_load_3d(sb + 32768 + L.kv_bytes, &tmKV, &v_full[st], 2 * D + head * ATT_D, 0, frame);
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (as v3) ------------------------------
    if (lane == 0) {
      const int my_units = (units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
      const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
      const uint32_t idesc_o = make_idesc_bf16_f32(128, ATT_D, false, true);
      int uj[2] = {0, 0};
      int phs[2] = {0, 0};
      long long t0 = clock64();
      while (uj[0] < my_units || uj[1] < my_units) {
        bool progressed = false;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const int j = uj[s];
          if (j >= my_units) continue;
          const int st = j & 1;
          const uint32_t sph = (j >> 1) & 1, uph = j & 1;
          const int b = s ^ (j & 1);
          uint8_t* sb = smem + st * L.stage_bytes;
          const uint32_t t_slot = tmem_base + s * 256;
          if (phs[s] == 0) {
            if (!mbar_test_wait(&kv_full[st], sph) || !mbar_test_wait(&slot_free[s], uph ^ 1u)) continue;
            tc_fence_after_sync();
            if (b < nblk) {
              const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
              const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
              for (int k = 0; k < ATT_D / 16; ++k)
                umma_bf16_ss(t_slot, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
            }
            umma_commit(&s_full[s]);
            phs[s] = 1;
          } else {
            if (!mbar_test_wait(&p_full[s], uph) || !mbar_test_wait(&v_full[st], sph)) continue;
            tc_fence_after_sync();
            if (b < nblk) {
              const uint32_t sv = smem_u32(sb + 32768 + L.kv_bytes);
              const uint64_t dv0 = make_smem_desc_sw128(sv, 1024, 1024);
              const uint32_t t_o_acc = t_slot + SK / 2;
              umma_bf16_ts(t_o_acc, t_slot, dv0, idesc_o, 0u);
#pragma unroll 4
              for (int kk = 1; kk < SK / 16; ++kk)
                umma_bf16_ts(t_o_acc, t_slot + kk * 8, dv0 + (uint64_t)(kk * 128), idesc_o, 1u);
            }
            umma_commit(&o_full[s]);
            phs[s] = 0;
            uj[s] = j + 1;
          }
          progressed = true;
        }
        if (progressed) t0 = clock64();
        else if (clock64() - t0 > 4000000000LL) {
          printf("vitk: attention(v5) MMA poll timeout block %d\n", (int)blockIdx.x);
          __trap();
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ------------------------------ loaders: S -> ring, then the previous block's epilogue ------
    const int wq = warp & 3;
    const int r = wq * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(wq * 32) << 16;
    uint8_t* ring = smem + L.off_ring + wq * 8192;
    uint64_t* rf = ring_full + wq * 2;
    uint64_t* re = ring_empty + wq * 2;
    uint8_t* sO = smem + L.off_o;
    uint32_t lc = 0;  // chunks this warp has pushed through the ring
    // epilogue of item (slot s, unit parity uph, ...): O from TMEM -> / rowsum -> bf16 -> TMA store
    auto epilogue = [&](int s, uint32_t uph, int st, int b, int frame, int head) {
      const bool blk_active = b < nblk;
      const int rows_valid = blk_active ? min(128, p.q_rows - b * 128) : 0;
      const bool warp_active = wq * 32 < rows_valid;
      const uint32_t t_o = tmem_base + s * 256 + lane_off + static_cast<uint32_t>(SK / 2);
      mbar_wait(&o_full[s], uph);
      tc_fence_after_sync();
      if (wq == 0 && lane == 0) {
        mbar_arrive(&stage_empty[st]);   // O = P V has retired: this block no longer needs Q / K / V
        tma_store_wait_read<0>();        // the previous block's store has drained the output tile
      }
      named_bar_sync(3, 128);
      if (warp_active) {
        mbar_wait(&sum_ready[s], uph);
        const float inv = 1.0f / s_sum[s * 128 + r];
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(t_o, v0);
        tmem_ld_32x32b_x32(t_o + 32, v1);
        tmem_ld_wait();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[s]);  // O is in registers: the slot can be refilled
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t* src = i < 4 ? v0 + 8 * i : v1 + 8 * (i - 4);
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(src[0]) * inv, __uint_as_float(src[1]) * inv);
          o.y = pack_bf16x2(__uint_as_float(src[2]) * inv, __uint_as_float(src[3]) * inv);
          o.z = pack_bf16x2(__uint_as_float(src[4]) * inv, __uint_as_float(src[5]) * inv);
          o.w = pack_bf16x2(__uint_as_float(src[6]) * inv, __uint_as_float(src[7]) * inv);
          *reinterpret_cast<uint4*>(sO + r * 128 + ((i ^ (r & 7)) << 4)) = o;
        }
        fence_proxy_async_smem();
      } else {
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[s]);
      }
      named_bar_sync(3, 128);
      if (wq == 0 && lane == 0 && blk_active) {
        tma_store_3d(&tmO, sO, head * ATT_D, b * 128, frame);
        tma_store_commit();
      }
    };
    bool have_prev = false;
    int ps = 0, pst = 0, pb = 0, pframe = 0, phead = 0;
    uint32_t puph = 0;
    int j = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
      const int st = j & 1;
      const uint32_t uph = j & 1;
      const int uu = p.reverse ? units - 1 - u : u;
      const int frame = uu / p.heads, head = uu % p.heads;
#pragma unroll 1
      for (int s = 0; s < 2; ++s) {
        const int b = s ^ (j & 1);
        const int rows_valid = b < nblk ? min(128, p.q_rows - b * 128) : 0;
        const bool warp_active = wq * 32 < rows_valid;
        const uint32_t t_slot = tmem_base + s * 256 + lane_off;
        const int item = 2 * j + s;
        const bool dbg_on = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && wq == 0 && lane == 0 && item < 16;
        long long* dbg = p.dbg + (0 * 16 + item) * 8;
        if (dbg_on) dbg[0] = clock64();
        mbar_wait(&s_full[s], uph);
        tc_fence_after_sync();
        if (dbg_on) dbg[1] = clock64();
        if (warp_active) {
#pragma unroll 1
          for (int c = 1; c < nchunks; ++c, ++lc) {   // chunk 0 is read by the math warp itself
            const uint32_t buf = lc & 1u;
            uint8_t* dst = ring + buf * 4096 + lane * 128;
            uint32_t v[32];
            if (c < n32) {
              tmem_ld_32x32b_x32(t_slot + c * 32, v);
              tmem_ld_wait();
              mbar_wait(&re[buf], ((lc >> 1) & 1u) ^ 1u);
#pragma unroll
              for (int g = 0; g < 8; ++g)
                *reinterpret_cast<uint4*>(dst + ((g ^ (lane & 7)) << 4)) =
                    make_uint4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
            } else {
              tmem_ld_32x32b_x16(t_slot + c * 32, *reinterpret_cast<uint32_t(*)[16]>(v));
              tmem_ld_wait();
              mbar_wait(&re[buf], ((lc >> 1) & 1u) ^ 1u);
#pragma unroll
              for (int g = 0; g < 4; ++g)
                *reinterpret_cast<uint4*>(dst + ((g ^ (lane & 7)) << 4)) =
                    make_uint4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&rf[buf]);
          }
        }
        if (dbg_on) dbg[2] = clock64();
        if (have_prev) epilogue(ps, puph, pst, pb, pframe, phead);
        if (dbg_on) dbg[3] = clock64();
        have_prev = true;
        ps = s; puph = uph; pst = st; pb = b; pframe = frame; phead = head;
      }
    }
    if (have_prev) epilogue(ps, puph, pst, pb, pframe, phead);
  } else if (warp >= 8) {
    // ------------------------------ math: exp2, row sums, P -> TMEM ------------------------------
    const int wq = warp & 3;
    const int r = wq * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(wq * 32) << 16;
    const uint8_t* ring = smem + L.off_ring + wq * 8192;
    uint64_t* rf = ring_full + wq * 2;
    uint64_t* re = ring_empty + wq * 2;
    uint32_t mc = 0;  // chunks this warp has taken out of the ring
    uint32_t redo_phase[2] = {0u, 0u};
    int j = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
      const int st = j & 1;
      const uint32_t uph = j & 1;
#pragma unroll 1
      for (int s = 0; s < 2; ++s) {
        const int b = s ^ (j & 1);
        const bool blk_active = b < nblk;
        const int rows_valid = blk_active ? min(128, p.q_rows - b * 128) : 0;
        const bool warp_active = wq * 32 < rows_valid;
        const uint32_t t_slot = tmem_base + s * 256 + lane_off;
        const int item = 2 * j + s;
        const bool dbg_on = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && wq == 0 && lane == 0 && item < 16;
        long long* dbg = p.dbg + (1 * 16 + item) * 8;
        if (dbg_on) dbg[0] = clock64();
        mbar_wait(&s_full[s], uph);
        tc_fence_after_sync();
        if (dbg_on) dbg[1] = clock64();
        float sum = 0.f;
        bool redo = false;
        // next chunk out of the ring into `dst` (floats as raw bits), releasing its buffer
        auto ring_take = [&](uint32_t* dst, bool full32) {
          const uint32_t buf = mc & 1u;
          const uint8_t* src = ring + buf * 4096 + lane * 128;
          mbar_wait(&rf[buf], (mc >> 1) & 1u);
          if (full32) {
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const uint4 t = *reinterpret_cast<const uint4*>(src + ((g ^ (lane & 7)) << 4));
              dst[4 * g] = t.x; dst[4 * g + 1] = t.y; dst[4 * g + 2] = t.z; dst[4 * g + 3] = t.w;
            }
          } else {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint4 t = *reinterpret_cast<const uint4*>(src + ((g ^ (lane & 7)) << 4));
              dst[4 * g] = t.x; dst[4 * g + 1] = t.y; dst[4 * g + 2] = t.z; dst[4 * g + 3] = t.w;
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&re[buf]);
          ++mc;
        };
        if (warp_active) {
          uint32_t va[32], vb[32];
          if (n32 > 0) tmem_ld_32x32b_x32(t_slot, va);
          else tmem_ld_32x32b_x16(t_slot, *reinterpret_cast<uint32_t(*)[16]>(va));
          tmem_ld_wait();
          reg_fence16(va);
          if (n32 > 0) reg_fence16(va + 16);
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
          const int first = n32 > 0 ? 32 : 16;
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (i < first && i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(va[i]));
          const float mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2e;
          float2 s4[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
          if (dbg_on) dbg[2] = clock64();
          // v3's chunk order and accumulation pattern: va / vb alternate, two 32-column chunks per turn
          int c = 0;
          for (; c + 2 <= n32; c += 2) {  // va holds chunk c
            ring_take(vb, true);
            att_exp_chunk<32>(va, c * 32, S, p.scale_log2e, mxs, s4, t_slot);
            if (c + 2 < n32) ring_take(va, true);
            else if (tail) ring_take(va, false);
            att_exp_chunk<32>(vb, (c + 1) * 32, S, p.scale_log2e, mxs, s4, t_slot);
          }
          if (c < n32) {  // odd number of 32-column chunks: va holds the last one
            if (tail) ring_take(vb, false);
            att_exp_chunk<32>(va, c * 32, S, p.scale_log2e, mxs, s4, t_slot);
            if (tail) att_exp_chunk<16>(vb, (c + 1) * 32, S, p.scale_log2e, mxs, s4, t_slot);
          } else if (tail) {  // va holds the 16-column tail
            att_exp_chunk<16>(va, n32 * 32, S, p.scale_log2e, mxs, s4, t_slot);
          }
          tmem_st_wait();
          sum = (s4[0].x + s4[0].y) + (s4[1].x + s4[1].y);
          redo = !(sum < 1.0e30f);
        }
        if (dbg_on) dbg[3] = clock64();
        // overflow of the single-pass shift: the whole block takes v3's exact path (S recomputed
        // into the slot, row maximum, second exp2 pass), run by the math warps straight from TMEM
        tc_fence_before_sync();
        if (named_bar_or(2, 128, redo)) {
          if (wq == 0 && lane == 0 && blk_active) {
            tc_fence_after_sync();
            uint8_t* sb = smem + st * L.stage_bytes;
            const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
            const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
            const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
            for (int k = 0; k < ATT_D / 16; ++k)
              umma_bf16_ss(tmem_base + s * 256, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
            umma_commit(&redo_bar[s]);
          }
          if (blk_active) {
            mbar_wait(&redo_bar[s], redo_phase[s]);
            redo_phase[s] ^= 1u;
            tc_fence_after_sync();
          }
          if (warp_active) {
            float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
            for (int c0 = 0; c0 < SK; c0 += 16) {
              uint32_t v[16];
              tmem_ld_32x32b_x16(t_slot + c0, v);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (c0 + i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
            }
            const float mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2e;
            float2 s4[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
            for (int c0 = 0; c0 < SK; c0 += 16) {
              uint32_t v[16];
              tmem_ld_32x32b_x16(t_slot + c0, v);
              tmem_ld_wait();
              reg_fence16(v);
              att_exp_chunk<16>(v, c0, S, p.scale_log2e, mxs, s4, t_slot);
            }
            tmem_st_wait();
            sum = (s4[0].x + s4[0].y) + (s4[1].x + s4[1].y);
          }
        }
        if (warp_active) s_sum[s * 128 + r] = sum;
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&sum_ready[s]);
          mbar_arrive(&p_full[s]);
        }
        if (dbg_on) dbg[4] = clock64();
      }
    }
  }

  if (warp == 4 && lane == 0) tma_store_wait_read<0>();  // before smem goes away
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

// =============================================================================================
// v6: v3 with the probabilities in SHARED memory instead of TMEM.
//
// What the round-2 timelines showed (tools/dev_attn_timeline.py): per (frame, head)
// the TMEM read port is busy ~5 550 of v3's ~7 000 cycles -- 2 x 1 300 for tcgen05.ld of the S
// blocks, 2 x 400 for O, and 2 x 1 075 for the 13 TS-form P.V instructions, which fetch their A
// operand (P) from TMEM at ~83 cycles each where the tensor pipe needs 32.  The S = Q K^T of the
// other slot queues behind them (wait_S ~700 cycles).  Overlap experiments (hand-over, v5, two
// issuing threads) cannot beat a saturated port; taking P out of TMEM can:
//   * the softmax warps write P (bf16) with st.shared into K-major 128B-swizzled 64-key tiles --
//     the layout of a Q tile -- and P.V becomes an SS-form tcgen05.mma (A = P tile, B = V);
//   * P.V is issued in two halves: keys 0-127 as soon as the first four chunks of the pass are
//     written (it runs under the rest of the pass), the remaining keys at the end of the pass;
//   * 48 KB hold the tiles (X: keys 0-127, Y: keys 128-191; the 16-key tail re-uses X tile 0 once the
//     first half has been consumed), shared by both slots because the two warpgroups take turns;
//     the space comes from staging O in the (by then dead) Q tile of its own block.
// Arithmetic per element is v3's (same shift, exp2 inputs, summation order, bf16 rounding of P);
// O accumulates the same 13 K-steps in the same order.  Needs S <= 208 like v5; longer sequences keep v3.
// =============================================================================================
struct Attn6SmemLayout {
  int sk, kv_bytes, stage_bytes, off_p, off_bar, total;
};
__host__ __device__ inline Attn6SmemLayout attn6_smem_layout(int S) {
  Attn6SmemLayout L;
  L.sk = (S + 15) & ~15;
  L.kv_bytes = ((L.sk * 128) + 1023) & ~1023;
  L.stage_bytes = 2 * 16384 + 2 * L.kv_bytes;   // Q0 | Q1 | K | V ; O of block b is staged over Q_b
  L.off_p = ATT3_STAGES * L.stage_bytes;        // X tile 0 | X tile 1 | Y : three 16 KB P tiles
  L.off_bar = L.off_p + 3 * 16384;
  L.total = L.off_bar + 512 + 1024;
  return L;
}

// exp2(s*scale - mxs) of NC = 32 or 16 scores starting at key c0 -> bf16, written into the P tile
// row `prow` (128 B = 64 keys, 16-byte granules XOR-swizzled with the row index); accumulates the
// row sum exactly like att_exp_chunk.  r7 = row & 7.
template <int NC>
__device__ __forceinline__ void att_exp_chunk_smem(const uint32_t* v, int c0, int S, float scale, float mxs,
                                                   float2 (&s2)[2], uint8_t* prow, int r7) {
  uint32_t pk[NC / 2];
  const bool full = c0 + NC <= S;
  const float2 sc2 = make_float2(scale, scale), nm2 = make_float2(-mxs, -mxs);
#pragma unroll
  for (int i = 0; i < NC / 2; ++i) {
    const float2 x = __ffma2_rn(make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), sc2, nm2);
    float2 e = att_exp2_pair(x, i);
    if (!full) {
      e.x = (c0 + 2 * i < S) ? e.x : 0.f;
      e.y = (c0 + 2 * i + 1 < S) ? e.y : 0.f;
    }
    s2[i & 1] = __fadd2_rn(s2[i & 1], e);
    pk[i] = pack_bf16x2(e.x, e.y);
  }
  const int g0 = ((c0 & 63) >> 3);   // first 16-byte granule of this chunk inside the 64-key row
#pragma unroll
  for (int i = 0; i < NC / 8; ++i)
    *reinterpret_cast<uint4*>(prow + (((g0 + i) ^ r7) << 4)) =
        make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
}

__global__ void __launch_bounds__(ATT3_THREADS, 1)
attention_smemp_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                       const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  // aligned by POINTER arithmetic on the __shared__ array: an integer round trip makes every access
  // through the result a generic LD / ST (with R2UR descriptor traffic) instead of LDS / STS
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const Attn6SmemLayout L = attn6_smem_layout(p.S);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* kv_full = bars + 0;      // [2] stage
  uint64_t* v_full = bars + 2;       // [2] stage
  uint64_t* stage_empty = bars + 4;  // [2] stage, count 2 (one per query block, after its O store)
  uint64_t* s_full = bars + 6;       // [2] slot
  uint64_t* p_half = bars + 8;       // [2] slot, count 4: keys 0-127 of P are in shared memory
  uint64_t* p_full = bars + 10;      // [2] slot, count 4: all of P is
  uint64_t* pv1_done = bars + 12;    // [2] slot: the first-half P.V instructions have retired
  uint64_t* o_full = bars + 14;      // [2] slot
  uint64_t* slot_free = bars + 16;   // [2] slot, count 4
  uint64_t* redo_bar = bars + 18;    // [2] slot
  uint64_t* turn_bar = bars + 20;    // [2] group, count 4
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 22);
  volatile int* redo_flag = reinterpret_cast<volatile int*>(bars + 23);  // [2] slot: P.V was issued by the exact path

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx(), lane = tid & 31;
  const int D = p.heads * ATT_D;
  const int S = p.S, SK = L.sk;
  const int units = p.frames * p.heads;
  const int nblk = p.nblk;
  const int nk = SK >> 4;                 // 16-key K-steps of P.V (<= 13)
  const int nk1 = nk < 8 ? nk : 8;        // first half: keys 0-127 (tiles X0, X1)
  uint8_t* sP = smem + L.off_p;
  // shared-memory address of K-step kk's P operand: tile kk/4 (tile 3, the 16-key tail, lives in X tile 0)
  auto p_tile_of = [&](int kk) { const int t = kk >> 2; return t == 3 ? 0 : t; };

  if (tid == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmO);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&stage_empty[i], 2);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_half[i], 4);
      mbar_init(&p_full[i], 4);
      mbar_init(&pv1_done[i], 1);
      mbar_init(&o_full[i], 1);
      mbar_init(&slot_free[i], 4);
      mbar_init(&redo_bar[i], 1);
      mbar_init(&turn_bar[i], 4);
      redo_flag[i] = 0;
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc_s = make_idesc_bf16_f32(128, (uint32_t)SK, false, false);
  const uint32_t idesc_o = make_idesc_bf16_f32(128, ATT_D, false, true);
  // P.V K-steps [k0, k1) of one block: A = P tile (K-major, like Q), B = V (MN-major as TMA delivered it)
  auto issue_pv = [&](uint32_t t_o_acc, uint8_t* sb, int k0, int k1, bool first_clears) {
    const uint32_t sv = smem_u32(sb + 32768 + L.kv_bytes);
    const uint64_t dv0 = make_smem_desc_sw128(sv, 1024, 1024);
    for (int kk = k0; kk < k1; ++kk) {
      const uint64_t dp = make_smem_desc_sw128(smem_u32(sP + p_tile_of(kk) * 16384), 1024, 0) +
                          static_cast<uint64_t>(2 * (kk & 3));
      umma_bf16_ss(t_o_acc, dp, dv0 + (uint64_t)(kk * 128), idesc_o, (kk == k0 && first_clears) ? 0u : 1u);
    }
  };

  if (warp == 0) {
    // ------------------------------ loader (as v3) ------------------------------
    if (lane == 0) {
      int j = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const int uu = p.reverse ? units - 1 - u : u;
        const int frame = uu / p.heads, head = uu % p.heads;
        uint8_t* sb = smem + st * L.stage_bytes;
        mbar_wait(&stage_empty[st], ph ^ 1u);
        mbar_arrive_expect_tx(&kv_full[st], 16384 * nblk + SK * 128);
        tma_load_3d(sb, &tmQ, &kv_full[st], head * ATT_D, 0, frame);
        if (nblk == 2) tma_load_3d(sb + 16384, &tmQ, &kv_full[st], head * ATT_D, 128, frame);
        tma_load_3d(sb + 32768, &tmKV, &kv_full[st], D + head * ATT_D, 0, frame);
        mbar_arrive_expect_tx(&v_full[st], SK * 128);
        tma_load_3d(sb + 32768 + L.kv_bytes, &tmKV, &v_full[st], 2 * D + head * ATT_D, 0, frame);
      }
    }
  } else if (warp == 1 || warp == 3) {
    // ------------------------------ MMA issuers: warp 1 = slot 0, warp 3 = slot 1 ------------------
    if (lane == 0) {
      const int my_units = (units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
      const int s = warp == 1 ? 0 : 1;
      const uint32_t t_slot = tmem_base + s * 256;
      for (int j = 0; j < my_units; ++j) {
        const int st = j & 1;
        const uint32_t sph = (j >> 1) & 1, uph = j & 1;
        const int b = s ^ (j & 1);
        uint8_t* sb = smem + st * L.stage_bytes;
        mbar_wait(&kv_full[st], sph);
        mbar_wait(&slot_free[s], uph ^ 1u);
        tc_fence_after_sync();
        if (b < nblk) {
          const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
          const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
          for (int k = 0; k < ATT_D / 16; ++k)
            umma_bf16_ss(t_slot, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
        }
        umma_commit(&s_full[s]);
        // first half of O = P V: keys 0-127, under the rest of the softmax pass
        mbar_wait(&p_half[s], uph);
        mbar_wait(&v_full[st], sph);
        tc_fence_after_sync();
        if (b < nblk) issue_pv(t_slot, sb, 0, nk1, true);   // O accumulates at columns 0-63 of the slot
        umma_commit(&pv1_done[s]);
        // second half at the end of the pass (unless the exact path has issued everything itself)
        mbar_wait(&p_full[s], uph);
        tc_fence_after_sync();
        if (redo_flag[s]) {
          redo_flag[s] = 0;
        } else {
          if (b < nblk) issue_pv(t_slot, sb, nk1, nk, false);
          umma_commit(&o_full[s]);
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------ softmax + epilogue ------------------------------
    const int g = (warp - 4) >> 2;  // warpgroup == TMEM slot
    const int wq = warp & 3;        // TMEM lane quarter
    const int r = wq * 32 + lane;   // row within the query block == TMEM lane
    const int r7 = r & 7;
    const uint32_t lane_off = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t t_slot = tmem_base + g * 256 + lane_off;
    const uint32_t t_o = t_slot;    // O accumulates over S columns 0-63 (read before the first-half P.V)
    const int n32 = SK >> 5;
    const bool tail = (SK & 16) != 0;
    const int half_cols = nk1 * 16;   // keys of the first-half P.V (128, or all of a short sequence)
    // row r of the P tile that holds keys [c0, c0 + 32): tiles X0, X1, Y, and X0 again for keys >= 192
    auto prow_of = [&](int c0) { return sP + p_tile_of(c0 >> 4) * 16384 + r * 128; };
    int j = 0;
    uint32_t redo_phase = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++j) {
      const int st = j & 1;
      const uint32_t uph = j & 1;
      const int b = g ^ (j & 1);
      const int uu = p.reverse ? units - 1 - u : u;
      const int frame = uu / p.heads, head = uu % p.heads;
      const bool blk_active = b < nblk;
      const int rows_valid = blk_active ? min(128, p.q_rows - b * 128) : 0;
      const bool warp_active = wq * 32 < rows_valid;
      uint8_t* sb = smem + st * L.stage_bytes;
      uint8_t* sO = sb + b * 16384;   // this block's Q tile: dead once S = Q K^T (and any redo) is done

      const bool dbg_on = (VITK_ATT_TIMELINE && p.dbg != nullptr) && blockIdx.x == 0 && wq == 0 && lane == 0 && j < 16;
      long long* dbg = p.dbg + (g * 16 + j) * 8;
      if (dbg_on) dbg[0] = clock64();
      mbar_wait(&s_full[g], uph);
      if (dbg_on) dbg[6] = clock64();
      // the two warpgroups take turns in the softmax (see v3); the turn also hands over the P tiles:
      // the other group's second-half P.V (which reads Y and the tail in X tile 0) must have retired
      if (g == 0) {
        if (j > 0) {
          mbar_wait(&turn_bar[0], (uint32_t)(j - 1) & 1u);
          mbar_wait(&o_full[1], (uint32_t)(j - 1) & 1u);
        }
      } else {
        mbar_wait(&turn_bar[1], uph);
        mbar_wait(&o_full[0], uph);
      }
      tc_fence_after_sync();
      if (dbg_on) dbg[1] = clock64();
      float sum = 0.f;
      bool redo = false;
      bool half_sent = false;
      // keys [0, 128) (or all of them when SK <= 128) are in shared memory: let the first-half P.V go
      auto send_half = [&]() {
        half_sent = true;
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_half[g]);
      };
      if (warp_active) {
        uint32_t va[32], vb[32];
        if (n32 > 0) tmem_ld_32x32b_x32(t_slot, va);
        else tmem_ld_32x32b_x16(t_slot, *reinterpret_cast<uint32_t(*)[16]>(va));
        tmem_ld_wait();
        reg_fence16(va);
        if (n32 > 0) reg_fence16(va + 16);
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        const int first = n32 > 0 ? 32 : 16;
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (i < first && i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(va[i]));
        const float mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2e;
        float2 s4[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        int c = 0;
        for (; c + 2 <= n32; c += 2) {  // va holds chunk c
          tmem_ld_32x32b_x32(t_slot + (c + 1) * 32, vb);
          att_exp_chunk_smem<32>(va, c * 32, S, p.scale_log2e, mxs, s4, prow_of(c * 32), r7);
          tmem_ld_wait();
          reg_fence16(vb);
          reg_fence16(vb + 16);
          if (c + 2 < n32) tmem_ld_32x32b_x32(t_slot + (c + 2) * 32, va);
          else if (tail) tmem_ld_32x32b_x16(t_slot + (c + 2) * 32, *reinterpret_cast<uint32_t(*)[16]>(va));
          att_exp_chunk_smem<32>(vb, (c + 1) * 32, S, p.scale_log2e, mxs, s4, prow_of((c + 1) * 32), r7);
          tmem_ld_wait();
          reg_fence16(va);
          reg_fence16(va + 16);
          if (!half_sent && (c + 2) * 32 >= half_cols) send_half();   // the first half's keys are written
          if ((c + 2) * 32 == 192 && tail) {    // the tail goes into X tile 0: first-half P.V must be done with it
            mbar_wait(&pv1_done[g], uph);
          }
        }
        if (c < n32) {  // odd number of 32-column chunks: va holds the last one
          if (tail) tmem_ld_32x32b_x16(t_slot + (c + 1) * 32, *reinterpret_cast<uint32_t(*)[16]>(vb));
          att_exp_chunk_smem<32>(va, c * 32, S, p.scale_log2e, mxs, s4, prow_of(c * 32), r7);
          tmem_ld_wait();
          reg_fence16(vb);
          if (!half_sent && (c + 1) * 32 >= half_cols) send_half();
          if (tail) {
            if ((c + 1) * 32 >= 192) mbar_wait(&pv1_done[g], uph);
            att_exp_chunk_smem<16>(vb, (c + 1) * 32, S, p.scale_log2e, mxs, s4, prow_of((c + 1) * 32), r7);
          }
        } else if (tail) {  // va holds the 16-column tail
          if (n32 * 32 >= 192) mbar_wait(&pv1_done[g], uph);
          att_exp_chunk_smem<16>(va, n32 * 32, S, p.scale_log2e, mxs, s4, prow_of(n32 * 32), r7);
        }
        sum = (s4[0].x + s4[0].y) + (s4[1].x + s4[1].y);
        redo = !(sum < 1.0e30f);
      }
      if (dbg_on) dbg[2] = clock64();
      if (!half_sent) send_half();   // inactive warps and short sequences: the arrival is still owed
      // Overflow of the single-pass shift -> the EXACT path for the whole block: S recomputed, row
      // maximum, second exp2 pass, and this group issues the block's P.V itself (the issuer thread
      // skips its second half when redo_flag is set).  Rare (never seen on real activations).
      tc_fence_before_sync();
      if (named_bar_or(2 + g, 128, redo)) {
        const bool leader = wq == 0 && lane == 0;
        // the first-half P.V issued on the overflowed probabilities must have retired before S is
        // recomputed over its accumulator columns (it is re-issued from scratch below)
        mbar_wait(&pv1_done[g], uph);
        if (leader && blk_active) {
          tc_fence_after_sync();
          const uint64_t dq = make_smem_desc_sw128(smem_u32(sb + b * 16384), 1024, 0);
          const uint64_t dk = make_smem_desc_sw128(smem_u32(sb + 32768), 1024, 0);
#pragma unroll
          for (int k = 0; k < ATT_D / 16; ++k)
            umma_bf16_ss(tmem_base + g * 256, dq + 2 * k, dk + 2 * k, idesc_s, k != 0 ? 1u : 0u);
          umma_commit(&redo_bar[g]);   // in order behind the (garbage) first-half P.V already issued
        }
        if (blk_active) {
          mbar_wait(&redo_bar[g], redo_phase);
          redo_phase ^= 1u;
          tc_fence_after_sync();
        }
        float mxs = 0.f;
        float2 s4[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        const int main_cols = SK < 192 ? SK : 192;     // keys that fit tiles X0, X1, Y
        if (warp_active) {
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
          for (int c0 = 0; c0 < SK; c0 += 16) {
            uint32_t v[16];
            tmem_ld_32x32b_x16(t_slot + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (c0 + i < S) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[i]));
          }
          mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2e;
          for (int c0 = 0; c0 < main_cols; c0 += 16) {
            uint32_t v[16];
            tmem_ld_32x32b_x16(t_slot + c0, v);
            tmem_ld_wait();
            reg_fence16(v);
            att_exp_chunk_smem<16>(v, c0, S, p.scale_log2e, mxs, s4, prow_of(c0), r7);
          }
        }
        uint32_t vt[16];
        if (warp_active && SK > 192) {   // the tail's scores leave TMEM before O = P V overwrites nothing of them
          tmem_ld_32x32b_x16(t_slot + 192, vt);
          tmem_ld_wait();
          reg_fence16(vt);
        }
        fence_proxy_async_smem();
        tc_fence_before_sync();
        named_bar_sync(2 + g, 128);
        if (leader && blk_active) {
          tc_fence_after_sync();
          issue_pv(tmem_base + g * 256, sb, 0, main_cols >> 4, true);
          umma_commit(&redo_bar[g]);
        }
        if (blk_active) {
          mbar_wait(&redo_bar[g], redo_phase);
          redo_phase ^= 1u;
          tc_fence_after_sync();
        }
        if (warp_active && SK > 192)
          att_exp_chunk_smem<16>(vt, 192, S, p.scale_log2e, mxs, s4, prow_of(192), r7);
        sum = (s4[0].x + s4[0].y) + (s4[1].x + s4[1].y);
        fence_proxy_async_smem();
        tc_fence_before_sync();
        named_bar_sync(2 + g, 128);
        if (leader) {
          if (blk_active) {
            tc_fence_after_sync();
            if (SK > 192) issue_pv(tmem_base + g * 256, sb, 12, nk, false);
          }
          redo_flag[g] = 1;
          umma_commit(&o_full[g]);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&p_full[g]);
        mbar_arrive(&turn_bar[g ^ 1]);  // hand the exp2 units (and the P tiles) to the other warpgroup
      }
      if (dbg_on) dbg[3] = clock64();

      mbar_wait(&o_full[g], uph);
      tc_fence_after_sync();
      if (dbg_on) dbg[4] = clock64();
      if (warp_active) {
        const float inv = 1.0f / sum;
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(t_o, v0);
        tmem_ld_32x32b_x32(t_o + 32, v1);
        tmem_ld_wait();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[g]);  // O is in registers: the slot can be refilled
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t* src = i < 4 ? v0 + 8 * i : v1 + 8 * (i - 4);
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(src[0]) * inv, __uint_as_float(src[1]) * inv);
          o.y = pack_bf16x2(__uint_as_float(src[2]) * inv, __uint_as_float(src[3]) * inv);
          o.z = pack_bf16x2(__uint_as_float(src[4]) * inv, __uint_as_float(src[5]) * inv);
          o.w = pack_bf16x2(__uint_as_float(src[6]) * inv, __uint_as_float(src[7]) * inv);
          *reinterpret_cast<uint4*>(sO + r * 128 + ((i ^ (r & 7)) << 4)) = o;
        }
        fence_proxy_async_smem();
      } else {
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[g]);
      }
      named_bar_sync(2 + g, 128);
      if (wq == 0 && lane == 0) {
        if (blk_active) {
          // rows of idle warps hold stale bytes: tokens >= S are clipped by the 3-D map, and with
          // q_rows < S (last-layer CLS shortcut) the extra rows are never read
          tma_store_3d(&tmO, sO, head * ATT_D, b * 128, frame);
          tma_store_commit();
          tma_store_wait_read<0>();   // the tile lives in the stage's Q block: read out before release
        }
        mbar_arrive(&stage_empty[st]);
      }
      if (dbg_on) dbg[5] = clock64();
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace vitk
