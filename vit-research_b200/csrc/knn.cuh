// knn.cuh — nearest-neighbour top-k + left/right/none vote against a replicated embedding bank.
//
// Reference behaviour being replaced:
//   exact cosine top-k     nba_proj/retrieval/ratt_chunk_retriever.py:74-75,115-116,123,144-148
//   Chroma L2 query        nba_proj/chroma.py:211-214, generate_clips_hmm.py:403-406 (squared L2)
//   vote                   nba_proj/chroma.py:102-134 (counts, first-max argmax)
// Pipeline: bf16 similarity GEMM (gemm_sm100.cuh) -> warp-shuffle candidate selection of kc >= k
// rows on the bf16 scores (small banks: register-resident select over the score matrix; large
// stores: threshold-filter GEMM epilogue + pool select, no score matrix) -> exact fp32 re-rank of
// the kc candidates -> ordered top-k -> class counts -> label.  The re-rank is what makes the indices match an fp32/fp64 brute force:
// bf16 similarities alone permute near-ties (SURVEY.md §7 "Top-k index exactness").
#pragma once
#include "vitk_common.cuh"
#include "gemm_sm100.cuh"  // KnnFilter

namespace vitk {

enum KnnMetric : int { METRIC_COSINE = 0, METRIC_L2 = 1 };

constexpr int KNN_MAX_CAND = 128;   // kc <= 128 (4 per lane): warp-extraction re-rank
constexpr int KNN_MAX_K = 1024;     // n_results of the clip-store queries reach 750 (config.py:10-11)
constexpr int KNN_SORT_MAX_CAND = 2048;  // kc of the sort-based re-rank (k + k/4 + 32 <= 1312)
constexpr int KNN_MAX_CLASSES = 8;

// f32 rows -> bf16 rows (+ optional x / (||x|| + 1e-8), the reference's query normalisation,
// ratt_chunk_retriever.py:75).  One warp per row.
__global__ void __launch_bounds__(256)
prep_queries_kernel(const float* __restrict__ q, float* __restrict__ q_out,
                    __nv_bfloat16* __restrict__ q_bf16, int n, int d, int normalise) {
  griddep_launch_dependents();
  griddep_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= n) return;
  const float* r = q + (size_t)warp * d;
  float inv = 1.f;
  if (normalise) {
    float s = 0.f;
    for (int i = lane; i < d; i += 32) s += r[i] * r[i];
    inv = 1.0f / (sqrtf(warp_sum(s)) + 1e-8f);
  }
  for (int i = lane; i < d; i += 32) {
    float v = r[i] * inv;
    q_out[(size_t)warp * d + i] = v;
    q_bf16[(size_t)warp * d + i] = __float2bfloat16_rn(v);
  }
}

// Bank rows -> library-owned f32 copy (+ optional b / (||b|| + 1e-8), ratt_chunk_retriever.py:116),
// bf16 copy for the similarity GEMM and ||b||^2 for the squared-L2 order.  One warp per row; the
// norm is accumulated in double like the host path of vitk_bank_set.
__global__ void __launch_bounds__(256)
bank_prep_kernel(const float* __restrict__ rows, float* __restrict__ out, __nv_bfloat16* __restrict__ out_bf16,
                 float* __restrict__ sqnorm, int m, int d, int normalise) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= m) return;
  const float* r = rows + (size_t)warp * d;
  float den = 1.f;
  if (normalise) {
    double s = 0.0;
    for (int i = lane; i < d; i += 32) s += (double)r[i] * r[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    den = (float)sqrt(s) + 1e-8f;
  }
  double s2 = 0.0;
  for (int i = lane; i < d; i += 32) {
    const float v = normalise ? r[i] / den : r[i];
    out[(size_t)warp * d + i] = v;
    out_bf16[(size_t)warp * d + i] = __float2bfloat16_rn(v);
    s2 += (double)v * v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  if (lane == 0) sqnorm[warp] = (float)s2;
}

// Per-call row penalty of a filtered query (Chroma `where=`): the metric's own per-row term for
// the rows in play, +inf for the others, so their keys sink below every real score.
__global__ void __launch_bounds__(256)
knn_penalty_kernel(const uint8_t* __restrict__ row_mask, const float* __restrict__ bank_sqnorm,
                   float* __restrict__ pen, int m, int metric) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  pen[i] = row_mask[i] ? (metric == METRIC_L2 ? bank_sqnorm[i] : 0.f) : INFINITY;
}

// Total order used everywhere: larger key first, ties -> smaller index first.
__device__ __forceinline__ bool key_before(float ka, int ia, float kb, int ib) {
  return ka > kb || (ka == kb && ia < ib);
}

// Warp-wide argmax under key_before; every lane returns the winner.
__device__ __forceinline__ void warp_argbest(float& k, int& i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ok = __shfl_xor_sync(0xffffffffu, k, o);
    int oi = __shfl_xor_sync(0xffffffffu, i, o);
    if (key_before(ok, oi, k, i)) { k = ok; i = oi; }
  }
}

// Candidate selection, register-resident fast path (m <= 32*KPL): one warp per query, lane l
// holds the keys of bank rows l, l+32, ... in registers.  The kc-th largest key is found by a
// 32-step bitwise binary search (count(key >= t) via compare + warp reduce_add: no memory
// traffic, no dependent rescans), then every row above the threshold — and rows equal to it in
// ascending index order until kc are taken — is compacted out with ballots.  The candidate SET
// equals the first kc rows of the (key desc, index asc) order; their order is irrelevant (the
// re-rank sorts them).
template <int KPL>
__global__ void __launch_bounds__(128)
knn_candidates_reg_kernel(const float* __restrict__ scores, int ld,
                          const float* __restrict__ bank_sqnorm, int nq, int m, int kc, int metric,
                          int32_t* __restrict__ cand, const float* __restrict__ pen, const KnnFilter filt) {
  griddep_launch_dependents();
  griddep_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nq) return;
  const float* row = scores + (size_t)warp * ld;
  // All KPL scores of the lane are requested first, unconditionally (index clamped), so that their
  // L2 round trips overlap; with the load inside the per-key `if (i < m)` / metric / predicate
  // branches the compiler issued them one after the other (48 serialised round trips: 31 us for a
  // 1 472-row bank).  The adjustments below are the same arithmetic as before.
  uint32_t key[KPL];
  float sc[KPL];
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const int i = lane + 32 * j;
    sc[j] = __ldg(row + (i < m ? i : m - 1));
  }
  // `pen` (masked query): ||b||^2 (L2) or 0 (cosine) for the rows in play, +inf otherwise
  if (pen != nullptr) {
    float pv[KPL];
#pragma unroll
    for (int j = 0; j < KPL; ++j) { const int i = lane + 32 * j; pv[j] = __ldg(pen + (i < m ? i : m - 1)); }
#pragma unroll
    for (int j = 0; j < KPL; ++j) sc[j] = (metric == METRIC_L2 ? 2.0f * sc[j] : sc[j]) - pv[j];
  } else if (metric == METRIC_L2) {
    float pv[KPL];
#pragma unroll
    for (int j = 0; j < KPL; ++j) { const int i = lane + 32 * j; pv[j] = __ldg(bank_sqnorm + (i < m ? i : m - 1)); }
#pragma unroll
    for (int j = 0; j < KPL; ++j) sc[j] = 2.0f * sc[j] - pv[j];
  }
  // per-query predicate (ratt_chunk_retriever.py:130-134): a rejected row sinks to the bottom
  if (filt.active) {
#pragma unroll
    for (int j = 0; j < KPL; ++j) {
      const int i = lane + 32 * j;
      if (i < m && !knn_filter_pass(filt, warp, i)) sc[j] = -INFINITY;
    }
  }
#pragma unroll
  for (int j = 0; j < KPL; ++j) key[j] = (lane + 32 * j) < m ? float_order_bits(sc[j]) : 0u;
  uint32_t t = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t tt = t | (1u << bit);
    int c = 0;
#pragma unroll
    for (int j = 0; j < KPL; ++j) c += key[j] >= tt ? 1 : 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if (c >= kc) t = tt;
  }
  int32_t* out = cand + (size_t)warp * kc;
  const uint32_t lt_mask = (1u << lane) - 1u;
  int base = 0;
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const bool pr = key[j] > t;
    const uint32_t b = __ballot_sync(0xffffffffu, pr);
    if (pr) out[base + __popc(b & lt_mask)] = lane + 32 * j;
    base += __popc(b);
  }
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const bool pr = key[j] == t && t != 0u;
    const uint32_t b = __ballot_sync(0xffffffffu, pr);
    const int pos = base + __popc(b & lt_mask);
    if (pr && pos < kc) out[pos] = lane + 32 * j;
    base += __popc(b);
  }
}

// ---------------------------------------------------------------------------------------------
// Large stores (m > 2048 rows; BASELINE configs[4]: 1 M-row clip store, top-10 cosine —
// ratt_chunk_retriever.py:123,144-148 pulls every candidate to the host and argpartitions).
// The score matrix [queries, m] is never materialised.  The store is walked in a few stages of
// geometrically growing size; each stage is ONE similarity GEMM whose epilogue
// (EPI_THRESH_POOL, gemm_sm100.cuh) compares every score with its query's running threshold —
// the kc-th best composite key seen so far — and appends the rare survivors to a per-query pool.
// Between stages knn_pool_select_kernel shrinks each pool back to its top kc and raises the
// threshold.  Composite key = key_bits << 32 | ~row: a strict total order (key descending, row
// ascending), so the final candidate SET is exactly the first kc rows of that order no matter
// how the atomics interleave.  A pool that overflows is detected on the host after the last
// stage and that stage is redone in quarters (a stage of <= cap - kc rows cannot overflow).
// ---------------------------------------------------------------------------------------------
constexpr int KNN_POOL_KPL = 16;     // pool entries per thread of the select kernel
constexpr int KNN_POOL_CAP = 2048;   // pool entries per query, kc <= 448 (128 select threads)
constexpr int KNN_POOL_CAP_BIG = 8192;  // larger k (n_results 100-750): 512 select threads

// Sum over the NW warps of a CTA; `buf` = NW ints of shared memory that no thread still reads
// from an earlier call (callers alternate two buffers).
template <int NW>
__device__ __forceinline__ int block_sum(int v, int* buf) {
  v = __reduce_add_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = v;
  __syncthreads();
  int t = 0;
#pragma unroll
  for (int i = 0; i < NW; ++i) t += buf[i];
  return t;
}

// One CTA (THREADS / 32 warps) per query; pool capacity = 16 * THREADS.  Loads the pool into registers (16 entries per thread), finds
// the kc-th largest composite — a 32-step bitwise binary search on the key bits (compare +
// block-wide count), and only when several rows share the kc-th key a second search on ~row
// among them — compacts the winners to the front of the pool, publishes the new threshold and,
// when `cand` is given, the candidate rows for the re-rank (-1 padded).  direct_rows >= 0: the
// pool was filled positionally by the first stage (entry i = row i of the stage) and holds
// exactly that many entries.  Rows in [ex_lo, ex_hi) are dropped first (a stage being redone
// after an overflow).  overflow[0] is set when a pool had been truncated.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
knn_pool_select_kernel(uint2* __restrict__ pool, int* __restrict__ pool_n,
                       unsigned long long* __restrict__ thresh, int nq, int kc, int direct_rows,
                       int ex_lo, int ex_hi, int32_t* __restrict__ cand, int* __restrict__ overflow) {
  constexpr int KPL = KNN_POOL_KPL;
  constexpr int NW = THREADS / 32;
  constexpr int CAP = KPL * THREADS;
  __shared__ int s_buf[2][NW];
  __shared__ uint32_t s_min[NW];
  const int qi = blockIdx.x;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  int count = direct_rows >= 0 ? direct_rows : pool_n[qi];
  if (count > CAP) {
    if (tid == 0) atomicExch(overflow, 1);
    count = CAP;
  }
  uint2* row = pool + (size_t)qi * CAP;
  uint32_t hi[KPL], lo[KPL];  // key bits | ~row; hi == 0 marks an empty slot
  // every slot of the thread is requested first (slots past `count` hold stale entries of the same
  // allocation and are ignored below): inside the `if (i < count)` the KPL loads went out one
  // memory round trip after the other
  uint2 ent[KPL];
#pragma unroll
  for (int j = 0; j < KPL; ++j) ent[j] = row[tid + THREADS * j];
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const int i = tid + THREADS * j;
    hi[j] = 0u;
    lo[j] = 0u;
    if (i < count) {
      const uint2 e = ent[j];
      const bool excluded = (int)e.x >= ex_lo && (int)e.x < ex_hi;
      if (!excluded && e.y != 0u) { hi[j] = e.y; lo[j] = ~e.x; }
    }
  }
  uint32_t t = 0u;
  int flip = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit, flip ^= 1) {
    const uint32_t tt = t | (1u << bit);
    int c = 0;
#pragma unroll
    for (int j = 0; j < KPL; ++j) c += hi[j] >= tt ? 1 : 0;
    if (block_sum<NW>(c, s_buf[flip]) >= kc) t = tt;
  }
  // t == the kc-th largest key (with multiplicity), or 0 when fewer than kc entries exist
  uint32_t t_lo = 0u;
  if (t != 0u) {  // block-uniform
    int gt = 0, eq = 0;
    uint32_t lo_min = 0xffffffffu;
#pragma unroll
    for (int j = 0; j < KPL; ++j) {
      gt += hi[j] > t ? 1 : 0;
      if (hi[j] == t) { ++eq; lo_min = min(lo_min, lo[j]); }
    }
    gt = block_sum<NW>(gt, s_buf[flip]); flip ^= 1;
    eq = block_sum<NW>(eq, s_buf[flip]); flip ^= 1;
    const int need = kc - gt;  // >= 1 rows with key == t are wanted, lowest rows first
    if (eq == need) {
      lo_min = __reduce_min_sync(0xffffffffu, lo_min);
      if (lane == 0) s_min[w] = lo_min;
      __syncthreads();
      t_lo = s_min[0];
#pragma unroll
      for (int i = 1; i < NW; ++i) t_lo = min(t_lo, s_min[i]);
    } else {
#pragma unroll 1
      for (int bit = 31; bit >= 0; --bit, flip ^= 1) {
        const uint32_t tt = t_lo | (1u << bit);
        int c = 0;
#pragma unroll
        for (int j = 0; j < KPL; ++j) c += (hi[j] == t && lo[j] >= tt) ? 1 : 0;
        if (block_sum<NW>(c, s_buf[flip]) >= need) t_lo = tt;
      }
    }
  }
  // compaction: winners of warp w go after those of warps < w (all reads of the pool are done:
  // every thread has passed at least one __syncthreads since its loads)
  int mine = 0;
#pragma unroll
  for (int j = 0; j < KPL; ++j)
    mine += (hi[j] != 0u && (hi[j] > t || (hi[j] == t && lo[j] >= t_lo))) ? 1 : 0;
  const int total = block_sum<NW>(mine, s_buf[flip]);
  int base = 0;
  for (int i = 0; i < w; ++i) base += s_buf[flip][i];
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const bool pr = hi[j] != 0u && (hi[j] > t || (hi[j] == t && lo[j] >= t_lo));
    const uint32_t b = __ballot_sync(0xffffffffu, pr);
    if (pr) {
      const int pos = base + __popc(b & lt_mask);
      const uint32_t bi = ~lo[j];
      row[pos] = make_uint2(bi, hi[j]);
      if (cand != nullptr && pos < kc) cand[(size_t)qi * kc + pos] = (int32_t)bi;
    }
    base += __popc(b);
  }
  if (cand != nullptr)
    for (int i = total + tid; i < kc; i += THREADS) cand[(size_t)qi * kc + i] = -1;
  if (tid == 0) {
    pool_n[qi] = total;
    // the old threshold may be the composite of a row that was just excluded: step below it so
    // the redo (strict >) can bring that row back
    unsigned long long told = thresh[qi];
    if (ex_hi > ex_lo && told > 0ull) told -= 1ull;
    const unsigned long long tn = (static_cast<unsigned long long>(t) << 32) | t_lo;
    thresh[qi] = tn > told ? tn : told;
  }
}

// Exact fp32 score of candidates [0, kc) of one query -> s_key / s_idx (key: larger == better;
// dropped candidates get (-inf, INT_MAX)).
//   cosine: score = <q, b>            (rows unit-normalised upstream), descending
//   L2    : score = sum (q-b)^2       (Chroma 'l2' space returns squared L2), ascending
// Each warp takes every NWARPS-th group of four candidates: the 32 lanes read the bank rows
// coalesced (float4 per lane per 128 columns), multiply against the query staged in shared memory
// and reduce with xor-shuffles (every lane ends with the same, order-deterministic sum).  Rows a
// `where=` mask or the per-query predicate rejects are dropped here, so they can never be returned.
template <int NWARPS>
__device__ __forceinline__ void knn_score_candidates(const float* q_smem, const float* __restrict__ bank,
                                                     const int32_t* __restrict__ cand_row, int qi, int m, int d,
                                                     int kc, int metric, const uint8_t* __restrict__ row_mask,
                                                     const KnnFilter& filt, float* s_key, int* s_idx) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c0 = w * 4; c0 < kc; c0 += 4 * NWARPS) {
    const float* br[4];
    int bi[4];
    float acc[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const int c = c0 + s;
      int b = c < kc ? __ldg(cand_row + c) : -1;
      if (b < 0 || b >= m) b = -1;
      if (b >= 0 && row_mask != nullptr && !__ldg(row_mask + b)) b = -1;  // filtered out (where=)
      if (b >= 0 && filt.active && !knn_filter_pass(filt, qi, b)) b = -1;  // per-query predicate
      bi[s] = b;
      br[s] = bank + (size_t)(b < 0 ? 0 : b) * d;
      acc[s] = 0.f;
    }
    // Three 128-column steps (12 float4 loads per lane) are issued before the first FMA: rolled one
    // step at a time every step waited for its own L2 / DRAM round trip (6 per candidate group at
    // d = 768) and the kernel was pure latency.  Same accumulation order as the rolled loop.
    auto step = [&](const float4& a, const float4& b, int s_) {
      if (metric == METRIC_L2) {
        const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, dw = a.w - b.w;
        acc[s_] += (dx * dx + dy * dy) + (dz * dz + dw * dw);
      } else {
        acc[s_] += (a.x * b.x + a.y * b.y) + (a.z * b.z + a.w * b.w);
      }
    };
    int i = lane * 4;
    for (; i + 256 < d; i += 384) {
      float4 a[3], b[3][4];
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        a[u] = *reinterpret_cast<const float4*>(q_smem + i + 128 * u);
#pragma unroll
        for (int s = 0; s < 4; ++s) b[u][s] = __ldg(reinterpret_cast<const float4*>(br[s] + i + 128 * u));
      }
#pragma unroll
      for (int u = 0; u < 3; ++u)
#pragma unroll
        for (int s = 0; s < 4; ++s) step(a[u], b[u][s], s);
    }
    for (; i < d; i += 128) {
      const float4 a = *reinterpret_cast<const float4*>(q_smem + i);
#pragma unroll
      for (int s = 0; s < 4; ++s) step(a, __ldg(reinterpret_cast<const float4*>(br[s] + i)), s);
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float tot = warp_sum(acc[s]);
      if (lane == 0 && c0 + s < kc) {
        s_key[c0 + s] = bi[s] < 0 ? -INFINITY : (metric == METRIC_L2 ? -tot : tot);
        s_idx[c0 + s] = bi[s] < 0 ? 0x7fffffff : bi[s];
      }
    }
  }
}

// Exact fp32 re-rank of kc <= 128 candidates, ordered top-k, vote.  One CTA (8 warps) per query:
// the four warps score the candidates, warp 0 then extracts the k best in (score, row) order and
// votes.
// Outputs: idx[nq,k], score[nq,k] in result order; counts[nq,n_classes]; label = first argmax
// of counts (np.argmax semantics, chroma.py:128: ties -> left > right > none).
constexpr int KNN_RERANK_THREADS = 256;
__global__ void __launch_bounds__(KNN_RERANK_THREADS)
knn_rerank_vote_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                       const int32_t* __restrict__ labels, const int32_t* __restrict__ cand,
                       int nq, int m, int d, int kc, int k, int metric, int n_classes,
                       int32_t* __restrict__ idx_out, float* __restrict__ score_out,
                       int32_t* __restrict__ label_out, int32_t* __restrict__ counts_out,
                       const uint8_t* __restrict__ row_mask, const KnnFilter filt) {
  extern __shared__ float knn_q_smem[];  // [d] query row
  __shared__ float s_key[KNN_MAX_CAND];
  __shared__ int s_idx[KNN_MAX_CAND];
  griddep_launch_dependents();
  griddep_wait();
  const int qi = blockIdx.x;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (qi >= nq) return;
  for (int i = threadIdx.x * 4; i < d; i += KNN_RERANK_THREADS * 4)
    *reinterpret_cast<float4*>(knn_q_smem + i) = *reinterpret_cast<const float4*>(q + (size_t)qi * d + i);
  __syncthreads();
  knn_score_candidates<KNN_RERANK_THREADS / 32>(knn_q_smem, bank, cand + (size_t)qi * kc, qi, m, d, kc, metric, row_mask, filt,
                          s_key, s_idx);
  __syncthreads();
  if (w != 0) return;
  float ck[KNN_MAX_CAND / 32];   // key (larger == better) of candidate slot lane + 32*s
  int ci[KNN_MAX_CAND / 32];
#pragma unroll
  for (int s = 0; s < KNN_MAX_CAND / 32; ++s) {
    const int c = lane + 32 * s;
    ck[s] = c < kc ? s_key[c] : -INFINITY;
    ci[s] = c < kc ? s_idx[c] : 0x7fffffff;
  }
  int cnt[KNN_MAX_CLASSES];
#pragma unroll
  for (int j = 0; j < KNN_MAX_CLASSES; ++j) cnt[j] = 0;
  for (int r = 0; r < k; ++r) {
    // lane-local best of <= 4 slots, then warp argmax
    float bk = -INFINITY;
    int bi = 0x7fffffff;
#pragma unroll
    for (int s = 0; s < KNN_MAX_CAND / 32; ++s)
      if (key_before(ck[s], ci[s], bk, bi)) { bk = ck[s]; bi = ci[s]; }
    float wk = bk;
    int wi = bi;
    warp_argbest(wk, wi);
    if (wi == 0x7fffffff) {  // fewer than k candidates: pad like an empty Chroma slot
      if (lane == 0) { idx_out[(size_t)qi * k + r] = -1; score_out[(size_t)qi * k + r] = 0.f; }
      continue;
    }
    if (wi == bi) {
#pragma unroll
      for (int s = 0; s < KNN_MAX_CAND / 32; ++s)
        if (ci[s] == wi) { ck[s] = -INFINITY; ci[s] = 0x7fffffff; }
    }
    const int lab = labels ? __ldg(labels + wi) : 0;
#pragma unroll
    for (int j = 0; j < KNN_MAX_CLASSES; ++j) cnt[j] += (lab == j) ? 1 : 0;
    if (lane == 0) {
      idx_out[(size_t)qi * k + r] = wi;
      score_out[(size_t)qi * k + r] = metric == METRIC_L2 ? -wk : wk;
    }
  }
  if (lane == 0) {
    int best = 0;
    for (int j = 0; j < n_classes; ++j) {
      if (counts_out) counts_out[(size_t)qi * n_classes + j] = cnt[j];
      if (cnt[j] > cnt[best]) best = j;
    }
    if (label_out) label_out[qi] = best;
  }
}

// In-place bitonic sort of n2 (a power of two) (key, row) pairs in shared memory into the
// (key descending, row ascending) order; all T threads of the CTA take part.
template <int T>
__device__ __forceinline__ void knn_bitonic_sort(float* s_key, int* s_idx, int n2) {
  for (int k2 = 2; k2 <= n2; k2 <<= 1) {
    for (int j = k2 >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += T) {
        const int l = i ^ j;
        if (l > i) {
          const float ka = s_key[i], kb = s_key[l];
          const int ia = s_idx[i], ib = s_idx[l];
          const bool up = (i & k2) == 0;  // this run is sorted best-first
          const bool swap = up ? key_before(kb, ib, ka, ia) : key_before(ka, ia, kb, ib);
          if (swap) { s_key[i] = kb; s_key[l] = ka; s_idx[i] = ib; s_idx[l] = ia; }
        }
      }
      __syncthreads();
    }
  }
}

// The same re-rank for the larger n_results of the clip-store queries (100-750: config.py:10-11,
// frame_retriever.py:41-43): kc up to KNN_SORT_MAX_CAND candidates are scored by the same code and
// ordered by a bitonic sort in shared memory.  One CTA (8 warps) per query.
__global__ void __launch_bounds__(256)
knn_rerank_sort_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                       const int32_t* __restrict__ labels, const int32_t* __restrict__ cand,
                       int nq, int m, int d, int kc, int k, int metric, int n_classes,
                       int32_t* __restrict__ idx_out, float* __restrict__ score_out,
                       int32_t* __restrict__ label_out, int32_t* __restrict__ counts_out,
                       const uint8_t* __restrict__ row_mask, const KnnFilter filt) {
  extern __shared__ float knn_q_smem[];  // [d] query row
  __shared__ float s_key[KNN_SORT_MAX_CAND];
  __shared__ int s_idx[KNN_SORT_MAX_CAND];
  __shared__ int s_cnt[KNN_MAX_CLASSES];
  const int qi = blockIdx.x;
  if (qi >= nq) return;
  for (int i = threadIdx.x * 4; i < d; i += 256 * 4)
    *reinterpret_cast<float4*>(knn_q_smem + i) = *reinterpret_cast<const float4*>(q + (size_t)qi * d + i);
  if (threadIdx.x < KNN_MAX_CLASSES) s_cnt[threadIdx.x] = 0;
  int n2 = 32;
  while (n2 < kc) n2 <<= 1;
  for (int i = kc + threadIdx.x; i < n2; i += 256) { s_key[i] = -INFINITY; s_idx[i] = 0x7fffffff; }
  __syncthreads();
  knn_score_candidates<8>(knn_q_smem, bank, cand + (size_t)qi * kc, qi, m, d, kc, metric, row_mask, filt,
                          s_key, s_idx);
  __syncthreads();
  knn_bitonic_sort<256>(s_key, s_idx, n2);
  for (int r = threadIdx.x; r < k; r += 256) {
    const int wi = r < n2 ? s_idx[r] : 0x7fffffff;
    if (wi == 0x7fffffff) {  // fewer than k rows in play: pad like an empty Chroma slot
      idx_out[(size_t)qi * k + r] = -1;
      score_out[(size_t)qi * k + r] = 0.f;
    } else {
      idx_out[(size_t)qi * k + r] = wi;
      score_out[(size_t)qi * k + r] = metric == METRIC_L2 ? -s_key[r] : s_key[r];
      if (labels != nullptr && n_classes > 0) atomicAdd(&s_cnt[__ldg(labels + wi)], 1);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int best = 0;
    for (int j = 0; j < n_classes; ++j) {
      if (counts_out) counts_out[(size_t)qi * n_classes + j] = s_cnt[j];
      if (s_cnt[j] > s_cnt[best]) best = j;
    }
    if (label_out) label_out[qi] = best;
  }
}

// Merge of per-shard result lists (the row-sharded clip store's exchange step):
// idx / score [nq, c] with GLOBAL row ids (-1 = empty slot) -> the k best per query in the
// kernels' total order (better score first, then the lower row id).  One CTA per query, bitonic
// sort in shared memory; c <= 4096.
__global__ void __launch_bounds__(256)
knn_merge_topk_kernel(const int32_t* __restrict__ idx_in, const float* __restrict__ score_in, int nq,
                      int c, int k, int descending, int32_t* __restrict__ idx_out,
                      float* __restrict__ score_out) {
  extern __shared__ float knn_merge_smem[];  // [n2] keys | [n2] rows
  const int qi = blockIdx.x;
  if (qi >= nq) return;
  int n2 = 32;
  while (n2 < c) n2 <<= 1;
  float* s_key = knn_merge_smem;
  int* s_idx = reinterpret_cast<int*>(knn_merge_smem + n2);
  for (int i = threadIdx.x; i < n2; i += 256) {
    int row = 0x7fffffff;
    float key = -INFINITY;
    if (i < c) {
      const int r = __ldg(idx_in + (size_t)qi * c + i);
      if (r >= 0) {
        row = r;
        const float sc = __ldg(score_in + (size_t)qi * c + i);
        key = descending ? sc : -sc;
      }
    }
    s_key[i] = key;
    s_idx[i] = row;
  }
  __syncthreads();
  knn_bitonic_sort<256>(s_key, s_idx, n2);
  for (int r = threadIdx.x; r < k; r += 256) {
    const int wi = r < n2 ? s_idx[r] : 0x7fffffff;
    idx_out[(size_t)qi * k + r] = wi == 0x7fffffff ? -1 : wi;
    score_out[(size_t)qi * k + r] = wi == 0x7fffffff ? 0.f : (descending ? s_key[r] : -s_key[r]);
  }
}

}  // namespace vitk
