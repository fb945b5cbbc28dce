// knn.cuh — nearest-neighbour top-k + left/right/none vote against a replicated embedding bank.
//
// Reference behaviour being replaced:
//   exact cosine top-k     nba_proj/retrieval/ratt_chunk_retriever.py:74-75,115-116,123,144-148
//   Chroma L2 query        nba_proj/chroma.py:211-214, generate_clips_hmm.py:403-406 (squared L2)
//   vote                   nba_proj/chroma.py:102-134 (counts, first-max argmax)
// Pipeline: bf16 similarity GEMM (gemm_sm100.cuh, EPI_F32) -> warp-shuffle candidate selection of
// kc >= k rows on the bf16 scores -> exact fp32 re-rank of the kc candidates -> ordered top-k ->
// class counts -> label.  The re-rank is what makes the indices match an fp32/fp64 brute force:
// bf16 similarities alone permute near-ties (SURVEY.md §7 "Top-k index exactness").
#pragma once
#include "vitk_common.cuh"

namespace vitk {

enum KnnMetric : int { METRIC_COSINE = 0, METRIC_L2 = 1 };

constexpr int KNN_MAX_CAND = 128;   // kc <= 128 (4 per lane)
constexpr int KNN_MAX_CLASSES = 8;

// f32 rows -> bf16 rows (+ optional x / (||x|| + 1e-8), the reference's query normalisation,
// ratt_chunk_retriever.py:75).  One warp per row.
__global__ void __launch_bounds__(256)
prep_queries_kernel(const float* __restrict__ q, float* __restrict__ q_out,
                    __nv_bfloat16* __restrict__ q_bf16, int n, int d, int normalise) {
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

// Order-preserving map float -> uint32 (larger float <=> larger integer); 0 is below every
// real value and marks "no row".
__device__ __forceinline__ uint32_t float_order_bits(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
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
                          int32_t* __restrict__ cand) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nq) return;
  const float* row = scores + (size_t)warp * ld;
  uint32_t key[KPL];
#pragma unroll
  for (int j = 0; j < KPL; ++j) {
    const int i = lane + 32 * j;
    float s = -INFINITY;
    if (i < m) {
      s = row[i];
      if (metric == METRIC_L2) s = 2.0f * s - __ldg(bank_sqnorm + i);
    }
    key[j] = i < m ? float_order_bits(s) : 0u;
  }
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

// Candidate selection, general path (any m).  One warp per query; lane l owns bank rows
// l, l+32, ...  Each round every lane offers its best not-yet-taken row; the warp-shuffle
// argmax picks the winner; only the winning lane rescans its rows for its next-best.
// key = score (cosine) or 2*score - ||b||^2 (L2: larger == closer).
__global__ void __launch_bounds__(128)
knn_candidates_scan_kernel(const float* __restrict__ scores, int ld, const float* __restrict__ bank_sqnorm,
                      int nq, int m, int kc, int metric, int32_t* __restrict__ cand) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nq) return;
  const float* row = scores + (size_t)warp * ld;
  auto key_of = [&](int i) -> float {
    float s = row[i];
    return metric == METRIC_L2 ? 2.0f * s - __ldg(bank_sqnorm + i) : s;
  };
  // lane-local best
  float bk = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = lane; i < m; i += 32) {
    float k = key_of(i);
    if (key_before(k, i, bk, bi)) { bk = k; bi = i; }
  }
  for (int c = 0; c < kc; ++c) {
    float wk = bk;
    int wi = bi;
    warp_argbest(wk, wi);
    if (lane == 0) cand[(size_t)warp * kc + c] = wi;
    if (wi == bi && bi != 0x7fffffff) {
      // this lane won: its next best is the best row strictly after (wk, wi) in the order
      float nk = -INFINITY;
      int ni = 0x7fffffff;
      for (int i = lane; i < m; i += 32) {
        float k = key_of(i);
        if (key_before(wk, wi, k, i) && key_before(k, i, nk, ni)) { nk = k; ni = i; }
      }
      bk = nk;
      bi = ni;
    }
  }
}

// Exact fp32 re-rank of the kc candidates, ordered top-k, vote.  One warp per query.
//   cosine: score = <q, b>            (rows unit-normalised upstream), descending
//   L2    : score = sum (q-b)^2       (Chroma 'l2' space returns squared L2), ascending
// Outputs: idx[nq,k], score[nq,k] in result order; counts[nq,n_classes]; label = first argmax
// of counts (np.argmax semantics, chroma.py:128: ties -> left > right > none).
__global__ void __launch_bounds__(128)
knn_rerank_vote_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                       const int32_t* __restrict__ labels, const int32_t* __restrict__ cand,
                       int nq, int m, int d, int kc, int k, int metric, int n_classes,
                       int32_t* __restrict__ idx_out, float* __restrict__ score_out,
                       int32_t* __restrict__ label_out, int32_t* __restrict__ counts_out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nq) return;
  // Each lane owns candidate slots lane, lane+32, ... and computes their full fp32 dot /
  // squared distance itself: no shuffles, (KNN_MAX_CAND/32) x 4 independent loads in flight per
  // lane, and the keys land exactly where the selection below expects them.  The query row is
  // staged in shared memory (broadcast reads).
  extern __shared__ float knn_q_smem[];
  float* qs = knn_q_smem + (threadIdx.x >> 5) * d;
  for (int i = lane * 4; i < d; i += 128)
    *reinterpret_cast<float4*>(qs + i) = *reinterpret_cast<const float4*>(q + (size_t)warp * d + i);
  __syncwarp();
  float ck[KNN_MAX_CAND / 32];   // key (larger == better) of candidate slot lane + 32*s
  int ci[KNN_MAX_CAND / 32];
  const float* br[KNN_MAX_CAND / 32];
  float acc[KNN_MAX_CAND / 32];
#pragma unroll
  for (int s = 0; s < KNN_MAX_CAND / 32; ++s) {
    const int c = lane + 32 * s;
    int bi = c < kc ? cand[(size_t)warp * kc + c] : -1;
    if (bi < 0 || bi >= m) bi = -1;
    ci[s] = bi < 0 ? 0x7fffffff : bi;
    br[s] = bank + (size_t)(bi < 0 ? 0 : bi) * d;
    acc[s] = 0.f;
  }
  const int nslots = (kc + 31) >> 5;  // warp-uniform
#pragma unroll 2
  for (int i = 0; i < d; i += 4) {
    const float4 a = *reinterpret_cast<const float4*>(qs + i);
#pragma unroll
    for (int s = 0; s < KNN_MAX_CAND / 32; ++s) {
      if (s < nslots) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(br[s] + i));
        if (metric == METRIC_L2) {
          const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, dw = a.w - b.w;
          acc[s] += (dx * dx + dy * dy) + (dz * dz + dw * dw);
        } else {
          acc[s] += (a.x * b.x + a.y * b.y) + (a.z * b.z + a.w * b.w);
        }
      }
    }
  }
#pragma unroll
  for (int s = 0; s < KNN_MAX_CAND / 32; ++s)
    ck[s] = ci[s] == 0x7fffffff ? -INFINITY : (metric == METRIC_L2 ? -acc[s] : acc[s]);
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
      if (lane == 0) { idx_out[(size_t)warp * k + r] = -1; score_out[(size_t)warp * k + r] = 0.f; }
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
      idx_out[(size_t)warp * k + r] = wi;
      score_out[(size_t)warp * k + r] = metric == METRIC_L2 ? -wk : wk;
    }
  }
  if (lane == 0) {
    int best = 0;
    for (int j = 0; j < n_classes; ++j) {
      if (counts_out) counts_out[(size_t)warp * n_classes + j] = cnt[j];
      if (cnt[j] > cnt[best]) best = j;
    }
    if (label_out) label_out[warp] = best;
  }
}

}  // namespace vitk
