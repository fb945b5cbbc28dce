// vitk — ChunkEncoder pieces: the temporal transformer that consumes the frame
// embeddings (nba_proj/models/chunk_encoder.py:9-97: CLS + T <= 24 frame tokens, pre-LN blocks with
// 8 heads of 96, LayerNorm, CLS row out, Dense(256, relu) -> Dense(1) proxy logit).
// Its linear layers run on the tcgen05 GEMM of gemm_sm100.cuh; what is here is the part that GEMM
// cannot do: attention over a sequence of at most 64 tokens with a head width that is not 64
// (CUDA-core math: 25 x 25 x 96 per head is ~0.24 MFLOP), an fp32 LayerNorm of strided rows (the
// CLS rows), and the two-layer head.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "vitk_common.cuh"

namespace vitk {

constexpr int SATT_MAX_S = 64;    // tokens per sequence (two keys per lane)
constexpr int SATT_MAX_DH = 128;  // head width, a multiple of 32

// One CTA (4 warps) per (sequence, head).  qkv [n_seq*S, 3*H*dh] bf16 as the fused projection
// leaves it (q | k | v, head-major inside each third); out [n_seq*S, H*dh] bf16.
// Keras MultiHeadAttention (the block's attention layer): scores = (q * dh^-1/2) k^T, softmax over
// the keys, out = probs v.  probs_out (optional) [n_seq, H, S, S] f32 = the attention scores the
// block returns (chunk_encoder.py:29 return_attention_scores=True).
__global__ void __launch_bounds__(128)
small_attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                       float* __restrict__ probs_out, int S, int heads, int dh, float scale) {
  extern __shared__ float satt_smem[];  // K [S][dh+1] | V [S][dh] | Q [S][dh]
  const int seq = blockIdx.x / heads, head = blockIdx.x % heads;
  const int D = heads * dh;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* sK = satt_smem;
  float* sV = sK + S * (dh + 1);
  float* sQ = sV + S * dh;
  const __nv_bfloat16* base = qkv + (size_t)seq * S * 3 * D + head * dh;
  for (int i = threadIdx.x; i < S * dh; i += 128) {
    const int t = i / dh, c = i % dh;
    const __nv_bfloat16* row = base + (size_t)t * 3 * D + c;
    sQ[t * dh + c] = __bfloat162float(row[0]) * scale;
    sK[t * (dh + 1) + c] = __bfloat162float(row[D]);
    sV[t * dh + c] = __bfloat162float(row[2 * D]);
  }
  __syncthreads();
  for (int qi = warp; qi < S; qi += 4) {
    // lane j scores keys j and j + 32
    float s0 = -INFINITY, s1 = -INFINITY;
    if (lane < S) {
      float a = 0.f;
      for (int c = 0; c < dh; ++c) a = fmaf(sQ[qi * dh + c], sK[lane * (dh + 1) + c], a);
      s0 = a;
    }
    if (lane + 32 < S) {
      float a = 0.f;
      for (int c = 0; c < dh; ++c) a = fmaf(sQ[qi * dh + c], sK[(lane + 32) * (dh + 1) + c], a);
      s1 = a;
    }
    float mx = fmaxf(s0, s1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float e0 = lane < S ? expf(s0 - mx) : 0.f;
    const float e1 = lane + 32 < S ? expf(s1 - mx) : 0.f;
    const float inv = 1.0f / warp_sum(e0 + e1);
    const float p0 = e0 * inv, p1 = e1 * inv;
    if (probs_out != nullptr) {
      float* pr = probs_out + (((size_t)seq * heads + head) * S + qi) * S;
      if (lane < S) pr[lane] = p0;
      if (lane + 32 < S) pr[lane + 32] = p1;
    }
    // out[c] = sum_j p_j V[j][c]; lane owns columns lane, lane+32, ...
    float acc[SATT_MAX_DH / 32];
#pragma unroll
    for (int u = 0; u < SATT_MAX_DH / 32; ++u) acc[u] = 0.f;
    for (int j = 0; j < S; ++j) {
      const float pj = __shfl_sync(0xffffffffu, j < 32 ? p0 : p1, j & 31);
#pragma unroll
      for (int u = 0; u < SATT_MAX_DH / 32; ++u)
        if (lane + 32 * u < dh) acc[u] = fmaf(pj, sV[j * dh + lane + 32 * u], acc[u]);
    }
    __nv_bfloat16* orow = out + ((size_t)seq * S + qi) * D + head * dh;
#pragma unroll
    for (int u = 0; u < SATT_MAX_DH / 32; ++u)
      if (lane + 32 * u < dh) orow[lane + 32 * u] = __float2bfloat16_rn(acc[u]);
  }
}

// fp32 LayerNorm of `rows` rows of width d that sit `row_stride` floats apart (the CLS row of every
// sequence: chunk_encoder.py:83-86 norm(x)[:, 0, :]); y [rows, d] f32, dense.  One warp per row,
// two-pass variance like tf.keras LayerNormalization.
__global__ void __launch_bounds__(256)
layernorm_rows_f32_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                          const float* __restrict__ beta, float* __restrict__ y, int rows,
                          long long row_stride, int d, float eps) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* r = x + (size_t)row * row_stride;
  float s = 0.f;
  for (int i = lane; i < d; i += 32) s += r[i];
  const float mean = warp_sum(s) / d;
  float ss = 0.f;
  for (int i = lane; i < d; i += 32) { const float t = r[i] - mean; ss = fmaf(t, t, ss); }
  const float rstd = rsqrtf(warp_sum(ss) / d + eps);
  for (int i = lane; i < d; i += 32) y[(size_t)row * d + i] = (r[i] - mean) * rstd * gamma[i] + beta[i];
}

// class_head (chunk_encoder.py:34-38): logit = w2 . relu(W1 x + b1) + b2, dropout = identity at
// inference.  One CTA per chunk; W1 [hid, d] row-major f32, one warp per hidden unit in turn.
__global__ void __launch_bounds__(256)
chunk_head_kernel(const float* __restrict__ x, const float* __restrict__ w1, const float* __restrict__ b1,
                  const float* __restrict__ w2, const float* __restrict__ b2, float* __restrict__ logit,
                  int d, int hid) {
  extern __shared__ float head_smem[];  // x [d] | partial [8]
  float* sx = head_smem;
  float* part = head_smem + d;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < d; i += 256) sx[i] = x[(size_t)blockIdx.x * d + i];
  __syncthreads();
  float acc = 0.f;
  for (int h = warp; h < hid; h += 8) {
    float a = 0.f;
    for (int i = lane; i < d; i += 32) a = fmaf(w1[(size_t)h * d + i], sx[i], a);
    a = warp_sum(a) + b1[h];
    acc = fmaf(w2[h], fmaxf(a, 0.f), acc);
  }
  if (lane == 0) part[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = b2[0];
    for (int w = 0; w < 8; ++w) t += part[w];
    logit[blockIdx.x] = t;
  }
}

}  // namespace vitk
