// elementwise.cuh — the HBM-bound kernels between the GEMMs: im2col ("patchify"), LayerNorm,
// and the CLS tail (final LayerNorm on the CLS row + L2 normalise).
#pragma once
#include "vitk_common.cuh"

namespace vitk {

enum PixelFormat : int {
  PIX_F32_NCHW = 0,  // HF pixel_values (N,3,H,W) float32, already rescaled/normalised
  PIX_BF16_NCHW = 1, // same, bf16
  PIX_U8_NHWC = 2,   // raw frames (N,H,W,3) uint8; the kernel applies (x - 127.5) / 127.5
};

struct PatchifyParams {
  const void* pix;
  __nv_bfloat16* out;  // [n*(P+1), K]   row f*(P+1) is the all-zero CLS row
  int n, C, H, W, ps, gh, gw, K;
  int fmt;
  float mean, std;  // u8 path: v = (x - mean) / std   (HF fused rescale+normalise: 127.5, 127.5)
};

// im2col for a stride==kernel convolution (HF modeling_vit.py:151,167): K index = (c, ky, kx).
// One thread produces 8 consecutive kx of one (row, c, ky): one 16 B store.  One CTA per
// PATCHIFY_ROWS consecutive output rows: the (frame, patch) split is block-uniform 32-bit
// arithmetic, the (c, ky, kx) split of a thread's K position is done once for all of its rows,
// and the rows' loads are independent (the flat-index form spent more time in 64-bit divisions
// than moving data; one row per CTA was still bound by index arithmetic).
constexpr int PATCHIFY_ROWS = 4;   // <= tokens per frame (>= 5 for every supported geometry)
__device__ __forceinline__ uint4 patchify_load8(const PatchifyParams& p, int f, int t, int c, int ky, int kx) {
  uint4 o = make_uint4(0u, 0u, 0u, 0u);
  if (t > 0) {
    const int pidx = t - 1;
    const int py = pidx / p.gw, px = pidx % p.gw;
    const int y = py * p.ps + ky, x = px * p.ps + kx;
    float v[8];
    if (p.fmt == PIX_F32_NCHW) {
      const float* s = reinterpret_cast<const float*>(p.pix) + (((size_t)f * p.C + c) * p.H + y) * p.W + x;
      float4 a = *reinterpret_cast<const float4*>(s);
      float4 b = *reinterpret_cast<const float4*>(s + 4);
      o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
      o.z = pack_bf16x2(b.x, b.y); o.w = pack_bf16x2(b.z, b.w);
    } else if (p.fmt == PIX_BF16_NCHW) {
      const __nv_bfloat16* s = reinterpret_cast<const __nv_bfloat16*>(p.pix) +
                               (((size_t)f * p.C + c) * p.H + y) * p.W + x;
      o = *reinterpret_cast<const uint4*>(s);
    } else {
      const uint8_t* s = reinterpret_cast<const uint8_t*>(p.pix) + (((size_t)f * p.H + y) * p.W + x) * p.C + c;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        // bit-identical to the HF processor's fused float32 rescale+normalise
        // (transformers 5.5 image_processing_backends.py: tvF.normalize(x, 127.5, 127.5))
        v[j] = __fdiv_rn(__fsub_rn((float)s[j * p.C], p.mean), p.std);
      }
      o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
      o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    }
  }
  return o;
}
__global__ void __launch_bounds__(256) patchify_kernel(const PatchifyParams p) {
  griddep_launch_dependents();
  griddep_wait();   // the previous step's kernels may still read `out` (the activation buffer)
  const int kv_per_row = p.K >> 3;
  const int P1 = p.gh * p.gw + 1;
  const unsigned rows = (unsigned)p.n * (unsigned)P1;
  const unsigned row0 = blockIdx.x * (unsigned)PATCHIFY_ROWS;
  const int f0 = (int)(row0 / (unsigned)P1), t0 = (int)(row0 % (unsigned)P1);
  const int pp = p.ps * p.ps;
  for (int kv = threadIdx.x; kv < kv_per_row; kv += blockDim.x) {
    const int k0 = kv << 3;
    const int c = k0 / pp, rem = k0 % pp;
    const int ky = rem / p.ps, kx = rem % p.ps;
    uint4 o[PATCHIFY_ROWS];
#pragma unroll
    for (int r = 0; r < PATCHIFY_ROWS; ++r) {
      int t = t0 + r, f = f0;
      if (t >= P1) { t -= P1; f += 1; }
      o[r] = row0 + r < rows ? patchify_load8(p, f, t, c, ky, kx) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int r = 0; r < PATCHIFY_ROWS; ++r)
      if (row0 + r < rows)
        *reinterpret_cast<uint4*>(p.out + (size_t)(row0 + r) * p.K + ((size_t)kv << 3)) = o[r];
  }
}

// LayerNorm over the last dim (HF modeling_vit.py:325-326, nn.LayerNorm: biased variance,
// y = (x-mean)*rsqrt(var+eps)*gamma+beta).  One warp per row, row held in registers
// (two-pass mean / variance in fp32), bf16 output feeds the next GEMM's A operand.
template <int VEC>  // D = 128 * VEC
__global__ void __launch_bounds__(256)
layernorm_bf16_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                      const float* __restrict__ beta, __nv_bfloat16* __restrict__ y, int rows,
                      float eps) {
  constexpr int D = 128 * VEC;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)warp * D);
  float4 v[VEC];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    v[j] = xr[lane + 32 * j];
    s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  const float mean = warp_sum(s) * (1.0f / D);
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
    ss += (a * a + b * b) + (c * c + d * d);
  }
  const float rstd = rsqrtf(warp_sum(ss) * (1.0f / D) + eps);
  uint2* yr = reinterpret_cast<uint2*>(y + (size_t)warp * D);
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * j);
    float4 b = __ldg(reinterpret_cast<const float4*>(beta) + lane + 32 * j);
    uint2 o;
    o.x = pack_bf16x2((v[j].x - mean) * rstd * g.x + b.x, (v[j].y - mean) * rstd * g.y + b.y);
    o.y = pack_bf16x2((v[j].z - mean) * rstd * g.z + b.z, (v[j].w - mean) * rstd * g.w + b.w);
    yr[lane + 32 * j] = o;
  }
}

// Row statistics + bf16 cast: seeds the folded LayerNorm (gemm_sm100.cuh) of layer 0 from the
// patch-embedding output.  xb = bf16(x); stats[row] = `parts` (sum, sumsq) pairs, pair 0 holds
// the full-row sums and the others are zero (same layout the residual GEMM epilogue writes).
template <int VEC>  // D = 128 * VEC
__global__ void __launch_bounds__(256)
rowstats_cast_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xb,
                     float* __restrict__ stats, int rows, int parts) {
  constexpr int D = 128 * VEC;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)warp * D);
  uint2* yr = reinterpret_cast<uint2*>(xb + (size_t)warp * D);
  float s = 0.f, ss = 0.f;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    const float4 v = xr[lane + 32 * j];
    s += (v.x + v.y) + (v.z + v.w);
    ss += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    uint2 o;
    o.x = pack_bf16x2(v.x, v.y);
    o.y = pack_bf16x2(v.z, v.w);
    yr[lane + 32 * j] = o;
  }
  s = warp_sum(s);
  ss = warp_sum(ss);
  float2* sp = reinterpret_cast<float2*>(stats + (size_t)warp * parts * 2);
  if (lane < parts) sp[lane] = lane == 0 ? make_float2(s, ss) : make_float2(0.f, 0.f);
}

// CLS tail: final LayerNorm on token 0 of every frame (HF modeling_vit.py:455 then
// last_hidden_state[:,0,:], ratt_write_clips_to_ragdb.py:148) and the reference's row
// normalisation  cls / (||cls|| + 1e-8)  (ratt_write_clips_to_ragdb.py:152).
// The other 196 rows of the final LayerNorm are never read by the reference and are skipped.
template <int VEC>
__global__ void __launch_bounds__(128)
cls_tail_kernel(const float* __restrict__ resid, const __nv_bfloat16* __restrict__ resid_bf16,
                const float* __restrict__ gamma,
                const float* __restrict__ beta, float* __restrict__ out_f32,
                __nv_bfloat16* __restrict__ out_bf16, int n, int tokens, float eps, int l2norm) {
  constexpr int D = 128 * VEC;
  griddep_launch_dependents();
  griddep_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= n) return;
  float4 v[VEC];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    if (resid_bf16 != nullptr) {  // bf16 residual stream (the default): 4 elements = 8 bytes
      const uint2 u = reinterpret_cast<const uint2*>(resid_bf16 + (size_t)warp * tokens * D)[lane + 32 * j];
      v[j] = make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u),
                         __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
    } else {
      v[j] = reinterpret_cast<const float4*>(resid + (size_t)warp * tokens * D)[lane + 32 * j];
    }
    s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  const float mean = warp_sum(s) * (1.0f / D);
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
    ss += (a * a + b * b) + (c * c + d * d);
  }
  const float rstd = rsqrtf(warp_sum(ss) * (1.0f / D) + eps);
  float n2 = 0.f;
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * j);
    float4 b = __ldg(reinterpret_cast<const float4*>(beta) + lane + 32 * j);
    v[j].x = (v[j].x - mean) * rstd * g.x + b.x;
    v[j].y = (v[j].y - mean) * rstd * g.y + b.y;
    v[j].z = (v[j].z - mean) * rstd * g.z + b.z;
    v[j].w = (v[j].w - mean) * rstd * g.w + b.w;
    n2 += (v[j].x * v[j].x + v[j].y * v[j].y) + (v[j].z * v[j].z + v[j].w * v[j].w);
  }
  float inv = 1.0f;
  if (l2norm) inv = 1.0f / (sqrtf(warp_sum(n2)) + 1e-8f);
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    float4 o = make_float4(v[j].x * inv, v[j].y * inv, v[j].z * inv, v[j].w * inv);
    reinterpret_cast<float4*>(out_f32 + (size_t)warp * D)[lane + 32 * j] = o;
    if (out_bf16) {
      uint2 ob;
      ob.x = pack_bf16x2(o.x, o.y);
      ob.y = pack_bf16x2(o.z, o.w);
      reinterpret_cast<uint2*>(out_bf16 + (size_t)warp * D)[lane + 32 * j] = ob;
    }
  }
}

}  // namespace vitk
