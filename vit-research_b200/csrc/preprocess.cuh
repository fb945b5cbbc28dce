// preprocess.cuh — the step right before the hot path: raw uint8 HWC frames of any size ->
// uint8 frames at the model's resolution, bit-identical to what the reference's processor feeds
// the model.  Reference: `processor(images=frames_list, return_tensors="pt")` inside
// hf_vit_embed_batch (nba_proj/ratt_write_clips_to_ragdb.py:143-146) = ViTImageProcessor
// [3P transformers 5.5 image_processing_vit.py:20-27] -> torchvision resize (BILINEAR,
// antialias=True) ON UINT8 -> ATen's separable fixed-point kernel (the Pillow algorithm):
// horizontal pass, uint8 intermediate, vertical pass, int16 weights with a per-dimension
// precision.  The weight tables are built on the host in double exactly as ATen does
// (vitk_api.cu: build_resize_axis); this kernel only does the integer arithmetic, so the result
// is bit-exact by construction.  The fused rescale+normalise happens in patchify_kernel.
//
// HBM-bound byte work: one CTA produces TY output rows of one frame.  It runs the horizontal
// pass for the input rows those outputs need into shared memory (uint8, out_w*C bytes per
// row), then the vertical pass out of shared memory, so the intermediate image never touches
// HBM and each input byte is read from HBM once (plus the vertical-support overlap between
// neighbouring CTAs, which hits L2).
#pragma once
#include "vitk_common.cuh"

namespace vitk {

struct ResizeAxis {     // device pointers into one table blob
  const int* xmin;      // [out]
  const int* xsize;     // [out]
  const short* w;       // [out, taps]
  int taps;             // max_interp_size
  int prec;             // weights precision (bits)
  int in_size, out_size;
};

struct ResizeParams {
  const uint8_t* in;    // (n, in_h, in_w, C)
  uint8_t* out;         // (n, out_h, out_w, C)
  int n, C;
  ResizeAxis hx, vy;    // horizontal (width) and vertical (height) tables
  int ty;               // output rows per CTA
  int max_rows;         // input rows a CTA may need (shared-memory rows)
};

__device__ __forceinline__ uint8_t clip8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

__global__ void __launch_bounds__(256) resize_u8_aa_kernel(const ResizeParams p) {
  extern __shared__ uint8_t rs_smem[];  // [max_rows][out_w*C] horizontally resized rows
  const int f = blockIdx.y;
  const int y0 = blockIdx.x * p.ty;
  const int y1 = min(y0 + p.ty, p.vy.out_size);
  const int in_w = p.hx.in_size, out_w = p.hx.out_size, in_h = p.vy.in_size, out_h = p.vy.out_size;
  const int C = p.C;
  const int row_elems = out_w * C;
  // input rows [r0, r1) feed output rows [y0, y1)
  const int r0 = p.vy.xmin[y0];
  int r1 = r0;
  for (int y = y0; y < y1; ++y) r1 = max(r1, p.vy.xmin[y] + p.vy.xsize[y]);
  const int rows = min(r1 - r0, p.max_rows);
  const uint8_t* src = p.in + (size_t)f * in_h * in_w * C;
  // ---- horizontal pass (ATen: ImagingResampleHorizontal) ----
  const bool h_identity = in_w == out_w;
  for (int i = threadIdx.x; i < rows * row_elems; i += blockDim.x) {
    const int r = i / row_elems, e = i - r * row_elems;
    const int x = e / C, c = e - x * C;
    const uint8_t* line = src + (size_t)(r0 + r) * in_w * C;
    uint8_t v;
    if (h_identity) {
      v = line[e];
    } else {
      const int lo = p.hx.xmin[x], cnt = p.hx.xsize[x];
      const short* w = p.hx.w + (size_t)x * p.hx.taps;
      int acc = 1 << (p.hx.prec - 1);
      for (int j = 0; j < cnt; ++j) acc += (int)w[j] * (int)line[(lo + j) * C + c];
      v = clip8(acc >> p.hx.prec);
    }
    rs_smem[i] = v;
  }
  __syncthreads();
  // ---- vertical pass (ATen: ImagingResampleVertical) ----
  const bool v_identity = in_h == out_h;
  uint8_t* dst = p.out + (size_t)f * out_h * row_elems;
  for (int i = threadIdx.x; i < (y1 - y0) * row_elems; i += blockDim.x) {
    const int yy = i / row_elems, e = i - yy * row_elems;
    const int y = y0 + yy;
    uint8_t v;
    if (v_identity) {
      v = rs_smem[(size_t)(y - r0) * row_elems + e];
    } else {
      const int lo = p.vy.xmin[y] - r0, cnt = p.vy.xsize[y];
      const short* w = p.vy.w + (size_t)y * p.vy.taps;
      int acc = 1 << (p.vy.prec - 1);
      for (int j = 0; j < cnt; ++j) acc += (int)w[j] * (int)rs_smem[(size_t)(lo + j) * row_elems + e];
      v = clip8(acc >> p.vy.prec);
    }
    dst[(size_t)y * row_elems + e] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// cv2.resize(frame, (768, 432), interpolation=cv2.INTER_AREA) -- the TF path's front-end
// (generate_clips_hmm.py:392, chroma.py:204-205, write_embeddings.py:225-226).  OpenCV [3P]
// imgproc/resize.cpp, shrinking case: per axis a table of (source index, weight) entries, weight =
// covered fraction of the source pixel / cell width (computeResizeAreaTab, built on the host in
// double); per destination row the source rows are reduced horizontally into a float buffer
// (buf += S * alpha), accumulated vertically (sum = beta * buf, then sum += beta * buf), and
// written with saturate_cast<uchar> (round half to even).  Multiplies and adds are kept separate
// (the C++ loop is not FMA-contracted in OpenCV's baseline x86 build).  When both scale factors
// are integers OpenCV takes ResizeAreaFast instead: integer box sum, then sum * (1.f / area) --
// except at exactly 2 x 2, where its vector form computes (a + b + c + d + 2) >> 2.
// Pinned byte for byte against cv2 4.13 (tests/golden/preprocess_cv2_golden.npz).
// ---------------------------------------------------------------------------------------------
struct AreaAxis {
  const int* off;      // [out + 1] first table entry of each destination index
  const int* si;       // source index of every entry
  const float* alpha;  // its weight
};
struct AreaParams {
  const uint8_t* in;   // [n, in_h, in_w, C]
  uint8_t* out;        // [n, out_h, out_w, C]
  AreaAxis x, y;
  int n, C, in_h, in_w, out_h, out_w;
  int fast_sx, fast_sy;  // > 0: both scales are integers (ResizeAreaFast)
  float fast_scale;      // 1.f / (fast_sx * fast_sy)
};

__global__ void __launch_bounds__(256) resize_area_u8_kernel(const AreaParams p) {
  const int dy = blockIdx.x, f = blockIdx.y;
  const int C = p.C, row_elems = p.out_w * C;
  const uint8_t* src = p.in + (size_t)f * p.in_h * p.in_w * C;
  uint8_t* dst = p.out + ((size_t)f * p.out_h + dy) * row_elems;
  for (int e = threadIdx.x; e < row_elems; e += blockDim.x) {
    const int dx = e / C, c = e - dx * C;
    float sum = 0.f;
    if (p.fast_sx > 0) {
      int acc = 0;
      for (int yy = 0; yy < p.fast_sy; ++yy) {
        const uint8_t* srow = src + ((size_t)(dy * p.fast_sy + yy) * p.in_w + (size_t)dx * p.fast_sx) * C + c;
        for (int xx = 0; xx < p.fast_sx; ++xx) acc += srow[xx * C];
      }
      if (p.fast_sx == 2 && p.fast_sy == 2) {
        // OpenCV's ResizeAreaFastVec: exactly 2 x 2 is the INTEGER mean, round half up
        dst[e] = (uint8_t)((acc + 2) >> 2);
        continue;
      }
      sum = __fmul_rn((float)acc, p.fast_scale);
    } else {
      bool first = true;
      for (int yi = p.y.off[dy]; yi < p.y.off[dy + 1]; ++yi) {
        const uint8_t* srow = src + (size_t)p.y.si[yi] * p.in_w * C + c;
        float buf = 0.f;
        for (int xi = p.x.off[dx]; xi < p.x.off[dx + 1]; ++xi)
          buf = __fadd_rn(buf, __fmul_rn((float)srow[p.x.si[xi] * C], p.x.alpha[xi]));
        const float t = __fmul_rn(p.y.alpha[yi], buf);
        sum = first ? t : __fadd_rn(sum, t);
        first = false;
      }
    }
    int r = __float2int_rn(sum);
    r = r < 0 ? 0 : (r > 255 ? 255 : r);
    dst[e] = (uint8_t)r;
  }
}

// ---------------------------------------------------------------------------------------------
// cv2.resize(im, (768, 432)) with the default interpolation, INTER_LINEAR -- the resize
// extract_embeddings applies before the TF model (smarter_generate_clips.py:81).  OpenCV's 8-bit
// bilinear path: two taps per axis (no antialiasing), short coefficients scaled by 2^11 built on
// the host exactly as OpenCV builds them (float arithmetic, cvRound), horizontal pass in int,
// vertical pass ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2.  HBM-bound byte work:
// one CTA per destination row, a thread per destination byte gathers its 2 x 2 source bytes (the
// two source rows of a destination row are read once per CTA and stay in L1).
// ---------------------------------------------------------------------------------------------
struct LinearParams {
  const uint8_t* in;   // [n, in_h, in_w, C]
  uint8_t* out;        // [n, out_h, out_w, C]
  const int* xs;       // [out_w] first source column (fraction already zeroed at the borders)
  const int* xa;       // [out_w][2] coefficients
  const int* ys;       // [out_h] first source row (may be -1 / in_h - 1: rows are clamped)
  const int* ya;       // [out_h][2]
  int n, C, in_h, in_w, out_h, out_w;
};

__global__ void __launch_bounds__(256) resize_linear_u8_kernel(const LinearParams p) {
  const int dy = blockIdx.x, f = blockIdx.y;
  const int C = p.C, row_elems = p.out_w * C;
  const uint8_t* src = p.in + (size_t)f * p.in_h * p.in_w * C;
  uint8_t* dst = p.out + ((size_t)f * p.out_h + dy) * row_elems;
  const int sy = p.ys[dy];
  const int y0 = min(max(sy, 0), p.in_h - 1), y1 = min(max(sy + 1, 0), p.in_h - 1);
  const int b0 = p.ya[2 * dy], b1 = p.ya[2 * dy + 1];
  const uint8_t* r0 = src + (size_t)y0 * p.in_w * C;
  const uint8_t* r1 = src + (size_t)y1 * p.in_w * C;
  for (int e = threadIdx.x; e < row_elems; e += blockDim.x) {
    const int dx = e / C, c = e - dx * C;
    const int x0 = p.xs[dx], x1 = min(x0 + 1, p.in_w - 1);
    const int a0 = p.xa[2 * dx], a1 = p.xa[2 * dx + 1];
    const int s0 = (int)r0[x0 * C + c] * a0 + (int)r0[x1 * C + c] * a1;
    const int s1 = (int)r1[x0 * C + c] * a0 + (int)r1[x1 * C + c] * a1;
    int v = (((b0 * (s0 >> 4)) >> 16) + ((b1 * (s1 >> 4)) >> 16) + 2) >> 2;
    v = v < 0 ? 0 : (v > 255 ? 255 : v);
    dst[e] = (uint8_t)v;
  }
}

}  // namespace vitk
