// vitk_api.cu — C ABI (include/vitk.h) over the sm_100a kernels: handle, weights, workspace,
// TMA tensor maps, launch orchestration of the ViT forward and the k-NN vote.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "../../include/vitk.h"
#include "attention_sm100.cuh"
#include "elementwise.cuh"
#include "gemm_sm100.cuh"
#include "knn.cuh"
#include "hmm.cuh"
#include "preprocess.cuh"
#include "chunk.cuh"

using namespace vitk;

namespace {

thread_local std::string g_create_error;

struct Err {
  std::string msg;
  void set(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    msg = buf;
  }
};

thread_local Err g_op_err;

#define CK(e, call)                                                                        \
  do {                                                                                     \
    cudaError_t _s = (call);                                                               \
    if (_s != cudaSuccess) {                                                               \
      (e).set("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_s));        \
      return 1;                                                                            \
    }                                                                                      \
  } while (0)

// ------------------------------------------------------------------------------------------
// TMA tensor maps (driver entry point resolved at run time: the .so links only libcudart)
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// rank-2 or rank-3 row-major tensor, innermost dim contiguous, 128B swizzle, zero OOB fill.
// dims/box are innermost-first. strides_bytes has rank-1 entries (dims 1..rank-1).
int make_tmap(Err& e, CUtensorMap* m, CUtensorMapDataType dt, int rank, void* base,
              const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    e.set("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return 1;
  }
  cuuint64_t gd[5];
  cuuint64_t gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
  CUresult r = fn(m, dt, (cuuint32_t)rank, base, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    e.set("cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu] box [%u,%u,%u]",
          (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
          (unsigned long long)(rank > 2 ? dims[2] : 0), box[0], rank > 1 ? box[1] : 0,
          rank > 2 ? box[2] : 0);
    return 1;
  }
  return 0;
}

// [rows, cols] bf16 row-major operand (A or W): box {64, box_rows}
int tmap_operand(Err& e, CUtensorMap* m, const void* base, int rows, int cols, int box_rows,
                 long long pitch_elems = 0) {
  uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
  uint64_t str[1] = {(uint64_t)(pitch_elems ? pitch_elems : cols) * 2};
  uint32_t box[2] = {64u, (uint32_t)box_rows};
  return make_tmap(e, m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, str,
                   box);
}
// [rows, cols] output: 32-row boxes, 128 B wide (64 bf16 / 32 f32)
int tmap_output(Err& e, CUtensorMap* m, void* base, int rows, int cols, bool bf16,
                long long pitch_elems = 0) {
  uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
  uint64_t str[1] = {(uint64_t)(pitch_elems ? pitch_elems : cols) * (bf16 ? 2u : 4u)};
  uint32_t box[2] = {bf16 ? 64u : 32u, 32u};
  return make_tmap(e, m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                   2, base, dims, str, box);
}
// [frames, S, cols] bf16: box {64, box_rows, 1}
int tmap_tokens(Err& e, CUtensorMap* m, const void* base, int frames, int S, int cols,
                int box_rows) {
  uint64_t dims[3] = {(uint64_t)cols, (uint64_t)S, (uint64_t)frames};
  uint64_t str[2] = {(uint64_t)cols * 2, (uint64_t)S * cols * 2};
  uint32_t box[3] = {64u, (uint32_t)box_rows, 1u};
  return make_tmap(e, m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, str,
                   box);
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
int g_num_sms = 0;
int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

// Programmatic dependent launch for the GEMM and attention kernels of the forward (they call
// griddepcontrol.wait before their first global access): the next kernel's CTAs take over each SM as
// the previous kernel's leave it.  VITK_PDL=0 restores plain stream order (the parity tests run both).
bool pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* v = getenv("VITK_PDL"); on = (v && atoi(v) == 0) ? 0 : 1; }
  return on != 0;
}

// <<<grid, block, smem, st>>> with the programmatic-dependent-launch attribute (kernels that call
// griddep_launch_dependents() / griddep_wait() at their top: the small kernels around the forward)
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

template <int BN, int EPI, int CG, bool S6 = false>
int launch_gemm_t(Err& e, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c,
                  const CUtensorMap& r, const GemmParams& p, cudaStream_t st) {
  auto kern = gemm_bf16_tn_kernel<BN, EPI, CG, S6>;
  using Cfg = GemmCfg<BN, CG, S6>;
  static bool attr_set = false;
  if (!attr_set) {
    CK(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  const int m_tiles = (p.M + GEMM_BM - 1) / GEMM_BM, n_tiles = (p.N + BN - 1) / BN;
  const int units = ((m_tiles + CG - 1) / CG) * n_tiles;
  const int slots = num_sms() / CG;  // CTA pairs occupy the two SMs of a TPC
  const int grid = CG * (units < slots ? units : slots);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(GEMM_THREADS, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  CK(e, cudaLaunchKernelEx(&cfg, kern, a, b, c, r, p));
  return 0;
}

// 0 = choose (CTA pair whenever there are at least two 128-row blocks), 1 / 2 = forced
// (VITK_GEMM_CG environment variable; the parity tests run both).
int gemm_cta_group(int M) {
  static int forced = -1;
  if (forced < 0) {
    const char* v = getenv("VITK_GEMM_CG");
    forced = v ? atoi(v) : 0;
    if (forced != 1 && forced != 2) forced = 0;
  }
  if (forced) return forced;
  return M > GEMM_BM ? 2 : 1;
}

// Which CTA-pair GEMMs take the six-stage / single-staging-tile form: bit 0 = bf16-output
// epilogues (qkv, fc1), bit 1 = residual epilogue with K > 1024 (fc2).  Timed alone at the bench
// shape it is 2.4-4.6 % faster on all three (qkv 126 -> 122 us, fc1 180 -> 176, fc2 181 -> 173);
// the short-K out-projection is bound by its residual stream and keeps two staging tiles.
// VITK_GEMM_S6 overrides the mask (the parity tests run both forms).
int gemm_s6_mask() {
  static int mask = -1;
  if (mask < 0) {
    const char* v = getenv("VITK_GEMM_S6");
    mask = v ? atoi(v) & 3 : 3;
  }
  return mask;
}

template <int EPI>
int launch_gemm_e(Err& e, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c,
                  const CUtensorMap& r, const GemmParams& p, cudaStream_t st) {
  if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_GELU_BF16 || EPI == EPI_BIAS_RESID_F32 ||
                EPI == EPI_BIAS_RESID_BF16) {
    const int want = (EPI == EPI_BIAS_RESID_F32 || EPI == EPI_BIAS_RESID_BF16)
                         ? ((gemm_s6_mask() & 2) && p.K > 1024) : (gemm_s6_mask() & 1);
    if (want && gemm_cta_group(p.M) == 2) return launch_gemm_t<256, EPI, 2, true>(e, a, b, c, r, p, st);
  }
  return gemm_cta_group(p.M) == 2 ? launch_gemm_t<256, EPI, 2>(e, a, b, c, r, p, st)
                                  : launch_gemm_t<256, EPI, 1>(e, a, b, c, r, p, st);
}

// `r` is the tensor map the residual epilogue loads from (pass `c` when the residual is
// updated in place or for the epilogues that do not read one).
int launch_gemm(Err& e, int epi, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c,
                const CUtensorMap& r, const GemmParams& p, cudaStream_t st) {
  if (p.K % GEMM_BK != 0 || p.K <= 0) {
    e.set("gemm: K=%d must be a positive multiple of %d", p.K, GEMM_BK);
    return 1;
  }
  if (p.stats != nullptr && (p.c1 == nullptr || p.stats_parts < 2 || p.stats_parts % 2 ||
                             p.stats_parts > GEMM_MAX_STAT_PARTS || p.stats_stride % 4)) {
    e.set("gemm: LayerNorm fold needs c1, an even stats_parts in [2,%d] and 16-byte aligned rows",
          GEMM_MAX_STAT_PARTS);
    return 1;
  }
  if (epi == EPI_BIAS_RESID_F32 && p.xb != nullptr &&
      (p.N % 32 != 0 || p.xb_ld % 16 != 0 || (reinterpret_cast<uintptr_t>(p.xb) & 31) != 0)) {
    e.set("gemm: bf16 residual copy needs N %% 32 == 0, a 32-byte aligned base and row pitch");
    return 1;
  }
  switch (epi) {
    case EPI_BIAS_BF16: return launch_gemm_e<EPI_BIAS_BF16>(e, a, b, c, r, p, st);
    case EPI_BIAS_GELU_BF16: return launch_gemm_e<EPI_BIAS_GELU_BF16>(e, a, b, c, r, p, st);
    case EPI_BIAS_GELUT_BF16: return launch_gemm_e<EPI_BIAS_GELUT_BF16>(e, a, b, c, r, p, st);
    case EPI_BIAS_RESID_F32: return launch_gemm_e<EPI_BIAS_RESID_F32>(e, a, b, c, r, p, st);
    case EPI_TABLE_F32: return launch_gemm_e<EPI_TABLE_F32>(e, a, b, c, r, p, st);
    case EPI_F32: return launch_gemm_e<EPI_F32>(e, a, b, c, r, p, st);
    case EPI_THRESH_POOL: return launch_gemm_e<EPI_THRESH_POOL>(e, a, b, c, r, p, st);
    case EPI_THRESH_POOL_FILT: return launch_gemm_e<EPI_THRESH_POOL_FILT>(e, a, b, c, r, p, st);
    case EPI_BIAS_RESID_BF16: return launch_gemm_e<EPI_BIAS_RESID_BF16>(e, a, b, c, r, p, st);
    case EPI_TABLE_BF16: return launch_gemm_e<EPI_TABLE_BF16>(e, a, b, c, r, p, st);
  }
  e.set("gemm: unknown epilogue %d", epi);
  return 1;
}

long long* g_attn_dbg = nullptr;  // developer timeline buffer (vitk_dev_attn_timeline)
long long* g_gemm_dbg = nullptr;  // same for the GEMM of vitk_op_gemm* (vitk_dev_gemm_timeline)

// q_rows > 0: only the first q_rows queries of every frame are needed (last layer: the CLS row)
int launch_attention(Err& e, const CUtensorMap& tq, const CUtensorMap& tkv, const CUtensorMap& to,
                     int frames, int S, int heads, cudaStream_t st, int q_rows = 0, int reverse = 0) {
  if (S < 1 || S > ATT_MAX_S) {
    e.set("attention: S=%d unsupported (1..%d)", S, ATT_MAX_S);
    return 1;
  }
  static int version = -1;  // VITK_ATTN=2: the simpler non-persistent v2 kernel; 3: v3; 5: the loader / math split
  if (version < 0) {
    const char* v = getenv("VITK_ATTN");
    version = v ? atoi(v) : 3;
    if (version != 2 && version != 3 && version != 5 && version != 6) version = 3;
  }
  AttnParams p;
  p.S = S;
  p.heads = heads;
  p.frames = frames;
  p.q_rows = q_rows > 0 && q_rows < S ? q_rows : S;
  p.nblk = (p.q_rows + 127) / 128;
  p.reverse = reverse;
  {
    static int handover = -1;  // VITK_ATTN_HANDOVER: chunk after which the exp2 turn is passed on (default 4 of 6.5;
                               // 1000 = at the end of the pass, the round-1 behaviour)
    if (handover < 0) { const char* hv = getenv("VITK_ATTN_HANDOVER"); handover = hv ? atoi(hv) : 4; if (handover < 0) handover = 0; }
    p.handover = handover;
    static int early_pv = -1;  // VITK_ATTN_EARLY_PV=0: all thirteen P.V instructions after the softmax pass (the in-order pass)
    if (early_pv < 0) { const char* ev = getenv("VITK_ATTN_EARLY_PV"); early_pv = (ev && atoi(ev) == 0) ? 0 : 1; }
    p.early_pv = early_pv;
    static int one_issuer = -1;  // VITK_ATTN_ISSUER=1: the round-1 single MMA-issuing thread
    if (one_issuer < 0) { const char* iv = getenv("VITK_ATTN_ISSUER"); one_issuer = (iv && atoi(iv) == 1) ? 1 : 0; }
    p.one_issuer = one_issuer;
    if (one_issuer) p.early_pv = 0;
  }
  p.scale_log2e = 0.125f * 1.4426950408889634f;  // head dim 64 -> 1/sqrt(64)
  p.dbg = g_attn_dbg;
  if (S > 256) {
    // the TF-path geometry (313 tokens): one CTA per (frame, head, 128-query block), all 512 TMEM
    // columns, keys in two halves
    const AttnSmemLayout L = attn_smem_layout(S);
    static int attr_bytes_w = 0;
    if (L.total > attr_bytes_w) {
      CK(e, cudaFuncSetAttribute(attention_v2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 L.total));
      attr_bytes_w = L.total;
    }
    attention_v2_kernel<true><<<frames * heads * p.nblk, ATT_THREADS, L.total, st>>>(tq, tkv, to, p);
  } else if (version == 2) {
    const AttnSmemLayout L = attn_smem_layout(S);
    static int attr_bytes = 0;
    if (L.total > attr_bytes) {
      CK(e, cudaFuncSetAttribute(attention_v2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 L.total));
      attr_bytes = L.total;
    }
    attention_v2_kernel<false><<<frames * heads * p.nblk, ATT_THREADS, L.total, st>>>(tq, tkv, to, p);
  } else if (version == 6 && ((S + 15) & ~15) <= ATT5_MAX_SK) {
    const Attn6SmemLayout L = attn6_smem_layout(S);
    static int attr_bytes6 = 0;
    if (L.total > attr_bytes6) {
      CK(e, cudaFuncSetAttribute(attention_smemp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 L.total));
      attr_bytes6 = L.total;
    }
    const int units = frames * heads;
    const int grid = units < num_sms() ? units : num_sms();
    attention_smemp_kernel<<<grid, ATT3_THREADS, L.total, st>>>(tq, tkv, to, p);
  } else if (version == 5 && ((S + 15) & ~15) <= ATT5_MAX_SK) {
    const Attn5SmemLayout L = attn5_smem_layout(S);
    static int attr_bytes5 = 0;
    if (L.total > attr_bytes5) {
      CK(e, cudaFuncSetAttribute(attention_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 L.total));
      attr_bytes5 = L.total;
    }
    const int units = frames * heads;
    const int grid = units < num_sms() ? units : num_sms();
    attention_split_kernel<<<grid, ATT3_THREADS, L.total, st>>>(tq, tkv, to, p);
  } else {
    const Attn3SmemLayout L = attn3_smem_layout(S);
    static int attr_bytes3 = 0;
    if (L.total > attr_bytes3) {
      CK(e, cudaFuncSetAttribute(attention_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
      CK(e, cudaFuncSetAttribute(attention_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
      attr_bytes3 = L.total;
    }
    const int units = frames * heads;
    const int grid = units < num_sms() ? units : num_sms();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid, 1, 1);
    cfg.blockDim = dim3(ATT3_THREADS, 1, 1);
    cfg.dynamicSmemBytes = (size_t)L.total;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    if (p.early_pv && L.sk == 208) CK(e, cudaLaunchKernelEx(&cfg, attention_kernel<true>, tq, tkv, to, p));
    else CK(e, cudaLaunchKernelEx(&cfg, attention_kernel<false>, tq, tkv, to, p));
  }
  CK(e, cudaGetLastError());
  return 0;
}

int launch_layernorm(Err& e, const float* x, const float* g, const float* b, __nv_bfloat16* y,
                     int rows, int d, float eps, cudaStream_t st) {
  const int blocks = (rows + 7) / 8;
  if (d == 768)
    layernorm_bf16_kernel<6><<<blocks, 256, 0, st>>>(x, g, b, y, rows, eps);
  else if (d == 1024)
    layernorm_bf16_kernel<8><<<blocks, 256, 0, st>>>(x, g, b, y, rows, eps);
  else {
    e.set("layernorm: hidden=%d unsupported (768 or 1024)", d);
    return 1;
  }
  CK(e, cudaGetLastError());
  return 0;
}

int launch_rowstats(Err& e, const float* x, __nv_bfloat16* xb, float* stats, int rows, int d,
                    int parts, cudaStream_t st) {
  if (parts < 1 || parts > 32) {
    e.set("rowstats: parts=%d outside [1,32]", parts);
    return 1;
  }
  const int blocks = (rows + 7) / 8;
  if (d == 768)
    rowstats_cast_kernel<6><<<blocks, 256, 0, st>>>(x, xb, stats, rows, parts);
  else if (d == 1024)
    rowstats_cast_kernel<8><<<blocks, 256, 0, st>>>(x, xb, stats, rows, parts);
  else {
    e.set("rowstats: hidden=%d unsupported (768 or 1024)", d);
    return 1;
  }
  CK(e, cudaGetLastError());
  return 0;
}

int launch_patchify(Err& e, const void* pix, int fmt, __nv_bfloat16* out, int n, int C, int H,
                    int W, int ps, cudaStream_t st, float u8_mean = 127.5f, float u8_std = 127.5f) {
  // stride == kernel convolution with 'valid' padding: trailing rows / columns that do not fill
  // a patch are ignored (432 rows at patch 32 -> 13 patch rows, like Conv2d / Conv2D 'valid')
  if (ps % 8 != 0 || H < ps || W < ps || W % 8 != 0) {
    e.set("patchify: patch %d must be a multiple of 8 and fit %dx%d (width a multiple of 8)", ps, H, W);
    return 1;
  }
  if (fmt < 0 || fmt > 2) {
    e.set("patchify: unknown pixel format %d", fmt);
    return 1;
  }
  PatchifyParams p;
  p.pix = pix;
  p.out = out;
  p.n = n; p.C = C; p.H = H; p.W = W; p.ps = ps;
  p.gh = H / ps; p.gw = W / ps;
  p.K = C * ps * ps;
  p.fmt = fmt;
  p.mean = u8_mean;  // default: HF ViTImageProcessor (rescale 1/255, mean .5, std .5, fused)
  p.std = u8_std;
  const long long rows = (long long)n * (p.gh * p.gw + 1);   // one CTA per PATCHIFY_ROWS output rows
  if (rows > 0x7fffffffLL) { e.set("patchify: too many rows"); return 1; }
  const int kv_per_row = p.K / 8;
  const int threads = kv_per_row >= 256 ? 256 : ((kv_per_row + 31) / 32) * 32;
  CK(e, launch_pdl(patchify_kernel, dim3((unsigned)((rows + PATCHIFY_ROWS - 1) / PATCHIFY_ROWS)), dim3(threads), 0, st, p));
  CK(e, cudaGetLastError());
  return 0;
}

uint16_t f32_to_bf16_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);  // round to nearest even
  return (uint16_t)(u >> 16);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// handle
// ------------------------------------------------------------------------------------------
struct LayerW {
  // w_qkv / w_fc1 hold the LayerNorm-folded weights gamma∘W (bf16); b_qkv / b_fc1 hold
  // c2[n] = b[n] + sum_k beta[k] W[n,k]; c1_* hold sum_k bf16(gamma[k] W[n,k]).
  __nv_bfloat16 *w_qkv = nullptr, *w_o = nullptr, *w_fc1 = nullptr, *w_fc2 = nullptr;
  float *b_qkv = nullptr, *b_o = nullptr, *b_fc1 = nullptr, *b_fc2 = nullptr;
  float *c1_qkv = nullptr, *c1_fc1 = nullptr;
  CUtensorMap tm_qkv, tm_o, tm_fc1, tm_fc2;
  CUtensorMap tm_q, tm_kv;   // rows [0,D) and [D,3D) of w_qkv (last layer: Q for the CLS rows only)
  // host fp32 copies kept from vitk_set_weight until vitk_finalize_weights folds them
  std::vector<float> h_wqkv, h_bqkv, h_wfc1, h_bfc1, h_ln1g, h_ln1b, h_ln2g, h_ln2b;
  bool dirty = true;
};

struct BatchMaps {  // tensor maps that depend on the number of frames in the call
  CUtensorMap a_patch, a_act, a_h, a_xb;    // GEMM A operands
  CUtensorMap c_resid, c_qkv, c_h;          // GEMM outputs
  CUtensorMap att_q, att_kv, att_o;         // attention
  // last layer, CLS rows only (row f of these maps = token 0 of frame f: pitch S*D)
  CUtensorMap a_act_cls, a_xb_cls, a_h_cls, c_resid_cls, c_h_cls;
  // bf16 residual stream (default): the residual GEMMs update xb in place
  CUtensorMap c_xb, c_xb_cls;
  CUtensorMap c_kv, c_q_cls;                // last layer: K|V columns of every token, Q of the CLS rows
  CUtensorMap c_patch;                      // fp32 patch-embedding output (scratch in hbuf)
};

struct vitk_ctx {
  vitk_config cfg;
  int device = 0;
  int S = 0, P = 0, Kp = 0, D = 0, F = 0;
  Err err;
  std::vector<LayerW> layers;
  __nv_bfloat16* w_patch = nullptr;
  CUtensorMap tm_patch;
  float *posb = nullptr, *lnf_g = nullptr, *lnf_b = nullptr;
  std::vector<float> h_cls, h_pos, h_pbias;
  std::set<std::string> have;
  bool finalized = false;
  // workspace
  __nv_bfloat16 *act = nullptr, *qkv = nullptr, *hbuf = nullptr, *emb_bf16 = nullptr;
  __nv_bfloat16* xb = nullptr;        // bf16 copy of the residual stream (GEMM A operand)
  float *stats[2] = {nullptr, nullptr};  // per-row LayerNorm partial sums, ping-pong
  int stat_parts = 0;                 // 2 * hidden / 256
  bool cls_only_last = true;          // VITK_FULL_LAST_LAYER=1 computes all tokens in the last layer
  bool resid_f32 = false;             // VITK_RESID_F32=1: fp32 master residual stream + bf16 copy (round-1 form)
  bool fused_patch = true;            // VITK_FUSED_PATCH=0: patch GEMM -> fp32 scratch -> cast + row-stats pass
  bool split_last_qkv = true;         // VITK_SPLIT_LAST_QKV=0: the last layer projects Q for every token too
  float *resid = nullptr, *emb = nullptr;
  uint8_t* pix_stage = nullptr;  // device staging for vitk_embed_host_u8
  std::map<int, BatchMaps> maps;
  // bank
  float *bank = nullptr, *bank_sqnorm = nullptr;
  __nv_bfloat16* bank_bf16 = nullptr;
  int32_t* bank_labels = nullptr;
  int bank_m = 0, bank_d = 0, bank_classes = 0;
  int bank_cap = 0;                 // rows allocated (vitk_bank_write grows geometrically)
  float* knn_pen = nullptr;         // [knn_pen_cap] row penalty of a filtered query
  int knn_pen_cap = 0;
  const float* knn_pen_cur = nullptr;  // == knn_pen while a filtered query is being launched
  CUtensorMap tm_bank;
  // knn workspace (grown on demand)
  float *knn_q = nullptr, *knn_scores = nullptr;
  __nv_bfloat16* knn_qb = nullptr;
  int32_t* knn_cand = nullptr;
  int knn_cap_q = 0, knn_cap_m = 0, knn_cap_kc = 0;
  // large-store path (bank_m > KNN_DIRECT_MAX): per-query candidate pools instead of scores
  uint2* knn_pool = nullptr;
  int* knn_pool_n = nullptr;
  unsigned long long* knn_thresh = nullptr;
  int* knn_overflow = nullptr;      // device, KNN_MAX_STAGES flags
  int* knn_overflow_host = nullptr; // pinned
  int knn_pool_cap_q = 0, knn_pool_cap_e = 0;   // queries / entries per query the pools hold
  int64_t knn_redo_stages = 0;      // stages redone after a pool overflow (diagnostic)
  int32_t *res_idx = nullptr, *res_label = nullptr, *res_counts = nullptr;
  float* res_score = nullptr;
  // resize plans: weight tables per input geometry (built once, kept on the device)
  struct ResizePlan {
    void* blob = nullptr;
    ResizeAxis hx{}, vy{};
    int ty = 0, max_rows = 0;
    size_t smem = 0;
  };
  std::map<std::pair<int, int>, ResizePlan> resize_plans;
  struct AreaPlan {
    void* blob = nullptr;
    AreaAxis x{}, y{};
    int fast_sx = 0, fast_sy = 0;
  };
  std::map<std::pair<int, int>, AreaPlan> area_plans;
  struct LinearPlan {
    void* blob = nullptr;
    const int *xs = nullptr, *xa = nullptr, *ys = nullptr, *ya = nullptr;
  };
  std::map<std::pair<int, int>, LinearPlan> linear_plans;
  uint8_t* raw_stage = nullptr;   // device staging for raw frames of vitk_embed_resize_host_u8
  size_t raw_stage_bytes = 0;
  // Viterbi workspace (grown on demand)
  double *hmm_loge = nullptr, *hmm_table = nullptr;
  int8_t* hmm_bp = nullptr;
  int* hmm_offsets = nullptr;
  int hmm_cap_t = 0, hmm_cap_seq = 0, hmm_cap_k = -1;
  // streamed Viterbi (vitk_hmm_stream_*): its own buffers, so it may run beside the forward
  double *hs_loge = nullptr, *hs_carry = nullptr;
  int8_t* hs_bp = nullptr;
  long long hs_cap = 0, hs_limit = 0, hs_t = -1;   // hs_t < 0: no stream open; hs_limit = capacity of the open stream
  int hs_seg_cap = 0;
  // several banks per handle (vitk_bank_new / vitk_bank_select): the fields above are the
  // CURRENT bank, the others are parked here
  struct BankState {
    float *bank = nullptr, *sqnorm = nullptr;
    __nv_bfloat16* bf16 = nullptr;
    int32_t* labels = nullptr;
    int m = 0, d = 0, classes = 0, cap = 0;
    CUtensorMap tm;
  };
  std::map<int, BankState> parked;
  int bank_id = 0, next_bank_id = 1;
  cudaStream_t own_stream = nullptr;
  // ordering between the caller's streams and own_stream (both work on the same workspace)
  cudaEvent_t ev_own = nullptr, ev_user = nullptr;
  bool own_pending = false, user_pending = false;
  cudaStream_t last_user = nullptr;
  int64_t launches = 0;
  // vitk_profile_fc1: CUDA-event pairs around every fc1 GEMM launch of the forward
  bool prof_fc1 = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
  // streaming host API (vitk_stream_submit / vitk_stream_wait): two batches in flight
  struct StreamSlot {
    uint8_t* pix = nullptr;
    uint8_t* raw = nullptr;     // raw-size frames of vitk_stream_submit_raw (grown on demand)
    size_t raw_bytes = 0;
    float *emb = nullptr, *score = nullptr;
    int32_t *idx = nullptr, *label = nullptr, *counts = nullptr;
    cudaEvent_t ev_h2d = nullptr, ev_done = nullptr, ev_out = nullptr;
    bool busy = false;
  } slots[2];
  cudaStream_t in_stream = nullptr, out_stream = nullptr;
  bool stream_ready = false;
};

namespace {

template <typename T>
int dev_alloc(Err& e, T** p, size_t count) {
  CK(e, cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T)));
  return 0;
}

// Every entry point works on the handle's one workspace (activations, k-NN buffers), some on the
// caller's stream, the host-buffer / streaming ones on own_stream.  order_enter makes the stream
// about to be used wait for the work last queued on the other side; order_leave marks the work
// just queued.  Calls on one stream need nothing; calls that alternate streams are serialised on
// the device instead of racing.
int order_enter(vitk_ctx* h, cudaStream_t st) {
  Err& e = h->err;
  CK(e, cudaSetDevice(h->device));
  if (st == h->own_stream) {
    if (h->user_pending) CK(e, cudaStreamWaitEvent(st, h->ev_user, 0));
  } else {
    if (h->own_pending) CK(e, cudaStreamWaitEvent(st, h->ev_own, 0));
    if (h->user_pending && h->last_user != st) CK(e, cudaStreamWaitEvent(st, h->ev_user, 0));
  }
  return 0;
}
int order_leave(vitk_ctx* h, cudaStream_t st) {
  Err& e = h->err;
  if (st == h->own_stream) {
    CK(e, cudaEventRecord(h->ev_own, st));
    h->own_pending = true;
  } else {
    CK(e, cudaEventRecord(h->ev_user, st));
    h->user_pending = true;
    h->last_user = st;
  }
  return 0;
}

void bank_park(vitk_ctx* h) {
  vitk_ctx::BankState b;
  b.bank = h->bank; b.sqnorm = h->bank_sqnorm; b.bf16 = h->bank_bf16; b.labels = h->bank_labels;
  b.m = h->bank_m; b.d = h->bank_d; b.classes = h->bank_classes; b.cap = h->bank_cap; b.tm = h->tm_bank;
  h->parked[h->bank_id] = b;
}
void bank_unpark(vitk_ctx* h, int id) {
  vitk_ctx::BankState b = h->parked[id];
  h->parked.erase(id);
  h->bank = b.bank; h->bank_sqnorm = b.sqnorm; h->bank_bf16 = b.bf16; h->bank_labels = b.labels;
  h->bank_m = b.m; h->bank_d = b.d; h->bank_classes = b.classes; h->bank_cap = b.cap; h->tm_bank = b.tm;
  h->bank_id = id;
}

int upload_bf16(Err& e, __nv_bfloat16* dst, const float* src, size_t count) {
  std::vector<uint16_t> tmp(count);
  for (size_t i = 0; i < count; ++i) tmp[i] = f32_to_bf16_bits(src[i]);
  CK(e, cudaMemcpy(dst, tmp.data(), count * 2, cudaMemcpyHostToDevice));
  return 0;
}
int upload_f32(Err& e, float* dst, const float* src, size_t count) {
  CK(e, cudaMemcpy(dst, src, count * 4, cudaMemcpyHostToDevice));
  return 0;
}

bool shape_is(const int64_t* shape, int ndim, std::initializer_list<int64_t> want) {
  // compare after dropping leading 1s on both sides
  std::vector<int64_t> a(shape, shape + ndim), b(want);
  while (a.size() > 1 && a.front() == 1) a.erase(a.begin());
  while (b.size() > 1 && b.front() == 1) b.erase(b.begin());
  return a == b;
}

void free_ctx(vitk_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  for (auto& L : c->layers) {
    cudaFree(L.w_qkv); cudaFree(L.w_o); cudaFree(L.w_fc1); cudaFree(L.w_fc2);
    cudaFree(L.b_qkv); cudaFree(L.b_o); cudaFree(L.b_fc1); cudaFree(L.b_fc2);
    cudaFree(L.c1_qkv); cudaFree(L.c1_fc1);
  }
  cudaFree(c->w_patch); cudaFree(c->posb); cudaFree(c->lnf_g); cudaFree(c->lnf_b);
  cudaFree(c->act); cudaFree(c->qkv); cudaFree(c->hbuf); cudaFree(c->emb_bf16);
  cudaFree(c->resid); cudaFree(c->emb); cudaFree(c->pix_stage);
  cudaFree(c->xb); cudaFree(c->stats[0]); cudaFree(c->stats[1]);
  cudaFree(c->bank); cudaFree(c->bank_sqnorm); cudaFree(c->bank_bf16); cudaFree(c->bank_labels);
  for (auto& kv : c->parked) {
    cudaFree(kv.second.bank); cudaFree(kv.second.sqnorm); cudaFree(kv.second.bf16); cudaFree(kv.second.labels);
  }
  if (c->ev_own) cudaEventDestroy(c->ev_own);
  if (c->ev_user) cudaEventDestroy(c->ev_user);
  cudaFree(c->knn_pen);
  cudaFree(c->res_idx); cudaFree(c->res_label); cudaFree(c->res_counts); cudaFree(c->res_score);
  cudaFree(c->knn_q); cudaFree(c->knn_scores); cudaFree(c->knn_qb); cudaFree(c->knn_cand);
  cudaFree(c->knn_pool); cudaFree(c->knn_pool_n); cudaFree(c->knn_thresh); cudaFree(c->knn_overflow);
  if (c->knn_overflow_host) cudaFreeHost(c->knn_overflow_host);
  cudaFree(c->hmm_loge); cudaFree(c->hmm_table); cudaFree(c->hmm_bp); cudaFree(c->hmm_offsets);
  cudaFree(c->hs_loge); cudaFree(c->hs_carry); cudaFree(c->hs_bp);
  for (auto& kv : c->resize_plans) cudaFree(kv.second.blob);
  for (auto& kv : c->area_plans) cudaFree(kv.second.blob);
  for (auto& kv : c->linear_plans) cudaFree(kv.second.blob);
  for (auto& pr : c->prof_events) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  cudaFree(c->raw_stage);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  for (auto& sl : c->slots) {
    cudaFree(sl.pix); cudaFree(sl.raw); cudaFree(sl.emb); cudaFree(sl.score); cudaFree(sl.idx); cudaFree(sl.label);
    cudaFree(sl.counts);
    if (sl.ev_h2d) cudaEventDestroy(sl.ev_h2d);
    if (sl.ev_done) cudaEventDestroy(sl.ev_done);
    if (sl.ev_out) cudaEventDestroy(sl.ev_out);
  }
  if (c->in_stream) cudaStreamDestroy(c->in_stream);
  if (c->out_stream) cudaStreamDestroy(c->out_stream);
  delete c;
}

}  // namespace

extern "C" const char* vitk_version(void) { return "vitk 0.1 (sm_100a, tcgen05/TMA)"; }

extern "C" int vitk_device_ok(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
    cudaGetLastError();
    return 0;
  }
  int major = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  return major == 10 ? 1 : 0;
}

extern "C" const char* vitk_last_error(vitk_handle h) {
  return h ? h->err.msg.c_str() : g_create_error.c_str();
}

extern "C" int64_t vitk_launch_count(vitk_handle h) { return h ? h->launches : 0; }
extern "C" int64_t vitk_knn_redo_count(vitk_handle h) { return h ? h->knn_redo_stages : 0; }

extern "C" int vitk_profile_fc1(vitk_handle h, int enable) {
  if (!h) return 1;
  h->prof_fc1 = enable != 0;
  return 0;
}

extern "C" int vitk_profile_fc1_read(vitk_handle h, double* total_ms, int* launches) {
  if (!h || !total_ms || !launches) return 1;
  Err& e = h->err;
  CK(e, cudaSetDevice(h->device));
  double tot = 0.0;
  int n = 0;
  for (auto& pr : h->prof_events) {
    CK(e, cudaEventSynchronize(pr.second));
    float ms = 0.f;
    CK(e, cudaEventElapsedTime(&ms, pr.first, pr.second));
    tot += ms;
    ++n;
    cudaEventDestroy(pr.first);
    cudaEventDestroy(pr.second);
  }
  h->prof_events.clear();
  *total_ms = tot;
  *launches = n;
  return 0;
}

extern "C" int vitk_create(const vitk_config* cfg, int device, vitk_handle* out) {
  Err e;
  auto fail = [&](vitk_ctx* c) {
    g_create_error = e.msg.empty() && c ? c->err.msg : e.msg;
    free_ctx(c);
    return 1;
  };
  if (!cfg || !out) { e.set("vitk_create: null argument"); return fail(nullptr); }
  *out = nullptr;
  if (!vitk_device_ok(device)) {
    e.set("vitk_create: device %d is not a visible sm_100 GPU (no CPU fallback exists)", device);
    return fail(nullptr);
  }
  if (cfg->hidden != cfg->heads * 64) { e.set("hidden must be heads*64"); return fail(nullptr); }
  if (cfg->hidden != 768 && cfg->hidden != 1024) { e.set("hidden must be 768 or 1024"); return fail(nullptr); }
  // the patch convolution is 'valid': rows / columns that do not fill a patch are dropped (the
  // reference's TF path feeds 432x768 frames to patch 32: 13 x 24 patches, generate_clips_hmm.py:48-56)
  if (cfg->patch < 8 || cfg->patch % 8 || cfg->image_h < cfg->patch || cfg->image_w < cfg->patch ||
      cfg->image_w % 8) {
    e.set("patch must be a multiple of 8 that fits the image; image width a multiple of 8"); return fail(nullptr);
  }
  if (cfg->hidden % 256 || cfg->mlp % 256 || cfg->max_batch < 1) {
    e.set("hidden/mlp must be multiples of 256 and max_batch >= 1"); return fail(nullptr);
  }
  vitk_ctx* c = new vitk_ctx();
  c->cfg = *cfg;
  c->device = device;
  c->P = (cfg->image_h / cfg->patch) * (cfg->image_w / cfg->patch);
  c->S = c->P + 1;
  c->Kp = cfg->channels * cfg->patch * cfg->patch;
  c->D = cfg->hidden;
  c->F = cfg->mlp;
  c->stat_parts = 2 * cfg->hidden / 256;
  {
    const char* v = getenv("VITK_FULL_LAST_LAYER");
    c->cls_only_last = !(v && atoi(v) == 1);
  }
  {
    const char* v = getenv("VITK_RESID_F32");
    c->resid_f32 = v && atoi(v) == 1;
    v = getenv("VITK_FUSED_PATCH");
    c->fused_patch = !(v && atoi(v) == 0);
    v = getenv("VITK_SPLIT_LAST_QKV");
    c->split_last_qkv = !(v && atoi(v) == 0);
  }
  if (c->S > ATT_MAX_S) { e.set("tokens per frame %d > %d unsupported by the attention kernel", c->S, ATT_MAX_S); return fail(c); }
  if (c->Kp % 64) { e.set("channels*patch^2 must be a multiple of 64"); return fail(c); }
  if (cudaSetDevice(device) != cudaSuccess) { e.set("cudaSetDevice failed"); return fail(c); }
  const size_t D = c->D, F = c->F, Kp = c->Kp;
  c->layers.resize(cfg->layers);
  int rc = 0;
  for (auto& L : c->layers) {
    rc |= dev_alloc(e, &L.w_qkv, 3 * D * D) | dev_alloc(e, &L.w_o, D * D) |
          dev_alloc(e, &L.w_fc1, F * D) | dev_alloc(e, &L.w_fc2, D * F) |
          dev_alloc(e, &L.b_qkv, 3 * D) | dev_alloc(e, &L.b_o, D) | dev_alloc(e, &L.b_fc1, F) |
          dev_alloc(e, &L.b_fc2, D) | dev_alloc(e, &L.c1_qkv, 3 * D) | dev_alloc(e, &L.c1_fc1, F);
    if (rc) return fail(c);
    rc |= tmap_operand(e, &L.tm_qkv, L.w_qkv, 3 * D, D, 128) |
          tmap_operand(e, &L.tm_q, L.w_qkv, D, D, 128) |
          tmap_operand(e, &L.tm_kv, L.w_qkv + D * D, 2 * D, D, 128) |
          tmap_operand(e, &L.tm_o, L.w_o, D, D, 128) |
          tmap_operand(e, &L.tm_fc1, L.w_fc1, F, D, 128) |
          tmap_operand(e, &L.tm_fc2, L.w_fc2, D, F, 128);
    if (rc) return fail(c);
  }
  const size_t M = (size_t)cfg->max_batch * c->S;
  const size_t actw = D > Kp ? D : Kp;
  rc |= dev_alloc(e, &c->w_patch, D * Kp) | dev_alloc(e, &c->posb, (size_t)c->S * D) |
        dev_alloc(e, &c->lnf_g, D) | dev_alloc(e, &c->lnf_b, D) | dev_alloc(e, &c->act, M * actw) |
        dev_alloc(e, &c->qkv, M * 3 * D) | dev_alloc(e, &c->hbuf, M * F) |
        ((c->resid_f32 || !c->fused_patch) ? dev_alloc(e, &c->resid, M * D) : 0) | dev_alloc(e, &c->xb, M * D) |
        dev_alloc(e, &c->stats[0], M * 2 * (2 * D / 256)) |
        dev_alloc(e, &c->stats[1], M * 2 * (2 * D / 256)) |
        dev_alloc(e, &c->emb, (size_t)cfg->max_batch * D) |
        dev_alloc(e, &c->emb_bf16, (size_t)cfg->max_batch * D) |
        dev_alloc(e, &c->pix_stage,
                  (size_t)cfg->max_batch * cfg->image_h * cfg->image_w * cfg->channels);
  if (rc) return fail(c);
  rc |= dev_alloc(e, &c->res_idx, (size_t)cfg->max_batch * 64) |
        dev_alloc(e, &c->res_score, (size_t)cfg->max_batch * 64) |
        dev_alloc(e, &c->res_label, (size_t)cfg->max_batch) |
        dev_alloc(e, &c->res_counts, (size_t)cfg->max_batch * KNN_MAX_CLASSES);
  if (rc) return fail(c);
  if (tmap_operand(e, &c->tm_patch, c->w_patch, D, Kp, 128)) return fail(c);
  if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_own, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_user, cudaEventDisableTiming) != cudaSuccess) {
    e.set("cudaStreamCreate / cudaEventCreate failed");
    return fail(c);
  }
  *out = c;
  return 0;
}

extern "C" int vitk_destroy(vitk_handle h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  free_ctx(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// weights (HF ViTModel state-dict keys, transformers modeling_vit.py)
// ------------------------------------------------------------------------------------------
extern "C" int vitk_set_weight(vitk_handle h, const char* name_c, const float* host,
                               const int64_t* shape, int ndim) {
  if (!h || !name_c || !host || !shape) return 1;
  Err& e = h->err;
  CK(e, cudaSetDevice(h->device));
  std::string name(name_c);
  if (name.rfind("vit.", 0) == 0) name = name.substr(4);
  if (name.rfind("pooler.", 0) == 0) return 0;  // never read by the reference
  const int64_t D = h->D, F = h->F, S = h->S, Kp = h->Kp;
  const int64_t ps = h->cfg.patch, C = h->cfg.channels;
  auto bad_shape = [&]() {
    e.set("vitk_set_weight: unexpected shape for %s", name.c_str());
    return 1;
  };
  int rc = 0;
  if (name == "embeddings.cls_token") {
    if (!shape_is(shape, ndim, {D})) return bad_shape();
    h->h_cls.assign(host, host + D);
  } else if (name == "embeddings.position_embeddings") {
    if (!shape_is(shape, ndim, {S, D})) return bad_shape();
    h->h_pos.assign(host, host + S * D);
  } else if (name == "embeddings.patch_embeddings.projection.weight") {
    if (!shape_is(shape, ndim, {D, C, ps, ps})) return bad_shape();
    rc = upload_bf16(e, h->w_patch, host, D * Kp);
  } else if (name == "embeddings.patch_embeddings.projection.bias") {
    if (!shape_is(shape, ndim, {D})) return bad_shape();
    h->h_pbias.assign(host, host + D);
  } else if (name == "layernorm.weight") {
    if (!shape_is(shape, ndim, {D})) return bad_shape();
    rc = upload_f32(e, h->lnf_g, host, D);
  } else if (name == "layernorm.bias") {
    if (!shape_is(shape, ndim, {D})) return bad_shape();
    rc = upload_f32(e, h->lnf_b, host, D);
  } else if (name.rfind("encoder.layer.", 0) == 0) {
    size_t pos = strlen("encoder.layer.");
    size_t dot = name.find('.', pos);
    if (dot == std::string::npos) { e.set("bad weight name %s", name.c_str()); return 1; }
    int li = atoi(name.substr(pos, dot - pos).c_str());
    if (li < 0 || li >= (int)h->layers.size()) { e.set("layer index out of range: %s", name.c_str()); return 1; }
    LayerW& L = h->layers[li];
    std::string rest = name.substr(dot + 1);
    // Tensors that take part in the LayerNorm fold are kept on the host until finalize.
    auto host_part = [&](std::vector<float>& dst, size_t total, size_t off, size_t cnt) {
      if (dst.size() != total) dst.assign(total, 0.f);
      memcpy(dst.data() + off, host, cnt * sizeof(float));
      L.dirty = true;
      return 0;
    };
    auto qkv_part = [&](int part, bool is_w) {
      if (is_w) {
        if (!shape_is(shape, ndim, {D, D})) return bad_shape();
        return host_part(L.h_wqkv, (size_t)3 * D * D, (size_t)part * D * D, (size_t)D * D);
      }
      if (!shape_is(shape, ndim, {D})) return bad_shape();
      return host_part(L.h_bqkv, (size_t)3 * D, (size_t)part * D, (size_t)D);
    };
    auto mat = [&](__nv_bfloat16* dst, int64_t r, int64_t c) {
      if (!shape_is(shape, ndim, {r, c})) return bad_shape();
      return upload_bf16(e, dst, host, r * c);
    };
    auto vec = [&](float* dst, int64_t n) {
      if (!shape_is(shape, ndim, {n})) return bad_shape();
      return upload_f32(e, dst, host, n);
    };
    auto hvec = [&](std::vector<float>& dst, int64_t n) {
      if (!shape_is(shape, ndim, {n})) return bad_shape();
      return host_part(dst, (size_t)n, 0, (size_t)n);
    };
    if (rest == "attention.attention.query.weight") rc = qkv_part(0, true);
    else if (rest == "attention.attention.query.bias") rc = qkv_part(0, false);
    else if (rest == "attention.attention.key.weight") rc = qkv_part(1, true);
    else if (rest == "attention.attention.key.bias") rc = qkv_part(1, false);
    else if (rest == "attention.attention.value.weight") rc = qkv_part(2, true);
    else if (rest == "attention.attention.value.bias") rc = qkv_part(2, false);
    else if (rest == "attention.output.dense.weight") rc = mat(L.w_o, D, D);
    else if (rest == "attention.output.dense.bias") rc = vec(L.b_o, D);
    else if (rest == "layernorm_before.weight") rc = hvec(L.h_ln1g, D);
    else if (rest == "layernorm_before.bias") rc = hvec(L.h_ln1b, D);
    else if (rest == "layernorm_after.weight") rc = hvec(L.h_ln2g, D);
    else if (rest == "layernorm_after.bias") rc = hvec(L.h_ln2b, D);
    else if (rest == "intermediate.dense.weight") {
      if (!shape_is(shape, ndim, {F, D})) return bad_shape();
      rc = host_part(L.h_wfc1, (size_t)F * D, 0, (size_t)F * D);
    } else if (rest == "intermediate.dense.bias") rc = hvec(L.h_bfc1, F);
    else if (rest == "output.dense.weight") rc = mat(L.w_fc2, D, F);
    else if (rest == "output.dense.bias") rc = vec(L.b_fc2, D);
    else { e.set("vitk_set_weight: unknown key %s", name.c_str()); return 1; }
  } else {
    e.set("vitk_set_weight: unknown key %s", name.c_str());
    return 1;
  }
  if (rc) return rc;
  h->have.insert(name);
  h->finalized = false;
  return 0;
}

extern "C" int vitk_finalize_weights(vitk_handle h) {
  if (!h) return 1;
  Err& e = h->err;
  const size_t want = 6 + 16 * h->layers.size();
  if (h->have.size() != want) {
    e.set("vitk_finalize_weights: %zu of %zu tensors supplied", h->have.size(), want);
    return 1;
  }
  CK(e, cudaSetDevice(h->device));
  // table[0] = cls + pos[0] (the CLS row of the patch matrix is all zeros, so the GEMM adds
  // nothing); table[t>0] = pos[t] + conv bias.  HF modeling_vit.py:117-124.
  const int S = h->S, D = h->D;
  std::vector<float> tab((size_t)S * D);
  for (int t = 0; t < S; ++t)
    for (int j = 0; j < D; ++j)
      tab[(size_t)t * D + j] =
          h->h_pos[(size_t)t * D + j] + (t == 0 ? h->h_cls[j] : h->h_pbias[j]);
  if (upload_f32(e, h->posb, tab.data(), tab.size())) return 1;
  // LayerNorm fold (gemm_sm100.cuh): W' = bf16(gamma∘W), c1[n] = sum_k W'[n,k] (from the
  // rounded values, so a constant row cancels exactly), c2[n] = b[n] + sum_k beta[k] W[n,k].
  for (size_t li = 0; li < h->layers.size(); ++li) {
    LayerW& L = h->layers[li];
    if (!L.dirty) continue;
    const size_t Dz = (size_t)D, Fz = (size_t)h->F;
    if (L.h_wqkv.size() != 3 * Dz * Dz || L.h_bqkv.size() != 3 * Dz || L.h_wfc1.size() != Fz * Dz ||
        L.h_bfc1.size() != Fz || L.h_ln1g.size() != Dz || L.h_ln1b.size() != Dz ||
        L.h_ln2g.size() != Dz || L.h_ln2b.size() != Dz) {
      e.set("vitk_finalize_weights: layer %zu changed since the last finalize — re-supply its "
            "layernorm_before/after, query/key/value and intermediate.dense tensors", li);
      return 1;
    }
    auto fold = [&](const std::vector<float>& W, const std::vector<float>& b,
                    const std::vector<float>& g, const std::vector<float>& be, size_t rows,
                    __nv_bfloat16* w_dev, float* c1_dev, float* c2_dev) {
      std::vector<uint16_t> wq(rows * Dz);
      std::vector<float> c1(rows), c2(rows);
      for (size_t n = 0; n < rows; ++n) {
        double s1 = 0, s2 = 0;
        const float* wr = W.data() + n * Dz;
        for (size_t k = 0; k < Dz; ++k) {
          const uint16_t q = f32_to_bf16_bits(g[k] * wr[k]);
          wq[n * Dz + k] = q;
          uint32_t u = (uint32_t)q << 16;
          float f;
          memcpy(&f, &u, 4);
          s1 += f;
          s2 += (double)be[k] * wr[k];
        }
        c1[n] = (float)s1;
        c2[n] = (float)(s2 + b[n]);
      }
      if (cudaMemcpy(w_dev, wq.data(), wq.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess ||
          cudaMemcpy(c1_dev, c1.data(), rows * 4, cudaMemcpyHostToDevice) != cudaSuccess ||
          cudaMemcpy(c2_dev, c2.data(), rows * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
        e.set("vitk_finalize_weights: upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        return 1;
      }
      return 0;
    };
    if (fold(L.h_wqkv, L.h_bqkv, L.h_ln1g, L.h_ln1b, 3 * Dz, L.w_qkv, L.c1_qkv, L.b_qkv) ||
        fold(L.h_wfc1, L.h_bfc1, L.h_ln2g, L.h_ln2b, Fz, L.w_fc1, L.c1_fc1, L.b_fc1))
      return 1;
    // the big host copies are only needed for the fold
    std::vector<float>().swap(L.h_wqkv);
    std::vector<float>().swap(L.h_wfc1);
    L.dirty = false;
  }
  h->finalized = true;
  return 0;
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
namespace {

int get_maps(vitk_ctx* c, int n, BatchMaps** out) {
  auto it = c->maps.find(n);
  if (it != c->maps.end()) { *out = &it->second; return 0; }
  Err& e = c->err;
  BatchMaps m;
  const int M = n * c->S, D = c->D, F = c->F;
  const int sk = attn_smem_layout(c->S).kv_box;
  int rc = tmap_operand(e, &m.a_patch, c->act, M, c->Kp, GEMM_BM) |
           tmap_operand(e, &m.a_act, c->act, M, D, GEMM_BM) |
           tmap_operand(e, &m.a_h, c->hbuf, M, F, GEMM_BM) |
           tmap_operand(e, &m.a_xb, c->xb, M, D, GEMM_BM) |
           (c->resid_f32 ? tmap_output(e, &m.c_resid, c->resid, M, D, false) : 0) |
           (c->resid ? tmap_output(e, &m.c_patch, c->resid, M, D, false) : 0) |
           tmap_output(e, &m.c_xb, c->xb, M, D, true) |
           tmap_output(e, &m.c_xb_cls, c->xb, n, D, true, (long long)c->S * D) |
           tmap_output(e, &m.c_qkv, c->qkv, M, 3 * D, true) |
           tmap_output(e, &m.c_kv, c->qkv + D, M, 2 * D, true, 3LL * D) |
           tmap_output(e, &m.c_q_cls, c->qkv, n, D, true, (long long)c->S * 3 * D) |
           tmap_output(e, &m.c_h, c->hbuf, M, F, true) |
           tmap_tokens(e, &m.att_q, c->qkv, n, c->S, 3 * D, 128) |
           tmap_tokens(e, &m.att_kv, c->qkv, n, c->S, 3 * D, sk) |
           tmap_tokens(e, &m.att_o, c->act, n, c->S, D, 128) |
           tmap_operand(e, &m.a_act_cls, c->act, n, D, GEMM_BM, (long long)c->S * D) |
           tmap_operand(e, &m.a_xb_cls, c->xb, n, D, GEMM_BM, (long long)c->S * D) |
           tmap_operand(e, &m.a_h_cls, c->hbuf, n, F, GEMM_BM) |
           (c->resid_f32 ? tmap_output(e, &m.c_resid_cls, c->resid, n, D, false, (long long)c->S * D) : 0) |
           tmap_output(e, &m.c_h_cls, c->hbuf, n, F, true);
  if (rc) return 1;
  *out = &c->maps.emplace(n, m).first->second;
  return 0;
}

// One chunk of n <= max_batch frames, everything asynchronous on `st`.
int forward_chunk(vitk_ctx* c, const void* pix, int fmt, int n, float* emb_f32,
                  __nv_bfloat16* emb_bf16, int l2norm, cudaStream_t st) {
  Err& e = c->err;
  BatchMaps* mp = nullptr;
  if (get_maps(c, n, &mp)) return 1;
  const int M = n * c->S, D = c->D, F = c->F;
  const vitk_config& cfg = c->cfg;
  // (a3) im2col, (a3/a4) patch GEMM + CLS/pos/bias table -> fp32 residual stream
  if (launch_patchify(e, pix, fmt, c->act, n, cfg.channels, cfg.image_h, cfg.image_w, cfg.patch, st,
                      cfg.u8_std != 0.f ? cfg.u8_mean : 127.5f, cfg.u8_std != 0.f ? cfg.u8_std : 127.5f))
    return 1;
  GemmParams gp{};
  gp.M = M; gp.N = D; gp.K = c->Kp;
  gp.table = c->posb; gp.table_rows = c->S;
  const int parts = c->stat_parts;
  const long long sstride = 2LL * parts;
  if (!c->resid_f32 && c->fused_patch) {
    // bf16 residual stream: the patch GEMM's epilogue writes the stream and the row sums itself
    gp.stats_out = c->stats[0];
    gp.stats_out_stride = sstride;
    if (launch_gemm(e, EPI_TABLE_BF16, mp->a_patch, c->tm_patch, mp->c_xb, mp->c_xb, gp, st)) return 1;
    c->launches -= 1;   // one launch fewer than the fp32 form counted below
  } else {
    // fp32 master residual (VITK_RESID_F32=1) or VITK_FUSED_PATCH=0: fp32 tokens, then one pass for
    // the bf16 stream / copy + the row statistics that feed layer 0's folded LayerNorm
    if (launch_gemm(e, EPI_TABLE_F32, mp->a_patch, c->tm_patch, mp->c_patch, mp->c_patch, gp, st))
      return 1;
    if (launch_rowstats(e, c->resid, c->xb, c->stats[0], M, D, parts, st)) return 1;
  }
  // Residual stream: bf16 in place in `xb` (default; the residual GEMMs read and write the tile
  // their own epilogue owns) or the round-1 form, an fp32 master copy + a bf16 copy (VITK_RESID_F32=1).
  const bool rf32 = c->resid_f32;
  const int epi_res = rf32 ? EPI_BIAS_RESID_F32 : EPI_BIAS_RESID_BF16;
  const CUtensorMap& c_res = rf32 ? mp->c_resid : mp->c_xb;
  const CUtensorMap& c_res_cls = rf32 ? mp->c_resid_cls : mp->c_xb_cls;
  c->launches += 3;
  const int epi_fc1 = cfg.gelu_tanh ? EPI_BIAS_GELUT_BF16 : EPI_BIAS_GELU_BF16;
  const size_t nl = c->layers.size();
  // "Snake" order: every big kernel walks its rows in the direction opposite to the kernel that
  // produced its input, so it starts on the ~100 MB its producer wrote LAST -- still in the
  // 126 MB L2 -- instead of on rows that were evicted long ago (each activation is 77-310 MB).
  // Results do not depend on the order.  VITK_SNAKE=0 keeps every kernel ascending.
  static int snake = -1;
  if (snake < 0) { const char* v = getenv("VITK_SNAKE"); snake = (v && atoi(v) == 0) ? 0 : 1; }
  int dir = snake;  // rowstats_cast wrote xb ascending: the first QKV GEMM goes down
  auto next_dir = [&]() { const int d = dir; dir ^= snake; return d; };
  for (size_t li = 0; li < nl; ++li) {
    LayerW& L = c->layers[li];
    // (a5.1+a5.2) LN1 folded into the fused QKV GEMM: A = raw residual (bf16), stats[0]
    GemmParams q{};
    q.M = M; q.N = 3 * D; q.K = D; q.bias = L.b_qkv; q.c1 = L.c1_qkv;
    q.stats = c->stats[0]; q.stats_parts = parts; q.stats_stride = sstride;
    q.inv_d = 1.0f / D; q.eps = cfg.ln_eps;
    q.reverse_m = next_dir();
    // Last layer: the reference reads only last_hidden_state[:,0,:] (ratt_write_clips_to_ragdb.py
    // :148), and after the attention every remaining op is row-wise, so only the CLS query and
    // the CLS rows of proj / MLP are computed (bit-identical for those rows).
    const bool cls_only = (li + 1 == nl) && c->cls_only_last && c->S > 1;
    if (cls_only && c->split_last_qkv) {
      // ... which also means only the CLS rows need a QUERY: K and V for every token (the 2D
      // right-hand columns of the fused projection), Q for one row per frame
      GemmParams kv = q;
      kv.N = 2 * D; kv.bias = L.b_qkv + D; kv.c1 = L.c1_qkv + D;
      if (launch_gemm(e, EPI_BIAS_BF16, mp->a_xb, L.tm_kv, mp->c_kv, mp->c_kv, kv, st)) return 1;
      GemmParams qc = q;
      qc.M = n; qc.N = D; qc.stats_stride = (long long)c->S * sstride; qc.reverse_m = 0;
      if (launch_gemm(e, EPI_BIAS_BF16, mp->a_xb_cls, L.tm_q, mp->c_q_cls, mp->c_q_cls, qc, st)) return 1;
      c->launches += 1;
    } else if (launch_gemm(e, EPI_BIAS_BF16, mp->a_xb, L.tm_qkv, mp->c_qkv, mp->c_qkv, q, st)) {
      return 1;
    }
    // (a5.3) attention
    if (launch_attention(e, mp->att_q, mp->att_kv, mp->att_o, n, c->S, cfg.heads, st,
                         cls_only ? 1 : 0, next_dir()))
      return 1;
    if (cls_only) {
      const long long pitch = (long long)c->S;
      GemmParams o{};
      o.M = n; o.N = D; o.K = D; o.bias = L.b_o;
      if (rf32) { o.xb = c->xb; o.xb_ld = pitch * D; }
      o.stats_out = c->stats[1]; o.stats_out_stride = pitch * sstride;
      if (launch_gemm(e, epi_res, mp->a_act_cls, L.tm_o, c_res_cls, c_res_cls, o, st))
        return 1;
      GemmParams f1{};
      f1.M = n; f1.N = F; f1.K = D; f1.bias = L.b_fc1; f1.c1 = L.c1_fc1;
      f1.stats = c->stats[1]; f1.stats_parts = parts; f1.stats_stride = pitch * sstride;
      f1.inv_d = 1.0f / D; f1.eps = cfg.ln_eps;
      if (launch_gemm(e, epi_fc1, mp->a_xb_cls, L.tm_fc1, mp->c_h_cls, mp->c_h_cls, f1, st)) return 1;
      GemmParams f2{};
      f2.M = n; f2.N = D; f2.K = F; f2.bias = L.b_fc2;
      if (launch_gemm(e, epi_res, mp->a_h_cls, L.tm_fc2, c_res_cls, c_res_cls, f2, st))
        return 1;
      c->launches += 5;
      continue;
    }
    // (a5.4) out-proj + residual; epilogue also emits the bf16 copy and LN2's row sums
    GemmParams o{};
    o.M = M; o.N = D; o.K = D; o.bias = L.b_o;
    if (rf32) { o.xb = c->xb; o.xb_ld = D; }
    o.stats_out = c->stats[1]; o.stats_out_stride = sstride;
    o.reverse_m = next_dir();
    if (launch_gemm(e, epi_res, mp->a_act, L.tm_o, c_res, c_res, o, st))
      return 1;
    // (a5.1+a5.5) LN2 folded into fc1 + GELU, then fc2 + residual (+ next layer's LN1 inputs)
    GemmParams f1{};
    f1.M = M; f1.N = F; f1.K = D; f1.bias = L.b_fc1; f1.c1 = L.c1_fc1;
    f1.stats = c->stats[1]; f1.stats_parts = parts; f1.stats_stride = sstride;
    f1.inv_d = 1.0f / D; f1.eps = cfg.ln_eps;
    f1.reverse_m = next_dir();
    cudaEvent_t pe0 = nullptr, pe1 = nullptr;
    if (c->prof_fc1) {
      CK(e, cudaEventCreate(&pe0));
      CK(e, cudaEventCreate(&pe1));
      CK(e, cudaEventRecord(pe0, st));
    }
    if (launch_gemm(e, epi_fc1, mp->a_xb, L.tm_fc1, mp->c_h, mp->c_h, f1, st)) return 1;
    if (c->prof_fc1) {
      CK(e, cudaEventRecord(pe1, st));
      c->prof_events.emplace_back(pe0, pe1);
    }
    GemmParams f2{};
    f2.M = M; f2.N = D; f2.K = F; f2.bias = L.b_fc2;
    f2.reverse_m = next_dir();
    if (li + 1 < nl) {  // nobody reads the copy / sums after the last layer
      if (rf32) { f2.xb = c->xb; f2.xb_ld = D; }
      f2.stats_out = c->stats[0]; f2.stats_out_stride = sstride;
    }
    if (launch_gemm(e, epi_res, mp->a_h, L.tm_fc2, c_res, c_res, f2, st))
      return 1;
    c->launches += 5;
  }
  // (a6) final LN on the CLS rows only, (a7) row normalisation
  const int blocks = (n + 3) / 4;
  const float* rs_f32 = rf32 ? c->resid : nullptr;
  const __nv_bfloat16* rs_bf16 = rf32 ? nullptr : c->xb;
  if (D == 768)
    CK(e, launch_pdl(cls_tail_kernel<6>, dim3(blocks), dim3(128), 0, st, rs_f32, rs_bf16, c->lnf_g, c->lnf_b, emb_f32,
                     emb_bf16, n, c->S, cfg.ln_eps, l2norm));
  else
    CK(e, launch_pdl(cls_tail_kernel<8>, dim3(blocks), dim3(128), 0, st, rs_f32, rs_bf16, c->lnf_g, c->lnf_b, emb_f32,
                     emb_bf16, n, c->S, cfg.ln_eps, l2norm));
  CK(e, cudaGetLastError());
  c->launches += 1;
  return 0;
}

size_t pixel_bytes_per_frame(const vitk_config& cfg, int fmt) {
  size_t px = (size_t)cfg.channels * cfg.image_h * cfg.image_w;
  return fmt == PIX_F32_NCHW ? px * 4 : fmt == PIX_BF16_NCHW ? px * 2 : px;
}

}  // namespace

extern "C" int vitk_embed(vitk_handle h, const void* pixels_dev, int pixel_format, int n,
                          float* emb_f32_dev, void* emb_bf16_dev, int l2_normalise, void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (!h->finalized) { e.set("vitk_embed: weights not finalized"); return 1; }
  if (n == 0) return 0;  // an empty batch is not an error (empty tensors have null pointers)
  if (!pixels_dev || !emb_f32_dev || n < 0) { e.set("vitk_embed: bad argument"); return 1; }
  if (pixel_format < 0 || pixel_format > 2) { e.set("vitk_embed: bad pixel format"); return 1; }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (order_enter(h, st)) return 1;
  const size_t fb = pixel_bytes_per_frame(h->cfg, pixel_format);
  for (int off = 0; off < n; off += h->cfg.max_batch) {
    const int cn = n - off < h->cfg.max_batch ? n - off : h->cfg.max_batch;
    __nv_bfloat16* eb = emb_bf16_dev
                            ? reinterpret_cast<__nv_bfloat16*>(emb_bf16_dev) + (size_t)off * h->D
                            : nullptr;
    if (forward_chunk(h, reinterpret_cast<const uint8_t*>(pixels_dev) + (size_t)off * fb,
                      pixel_format, cn, emb_f32_dev + (size_t)off * h->D, eb, l2_normalise, st))
      return 1;
  }
  if (order_leave(h, st)) return 1;
  return 0;
}

extern "C" int vitk_embed_host_u8(vitk_handle h, const uint8_t* frames_host, int n, float* emb_host,
                                  int l2_normalise) {
  if (!h) return 1;
  Err& e = h->err;
  if (!h->finalized) { e.set("vitk_embed_host_u8: weights not finalized"); return 1; }
  if (n == 0) return 0;
  if (!frames_host || !emb_host || n < 0) { e.set("vitk_embed_host_u8: bad argument"); return 1; }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = h->own_stream;
  if (order_enter(h, st)) return 1;
  const size_t fb = pixel_bytes_per_frame(h->cfg, PIX_U8_NHWC);
  for (int off = 0; off < n; off += h->cfg.max_batch) {
    const int cn = n - off < h->cfg.max_batch ? n - off : h->cfg.max_batch;
    CK(e, cudaMemcpyAsync(h->pix_stage, frames_host + (size_t)off * fb, (size_t)cn * fb,
                          cudaMemcpyHostToDevice, st));
    if (forward_chunk(h, h->pix_stage, PIX_U8_NHWC, cn, h->emb, nullptr, l2_normalise, st)) return 1;
    CK(e, cudaMemcpyAsync(emb_host + (size_t)off * h->D, h->emb, (size_t)cn * h->D * 4,
                          cudaMemcpyDeviceToHost, st));
  }
  if (order_leave(h, st)) return 1;
  CK(e, cudaStreamSynchronize(st));
  return 0;
}

// ------------------------------------------------------------------------------------------
// bank + k-NN vote
// ------------------------------------------------------------------------------------------
extern "C" int vitk_bank_set(vitk_handle h, const float* rows_host, int m, int d,
                             const int32_t* labels_host, int n_classes, int normalise_rows) {
  if (!h) return 1;
  Err& e = h->err;
  if (!rows_host || m < 1 || d < 64 || d % 64 != 0) {
    e.set("vitk_bank_set: need m >= 1 rows and d a multiple of 64 (got m=%d d=%d)", m, d);
    return 1;
  }
  if (n_classes < 0 || n_classes > KNN_MAX_CLASSES) {
    e.set("vitk_bank_set: n_classes must be in [0,%d]", KNN_MAX_CLASSES);
    return 1;
  }
  CK(e, cudaSetDevice(h->device));
  CK(e, cudaDeviceSynchronize());
  cudaFree(h->bank); cudaFree(h->bank_sqnorm); cudaFree(h->bank_bf16); cudaFree(h->bank_labels);
  h->bank = nullptr; h->bank_sqnorm = nullptr; h->bank_bf16 = nullptr; h->bank_labels = nullptr;
  h->bank_m = 0;
  std::vector<float> rows(rows_host, rows_host + (size_t)m * d);
  std::vector<float> sq(m);
  for (int i = 0; i < m; ++i) {
    float* r = rows.data() + (size_t)i * d;
    if (normalise_rows) {
      // numpy: np.linalg.norm (float32 in, float32 out) then x /= (norm + 1e-8)
      double s = 0;
      for (int j = 0; j < d; ++j) s += (double)r[j] * r[j];
      const float inv_den = (float)std::sqrt(s) + 1e-8f;
      for (int j = 0; j < d; ++j) r[j] = r[j] / inv_den;
    }
    double s2 = 0;
    for (int j = 0; j < d; ++j) s2 += (double)r[j] * r[j];
    sq[i] = (float)s2;
  }
  if (dev_alloc(e, &h->bank, (size_t)m * d) || dev_alloc(e, &h->bank_bf16, (size_t)m * d) ||
      dev_alloc(e, &h->bank_sqnorm, (size_t)m) || dev_alloc(e, &h->bank_labels, (size_t)m))
    return 1;
  if (upload_f32(e, h->bank, rows.data(), rows.size()) ||
      upload_bf16(e, h->bank_bf16, rows.data(), rows.size()) ||
      upload_f32(e, h->bank_sqnorm, sq.data(), sq.size()))
    return 1;
  std::vector<int32_t> lab(m, 0);
  if (labels_host) {
    for (int i = 0; i < m; ++i) {
      if (labels_host[i] < 0 || labels_host[i] >= (n_classes > 0 ? n_classes : 1)) {
        e.set("vitk_bank_set: label %d of row %d outside [0,%d)", labels_host[i], i, n_classes);
        return 1;
      }
      lab[i] = labels_host[i];
    }
  }
  CK(e, cudaMemcpy(h->bank_labels, lab.data(), (size_t)m * 4, cudaMemcpyHostToDevice));
  if (tmap_operand(e, &h->tm_bank, h->bank_bf16, m, d, 128)) return 1;
  h->bank_m = m;
  h->bank_d = d;
  h->bank_classes = labels_host ? n_classes : 0;
  h->bank_cap = m;
  h->knn_cap_m = 0;  // force workspace re-check
  return 0;
}

// Device-resident variant of vitk_bank_set: rows_dev (m,d) f32 and labels_dev (m) int32 (or
// NULL) already live on this handle's device (a store built by vitk_embed, or synthetic).  The
// same row normalisation / bf16 copy / ||b||^2 are produced by one kernel pass.
extern "C" int vitk_bank_set_device(vitk_handle h, const float* rows_dev, int m, int d,
                                    const int32_t* labels_dev, int n_classes, int normalise_rows,
                                    void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (!rows_dev || m < 1 || d < 64 || d % 64 != 0) {
    e.set("vitk_bank_set_device: need m >= 1 rows and d a multiple of 64 (got m=%d d=%d)", m, d);
    return 1;
  }
  if (n_classes < 0 || n_classes > KNN_MAX_CLASSES) {
    e.set("vitk_bank_set_device: n_classes must be in [0,%d]", KNN_MAX_CLASSES);
    return 1;
  }
  CK(e, cudaSetDevice(h->device));
  CK(e, cudaDeviceSynchronize());
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (order_enter(h, st)) return 1;
  cudaFree(h->bank); cudaFree(h->bank_sqnorm); cudaFree(h->bank_bf16); cudaFree(h->bank_labels);
  h->bank = nullptr; h->bank_sqnorm = nullptr; h->bank_bf16 = nullptr; h->bank_labels = nullptr;
  h->bank_m = 0;
  if (dev_alloc(e, &h->bank, (size_t)m * d) || dev_alloc(e, &h->bank_bf16, (size_t)m * d) ||
      dev_alloc(e, &h->bank_sqnorm, (size_t)m) || dev_alloc(e, &h->bank_labels, (size_t)m))
    return 1;
  bank_prep_kernel<<<(m + 7) / 8, 256, 0, st>>>(rows_dev, h->bank, h->bank_bf16, h->bank_sqnorm, m, d,
                                                normalise_rows);
  CK(e, cudaGetLastError());
  if (labels_dev)
    CK(e, cudaMemcpyAsync(h->bank_labels, labels_dev, (size_t)m * 4, cudaMemcpyDeviceToDevice, st));
  else
    CK(e, cudaMemsetAsync(h->bank_labels, 0, (size_t)m * 4, st));
  CK(e, cudaStreamSynchronize(st));
  if (tmap_operand(e, &h->tm_bank, h->bank_bf16, m, d, 128)) return 1;
  h->bank_m = m;
  h->bank_d = d;
  h->bank_classes = labels_dev ? n_classes : 0;
  h->bank_cap = m;
  h->launches += 1;
  return 0;
}

// Mutable bank (the confident-upsert loop, generate_clips_hmm.py:456-470 / chroma.py:257-261):
// rows [at_row, at_row + m_new) are written -- at_row == bank rows appends, smaller overwrites
// (Chroma upsert of an existing id).  Storage grows geometrically, only the new rows go through
// the prep kernel, so a 16-row upsert costs microseconds whatever the bank size.
extern "C" int vitk_bank_write(vitk_handle h, int at_row, const float* rows, int m_new,
                               const int32_t* labels, int on_device, int normalise_rows,
                               void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (h->bank_m == 0) { e.set("vitk_bank_write: no bank set (call vitk_bank_set first)"); return 1; }
  if (m_new == 0) return 0;
  if (!rows || m_new < 0 || at_row < 0 || at_row > h->bank_m) {
    e.set("vitk_bank_write: bad argument (at_row=%d, m_new=%d, bank rows=%d)", at_row, m_new, h->bank_m);
    return 1;
  }
  if ((labels != nullptr) != (h->bank_classes > 0)) {
    e.set("vitk_bank_write: labels must be given exactly when the bank has labels");
    return 1;
  }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (order_enter(h, st)) return 1;
  const int d = h->bank_d;
  const int m_after = at_row + m_new > h->bank_m ? at_row + m_new : h->bank_m;
  if (m_after > h->bank_cap) {
    CK(e, cudaDeviceSynchronize());  // nothing may still be reading the old arrays
    int cap = 2 * h->bank_cap;
    if (cap < m_after) cap = m_after;
    cap = (cap + 127) & ~127;
    float *nb = nullptr, *nsq = nullptr;
    __nv_bfloat16* nbb = nullptr;
    int32_t* nl = nullptr;
    if (dev_alloc(e, &nb, (size_t)cap * d) || dev_alloc(e, &nbb, (size_t)cap * d) ||
        dev_alloc(e, &nsq, (size_t)cap) || dev_alloc(e, &nl, (size_t)cap)) {
      cudaFree(nb); cudaFree(nbb); cudaFree(nsq); cudaFree(nl);
      return 1;
    }
    const size_t m0 = (size_t)h->bank_m;
    CK(e, cudaMemcpy(nb, h->bank, m0 * d * 4, cudaMemcpyDeviceToDevice));
    CK(e, cudaMemcpy(nbb, h->bank_bf16, m0 * d * 2, cudaMemcpyDeviceToDevice));
    CK(e, cudaMemcpy(nsq, h->bank_sqnorm, m0 * 4, cudaMemcpyDeviceToDevice));
    CK(e, cudaMemcpy(nl, h->bank_labels, m0 * 4, cudaMemcpyDeviceToDevice));
    cudaFree(h->bank); cudaFree(h->bank_bf16); cudaFree(h->bank_sqnorm); cudaFree(h->bank_labels);
    h->bank = nb; h->bank_bf16 = nbb; h->bank_sqnorm = nsq; h->bank_labels = nl;
    h->bank_cap = cap;
  }
  const float* src = rows;
  float* tmp = nullptr;
  if (!on_device) {
    if (labels) {
      for (int i = 0; i < m_new; ++i)
        if (labels[i] < 0 || labels[i] >= h->bank_classes) {
          e.set("vitk_bank_write: label %d of row %d outside [0,%d)", labels[i], i, h->bank_classes);
          return 1;
        }
    }
    if (dev_alloc(e, &tmp, (size_t)m_new * d)) return 1;
    src = tmp;
  }
  // every CUDA call below reports through `st_err` so that the staging copy is released on any path
  cudaError_t st_err = cudaSuccess;
  auto ok = [&](cudaError_t r) { if (st_err == cudaSuccess) st_err = r; return st_err == cudaSuccess; };
  if (!on_device) ok(cudaMemcpyAsync(tmp, rows, (size_t)m_new * d * 4, cudaMemcpyHostToDevice, st));
  if (labels && st_err == cudaSuccess)
    ok(cudaMemcpyAsync(h->bank_labels + at_row, labels, (size_t)m_new * 4,
                       on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
  if (!labels && st_err == cudaSuccess) ok(cudaMemsetAsync(h->bank_labels + at_row, 0, (size_t)m_new * 4, st));
  if (st_err == cudaSuccess) {
    bank_prep_kernel<<<(m_new + 7) / 8, 256, 0, st>>>(src, h->bank + (size_t)at_row * d,
                                                      h->bank_bf16 + (size_t)at_row * d,
                                                      h->bank_sqnorm + at_row, m_new, d, normalise_rows);
    ok(cudaGetLastError());
  }
  ok(cudaStreamSynchronize(st));  // host rows / the staging copy may be released on return
  cudaFree(tmp);
  if (st_err != cudaSuccess) {
    e.set("vitk_bank_write: CUDA error: %s", cudaGetErrorString(st_err));
    return 1;
  }
  if (tmap_operand(e, &h->tm_bank, h->bank_bf16, m_after, d, 128)) return 1;
  h->bank_m = m_after;
  h->launches += 1;
  return 0;
}

extern "C" int vitk_bank_rows(vitk_handle h) { return h ? h->bank_m : -1; }

extern "C" int vitk_bank_new(vitk_handle h) {
  if (!h) return -1;
  bank_park(h);
  h->bank = nullptr; h->bank_sqnorm = nullptr; h->bank_bf16 = nullptr; h->bank_labels = nullptr;
  h->bank_m = 0; h->bank_d = 0; h->bank_classes = 0; h->bank_cap = 0;
  h->bank_id = h->next_bank_id++;
  return h->bank_id;
}

extern "C" int vitk_bank_current(vitk_handle h) { return h ? h->bank_id : -1; }

extern "C" int vitk_bank_select(vitk_handle h, int bank_id) {
  if (!h) return 1;
  if (bank_id == h->bank_id) return 0;
  if (h->parked.find(bank_id) == h->parked.end()) {
    h->err.set("vitk_bank_select: no bank %d on this handle", bank_id);
    return 1;
  }
  bank_park(h);
  bank_unpark(h, bank_id);
  return 0;
}

extern "C" int vitk_bank_free(vitk_handle h, int bank_id) {
  if (!h) return 1;
  Err& e = h->err;
  CK(e, cudaSetDevice(h->device));
  if (bank_id == h->bank_id) {
    CK(e, cudaDeviceSynchronize());
    cudaFree(h->bank); cudaFree(h->bank_sqnorm); cudaFree(h->bank_bf16); cudaFree(h->bank_labels);
    h->bank = nullptr; h->bank_sqnorm = nullptr; h->bank_bf16 = nullptr; h->bank_labels = nullptr;
    h->bank_m = 0; h->bank_d = 0; h->bank_classes = 0; h->bank_cap = 0;
    return 0;
  }
  auto it = h->parked.find(bank_id);
  if (it == h->parked.end()) return 0;   // already gone
  CK(e, cudaDeviceSynchronize());
  cudaFree(it->second.bank); cudaFree(it->second.sqnorm); cudaFree(it->second.bf16); cudaFree(it->second.labels);
  h->parked.erase(it);
  return 0;
}

namespace {
int knn_candidates_for(int k, int m) {
  // bf16 similarity error on unit vectors is ~1e-4..1e-3; the slack keeps every true top-k
  // row inside the candidate set so the fp32 re-rank can restore the exact order.  Small k (the
  // vote: 5 / 25 / 50): 2k + 32, at most 128 (warp-extraction re-rank).  Larger k (the clip-store
  // queries, n_results 100-750): k + k/4 + 32 -- at rank 750 of a 1 M-row store of unit vectors
  // about 140 rows lie within 1e-3 of the k-th score, 28 within 2e-4.
  int kc;
  if (k <= 64) {
    kc = 2 * k + 32;
    if (kc > KNN_MAX_CAND) kc = KNN_MAX_CAND;
  } else {
    kc = (k + k / 4 + 32 + 31) & ~31;
    if (kc < KNN_MAX_CAND) kc = KNN_MAX_CAND;
  }
  if (kc > m) kc = m;
  if (kc < k) kc = k < m ? k : m;
  return kc;
}

constexpr int KNN_DIRECT_MAX = 2048;  // banks up to this many rows: score matrix + register select
constexpr int KNN_MAX_STAGES = 32;

int knn_pool_cap_for(int kc) { return kc <= 448 ? KNN_POOL_CAP : KNN_POOL_CAP_BIG; }

int launch_pool_select(vitk_ctx* h, int cap, int n, int kc, int direct_rows, int ex_lo, int ex_hi,
                       int32_t* cand, int* overflow, cudaStream_t st) {
  Err& e = h->err;
  if (cap == KNN_POOL_CAP)
    knn_pool_select_kernel<KNN_POOL_CAP / KNN_POOL_KPL><<<n, KNN_POOL_CAP / KNN_POOL_KPL, 0, st>>>(
        h->knn_pool, h->knn_pool_n, h->knn_thresh, n, kc, direct_rows, ex_lo, ex_hi, cand, overflow);
  else
    knn_pool_select_kernel<KNN_POOL_CAP_BIG / KNN_POOL_KPL><<<n, KNN_POOL_CAP_BIG / KNN_POOL_KPL, 0, st>>>(
        h->knn_pool, h->knn_pool_n, h->knn_thresh, n, kc, direct_rows, ex_lo, ex_hi, cand, overflow);
  CK(e, cudaGetLastError());
  h->launches += 1;
  return 0;
}

// One stage of the large-store walk: bank rows [lo, hi) through the threshold-filter GEMM, then
// the pool select (flag = overflow slot).
int knn_stage(vitk_ctx* h, int n, int kc, int cap, int metric, int lo, int hi, int flag, bool direct,
              const KnnFilter& filt, cudaStream_t st, int32_t* cand = nullptr) {
  Err& e = h->err;
  const int d = h->bank_d;
  CUtensorMap ta, tb;
  if (tmap_operand(e, &ta, h->knn_qb, n, d, GEMM_BM) ||
      tmap_operand(e, &tb, h->bank_bf16 + (size_t)lo * d, hi - lo, d, 128))
    return 1;
  GemmParams gp{};
  gp.M = n; gp.N = hi - lo; gp.K = d;
  gp.bias = h->knn_pen_cur ? h->knn_pen_cur + lo : (metric == METRIC_L2 ? h->bank_sqnorm + lo : nullptr);
  gp.pool_thresh = h->knn_thresh; gp.pool = h->knn_pool; gp.pool_n = h->knn_pool_n;
  gp.pool_cap = cap; gp.col_base = lo;
  gp.pool_l2 = metric == METRIC_L2 ? 1 : (h->knn_pen_cur ? 2 : 0);
  gp.pool_direct = direct ? 1 : 0;
  gp.filt = filt;
  gp.dbg = g_gemm_dbg;   // developer timeline (tools/dev_store_timeline.py); null in production
  if (launch_gemm(e, filt.active ? EPI_THRESH_POOL_FILT : EPI_THRESH_POOL, ta, tb, ta, ta, gp, st)) return 1;
  h->launches += 1;
  // cand != null: this is the walk's last stage -- its select also writes the candidate lists, which
  // stand if no stage overflowed (the separate final select is then skipped)
  return launch_pool_select(h, cap, n, kc, direct ? hi - lo : -1, 0, 0, cand, h->knn_overflow + flag, st);
}

// Candidate rows (h->knn_cand [n, kc]) for a bank of more than KNN_DIRECT_MAX rows.  All stages
// are launched back to back; ONE stream synchronisation at the end reads the overflow flags.  A
// flagged stage (its pool was truncated; thresholds stay valid lower bounds, so later stages are
// unaffected) is redone in quarters with the final thresholds.
int knn_large_candidates(vitk_ctx* h, int n, int kc, int metric, const KnnFilter& filt, cudaStream_t st) {
  Err& e = h->err;
  const int m = h->bank_m;
  const int cap = knn_pool_cap_for(kc);
  if (n > h->knn_pool_cap_q || cap > h->knn_pool_cap_e) {
    CK(e, cudaStreamSynchronize(st));
    cudaFree(h->knn_pool); cudaFree(h->knn_pool_n); cudaFree(h->knn_thresh);
    h->knn_pool = nullptr; h->knn_pool_n = nullptr; h->knn_thresh = nullptr;
    const int cq = n > h->knn_pool_cap_q ? n : h->knn_pool_cap_q;
    const int ce = cap > h->knn_pool_cap_e ? cap : h->knn_pool_cap_e;
    h->knn_pool_cap_q = 0;
    h->knn_pool_cap_e = 0;
    if (dev_alloc(e, &h->knn_pool, (size_t)cq * ce) || dev_alloc(e, &h->knn_pool_n, (size_t)cq) ||
        dev_alloc(e, &h->knn_thresh, (size_t)cq))
      return 1;
    if (!h->knn_overflow && dev_alloc(e, &h->knn_overflow, (size_t)KNN_MAX_STAGES)) return 1;
    if (!h->knn_overflow_host)
      CK(e, cudaMallocHost(reinterpret_cast<void**>(&h->knn_overflow_host), KNN_MAX_STAGES * sizeof(int)));
    h->knn_pool_cap_q = cq;
    h->knn_pool_cap_e = ce;
  }
  CK(e, cudaMemsetAsync(h->knn_pool_n, 0, (size_t)n * sizeof(int), st));
  CK(e, cudaMemsetAsync(h->knn_thresh, 0, (size_t)n * sizeof(unsigned long long), st));
  // stage sizes: the first fills the empty pools (threshold 0: everything is appended), later
  // ones grow by g, chosen so the EXPECTED survivors per pool (kc * g on exchangeable rows) stay
  // 4x below the pool capacity
  int g = (cap - kc) / (4 * kc);
  g = g < 2 ? 2 : g > 8 ? 8 : g;
  struct Stage { int lo, hi; };
  std::vector<Stage> todo;
  for (int lo = 0, step = cap; lo < m;) {
    const long long want = (long long)lo + step;
    const int hi = want < m ? (int)want : m;
    todo.push_back({lo, hi});
    lo = hi;
    step = g * hi;
    if ((int)todo.size() == KNN_MAX_STAGES - 1) step = m;  // cannot happen for m < 2^31, g >= 2
  }
  bool first_pass = true;  // its first stage meets empty pools: filled positionally, no atomics
  while (!todo.empty()) {
    std::vector<Stage> next;
    for (size_t base = 0; base < todo.size(); base += KNN_MAX_STAGES) {
      const size_t cnt = todo.size() - base < (size_t)KNN_MAX_STAGES ? todo.size() - base : KNN_MAX_STAGES;
      CK(e, cudaMemsetAsync(h->knn_overflow, 0, KNN_MAX_STAGES * sizeof(int), st));
      for (size_t i = 0; i < cnt; ++i)
        if (knn_stage(h, n, kc, cap, metric, todo[base + i].lo, todo[base + i].hi, (int)i,
                      first_pass && base + i == 0, filt, st,
                      base + i + 1 == todo.size() ? h->knn_cand : nullptr))
          return 1;
      CK(e, cudaMemcpyAsync(h->knn_overflow_host, h->knn_overflow, KNN_MAX_STAGES * sizeof(int),
                            cudaMemcpyDeviceToHost, st));
      CK(e, cudaStreamSynchronize(st));
      for (size_t i = 0; i < cnt; ++i) {
        if (!h->knn_overflow_host[i]) continue;
        const Stage s = todo[base + i];
        const int rows = s.hi - s.lo;
        if (rows <= cap - kc) {
          e.set("vitk_knn_vote: candidate pool overflow in a %d-row stage (internal error)", rows);
          return 1;
        }
        h->knn_redo_stages += 1;
        // forget whatever the truncated stage left in the pools, keep the thresholds
        if (launch_pool_select(h, cap, n, kc, -1, s.lo, s.hi, nullptr, h->knn_overflow, st)) return 1;
        const int q4 = (rows + 3) / 4;
        for (int lo = s.lo; lo < s.hi; lo += q4) next.push_back({lo, lo + q4 < s.hi ? lo + q4 : s.hi});
      }
    }
    if (next.empty()) return 0;   // the last stage's select has already written the candidates
    todo.swap(next);
    first_pass = false;
  }
  // (only reached with an empty walk) final select: candidates for the fp32 re-rank
  return launch_pool_select(h, cap, n, kc, -1, 0, 0, h->knn_cand, h->knn_overflow, st);
}

int knn_vote_impl(vitk_ctx* h, const float* q_dev, int n, int k, int metric, int normalise_queries,
                  const uint8_t* row_mask_dev, const vitk_row_filter* rf, int32_t* idx_out,
                  float* score_out, int32_t* label_out, int32_t* counts_out, cudaStream_t st) {
  Err& e = h->err;
  if (h->bank_m == 0) { e.set("vitk_knn_vote: no bank set"); return 1; }
  if (n == 0) return 0;
  if (!q_dev || !idx_out || !score_out || n < 0) { e.set("vitk_knn_vote: bad argument"); return 1; }
  if (k < 1 || k > KNN_MAX_K) { e.set("vitk_knn_vote: k=%d outside [1,%d]", k, KNN_MAX_K); return 1; }
  if (metric != METRIC_COSINE && metric != METRIC_L2) { e.set("vitk_knn_vote: bad metric"); return 1; }
  KnnFilter filt{};
  if (rf != nullptr) {
    if ((rf->ne_col && !rf->q_ne) || (rf->eq_col && !rf->q_eq) || (rf->rng_col && (!rf->q_lo || !rf->q_hi)) ||
        ((rf->q_lo2 != nullptr) != (rf->q_hi2 != nullptr)) || (rf->q_lo2 && !rf->rng_col)) {
      e.set("vitk_knn_vote_filtered: every column needs its per-query operand(s)");
      return 1;
    }
    filt.ne_col = rf->ne_col; filt.q_ne = rf->q_ne; filt.eq_col = rf->eq_col; filt.q_eq = rf->q_eq;
    filt.rng_col = rf->rng_col; filt.q_lo = rf->q_lo; filt.q_hi = rf->q_hi;
    filt.q_lo2 = rf->q_lo2; filt.q_hi2 = rf->q_hi2; filt.rng_f32 = rf->rng_round_f32;
    filt.active = (rf->ne_col || rf->eq_col || rf->rng_col) ? 1 : 0;
  }
  CK(e, cudaSetDevice(h->device));
  const int m = h->bank_m, d = h->bank_d;
  if (d % 4 != 0 || (size_t)d * sizeof(float) > 28 * 1024) {
    e.set("vitk_knn_vote: d=%d must be a multiple of 4 and at most 7168", d);
    return 1;
  }
  const bool large = m > KNN_DIRECT_MAX;
  const int ld = large ? 4 : (m + 3) & ~3;  // TMA row pitch must be a multiple of 16 B
  const int kc = knn_candidates_for(k, m);
  if (n > h->knn_cap_q || ld > h->knn_cap_m || kc > h->knn_cap_kc) {
    CK(e, cudaStreamSynchronize(st));
    cudaFree(h->knn_q); cudaFree(h->knn_qb); cudaFree(h->knn_scores); cudaFree(h->knn_cand);
    h->knn_q = nullptr; h->knn_qb = nullptr; h->knn_scores = nullptr; h->knn_cand = nullptr;
    const int cq = n > h->knn_cap_q ? n : h->knn_cap_q;
    // a growing (mutable) bank: leave headroom so that a 16-row upsert does not reallocate
    int cm = h->knn_cap_m;
    if (ld > cm) {
      cm = ld + (ld >> 1);
      const int top = (KNN_DIRECT_MAX + 3) & ~3;
      if (cm > top) cm = top;
      if (cm < ld) cm = ld;
    }
    int ckc = h->knn_cap_kc > KNN_MAX_CAND ? h->knn_cap_kc : KNN_MAX_CAND;
    if (kc > ckc) ckc = kc;
    h->knn_cap_q = 0;
    if (dev_alloc(e, &h->knn_q, (size_t)cq * d) || dev_alloc(e, &h->knn_qb, (size_t)cq * d) ||
        dev_alloc(e, &h->knn_scores, (size_t)cq * cm) ||
        dev_alloc(e, &h->knn_cand, (size_t)cq * ckc))
      return 1;
    h->knn_cap_q = cq;
    h->knn_cap_m = cm;
    h->knn_cap_kc = ckc;
  }
  CK(e, launch_pdl(prep_queries_kernel, dim3((n + 7) / 8), dim3(256), 0, st, q_dev, h->knn_q, h->knn_qb, n, d,
                   normalise_queries));
  CK(e, cudaGetLastError());
  h->launches += 1;
  const float* pen = nullptr;
  if (row_mask_dev != nullptr) {
    if (m > h->knn_pen_cap) {
      CK(e, cudaStreamSynchronize(st));
      cudaFree(h->knn_pen);
      h->knn_pen = nullptr;
      h->knn_pen_cap = 0;
      const int cap = (m + (m >> 1) + 127) & ~127;
      if (dev_alloc(e, &h->knn_pen, (size_t)cap)) return 1;
      h->knn_pen_cap = cap;
    }
    knn_penalty_kernel<<<(m + 255) / 256, 256, 0, st>>>(row_mask_dev, h->bank_sqnorm, h->knn_pen, m, metric);
    CK(e, cudaGetLastError());
    h->launches += 1;
    pen = h->knn_pen;
  }
  if (large) {
    h->knn_pen_cur = pen;
    const int rc = knn_large_candidates(h, n, kc, metric, filt, st);
    h->knn_pen_cur = nullptr;
    if (rc) return 1;
  } else {
    CUtensorMap ta, tc;
    if (tmap_operand(e, &ta, h->knn_qb, n, d, GEMM_BM) ||
        tmap_output(e, &tc, h->knn_scores, n, ld, false))
      return 1;
    GemmParams gp{};
    gp.M = n; gp.N = m; gp.K = d;
    if (launch_gemm(e, EPI_F32, ta, h->tm_bank, tc, tc, gp, st)) return 1;
    const int blocks = (n + 3) / 4;
    const float* sc = h->knn_scores;
    const float* sq = h->bank_sqnorm;
    int32_t* cd = h->knn_cand;
    if (m <= 256) CK(e, launch_pdl(knn_candidates_reg_kernel<8>, dim3(blocks), dim3(128), 0, st, sc, ld, sq, n, m, kc, metric, cd, pen, filt));
    else if (m <= 512) CK(e, launch_pdl(knn_candidates_reg_kernel<16>, dim3(blocks), dim3(128), 0, st, sc, ld, sq, n, m, kc, metric, cd, pen, filt));
    else if (m <= 1024) CK(e, launch_pdl(knn_candidates_reg_kernel<32>, dim3(blocks), dim3(128), 0, st, sc, ld, sq, n, m, kc, metric, cd, pen, filt));
    else if (m <= 1536) CK(e, launch_pdl(knn_candidates_reg_kernel<48>, dim3(blocks), dim3(128), 0, st, sc, ld, sq, n, m, kc, metric, cd, pen, filt));
    else CK(e, launch_pdl(knn_candidates_reg_kernel<64>, dim3(blocks), dim3(128), 0, st, sc, ld, sq, n, m, kc, metric, cd, pen, filt));
    CK(e, cudaGetLastError());
    h->launches += 2;
  }
  const int32_t* lab = h->bank_classes ? h->bank_labels : nullptr;
  if (kc <= KNN_MAX_CAND)
    CK(e, launch_pdl(knn_rerank_vote_kernel, dim3(n), dim3(KNN_RERANK_THREADS), (size_t)d * sizeof(float), st,
                     h->knn_q, h->bank, lab, h->knn_cand, n, m, d, kc, k, metric, h->bank_classes, idx_out,
                     score_out, label_out, counts_out, row_mask_dev, filt));
  else
    knn_rerank_sort_kernel<<<n, 256, (size_t)d * sizeof(float), st>>>(
        h->knn_q, h->bank, lab, h->knn_cand, n, m, d, kc, k, metric, h->bank_classes, idx_out,
        score_out, label_out, counts_out, row_mask_dev, filt);
  CK(e, cudaGetLastError());
  h->launches += 1;
  return 0;
}
}  // namespace

extern "C" int vitk_knn_vote(vitk_handle h, const float* q_dev, int n, int k, int metric,
                             int normalise_queries, int32_t* idx_out, float* score_out,
                             int32_t* label_out, int32_t* counts_out, void* stream) {
  return vitk_knn_vote_filtered(h, q_dev, n, k, metric, normalise_queries, nullptr, nullptr, idx_out,
                                score_out, label_out, counts_out, stream);
}

extern "C" int vitk_knn_vote_masked(vitk_handle h, const float* q_dev, int n, int k, int metric,
                                    int normalise_queries, const uint8_t* row_mask_dev,
                                    int32_t* idx_out, float* score_out, int32_t* label_out,
                                    int32_t* counts_out, void* stream) {
  return vitk_knn_vote_filtered(h, q_dev, n, k, metric, normalise_queries, row_mask_dev, nullptr,
                                idx_out, score_out, label_out, counts_out, stream);
}

extern "C" int vitk_knn_vote_filtered(vitk_handle h, const float* q_dev, int n, int k, int metric,
                                      int normalise_queries, const uint8_t* row_mask_dev,
                                      const vitk_row_filter* filter, int32_t* idx_out, float* score_out,
                                      int32_t* label_out, int32_t* counts_out, void* stream) {
  if (!h) return 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (order_enter(h, st)) return 1;
  if (knn_vote_impl(h, q_dev, n, k, metric, normalise_queries, row_mask_dev, filter, idx_out, score_out,
                    label_out, counts_out, st))
    return 1;
  return order_leave(h, st);
}

extern "C" int vitk_op_merge_topk(const int32_t* idx_in_dev, const float* score_in_dev, int n, int c, int k,
                                  int descending, int32_t* idx_out_dev, float* score_out_dev, void* stream) {
  Err& e = g_op_err;
  auto fail = [&]() { g_create_error = e.msg; return 1; };
  if (n == 0) return 0;
  if (!idx_in_dev || !score_in_dev || !idx_out_dev || !score_out_dev || n < 0 || c < 1 || c > 4096 || k < 1) {
    e.set("vitk_op_merge_topk: need 1 <= c <= 4096 candidates per query and k >= 1 (c=%d k=%d)", c, k);
    return fail();
  }
  int n2 = 32;
  while (n2 < c) n2 <<= 1;
  knn_merge_topk_kernel<<<n, 256, (size_t)n2 * 8, reinterpret_cast<cudaStream_t>(stream)>>>(
      idx_in_dev, score_in_dev, n, c, k, descending, idx_out_dev, score_out_dev);
  if (cudaGetLastError() != cudaSuccess) { e.set("vitk_op_merge_topk: launch failed"); return fail(); }
  return 0;
}

// ------------------------------------------------------------------------------------------
// preprocessing front-end: antialiased bilinear uint8 resize (preprocess.cuh)
// ------------------------------------------------------------------------------------------
namespace {
struct HostAxis {
  std::vector<int> xmin, xsize;
  std::vector<short> w;
  int taps = 0, prec = 0;
};

// ATen UpSampleKernel.cpp: _compute_indices_min_size_weights_aa + _compute_index_ranges_int16_weights
// for the bilinear filter (interp_size 2, align_corners false), all in double.
HostAxis build_resize_axis(int in_size, int out_size) {
  HostAxis a;
  const double scale = (double)in_size / (double)out_size;
  const double support = scale >= 1.0 ? 1.0 * scale : 1.0;
  a.taps = (int)std::ceil(support) * 2 + 1;
  const double invscale = scale >= 1.0 ? 1.0 / scale : 1.0;
  a.xmin.resize(out_size);
  a.xsize.resize(out_size);
  std::vector<double> wd((size_t)out_size * a.taps, 0.0);
  double wt_max = 0.0;
  for (int i = 0; i < out_size; ++i) {
    const double center = scale * (i + 0.5);
    long long lo = (long long)(center - support + 0.5);
    if (lo < 0) lo = 0;
    long long hi = (long long)(center + support + 0.5);
    if (hi > in_size) hi = in_size;
    long long n = hi - lo;
    if (n < 0) n = 0;
    if (n > a.taps) n = a.taps;
    double tot = 0.0;
    double* w = wd.data() + (size_t)i * a.taps;
    for (long long j = 0; j < n; ++j) {
      double x = ((double)(j + lo) - center + 0.5) * invscale;
      if (x < 0.0) x = -x;
      w[j] = x < 1.0 ? 1.0 - x : 0.0;
      tot += w[j];
    }
    if (tot != 0.0)
      for (long long j = 0; j < n; ++j) {
        w[j] /= tot;
        if (w[j] > wt_max) wt_max = w[j];
      }
    a.xmin[i] = (int)lo;
    a.xsize[i] = (int)n;
  }
  int prec = 0;
  for (; prec < 22; ++prec) {
    const int next = (int)(0.5 + wt_max * (double)(1 << (prec + 1)));
    if (next >= (1 << 15)) break;
  }
  a.prec = prec;
  a.w.resize(wd.size());
  for (size_t i = 0; i < wd.size(); ++i)
    a.w[i] = (short)(wd[i] < 0 ? (int)(-0.5 + wd[i] * (double)(1 << prec)) : (int)(0.5 + wd[i] * (double)(1 << prec)));
  return a;
}

int get_resize_plan(vitk_ctx* h, int in_h, int in_w, vitk_ctx::ResizePlan** out) {
  Err& e = h->err;
  auto key = std::make_pair(in_h, in_w);
  auto it = h->resize_plans.find(key);
  if (it != h->resize_plans.end()) { *out = &it->second; return 0; }
  const int out_h = h->cfg.image_h, out_w = h->cfg.image_w, C = h->cfg.channels;
  HostAxis hx = build_resize_axis(in_w, out_w), vy = build_resize_axis(in_h, out_h);
  // one blob: hx.xmin | hx.xsize | vy.xmin | vy.xsize | hx.w | vy.w
  const size_t ints = (size_t)2 * out_w + 2 * out_h;
  const size_t bytes = ints * 4 + (hx.w.size() + vy.w.size()) * 2;
  std::vector<uint8_t> blob(bytes);
  int* ip = reinterpret_cast<int*>(blob.data());
  memcpy(ip, hx.xmin.data(), (size_t)out_w * 4);
  memcpy(ip + out_w, hx.xsize.data(), (size_t)out_w * 4);
  memcpy(ip + 2 * out_w, vy.xmin.data(), (size_t)out_h * 4);
  memcpy(ip + 2 * out_w + out_h, vy.xsize.data(), (size_t)out_h * 4);
  short* sp = reinterpret_cast<short*>(blob.data() + ints * 4);
  memcpy(sp, hx.w.data(), hx.w.size() * 2);
  memcpy(sp + hx.w.size(), vy.w.data(), vy.w.size() * 2);
  vitk_ctx::ResizePlan pl;
  uint8_t* dev = nullptr;
  if (dev_alloc(e, &dev, bytes)) return 1;
  CK(e, cudaMemcpy(dev, blob.data(), bytes, cudaMemcpyHostToDevice));
  pl.blob = dev;
  const int* dip = reinterpret_cast<const int*>(dev);
  const short* dsp = reinterpret_cast<const short*>(dev + ints * 4);
  pl.hx = ResizeAxis{dip, dip + out_w, dsp, hx.taps, hx.prec, in_w, out_w};
  pl.vy = ResizeAxis{dip + 2 * out_w, dip + 2 * out_w + out_h, dsp + hx.w.size(), vy.taps, vy.prec, in_h, out_h};
  // output rows per CTA: as many as keep the horizontally resized input rows within 96 KB
  const size_t row_bytes = (size_t)out_w * C;
  int ty = 16;
  int max_rows = 0;
  for (; ty >= 1; ty /= 2) {
    max_rows = 0;
    for (int y0 = 0; y0 < out_h; y0 += ty) {
      int r1 = 0;
      const int y1 = y0 + ty < out_h ? y0 + ty : out_h;
      for (int y = y0; y < y1; ++y) r1 = std::max(r1, vy.xmin[y] + vy.xsize[y]);
      max_rows = std::max(max_rows, r1 - vy.xmin[y0]);
    }
    if ((size_t)max_rows * row_bytes <= 96 * 1024) break;
  }
  if (ty < 1) { cudaFree(dev); e.set("resize: %dx%d -> %dx%d needs too much shared memory", in_h, in_w, out_h, out_w); return 1; }
  pl.ty = ty;
  pl.max_rows = max_rows;
  pl.smem = (size_t)max_rows * row_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    CK(e, cudaFuncSetAttribute(resize_u8_aa_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set = true;
  }
  *out = &h->resize_plans.emplace(key, pl).first->second;
  return 0;
}

int launch_resize(vitk_ctx* h, const uint8_t* in_dev, int n, int in_h, int in_w, uint8_t* out_dev,
                  cudaStream_t st) {
  Err& e = h->err;
  vitk_ctx::ResizePlan* pl = nullptr;
  if (get_resize_plan(h, in_h, in_w, &pl)) return 1;
  ResizeParams p;
  p.in = in_dev; p.out = out_dev; p.n = n; p.C = h->cfg.channels;
  p.hx = pl->hx; p.vy = pl->vy; p.ty = pl->ty; p.max_rows = pl->max_rows;
  dim3 grid((unsigned)((h->cfg.image_h + pl->ty - 1) / pl->ty), (unsigned)n, 1);
  resize_u8_aa_kernel<<<grid, 256, pl->smem, st>>>(p);
  CK(e, cudaGetLastError());
  h->launches += 1;
  return 0;
}
}  // namespace

extern "C" int vitk_preprocess_resize_u8(vitk_handle h, const uint8_t* frames_dev, int n, int in_h,
                                         int in_w, uint8_t* out_dev, void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (!frames_dev || !out_dev || n < 0 || in_h < 1 || in_w < 1) { e.set("vitk_preprocess_resize_u8: bad argument"); return 1; }
  if (n == 0) return 0;
  if (n > 65535) { e.set("vitk_preprocess_resize_u8: at most 65535 frames per call"); return 1; }
  CK(e, cudaSetDevice(h->device));
  return launch_resize(h, frames_dev, n, in_h, in_w, out_dev, reinterpret_cast<cudaStream_t>(stream));
}

namespace {
struct HostAreaAxis {
  std::vector<int> off, si;
  std::vector<float> alpha;
};
// OpenCV computeResizeAreaTab (imgproc/resize.cpp), scale = ssize / dsize >= 1, in double
HostAreaAxis build_area_axis(int ssize, int dsize) {
  HostAreaAxis a;
  const double scale = (double)ssize / dsize;
  a.off.push_back(0);
  for (int dx = 0; dx < dsize; ++dx) {
    const double fsx1 = dx * scale, fsx2 = fsx1 + scale;
    const double cell = std::min(scale, ssize - fsx1);
    int sx1 = (int)std::ceil(fsx1), sx2 = (int)std::floor(fsx2);
    sx2 = std::min(sx2, ssize - 1);
    sx1 = std::min(sx1, sx2);
    if (sx1 - fsx1 > 1e-3) { a.si.push_back(sx1 - 1); a.alpha.push_back((float)((sx1 - fsx1) / cell)); }
    for (int sx = sx1; sx < sx2; ++sx) { a.si.push_back(sx); a.alpha.push_back((float)(1.0 / cell)); }
    if (fsx2 - sx2 > 1e-3) {
      a.si.push_back(sx2);
      a.alpha.push_back((float)(std::min(std::min(fsx2 - sx2, 1.0), cell) / cell));
    }
    a.off.push_back((int)a.si.size());
  }
  return a;
}

int get_area_plan(vitk_ctx* h, int in_h, int in_w, vitk_ctx::AreaPlan** out) {
  Err& e = h->err;
  auto key = std::make_pair(in_h, in_w);
  auto it = h->area_plans.find(key);
  if (it != h->area_plans.end()) { *out = &it->second; return 0; }
  const int out_h = h->cfg.image_h, out_w = h->cfg.image_w;
  vitk_ctx::AreaPlan pl;
  if (in_w % out_w == 0 && in_h % out_h == 0) {  // both scales integer: ResizeAreaFast
    pl.fast_sx = in_w / out_w;
    pl.fast_sy = in_h / out_h;
  } else {
    HostAreaAxis ax = build_area_axis(in_w, out_w), ay = build_area_axis(in_h, out_h);
    const size_t nx = ax.si.size(), ny = ay.si.size();
    const size_t bytes = 4 * ((out_w + 1) + nx + nx + (out_h + 1) + ny + ny);
    std::vector<uint8_t> blob(bytes);
    int* ip = reinterpret_cast<int*>(blob.data());
    size_t o = 0;
    memcpy(ip + o, ax.off.data(), 4 * (size_t)(out_w + 1)); const size_t o_xoff = o; o += out_w + 1;
    memcpy(ip + o, ax.si.data(), 4 * nx); const size_t o_xsi = o; o += nx;
    memcpy(ip + o, ax.alpha.data(), 4 * nx); const size_t o_xa = o; o += nx;
    memcpy(ip + o, ay.off.data(), 4 * (size_t)(out_h + 1)); const size_t o_yoff = o; o += out_h + 1;
    memcpy(ip + o, ay.si.data(), 4 * ny); const size_t o_ysi = o; o += ny;
    memcpy(ip + o, ay.alpha.data(), 4 * ny); const size_t o_ya = o; o += ny;
    uint8_t* dev = nullptr;
    if (dev_alloc(e, &dev, bytes)) return 1;
    CK(e, cudaMemcpy(dev, blob.data(), bytes, cudaMemcpyHostToDevice));
    pl.blob = dev;
    const int* dip = reinterpret_cast<const int*>(dev);
    const float* dfp = reinterpret_cast<const float*>(dev);
    pl.x = AreaAxis{dip + o_xoff, dip + o_xsi, dfp + o_xa};
    pl.y = AreaAxis{dip + o_yoff, dip + o_ysi, dfp + o_ya};
  }
  *out = &h->area_plans.emplace(key, pl).first->second;
  return 0;
}
}  // namespace

namespace {
int launch_resize_area(vitk_ctx* h, const uint8_t* frames_dev, int n, int in_h, int in_w, uint8_t* out_dev,
                       cudaStream_t st) {
  Err& e = h->err;
  if (in_h < h->cfg.image_h || in_w < h->cfg.image_w) {
    e.set("vitk_preprocess_resize_area_u8: INTER_AREA is built for shrinking only (%dx%d -> %dx%d)", in_w, in_h,
          h->cfg.image_w, h->cfg.image_h);
    return 1;
  }
  vitk_ctx::AreaPlan* pl = nullptr;
  if (get_area_plan(h, in_h, in_w, &pl)) return 1;
  AreaParams p;
  p.in = frames_dev; p.out = out_dev; p.x = pl->x; p.y = pl->y;
  p.n = n; p.C = h->cfg.channels; p.in_h = in_h; p.in_w = in_w;
  p.out_h = h->cfg.image_h; p.out_w = h->cfg.image_w;
  p.fast_sx = pl->fast_sx; p.fast_sy = pl->fast_sy;
  p.fast_scale = pl->fast_sx > 0 ? 1.f / (float)(pl->fast_sx * pl->fast_sy) : 0.f;
  resize_area_u8_kernel<<<dim3((unsigned)p.out_h, (unsigned)n, 1), 256, 0, st>>>(p);
  CK(e, cudaGetLastError());
  h->launches += 1;
  return 0;
}

// OpenCV resizeGeneric_ coefficient tables for INTER_LINEAR on 8-bit data (float arithmetic and
// cvRound exactly as OpenCV; oracle/preprocess_ref.linear_tab is the numpy twin)
void build_linear_axis(int ssize, int dsize, bool clamp_fraction, std::vector<int>& idx, std::vector<int>& coef) {
  const double scale = 1.0 / ((double)dsize / (double)ssize);
  idx.resize(dsize);
  coef.resize(2 * (size_t)dsize);
  for (int d = 0; d < dsize; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int sx = (int)std::floor(f);
    f -= (float)sx;
    if (clamp_fraction) {
      if (sx < 0) { f = 0.f; sx = 0; }
      if (sx >= ssize - 1) { f = 0.f; sx = ssize - 1; }
    }
    idx[d] = sx;
    coef[2 * d] = (int)lrintf((1.f - f) * 2048.f);
    coef[2 * d + 1] = (int)lrintf(f * 2048.f);
  }
}

int launch_resize_linear(vitk_ctx* h, const uint8_t* frames_dev, int n, int in_h, int in_w, uint8_t* out_dev,
                         cudaStream_t st) {
  Err& e = h->err;
  const int out_h = h->cfg.image_h, out_w = h->cfg.image_w;
  auto key = std::make_pair(in_h, in_w);
  auto it = h->linear_plans.find(key);
  if (it == h->linear_plans.end()) {
    std::vector<int> xs, xa, ys, ya;
    build_linear_axis(in_w, out_w, true, xs, xa);    // x: fraction zeroed at the borders
    build_linear_axis(in_h, out_h, false, ys, ya);   // y: rows are clamped instead
    std::vector<int> blob;
    blob.insert(blob.end(), xs.begin(), xs.end());
    blob.insert(blob.end(), xa.begin(), xa.end());
    blob.insert(blob.end(), ys.begin(), ys.end());
    blob.insert(blob.end(), ya.begin(), ya.end());
    int* dev = nullptr;
    if (dev_alloc(e, &dev, blob.size())) return 1;
    CK(e, cudaMemcpy(dev, blob.data(), blob.size() * sizeof(int), cudaMemcpyHostToDevice));
    vitk_ctx::LinearPlan pl;
    pl.blob = dev;
    pl.xs = dev; pl.xa = dev + out_w; pl.ys = dev + 3 * (size_t)out_w; pl.ya = dev + 3 * (size_t)out_w + out_h;
    it = h->linear_plans.emplace(key, pl).first;
  }
  LinearParams p;
  p.in = frames_dev; p.out = out_dev;
  p.xs = it->second.xs; p.xa = it->second.xa; p.ys = it->second.ys; p.ya = it->second.ya;
  p.n = n; p.C = h->cfg.channels; p.in_h = in_h; p.in_w = in_w; p.out_h = out_h; p.out_w = out_w;
  resize_linear_u8_kernel<<<dim3((unsigned)out_h, (unsigned)n, 1), 256, 0, st>>>(p);
  CK(e, cudaGetLastError());
  h->launches += 1;
  return 0;
}

// interp: VITK_RESIZE_BILINEAR_AA (the HF processor), VITK_RESIZE_AREA, VITK_RESIZE_LINEAR (cv2)
int launch_resize_interp(vitk_ctx* h, int interp, const uint8_t* in_dev, int n, int in_h, int in_w,
                         uint8_t* out_dev, cudaStream_t st) {
  switch (interp) {
    case VITK_RESIZE_BILINEAR_AA: return launch_resize(h, in_dev, n, in_h, in_w, out_dev, st);
    case VITK_RESIZE_AREA: return launch_resize_area(h, in_dev, n, in_h, in_w, out_dev, st);
    case VITK_RESIZE_LINEAR: return launch_resize_linear(h, in_dev, n, in_h, in_w, out_dev, st);
  }
  h->err.set("unknown resize interpolation %d", interp);
  return 1;
}
}  // namespace

extern "C" int vitk_preprocess_resize_area_u8(vitk_handle h, const uint8_t* frames_dev, int n, int in_h,
                                              int in_w, uint8_t* out_dev, void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (n == 0) return 0;
  if (!frames_dev || !out_dev || n < 0 || in_h < 1 || in_w < 1) { e.set("vitk_preprocess_resize_area_u8: bad argument"); return 1; }
  if (n > 65535) { e.set("vitk_preprocess_resize_area_u8: at most 65535 frames per call"); return 1; }
  CK(e, cudaSetDevice(h->device));
  return launch_resize_area(h, frames_dev, n, in_h, in_w, out_dev, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int vitk_preprocess_resize_linear_u8(vitk_handle h, const uint8_t* frames_dev, int n, int in_h,
                                                int in_w, uint8_t* out_dev, void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (n == 0) return 0;
  if (!frames_dev || !out_dev || n < 0 || in_h < 1 || in_w < 1) { e.set("vitk_preprocess_resize_linear_u8: bad argument"); return 1; }
  if (n > 65535) { e.set("vitk_preprocess_resize_linear_u8: at most 65535 frames per call"); return 1; }
  CK(e, cudaSetDevice(h->device));
  return launch_resize_linear(h, frames_dev, n, in_h, in_w, out_dev, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int vitk_embed_resize_host_u8(vitk_handle h, const uint8_t* frames_host, int n, int in_h,
                                         int in_w, float* emb_host, int l2_normalise) {
  if (!h) return 1;
  Err& e = h->err;
  if (!h->finalized) { e.set("vitk_embed_resize_host_u8: weights not finalized"); return 1; }
  if (n == 0) return 0;
  if (!frames_host || !emb_host || n < 0 || in_h < 1 || in_w < 1) { e.set("vitk_embed_resize_host_u8: bad argument"); return 1; }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = h->own_stream;
  if (order_enter(h, st)) return 1;
  const size_t fb = (size_t)in_h * in_w * h->cfg.channels;
  const int mb = h->cfg.max_batch;
  if (h->raw_stage_bytes < fb * mb) {
    CK(e, cudaStreamSynchronize(st));
    cudaFree(h->raw_stage);
    h->raw_stage = nullptr;
    h->raw_stage_bytes = 0;
    if (dev_alloc(e, &h->raw_stage, fb * mb)) return 1;
    h->raw_stage_bytes = fb * mb;
  }
  for (int off = 0; off < n; off += mb) {
    const int cn = n - off < mb ? n - off : mb;
    CK(e, cudaMemcpyAsync(h->raw_stage, frames_host + (size_t)off * fb, (size_t)cn * fb,
                          cudaMemcpyHostToDevice, st));
    if (launch_resize(h, h->raw_stage, cn, in_h, in_w, h->pix_stage, st)) return 1;
    if (forward_chunk(h, h->pix_stage, PIX_U8_NHWC, cn, h->emb, nullptr, l2_normalise, st)) return 1;
    CK(e, cudaMemcpyAsync(emb_host + (size_t)off * h->D, h->emb, (size_t)cn * h->D * 4,
                          cudaMemcpyDeviceToHost, st));
  }
  if (order_leave(h, st)) return 1;
  CK(e, cudaStreamSynchronize(st));
  return 0;
}

// ------------------------------------------------------------------------------------------
// vote -> HMM smoothing (hmm.py:5-135)
// ------------------------------------------------------------------------------------------
namespace {
HmmParams hmm_params() {
  HmmParams hp;
  hp.log_stay = std::log(0.985);
  hp.log_to_none = std::log(0.015);
  hp.log_from_none = std::log(0.15);
  hp.log_none_none = std::log(0.7);
  hp.log_third = std::log(1.0 / 3.0);
  return hp;
}

int hmm_prepare(vitk_ctx* h, const int32_t* seq_offsets_host, int n_seq, int* total, cudaStream_t st) {
  Err& e = h->err;
  if (!seq_offsets_host || n_seq < 1) { e.set("vitk_hmm: need at least one sequence"); return 1; }
  if (seq_offsets_host[0] != 0) { e.set("vitk_hmm: seq_offsets[0] must be 0"); return 1; }
  for (int i = 0; i < n_seq; ++i)
    if (seq_offsets_host[i + 1] < seq_offsets_host[i]) { e.set("vitk_hmm: seq_offsets must not decrease"); return 1; }
  const int T = seq_offsets_host[n_seq];
  if (T > h->hmm_cap_t || n_seq > h->hmm_cap_seq) {
    CK(e, cudaStreamSynchronize(st));
    cudaFree(h->hmm_loge); cudaFree(h->hmm_bp); cudaFree(h->hmm_offsets);
    h->hmm_loge = nullptr; h->hmm_bp = nullptr; h->hmm_offsets = nullptr;
    h->hmm_cap_t = h->hmm_cap_seq = 0;
    const int ct = T > 1 ? T : 1;
    if (dev_alloc(e, &h->hmm_loge, (size_t)ct * 3) || dev_alloc(e, &h->hmm_bp, (size_t)ct * 3) ||
        dev_alloc(e, &h->hmm_offsets, (size_t)n_seq + 1))
      return 1;
    h->hmm_cap_t = ct;
    h->hmm_cap_seq = n_seq;
  }
  CK(e, cudaMemcpyAsync(h->hmm_offsets, seq_offsets_host, ((size_t)n_seq + 1) * sizeof(int),
                        cudaMemcpyHostToDevice, st));
  *total = T;
  return 0;
}
// emission table on the host: std::log is the libm CPython's math.log calls
int hmm_ensure_table(vitk_ctx* h, int k, cudaStream_t st) {
  Err& e = h->err;
  if (h->hmm_cap_k == k) return 0;
  std::vector<double> tab(2 * ((size_t)k + 1));
  for (int c = 0; c <= k; ++c) {
    const double p = (double)c / (double)k;
    tab[c] = std::log(p + 0.000001);                       // hmm.py:36-38
    tab[(size_t)k + 1 + c] = std::log(c == 0 ? 0.000001 : p);  // hmm.py:45-50
  }
  CK(e, cudaDeviceSynchronize());   // a streamed decode on another stream may still read the old table
  cudaFree(h->hmm_table);
  h->hmm_table = nullptr;
  if (dev_alloc(e, &h->hmm_table, tab.size())) return 1;
  CK(e, cudaMemcpy(h->hmm_table, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice));
  h->hmm_cap_k = k;
  return 0;
}
}  // namespace

// ---- streamed form: a sequence whose vote counts arrive segment by segment, in time order -------
extern "C" int vitk_hmm_stream_begin(vitk_handle h, long long capacity_frames) {
  if (!h) return 1;
  Err& e = h->err;
  if (capacity_frames < 1) { e.set("vitk_hmm_stream_begin: capacity must be >= 1 frame"); return 1; }
  CK(e, cudaSetDevice(h->device));
  if (capacity_frames > h->hs_cap) {
    CK(e, cudaDeviceSynchronize());
    cudaFree(h->hs_bp);
    h->hs_bp = nullptr;
    h->hs_cap = 0;
    if (dev_alloc(e, &h->hs_bp, (size_t)capacity_frames * 3)) return 1;
    h->hs_cap = capacity_frames;
  }
  if (!h->hs_carry && dev_alloc(e, &h->hs_carry, (size_t)8)) return 1;
  h->hs_limit = capacity_frames;
  h->hs_t = 0;
  return 0;
}

extern "C" int vitk_hmm_stream_append_counts(vitk_handle h, const int32_t* counts_dev, int n, int k,
                                             void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (h->hs_t < 0) { e.set("vitk_hmm_stream_append_counts: no stream open (call vitk_hmm_stream_begin)"); return 1; }
  if (n == 0) return 0;
  if (!counts_dev || n < 0 || k < 1 || k > 4096) { e.set("vitk_hmm_stream_append_counts: bad argument"); return 1; }
  if (h->hs_t + n > h->hs_limit) {
    e.set("vitk_hmm_stream_append_counts: %lld + %d frames exceed the capacity %lld given to begin", h->hs_t, n, h->hs_limit);
    return 1;
  }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (hmm_ensure_table(h, k, st)) return 1;
  if (n > h->hs_seg_cap) {
    CK(e, cudaDeviceSynchronize());
    cudaFree(h->hs_loge);
    h->hs_loge = nullptr;
    h->hs_seg_cap = 0;
    const int cap = n < 4096 ? 4096 : n;
    if (dev_alloc(e, &h->hs_loge, (size_t)cap * 3)) return 1;
    h->hs_seg_cap = cap;
  }
  const int blocks = (n * 3 + 255) / 256;
  hmm_emission_segment_kernel<<<blocks < 64 ? blocks : 64, 256, 0, st>>>(counts_dev, h->hs_t, n, k, h->hmm_table,
                                                                     h->hs_loge);
  CK(e, cudaGetLastError());
  hmm_viterbi_segment_kernel<<<1, 32, 0, st>>>(h->hs_loge, h->hs_t, n, hmm_params(), h->hs_carry, h->hs_bp);
  CK(e, cudaGetLastError());
  h->hs_t += n;
  h->launches += 2;
  return 0;
}

extern "C" int vitk_hmm_stream_finish(vitk_handle h, int32_t* states_out_dev, long long* frames_out,
                                      void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (h->hs_t < 0) { e.set("vitk_hmm_stream_finish: no stream open"); return 1; }
  if (frames_out) *frames_out = h->hs_t;
  const long long T = h->hs_t;
  h->hs_t = -1;
  if (T == 0) return 0;
  if (!states_out_dev) { e.set("vitk_hmm_stream_finish: states_out is NULL"); return 1; }
  CK(e, cudaSetDevice(h->device));
  hmm_backtrace_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(h->hs_bp, T, h->hs_carry, states_out_dev);
  CK(e, cudaGetLastError());
  h->launches += 1;
  return 0;
}

extern "C" int vitk_hmm_decode_counts(vitk_handle h, const int32_t* counts_dev, int k,
                                      const int32_t* seq_offsets_host, int n_seq,
                                      int32_t* states_out_dev, double* dp_out_dev, void* stream) {
  if (!h) return 1;
  Err& e = h->err;
  if (seq_offsets_host && n_seq >= 1 && seq_offsets_host[n_seq] == 0) return 0;  // no frames at all
  if (!counts_dev || !states_out_dev || k < 1 || k > 4096) { e.set("vitk_hmm_decode_counts: bad argument"); return 1; }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (order_enter(h, st)) return 1;
  int T = 0;
  if (hmm_prepare(h, seq_offsets_host, n_seq, &T, st)) return 1;
  if (T == 0) return 0;
  if (hmm_ensure_table(h, k, st)) return 1;
  hmm_emission_from_counts_kernel<<<n_seq, 256, 0, st>>>(counts_dev, h->hmm_offsets, n_seq, k,
                                                         h->hmm_table, h->hmm_loge);
  CK(e, cudaGetLastError());
  hmm_viterbi_kernel<<<n_seq, 32, 0, st>>>(h->hmm_loge, h->hmm_offsets, n_seq, hmm_params(), h->hmm_bp,
                                           states_out_dev, dp_out_dev);
  CK(e, cudaGetLastError());
  if (order_leave(h, st)) return 1;
  h->launches += 2;
  return 0;
}

extern "C" int vitk_hmm_decode_probs_host(vitk_handle h, const double* probs_host,
                                          const int32_t* seq_offsets_host, int n_seq,
                                          int32_t* states_out_host, double* dp_out_host) {
  if (!h) return 1;
  Err& e = h->err;
  if (!probs_host || !states_out_host) { e.set("vitk_hmm_decode_probs_host: bad argument"); return 1; }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = h->own_stream;
  int T = 0;
  if (hmm_prepare(h, seq_offsets_host, n_seq, &T, st)) return 1;
  if (T == 0) return 0;
  std::vector<double> loge((size_t)T * 3);
  for (int sq = 0; sq < n_seq; ++sq) {
    for (int f = seq_offsets_host[sq]; f < seq_offsets_host[sq + 1]; ++f) {
      const bool first = f == seq_offsets_host[sq];
      for (int j = 0; j < 3; ++j) {
        const double p = probs_host[(size_t)f * 3 + j];
        loge[(size_t)f * 3 + j] = first ? std::log(p + 0.000001) : std::log(p == 0.0 ? 0.000001 : p);
      }
    }
  }
  int32_t* states_dev = nullptr;
  double* dp_dev = nullptr;
  if (dev_alloc(e, &states_dev, (size_t)T)) return 1;
  if (dp_out_host && dev_alloc(e, &dp_dev, (size_t)T * 3)) { cudaFree(states_dev); return 1; }
  int rc = 0;
  do {
    if (cudaMemcpyAsync(h->hmm_loge, loge.data(), loge.size() * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess) { rc = 1; break; }
    hmm_viterbi_kernel<<<n_seq, 32, 0, st>>>(h->hmm_loge, h->hmm_offsets, n_seq, hmm_params(), h->hmm_bp,
                                             states_dev, dp_dev);
    if (cudaGetLastError() != cudaSuccess) { rc = 1; break; }
    h->launches += 1;
    if (cudaMemcpyAsync(states_out_host, states_dev, (size_t)T * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) { rc = 1; break; }
    if (dp_out_host && cudaMemcpyAsync(dp_out_host, dp_dev, (size_t)T * 24, cudaMemcpyDeviceToHost, st) != cudaSuccess) { rc = 1; break; }
    if (cudaStreamSynchronize(st) != cudaSuccess) { rc = 1; break; }
  } while (false);
  if (rc) e.set("vitk_hmm_decode_probs_host: CUDA error: %s", cudaGetErrorString(cudaGetLastError()));
  cudaFree(states_dev);
  cudaFree(dp_dev);
  return rc;
}

// ------------------------------------------------------------------------------------------
// Viterbi decode that allocates nothing (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving
// loops).  The scratch vitk_hmm_decode_counts keeps in per-handle cudaMalloc'd buffers -- the
// log-emission table [T,3] f64, the backpointers [T,3] i8 and the sequence offsets -- lives in ONE
// file-scope device array of fixed size that every GPU loading the library holds until exit.
// ------------------------------------------------------------------------------------------
#define VITK_HMM_STATIC_MAX_FRAMES_ 131072
#define VITK_HMM_STATIC_MAX_K_ 128
namespace {
constexpr size_t kHmmStaticLogeBytes = (size_t)VITK_HMM_STATIC_MAX_FRAMES_ * 3 * sizeof(double);
constexpr size_t kHmmStaticBpBytes = (size_t)VITK_HMM_STATIC_MAX_FRAMES_ * 3;
constexpr size_t kHmmStaticBytes = kHmmStaticLogeBytes + kHmmStaticBpBytes + 256;  // + offsets {0, T}
}  // namespace
static __device__ __align__(256) unsigned char g_hmm_static_scratch[kHmmStaticBytes];  // 3.4 MB

namespace {
// the emission table travels as a kernel argument (2 * (k + 1) doubles <= 2 064 B): filled on the
// host with the libm log the reference's math.log calls, and no copy that would need a staging buffer
struct HmmStaticTable { double v[2 * (VITK_HMM_STATIC_MAX_K_ + 1)]; };

__global__ void __launch_bounds__(256)
hmm_emission_static_kernel(const int32_t* __restrict__ counts, int n, int k, const HmmStaticTable tab,
                           double* __restrict__ loge, int* __restrict__ offsets) {
  if (blockIdx.x == 0 && threadIdx.x == 0) { offsets[0] = 0; offsets[1] = n; }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)n * 3;
       i += (long long)gridDim.x * blockDim.x) {
    int c = counts[i];
    c = c < 0 ? 0 : (c > k ? k : c);
    loge[i] = tab.v[(i < 3 ? 0 : k + 1) + c];   // same two rows as hmm_emission_from_counts_kernel
  }
}

// device address of the scratch array, looked up once per device with that device current (in
// host code the array's NAME is a host-side stand-in, not a device pointer)
constexpr int kMaxDevices = 64;
std::mutex g_hmm_static_mu;
unsigned char* g_hmm_static_ptr[kMaxDevices] = {};
}  // namespace

extern "C" int vitk_hmm_decode_counts_static(const int32_t* counts_dev, int n_frames, int k,
                                             int32_t* states_out_dev, void* stream) {
  Err& e = g_op_err;
  auto fail = [&]() { g_create_error = e.msg; return 1; };
  if (n_frames == 0) return 0;
  if (!counts_dev || !states_out_dev || n_frames < 0) { e.set("vitk_hmm_decode_counts_static: bad argument"); return fail(); }
  if (n_frames > VITK_HMM_STATIC_MAX_FRAMES_) {
    e.set("vitk_hmm_decode_counts_static: %d frames exceed the static scratch (at most %d frames per call; "
          "use vitk_hmm_decode_counts or the streamed form)", n_frames, VITK_HMM_STATIC_MAX_FRAMES_);
    return fail();
  }
  if (k < 1 || k > VITK_HMM_STATIC_MAX_K_) {
    e.set("vitk_hmm_decode_counts_static: k = %d outside 1..%d", k, VITK_HMM_STATIC_MAX_K_);
    return fail();
  }
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    e.set("vitk_hmm_decode_counts_static: no current CUDA device");
    return fail();
  }
  unsigned char* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_hmm_static_mu);
    if (!g_hmm_static_ptr[dev]) {
      void* sym = nullptr;
      const cudaError_t rc = cudaGetSymbolAddress(&sym, g_hmm_static_scratch);
      if (rc != cudaSuccess || !sym) {
        e.set("vitk_hmm_decode_counts_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(rc));
        return fail();
      }
      g_hmm_static_ptr[dev] = static_cast<unsigned char*>(sym);
    }
    base = g_hmm_static_ptr[dev];
  }
  double* loge = reinterpret_cast<double*>(base);
  int8_t* bp = reinterpret_cast<int8_t*>(base + kHmmStaticLogeBytes);
  int* offsets = reinterpret_cast<int*>(base + kHmmStaticLogeBytes + kHmmStaticBpBytes);
  HmmStaticTable tab;
  for (int c = 0; c <= k; ++c) {
    const double pr = (double)c / (double)k;
    tab.v[c] = std::log(pr + 0.000001);                        // hmm.py:36-38
    tab.v[k + 1 + c] = std::log(c == 0 ? 0.000001 : pr);      // hmm.py:45-50
  }
  for (int c = 2 * (k + 1); c < 2 * (VITK_HMM_STATIC_MAX_K_ + 1); ++c) tab.v[c] = 0.0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // nothing in the scratch has to start at zero: every entry the Viterbi kernel reads (emissions
  // [0, 3T), offsets, backpointers of columns it has written) is rewritten by this call first
  const int blocks = (n_frames * 3 + 255) / 256;
  hmm_emission_static_kernel<<<blocks < 64 ? blocks : 64, 256, 0, st>>>(counts_dev, n_frames, k, tab, loge, offsets);
  if (cudaGetLastError() != cudaSuccess) { e.set("vitk_hmm_decode_counts_static: emission launch failed"); return fail(); }
  hmm_viterbi_kernel<<<1, 32, 0, st>>>(loge, offsets, 1, hmm_params(), bp, states_out_dev, nullptr);
  if (cudaGetLastError() != cudaSuccess) { e.set("vitk_hmm_decode_counts_static: viterbi launch failed"); return fail(); }
  return 0;
}

// ------------------------------------------------------------------------------------------
// operator-level entry points
// ------------------------------------------------------------------------------------------

namespace {
int op_gemm_common(const void* a, const void* w, void* out, int M, int N, int K, int epi,
                   const float* bias, const float* c1, const float* stats, int stats_parts,
                   float ln_eps, const float* resid, void* xb_out, float* stats_out,
                   const float* table, int table_rows, void* stream) {
  Err& e = g_op_err;
  auto fail = [&]() { g_create_error = e.msg; return 1; };
  if (K < GEMM_BK || K % GEMM_BK) {
    e.set("vitk_op_gemm: K=%d must be a positive multiple of %d", K, GEMM_BK);
    return fail();
  }
  if (M < 1 || N < 4 || N % 4) { e.set("vitk_op_gemm: bad shape M=%d N=%d", M, N); return fail(); }
  const bool is_resid = epi == EPI_BIAS_RESID_F32 || epi == EPI_BIAS_RESID_BF16;
  const bool bf16_out = epi == EPI_BIAS_BF16 || epi == EPI_BIAS_GELU_BF16 || epi == EPI_BIAS_GELUT_BF16 ||
                        epi == EPI_BIAS_RESID_BF16 || epi == EPI_TABLE_BF16;
  if (epi == EPI_THRESH_POOL || epi == EPI_THRESH_POOL_FILT) { e.set("vitk_op_gemm: epilogue %d is internal to the k-NN path", epi); return fail(); }
  if (bf16_out && N % 8) { e.set("vitk_op_gemm: bf16 output needs N %% 8 == 0"); return fail(); }
  if ((bf16_out || is_resid) && epi != EPI_TABLE_BF16 && !bias) { e.set("vitk_op_gemm: bias required"); return fail(); }
  if ((epi == EPI_TABLE_F32 || epi == EPI_TABLE_BF16) && !table) { e.set("vitk_op_gemm: table required"); return fail(); }
  if (is_resid && !resid) { e.set("vitk_op_gemm: resid required"); return fail(); }
  CUtensorMap ta, tb, tc, tr;
  GemmParams p{};
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.table = table;
  p.table_rows = table_rows > 0 ? table_rows : 1;
  if (stats) {
    p.c1 = c1; p.stats = stats; p.stats_parts = stats_parts; p.stats_stride = 2LL * stats_parts;
    p.inv_d = 1.0f / K; p.eps = ln_eps;
  }
  if (is_resid || epi == EPI_TABLE_BF16) {
    if (epi == EPI_BIAS_RESID_F32) { p.xb = reinterpret_cast<__nv_bfloat16*>(xb_out); p.xb_ld = N; }
    p.stats_out = stats_out; p.stats_out_stride = 2LL * (2 * ((N + 255) / 256));
  }
  int rc = tmap_operand(e, &ta, a, M, K, GEMM_BM) || tmap_operand(e, &tb, w, N, K, 128) ||
           tmap_output(e, &tc, out, M, N, bf16_out);
  if (!rc) {
    if (is_resid && static_cast<const void*>(resid) != out)
      rc = tmap_output(e, &tr, const_cast<float*>(resid), M, N, epi == EPI_BIAS_RESID_BF16);
    else
      tr = tc;
  }
  p.dbg = g_gemm_dbg;
  rc = rc || launch_gemm(e, epi, ta, tb, tc, tr, p, reinterpret_cast<cudaStream_t>(stream));
  if (rc) return fail();
  return 0;
}
}  // namespace

extern "C" int vitk_op_gemm(const void* a, const void* w, void* out, int M, int N, int K, int epi,
                            const float* bias, const float* resid, const float* table,
                            int table_rows, void* stream) {
  return op_gemm_common(a, w, out, M, N, K, epi, bias, nullptr, nullptr, 0, 0.f, resid, nullptr,
                        nullptr, table, table_rows, stream);
}

extern "C" int vitk_op_gemm_fused(const void* a, const void* w, void* out, int M, int N, int K,
                                  int epi, const float* bias, const float* c1, const float* stats,
                                  int stats_parts, float ln_eps, const float* resid, void* xb_out,
                                  float* stats_out, void* stream) {
  return op_gemm_common(a, w, out, M, N, K, epi, bias, c1, stats, stats_parts, ln_eps, resid, xb_out,
                        stats_out, nullptr, 1, stream);
}

extern "C" int vitk_op_rowstats(const float* x, void* xb, float* stats, int rows, int d, int parts,
                                void* stream) {
  Err& e = g_op_err;
  int rc = launch_rowstats(e, x, reinterpret_cast<__nv_bfloat16*>(xb), stats, rows, d, parts,
                           reinterpret_cast<cudaStream_t>(stream));
  if (rc) g_create_error = e.msg;
  return rc;
}

extern "C" int vitk_op_layernorm(const float* x, const float* gamma, const float* beta, void* y,
                                 int rows, int d, float eps, void* stream) {
  Err& e = g_op_err;
  int rc = launch_layernorm(e, x, gamma, beta, reinterpret_cast<__nv_bfloat16*>(y), rows, d, eps,
                            reinterpret_cast<cudaStream_t>(stream));
  if (rc) g_create_error = e.msg;
  return rc;
}

extern "C" int vitk_op_attention(const void* qkv, void* out, int frames, int S, int heads,
                                 void* stream) {
  Err& e = g_op_err;
  const int D = heads * 64;
  CUtensorMap tq, tkv, to;
  int rc = (S < 1 || S > ATT_MAX_S) ? (e.set("vitk_op_attention: S outside [1,%d]", ATT_MAX_S), 1) : 0;
  rc = rc || tmap_tokens(e, &tq, qkv, frames, S, 3 * D, 128) ||
       tmap_tokens(e, &tkv, qkv, frames, S, 3 * D, attn_smem_layout(S).kv_box) ||
       tmap_tokens(e, &to, out, frames, S, D, 128) ||
       launch_attention(e, tq, tkv, to, frames, S, heads, reinterpret_cast<cudaStream_t>(stream));
  if (rc) g_create_error = e.msg;
  return rc;
}

// ---- ChunkEncoder pieces (chunk.cuh) ----------------------------------------------------------
extern "C" int vitk_op_attention_small(const void* qkv, void* out, float* probs_out, int n_seq, int S,
                                       int heads, int d_head, void* stream) {
  Err& e = g_op_err;
  auto fail = [&]() { g_create_error = e.msg; return 1; };
  if (n_seq == 0) return 0;
  if (!qkv || !out || n_seq < 0 || S < 1 || S > SATT_MAX_S || heads < 1 || d_head < 32 ||
      d_head > SATT_MAX_DH || d_head % 32 != 0) {
    e.set("vitk_op_attention_small: need 1 <= S <= %d and d_head a multiple of 32 in [32,%d] "
          "(S=%d d_head=%d heads=%d)", SATT_MAX_S, SATT_MAX_DH, S, d_head, heads);
    return fail();
  }
  const size_t smem = ((size_t)S * (d_head + 1) + 2 * (size_t)S * d_head) * sizeof(float);
  static size_t attr_smem = 48 * 1024;
  if (smem > attr_smem) {
    if (cudaFuncSetAttribute(small_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem) != cudaSuccess) {
      e.set("vitk_op_attention_small: cannot reserve %zu B of shared memory", smem);
      return fail();
    }
    attr_smem = smem;
  }
  small_attention_kernel<<<n_seq * heads, 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(out), probs_out, S,
      heads, d_head, 1.0f / sqrtf((float)d_head));
  if (cudaGetLastError() != cudaSuccess) { e.set("vitk_op_attention_small: launch failed"); return fail(); }
  return 0;
}

extern "C" int vitk_op_layernorm_rows_f32(const float* x, const float* gamma, const float* beta, float* y,
                                          int rows, long long row_stride, int d, float eps, void* stream) {
  Err& e = g_op_err;
  if (rows == 0) return 0;
  if (!x || !gamma || !beta || !y || rows < 0 || d < 1 || row_stride < d) {
    e.set("vitk_op_layernorm_rows_f32: bad argument");
    g_create_error = e.msg;
    return 1;
  }
  layernorm_rows_f32_kernel<<<(rows + 7) / 8, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, gamma, beta, y, rows, row_stride, d, eps);
  if (cudaGetLastError() != cudaSuccess) {
    e.set("vitk_op_layernorm_rows_f32: launch failed");
    g_create_error = e.msg;
    return 1;
  }
  return 0;
}

extern "C" int vitk_op_chunk_head(const float* x, const float* w1, const float* b1, const float* w2,
                                  const float* b2, float* logit, int n, int d, int hidden, void* stream) {
  Err& e = g_op_err;
  if (n == 0) return 0;
  if (!x || !w1 || !b1 || !w2 || !b2 || !logit || n < 0 || d < 1 || d > 8192 || hidden < 1) {
    e.set("vitk_op_chunk_head: bad argument");
    g_create_error = e.msg;
    return 1;
  }
  chunk_head_kernel<<<n, 256, (size_t)(d + 8) * sizeof(float), reinterpret_cast<cudaStream_t>(stream)>>>(
      x, w1, b1, w2, b2, logit, d, hidden);
  if (cudaGetLastError() != cudaSuccess) {
    e.set("vitk_op_chunk_head: launch failed");
    g_create_error = e.msg;
    return 1;
  }
  return 0;
}

// Developer hook (not in include/vitk.h): device buffer of 3*16*8 int64 clock samples written by
// CTA 0 of the attention kernel; pass NULL to switch the instrumentation off.
extern "C" int vitk_dev_attn_timeline(long long* buf_dev) {
  g_attn_dbg = buf_dev;
  return 0;
}

// Same for the stand-alone GEMM entry points (vitk_op_gemm / vitk_op_gemm_fused): 40 * 8 int64 samples
// written by CTA 0 (MMA thread, first epilogue warp of each column half, producer), per work unit.
extern "C" int vitk_dev_gemm_timeline(long long* buf_dev) {
  g_gemm_dbg = buf_dev;
  return 0;
}

extern "C" int vitk_op_patchify(const void* pixels, int fmt, void* out, int n, int C, int H, int W,
                                int patch, void* stream) {
  Err& e = g_op_err;
  int rc = launch_patchify(e, pixels, fmt, reinterpret_cast<__nv_bfloat16*>(out), n, C, H, W, patch,
                           reinterpret_cast<cudaStream_t>(stream));
  if (rc) g_create_error = e.msg;
  return rc;
}

// ------------------------------------------------------------------------------------------
// host-buffer end-to-end call: frames in, embeddings + labels (+ neighbours) out
// ------------------------------------------------------------------------------------------
extern "C" int vitk_embed_vote_host_u8(vitk_handle h, const uint8_t* frames_host, int n, int k,
                                       int metric, float* emb_host, int32_t* label_host,
                                       int32_t* idx_host, float* score_host,
                                       int32_t* counts_host) {
  if (!h) return 1;
  Err& e = h->err;
  if (!h->finalized) { e.set("vitk_embed_vote_host_u8: weights not finalized"); return 1; }
  if (h->bank_m == 0) { e.set("vitk_embed_vote_host_u8: no bank set"); return 1; }
  if (!frames_host || !emb_host || n < 0 || k < 1 || k > 64) {
    e.set("vitk_embed_vote_host_u8: bad argument");
    return 1;
  }
  if (h->bank_d != h->D) { e.set("bank width %d != hidden %d", h->bank_d, h->D); return 1; }
  CK(e, cudaSetDevice(h->device));
  cudaStream_t st = h->own_stream;
  if (order_enter(h, st)) return 1;
  const size_t fb = pixel_bytes_per_frame(h->cfg, PIX_U8_NHWC);
  const int nc = h->bank_classes > 0 ? h->bank_classes : 1;
  for (int off = 0; off < n; off += h->cfg.max_batch) {
    const int cn = n - off < h->cfg.max_batch ? n - off : h->cfg.max_batch;
    CK(e, cudaMemcpyAsync(h->pix_stage, frames_host + (size_t)off * fb, (size_t)cn * fb,
                          cudaMemcpyHostToDevice, st));
    // the reference stores / queries L2-normalised rows (ratt_write_clips_to_ragdb.py:152)
    if (forward_chunk(h, h->pix_stage, PIX_U8_NHWC, cn, h->emb, nullptr, 1, st)) return 1;
    if (vitk_knn_vote(h, h->emb, cn, k, metric, 0, h->res_idx, h->res_score, h->res_label,
                      h->res_counts, st))
      return 1;
    CK(e, cudaMemcpyAsync(emb_host + (size_t)off * h->D, h->emb, (size_t)cn * h->D * 4,
                          cudaMemcpyDeviceToHost, st));
    if (label_host)
      CK(e, cudaMemcpyAsync(label_host + off, h->res_label, (size_t)cn * 4, cudaMemcpyDeviceToHost, st));
    if (idx_host)
      CK(e, cudaMemcpyAsync(idx_host + (size_t)off * k, h->res_idx, (size_t)cn * k * 4,
                            cudaMemcpyDeviceToHost, st));
    if (score_host)
      CK(e, cudaMemcpyAsync(score_host + (size_t)off * k, h->res_score, (size_t)cn * k * 4,
                            cudaMemcpyDeviceToHost, st));
    if (counts_host)
      CK(e, cudaMemcpyAsync(counts_host + (size_t)off * nc, h->res_counts, (size_t)cn * nc * 4,
                            cudaMemcpyDeviceToHost, st));
  }
  if (order_leave(h, st)) return 1;
  CK(e, cudaStreamSynchronize(st));
  return 0;
}

// ------------------------------------------------------------------------------------------
// streaming host API: H2D of batch i+1 and D2H of batch i-1 overlap the forward of batch i
// ------------------------------------------------------------------------------------------
namespace {
int stream_init(vitk_ctx* h) {
  if (h->stream_ready) return 0;
  Err& e = h->err;
  const vitk_config& cfg = h->cfg;
  const size_t mb = (size_t)cfg.max_batch;
  for (auto& sl : h->slots) {
    if (dev_alloc(e, &sl.pix, mb * cfg.image_h * cfg.image_w * cfg.channels) ||
        dev_alloc(e, &sl.emb, mb * h->D) || dev_alloc(e, &sl.score, mb * 64) ||
        dev_alloc(e, &sl.idx, mb * 64) || dev_alloc(e, &sl.label, mb) ||
        dev_alloc(e, &sl.counts, mb * KNN_MAX_CLASSES))
      return 1;
    CK(e, cudaEventCreateWithFlags(&sl.ev_h2d, cudaEventDisableTiming));
    CK(e, cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
    CK(e, cudaEventCreateWithFlags(&sl.ev_out, cudaEventDisableTiming));
  }
  CK(e, cudaStreamCreateWithFlags(&h->in_stream, cudaStreamNonBlocking));
  CK(e, cudaStreamCreateWithFlags(&h->out_stream, cudaStreamNonBlocking));
  h->stream_ready = true;
  return 0;
}
}  // namespace

namespace {
// in_h == 0: frames are already at the model's resolution; otherwise raw (n,in_h,in_w,C) frames
// are uploaded into the slot's raw buffer and resized on the slot's stream before the forward.
int stream_submit_impl(vitk_ctx* h, int slot, const uint8_t* frames_host, int n, int in_h, int in_w,
                       int interp, int l2_normalise, int k, int metric, float* emb_host,
                       int32_t* label_host, int32_t* idx_host, float* score_host, int32_t* counts_host) {
  Err& e = h->err;
  if (!h->finalized) { e.set("vitk_stream_submit: weights not finalized"); return 1; }
  if (slot < 0 || slot > 1) { e.set("vitk_stream_submit: slot must be 0 or 1"); return 1; }
  if (!frames_host || !emb_host || n < 1 || n > h->cfg.max_batch) {
    e.set("vitk_stream_submit: need 1 <= n <= max_batch (%d) frames per submit", h->cfg.max_batch);
    return 1;
  }
  const bool vote = k > 0;
  if (vote && (h->bank_m == 0 || k > 64 || h->bank_d != h->D)) {
    e.set("vitk_stream_submit: k=%d needs a bank of width %d and k <= 64", k, h->D);
    return 1;
  }
  // frames already at the model's resolution need no resize whatever the caller asked for
  const bool raw = in_h > 0 && !(in_h == h->cfg.image_h && in_w == h->cfg.image_w);
  if (raw && (in_w < 1 || interp < 0 || interp > 2)) { e.set("vitk_stream_submit_raw: bad frame size / interpolation"); return 1; }
  CK(e, cudaSetDevice(h->device));
  if (stream_init(h)) return 1;
  vitk_ctx::StreamSlot& sl = h->slots[slot];
  if (sl.busy) { e.set("vitk_stream_submit: slot %d still in flight (call vitk_stream_wait)", slot); return 1; }
  const size_t fb = raw ? (size_t)in_h * in_w * h->cfg.channels : pixel_bytes_per_frame(h->cfg, PIX_U8_NHWC);
  const int nc = h->bank_classes > 0 ? h->bank_classes : 1;
  cudaStream_t st = h->own_stream;
  if (raw && sl.raw_bytes < fb * (size_t)h->cfg.max_batch) {
    // the slot is idle (its last use was waited for), but own_stream may still run the OTHER
    // slot's resize out of its own buffer: cudaFree synchronises the device anyway
    cudaFree(sl.raw);
    sl.raw = nullptr;
    sl.raw_bytes = 0;
    if (dev_alloc(e, &sl.raw, fb * (size_t)h->cfg.max_batch)) return 1;
    sl.raw_bytes = fb * (size_t)h->cfg.max_batch;
  }
  if (order_enter(h, st)) return 1;
  CK(e, cudaMemcpyAsync(raw ? sl.raw : sl.pix, frames_host, (size_t)n * fb, cudaMemcpyHostToDevice, h->in_stream));
  CK(e, cudaEventRecord(sl.ev_h2d, h->in_stream));
  CK(e, cudaStreamWaitEvent(st, sl.ev_h2d, 0));
  if (raw && launch_resize_interp(h, interp, sl.raw, n, in_h, in_w, sl.pix, st)) return 1;
  if (forward_chunk(h, sl.pix, PIX_U8_NHWC, n, sl.emb, nullptr, l2_normalise, st)) return 1;
  if (vote && knn_vote_impl(h, sl.emb, n, k, metric, 0, nullptr, nullptr, sl.idx, sl.score, sl.label, sl.counts, st))
    return 1;
  CK(e, cudaEventRecord(sl.ev_done, st));
  if (order_leave(h, st)) return 1;
  cudaStream_t so = h->out_stream;
  CK(e, cudaStreamWaitEvent(so, sl.ev_done, 0));
  CK(e, cudaMemcpyAsync(emb_host, sl.emb, (size_t)n * h->D * 4, cudaMemcpyDeviceToHost, so));
  if (vote) {
    if (label_host) CK(e, cudaMemcpyAsync(label_host, sl.label, (size_t)n * 4, cudaMemcpyDeviceToHost, so));
    if (idx_host) CK(e, cudaMemcpyAsync(idx_host, sl.idx, (size_t)n * k * 4, cudaMemcpyDeviceToHost, so));
    if (score_host) CK(e, cudaMemcpyAsync(score_host, sl.score, (size_t)n * k * 4, cudaMemcpyDeviceToHost, so));
    if (counts_host) CK(e, cudaMemcpyAsync(counts_host, sl.counts, (size_t)n * nc * 4, cudaMemcpyDeviceToHost, so));
  }
  CK(e, cudaEventRecord(sl.ev_out, so));
  sl.busy = true;
  return 0;
}
}  // namespace

extern "C" int vitk_stream_submit(vitk_handle h, int slot, const uint8_t* frames_host, int n, int k,
                                  int metric, float* emb_host, int32_t* label_host,
                                  int32_t* idx_host, float* score_host, int32_t* counts_host) {
  if (!h) return 1;
  return stream_submit_impl(h, slot, frames_host, n, 0, 0, 0, 1, k, metric, emb_host, label_host, idx_host,
                            score_host, counts_host);
}

extern "C" int vitk_stream_submit_raw(vitk_handle h, int slot, const uint8_t* frames_host, int n, int in_h,
                                      int in_w, int interp, int l2_normalise, int k, int metric,
                                      float* emb_host, int32_t* label_host, int32_t* idx_host,
                                      float* score_host, int32_t* counts_host) {
  if (!h) return 1;
  if (in_h < 1 || in_w < 1) { h->err.set("vitk_stream_submit_raw: bad frame size"); return 1; }
  return stream_submit_impl(h, slot, frames_host, n, in_h, in_w, interp, l2_normalise, k, metric, emb_host,
                            label_host, idx_host, score_host, counts_host);
}

extern "C" int vitk_stream_wait(vitk_handle h, int slot) {
  if (!h) return 1;
  Err& e = h->err;
  if (slot < 0 || slot > 1) { e.set("vitk_stream_wait: slot must be 0 or 1"); return 1; }
  vitk_ctx::StreamSlot& sl = h->slots[slot];
  if (!sl.busy) return 0;
  CK(e, cudaSetDevice(h->device));
  sl.busy = false;
  CK(e, cudaEventSynchronize(sl.ev_out));
  return 0;
}
