/* vitk.h — C ABI of libvitk.so: B200 (sm_100a) ViT embed + k-NN vote hot path.
 *
 * The reference (Gurinder-Vasanta/vit-research) is pure Python and has no FFI of its own; the
 * de-facto boundary of this path is a handful of Python callables.  Each entry
 * point below names the reference call it stands behind.  A maintainer binds this header with
 * ctypes (INTEGRATION.md shows the stub); `vit-research_b200/_lib.py` is that binding.
 *
 * Conventions: every function returns 0 on success, non-zero on error (message via
 * vitk_last_error); no C++ exceptions cross the ABI; device pointers are caller-owned
 * (torch tensor .data_ptr()); the library owns only its weight copy, bank copy and workspace;
 * calls are asynchronous on the given cudaStream_t (passed as void*) unless named *_host;
 * one handle per (device, stream) user, not thread-safe per handle.  There is no CPU fallback:
 * without a CUDA device every compute entry point fails.
 */
#ifndef VITK_H_
#define VITK_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vitk_ctx* vitk_handle;

/* Model geometry.  Defaults of HF ViTConfig() == google/vit-base-patch16-224
 * (nba_proj/ratt_write_clips_to_ragdb.py:58-60). Runtime, not compile-time. */
typedef struct vitk_config {
  int32_t image_h, image_w;   /* 224, 224 */
  int32_t patch;              /* 16 */
  int32_t channels;           /* 3 */
  int32_t hidden;             /* 768  (ViT-L: 1024) */
  int32_t heads;              /* 12   (ViT-L: 16); head dim must be 64 */
  int32_t mlp;                /* 3072 (ViT-L: 4096) */
  int32_t layers;             /* 12   (ViT-L: 24) */
  float ln_eps;               /* HF 1e-12; TF Model Garden 1e-6 */
  int32_t gelu_tanh;          /* 0: exact erf GELU (HF "gelu"); 1: tanh approximation (TF) */
  int32_t max_batch;          /* frames per vitk_embed call the workspace is sized for */
  float u8_mean, u8_std;      /* VITK_PIX_U8_NHWC pixels enter the model as (x - u8_mean) / u8_std.
                               * 0,0 = the HF processor (127.5, 127.5: rescale 1/255, mean .5, std .5
                               * fused); the reference's TF scripts feed RAW pixels: 0, 1
                               * (write_embeddings.py:185, generate_clips_hmm.py:397-401) */
} vitk_config;

enum { VITK_PIX_F32_NCHW = 0, VITK_PIX_BF16_NCHW = 1, VITK_PIX_U8_NHWC = 2 };
enum { VITK_METRIC_COSINE = 0, VITK_METRIC_L2 = 1 };
/* resize of raw frames to the model resolution: the HF processor's antialiased bilinear
 * (image_processing_vit.py:20-27), cv2.INTER_AREA (generate_clips_hmm.py:392, loader.py:7) and
 * cv2's default INTER_LINEAR (smarter_generate_clips.py:81) */
enum { VITK_RESIZE_BILINEAR_AA = 0, VITK_RESIZE_AREA = 1, VITK_RESIZE_LINEAR = 2 };
enum {
  VITK_EPI_BIAS_BF16 = 0, VITK_EPI_BIAS_GELU_BF16 = 1, VITK_EPI_BIAS_GELUT_BF16 = 2,
  VITK_EPI_BIAS_RESID_F32 = 3, VITK_EPI_TABLE_F32 = 4, VITK_EPI_F32 = 5,
  VITK_EPI_BIAS_RESID_BF16 = 7   /* 6 is internal (k-NN threshold filter) */
};

/* Library / build identification (no device needed). */
const char* vitk_version(void);
/* 1 if a CUDA device of compute capability 10.x is visible, else 0. */
int vitk_device_ok(int device);

/* ---- lifecycle ------------------------------------------------------------------------ */
/* Stands behind `ViTModel.from_pretrained(...).to(device).eval()`
 * (ratt_write_clips_to_ragdb.py:58-60). Allocates weights + workspace on `device`. */
int vitk_create(const vitk_config* cfg, int device, vitk_handle* out);
int vitk_destroy(vitk_handle h);
const char* vitk_last_error(vitk_handle h); /* h may be NULL: last create error */

/* Load one tensor by its HF ViTModel state-dict key (e.g.
 * "encoder.layer.0.attention.attention.query.weight"); host float32, row-major, shape checked.
 * Stands behind `model.load_weights` / `from_pretrained` (write_embeddings.py:243, chroma.py:159).
 * "pooler.*" keys are accepted and ignored (the reference never reads pooler_output). */
int vitk_set_weight(vitk_handle h, const char* name, const float* host, const int64_t* shape,
                    int ndim);
/* Checks every tensor was supplied, builds the fused tables (pos-emb + conv bias / CLS row). */
int vitk_finalize_weights(vitk_handle h);

/* ---- embedding ------------------------------------------------------------------------ */
/* Device-resident forward: pixels -> CLS embedding.  Stands behind
 * `vit_model(pixel_values=...).last_hidden_state[:,0,:]` (+ the numpy row normalisation
 * `cls / (norm + 1e-8)` when l2_normalise != 0) — ratt_write_clips_to_ragdb.py:146-153.
 *   pixels_dev : VITK_PIX_F32_NCHW / BF16_NCHW: (n,C,H,W) already rescaled+normalised;
 *                VITK_PIX_U8_NHWC: (n,H,W,C) raw uint8, the HF processor's rescale (1/255) and
 *                normalise ((x-.5)/.5) are applied on the fly, fused as transformers 5.5 does:
 *                (x - 127.5) / 127.5 in float32 (image_processing_vit.py:20-27).
 *   emb_f32_dev: (n, hidden) float32 out.  emb_bf16_dev: optional (n, hidden) bf16 out or NULL. */
int vitk_embed(vitk_handle h, const void* pixels_dev, int pixel_format, int n, float* emb_f32_dev,
               void* emb_bf16_dev, int l2_normalise, void* stream);

/* Host-buffer convenience used by `hf_vit_embed_batch(frames_np)`
 * (ratt_write_clips_to_ragdb.py:135-159): H2D copy, forward in max_batch chunks, D2H copy,
 * synchronous.  frames_host: (n,H,W,C) uint8 (pinned or pageable). emb_host: (n, hidden) f32. */
int vitk_embed_host_u8(vitk_handle h, const uint8_t* frames_host, int n, float* emb_host,
                       int l2_normalise);

/* ---- preprocessing front-end (the step right before the path) -------------------------- */
/* Stands behind `processor(images=frames_list, return_tensors="pt")` in hf_vit_embed_batch
 * (ratt_write_clips_to_ragdb.py:143-146): ViTImageProcessor's resize to the model resolution —
 * torchvision BILINEAR, antialias=True, on uint8 (image_processing_vit.py:20-27 [3P]) — as one
 * CUDA kernel, bit-identical (ATen's fixed-point separable algorithm: horizontal pass, uint8
 * intermediate kept in shared memory, vertical pass).  frames_dev (n,in_h,in_w,C) uint8 ->
 * out_dev (n,image_h,image_w,C) uint8, which vitk_embed takes as VITK_PIX_U8_NHWC (the fused
 * rescale+normalise happens there).  Asynchronous on `stream`. */
int vitk_preprocess_resize_u8(vitk_handle h, const uint8_t* frames_dev, int n, int in_h, int in_w,
                              uint8_t* out_dev, void* stream);
/* The TF path's front-end: cv2.resize(frame, (image_w, image_h), interpolation=cv2.INTER_AREA)
 * (generate_clips_hmm.py:392, chroma.py:204-205, write_embeddings.py:225-226), shrinking only:
 * OpenCV's area tables (covered fraction / cell width per source pixel, float accumulation,
 * round half to even), the integer box sum * (1.f / area) when both scale factors are integers,
 * and (a + b + c + d + 2) >> 2 at exactly 2 x 2 -- byte-identical to cv2 4.13.
 * frames_dev (n,in_h,in_w,C) uint8 -> out_dev (n,image_h,image_w,C) uint8.  Asynchronous. */
int vitk_preprocess_resize_area_u8(vitk_handle h, const uint8_t* frames_dev, int n, int in_h,
                                   int in_w, uint8_t* out_dev, void* stream);
/* cv2.resize(im, (image_w, image_h)) with the default interpolation, INTER_LINEAR: the resize
 * extract_embeddings applies (smarter_generate_clips.py:81).  OpenCV's 8-bit fixed-point bilinear
 * (11-bit coefficients, no antialiasing), shrinking or growing, byte-identical to cv2 4.13.
 * frames_dev (n,in_h,in_w,C) uint8 -> out_dev (n,image_h,image_w,C) uint8.  Asynchronous. */
int vitk_preprocess_resize_linear_u8(vitk_handle h, const uint8_t* frames_dev, int n, int in_h,
                                     int in_w, uint8_t* out_dev, void* stream);
/* hf_vit_embed_batch(frames_np) for host frames of ANY size: H2D of the raw frames, resize,
 * forward in max_batch chunks, D2H of the embeddings; synchronous. */
int vitk_embed_resize_host_u8(vitk_handle h, const uint8_t* frames_host, int n, int in_h, int in_w,
                              float* emb_host, int l2_normalise);

/* ---- banks + k-NN vote ---------------------------------------------------------------- */
/* Replicated embedding bank.  Stands behind np.load('*_embeddings.npz') + the Chroma bank
 * build (clustering.py:26-36, write_per_vid_embeddings_chroma.py:229-265).
 * rows_host (m,d) f32; labels_host (m) int32 in [0,n_classes) or NULL; normalise_rows != 0
 * applies b / (||b|| + 1e-8) (ratt_chunk_retriever.py:116). */
int vitk_bank_set(vitk_handle h, const float* rows_host, int m, int d, const int32_t* labels_host,
                  int n_classes, int normalise_rows);

/* Same, for rows (m,d) f32 / labels (m) int32 or NULL that already live on the handle's device
 * (a clip store produced by vitk_embed — build_frame_store, ratt_write_clips_to_ragdb.py:165-254 —
 * never has to visit the host).  Synchronous on `stream`. */
int vitk_bank_set_device(vitk_handle h, const float* rows_dev, int m, int d,
                         const int32_t* labels_dev, int n_classes, int normalise_rows, void* stream);

/* Mutable bank: write m_new rows at [at_row, at_row + m_new).  at_row == current row count
 * appends (`collection.upsert(embeddings, ids, metadatas)` of new ids: the confident-upsert loop,
 * generate_clips_hmm.py:456-470, chroma.py:257-261, generate_clips.py:348-370); a smaller at_row
 * overwrites (upsert of an existing id).  rows (m_new,d) f32 and labels (m_new) int32 (NULL
 * exactly when the bank was set without labels) are host pointers, or device pointers when
 * on_device != 0.  Storage grows geometrically; only the new rows are normalised / cast.
 * Synchronous on `stream`. */
int vitk_bank_write(vitk_handle h, int at_row, const float* rows, int m_new, const int32_t* labels,
                    int on_device, int normalise_rows, void* stream);
/* Rows currently in the bank (`collection.count()`), -1 for a NULL handle. */
int vitk_bank_rows(vitk_handle h);

/* Several banks on one handle -- the reference keeps several Chroma collections side by side (the
 * left/right/none label bank, chroma.py:145; the clip / chunk stores, ratt_write_clips_to_ragdb.py:
 * 302-310).  A handle starts with bank 0 current.  vitk_bank_new parks the current bank and makes
 * a fresh empty one current (returns its id >= 1, -1 on error); vitk_bank_select switches the
 * current bank (every vitk_bank_* / vitk_knn_* / vote call works on the current one);
 * vitk_bank_free releases a bank's device memory (the current one becomes empty). */
int vitk_bank_new(vitk_handle h);
int vitk_bank_select(vitk_handle h, int bank_id);
int vitk_bank_current(vitk_handle h);
int vitk_bank_free(vitk_handle h, int bank_id);

/* Top-k + vote.  Stands behind `collection.query(query_embeddings, n_results=k)`
 * (chroma.py:211-214; 'l2' space = squared L2 ascending) / RattChunkRetriever's
 * `sims = Q @ C.T; argpartition; argsort` (ratt_chunk_retriever.py:123,144-148; cosine
 * descending) followed by determine_class_pass2 (chroma.py:102-134).
 *   q_dev (n,d) f32; normalise_queries: q / (||q|| + 1e-8) first.
 *   idx_out (n,k) int32 bank rows in result order; score_out (n,k) f32 (cosine similarity, or
 *   squared L2 distance); label_out (n) int32 (first-max of counts); counts_out (n,n_classes).
 *   Any out pointer except idx_out/score_out may be NULL.  1 <= k <= 1024 (the vote uses 5 / 25 /
 *   50; the clip-store queries ask for n_results = search_k = 100-750, frame_retriever.py:41-43,
 *   config.py:10-11).  k <= 64 re-ranks 2k+32 candidates, larger k re-ranks k + k/4 + 32.
 * Banks of up to 2048 rows: fully asynchronous on `stream`.  Larger stores (the 1 M-row clip
 * store of rag_vit / RattChunkRetriever) never materialise the (n,m) score matrix: the
 * similarity GEMM's epilogue filters scores against a running per-query threshold into small
 * candidate pools; that path synchronises `stream` once per call (pool-overflow check). */
int vitk_knn_vote(vitk_handle h, const float* q_dev, int n, int k, int metric,
                  int normalise_queries, int32_t* idx_out, float* score_out, int32_t* label_out,
                  int32_t* counts_out, void* stream);
/* Same under a metadata filter: `collection.query(..., where={...})` (frame_retriever.py:41-53,
 * ratt_chunk_retriever.py:98-107).  The host evaluates the where-clause on the metadata and
 * passes row_mask_dev (bank rows) uint8, 1 = row in play; NULL = no filter.  Filtered rows can
 * never be returned; a query with fewer than k rows in play pads idx with -1 like vitk_knn_vote. */
int vitk_knn_vote_masked(vitk_handle h, const float* q_dev, int n, int k, int metric,
                         int normalise_queries, const uint8_t* row_mask_dev, int32_t* idx_out,
                         float* score_out, int32_t* label_out, int32_t* counts_out, void* stream);

/* Same with a PER-QUERY row predicate evaluated inside the kernels -- RattChunkRetriever's
 * `valid_mask = (cand_vids != vid) & (cand_tcenters >= q_start) & (cand_tcenters <= q_end)` within
 * one side (ratt_chunk_retriever.py:85-88,98-107,130-134) and FrameRetriever's
 * `where={"$and": [{"vid_num": {"$ne": vid}}, {"side": side}, {"t_norm": {"$gte": ..}}, {"t_norm":
 * {"$lte": ..}}]}` (frame_retriever.py:41-53): no (queries x rows) mask is ever built.
 * Columns are device arrays indexed by bank row, q_* device arrays indexed by query; a NULL column
 * switches its clause off.  Row r is in play for query i iff
 *     ne_col[r] != q_ne[i]  and  eq_col[r] == q_eq[i]  and  q_lo[i] <= x <= q_hi[i]
 * with x = rng_col[r], rounded to float32 first when rng_round_f32 != 0 (the reference compares
 * float32 arrays), and -- when q_lo2/q_hi2 are given -- also q_lo2[i] <= rng_col[r] <= q_hi2[i] on
 * the un-rounded double (the side-wide `collection.get` window).  row_mask_dev (may be NULL)
 * applies on top, as in vitk_knn_vote_masked.  Queries with fewer than k rows in play pad idx
 * with -1. */
typedef struct vitk_row_filter {
  const int32_t* ne_col; const int32_t* q_ne;
  const int32_t* eq_col; const int32_t* q_eq;
  const double* rng_col; const double* q_lo; const double* q_hi;
  const double* q_lo2;   const double* q_hi2;
  int32_t rng_round_f32;
} vitk_row_filter;
int vitk_knn_vote_filtered(vitk_handle h, const float* q_dev, int n, int k, int metric,
                           int normalise_queries, const uint8_t* row_mask_dev,
                           const vitk_row_filter* filter, int32_t* idx_out, float* score_out,
                           int32_t* label_out, int32_t* counts_out, void* stream);

/* Host-buffer end-to-end call: what a frame-labelling loop does per batch
 * (generate_clips_hmm.py:387-412 — embed, query, vote), on static banks.  H2D of the uint8
 * frames, forward (L2-normalised rows), top-k + vote, D2H of the results; synchronous.
 * emb_host (n,hidden) f32; label_host (n); idx_host/score_host (n,k); counts_host
 * (n,n_classes); every output except emb_host may be NULL. */
int vitk_embed_vote_host_u8(vitk_handle h, const uint8_t* frames_host, int n, int k, int metric,
                            float* emb_host, int32_t* label_host, int32_t* idx_host,
                            float* score_host, int32_t* counts_host);

/* Streaming form of the call above for frame STREAMS — the reference's batched drivers
 * (build_frame_store, ratt_write_clips_to_ragdb.py:197-209; extract_embeddings,
 * smarter_generate_clips.py:69-94) walk a long frame list batch by batch.  Two batches may be in
 * flight (slot 0 / 1): the H2D copy of batch i+1 and the D2H copy of batch i-1 overlap the
 * forward of batch i on separate streams.  submit() returns immediately; the host output
 * buffers (pinned for true overlap) are valid after wait(slot).  n <= max_batch per submit;
 * k == 0 skips the vote (embeddings only; the vote outputs are then ignored). */
int vitk_stream_submit(vitk_handle h, int slot, const uint8_t* frames_host, int n, int k, int metric,
                       float* emb_host, int32_t* label_host, int32_t* idx_host, float* score_host,
                       int32_t* counts_host);
int vitk_stream_wait(vitk_handle h, int slot);
/* Same for RAW frames of any size -- what build_frame_store feeds (PIL frames at the video's
 * resolution, ratt_write_clips_to_ragdb.py:197-209) and what extract_embeddings loads: the H2D
 * copy carries (n,in_h,in_w,C) uint8 frames, the resize (interp = VITK_RESIZE_*) runs on the
 * slot's stream in front of the forward, so raw frames take the same overlapped path.
 * l2_normalise: hf_vit_embed_batch normalises (1); Keras pre_logits are raw (0). */
int vitk_stream_submit_raw(vitk_handle h, int slot, const uint8_t* frames_host, int n, int in_h,
                           int in_w, int interp, int l2_normalise, int k, int metric,
                           float* emb_host, int32_t* label_host, int32_t* idx_host,
                           float* score_host, int32_t* counts_host);

/* ---- vote -> HMM smoothing (the step right after the path) ------------------------------ */
/* Stands behind class hmm (nba_proj/hmm.py:5-135: add_first, add_col_to_lattice,
 * decode_sequence) as driven by generate_clips_hmm.py:238-245,484-490: a 3-state
 * (left,right,none) log-space Viterbi with the reference's fixed transitions, its per-step
 * Python round(x,3), its tie rules and its backtrace as written (it starts one column early
 * and reads backpointers one column off; entries 0 and 1 of each sequence come back -1).
 * Sequences (videos) are independent: frame range [seq_offsets[i], seq_offsets[i+1]) each,
 * seq_offsets_host has n_seq+1 entries starting at 0.
 *   counts_dev (T,3) int32 = the vote counts vitk_knn_vote produced with n_classes == 3 and this
 *   k; with static one-hot banks the reference's emission means are count/k.
 *   states_out_dev (T) int32 in {0 left,1 right,2 none,-1}; dp_out_dev optional (T,3) float64
 *   rounded lattice.  Asynchronous on `stream`. */
int vitk_hmm_decode_counts(vitk_handle h, const int32_t* counts_dev, int k,
                           const int32_t* seq_offsets_host, int n_seq, int32_t* states_out_dev,
                           double* dp_out_dev, void* stream);
/* Streamed form for ONE long sequence whose vote counts arrive segment by segment in time order
 * (a game sharded over several GPUs by chunk: the lattice advances while the stream is still being
 * embedded, instead of after it).  begin(capacity) opens a stream; append_counts consumes the next
 * n frames (counts_dev (n,3) int32, same k throughout) and carries the lattice column over;
 * finish writes the states of all T frames consumed (same conventions as above) and closes the
 * stream.  Any segmentation yields bit for bit the states of vitk_hmm_decode_counts on the whole
 * sequence.  These calls use their own buffers and are NOT ordered against the forward: they may
 * run on any stream beside vitk_stream_submit; consecutive appends must be issued on one stream
 * (or on streams ordered by the caller). */
int vitk_hmm_stream_begin(vitk_handle h, long long capacity_frames);
int vitk_hmm_stream_append_counts(vitk_handle h, const int32_t* counts_dev, int n, int k, void* stream);
int vitk_hmm_stream_finish(vitk_handle h, int32_t* states_out_dev, long long* frames_out, void* stream);
/* Viterbi decode that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arithmetic and output as vitk_hmm_decode_counts
 * for ONE sequence of n_frames frames (counts_dev (n_frames,3) int32, states_out_dev (n_frames)
 * int32, both caller-owned device buffers), but the scratch that entry point keeps in cudaMalloc'd
 * buffers of the handle (log emissions, backpointers, offsets) lives in a fixed array inside the
 * library image: 3.4 MB of device memory that every GPU which loads the library holds until exit.
 * Limits: n_frames <= VITK_HMM_STATIC_MAX_FRAMES and 1 <= k <= VITK_HMM_STATIC_MAX_K; anything
 * larger returns non-zero with the reason in vitk_last_error(NULL) -- nothing is truncated.
 * Needs no handle; runs on the CURRENT device, asynchronously on `stream`; no cudaMalloc, cudaFree
 * or other allocating call is made (the first call on a device resolves the array's address).
 * All calls on one GPU share the array, which keeps the last call's contents (no part of it has
 * to start at zero: every entry read is rewritten by the same call first); two calls that overlap
 * on the GPU (different streams, no ordering between them) RACE -- order them. */
#define VITK_HMM_STATIC_MAX_FRAMES 131072
#define VITK_HMM_STATIC_MAX_K 128
int vitk_hmm_decode_counts_static(const int32_t* counts_dev, int n_frames, int k,
                                  int32_t* states_out_dev, void* stream);
/* Same for arbitrary per-frame probabilities (T,3) float64 on the host (the dicts handed to
 * add_first / add_col_to_lattice); outputs on the host; synchronous. */
int vitk_hmm_decode_probs_host(vitk_handle h, const double* probs_host,
                               const int32_t* seq_offsets_host, int n_seq, int32_t* states_out_host,
                               double* dp_out_host);

/* ---- operator-level entry points (used by the parity tests, one kernel each) ----------- */
/* C[M,N] = A[M,K] (bf16) · W[N,K]^T (bf16) with epilogue `epi` (VITK_EPI_*).
 * out is bf16 for the *_BF16 epilogues, f32 otherwise (row pitch = N elements). */
int vitk_op_gemm(const void* a_bf16_dev, const void* w_bf16_dev, void* out_dev, int M, int N, int K,
                 int epi, const float* bias_dev, const float* resid_dev, const float* table_dev,
                 int table_rows, void* stream);
/* The two fused forms the forward actually runs (HF modeling_vit.py:325-346):
 *  - LayerNorm folded into the consuming GEMM (epi = VITK_EPI_BIAS[_GELU|_GELUT]_BF16):
 *      out = act( rstd[m] * (acc[m,n] - mean[m]*c1[n]) + bias[n] )
 *    where A is the RAW (un-normalised) row in bf16, W is pre-scaled by gamma, c1[n] =
 *    sum_k W[n,k], bias[n] = b[n] + sum_k beta[k]*W0[n,k], and mean/rstd come from
 *    stats_dev: per row `stats_parts` partial (sum, sumsq) float pairs over the K = d columns.
 *  - residual GEMM that also emits the next LayerNorm's inputs (epi = VITK_EPI_BIAS_RESID_F32):
 *      out f32 = acc + bias + resid;  xb_out (bf16 copy of out, may be NULL);  stats_out
 *      (may be NULL): per row 2*ceil(N/256) partial (sum, sumsq) pairs of out, pair index
 *      2*(n/256) + (n%256)/128.
 *    epi = VITK_EPI_BIAS_RESID_BF16 is the form the forward runs by default: resid_dev and
 *    out_dev are bf16 (pass the bf16 pointer through the float* parameter; they may alias),
 *      out bf16 = bf16(acc + bias + resid), stats_out from the fp32 value before rounding;
 *    xb_out is ignored (the output IS the next GEMM's bf16 operand).
 * NULL c1/stats/xb_out/stats_out reduce to vitk_op_gemm. */
int vitk_op_gemm_fused(const void* a_bf16_dev, const void* w_bf16_dev, void* out_dev, int M, int N,
                       int K, int epi, const float* bias_dev, const float* c1_dev,
                       const float* stats_dev, int stats_parts, float ln_eps,
                       const float* resid_dev, void* xb_out_bf16_dev, float* stats_out_dev,
                       void* stream);
/* Row statistics + bf16 cast of an f32 matrix [rows, d] (d in {768,1024}): xb = bf16(x),
 * stats[row] = `parts` (sum, sumsq) pairs, pair 0 = the full-row sums, the rest zero.  Seeds
 * the folded LayerNorm of layer 0 from the patch-embedding output. */
int vitk_op_rowstats(const float* x_dev, void* xb_bf16_dev, float* stats_dev, int rows, int d,
                     int parts, void* stream);
/* y(bf16) = LayerNorm(x f32) over the last dim d (d in {768, 1024}). */
int vitk_op_layernorm(const float* x_dev, const float* gamma_dev, const float* beta_dev,
                      void* y_bf16_dev, int rows, int d, float eps, void* stream);
/* out[frames,S,heads*64] (bf16) = softmax(q k^T / 8) v from qkv[frames,S,3*heads*64] (bf16). */
int vitk_op_attention(const void* qkv_bf16_dev, void* out_bf16_dev, int frames, int S, int heads,
                      void* stream);
/* ---- ChunkEncoder pieces (models/chunk_encoder.py:9-97) ----------------
 * The temporal transformer over [CLS, frame_1..frame_T] embeddings runs its linear layers on
 * vitk_op_gemm; these three cover the rest.
 * Attention over short sequences with any head width (Keras MultiHeadAttention inside
 * nn_blocks.TransformerEncoderBlock, chunk_encoder.py:17-31: 8 heads of 96):
 *   out[n_seq,S,heads*d_head] (bf16) = softmax((q / sqrt(d_head)) k^T) v from
 *   qkv[n_seq,S,3*heads*d_head] (bf16);  S <= 64, d_head a multiple of 32 up to 128;
 *   probs_out (may be NULL) f32 [n_seq,heads,S,S] = the attention scores the block returns. */
int vitk_op_attention_small(const void* qkv_bf16_dev, void* out_bf16_dev, float* probs_out_dev,
                            int n_seq, int S, int heads, int d_head, void* stream);
/* y f32 [rows,d] = LayerNorm of `rows` f32 rows that sit row_stride floats apart (the CLS row of
 * every sequence: chunk_encoder.py:83-86). */
int vitk_op_layernorm_rows_f32(const float* x_dev, const float* gamma_dev, const float* beta_dev,
                               float* y_dev, int rows, long long row_stride, int d, float eps,
                               void* stream);
/* class_head (chunk_encoder.py:34-38): logit[n] = w2 . relu(W1 x + b1) + b2; W1 [hidden,d] f32. */
int vitk_op_chunk_head(const float* x_dev, const float* w1_dev, const float* b1_dev,
                       const float* w2_dev, const float* b2_dev, float* logit_dev, int n, int d,
                       int hidden, void* stream);
/* Merge of per-shard top-k lists (the row-sharded clip store's exchange step): idx_in / score_in
 * [n, c] with GLOBAL row ids (-1 = empty slot), c <= 4096 -> the k best per query, better score
 * first (descending != 0: larger is better), ties to the lower row id -- the kernels' own total
 * order; missing entries come back as (-1, 0). */
int vitk_op_merge_topk(const int32_t* idx_in_dev, const float* score_in_dev, int n, int c, int k,
                       int descending, int32_t* idx_out_dev, float* score_out_dev, void* stream);
/* im2col + pixel-format conversion: out bf16 [n*(P+1), C*patch*patch] with a zero CLS row. */
int vitk_op_patchify(const void* pixels_dev, int pixel_format, void* out_bf16_dev, int n, int C,
                     int H, int W, int patch, void* stream);

/* In-step timing of the dominant kernel: while enabled, every fc1 GEMM launch of the forward
 * (HF modeling_vit.py:287-299; 26 % of the step) is bracketed by a CUDA-event pair on the
 * launching stream.  vitk_profile_fc1_read waits for them, returns the summed duration and the
 * number of launches, and clears the list (bench.py's roofline.in_step). */
int vitk_profile_fc1(vitk_handle h, int enable);
int vitk_profile_fc1_read(vitk_handle h, double* total_ms, int* launches);
/* Kernel launches issued by this handle since creation (bench.py's gpu_launches). */
int64_t vitk_launch_count(vitk_handle h);
/* Large-store k-NN stages that had to be redone because a candidate pool overflowed (0 on
 * exchangeable stores; the parity tests provoke it with a store sorted by similarity). */
int64_t vitk_knn_redo_count(vitk_handle h);

#ifdef __cplusplus
}
#endif
#endif /* VITK_H_ */
