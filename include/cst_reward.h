/*
 * cst_reward.h -- C ABI of the B200 (sm_100a) CIDEr-D reward path.
 *
 * Drop-in boundary for the one hot path of AislingGui/cst_captioning: the
 * CIDEr-D reward of the RL loop.  Every entry point names the reference
 * interface it replaces (file:line under the reference checkout).
 *
 * Conventions
 *   - plain C symbols, no exceptions cross the boundary;
 *   - return value: 0 = OK, negative = error (CST_E_*); the message of the last
 *     error on the calling thread is cst_last_error();
 *   - the caller owns every buffer; the library owns only df-table handles;
 *   - pointers marked [dev] are device pointers on the table's device, [host]
 *     are host pointers;
 *   - work is enqueued on the caller's stream (a cudaStream_t passed as void*,
 *     NULL = legacy default stream) and the call returns WITHOUT synchronising,
 *     except where a function says "synchronous";
 *   - the product has no CPU fallback: without a CUDA device every compute entry
 *     point fails with CST_E_CUDA.
 *
 * Token rows
 *   A caption is a row of token ids (int16, int32 or int64, row-major, `stride` elements
 *   between rows).  Canonicalisation follows utils.py:135-152 (array_to_str):
 *   the first 0 (<end>) ends the caption and is kept only if use_eos != 0; id 1
 *   (<start>) is dropped wherever it occurs.  Rows may be at most 32 tokens wide
 *   (reference: 30 for GT rows, <= 28 for sampled rows).  Ids must lie in
 *   [0, 65534]; the id -1 is padding (it ends the caption and is never emitted,
 *   whatever use_eos says).  Rows holding any other id are scored as if the bad
 *   token ended the caption and are counted in stats[4].
 *
 * Packed n-gram keys
 *   The n-gram (t0 .. t(k-1)), k = 1..4, is the 64-bit key
 *       key = sum_i (t_i + 1) << (16 * i)
 *   (a zero 16-bit lane means "no token", so <end> = id 0 stays representable).
 *   This replaces the tuple-of-decimal-strings keys of compute_ciderdf.py:32,84.
 */
#ifndef CST_REWARD_H_
#define CST_REWARD_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CST_OK 0
#define CST_E_ARG (-1)     /* bad argument (NULL, size, range, mode) */
#define CST_E_CUDA (-2)    /* CUDA runtime error, or no device */
#define CST_E_NOMEM (-3)   /* host or device allocation failed */
#define CST_E_RANGE (-4)   /* value outside what the kernels support */

#define CST_DTYPE_I32 0
#define CST_DTYPE_I64 1
#define CST_DTYPE_I16 2 /* token rows only (ids <= 32767): halves an upload from the host */

/* baseline modes of cst_reward / cst_scb_apply */
#define CST_BASE_NONE 0        /* reward = score               (utils.py:317-319) */
#define CST_BASE_GREEDY 1      /* reward = sampled - greedy    (utils.py:219-222) */
#define CST_BASE_SCB_GT 2      /* scb_baseline == 1            (utils.py:296-300, 309-310) */
#define CST_BASE_SCB_SAMPLE 3  /* scb_baseline == 2            (utils.py:301-304, 311-312) */

#define CST_MAX_ROW_WIDTH 32
#define CST_MAX_TOKEN_ID 65534
#define CST_NGRAM_SLOTS_PER_ROW 128 /* 4 orders x 32 positions */
#define CST_N_STATS 7

typedef struct cst_df_table cst_df_table;

/* Message of the last failing call on this thread ("" if none). */
const char* cst_last_error(void);

/* ABI version of the loaded library. */
int cst_abi_version(void);

/* Number of CUDA kernels this library has launched in this process so far
 * (instrumentation for bench.py's gpu_launches; no counterpart in the reference). */
int64_t cst_kernel_launches(void);

/* ---- document-frequency table ------------------------------------------------
 * Replaces the unpickled {'document_frequency': {ngram: df}, 'ref_len': n} dict
 * that CiderD(df=path) holds (construction site train.py:121; pickle schema
 * compute_ciderdf.py:123-129) with a device-resident open-addressing table
 * keyed by packed n-grams.
 *
 * keys[n], df[n] are [host]; df values are the integer-valued floats of the
 * pickle.  idf_by_df [host] has n_documents + 1 entries,
 *     idf_by_df[d] = log(n_documents) - log(max(1, d)),
 * evaluated by the caller in float64 exactly as the scorer's counts2vec does (the
 * Python wrapper uses numpy, so the values are bit-identical to the reference's);
 * entry 0 is the idf of an n-gram the table does not hold.  All entries must be
 * >= 0 (df <= n_documents, which compute_ciderdf guarantees).
 * Synchronous (one-off, like the pickle load it replaces). */
int cst_df_create(const uint64_t* keys, const double* df, int64_t n,
                  const double* idf_by_df, int64_t n_documents, int device,
                  cst_df_table** out);

/* Build the table on the device from token rows: the device-side form of
 * compute_ciderdf.py:56-93 (compute_doc_freq over per-video ref sets) and of the
 * corpus-mode df the coco-caption scorer recomputes per call
 * (compute_scores.py:92).  rows [dev] are canonicalised with `use_eos` like any
 * other row; video v owns rows [video_offsets[v], video_offsets[v+1]) ([host],
 * n_videos + 1 entries).  idf_by_df [host] has n_videos + 1 entries:
 * idf_by_df[d] = log(n_videos) - log(max(1, d)) evaluated by the caller.
 * Synchronous. */
int cst_df_build_from_rows(const void* rows, int32_t dtype, int64_t n_rows,
                           int32_t width, int64_t stride,
                           const int64_t* video_offsets, int64_t n_videos,
                           int32_t use_eos, const double* idf_by_df, int device,
                           cst_df_table** out);

int cst_df_destroy(cst_df_table* table);

/* Number of stored n-grams / slots / device of a table. */
int64_t cst_df_size(const cst_df_table* table);
int64_t cst_df_capacity(const cst_df_table* table);
int cst_df_device(const cst_df_table* table);

/* Copy the table's content back: keys_out[n], df_out[n] are [host], n =
 * cst_df_size().  Order is unspecified.  Synchronous.  (Lets a device-built
 * table be written as the reference's pickle, compute_ciderdf.py:123-129.) */
int cst_df_export(const cst_df_table* table, uint64_t* keys_out, double* df_out);

/* Parity probe: df_out[i] = document frequency of keys[i] (0.0 if absent),
 * idf_out[i] its idf (idf_unseen if absent; may be NULL).  keys, df_out, idf_out
 * are [dev].  The dict lookup it replaces is document_frequency[ngram] in the
 * scorer's counts2vec. */
int cst_df_lookup(const cst_df_table* table, const uint64_t* keys, int64_t n,
                  double* df_out, double* idf_out, void* stream);

/* ---- n-gram extraction probe ---------------------------------------------------
 * Replaces array_to_str (utils.py:135-152) followed by precook
 * (compute_ciderdf.py:19-34) for each row.  For row i the distinct n-grams are
 * written to keys_out[i*128 ...] with their term frequencies in counts_out, in the
 * insertion order of precook's dict (order 1..4 outer, first position inner);
 * n_out[i] = how many, len_out[i] = canonical token count.  All outputs [dev].
 * `device` selects the GPU (no table is involved). */
int cst_ngram_counts(const void* rows, int32_t dtype, int64_t n_rows,
                     int32_t width, int64_t stride, int32_t use_eos,
                     uint64_t* keys_out, int32_t* counts_out, int32_t* n_out,
                     int32_t* len_out, int device, void* stream);

/* ---- scoring + reward ------------------------------------------------------------
 * Replaces CiderD.compute_score(gts, res) as called from utils.py:208,274 together
 * with the string building around it (utils.py:179-206, 249-272) and, for
 * cst_reward, the baseline block (utils.py:219-224, 292-322).
 *
 *   cand        [dev] n_cand sampled rows (cand_width tokens, cand_stride apart)
 *   greedy      [dev] n_cand greedy rows of the same geometry, or NULL
 *               (required for CST_BASE_GREEDY; scored like utils.py:180-183)
 *   refs        [dev] all reference rows, video-major
 *   ref_offsets [dev] n_videos + 1 row offsets into refs (ragged ref counts), or
 *               NULL when every video has exactly refs_per_video rows; with
 *               ref_offsets, refs_per_video is the LARGEST ref count of any video
 *               (it sizes the shared-memory staging; videos with more than 32
 *               refs are staged in passes of 32)
 *   cands_per_video   candidate i is scored against video i / cands_per_video:
 *               seq_per_img for expand_feat == 1 (utils.py:266-270), 1 for
 *               expand_feat == 0 (utils.py:271-272).  n_cand must equal
 *               n_videos * cands_per_video.
 *   scores_out  [dev] float64, n_cand entries (2 * n_cand when greedy != NULL:
 *               sampled first, then greedy, the order of utils.py:180-183);
 *               may be NULL in cst_reward.
 */
int cst_cider_scores(const cst_df_table* table,
                     const void* cand, const void* greedy, int32_t cand_dtype,
                     int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                     const void* refs, int32_t ref_dtype, int32_t ref_width,
                     int64_t ref_stride, const int64_t* ref_offsets,
                     int32_t refs_per_video, int64_t n_videos,
                     int32_t cands_per_video, int32_t use_eos,
                     double* scores_out, void* stream);

/* Bytes of scratch cst_reward / cst_scb_apply need for n_groups baseline groups
 * (for cst_reward: max(n_videos, n_cand / seq_per_img)).  The scratch buffer [dev] must be zero-filled once before its first
 * use; the kernels leave it ready for the next call.  Calls sharing one scratch
 * buffer must not overlap on the device. */
int64_t cst_reward_scratch_bytes(int64_t n_groups);

/* Scores + baseline in one launch.  seq_per_img is the row length of the
 * reference's scores.reshape(-1, seq_per_img) (utils.py:284); the fused path needs
 * cands_per_video == seq_per_img (expand_feat == 1, what train.py uses); for
 * cands_per_video == 1 the baseline runs as a second small kernel over the scores.
 *
 *   mode        CST_BASE_*
 *   scb_k       scb_captions (1..seq_per_img) for the two SCB modes
 *   bcmrscores  [dev] float64 (n_cand / seq_per_img, seq_per_img) precomputed GT
 *               consensus scores (dataloader.py:142-144), CST_BASE_SCB_GT only
 *   reward_out  [dev] float32 n_cand: the column the reference repeats T times
 *               (utils.py:224, 321-322) and casts to float (train.py:209)
 *   stats_out   [dev] float64[CST_N_STATS] = { sum_m, sum_b, n_m, n_b, n_bad_rows,
 *               m_score, b_score }:
 *               m_score = sum_m / n_m, b_score = sum_b / n_b as returned at
 *               utils.py:226, 324 (sums, so ranks can all-reduce them);
 *               n_bad_rows counts rows with ids outside [0, 65534].
 */
int cst_reward(const cst_df_table* table,
               const void* cand, const void* greedy, int32_t cand_dtype,
               int64_t n_cand, int32_t cand_width, int64_t cand_stride,
               const void* refs, int32_t ref_dtype, int32_t ref_width,
               int64_t ref_stride, const int64_t* ref_offsets,
               int32_t refs_per_video, int64_t n_videos,
               int32_t cands_per_video, int32_t use_eos,
               int32_t mode, int32_t seq_per_img, int32_t scb_k,
               const double* bcmrscores,
               float* reward_out, double* scores_out, double* stats_out,
               void* scratch, int64_t scratch_bytes, void* stream);

/* The same launch as cst_reward for hosts that cannot allocate device memory at the
 * time of the call: hosts that gave all of HBM to their own pool at start-up (so a
 * plugin's cudaMalloc fails), or whose step loop bans allocation because cudaMalloc
 * and cudaFree stall the whole device.  The statistics scratch is a fixed array
 * inside the library instead of a caller buffer; inputs and outputs are still the
 * caller's.  It has no counterpart in the reference.
 *   - Limit: max(n_videos, n_cand / seq_per_img) <= CST_NOALLOC_MAX_GROUPS; larger
 *     batches are refused with CST_E_RANGE (split them, or use cst_reward).
 *   - Cost: about 256 KiB of device memory on every GPU that loads the library,
 *     held until the process exits.
 *   - The first call on a device looks the array's address up (one-time set-up);
 *     make that first call before any stream capture begins.
 *   - All calls on a device share the one array: two calls whose kernels overlap on
 *     the GPU (different streams) race on it.  Serialise them.
 *   - It calls no cudaMalloc / cudaFree or other allocating routine; its counters are
 *     cleared on the caller's stream at the start of every call.
 *   - One thing still allocates, and not in this library: under CUDA's default lazy module
 *     loading the DRIVER loads a kernel's code at its first launch, which takes device memory.
 *     The entry point launches the scoring kernel, the statistics fold (batches of 2,368 videos
 *     or more on a B200) and, when cands_per_video != seq_per_img, the baseline kernel: make one
 *     small call of each shape you will use BEFORE the pool takes the memory, or run with
 *     CUDA_MODULE_LOADING=EAGER. */
#define CST_NOALLOC_MAX_GROUPS 16384
int cst_reward_noalloc(const cst_df_table* table,
                       const void* cand, const void* greedy, int32_t cand_dtype,
                       int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                       const void* refs, int32_t ref_dtype, int32_t ref_width,
                       int64_t ref_stride, const int64_t* ref_offsets,
                       int32_t refs_per_video, int64_t n_videos,
                       int32_t cands_per_video, int32_t use_eos,
                       int32_t mode, int32_t seq_per_img, int32_t scb_k,
                       const double* bcmrscores,
                       float* reward_out, double* scores_out, double* stats_out,
                       void* stream);

/* ---- staged references (the producer side) -----------------------------
 * The reference rebuilds every video's reference vectors on every call (cook_refs inside the
 * scorer's cook_append, reached from utils.py:208,274 with the gts of dataloader.py:138-139),
 * although the ground-truth captions never change during training.  cst_refs_cook runs that
 * half of the scoring kernel ONCE per video and keeps the result in device memory -- the
 * per-video n-gram table with its presence masks, the refs' inverse norms and lengths: one
 * image of cst_refs_cooked_bytes() = 21,248 bytes per video whatever the row width and count
 * (MSR-VTT's 10 k videos: 0.21 GB of the 180 GB).  cst_reward_cooked then scores a batch against the
 * images of the videos named by image_index (device int64, one per video of the batch; NULL =
 * images 0..n_videos-1 in order): a step passes video indices instead of reference rows.
 * The indices are not range-checked (that would need a host read): every one must name an
 * image inside the buffer.
 * Images depend on the df table, use_eos, ref_width and refs_per_video they were cooked with;
 * an image of another layout, table geometry or use_eos makes the rewards of that video NaN
 * and is counted in stats_out[4].  An image is ONE table: at most 32 refs per video and at most
 * 1,100 n-gram instances over them (sum over the refs of 4 L - 6 for L tokens: 20 refs of up to 15
 * tokens on average); a video beyond that gets an invalid image -- score it from raw rows, where
 * the kernel works through the references in chunks.  bad_rows_out (device, optional, caller zeroes it): += reference rows that
 * held ids outside [-1, 65534], += 1 per invalid image. */
int64_t cst_refs_cooked_bytes(int32_t ref_width, int32_t refs_per_video);   /* -1: out of range */
int cst_refs_cook(const cst_df_table* table,
                  const void* refs, int32_t ref_dtype, int64_t n_rows, int32_t ref_width,
                  int64_t ref_stride, const int64_t* ref_offsets /* NULL: uniform */,
                  int32_t refs_per_video /* exact, or the maximum with ref_offsets */,
                  int64_t n_videos, int32_t use_eos,
                  void* images_out, int64_t image_stride, uint32_t* bad_rows_out,
                  void* stream);
/* cst_reward with the refs given as images; every other argument as in cst_reward. */
int cst_reward_cooked(const cst_df_table* table,
                      const void* cand, const void* greedy, int32_t cand_dtype,
                      int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                      const void* images, int64_t image_stride, const int64_t* image_index,
                      int32_t ref_width, int32_t refs_per_video, int64_t n_videos,
                      int32_t cands_per_video, int32_t use_eos,
                      int32_t mode, int32_t seq_per_img, int32_t scb_k,
                      const double* bcmrscores,
                      float* reward_out, double* scores_out, double* stats_out,
                      void* scratch, int64_t scratch_bytes, void* stream);

/* ---- leave-one-out document frequencies (GT consensus scores) ----------------------------
 * compute_scores.py:75-103 (make compute_evalscores, --remove_in_ref) scores caption i of every video
 * against the video's other captions with coco-caption's Cider in corpus mode: the document
 * frequency is recomputed from the references of that call, i.e. from all captions but the i-th
 * of every video -- twenty rebuilds of the table in the reference.  They differ from the df of ALL
 * captions only where caption i is a video's sole holder of an n-gram:
 *     df_i(g) = df_all(g) - #{videos whose caption i alone holds g}.
 * cst_df_sole_holders counts those per (caption, table slot) in one pass over the rows the table was
 * built from (cst_df_build_from_rows; `per_video` <= 32 captions per video, uniform):
 *     excl_out [dev] uint32 [per_video][cst_df_capacity(table)], zeroed by the caller;
 *     too_large_out [dev] uint32, optional, zeroed by the caller: += 1 per video whose captions hold
 *     more than 1,100 n-gram instances (not counted: fall back to one table per pass).
 * cst_df_leave_one_out then switches the table's idf to pass `caption` (idf = log N - log max(1,
 * df_all - excl[caption])) in one streaming kernel; scoring calls on the same stream afterwards see
 * that pass.  The table stays in that state: it belongs to the consensus computation. */
int cst_df_sole_holders(const cst_df_table* table, const void* rows, int32_t dtype, int64_t n_rows,
                        int32_t width, int64_t stride, int32_t per_video, int64_t n_videos,
                        int32_t use_eos, uint32_t* excl_out, uint32_t* too_large_out, void* stream);
int cst_df_leave_one_out(cst_df_table* table, const uint32_t* excl, int32_t caption, void* stream);

/* Baseline alone on given scores: utils.py:286-288 (WXE: scores are the
 * precomputed consensus scores) and utils.py:292-322.  scores [dev] float64
 * (n_rows, seq_per_img); for CST_BASE_GREEDY `other` holds the greedy scores, for
 * CST_BASE_SCB_GT the bcmrscores; `device` selects the GPU. */
int cst_scb_apply(const double* scores, const double* other, int64_t n_rows,
                  int32_t seq_per_img, int32_t mode, int32_t scb_k,
                  float* reward_out, double* stats_out, void* scratch,
                  int64_t scratch_bytes, int device, void* stream);

/* ---- the other reward scorers ------------------------------------------
 * train.py:119-124 picks the reward scorer by --eval_metric: 'CIDEr' (everything above), 'Bleu_4',
 * 'ROUGE_L' or 'METEOR'.  These two entry points replace Bleu().compute_score / Rouge().compute_score
 * of pycocoevalcap as called at utils.py:208, 274 (dict protocol: res[i] = [candidate string]) together
 * with the string building around them; rows, canonicalisation, the candidate -> video map and the
 * greedy row set are those of cst_cider_scores.  The rewards follow with cst_scb_apply on the score
 * vector (BLEU: row 3 of bleu_out, the Bleu_4 list -- utils.py:210-213, 276-279).  METEOR is a Java
 * program in the reference's checkout: no counterpart.
 *
 *   bleu_out    [dev] float64 [4][n_total], n_total = n_cand (2 n_cand with greedy rows, sampled first):
 *               per-sentence BLEU-1..4 of BleuScorer.compute_score(option='closest')
 *   comps_out   [dev] int32 [n_total][6] or NULL: testlen, closest reflen, clipped-correct counts of
 *               orders 1..4 -- the sums the corpus-level score is made of (guess_k = max(testlen-k+1, 0))
 *   scores_out  [dev] float64 [n_total]: Rouge.calc_score per candidate (beta = 1.2), bit-exact
 *   bad_rows_out [dev] uint32 or NULL: incremented by the number of rows holding ids outside
 *               [-1, 65534] (scored as if the bad id ended the caption)
 * Any number of references per video (they stream through in batches of 32); rows of at most 32
 * tokens.  A video without references scores as if its reference lengths were 0 (the reference
 * scorers raise there). */
int cst_bleu_scores(const void* cand, const void* greedy, int32_t cand_dtype,
                    int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                    const void* refs, int32_t ref_dtype, int32_t ref_width,
                    int64_t ref_stride, const int64_t* ref_offsets,
                    int32_t refs_per_video, int64_t n_videos,
                    int32_t cands_per_video, int32_t use_eos,
                    double* bleu_out, int32_t* comps_out, uint32_t* bad_rows_out,
                    int32_t device, void* stream);
int cst_rouge_scores(const void* cand, const void* greedy, int32_t cand_dtype,
                     int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                     const void* refs, int32_t ref_dtype, int32_t ref_width,
                     int64_t ref_stride, const int64_t* ref_offsets,
                     int32_t refs_per_video, int64_t n_videos,
                     int32_t cands_per_video, int32_t use_eos,
                     double* scores_out, uint32_t* bad_rows_out,
                     int32_t device, void* stream);

/* ---- fused consumer of the reward -------------------------------
 * Replaces RewardCriterion.forward (model.py:12-23) and its autograd backward as
 * called at train.py:206-210: loss = sum(-logprobs * reward * mask) / sum(mask) with
 * mask[i][0] = 1, mask[i][t] = (seq[i][t-1] > 0).  The reward is read through two
 * strides, so the per-caption column of cst_reward (row stride 1, column stride 0) is
 * consumed directly and the (N, T) broadcast of utils.py:224,322 never exists; a full
 * (N, T) reward works too.  seq [dev] int32/int64 (n_rows, width), logprobs [dev]
 * float32, loss_out / den_out [dev] one float each (den = sum(mask), kept for the
 * backward).  Scratch as for cst_reward: cst_criterion_scratch_bytes() bytes, zeroed
 * once.  Deterministic. */
int64_t cst_criterion_scratch_bytes(void);
int cst_reward_criterion_fwd(const void* seq, int32_t seq_dtype, int64_t seq_stride,
                             const float* logprobs, int64_t logp_stride,
                             const float* reward, int64_t reward_row_stride,
                             int64_t reward_col_stride, int64_t n_rows, int32_t width,
                             float* loss_out, float* den_out, void* scratch,
                             int64_t scratch_bytes, int device, void* stream);
/* grad_logprobs[i][t] = -grad_out * reward[i][t] * mask[i][t] / den (what autograd
 * derives from model.py:20-21); grad_out [dev] is the scalar upstream gradient. */
int cst_reward_criterion_bwd(const void* seq, int32_t seq_dtype, int64_t seq_stride,
                             const float* reward, int64_t reward_row_stride,
                             int64_t reward_col_stride, const float* den,
                             const float* grad_out, int64_t n_rows, int32_t width,
                             float* grad_logprobs, int64_t grad_stride, int device,
                             void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CST_REWARD_H_ */
