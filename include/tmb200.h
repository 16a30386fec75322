/*
 * tmb200.h — C ABI of libtmb200.so, the sm_100a kernels behind the classification-metric
 * update path (torchmetrics 1.3.0dev; citations are relative to /root/reference/src/torchmetrics/).
 *
 * Conventions shared by every entry point
 *   - plain pointers and sizes only; no torch types, nothing here allocates or frees device memory
 *     (scratch is caller-owned and passed in; the *_scratch_bytes functions are pure host code);
 *   - all data pointers are device pointers on the current device unless a comment says "host";
 *   - kernels are enqueued on `stream` (a cudaStream_t passed as void*) and the call returns without
 *     synchronising;
 *   - return value: 0 = enqueued, negative = TMB_E_* argument error (nothing enqueued),
 *     positive = cudaError_t of a failed launch;
 *   - integer metric states are int64 and are ACCUMULATED IN PLACE (`state += batch`), which is what
 *     lets MetricCollection compute groups keep sharing one tensor (collections.py:289-307);
 *   - data errors (a label outside [0, num_classes)) do not abort the kernel: the offending sample is
 *     skipped and a bit is OR-ed into the caller's device error word `err` (int32[1]); the host layer
 *     turns a non-zero word into RuntimeError like the reference's failing reshape/bincount would
 *     (functional/classification/confusion_matrix.py:335-337, utilities/data.py:199).
 */
#ifndef TMB200_H_
#define TMB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enums ------------------------------------------------------------------------------------ */

/* floating dtypes of `preds` */
#define TMB_F32 0
#define TMB_F16 1
#define TMB_BF16 2

/* integer dtypes of `target` (and of label-valued `preds`) */
#define TMB_I64 0
#define TMB_I32 1
#define TMB_I16 2
#define TMB_U8 3 /* also torch.bool */
#define TMB_I8 4

/* argument errors */
#define TMB_E_DTYPE (-1)
#define TMB_E_SHAPE (-2)
#define TMB_E_NULL (-3)
#define TMB_E_MODE (-4)
#define TMB_E_SCRATCH (-5)
#define TMB_E_LIMIT (-6)

/* bits of the device error word */
#define TMB_ERR_TARGET_RANGE 1 /* a non-ignored target outside [0, num_classes)            */
#define TMB_ERR_PREDS_RANGE 2  /* a label-valued pred outside [0, num_classes)              */
#define TMB_ERR_TARGET_BINARY 4 /* a non-ignored binary/multilabel target that is not 0 or 1 */
#define TMB_ERR_SCORE_RANGE 8   /* a negative score reached the exact-curve sort                */

/* what the multiclass kernels accumulate */
#define TMB_MC_CONFMAT 0 /* out = int64 state [C, C], row = target, col = pred                    */
#define TMB_MC_STATS 1   /* out = int64 scratch [3C]: hist_target[C] hist_pred[C] match[C]        */
#define TMB_MC_MICRO 2   /* out = int64 scratch [2]: number of matches, number of counted samples */
#define TMB_MC_BOTH 3    /* out = int64 scratch [C*C + 3C]: the batch's confusion matrix, then the STATS layout */

/* Library / build identification: returns the compiled-for SM (100) and writes nothing. */
int tmb_build_arch(void);

/* ---- K1: multiclass argmax + histogram ---------------------------------------------------------
 *
 * tmb_mc_argmax_hist replaces, in one pass over `preds`:
 *   _multiclass_confusion_matrix_format  functional/classification/confusion_matrix.py:306-330
 *   _multiclass_confusion_matrix_update  functional/classification/confusion_matrix.py:333-337
 *   _multiclass_stat_scores_format       functional/classification/stat_scores.py:325-341
 *   _multiclass_stat_scores_update       functional/classification/stat_scores.py:394-418 (global, top_k=1)
 *   _bincount                            utilities/data.py:169-199
 *
 * preds  : float tensor [n_outer, num_classes, n_inner] contiguous (n_inner = 1 for [N, C]).
 * target : integer tensor [n_outer, n_inner] contiguous.
 * Per sample (n, s): skip if has_ignore && target == ignore_index; p = argmax_c preds[n, c, s] with
 * torch.argmax's rule (first maximal index, NaN is maximal); then accumulate according to `mode`.
 */
int tmb_mc_argmax_hist(const void* preds, int preds_dtype, const void* target, int target_dtype,
                       int64_t n_outer, int64_t num_classes, int64_t n_inner, int has_ignore,
                       int64_t ignore_index, int mode, int64_t* out, int32_t* err, void* stream);

/* Same accumulation for label-valued preds (preds.ndim == target.ndim): the int branch of
 * _multiclass_confusion_matrix_format / _multiclass_stat_scores_format (no argmax). */
int tmb_mc_label_hist(const void* preds, int preds_dtype, const void* target, int target_dtype,
                      int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index, int mode,
                      int64_t* out, int32_t* err, void* stream);

/* tmb_mc_topk_hist: the top_k > 1 branch of
 *   _multiclass_stat_scores_update      functional/classification/stat_scores.py:363-393
 *   select_topk (torch.topk + scatter)  utilities/data.py:115-139
 * preds [n_rows, num_classes] row-major float scores (num_classes >= 2; up to 1024 classes a warp keeps the row
 * in registers, wider rows are walked once per selection round), target one integer per row.  Adds into `scratch` (int64 [3 * num_classes], the TMB_MC_STATS layout: rows per target class |
 * rows with the class among the top_k scores | rows whose target is among its top_k); follow with
 * tmb_mc_stats_finalize(mode TMB_MC_STATS).  Ties at the k-th score go to the lower class index, NaN ranks highest
 * (torch.topk's CUDA order).  Rows with target == ignore_index are dropped. */
int tmb_mc_topk_hist(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n_rows,
                     int64_t num_classes, int64_t top_k, int has_ignore, int64_t ignore_index, int64_t* scratch,
                     int32_t* err, void* stream);

/* Turns the TMB_MC_STATS / TMB_MC_MICRO scratch into `tp/fp/tn/fn += ...`
 * (functional/classification/stat_scores.py:401-404 and :414-418) and re-zeroes the scratch so the
 * next update needs no memset.  tp/fp/tn/fn are int64 [C] (STATS) or [1] (MICRO). */
int tmb_mc_stats_finalize(int64_t* scratch, int64_t num_classes, int mode, int64_t* tp, int64_t* fp,
                          int64_t* tn, int64_t* fn, void* stream);

/* multidim_average="samplewise" for multiclass stat scores (the other half of SURVEY §8 row f3):
 *   _multiclass_stat_scores_update with sum_dim = [1]   functional/classification/stat_scores.py:363-393
 * preds [n_samples, num_classes, n_inner] float scores (argmax fused) or, with preds_is_label, [n_samples, n_inner]
 * integer labels; target [n_samples, n_inner].  Fills scratch (int64 [n_samples, 3 * num_classes], zero on entry)
 * with the TMB_MC_STATS histograms of every sample; tmb_mc_stats_finalize_samplewise WRITES tp/fp/tn/fn (int64
 * [n_samples, num_classes]) from them and re-zeroes the scratch. */
int tmb_mc_hist_samplewise(const void* preds, int preds_dtype, int preds_is_label, const void* target,
                           int target_dtype, int64_t n_samples, int64_t num_classes, int64_t n_inner,
                           int has_ignore, int64_t ignore_index, int64_t* scratch, int32_t* err, void* stream);
int tmb_mc_stats_finalize_samplewise(int64_t* scratch, int64_t n_samples, int64_t num_classes, int64_t* tp,
                                     int64_t* fp, int64_t* tn, int64_t* fn, void* stream);

/* tmb_mc_topk_hist_samplewise: top_k > 1 together with multidim_average="samplewise" — the `sum_dim = [1]` form of
 * the same branch (functional/classification/stat_scores.py:363-393).  preds [n_samples * rows_per_sample,
 * num_classes] row-major (the extra dimensions of a sample folded into its rows), target one integer per row; the
 * TMB_MC_STATS histograms of sample n are ADDED into scratch[n * 3 * num_classes ...] (int64 [n_samples, 3 *
 * num_classes], zero-initialised by the caller); follow with tmb_mc_stats_finalize_samplewise. */
int tmb_mc_topk_hist_samplewise(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                int64_t n_samples, int64_t rows_per_sample, int64_t num_classes, int64_t top_k,
                                int has_ignore, int64_t ignore_index, int64_t* scratch, int32_t* err, void* stream);

/* One argmax for a whole MetricCollection (collections.py:200-226 calls every compute group's
 * update, so Accuracy + ConfusionMatrix read the logits twice and run argmax twice): tmb_mc_argmax_hist /
 * tmb_mc_label_hist with mode TMB_MC_BOTH fill one scratch, then
 *   tmb_mc_shared_confmat  adds scratch[0 .. C*C) into the int64 [C, C] confusion-matrix state and zeroes it;
 *   tmb_mc_shared_stats    adds the stat scores into tp/fp/tn/fn — int64 [C], or [1] summed over classes when
 *                          micro != 0 (functional/classification/stat_scores.py:394-418) — once per subscribing
 *                          metric; pass clear != 0 on the last call to zero the histograms. */
int tmb_mc_shared_confmat(int64_t* scratch, int64_t num_classes, int64_t* confmat, void* stream);
int tmb_mc_shared_stats(int64_t* scratch, int64_t num_classes, int micro, int clear, int64_t* tp, int64_t* fp,
                        int64_t* tn, int64_t* fn, void* stream);

/* Scratch size in bytes for a mode (0 for TMB_MC_CONFMAT). Host only. */
int64_t tmb_mc_scratch_bytes(int64_t num_classes, int mode);

/* ---- K2 / K1b: one-pass threshold binning ----------------------------------------------------------
 *
 * tmb_thresh_bin_update + tmb_thresh_bin_finalize replace
 *   _binary_precision_recall_curve_format/_update (+_vectorized, _loop)
 *                                       functional/classification/precision_recall_curve.py:162-250
 *   _multilabel_precision_recall_curve_format/_update              same file :739-793
 *   _binary/_multilabel_stat_scores_format/_update     functional/classification/stat_scores.py:90-131, 647-681
 *   _binary/_multilabel_confusion_matrix_format/_update
 *                                       functional/classification/confusion_matrix.py:118-153, 486-526
 *
 * preds  : float (or 0/1 integer labels) [n_outer, num_labels, n_inner] contiguous; binary = num_labels 1.
 * target : integers of the same shape.
 * thr    : T float32 thresholds in ASCENDING order, device memory.
 * Per non-ignored element: q = preds, or sigmoid(preds) when the batch has any value outside [0,1]
 * (decided on the device; `flag_counts_ignored` says whether ignored positions take part in that test:
 * 0 for binary curves / binary confusion matrix, 1 for stat scores and every multilabel metric);
 * k = #{j : thr[j] <= q} (`cmp_gt` = 0, curves) or #{j : thr[j] < q} (`cmp_gt` = 1, the strict `preds >
 * threshold` of stat scores / confusion matrices); hist[label][target][k] += 1.  NaN gives k = 0.
 * The histogram lives in the caller's int64 scratch, which must be zero before the first call and is
 * re-zeroed by tmb_thresh_bin_finalize.  (Where the reference compares in a narrow dtype, the host
 * rounds thr to that dtype before the call.)
 */
#define TMB_TB_CURVE 0   /* out0 = int64 state [T, L, 2, 2] (binary: [T, 2, 2])                         */
#define TMB_TB_CONFMAT 1 /* T == 1: out0 = int64 state [L, 2, 2] (binary: [2, 2]), [l][target][pred]     */
#define TMB_TB_STATS 2   /* T == 1: out0..out3 = tp, fp, tn, fn, int64 [L] each                          */

/* int64 elements of scratch for (num_labels, T). Host only. */
int64_t tmb_thresh_bin_scratch_elems(int64_t num_labels, int64_t num_thresholds);

int tmb_thresh_bin_update(const void* preds, int preds_dtype, int preds_is_label, const void* target,
                          int target_dtype, int64_t n_outer, int64_t num_labels, int64_t n_inner,
                          const float* thr, int64_t num_thresholds, int cmp_gt, int has_ignore,
                          int64_t ignore_index, int flag_counts_ignored, int64_t* scratch,
                          int64_t scratch_elems, int32_t* err, void* stream);

/* Picks the raw or the sigmoid histogram, forms the per-threshold counts
 * (TP_j = sum_{k>j} hist[l][1][k], FP_j likewise from hist[l][0], FN_j = n1 - TP_j, TN_j = n0 - FP_j;
 * precision_recall_curve.py:246-249) and ADDS them into the state in the requested layout.
 * `perm` (int32[T], device, may be NULL) maps the sorted threshold position to the row of the state,
 * for thresholds the user did not give in ascending order.  Re-zeroes the scratch. */
int tmb_thresh_bin_finalize(int64_t* scratch, int64_t scratch_elems, int64_t num_labels,
                            int64_t num_thresholds, const int32_t* perm, int layout, int64_t* out0,
                            int64_t* out1, int64_t* out2, int64_t* out3, void* stream);

/* tmb_binary_curve_update_noalloc — no counterpart in the reference (which allocates temporaries freely
 * through torch); it exists for callers of this C ABI that have no way to provide scratch memory:
 *   - a host binding the library over a bare FFI (ctypes without torch, cgo, JNI) that owns no device
 *     allocator and only holds the input / state pointers it was handed;
 *   - a service that gives all device memory to one pool at start-up, where a later cudaMalloc fails or
 *     fragments that pool;
 *   - a latency-bound evaluation loop: cudaMalloc and cudaFree synchronise the device.
 * It performs one whole binned binary-curve update — tmb_thresh_bin_update (num_labels 1, `>=` compare,
 * ignored samples excluded from the sigmoid test) followed by tmb_thresh_bin_finalize (TMB_TB_CURVE) — and
 * ADDS the result into `confmat` (int64 [num_thresholds, 2, 2]).  The caller still passes inputs, thresholds,
 * state and error word; the histogram scratch is a static __device__ array inside the library (about 36 KB on
 * every GPU that loads it, kept until the process exits).  The call itself allocates and frees nothing.
 *   Thresholds: `thr` is a device array of float32 in ASCENDING order (as for tmb_thresh_bin_update; an unsorted
 *     user list has to be sorted by the caller, who then permutes the rows of `confmat` back).
 *   Limit: num_thresholds <= 1024; more returns TMB_E_LIMIT and launches nothing.  All argument checks (dtype
 *     codes, sizes, NULL pointers) run before anything is enqueued: a negative return leaves the array untouched.
 *   Threads: the per-device address cache is an atomic pointer, so host threads may make the first call together.
 *   Set-up: the first call on each device looks up the array's device address (cudaGetSymbolAddress) with that
 *     device current.  Make that first call outside any stream capture.
 *   Sharing: every call on a device uses the same array, and the array keeps the previous call's contents.
 *     The call clears what has to start at zero before using it, but two calls whose kernels overlap on the GPU
 *     (different streams, or different host threads) race on it — serialise them on one stream.
 */
int tmb_binary_curve_update_noalloc(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                    int64_t n, const float* thr, int64_t num_thresholds, int has_ignore,
                                    int64_t ignore_index, int64_t* confmat, int32_t* err, void* stream);

/* ---- K3: exact curves (thresholds=None) ---------------------------------------------------------------
 *
 * tmb_sigmoid_normalize is the exact-mode half of
 *   _binary_precision_recall_curve_format      functional/classification/precision_recall_curve.py:176-186
 *   _multilabel_precision_recall_curve_format  same file :752-757
 * out[i] = sigmoid(preds[i]) if ANY element taking part in the test lies outside [0,1] (NaN included),
 * else out[i] = preds[i]; same dtype as preds.  One read of the data either way: the testing pass copies as it goes
 * and stops at the first outside value, the sigmoid pass runs only if there was one.  With `target` non-NULL and has_ignore, elements whose target
 * equals ignore_index take no part in the test.  `flag` is one caller-owned int64 that must start at 0 and
 * `generation` a strictly increasing positive call counter for that flag (it replaces clearing the flag).
 */
int tmb_sigmoid_normalize(const void* preds, int preds_dtype, int64_t n, const void* target,
                          int target_dtype, int has_ignore, int64_t ignore_index, void* out, int64_t* flag,
                          int64_t generation, void* stream);

/* ---- K5: multiclass curves --------------------------------------------------------------
 *
 * The conditional row softmax of
 *   _multiclass_precision_recall_curve_format   functional/classification/precision_recall_curve.py:423-458
 * ("if not torch.all((preds >= 0) * (preds <= 1)): preds = preds.softmax(1)", evaluated after the rows with
 * target == ignore_index were dropped), in the two-call flag protocol of tmb_sigmoid_normalize:
 *
 * tmb_rows_outside_flag: preds [n_rows, num_classes] row-major; raises *flag to `generation` if any element of a
 *   row whose target (one integer per row; NULL = every row counts) differs from ignore_index lies outside [0,1]
 *   or is NaN.  Any num_classes.
 * tmb_softmax_rows: out = softmax(preds, dim=1) if *flag == generation; otherwise `out` is left as it is (pass B of
 *   tmb_softmax_normalize: pass A has already copied the rows).  Same dtype as preds.  The result is bit-identical to
 *   torch.softmax on the device for num_classes <= 1024 (its persistent warp kernel's lane layout, reduction order,
 *   expf and IEEE division are reproduced); wider rows return TMB_E_LIMIT and the caller keeps torch's own kernel.
 * tmb_softmax_normalize: the whole conditional normalisation with ONE read of the data: pass A copies preds to out
 *   while it tests the rows that count and stops at the first value outside [0,1] (for logits: after the first
 *   wave of CTAs); pass B (tmb_softmax_rows) overwrites out with the softmax only if pass A raised the flag.
 *   num_classes <= 1024.  `flag` / `generation` as for tmb_sigmoid_normalize.
 */
int tmb_rows_outside_flag(const void* preds, int preds_dtype, int64_t n_rows, int64_t num_classes,
                          const void* target, int target_dtype, int has_ignore, int64_t ignore_index,
                          int64_t* flag, int64_t generation, void* stream);
int tmb_softmax_normalize(const void* preds, int preds_dtype, int64_t n_rows, int64_t num_classes,
                          const void* target, int target_dtype, int has_ignore, int64_t ignore_index, void* out,
                          int64_t* flag, int64_t generation, void* stream);
int tmb_softmax_rows(const void* preds, int preds_dtype, int64_t n_rows, int64_t num_classes, void* out,
                     const int64_t* flag, int64_t generation, void* stream);

/* tmb_mc_curve_update: the binned multiclass curve update,
 *   _multiclass_precision_recall_curve_format + _update (+_vectorized / _loop)
 *       functional/classification/precision_recall_curve.py:423-527
 * in one pass over the logits: conditional softmax (as above, ignored rows excluded from the range test), every
 * (row, class) probability binned against thr[0..T) with `>=`, one-hot target `class == target[row]`.
 * preds [n_rows, num_classes] row-major (2 <= num_classes <= 1024, else TMB_E_LIMIT), target one integer per row
 * (outside [0, num_classes) and not ignored: TMB_ERR_TARGET_RANGE).  micro != 0 folds all classes into one binary
 * task (average="micro").  scratch: tmb_thresh_bin_scratch_elems(micro ? 1 : num_classes, T), zero before first
 * use; follow with tmb_thresh_bin_finalize(num_labels = micro ? 1 : num_classes, layout TMB_TB_CURVE), which adds
 * into the int64 [T, num_classes, 2, 2] (micro: [T, 2, 2]) state and re-zeroes the scratch. */
int tmb_mc_curve_update(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n_rows,
                        int64_t num_classes, const float* thr, int64_t num_thresholds, int has_ignore,
                        int64_t ignore_index, int micro, int64_t* scratch, int64_t scratch_elems, int32_t* err,
                        void* stream);

/* tmb_clf_curves replaces, for all labels at once,
 *   _binary_clf_curve                          functional/classification/precision_recall_curve.py:28-80
 *   the per-label loops that call it           same file :819-829, functional/classification/roc.py:345-355
 * preds [n, num_labels] row-major scores (non-negative or NaN — what the update stores; a negative score sets
 * TMB_ERR_SCORE_RANGE), target [n, num_labels] integers; entries with target == ignore_index are dropped
 * per label.  For label l with D_l distinct scores the first D_l entries of fps/tps/thresholds[l * n ...]
 * receive, in order of DESCENDING score: false positives, true positives (both exact counts as float32,
 * positives = target == pos_label) and the score; counts[l] = D_l (int64, device).  n < 2^31.
 * scratch: tmb_clf_curve_scratch_bytes(n, num_labels) bytes of device memory, contents irrelevant. */
int64_t tmb_clf_curve_scratch_bytes(int64_t n, int64_t num_labels);
int tmb_clf_curves(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n,
                   int64_t num_labels, int has_ignore, int64_t ignore_index, int64_t pos_label, void* scratch,
                   int64_t scratch_bytes, float* fps, float* tps, float* thresholds, int64_t* counts,
                   int32_t* err, void* stream);

/* The same, reading the scores / targets from the per-update buffers of the metric state instead of one
 * concatenated tensor — it also replaces the `dim_zero_cat(self.preds)`, `dim_zero_cat(self.target)` of
 *   the exact-mode state assembly in compute()   classification/precision_recall_curve.py:175, :373, :563,
 *                                                classification/auroc.py:122, :425, classification/average_precision.py:116, :428
 * chunk_table (device memory, 3 * num_chunks + 1 int64):
 *   [0 .. C]       first row of chunk c, ascending, table[0] = 0 and table[C] = n
 *   [C+1 .. 2C]    device address of chunk c's scores  ([rows_c, num_labels] row-major, preds_dtype)
 *   [2C+1 .. 3C]   device address of chunk c's targets ([rows_c, num_labels] row-major, target_dtype)
 * Empty chunks are allowed.  class_targets != 0 is the multiclass form (the per-class loop of
 * _multiclass_precision_recall_curve_compute, precision_recall_curve.py:548-555, `pos_label=i`): the target chunks
 * then hold ONE class index per row ([rows_c]) and label l is positive where target == l (pos_label unused).
 * Everything else as tmb_clf_curves. */
int tmb_clf_curves_chunked(const int64_t* chunk_table, int64_t num_chunks, int preds_dtype, int target_dtype,
                           int64_t n, int64_t num_labels, int has_ignore, int64_t ignore_index, int64_t pos_label,
                           int class_targets, void* scratch, int64_t scratch_bytes, float* fps, float* tps,
                           float* thresholds, int64_t* counts, int32_t* err, void* stream);

/* Chunks -> the four per-label area reductions tmb_curve_reduce documents (scores: float64 [num_labels, 4],
 * device), without writing the curves: pack, sort, then one pass over the sorted keys.  Replaces, for the
 * exact-mode AUROC / average-precision metrics, tmb_clf_curves + tmb_curve_reduce, i.e.
 *   _binary_auroc_compute / _multilabel_auroc_compute               functional/classification/auroc.py:82-106, :307-332
 *   _binary_ / _multilabel_average_precision_compute                functional/classification/average_precision.py:70-75, :284-311
 * want_ap = 0 skips the average-precision sum (column 1 stays 0; it costs a float64 division per curve point
 * that AUROC does not need).  Same scratch size as tmb_clf_curves. */
int tmb_clf_curve_scores_chunked(const int64_t* chunk_table, int64_t num_chunks, int preds_dtype, int target_dtype,
                                 int64_t n, int64_t num_labels, int has_ignore, int64_t ignore_index,
                                 int64_t pos_label, int class_targets, int want_ap, void* scratch,
                                 int64_t scratch_bytes, double* scores, int32_t* err, void* stream);

/* ---- label-sharded exact compute over several GPUs (K4 for "cat" states) -----------------------------------
 *
 * Replaces, for the scalar exact-mode metrics (AUROC / AveragePrecision with thresholds=None), the reference's
 *   Metric._sync_dist                     metric.py:423-453   (pre-concatenate the list states, all_gather each)
 *   gather_all_tensors                    utilities/distributed.py:97-147
 * followed by every rank sorting every label
 *   _multilabel_roc_compute / _binary_clf_curve   functional/classification/roc.py:345-355, precision_recall_curve.py:28-80
 * The labels are dealt out in blocks of `labels_per_rank` (rank r owns [r * labels_per_rank, ...)).
 *
 * tmb_clf_pack_keys_sharded: rank s turns its local chunks (same chunk table as tmb_clf_curves_chunked) into sort keys
 * and stores the keys of label l directly into the OWNER's key buffer, rows [row_offset, row_offset + n_local) of that
 * label's n_total-row segment: peer_keys[r] / peer_nvalid[r] (device arrays of `world` addresses) are rank r's key
 * buffer ([labels_per_rank, n_total] uint32) and its n_valid array (uint64 [labels_per_rank], zeroed by its owner
 * before any rank packs) as mapped into THIS process (peer / symmetric memory; ordinary stores and atomics over
 * NVLink).  The caller separates the calls of different ranks from the owner's use of the buffer with barriers.
 *
 * tmb_clf_sorted_scores: sort `num_labels` segments of `n` packed keys in place and reduce them to the same
 * [num_labels, 4] float64 rows as tmb_clf_curve_scores_chunked.  all_rows_valid != 0: no entry was ignored, n_valid is
 * filled with n here.  Scratch: tmb_clf_sorted_scores_scratch_bytes(n, num_labels).
 */
int tmb_clf_pack_keys_sharded(const int64_t* chunk_table, int64_t num_chunks, int preds_dtype, int target_dtype,
                              int64_t n_local, int64_t num_labels, int has_ignore, int64_t ignore_index,
                              int64_t pos_label, int class_targets, const uint64_t* peer_keys,
                              const uint64_t* peer_nvalid, int64_t world, int64_t labels_per_rank, int64_t n_total,
                              int64_t row_offset, int32_t* err, void* stream);
int64_t tmb_clf_sorted_scores_scratch_bytes(int64_t n, int64_t num_labels);
int tmb_clf_sorted_scores(uint32_t* keys, uint64_t* n_valid, int64_t n, int64_t num_labels, int all_rows_valid,
                          int want_ap, void* scratch, int64_t scratch_bytes, double* scores, void* stream);

/* Area reductions over the curves tmb_clf_curves emitted (same fps / tps buffers, label stride n, counts[l] points
 * each), replacing the per-label torch chains of
 *   _binary_roc_compute + _auc_compute_without_check   functional/classification/roc.py:53-78, utilities/compute.py:88-96
 *   _binary_precision_recall_curve_compute + AP sum     precision_recall_curve.py:274-283, average_precision.py:43-53
 * out (float64 [num_labels, 4], device) receives per label: the un-normalised trapezoid area under (fps, tps)
 * starting from (0,0); sum_b (tps_b - tps_{b-1}) * tps_b / (tps_b + fps_b); the last fps; the last tps.
 * AUROC = out0 / (out2 * out3) (0 when either total is 0), AP = out1 / out3. */
int tmb_curve_reduce(const float* fps, const float* tps, int64_t n, int64_t num_labels, const int64_t* counts,
                     double* out, void* stream);

/* Order-preserving removal of the samples whose target equals ignore_index — the boolean-mask indexing of
 * precision_recall_curve.py:178-181 (binary) and auroc.py:317-322 (micro average).  out_* must hold n
 * elements; *count (int64, device) receives how many were kept.  preds is copied bytewise
 * (preds_elem_bytes in {1,2,4,8}).  n < 2^32. */
int64_t tmb_compact_scratch_bytes(int64_t n);
int tmb_compact_not_ignored(const void* preds, int preds_elem_bytes, const void* target, int target_dtype,
                            int64_t n, int64_t ignore_index, void* out_preds, void* out_target,
                            int64_t* count, void* scratch, int64_t scratch_bytes, void* stream);

/* ---- utilities ------------------------------------------------------------------------------------
 * tmb_bincount: out[v] += 1 for every v = x[i] (integer dtype), 0 <= v < length; other values set
 * TMB_ERR_PREDS_RANGE.  Replaces torch.bincount inside _bincount (utilities/data.py:169-199). */
int tmb_bincount(const void* x, int x_dtype, int64_t n, int64_t length, int64_t* out, int32_t* err,
                 void* stream);
/* ---- K4: state sync helpers ------------------------------------------------------------------------------------
 * Metric._sync_dist (metric.py:423-453) calls gather_all_tensors (utilities/distributed.py:97-147) once per state:
 * a barrier and two all-gathers each, then a stack and a sum.  Here every "sum" state of a metric travels in ONE
 * all-reduce over a packed buffer (the collective itself is the host's: NCCL through torch.distributed).
 *   tmb_pack_states    copies the k states into `packed` (one launch, any k);
 *   tmb_unpack_states  copies `packed` back into the k states (hosts without tensor views; the Python side simply
 *                      views the reduced buffer).
 * table: device int64 [2k + 1] = byte offsets of the states inside `packed` (k + 1 entries, the last = total_bytes,
 * all multiples of 4), then the k device addresses of the states. */
int tmb_pack_states(const int64_t* table, int64_t num_states, int64_t total_bytes, void* packed, void* stream);
int tmb_unpack_states(const int64_t* table, int64_t num_states, int64_t total_bytes, const void* packed, void* stream);

/* float32 sum of x[0..n) in index order, one rounding per addition -> *out.  Replaces Python's `sum(tk * pk)` over a
 * device tensor in _matthews_corrcoef_reduce (functional/classification/matthews_corrcoef.py:59-61): the covariances
 * subtract numbers of size N^2, so the order of the additions decides bits that the cancellation magnifies. */
int tmb_sum_sequential_f32(const float* x, int64_t n, float* out, void* stream);
/* Largest value of x (int64 result, INT64_MIN for n == 0) -> *out_max, used to size a bincount. */
int tmb_label_max(const void* x, int x_dtype, int64_t n, int64_t* out_max, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TMB200_H_ */
