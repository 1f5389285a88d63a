/* mlc.h - C ABI of libmlcomp_b200.so
 *
 * B200 (sm_100a) implementation of MLComp's post-inference scoring and
 * post-processing hot path.  The reference (deepalcoholic/mlcomp 20.3.1.1) is
 * pure Python: this path has no FFI there.  Each entry point therefore names
 * the Python function(s) of the reference whose arithmetic it replaces
 * (paths relative to the reference root); INTEGRATION.md shows the ctypes
 * stub a maintainer would add on the reference side.
 *
 * Conventions
 *  - every function returns MLC_OK (0) or a negative MLC_ERR_* code; the text
 *    of the last error on the calling thread is mlc_last_error();
 *  - "dev" pointers are device memory owned by the caller, "host" pointers
 *    are ordinary host memory read before the call returns;
 *  - nothing here allocates or frees device memory; scratch is passed in;
 *  - `stream` is a cudaStream_t (NULL = legacy default stream); all work is
 *    enqueued asynchronously on it; no call synchronises unless it says so;
 *  - planes are (image, class) pairs, plane index p = image * n_cls + class;
 *    logits are row-major [n_img][n_cls][H][W] per source.
 */
#ifndef MLC_H_
#define MLC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MLC_OK               0
#define MLC_ERR_ARG        (-1)  /* bad argument (null, range, alignment)   */
#define MLC_ERR_CUDA       (-2)  /* a CUDA runtime call or launch failed     */
#define MLC_ERR_UNSUPPORTED (-3) /* valid request this build cannot serve    */
#define MLC_ERR_CAPACITY   (-4)  /* caller-provided buffer too small         */

/* element type of logits */
#define MLC_F16 0
#define MLC_F32 1

/* activation: utils/torch.py:11-18 apply_activation */
#define MLC_ACT_NONE    0
#define MLC_ACT_SIGMOID 1
#define MLC_ACT_SOFTMAX 2  /* over the class dim (dim 1) of each source */

/* Composed inverse test-time transform of one source, i.e. what
 * TtaWrap.inverse (contrib/transform/tta.py:20-31) does to it:
 *   out[y][x] = src[ys][xs],  (u, v) = TRANSPOSE ? (x, y) : (y, x),
 *   ys = VFLIP ? Hs-1-u : u,  xs = HFLIP ? Ws-1-v : v,
 * (Hs, Ws) = the source's own plane shape = TRANSPOSE ? (W, H) : (H, W). */
#define MLC_XF_HFLIP     1
#define MLC_XF_VFLIP     2
#define MLC_XF_TRANSPOSE 4

#define MLC_MAX_SOURCES  32   /* TTA variants x ensemble members            */
#define MLC_MAX_THR      255  /* thresholds in one search grid              */
#define MLC_MAX_AREAS    8    /* min-area candidates per class              */
#define MLC_MAX_CLASSES_SEG 64
#define MLC_LUT_BYTES    69632 /* 64 KiB byte table + 4 KiB cell table   */

int mlc_version(void);
const char *mlc_last_error(void);

/* ---- classification ---------------------------------------------------
 * Replaces  preds.argmax(axis=1)  +  sklearn confusion_matrix(targets,
 * argmax, labels=arange(C))  (worker/reports/classification.py:133-138) and
 * the per-sample hit (examples/digit-recognizer/executors/valid.py:47).
 *   preds_dev   [n][n_cls] f16/f32 row-major (logits or probabilities)
 *   labels_dev  [n] int64; rows whose label is outside [0,n_cls) are dropped
 *               from cm (sklearn's rule) and counted in stats[1]
 *   cm_dev      [n_cls][n_cls] int64, rows = truth, cols = prediction;
 *               ADDED to (zero it for a fresh matrix) so chunks accumulate
 *   stats_dev   [2] int64, ADDED to: {#argmax==label, #dropped}
 *   argmax_dev  [n] int32 or NULL: first index of the row maximum (NaN
 *               counts as maximal, like numpy)
 */
int mlc_confusion(const void *preds_dev, int dtype, int64_t n, int n_cls,
                  const int64_t *labels_dev, int64_t *cm_dev,
                  int64_t *stats_dev, int32_t *argmax_dev, void *stream);

/* ---- segmentation: fused TTA / ensemble mean + threshold histogram -----
 * Replaces, for every pixel, apply_activation (utils/torch.py:11-18) ->
 * TtaWrap.inverse / flip (contrib/transform/tta.py:20-31,
 * contrib/torch/tensors.py:4-8) -> mean over sources -> the `>=`, `.sum()`
 * and `targets*outputs` passes of dice_numpy (contrib/metrics/dice.py:6-11,
 * 20) for ALL thresholds at once.
 *
 * cnt(p) = #{k : p >= thr[k]} (fp32 compare, thr ascending); NaN -> 0.
 *   hist_dev [n_img*n_cls][2][n_thr+1] uint32: hist[p][g][j] = pixels of
 *   plane p with ground truth g (gt byte != 0) and cnt == j.  Overwritten.
 *
 *   src_dev_ptrs host array of n_src device pointers, each [n_img][n_cls]
 *               [Hs][Ws] of `dtype`; xform host array of MLC_XF_* masks
 *   gt_dev      [n_img][n_cls][H][W] uint8 one-hot planes
 *               (contrib/dataset/segment.py:36-56 builds them as fp32)
 *   thr_host    [n_thr] fp32 ascending, n_thr <= MLC_MAX_THR
 *   lut_dev     NULL, or MLC_LUT_BYTES made by mlc_seg_lut_build with the
 *               same thr/act (NONE or SIGMOID): enables the table path for
 *               a single untransformed fp16 source (results are identical).  Holds
 *               cnt(act(x)) for each of the 65 536 fp16 codes, then a
 *               2 048-entry summary of it (one entry per 32 codes)
 */
int mlc_seg_lut_build(const float *thr_host, int n_thr, int act,
                      uint8_t *lut_dev, void *stream);

int mlc_seg_hist(const void *const *src_dev_ptrs, const int32_t *xform,
                 int n_src, int dtype, const uint8_t *gt_dev, int64_t n_img,
                 int n_cls, int H, int W, const float *thr_host, int n_thr,
                 int act, const uint8_t *lut_dev, uint32_t *hist_dev,
                 void *stream);

/* mlc_seg_hist with a second output: the bin index cnt(p) of every pixel.
 * No reference counterpart as a function - it is the same pass (the lines
 * cited above), keeping what the reference would re-compute: once the search
 * has chosen thr[k], (p >= thr[k]) is (bin > k), so the masks for
 * mlc_rle_encode (dtype MLC_U8, act NONE, thr = k + 0.5) or for a saved
 * prediction come from 1 byte per pixel instead of all the sources again.
 *   bins_dev   [n_img][n_cls][H][W] uint8 (n_thr <= 255), overwritten
 * Always served by the multi-source kernels (no lut). */
int mlc_seg_hist_bins(const void *const *src_dev_ptrs, const int32_t *xform,
                      int n_src, int dtype, const uint8_t *gt_dev,
                      int64_t n_img, int n_cls, int H, int W,
                      const float *thr_host, int n_thr, int act,
                      uint32_t *hist_dev, uint8_t *bins_dev, void *stream);

/* Same fusion, writing the fp32 probabilities [n_img][n_cls][H][W] instead
 * (what utils/torch.py:24-29 `_infer` returns after the mean). */
int mlc_seg_prob(const void *const *src_dev_ptrs, const int32_t *xform,
                 int n_src, int dtype, int64_t n_img, int n_cls, int H, int W,
                 int act, float *prob_dev, void *stream);

/* ---- segmentation: arg-max masks ----------------------------------------
 * The multiclass deployment: per pixel the first maximal class of the fused
 * class scores.  Replaces apply_activation(x, 'softmax') (utils/torch.py:16-17)
 * -> TtaWrap.inverse (contrib/transform/tta.py:20-31) -> mean over sources ->
 * numpy arg-max over the class axis (the rule of worker/reports/
 * classification.py:111,136: a NaN beats everything, ties go to the lowest
 * index) -> the sums of dice_numpy (contrib/metrics/dice.py:10-11,20) on the
 * resulting one-hot masks.  Scores are computed exactly as mlc_seg_prob does,
 * so label == argmax(mlc_seg_prob) bit for bit.  act: NONE, SIGMOID, SOFTMAX.
 *   gt_onehot_dev [n_img][n_cls][H][W] uint8 or NULL (non-zero = set)
 *   gt_label_dev  [n_img][H][W] uint8 or NULL; values >= n_cls are ignored
 *                 (at most one of the two)
 *   label_dev     [n_img][H][W] uint8 class index (n_cls <= 255), or NULL
 *   onehot_dev    [n_img][n_cls][H][W] uint8 {0,1} = (label == class), or
 *                 NULL: the masks mlc_rle_encode (MLC_U8, thr 0.5) and the
 *                 report builders take
 *   counts_dev    [n_img][n_cls][3] int64 {T, P, I} of those masks, or NULL
 *                 (T and I are 0 without a truth), overwritten
 *   cm_dev        [n_cls][n_cls] int64 pixel confusion, rows = truth, cols =
 *                 prediction, or NULL (needs gt_label_dev), overwritten */
int mlc_seg_argmax(const void *const *src_dev_ptrs, const int32_t *xform,
                   int n_src, int dtype, int64_t n_img, int n_cls, int H,
                   int W, int act, const uint8_t *gt_onehot_dev,
                   const uint8_t *gt_label_dev, uint8_t *label_dev,
                   uint8_t *onehot_dev, int64_t *counts_dev, int64_t *cm_dev,
                   void *stream);

/* ---- segmentation: histograms -> dice for the whole search grid --------
 * Replaces the loop of dice_numpy calls (contrib/metrics/dice.py:4-23) over
 * thresholds x min-areas x planes.  From hist: T = sum_j hist[1][j],
 * P[k] = sum_{j>k} (hist[0][j]+hist[1][j]), I[k] = sum_{j>k} hist[1][j];
 * min-area rule dice.py:13-15, empty rule :17-18, 2I/(T+P+eps) in fp64 :20-23.
 *   min_area_host [n_cls][n_area] int32 (0 disables), n_area <= MLC_MAX_AREAS
 *   dice_sum_dev  [n_cls][n_thr][n_area] fp64: sum over images (overwritten;
 *                 summed in a fixed order: bit-reproducible)
 *   micro_dev     [n_cls][1+2*n_thr] int64 or NULL: {T, P[0..], I[0..]}
 *                 summed over images (overwritten)
 *   work_dev      scratch of at least mlc_dice_search_workspace(...) bytes
 */
size_t mlc_dice_search_workspace(int64_t n_img, int n_cls, int n_thr,
                                 int n_area);
int mlc_dice_search(const uint32_t *hist_dev, int64_t n_img, int n_cls,
                    int n_thr, const int32_t *min_area_host, int n_area,
                    int empty_one, double eps, double *dice_sum_dev,
                    int64_t *micro_dev, void *work_dev, size_t work_bytes,
                    void *stream);

/* mlc_dice_search that also sums the intersection-over-union of the same
 * masks: iou = I / (T + P - I + eps) per (plane, threshold, min-area), with
 * the min-area rule (an emptied prediction has P = I = 0) and the empty rule
 * (T = P = 0 -> 1 when empty_one) of dice.py:13-18 applied alike.  The
 * reference has no IoU (contrib/metrics/dice.py:4-23 is dice only): this
 * output is builder-defined and pinned only against this repo's oracle.
 *   iou_sum_dev [n_cls][n_thr][n_area] fp64, overwritten
 *   work_dev    at least 2 x mlc_dice_search_workspace(...) bytes */
int mlc_dice_iou_search(const uint32_t *hist_dev, int64_t n_img, int n_cls,
                        int n_thr, const int32_t *min_area_host, int n_area,
                        int empty_one, double eps, double *dice_sum_dev,
                        double *iou_sum_dev, int64_t *micro_dev,
                        void *work_dev, size_t work_bytes, void *stream);

/* mlc_dice_search without a workspace argument, for callers that cannot
 * allocate device memory: a scoring loop sharing the GPU with an inference
 * engine that reserved the whole card at start-up, or a thin ctypes binding
 * from the reference's worker that is handed its input/output buffers and has
 * no device allocator of its own.  (No counterpart in the reference: it is an
 * addition of this library, same arithmetic and same results as
 * mlc_dice_search.)
 *  - The scratch is a 4 MiB array built into the library image; every GPU
 *    that loads the library holds it until the process exits.
 *  - Limit: mlc_dice_search_workspace(n_img, n_cls, n_thr, n_area) must be
 *    <= 4 MiB (4 194 304 bytes); for a large image count that means about
 *    n_thr * n_area <= 880 (the C2 grid, 100 thresholds x 5 areas over 4
 *    classes, needs 2.4 MB).  Larger searches return MLC_ERR_CAPACITY and
 *    must use mlc_dice_search.
 *  - The first call on each device looks up the array's device address
 *    (cudaGetSymbolAddress) before enqueueing work; make that first call
 *    outside any stream capture.  No call allocates or frees anything.
 *  - All calls on a device share the one array, and it keeps the previous
 *    call's partial sums (each call overwrites what it reads, nothing needs
 *    clearing).  Two calls whose kernels can overlap on the GPU (different
 *    streams, different host threads) race on it: serialise them. */
int mlc_dice_search_static(const uint32_t *hist_dev, int64_t n_img, int n_cls,
                           int n_thr, const int32_t *min_area_host, int n_area,
                           int empty_one, double eps, double *dice_sum_dev,
                           int64_t *micro_dev, void *stream);

/* Fold the all-gathered partial results of the ranks of a sharded search
 * (images split over GPUs).  No reference counterpart: the reference scores
 * in one process (worker/tasks.py:106-131 pins one task to one GPU); this is
 * the only step in which ranks exchange data, and it moves KBs.
 *   gathered_dev [world][n_f64 + n_i64] int64: row r = rank r's buffer - the
 *               first n_f64 entries are fp64 bit patterns (dice_sum), the
 *               rest plain integers (micro counts, image count)
 *   out_dev     [n_f64 + n_i64] int64: fp64 entries summed in rank order
 *               (bit-identical on every rank), integers summed exactly */
int mlc_combine_ranks(const int64_t *gathered_dev, int world, int64_t n_f64,
                      int64_t n_i64, int64_t *out_dev, void *stream);

/* Per-plane dice at one chosen (threshold index, min_area) per class:
 * dice_dev [n_img*n_cls] fp64 (the per-image `scores` the report builders
 * take, worker/reports/segmenation.py:148-150). */
int mlc_plane_dice(const uint32_t *hist_dev, int64_t n_img, int n_cls,
                   int n_thr, const int32_t *k_host, const int32_t *min_area_host,
                   int empty_one, double eps, double *dice_dev, void *stream);

/* Mask area of every plane at one threshold index per class: area[p] =
 * P(k_c) = sum_{j > k_c} (hist[p][0][j] + hist[p][1][j]) - `outputs.sum()` of
 * contrib/metrics/dice.py:11 for thr[k_c], before the min-area rule.
 *   area_dev [n_img*n_cls] int32 */
int mlc_plane_area(const uint32_t *hist_dev, int64_t n_img, int n_cls,
                   int n_thr, const int32_t *k_host, int32_t *area_dev,
                   void *stream);

/* ---- run-length encoding ------------------------------------------------
 * Replaces (prob >= t) + min-area reject + mask2rle
 * (contrib/transform/rle.py:4-13): column-major, 1-based (start, length).
 *   thr_host, min_area_host  [n_cls]
 *   runs_dev   int32[capacity]: (start, length) pairs, plane by plane in
 *              arbitrary order; plane p owns runs_dev[off[p] .. off[p]+2*cnt[p])
 *   off_dev    [planes] int64, cnt_dev [planes] int32 (#runs),
 *   area_dev   [planes] int32 mask area BEFORE the min-area rule (or NULL)
 *   total_dev  [1] int64: int32 elements needed; if > capacity the planes
 *              that did not fit have off = -1 and the caller retries bigger.
 *   src may be uint8 masks already: dtype MLC_U8 with act NONE and thr 0.5
 *   Planes of any size: up to ~1.6 M pixels (a column-major bit plane of at
 *   most 200 KB) are transposed in shared memory and read once; larger ones
 *   take a column-walking kernel that reads the plane twice (W <= 25 600;
 *   beyond that MLC_ERR_UNSUPPORTED).
 */
#define MLC_U8 2
int mlc_rle_encode(const void *const *src_dev_ptrs, const int32_t *xform,
                   int n_src, int dtype, int64_t n_img, int n_cls, int H, int W,
                   const float *thr_host, const int32_t *min_area_host, int act,
                   int32_t *runs_dev, int64_t capacity, int64_t *off_dev,
                   int32_t *cnt_dev, int32_t *area_dev, int64_t *total_dev,
                   void *stream);

/* mlc_rle_encode for masks whose AREA is already known - the usual case after
 * a threshold search: with thr = thr_grid[k], the area of plane p's mask is
 * P(k) of its histogram (mlc_plane_area).  Same reference lines as
 * mlc_rle_encode; the min-area rule (contrib/metrics/dice.py:13-15) and the
 * empty string of an empty mask (contrib/transform/rle.py:4-13) are decided
 * from the known area, so planes that end up without runs are not read at
 * all (most planes of a Severstal-like set).
 *   known_area_dev [planes] int32: the exact pixel count of (prob >= thr_c)
 *              for every plane; a wrong value gives a wrong result
 * Everything else as for mlc_rle_encode (skipped planes get cnt = 0, off = 0,
 * area = the known area). */
int mlc_rle_encode_known(const void *const *src_dev_ptrs, const int32_t *xform,
                         int n_src, int dtype, int64_t n_img, int n_cls, int H,
                         int W, const float *thr_host,
                         const int32_t *min_area_host, int act,
                         const int32_t *known_area_dev, int32_t *runs_dev,
                         int64_t capacity, int64_t *off_dev, int32_t *cnt_dev,
                         int32_t *area_dev, int64_t *total_dev, void *stream);

/* Text of the run-length strings, replacing the ' '.join(str(x) for x in
 * runs) of contrib/transform/rle.py:13 (and what
 * examples/digit-recognizer/executors/infer.py:76-78 then writes to the csv
 * the Submit executor reads, worker/executors/kaggle.py:100-110).
 *   runs_dev/off_dev/cnt_dev  the outputs of mlc_rle_encode
 *   chars_dev   char[capacity]: ASCII decimal values separated by single
 *               blanks, no terminator; plane p owns
 *               chars_dev[str_off[p] .. str_off[p]+str_len[p]) ('' -> len 0)
 *   total_dev   [1] int64 characters needed; planes that did not fit have
 *               str_off = -1 and the caller retries with a larger buffer */
int mlc_rle_format(const int32_t *runs_dev, const int64_t *off_dev,
                   const int32_t *cnt_dev, int64_t n_planes, char *chars_dev,
                   int64_t capacity, int64_t *str_off_dev, int32_t *str_len_dev,
                   int64_t *total_dev, void *stream);

/* Inverse of mlc_rle_encode, replacing rle2mask
 * (contrib/transform/rle.py:16-31): column-major 1-based (start, length)
 * pairs -> masks.  Used to turn the competition's RLE ground truth into the
 * uint8 planes mlc_seg_hist reads.
 *   runs_dev/off_dev/cnt_dev  as produced by mlc_rle_encode (or uploaded by
 *               the caller after parsing the strings); off < 0 = empty plane;
 *               runs may overlap; a run is clipped to the plane on both
 *               sides (start < 1, past the end, negative length: no write)
 *   mask_dev    [n_planes][H][W] uint8 {0,1}, overwritten */
int mlc_rle_decode(const int32_t *runs_dev, const int64_t *off_dev,
                   const int32_t *cnt_dev, int64_t n_planes, int H, int W,
                   uint8_t *mask_dev, void *stream);

/* Colour composite of predicted masks, replacing
 * SegmentationReportBuilder.encode_pred (worker/reports/segmenation.py:71-80):
 *   out[i][y][x][ch] = sum over classes of uint8(mask[i][c][y][x] * color[c][ch])
 * with numpy's uint8 wrap-around in the product and in the running sum.
 *   mask_dev    [n_img][n_cls][H][W] uint8 (the reference passes {0,1})
 *   colors_host [n_cls][3] uint8, or NULL = (255, 255, 255) for every class,
 *               the reference's default (:75-77)
 *   out_dev     [n_img][H][W][3] uint8, overwritten */
int mlc_mask_composite(const uint8_t *mask_dev, int64_t n_img, int n_cls, int H,
                       int W, const uint8_t *colors_host, uint8_t *out_dev,
                       void *stream);

/* Class outlines over images, replacing SegmentationReportBuilder.plot_mask
 * (worker/reports/segmenation.py:82-96: cv2.findContours(RETR_LIST,
 * CHAIN_APPROX_NONE) per class, then cv2.polylines(img, contour, True, color,
 * 2) on the POINT array of each contour).  Pixel-exact against OpenCV 4.13:
 * a pixel takes the colour of the last class that has a contour point (a
 * mask pixel with a zero 4-neighbour, frame = zero) on it or on one of its
 * four 4-neighbours.
 *   img_dev     [n_img][H][W][3] uint8, updated in place (grey images are
 *               replicated to 3 channels by the caller, :83-84)
 *   mask_dev    [n_img][n_cls][H][W] uint8, non-zero = set
 *   colors_host [n_cls][3] uint8 or NULL = (0, 255, 0), the default (:92) */
int mlc_mask_outline(uint8_t *img_dev, const uint8_t *mask_dev, int64_t n_img,
                     int n_cls, int H, int W, const uint8_t *colors_host,
                     void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MLC_H_ */
