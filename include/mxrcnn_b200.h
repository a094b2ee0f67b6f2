/*
 * mxrcnn_b200.h -- C ABI of libmxrcnn_b200.so
 *
 * B200 (sm_100a) implementation of the RPN proposal / box-processing hot path of
 * walkoncross/mx-rcnn.  The reference has no native code and no FFI on this path (it is
 * numpy inside MXNet CustomOp callbacks), so each entry point cites the reference
 * PYTHON function (file:line under the reference root) whose work it replaces; the
 * Python host layer in mxrcnn_b200/ binds these with ctypes and mirrors the reference's
 * own signatures.  INTEGRATION.md shows the reference-side stubs.
 *
 * Conventions
 *   - every function returns 0 on success or a negative mxr_status; the text of the
 *     last error of the calling thread is returned by mxr_last_error_string().
 *   - "dev" pointers are device pointers on the handle's device; "host" pointers are
 *     host pointers.  The caller owns every in/out buffer.  The library owns only the
 *     handle's workspace (device scratch + pinned staging), which grows on demand in
 *     mxr_reserve_*() -- call those at set-up to keep allocations out of the hot call.
 *   - `stream` is a cudaStream_t passed as void* (NULL = CUDA's legacy default stream, as everywhere in CUDA).
 *     Device-pointer calls only enqueue work and never synchronise.  *_host calls
 *     copy in, compute, copy out and return after the results are in the host buffers.
 *   - a handle is bound to one device and is not thread-safe; use one per thread/stream.
 *   - no random numbers are drawn inside the library: sampling entry points take the
 *     permutations / draws from the caller (the reference uses numpy's global RNG).
 *   - layouts and dtypes are the reference's: NCHW f32 maps,
 *     boxes [x1,y1,x2,y2] with inclusive pixel coordinates (+1 widths), f64 box math.
 */
#ifndef MXRCNN_B200_H_
#define MXRCNN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mxr_handle_s* mxr_handle_t;

typedef enum {
  MXR_OK = 0,
  MXR_ERR_INVALID = -1,   /* bad argument                          -> ValueError   */
  MXR_ERR_CUDA = -2,      /* CUDA runtime failure                  -> RuntimeError */
  MXR_ERR_NAN = -3,       /* NaN in scores/deltas (proposal.py:58-62) -> ValueError */
  MXR_ERR_LIMIT = -4,     /* size beyond a documented maximum      -> ValueError   */
  MXR_ERR_NOMEM = -5,     /* allocation failed                     -> MemoryError  */
  MXR_ERR_EMPTY = -6      /* empty candidate set where the reference raises (Q7) */
} mxr_status;

#define MXR_DTYPE_F32 0
#define MXR_DTYPE_F64 1

const char* mxr_last_error_string(void);
int mxr_version(void);

/* ---- handle ------------------------------------------------------------------------ */
int mxr_create(int device, mxr_handle_t* out);
int mxr_destroy(mxr_handle_t h);
/* Bytes of device workspace currently held by the handle. */
size_t mxr_workspace_bytes(mxr_handle_t h);
/* Wait for everything enqueued on `stream`. */
int mxr_synchronize(mxr_handle_t h, void* stream);

/* Optional per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg;
 * the reference's only instrumentation is the Speedometer log, rcnn/callback.py:14-36).
 * enable(1) starts recording every kernel launch of this handle; read() synchronises the device,
 * sums milliseconds and launch counts per stage id (< max_stages) and clears the records. */
int mxr_profile_enable(mxr_handle_t h, int on);
int mxr_profile_read(mxr_handle_t h, int max_stages, double* total_ms, int32_t* counts);
const char* mxr_profile_stage_name(int stage);
/* Kernels launched by the most recent call on this handle. */
int mxr_last_launch_count(mxr_handle_t h);
/* Phase durations (microseconds, %globaltimer stamps taken by the leader CTA) of the most recent fused
 * proposal launch: us_out[image*9 + i], i = 0 start-up barrier, 1 decode, 2 select, 3 compact, 4 sort,
 * 5 re-decode, 6 NMS stripes, 7 outputs; [8] = number of 256-box stripes walked.  Synchronises. */
int mxr_proposal_phase_times(mxr_handle_t h, int max_images, double* us_out, int32_t* images_out);
/* The raw stamps of the same launch: ns_out[image*48 + i].  i = 0..8 the phase boundaries above; fused kernel of
 * proposal_fused.cuh additionally: 10 first stripe staged, 11 its IoU bit blocks done, 12 barrier passed, 13 stripe
 * resolved by the leader; [9] stripes walked, [14] windows materialised, [15] fixed-point rounds of the first stripe;
 * 16..47: SM cycle counter (clock64) of the leader's first thread at fine-grained points of the first stripe
 * (tools/phase_probe2.py names them). */
int mxr_proposal_stamps(mxr_handle_t h, int max_images, uint64_t* ns_out, int32_t* images_out);
/* Instrumented (un-timed) runs for bench.py's ALU roofline: count_pairs(1) makes the following fused proposal calls
 * count the IoU pairs they evaluate (nms.py:25-36 evaluations, SURVEY 8d "pairs_evaluated"); pairs_read() returns the
 * per-image counters of the most recent such call: pairs_out[2*image] pairs evaluated, [2*image+1] pairs that the
 * float32 screen could not decide and the float64 arithmetic settled.  Synchronises. */
int mxr_proposal_count_pairs(mxr_handle_t h, int on);
int mxr_proposal_pairs_read(mxr_handle_t h, int max_images, uint32_t* pairs_out, int32_t* images_out);

/* Measured non-tensor ALU throughput of this GPU (Gop/s; streams of independent multiplies and adds,
 * no FMA contraction) -- the denominator of the NMS kernels' ALU roofline in bench.py. */
int mxr_bench_alu_peak(mxr_handle_t h, int dtype, double* gops_out);

/* ---- A0  helper/processing/generate_anchor.py:8-72 ---------------------------------- */
/* Host-side constant table (<= 64 rows); out_anchors: host, (n_ratios*n_scales, 4) f64,
 * ratio-major / scale-minor row order. */
int mxr_generate_anchors(double base_size, const double* ratios, int n_ratios, const double* scales, int n_scales,
                         double* out_anchors);

/* ---- A1  rcnn/rpn/proposal.py:39-149  ProposalOperator.forward ---------------------- */
typedef struct {
  int batch;            /* B images (the reference supports 1: proposal.py:195-196)          */
  int num_anchors;      /* A                                                                 */
  int height, width;    /* H, W of the score / delta maps                                    */
  int crop_h, crop_w;   /* proposal.py:70-71 TRAIN-mode crop; <=0 or >=H/W means no crop     */
  double feat_stride;   /* proposal.py:22                                                    */
  int pre_nms_top_n;    /* <=0: keep all (proposal.py:124)                                   */
  int post_nms_top_n;   /* <=0: keep all NMS survivors (proposal.py:132); output rows = this */
  double nms_thresh;    /* python float                                                      */
  float min_size;       /* RPN_MIN_SIZE; multiplied by im_info[2] in f32 (NEP 50)            */
  int first_stripe;     /* tuning: width of the first NMS stripe (0 = automatic)             */
} mxr_proposal_params;

/* Size the workspace for a configuration (optional; forward calls reserve on demand). */
int mxr_reserve_proposal(mxr_handle_t h, const mxr_proposal_params* p);

/*
 * cls_prob  dev (B, 2A, H, W) f32   only the foreground half [A:2A) is read
 * bbox_pred dev (B, 4A, H, W) f32
 * im_info   dev (B, 3) f32          [height, width, scale]
 * anchors   HOST (A, 4) f64         from mxr_generate_anchors
 * out_rois  dev (B, post, 5) f32    [batch_index, x1, y1, x2, y2]; rows >= nkeep are zero
 * out_scores dev (B, post) f32 or NULL
 * out_keep  dev (B, post) i32 or NULL   index into the sorted pre-NMS list (the reference's `keep`)
 * out_anchor_index dev (B, post) i32 or NULL   (h,w,a)-order row of the kept proposal
 * out_nkeep dev (B, 2) i32          [nkeep, flags]; flags bit0 = NaN seen in inputs
 * When post_nms_top_n <= 0 the row capacity of the outputs is `out_capacity` per image.
 * Rows beyond nkeep are padded by the caller (proposal.py:135-139 draws from numpy's RNG):
 * see mxr_proposal_pad.
 */
int mxr_proposal_forward(mxr_handle_t h, void* stream, const mxr_proposal_params* p, const float* cls_prob,
                         const float* bbox_pred, const float* im_info, const double* anchors, int out_capacity,
                         float* out_rois, float* out_scores, int32_t* out_keep, int32_t* out_anchor_index,
                         int32_t* out_nkeep);

/* Same, all pointers HOST; returns MXR_ERR_NAN like the reference's ValueError. */
int mxr_proposal_forward_host(mxr_handle_t h, const mxr_proposal_params* p, const float* cls_prob,
                              const float* bbox_pred, const float* im_info, const double* anchors, int out_capacity,
                              float* out_rois, float* out_scores, int32_t* out_keep, int32_t* out_anchor_index,
                              int32_t* out_nkeep);

/* Pipelined form of the host entry point (mxr_proposal_forward_host == submit + wait on slot 0).
 * submit enqueues, on the private stream of `slot` (0 or 1), the H2D copies of the inputs (pinned host
 * memory recommended), the kernels and the D2H copy of the results into the slot's pinned block, and
 * returns at once; wait blocks until that slot is finished and copies the results into the caller's
 * arrays.  Alternating the two slots overlaps the input copy of batch i+1 with the kernels of batch i,
 * which is how the reference's MXNet engine would feed the op (asnumpy/assign around the callback,
 * rcnn/rpn/proposal.py:57-63,147-149).  The input arrays must stay valid until wait returns.  Call
 * mxr_reserve_proposal first: the workspace cannot grow while the other slot is in flight. */
int mxr_proposal_submit_host(mxr_handle_t h, int slot, const mxr_proposal_params* p, const float* cls_prob,
                             const float* bbox_pred, const float* im_info, const double* anchors, int out_capacity);
/* The same with the inputs packed by the caller into ONE host block, so that they cross the bus in one copy:
 * [foreground scores (B, A, H, W)] [bbox deltas (B, 4A, H, W)] [im_info (B, 3)], float32, contiguous -- the arrays
 * of proposal.py:57-63 (cls_prob[:, A:], bbox_pred, im_info) laid end to end.  Results are fetched with
 * mxr_proposal_wait_host as usual. */
int mxr_proposal_submit_host_packed(mxr_handle_t h, int slot, const mxr_proposal_params* p, const float* packed,
                                    const double* anchors, int out_capacity);
int mxr_proposal_wait_host(mxr_handle_t h, int slot, float* out_rois, float* out_scores, int32_t* out_keep,
                           int32_t* out_anchor_index, int32_t* out_nkeep);

/* proposal.py:135-139: rows [nkeep, nkeep+n_pad) of image `image` := row pad_rows[t] (< nkeep).
 * pad_rows: HOST, drawn by the caller as randint(0, nkeep, n_pad). */
int mxr_proposal_pad(mxr_handle_t h, void* stream, int image, int rows_per_image, int nkeep, const int32_t* pad_rows,
                     int n_pad, float* out_rois, float* out_scores);

/* Stage-level views of the proposal op for parity tests (mxr_decode_clip; rcnn/rpn/proposal.py:110-120).
 * decode_clip: bbox_pred + clip_boxes + the min-size decision for EVERY anchor of the (cropped) map, device pointers,
 * rows in the reference's order (h, w, a): out_boxes (B, A*hc*wc, 4) float64 -- the values the op's outputs are cast
 * from -- and out_valid (B, A*hc*wc) int32 -- 1 where _filter_boxes (proposal.py:156-161) keeps the row, decided by
 * the same screened test the fused kernel uses.
 * counts: per image of the most recent proposal call [n_valid (= len(keep) of proposal.py:117), boxes entering NMS
 * (proposal.py:123-127), boxes kept, NaN flag].  Synchronises. */
int mxr_decode_clip(mxr_handle_t h, void* stream, const mxr_proposal_params* p, const float* bbox_pred, const float* im_info,
                    const double* anchors, double* out_boxes, int32_t* out_valid);
/* SURVEY 8b mxr_topk_sorted; rcnn/rpn/proposal.py:113-127: the complete pre-NMS candidate list of every image in
 * processing order (min-size filter, descending score, pre_nms_top_n cut, ties as documented), as anchor rows
 * (h*W + w)*A + a: out_order (B, out_capacity) int32 device, out_capacity >= min(pre_nms_top_n, A*hc*wc);
 * out_count (B, 2) int32 device: list length, NaN flag.  nms_thresh / post_nms_top_n of `p` are ignored.  A stage for
 * parity tests (it runs the op with a threshold nothing exceeds, on the float64 pair test: milliseconds). */
int mxr_topk_sorted(mxr_handle_t h, void* stream, const mxr_proposal_params* p, const float* cls_prob, const float* bbox_pred,
                    const float* im_info, const double* anchors, int out_capacity, int32_t* out_order, int32_t* out_count);
int mxr_proposal_counts(mxr_handle_t h, int max_images, int32_t* counts_out, int32_t* images_out);

/* ---- A5  helper/processing/nms.py:4-38  nms(dets, thresh) ---------------------------- */
/* dets dev (n,5) [x1,y1,x2,y2,score] f32 or f64 (arithmetic in that dtype, like numpy);
 * sorts by descending score (ties: higher index first), greedy-suppresses IoU > thresh.
 * keep_out dev (n) i32: original row indices in keep order; nkeep_out dev (2) i32.
 * max_keep > 0 stops after that many (prefix of the full result). */
int mxr_nms(mxr_handle_t h, void* stream, const void* dets, int n, int dtype, double thresh, int max_keep,
            int32_t* keep_out, int32_t* nkeep_out);
int mxr_nms_host(mxr_handle_t h, const void* dets, int n, int dtype, double thresh, int max_keep, int32_t* keep_out,
                 int32_t* nkeep_out);
/* Batched form for the per-class loop of rcnn/tester.py:50-56: problem c owns rows
 * [offsets[c], offsets[c+1]) of dets; keep_out rows are problem-local indices written at
 * keep_out + offsets[c]; nkeep_out (n_problems, 2).  offsets: HOST (n_problems+1). */
int mxr_nms_batched(mxr_handle_t h, void* stream, const void* dets, const int32_t* offsets, int n_problems, int dtype,
                    double thresh, int max_keep, int32_t* keep_out, int32_t* nkeep_out);
/* Same with HOST dets / keep_out / nkeep_out: one upload, one launch, one download. */
int mxr_nms_batched_host(mxr_handle_t h, const void* dets, const int32_t* offsets, int n_problems, int dtype, double thresh,
                         int max_keep, int32_t* keep_out, int32_t* nkeep_out);

/* ---- helper/dataset/imdb.py:153-170  the one-to-one matching loop of evaluate_recall ---- */
/* boxes (n,4) f64 proposals, gt_boxes (k,4) f64, both HOST; covered_out (k) f64: round j's IoU -- the gt box with the best
 * remaining cover (first maximum), matched to the proposal that covers it best (first maximum), both retired.  The IoU is
 * bbox_overlaps' (helper/processing/bbox_regression.py:11-31) in float64; the (n,k) matrix is never materialised.
 * A round with nothing left to match records -1 (the reference asserts on it). */
int mxr_recall_cover_host(mxr_handle_t h, const double* boxes, int n, const double* gt_boxes, int k, double* covered_out);

/* ---- helper/processing/nms.py:40-70  nest(dets, thresh) -------------- */
/* keep_flag[i] = 0 when some other box j has inter(i,j)/area(i) > thresh, else 1; dets (n,5) f32|f64,
 * arithmetic in that dtype.  The kept list of the reference is the ascending indices with flag 1. */
int mxr_nest(mxr_handle_t h, void* stream, const void* dets, int n, int dtype, double thresh, int32_t* keep_flag);
int mxr_nest_host(mxr_handle_t h, const void* dets, int n, int dtype, double thresh, int32_t* keep_flag);

/* ---- test-time box head ------------------------------------------------
 * rcnn/detector.py:73-74 (pred_boxes = clip_boxes(bbox_pred(rois, bbox_deltas), im_shape)) followed by the
 * class loop of rcnn/tester.py:50-65: for every foreground class j, rows with scores[:, j] > score_thresh
 * (compared in float32, as numpy does for a float32 array and a python float), boxes / scale, float64 greedy
 * NMS (nms.py:4-38) at nms_thresh, and finally the cap: when more than max_per_image detections survive over
 * all classes, only those with score >= the max_per_image-th largest score stay (ties stay, tester.py:60-65).
 * One launch for the whole batch. */
#define MXR_BOX_HEAD_MAX_ROIS 4096
typedef struct {
  int32_t n_images;      /* B; the reference handles one image per call */
  int32_t n_rois;        /* R rows per image, <= MXR_BOX_HEAD_MAX_ROIS */
  int32_t n_classes;     /* C, background (class 0) included */
  int32_t max_per_image; /* tester.py:24: 100; <= 0: no cap */
  double score_thresh;   /* tester.py:22: 0.05 */
  double nms_thresh;     /* config.TEST.NMS */
} mxr_box_head_params;
/* rois dev (B,R,4) f32 (batch column dropped, detector.py:68); bbox_deltas dev (B,R,4C) f32; scores dev (B,R,C) f32;
 * im dev (B,3) f64: height and width of the network input (im_array.shape[-2:]) and the scale the boxes are
 * divided by (im_info[0,2], tester.py:39,53).
 * dets_out dev (B,C,R,5) f64: rows [0, count) of class j are all_boxes[j][i] = [x1,y1,x2,y2,score] in
 * descending score; rows_out dev (B,C,R) i32: the roi row of each detection; count_out dev (B,C,2) i32:
 * [detections after the cap, detections after NMS].  Class 0 always has count 0. */
int mxr_box_head(mxr_handle_t h, void* stream, const mxr_box_head_params* p, const float* rois, const float* bbox_deltas,
                 const float* scores, const double* im, double* dets_out, int32_t* rows_out, int32_t* count_out);
/* Same with HOST arrays (copies staged through the workspace); only rows [0, count) of each class list are
 * guaranteed to be written. */
int mxr_box_head_host(mxr_handle_t h, const mxr_box_head_params* p, const float* rois, const float* bbox_deltas,
                      const float* scores, const double* im, double* dets_out, int32_t* rows_out, int32_t* count_out);

/* ---- A2  helper/processing/bbox_transform.py:35-78  bbox_pred(boxes, box_deltas) -------- */
/* boxes (n,4) f32|f64 (cast to f64 like :46), deltas (n,c4) f32|f64 (exp evaluated in that dtype),
 * out (n,c4) f64.  n == 0 is a no-op (:43-44).  *_host: all pointers host. */
int mxr_bbox_pred(mxr_handle_t h, void* stream, const void* boxes, int boxes_dtype, const void* deltas,
                  int deltas_dtype, int n, int c4, double* out);
int mxr_bbox_pred_host(mxr_handle_t h, const void* boxes, int boxes_dtype, const void* deltas, int deltas_dtype, int n,
                       int c4, double* out);

/* ---- A3  helper/processing/bbox_transform.py:81-96  clip_boxes(boxes, im_shape), in place -- */
/* lim_x = im_shape[1]-1, lim_y = im_shape[0]-1, evaluated by the caller with numpy's scalar rules. */
int mxr_clip_boxes(mxr_handle_t h, void* stream, void* boxes, int dtype, int n, int c4, double lim_x, double lim_y);
int mxr_clip_boxes_host(mxr_handle_t h, void* boxes, int dtype, int n, int c4, double lim_x, double lim_y);

/* ---- A10 helper/processing/bbox_transform.py:8-32  bbox_transform(ex_rois, gt_rois) -------- */
/* ex (n,4) and gt (n,4) each f32|f64; numpy computes each side in its own dtype and the targets in
 * the promoted one, so out (n,4) is f32 only when both inputs are f32, else f64. */
int mxr_bbox_transform(mxr_handle_t h, void* stream, const void* ex, int ex_dtype, const void* gt, int gt_dtype, int n,
                       void* out);
int mxr_bbox_transform_host(mxr_handle_t h, const void* ex, int ex_dtype, const void* gt, int gt_dtype, int n, void* out);

/* ---- A7  helper/processing/bbox_regression.py:11-31  bbox_overlaps(boxes, query_boxes) ----- */
/* f64 IoU with +1 widths; strides are in elements (>= 4; only columns 0..3 are read).  Fused with
 * the reductions its callers take (rcnn/minibatch.py:296-300, rcnn/rpn/proposal_target.py:143-144):
 * out (n,k) | row_argmax (n) first maximum | row_max (n) | col_max (k); any of them may be NULL. */
int mxr_bbox_overlaps(mxr_handle_t h, void* stream, const double* boxes, int n, int box_stride, const double* query,
                      int k, int query_stride, double* out, int32_t* row_argmax, double* row_max, double* col_max);
int mxr_bbox_overlaps_host(mxr_handle_t h, const double* boxes, int n, int box_stride, const double* query, int k,
                           int query_stride, double* out, int32_t* row_argmax, double* row_max, double* col_max);

/* ---- A9  helper/processing/bbox_regression.py:67-85  expand_bbox_regression_targets -------- */
/* data (k,5) f32 [cls,dx,dy,dw,dh] -> targets, weights (k,4*num_classes) f32.
 * inside_weights4: HOST, config.TRAIN.BBOX_INSIDE_WEIGHTS. */
int mxr_expand_bbox_targets(mxr_handle_t h, void* stream, const float* data, int k, int num_classes,
                            const float* inside_weights4, float* targets, float* weights);
int mxr_expand_bbox_targets_host(mxr_handle_t h, const float* data, int k, int num_classes, const float* inside_weights4,
                                 float* targets, float* weights);

/* ---- A6  rcnn/minibatch.py:204-395  assign_anchor ---------------------------------------- */
/* Two phases around the reference's RNG draws (minibatch.py:319-334).
 * phase1 (synchronous): inside filter, IoU vs gt, argmax / gt_argmax labels (quirk Q4), ordered fg / bg
 *   lists.  anchors: HOST (A,4) f64 from mxr_generate_anchors(16, ratios, float32 scales);
 *   gt_boxes: HOST (G,5) f32; counts_out: HOST [n_inside, n_fg, n_bg] (labels==1 / ==0 before subsampling).
 * The caller draws, in this order and only when over quota, the fg then bg ranks to disable:
 *   npr.choice(fg_inds, size=k, replace=False) == fg_inds[npr.permutation(n_fg)[:k]]  -> pass the RANKS.
 * phase2: applies the draws and writes, in the reference's output layouts,
 *   label (A*H*W) f32 in (a,h,w) order; bbox_target / inside / outside weights (4A,H,W) f32, channel a*4+k.
 *   positive_weight / negative_weight: the values of minibatch.py:343-354 computed by the caller.
 * State lives in the handle between the two calls (no other anchor-target call in between). */
typedef struct {
  int feat_height, feat_width, num_anchors;
  int feat_stride;      /* python int, minibatch.py:260 */
  int allowed_border;
  float im_height, im_width; /* im_info[0][0], im_info[0][1] */
  int num_gt;
  double negative_overlap, positive_overlap; /* RPN_NEGATIVE_OVERLAP, RPN_POSITIVE_OVERLAP */
  int clobber_positives;
} mxr_anchor_target_params;
int mxr_anchor_target_phase1(mxr_handle_t h, const mxr_anchor_target_params* p, const double* anchors,
                             const float* gt_boxes, int32_t* counts_out);
int mxr_anchor_target_phase2(mxr_handle_t h, void* stream, const int32_t* fg_disable_ranks, int n_fg_disable,
                             const int32_t* bg_disable_ranks, int n_bg_disable, double positive_weight,
                             double negative_weight, const float* inside_weights4, float* label, float* bbox_target,
                             float* bbox_inside_weight, float* bbox_outside_weight);
/* The same two phases for a BATCH of images that share the feature map (the loader pads the data first,
 * rcnn/loader.py:283-286): one launch and one synchronisation per phase instead of one per image.
 *   p: feat_*, num_anchors, feat_stride, allowed_border, overlaps thresholds, clobber (im_height / im_width / num_gt
 *      are per image and ignored here);
 *   im_heights_widths (n_images, 2): im_info[0][0], im_info[0][1] of every image;
 *   gt_boxes (sum G, 5) f32, all images' boxes one after the other; gt_offsets (n_images + 1): image b owns rows
 *      gt_offsets[b] .. gt_offsets[b+1];
 *   counts_out (n_images, 3): n_inside, n_fg, n_bg.
 * phase2: disable_ranks: image by image, the fg ranks then the bg ranks; disable_counts (n_images, 2);
 *   weights (n_images, 2): positive, negative weight; outputs are the batch tensors label (n_images, A*H*W) and
 *   bbox_target / inside / outside weights (n_images, 4A, H, W), device pointers, written on `stream`. */
int mxr_anchor_target_batch_phase1(mxr_handle_t h, const mxr_anchor_target_params* p, const double* anchors, int n_images,
                                   const float* im_heights_widths, const float* gt_boxes, const int32_t* gt_offsets,
                                   int32_t* counts_out);
int mxr_anchor_target_batch_phase2(mxr_handle_t h, void* stream, const int32_t* disable_ranks, const int32_t* disable_counts,
                                   const double* weights, const float* inside_weights4, float* label, float* bbox_target,
                                   float* bbox_inside_weight, float* bbox_outside_weight);
int mxr_anchor_target_phase2_host(mxr_handle_t h, const int32_t* fg_disable_ranks, int n_fg_disable,
                                  const int32_t* bg_disable_ranks, int n_bg_disable, double positive_weight,
                                  double negative_weight, const float* inside_weights4, float* label,
                                  float* bbox_target, float* bbox_inside_weight, float* bbox_outside_weight);

/* ---- A8/A11  rcnn/rpn/proposal_target.py:148,160-161 and rcnn/minibatch.py:166,174 ---------- */
/* fg_list = where(ov >= fg_thresh), bg_list = where((ov < bg_hi) & (ov >= bg_lo)), ascending, compared
 * in the dtype of `overlaps` (f64 from bbox_overlaps, f32 from a roidb).  counts: [n_fg, n_bg]. */
int mxr_fg_bg_lists(mxr_handle_t h, void* stream, const void* overlaps, int dtype, int n, double fg_thresh, double bg_hi,
                    double bg_lo, int32_t* fg_list, int32_t* bg_list, int32_t* counts);
int mxr_fg_bg_lists_host(mxr_handle_t h, const void* overlaps, int dtype, int n, double fg_thresh, double bg_hi,
                         double bg_lo, int32_t* fg_list, int32_t* bg_list, int32_t* counts);

/* ---- A8  rcnn/rpn/proposal_target.py:183-194,120-133 -------------------------------------- */
/* Given the sampled roi indices (drawn by the caller, proposal_target.py:153-180): gather rois, labels
 * (zeroed from fg_this on, quirk Q3), float32 bbox_transform targets, optional (t-mean)/std in f64,
 * class-expanded targets and inside / outside weights.  rois (R,5) f32, gt_boxes (G,5) f32,
 * gt_assignment (R) i32, keep_inds (K) i32 -> out_rois (K,5), labels (K), three (K,4C) maps. */
typedef struct {
  int num_rois, num_gt, num_keep, fg_this, num_classes;
  int normalize;        /* config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED */
  double means[4], stds[4];
  float inside_weights[4];
} mxr_proposal_target_params;
int mxr_proposal_target_gather(mxr_handle_t h, void* stream, const mxr_proposal_target_params* p, const float* rois,
                               const float* gt_boxes, const int32_t* gt_assignment, const int32_t* keep_inds,
                               float* out_rois, float* labels, float* bbox_targets, float* inside, float* outside);
int mxr_proposal_target_gather_host(mxr_handle_t h, const mxr_proposal_target_params* p, const float* rois,
                                    const float* gt_boxes, const int32_t* gt_assignment, const int32_t* keep_inds,
                                    float* out_rois, float* labels, float* bbox_targets, float* inside, float* outside);

/* _sample_rois (proposal_target.py:135-196) in two calls with host arrays.  phase1: ONE launch computes the float64
 * IoU of every roi with every gt box, argmax / max per roi (:140-144) and the ordered index lists the sampler draws
 * from -- fg (max >= fg_thresh, :148), bg (bg_lo <= max < bg_hi, :160-161) and the TRAIN-mode fallback bg
 * (fallback_lo <= max < fallback_hi, :162-166) -- and returns them in one copy (counts3 = their lengths;
 * gt_assignment / max_overlaps optional).  The rois, gt boxes and assignments stay on the device.  The caller draws
 * with numpy's RNG in the reference's order, then phase2 gathers the num_keep <= max_keep sampled rows exactly like
 * mxr_proposal_target_gather and returns all five outputs in one copy.  phase2 consumes the state of phase1. */
int mxr_proposal_target_phase1_host(mxr_handle_t h, const float* rois, int num_rois, const float* gt_boxes, int num_gt,
                                    int max_keep, int num_classes, double fg_thresh, double bg_hi, double bg_lo,
                                    double fallback_hi, double fallback_lo, int32_t* fg_list, int32_t* bg_list,
                                    int32_t* fallback_list, int32_t* counts3, int32_t* gt_assignment, double* max_overlaps);
int mxr_proposal_target_phase2_host(mxr_handle_t h, const mxr_proposal_target_params* p, const int32_t* keep_inds,
                                    float* out_rois, float* labels, float* bbox_targets, float* inside, float* outside);

/* ---- nms without any allocation (no counterpart in the reference) --------------------- */
/* Same result as mxr_nms on device arrays, for callers that cannot allocate GPU memory: a plug-in
 * inside a framework whose pool already owns the whole device (a cudaMalloc of ours would fail or
 * eat into that pool), or a request path that cannot take the device-wide stall of cudaMalloc /
 * cudaFree.  Takes no handle and allocates nothing: the scratch is a fixed 640 KiB array inside the
 * library image, held by every GPU that loads the library until the process exits.
 *   - n <= MXR_NMS_NOALLOC_MAX_BOXES, otherwise MXR_ERR_LIMIT (use mxr_nms for larger inputs).
 *   - runs on the CURRENT device; the first call on a device resolves the array's device address
 *     (one-time set-up): make that call before any stream capture starts.
 *   - every call on a device reuses the same array (it is cleared where needed at the start of each
 *     call), so two calls whose GPU work overlaps race: keep them on one stream or serialise them.
 * dets / keep_out / nkeep_out are device pointers as for mxr_nms; nothing is synchronised. */
#define MXR_NMS_NOALLOC_MAX_BOXES 2048
int mxr_nms_noalloc(void* stream, const void* dets, int n, int dtype, double thresh, int max_keep, int32_t* keep_out,
                    int32_t* nkeep_out);

#ifdef __cplusplus
}
#endif
#endif /* MXRCNN_B200_H_ */
