/*
 * ypp.h -- C ABI of libypp.so: B200 (sm_100a) YOLO post-processing hot path.
 *
 * Drop-in boundary for the reference's Python helpers (the reference has no FFI of
 * its own; these are the entry points a binding for this path would need -- see
 * INTEGRATION.md for the ctypes stub).  Paths below are relative to the reference
 * checkout (myungsanglee/PyTorch-Object-Detection).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller unless named host_*;
 *   - every call is asynchronous on the given stream (a cudaStream_t passed as void*),
 *     never synchronises, never allocates: scratch memory comes from the caller and its
 *     size is queried first (ypp_workspace_bytes / ypp_map_workspace_bytes);
 *   - returns 0 on success, a negative YPP_E_* otherwise; ypp_last_error() gives the
 *     text for the calling thread;  re-entrant per (workspace, stream); no global state.
 *   - fp32 throughout; heads are contiguous NCHW exactly as the detectors emit them
 *     (models/detector/yolov3.py:71-91, models/detector/yolov2.py).
 */
#ifndef YPP_H_
#define YPP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YPP_MAX_SCALES 4
#define YPP_MAX_ANCHORS 8
#define YPP_MAX_CLASSES 1024      /* class id field of the sort key is 12 bits; S1 smem table */
#define YPP_MAX_BOXES (1 << 20)   /* box index field of the sort key is 20 bits */

enum {
    YPP_OK = 0,
    YPP_E_BADARG = -1,        /* null pointer, non-positive size, bad enum */
    YPP_E_UNSUPPORTED = -2,   /* shape outside the limits above */
    YPP_E_WORKSPACE = -3,     /* workspace too small / misaligned */
    YPP_E_CUDA = -4           /* a CUDA call failed; text in ypp_last_error() */
};

/* score written to the rows / used for thresholding and ordering */
enum {
    YPP_SCORE_OBJ = 0,            /* sigmoid(objectness): what the reference computes
                                     (utils/yolov3_utils.py:50,59) */
    YPP_SCORE_OBJ_TIMES_CLS = 1   /* objectness x best class probability (the vendored,
                                     uncalled utils/general.py:684 flavour) */
};

/* how torchvision.ops.batched_nms separates classes (site-packages boxes.py, v0.26.0) */
enum {
    YPP_OFFSET_NONE = 0,       /* _batched_nms_vanilla: per class, unshifted boxes */
    YPP_OFFSET_TRICK = 1,      /* _batched_nms_coordinate_trick: + cls*(max+1), class-agnostic */
    YPP_OFFSET_AUTO_CPU = 2,   /* per image: 4*n_cand > 4000   ? vanilla : trick */
    YPP_OFFSET_AUTO_CUDA = 3   /* per image: 4*n_cand > 100000 ? vanilla : trick */
};

/* rounding of the union term inside torchvision::nms */
enum {
    YPP_UNION_SEPARATE = 0,    /* (Sa + Sb) - inter, both areas rounded (CPU kernel) */
    YPP_UNION_FMA = 1          /* fma(w_b, h_b, Sa) - inter (nvcc contraction of the CUDA kernel) */
};

/* One detection head (one scale).  Replaces the per-scale arguments of
 * decode_predictions: utils/yolov3_utils.py:12 / utils/yolov2_utils.py:12. */
typedef struct {
    const float* head;                  /* [B, A*(5+nc), H, W] */
    int A, H, W;
    float anchor_w[YPP_MAX_ANCHORS];    /* GRID units, rounded to fp32 once by the caller  */
    float anchor_h[YPP_MAX_ANCHORS];    /* (yolov3_utils.py:31,40; yolov2 passes them as is) */
    float stride_w, stride_h;           /* input_size / layer_w, input_size / layer_h      */
} ypp_scale;

typedef struct {
    int nc;               /* classes */
    int B;                /* images in the batch */
    int n_scales;         /* 1..YPP_MAX_SCALES, concatenated in this order (yolov3_utils.py:84) */
    float conf_thr;       /* keep score > conf_thr (strict, utils/yolo_utils.py:334) */
    float iou_thr;        /* suppress iou > iou_thr (strict fp32 compare; see ypp_effective_iou_thr) */
    int score_mode;       /* YPP_SCORE_* */
    int offset_mode;      /* YPP_OFFSET_* */
    int union_mode;       /* YPP_UNION_* */
} ypp_cfg;

/* ---- whole chain: decode -> threshold/compact -> sort -> per-class NMS -> rows -------- */

/* Bytes of scratch for a batch of cfg->B images of n_boxes boxes each (n_boxes is the sum
 * over scales of A*H*W, or N for ypp_nms_rows). */
size_t ypp_workspace_bytes(int B, int n_boxes, int nc);

/* torchvision's CPU kernel compares the fp32 IoU against a DOUBLE threshold, the CUDA
 * kernel against (float)thr.  Returns the fp32 value to put in cfg->iou_thr so that the
 * strict fp32 compare reproduces either (compare_in_double = 1 for the CPU behaviour). */
float ypp_effective_iou_thr(double iou_thr, int compare_in_double);

/* = decode_predictions on every scale + torch.cat(dim=1): out [B, n_boxes, 6] rows
 * cx,cy,w,h,score,cls(float).  utils/yolov3_utils.py:12-59,76-84; utils/yolov2_utils.py:12-58. */
int ypp_decode_dense(const ypp_cfg* cfg, const ypp_scale* scales, float* out_BN6, void* stream);

/* D1: fused decode + confidence threshold + stream compaction into the workspace
 * (replaces decode_predictions + the torch.where of nms_v3, utils/yolo_utils.py:334). */
int ypp_decode_compact(const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes,
                       void* stream);

/* Same compaction, but from already decoded rows [B, n_boxes, 6] (the nms_v3 entry point:
 * utils/yolo_utils.py:321-337). */
int ypp_rows_compact(const ypp_cfg* cfg, const float* rows_BN6, int n_boxes, void* ws,
                     size_t ws_bytes, void* stream);

/* S1: per image stable order by score (desc, ties -> lower box index) and per class
 * segmentation (scores.sort(stable=True, descending=True) inside torchvision::nms). */
int ypp_sort_segments(const ypp_cfg* cfg, int n_boxes, void* ws, size_t ws_bytes, void* stream);

/* N1+N2: bitmask IoU suppression + greedy keep scan per (image, class) segment
 * (torchvision batched_nms as called from utils/yolo_utils.py:346), then the row gather
 * (boxes[keep], utils/yolo_utils.py:348, and the img_idx column of
 * utils/yolov3_utils.py:126-129).
 *   cand_count [B] out (optional): boxes passing the threshold
 *   keep_count [B] out: boxes kept per image
 *   keep_idx   [B, n_boxes] out (optional): kept box indices n in NMS order, per image
 *   det_rows   out (optional): packed rows, image after image, row_width floats each:
 *              7 -> img,cx,cy,w,h,score,cls ; 6 -> cx,cy,w,h,score,cls
 *   det_total  [1] out (optional): total rows written
 *   det_capacity: rows that fit in det_rows (extra rows are dropped; det_total still counts them) */
int ypp_nms(const ypp_cfg* cfg, int n_boxes, void* ws, size_t ws_bytes,
            int* cand_count, int* keep_count, int* keep_idx,
            float* det_rows, int row_width, int64_t det_capacity, int* det_total,
            float img_idx_base, void* stream);

/* Optional companion of every append-mode call: the kernel that advances the device cursor also stores
 * (new cursor value << 24) | (seq & 0xffffff) - ONE 8-byte word - into pinned host memory (cudaHostAlloc / cudaMallocHost
 * memory is device-accessible under unified addressing).  The host then knows, without a copy, an event or a wait, how
 * many rows the accumulator held after append number `seq`, and sizes its buffer from that plus the worst case of the
 * appends issued since (utils/yolov3_utils.py:139-146 reads the count through torch.cat on every batch).  NULL = none. */
typedef struct {
    int64_t* host_word;   /* pinned host memory, 8-byte aligned */
    uint32_t seq;         /* the caller's number for this append (only the low 24 bits travel) */
} ypp_probe;

/* ypp_nms in append mode: the accumulation `self.all_pred_boxes_variable = torch.cat([self.all_pred_boxes_variable,
 * pred_boxes])` of MeanAveragePrecision.update_state (utils/yolov3_utils.py:139-144) done by the gather itself,
 * with no host read of the row count between batches.  det_rows is the whole accumulator ([det_capacity,
 * row_width]); this call's rows land after the *cursor_in rows already there and *cursor_out receives
 * *cursor_in + rows of this call.  cursor_in and cursor_out are device int64 words and must be DIFFERENT words
 * (blocks of the gather read the one while another writes the other): alternate between two.  Rows past
 * det_capacity are dropped while the cursor still counts them, so the caller sizes det_capacity for the worst
 * case (B * n_boxes rows per call) between two reads of the cursor. */
int ypp_nms_append(const ypp_cfg* cfg, int n_boxes, void* ws, size_t ws_bytes, int* cand_count, int* keep_count,
                   float* det_rows, int row_width, int64_t det_capacity, const int64_t* cursor_in,
                   int64_t* cursor_out, float img_idx_base, void* stream, const ypp_probe* probe);

/* ypp_postprocess with the gather in append mode (one call per batch of a validation loop). */
int ypp_postprocess_append(const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes, int* cand_count,
                           int* keep_count, float* det_rows, int row_width, int64_t det_capacity,
                           const int64_t* cursor_in, int64_t* cursor_out, float img_idx_base, void* stream,
                           const ypp_probe* probe);

/* The three stages above, back to back, from the heads (DecodeYoloV3/V2.forward and the body
 * of MeanAveragePrecision.update_state: utils/yolov3_utils.py:73-90,108-146). */
int ypp_postprocess(const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes,
                    int* cand_count, int* keep_count, int* keep_idx,
                    float* det_rows, int row_width, int64_t det_capacity, int* det_total,
                    float img_idx_base, void* stream);

/* ypp_postprocess for callers that cannot provide scratch memory (code under CUDA graph capture set-up, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The chain's workspace is a device array built into the library - YPP_STATIC_WS_BYTES (4 MiB) on every
 * GPU that loads libypp.so, held until the process exits - so the call allocates nothing: no cudaMalloc / cudaFree, no
 * allocating library routine.  Inputs and outputs are the caller's, with the meaning they have in ypp_postprocess.
 * Limit: ypp_workspace_bytes(cfg->B, n_boxes, cfg->nc) <= YPP_STATIC_WS_BYTES, else YPP_E_WORKSPACE with the sizes in
 * ypp_last_error() (one YOLOv3-416 image - 10 647 boxes, 80 classes - needs 0.63 MB; four of them, six with 20
 * classes, or one 608-pixel image of 22 743 boxes fit; a batch of 8 does not).  Nothing in the workspace has to start at zero (every kernel of the chain initialises what it
 * reads), so the array simply keeps the last call's contents.  ALL calls on a device share the one array: calls that
 * overlap on the GPU (two streams, two host threads) race - serialise them on one stream or with events. */
#define YPP_STATIC_WS_BYTES (4 << 20)
int ypp_postprocess_static(const ypp_cfg* cfg, const ypp_scale* scales, int* cand_count, int* keep_count, int* keep_idx,
                           float* det_rows, int row_width, int64_t det_capacity, int* det_total, float img_idx_base,
                           void* stream);

/* Any contiguous part of the chain as ONE call: stages is a mask of YPP_STAGE_* (DECODE = D1, SORT = S1,
 * SUPPRESS = N1 (+ the long-segment kernel), ROWS = the row gather G).  The state between stages lives in
 * the workspace, so a chain may be cut anywhere and continued by a later call on the same workspace (on any
 * stream that is ordered after the first).  What it is for: several batches in flight, each on its own
 * stream and workspace, with only the ROWS stages - which append behind one device cursor and therefore run
 * in batch order - chained by events (the validation loop of module/yolov3_detector.py:33-47 without the
 * host ever waiting for the device).  scales is read when DECODE is set (n_boxes may then be 0 = derive it);
 * the outputs and cursors when ROWS is set, with the meaning they have in ypp_nms / ypp_nms_append
 * (cursor_in == cursor_out == NULL: plain mode). */
#define YPP_STAGE_DECODE 1
#define YPP_STAGE_SORT 2
#define YPP_STAGE_SUPPRESS 4
#define YPP_STAGE_ROWS 8
int ypp_run_stages(const ypp_cfg* cfg, const ypp_scale* scales, int n_boxes, void* ws, size_t ws_bytes, int stages,
                   int* cand_count, int* keep_count, int* keep_idx, float* det_rows, int row_width,
                   int64_t det_capacity, int* det_total, const int64_t* cursor_in, int64_t* cursor_out,
                   float img_idx_base, void* stream, const ypp_probe* probe);

/* ---- the head's 1x1 convolution fused with the decode ---------------------------
 * Every detector ends each scale in nn.Conv2d(Cin, A*(5+nc), 1, 1, bias=False)
 * (models/detector/yolov3.py:56,62,68; yolov2.py; yolov4_tiny.py) and hands its output to
 * decode_predictions (utils/yolov3_utils.py:12).  These entry points take the INPUT of that
 * convolution instead and never materialise the head: one tcgen05 (TF32 products, fp32
 * accumulation - cuDNN's arithmetic for this layer under torch's default allow_tf32) kernel per scale
 * decodes, thresholds and compacts the accumulator tiles.  scales[i].head is ignored. */
typedef struct {
    const float* feature;   /* [B, H, W, Cin] fp32, channels-last (x.permute(0,2,3,1).contiguous() of the
                               conv input, or a torch.channels_last tensor's storage); 16-byte aligned */
    const float* weight;    /* [A*(5+nc), Cin] fp32: conv.weight.reshape(A*(5+nc), Cin); 16-byte aligned */
    int cin;                /* multiple of 4 */
    float* head_out;        /* optional [B, A*(5+nc), H, W]: the head tensor, for inspection / parity */
} ypp_head_src;

/* = the convolutions + ypp_decode_compact */
int ypp_head_decode_compact(const ypp_cfg* cfg, const ypp_scale* scales, const ypp_head_src* src,
                            void* ws, size_t ws_bytes, void* stream);
/* = the convolutions + ypp_postprocess */
int ypp_head_postprocess(const ypp_cfg* cfg, const ypp_scale* scales, const ypp_head_src* src,
                         void* ws, size_t ws_bytes, int* cand_count, int* keep_count, int* keep_idx,
                         float* det_rows, int row_width, int64_t det_capacity, int* det_total,
                         float img_idx_base, void* stream);

/* nms_v3 on decoded rows [B, n_boxes, 6] (utils/yolo_utils.py:321-348), batched. */
int ypp_nms_rows(const ypp_cfg* cfg, const float* rows_BN6, int n_boxes, void* ws, size_t ws_bytes,
                 int* cand_count, int* keep_count, int* keep_idx,
                 float* det_rows, int row_width, int64_t det_capacity, int* det_total,
                 float img_idx_base, void* stream);

/* ---- alternate semantics: the vendored Ultralytics non_max_suppression -----
 * utils/general.py:640-731.  pred [B, n_boxes, 5+nc] rows cx,cy,w,h,obj,cls_0.. (already decoded);
 * per image:  obj > conf_thr (:645)  ->  conf_j = cls_j * obj (:669)  ->  best class only: the first
 * maximum, kept when conf > conf_thr (:679-680); multi_label: every (box, j) with conf_j > conf_thr in
 * (box, j) order (:676-677)  ->  optional class filter (:683-684)  ->  the max_nms best by conf when
 * more remain (:694-695; the reference's argsort is unstable, here ties keep (box, j) order)  ->
 * torchvision.ops.nms on boxes + cls * max_wh (agnostic: + 0) (:698-700)  ->  the first max_det kept
 * (:701-702).  labels (autolabelling) and merge-NMS (off in the reference) are not provided.
 * Output rows are x1,y1,x2,y2,conf,cls (row_width 6) or img,x1,..,cls (7), image after image in NMS order.
 * keep_idx holds box numbers (best class) or box * nc + j (multi_label).  The workspace is
 * ypp_workspace_bytes(B, ypp_ultra_rows(n_boxes, nc, multi_label), nc). */
typedef struct {
    int nc, B;
    float conf_thr;       /* strict compares, fp32 */
    float iou_thr;        /* see ypp_effective_iou_thr */
    int multi_label;      /* forced to 0 when nc == 1, as the reference does (:659) */
    int agnostic;
    int max_det;          /* <= 0: no limit */
    int max_nms;          /* <= 0: no limit */
    float max_wh;         /* 4096 in the reference (:656) */
    const unsigned char* class_allow;   /* device [nc] 0/1, or NULL = every class (classes=None) */
    int union_mode;       /* YPP_UNION_* */
} ypp_ultra_cfg;

/* rows per image the workspace must be sized for: n_boxes, or n_boxes * nc with multi_label */
int64_t ypp_ultra_rows(int n_boxes, int nc, int multi_label);

int ypp_nms_ultralytics(const ypp_ultra_cfg* cfg, const float* pred_BNC, int n_boxes, void* ws, size_t ws_bytes,
                        int* cand_count, int* keep_count, int* keep_idx,
                        float* det_rows, int row_width, int64_t det_capacity, int* det_total,
                        float img_idx_base, void* stream);

/* decode_predictions of YOLO v1, utils/yolov1_utils.py:12-66: input [B, S*S*(5*num_boxes + nc)]
 * (cell-major, per cell: nc class scores, then num_boxes x (conf, x, y, w, h)) -> out [B, S*S, 6]
 * cx,cy,w,h,confidence,class with the box of the most confident predictor of each cell. */
int ypp_decode_yolov1(const float* input, int B, int S, int nc, int num_boxes, float input_size,
                      float* out_BS6, void* stream);

/* ---- mAP side ------------------------------------------------------------------------ */

/* get_target_boxes_for_map + the img_idx column (utils/yolo_utils.py:682-710,
 * utils/yolov3_utils.py:134-137): annot [B,T,5] -> packed rows [G,7] img,cx,cy,w,h,1,cls in
 * (image, slot) order; rows with sum <= 0 are dropped.  gt_count [1] out.  out needs B*T rows. */
int ypp_target_rows(const float* annot_BT5, int B, int T, float input_size, float img_idx_base,
                    float* out_G7, int* gt_count, void* stream);

/* The same in append mode (`self.all_true_boxes_variable = torch.cat(...)`, utils/yolov3_utils.py:141-145):
 * out_G7 is the whole accumulator [capacity,7]; rows go after the *cursor rows already there and *cursor
 * (device int64, in/out - the kernel is one block) is advanced.  A call adds at most B*T rows. */
int ypp_target_rows_append(const float* annot_BT5, int B, int T, float input_size, float img_idx_base,
                           float* out_G7, int64_t capacity, int64_t* cursor, void* stream, const ypp_probe* probe);

/* ---- detection writers --------------------------------------------------
 * The integer box columns the reference's detection writers put on disk, for K rows
 * [img,]cx,cy,w,h,score,cls (row_width 7 or 6):
 *   YPP_WRITER_PRED_FILE  make_pred_file_yolov3.py:73-86 (and the yolov2 twin):
 *       xmin,ymin,xmax,ymax = int(cx -/+ w/2), int(cy -/+ h/2)
 *   YPP_WRITER_COCO       yolo2coco_pred_file_yolov3.py:83-100: cx,cy,w,h are first multiplied by
 *       (width / input_size), (height / input_size) of the row's image; bbox = int(cx - w/2),
 *       int(cy - h/2), int(w), int(h)
 * int() truncates toward zero; values must fit int32 (NaN -> 0, out of range saturates).
 * scale_xy [n_img,2] (COCO only; fp32 of the Python-float ratios) is indexed by
 * (int)row[0] - img_idx_base for 7-column rows, and by 0 for 6-column rows. */
#define YPP_WRITER_PRED_FILE 0
#define YPP_WRITER_COCO 1
int ypp_detection_ints(const float* rows, int64_t K, int row_width, int mode, const float* scale_xy, int n_img,
                       float img_idx_base, int* out_K4, void* stream);

/* box_iou, utils/metrics.py:233-255: xyxy [N,4] x [M,4] -> [N,M]. */
int ypp_box_iou(const float* a_N4, int N, const float* b_M4, int M, float* out_NM, void* stream);

/* bbox_iou, utils/metrics.py:192-230 (the Ultralytics flavour: eps on the union, no abs; the mAP matching uses the
 * one of utils/yolo_utils.py, inside ypp_map_match): box1 [n1,4] against box2 [n2,4] element by element with
 * n1 == n2 or either equal to 1 (broadcast), out [max(n1,n2)] (the reference returns [n,1]).
 * xywh != 0: rows are cx,cy,w,h, else x1,y1,x2,y2.  mode: */
#define YPP_IOU 0
#define YPP_GIOU 1
#define YPP_DIOU 2
#define YPP_CIOU 3      /* atan through CUDA's atanf: within an ulp of torch's on any device */
int ypp_bbox_iou(const float* box1, int n1, const float* box2, int n2, int xywh, int mode, double eps, float* out,
                 void* stream);

/* ConfusionMatrix.process_batch, utils/metrics.py:122-160, for n_img images in one launch (the
 * reference is called once per image):
 *   det [K,6] x1,y1,x2,y2,conf,cls of all images, det_off [n_img+1] row offsets (det_off[n_img] = K)
 *   lab [G,5] cls,x1,y1,x2,y2, lab_off [n_img+1]
 *   matrix [(nc+1)*(nc+1)] int64, accumulated: matrix[predicted][true]; index nc = background
 *   scratch: K * 8 bytes.  Class ids outside [0, nc) are skipped (the reference raises).
 * Equal IoUs (unspecified in the reference: unstable argsort) go to the lower index. */
int ypp_confusion_update(const float* det_K6, const int* det_off, const float* lab_G5, const int* lab_off, int n_img,
                         int64_t K, int nc, float conf_thr, float iou_thr, void* scratch, size_t scratch_bytes,
                         int64_t* matrix, void* stream);

size_t ypp_map_workspace_bytes(int64_t G, int64_t K, int nc);

/* mean_average_precision / metrics_per_class, utils/yolo_utils.py:351-456,458-574.
 *   gt [G,7], det [K,7] rows img,cx,cy,w,h,score,cls.  Detections in any order.  Ground truths in any order too, but
 *   rows sorted by image (what the accumulators above produce, one process or gathered rank after rank) are found by
 *   binary search; any other order is detected on the device and handled by a scan of all G rows per detection
 *   tp [K] out (optional): 1 TP, 0 FP, 255 class without ground truth
 *   per_class [nc,4] out (optional): AP, TP, FP, FN
 *   map [1] out: mean AP over all nc classes */
int ypp_map_match(const float* gt_G7, int64_t G, const float* det_K7, int64_t K, int nc,
                  float iou_thr, void* ws, size_t ws_bytes,
                  unsigned char* tp, float* per_class, float* map, void* stream);

/* The same matching split for an image-sharded evaluation (the reference never gathers, its metric
 * state is rank-local: module/yolov3_detector.py:45-47).  An image's detections and ground truths live on ONE
 * rank, and the greedy rule of utils/yolo_utils.py:404-430 never looks across images, so
 *   M1  ypp_map_flags          on every rank, over its own rows: the TP flag of every detection (no sort of the
 *                              detections: "first detection to choose a ground truth" is a 64-bit atomic minimum
 *                              of (score descending, row order));
 *   --  all-gather             of the (score, cls, tp) triples - 12 bytes per detection instead of the 28-byte row,
 *                              and no ground-truth rows at all, only their per-class counts;
 *   M2  ypp_map_ap_from_flags  on the rank that owns classes [c_lo, c_hi): select them, stable sort by score, AP;
 *   --  all-gather             of the per-class results, then ypp_map_mean.
 * The result is bit-identical to ypp_map_match over the gathered rows when the segments are concatenated in the
 * order one process would have seen the images (ties in score keep that order).
 *   tp [K] out (optional): 1 TP, 0 FP, 255 class without ground truth ON THIS RANK (pass the global counts through
 *                          ypp_map_ap_from_flags for the real decision; the triples carry the plain flag)
 *   triples [K,3] out (optional): score, cls, tp (0/1) per detection, in row order
 *   gt_per_class [nc] out (optional): ground truths of every class among gt_G7 (device int64) */
int ypp_map_flags(const float* gt_G7, int64_t G, const float* det_K7, int64_t K, int nc, float iou_thr,
                  void* ws, size_t ws_bytes, unsigned char* tp, float* triples, int64_t* gt_per_class, void* stream);

/* triples: n_seg segments of seg_stride rows each (the padded all-gather buffer as it arrives); seg_count [n_seg]
 * device int64 = valid rows of every segment, or NULL = all.  gt_per_class [nc] device int64 (summed over the ranks).
 * per_class [nc,4]: rows c_lo .. c_hi-1 are written (AP, TP, FP, FN).  Workspace: ypp_map_workspace_bytes(0,
 * n_seg * seg_stride, nc). */
int ypp_map_ap_from_flags(const float* triples, int64_t n_seg, int64_t seg_stride, const int64_t* seg_count,
                          const int64_t* gt_per_class, int nc, int c_lo, int c_hi, void* ws, size_t ws_bytes,
                          float* per_class, void* stream);

/* mean of column 0 of per_class [nc,4] over all nc classes, fp32 in class order (torch.mean, utils/yolo_utils.py:456) */
int ypp_map_mean(const float* per_class_nc4, int nc, float* map, void* stream);

/* ---- several batches in flight (the validation loop of module/yolov3_detector.py:33-47) --------------------------
 * MeanAveragePrecision.update_state (utils/yolov3_utils.py:108-146) is called batch after batch, and a batch's
 * chain leaves most of the GPU idle for a third of its time (the per-image sort is one CTA per image).  A pipeline
 * owns `depth` side streams (and two events each; no device memory): ypp_pipeline_submit runs batch i on stream
 * i mod depth - it starts behind whatever the caller's stream has queued, its decode / sort / suppression overlap
 * the neighbouring batches', and only its row gather waits for the previous batch's, because both append behind one
 * device cursor (ypp_postprocess_append's contract: two different cursor words, alternate).  One call per batch
 * replaces ~10 runtime calls from the host language.  The caller provides a workspace PER SLOT
 * (ypp_pipeline_next_slot says which the next submit uses) and keeps inputs, workspaces and the accumulator alive
 * until the batch has completed.  probe (optional, see ypp_probe): the gather reports the cursor to pinned host
 * memory, so the host sizes its buffers from the newest completed count without ever blocking.
 * Not thread-safe: one pipeline per host thread.  Created on, and bound to, the current device. */
#define YPP_PIPELINE_MAX_DEPTH 8
typedef struct ypp_pipeline ypp_pipeline;
int ypp_pipeline_create(int depth, ypp_pipeline** out);
int ypp_pipeline_destroy(ypp_pipeline* p);
int ypp_pipeline_next_slot(const ypp_pipeline* p);
void* ypp_pipeline_stream(const ypp_pipeline* p, int slot);      /* cudaStream_t of a slot (allocator bookkeeping) */
int ypp_pipeline_submit(ypp_pipeline* p, const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes,
                        int* cand_count, int* keep_count, float* det_rows, int row_width, int64_t det_capacity,
                        const int64_t* cursor_in, int64_t* cursor_out, const ypp_probe* probe, float img_idx_base,
                        void* caller_stream, int64_t* ticket);
int ypp_pipeline_query(const ypp_pipeline* p, int64_t ticket);          /* 1 = that batch has completed */
int ypp_pipeline_wait_ticket(const ypp_pipeline* p, int64_t ticket);    /* the HOST waits for that batch only */
int ypp_pipeline_join(const ypp_pipeline* p, void* stream);             /* `stream` waits on the device for all batches */

/* Stream-ordered copy of a few bytes (<= 4096) of device memory - an append cursor, a count - into PINNED host memory,
 * without waiting for it.  The accumulation of MeanAveragePrecision.update_state (utils/yolov3_utils.py:139-146) needs
 * the row count on the host only to size its buffer; with a probe behind every append and an event behind the probe
 * the host reads the newest COMPLETED count and never blocks on the device.  Never synchronises. */
int ypp_peek_async(const void* device_src, void* host_dst, size_t bytes, void* stream);

/* ---- misc ---------------------------------------------------------------------------- */

/* Launch-width hint for the long-segment NMS kernel (segments > 512 boxes; no counterpart in the reference, where
 * torchvision sizes every launch after reading counts back).  Its CTAs take a whole SM each, so launching it at full
 * width for batches that hold no long segment costs every batch in flight a few per cent.  enable != 0: the library
 * pins one word per GPU of its OWN host memory (cudaHostRegister, once per process: nothing is allocated) which that
 * kernel sets the first time it finds a long segment on the device; until then configurations with many classes
 * (boxes per image / classes < 1024) launch it 8 CTAs wide, afterwards - for the rest of the process - at full width.
 * Results never depend on it; the first batch with long segments runs them on 8 CTAs.  A CUDA graph captured before
 * that keeps the narrow launch.  enable == 0 (the default of the C library): always full width.  The Python layer
 * turns it on.  Returns YPP_E_CUDA if the host memory cannot be pinned (the library then stays at full width). */
int ypp_adaptive_launch(int enable);

const char* ypp_last_error(void);
int ypp_version(void);
/* kernels launched by this library in the calling thread since the last reset (bench bookkeeping) */
int64_t ypp_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* YPP_H_ */
