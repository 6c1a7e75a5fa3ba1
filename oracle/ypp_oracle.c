/*
 * ypp_oracle.c -- CPU restatement of the reference YOLO post-processing path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product package may link, import or
 * call this file; it exists so that tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs have something to check the
 * CUDA path against (and to time on the host cores).
 *
 * Parity status: PINNED against outputs of the reference itself, generated in
 * the build container by tests/golden/make_golden.py (imports /root/reference
 * and torchvision 0.26.0 CPU) and committed under tests/golden/.  The reference
 * ships no tests / golden vectors of its own for this path.
 *
 * Each function cites the reference lines it follows (paths relative to the
 * reference checkout).  The suppression arithmetic itself lives in a third
 * party wheel that is not part of the reference tree (torchvision 0.26.0,
 * torchvision.ops.batched_nms / torch.ops.torchvision.nms); its published
 * algorithm is restated in ypo_nms_sorted() and pinned through the reference's
 * single call site utils/yolo_utils.py:346.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC
 * (-ffp-contract=off matters: the separate-rounding union must not be fused).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define YPO_API __attribute__((visibility("default")))

/* ---------------------------------------------------------------- decode -- */

static inline float ypo_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

/*
 * One scale of decode_predictions.
 *   utils/yolov3_utils.py:27-59 (anchors in pixels; the caller divides by the
 *   stride in float64 and rounds to fp32 once, as :31,:40 do) and
 *   utils/yolov2_utils.py:27-58 (anchors already in grid units).
 * head  [B, A*(5+nc), H, W] fp32 NCHW
 * out   [B, A*H*W, 6] rows cx,cy,w,h,score,cls ; box index = a*H*W + y*W + x
 * grid_x and grid_y are both built from arange(layer_w) in the reference
 * (:43-44) so the map is assumed square; y*W+x with y<H is what view() yields.
 */
YPO_API void ypo_decode_scale(const float* head, int B, int A, int nc, int H, int W,
                              const float* anchor_w, const float* anchor_h,
                              float stride_w, float stride_h,
                              float* out, int64_t out_image_stride, int64_t out_box_offset)
{
    const int C = 5 + nc;
    const int64_t HW = (int64_t)H * W;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b) {
        for (int a = 0; a < A; ++a) {
            const float* base = head + ((int64_t)b * A + a) * C * HW;
            for (int y = 0; y < H; ++y) {
                for (int x = 0; x < W; ++x) {
                    const int64_t cell = (int64_t)y * W + x;
                    float* row = out + (int64_t)b * out_image_stride
                                     + (out_box_offset + (int64_t)a * HW + cell) * 6;
                    /* pxy = sigmoid(t)+grid (:47); pbox *= stride (:54) */
                    float px = ypo_sigmoid(base[0 * HW + cell]) + (float)x;
                    float py = ypo_sigmoid(base[1 * HW + cell]) + (float)y;
                    /* pwh = exp(t)*scaled_anchor (:48), then *stride (:54): two roundings */
                    float pw = expf(base[2 * HW + cell]) * anchor_w[a];
                    float ph = expf(base[3 * HW + cell]) * anchor_h[a];
                    row[0] = px * stride_w;
                    row[1] = py * stride_h;
                    row[2] = pw * stride_w;
                    row[3] = ph * stride_h;
                    row[4] = ypo_sigmoid(base[4 * HW + cell]);               /* :50 */
                    /* argmax over sigmoid(cls) AFTER the sigmoid (:51,:57): first max wins */
                    int best = 0;
                    float best_v = ypo_sigmoid(base[5 * HW + cell]);
                    for (int c = 1; c < nc; ++c) {
                        float v = ypo_sigmoid(base[(5 + c) * HW + cell]);
                        if (v > best_v) { best_v = v; best = c; }
                    }
                    row[5] = (float)best;
                }
            }
        }
    }
}

/* ------------------------------------------------------------------- NMS -- */

typedef struct { float s; int i; } ypo_key;

/* stable descending order by score, ties -> lower index first
 * (scores.sort(stable=True, descending=True) inside torchvision::nms) */
static int ypo_key_cmp(const void* pa, const void* pb)
{
    const ypo_key* a = (const ypo_key*)pa;
    const ypo_key* b = (const ypo_key*)pb;
    if (a->s > b->s) return -1;
    if (a->s < b->s) return 1;
    return (a->i > b->i) - (a->i < b->i);
}

/*
 * Greedy NMS on xyxy boxes -- restatement of torch.ops.torchvision.nms
 * (torchvision 0.26.0).  union_mode 0: areas rounded separately,
 * union = (Sa + Sb) - inter (CPU kernel).  union_mode 1: union =
 * fma(w_b, h_b, Sa) - inter, Sa = row (higher score) box, b = column box
 * (what nvcc contracts the CUDA kernel's devIoU to).
 * No +1, no eps, no abs; suppress iff iou > thr (strict).  thr_strict is the
 * float threshold already adjusted by the caller for the CPU kernel's
 * double-precision compare (see ypo_effective_thr).
 * Returns number kept; keep[] receives input indices in descending-score order.
 */
static int ypo_nms_sorted(const float* box, const float* score, int n,
                          float thr_strict, int union_mode, int* keep)
{
    if (n <= 0) return 0;
    ypo_key* keys = (ypo_key*)malloc(sizeof(ypo_key) * (size_t)n);
    unsigned char* dead = (unsigned char*)calloc((size_t)n, 1);
    float* area = (float*)malloc(sizeof(float) * (size_t)n);
    for (int i = 0; i < n; ++i) {
        keys[i].s = score[i]; keys[i].i = i;
        area[i] = (box[4 * i + 2] - box[4 * i + 0]) * (box[4 * i + 3] - box[4 * i + 1]);
    }
    qsort(keys, (size_t)n, sizeof(ypo_key), ypo_key_cmp);
    int k = 0;
    for (int oi = 0; oi < n; ++oi) {
        const int i = keys[oi].i;
        if (dead[i]) continue;
        keep[k++] = i;
        const float ax1 = box[4 * i], ay1 = box[4 * i + 1], ax2 = box[4 * i + 2], ay2 = box[4 * i + 3];
        const float sa = area[i];
        for (int oj = oi + 1; oj < n; ++oj) {
            const int j = keys[oj].i;
            if (dead[j]) continue;
            const float bx1 = box[4 * j], by1 = box[4 * j + 1], bx2 = box[4 * j + 2], by2 = box[4 * j + 3];
            const float l = fmaxf(ax1, bx1), t = fmaxf(ay1, by1);
            const float r = fminf(ax2, bx2), bt = fminf(ay2, by2);
            const float w = fmaxf(r - l, 0.0f), h = fmaxf(bt - t, 0.0f);
            const float inter = w * h;
            float uni;
            if (union_mode == 1) uni = fmaf(bx2 - bx1, by2 - by1, sa) - inter;
            else                 uni = (sa + area[j]) - inter;
            const float iou = inter / uni;
            if (iou > thr_strict) dead[j] = 1;
        }
    }
    free(keys); free(dead); free(area);
    return k;
}

/*
 * The CPU kernel of torchvision::nms takes iou_threshold as a double and
 * compares the float IoU against it; the CUDA kernel compares against
 * (float)thr.  Both reduce to a strict float compare against an adjusted
 * threshold: if (double)(float)thr > thr then "iou > thr (double)" equals
 * "iou >= (float)thr" equals "iou > nextbelow((float)thr)".
 */
YPO_API float ypo_effective_thr(double thr, int compare_in_double)
{
    float tf = (float)thr;
    if (compare_in_double && (double)tf > thr) tf = nextafterf(tf, -INFINITY);
    return tf;
}

YPO_API int ypo_nms_xyxy(const float* box, const float* score, int n,
                         float thr_strict, int union_mode, int* keep)
{
    return ypo_nms_sorted(box, score, n, thr_strict, union_mode, keep);
}

/*
 * torchvision.ops.batched_nms (boxes.py, v0.26.0) as called from
 * utils/yolo_utils.py:346.
 *   offset_mode 1: _batched_nms_coordinate_trick -- off = cls * (max(boxes)+1)
 *                  in fp32, class-agnostic nms on boxes + off.
 *   offset_mode 0: _batched_nms_vanilla -- per class nms, final order by score
 *                  (non-stable in torch; canonicalised here: score desc, index asc).
 * cls are float class ids.  keep receives indices into the n inputs.
 */
YPO_API int ypo_batched_nms(const float* box, const float* score, const float* cls, int n,
                            float thr_strict, int offset_mode, int union_mode, int* keep)
{
    if (n <= 0) return 0;
    if (offset_mode == 1) {
        float m = box[0];
        for (int i = 1; i < 4 * n; ++i) if (box[i] > m) m = box[i];
        const float step = m + 1.0f;
        float* sb = (float*)malloc(sizeof(float) * 4 * (size_t)n);
        for (int i = 0; i < n; ++i) {
            const float off = cls[i] * step;
            for (int q = 0; q < 4; ++q) sb[4 * i + q] = box[4 * i + q] + off;
        }
        int k = ypo_nms_sorted(sb, score, n, thr_strict, union_mode, keep);
        free(sb);
        return k;
    }
    /* vanilla: loop over the distinct class ids */
    unsigned char* kept = (unsigned char*)calloc((size_t)n, 1);
    unsigned char* done = (unsigned char*)calloc((size_t)n, 1);
    int* idx = (int*)malloc(sizeof(int) * (size_t)n);
    int* sub_keep = (int*)malloc(sizeof(int) * (size_t)n);
    float* sb = (float*)malloc(sizeof(float) * 4 * (size_t)n);
    float* ss = (float*)malloc(sizeof(float) * (size_t)n);
    for (int s = 0; s < n; ++s) {
        if (done[s]) continue;
        const float c = cls[s];
        int m = 0;
        for (int i = s; i < n; ++i) {
            if (!done[i] && cls[i] == c) {
                done[i] = 1; idx[m] = i;
                memcpy(sb + 4 * m, box + 4 * i, 4 * sizeof(float));
                ss[m] = score[i]; ++m;
            }
        }
        int k = ypo_nms_sorted(sb, ss, m, thr_strict, union_mode, sub_keep);
        for (int q = 0; q < k; ++q) kept[idx[sub_keep[q]]] = 1;
    }
    ypo_key* keys = (ypo_key*)malloc(sizeof(ypo_key) * (size_t)n);
    int k = 0;
    for (int i = 0; i < n; ++i) if (kept[i]) { keys[k].s = score[i]; keys[k].i = i; ++k; }
    qsort(keys, (size_t)k, sizeof(ypo_key), ypo_key_cmp);
    for (int q = 0; q < k; ++q) keep[q] = keys[q].i;
    free(keys); free(kept); free(done); free(idx); free(sub_keep); free(sb); free(ss);
    return k;
}

/*
 * nms_v3, utils/yolo_utils.py:321-348, on one image's decoded rows [N,6].
 *   score > conf (strict, fp32, :334); xyxy = c -/+ wh/2 (:339-343);
 *   batched_nms (:346).  offset_mode: 0 vanilla, 1 trick, 2 auto as the CPU
 *   build of torchvision decides (numel > 4000 -> vanilla), 3 auto as the
 *   CUDA build decides (numel > 100000 -> vanilla).
 * keep_src receives the kept BOX indices n (into the N rows), NMS order.
 * Returns the number kept; *n_cand_out the number passing the threshold.
 */
YPO_API int ypo_nms_v3(const float* rows, int N, float conf, float thr_strict,
                       int offset_mode, int union_mode, int* keep_src, int* n_cand_out)
{
    int* src = (int*)malloc(sizeof(int) * (size_t)(N > 0 ? N : 1));
    int n = 0;
    for (int i = 0; i < N; ++i) if (rows[6 * i + 4] > conf) src[n++] = i;
    if (n_cand_out) *n_cand_out = n;
    if (n == 0) { free(src); return 0; }
    float* box = (float*)malloc(sizeof(float) * 4 * (size_t)n);
    float* sc = (float*)malloc(sizeof(float) * (size_t)n);
    float* cl = (float*)malloc(sizeof(float) * (size_t)n);
    int* keep = (int*)malloc(sizeof(int) * (size_t)n);
    for (int q = 0; q < n; ++q) {
        const float* r = rows + 6 * (int64_t)src[q];
        const float hw = r[2] / 2.0f, hh = r[3] / 2.0f;
        box[4 * q + 0] = r[0] - hw; box[4 * q + 1] = r[1] - hh;
        box[4 * q + 2] = r[0] + hw; box[4 * q + 3] = r[1] + hh;
        sc[q] = r[4]; cl[q] = r[5];
    }
    int mode = offset_mode;
    if (mode == 2) mode = (4 * (int64_t)n > 4000) ? 0 : 1;
    if (mode == 3) mode = (4 * (int64_t)n > 100000) ? 0 : 1;
    int k = ypo_batched_nms(box, sc, cl, n, thr_strict, mode, union_mode, keep);
    for (int q = 0; q < k; ++q) keep_src[q] = src[keep[q]];
    free(src); free(box); free(sc); free(cl); free(keep);
    return k;
}

/*
 * Batched driver used for timing and parity: rows [B,N,6] -> per image keep
 * lists (keep_src [B,N], keep_count [B], cand_count [B]).  Images in parallel.
 */
YPO_API void ypo_nms_v3_batch(const float* rows, int B, int N, float conf, float thr_strict,
                              int offset_mode, int union_mode,
                              int* keep_src, int* keep_count, int* cand_count)
{
#pragma omp parallel for schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        int nc_ = 0;
        keep_count[b] = ypo_nms_v3(rows + (int64_t)b * N * 6, N, conf, thr_strict,
                                   offset_mode, union_mode, keep_src + (int64_t)b * N, &nc_);
        cand_count[b] = nc_;
    }
}

/* ------------------------------------------------------------------- IoU -- */

/* bbox_iou, IoU branch, cxcywh inputs: utils/yolo_utils.py:103-123,143-144
 * (abs() on the areas, eps = 1e-6 added to the union). */
static inline float ypo_bbox_iou_cxcywh(const float* a, const float* b)
{
    const float ax1 = a[0] - a[2] / 2.0f, ay1 = a[1] - a[3] / 2.0f;
    const float ax2 = a[0] + a[2] / 2.0f, ay2 = a[1] + a[3] / 2.0f;
    const float bx1 = b[0] - b[2] / 2.0f, by1 = b[1] - b[3] / 2.0f;
    const float bx2 = b[0] + b[2] / 2.0f, by2 = b[1] + b[3] / 2.0f;
    const float ix1 = fmaxf(ax1, bx1), iy1 = fmaxf(ay1, by1);
    const float ix2 = fminf(ax2, bx2), iy2 = fminf(ay2, by2);
    const float inter = fmaxf(ix2 - ix1, 0.0f) * fmaxf(iy2 - iy1, 0.0f);
    const float sa = fabsf((ax2 - ax1) * (ay2 - ay1));
    const float sb = fabsf((bx2 - bx1) * (by2 - by1));
    const float uni = ((sa + sb) - inter) + 1e-6f;
    return inter / uni;
}

YPO_API void ypo_bbox_iou_cxcywh_pairs(const float* a, const float* b, int n, float* out)
{
    for (int i = 0; i < n; ++i) out[i] = ypo_bbox_iou_cxcywh(a + 4 * i, b + 4 * i);
}

/* box_iou, utils/metrics.py:233-255: N x M xyxy, no eps, clamp(0).prod */
YPO_API void ypo_box_iou(const float* a, int N, const float* b, int M, float* out)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        const float sa = (a[4 * i + 2] - a[4 * i]) * (a[4 * i + 3] - a[4 * i + 1]);
        for (int j = 0; j < M; ++j) {
            const float sb = (b[4 * j + 2] - b[4 * j]) * (b[4 * j + 3] - b[4 * j + 1]);
            const float w = fmaxf(fminf(a[4 * i + 2], b[4 * j + 2]) - fmaxf(a[4 * i], b[4 * j]), 0.0f);
            const float h = fmaxf(fminf(a[4 * i + 3], b[4 * j + 3]) - fmaxf(a[4 * i + 1], b[4 * j + 1]), 0.0f);
            const float inter = w * h;
            out[(int64_t)i * M + j] = inter / ((sa + sb) - inter);
        }
    }
}

/* ------------------------------------------------------------ GT decoding -- */

/*
 * get_target_boxes_for_map, utils/yolo_utils.py:682-710, flattened: for every
 * (b,t) with sum(row) > 0 emit [img_base+b, cx*S, cy*S, w*S, h*S, 1, cid]
 * (the img column is what update_state prepends, utils/yolov3_utils.py:134-137).
 * Returns the number of rows written to out [<= B*T, 7].
 */
YPO_API int ypo_target_rows(const float* annot, int B, int T, float input_size,
                            float img_base, float* out)
{
    int g = 0;
    for (int b = 0; b < B; ++b) {
        for (int t = 0; t < T; ++t) {
            const float* r = annot + ((int64_t)b * T + t) * 5;
            /* torch .sum() of 5 fp32 values: sequential fp32 adds */
            float s = r[0]; s += r[1]; s += r[2]; s += r[3]; s += r[4];
            if (s <= 0.0f) continue;
            float* o = out + 7 * (int64_t)g++;
            o[0] = img_base + (float)b;
            o[1] = r[0] * input_size; o[2] = r[1] * input_size;
            o[3] = r[2] * input_size; o[4] = r[3] * input_size;
            o[5] = 1.0f; o[6] = r[4];
        }
    }
    return g;
}

/* ------------------------------------------------------------------ mAP -- */

typedef struct { float s; int i; } ypo_det_key;

/*
 * mean_average_precision, utils/yolo_utils.py:351-456 (metrics_per_class
 * :458-574 shares the matching and adds TP/FP/FN sums).
 *   gt  [G,7], det [K,7] rows img,cx,cy,w,h,score,cls
 *   tp  [K]   out: 1 = true positive, 0 = false positive, 255 = class had no
 *             GT (reference skips the class entirely, :379-381)
 *   per_class [nc,4] out: AP, TP, FP, FN (zeros when the class has no GT)
 * Detections of a class are visited in descending score, ties in original
 * order (torch.sort on CPU is stable in practice; canonical choice).
 * Best-IoU GT only (strict >, first max wins, :412-417); TP iff best_iou > thr
 * and that GT unseen (:419-430).  AP: cumsum, eps 1e-6, (1,0) prepended,
 * right-to-left max envelope, sum over recall change points (:432-452).
 * Returns the mean AP over all nc classes (:456).
 */
YPO_API float ypo_mean_average_precision(const float* gt, int G, const float* det, int K,
                                         int nc, float iou_thr,
                                         unsigned char* tp_out, float* per_class)
{
    const float eps = 1e-6f;
    float ap_sum = 0.0f;
    unsigned char* seen = (unsigned char*)malloc((size_t)(G > 0 ? G : 1));
    int* gsel = (int*)malloc(sizeof(int) * (size_t)(G > 0 ? G : 1));
    ypo_key* dk = (ypo_key*)malloc(sizeof(ypo_key) * (size_t)(K > 0 ? K : 1));
    float* prec = (float*)malloc(sizeof(float) * (size_t)(K + 2));
    float* rec = (float*)malloc(sizeof(float) * (size_t)(K + 2));
    if (tp_out) memset(tp_out, 255, (size_t)K);
    for (int c = 0; c < nc; ++c) {
        const float cf = (float)c;
        int ng = 0, nd = 0;
        for (int g = 0; g < G; ++g) if (gt[7 * g + 6] == cf) gsel[ng++] = g;
        float* pc = per_class ? per_class + 4 * c : NULL;
        if (pc) { pc[0] = pc[1] = pc[2] = pc[3] = 0.0f; }
        if (ng == 0) continue;                       /* AP 0 appended */
        for (int d = 0; d < K; ++d) if (det[7 * d + 6] == cf) { dk[nd].s = det[7 * d + 5]; dk[nd].i = d; ++nd; }
        qsort(dk, (size_t)nd, sizeof(ypo_key), ypo_key_cmp);
        memset(seen, 0, (size_t)ng);
        float tpc = 0.0f, fpc = 0.0f;
        const float rec_den = (float)((double)ng + 1e-6);
        prec[0] = 1.0f; rec[0] = 0.0f;
        for (int q = 0; q < nd; ++q) {
            const float* d = det + 7 * (int64_t)dk[q].i;
            float best = 0.0f; int best_g = -1;
            for (int s = 0; s < ng; ++s) {
                const float* g = gt + 7 * (int64_t)gsel[s];
                if (g[0] != d[0]) continue;
                const float iou = ypo_bbox_iou_cxcywh(d + 1, g + 1);
                if (iou > best) { best = iou; best_g = s; }
            }
            int is_tp = 0;
            if (best > iou_thr && !seen[best_g]) { seen[best_g] = 1; is_tp = 1; }
            if (tp_out) tp_out[dk[q].i] = (unsigned char)is_tp;
            if (is_tp) tpc += 1.0f; else fpc += 1.0f;     /* cumsum in fp32 */
            rec[q + 1] = tpc / rec_den;   /* python float64 (G + 1e-6) rounded to fp32 by the tensor / scalar op */
            prec[q + 1] = tpc / ((tpc + fpc) + eps);
        }
        for (int i = nd; i > 0; --i) if (prec[i] > prec[i - 1]) prec[i - 1] = prec[i];
        float ap = 0.0f;
        for (int i = 0; i < nd; ++i)
            if (rec[i + 1] != rec[i]) ap += (rec[i + 1] - rec[i]) * prec[i + 1];
        ap_sum += ap;
        if (pc) { pc[0] = ap; pc[1] = tpc; pc[2] = fpc; pc[3] = (float)ng - tpc; }
    }
    free(seen); free(gsel); free(dk); free(prec); free(rec);
    /* torch.mean(torch.stack(...)): fp32 sum / nc */
    return ap_sum / (float)nc;
}

YPO_API int ypo_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

YPO_API void ypo_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
