// Kernel parameter blocks and host launcher prototypes (internal to libypp).
#pragma once
#include "ypp_common.cuh"
#include <cstdlib>

struct YppScaleParams {
    const float* head;
    int A, H, W, HW;
    int box_base;        // first box index of this scale inside the image (cat order)
    int tma;             // 1: planes staged with cp.async.bulk (H*W % 4 == 0, head 16 B aligned); 0: plain loads
    int tpp;             // tiles per anchor plane = ceil(HW / YPP_TILE_BOXES)
    long long elems;     // floats in the head tensor (bulk copies never read past it)
    float aw[YPP_MAX_ANCHORS], ah[YPP_MAX_ANCHORS];
    float sw, sh;
    float inv_w;         // 1 / W
};

struct YppDecodeParams {
    YppScaleParams sc[YPP_MAX_SCALES];
    int tile_base[YPP_MAX_SCALES + 1];   // blockIdx.x ranges per scale
    int n_scales, nc, C, N;
    float conf;
    int score_mode;
    int stage_bytes;     // dynamic shared memory of one stage of the staged path (0 when no scale uses it)
    int pitch;           // floats per channel row a stage must hold: 256, or 264 when some plane's rows are not 16 B-aligned
    YppWs ws;
};

struct YppOrgParams {
    YppWs ws;
    int B, N, nc;
    int offset_mode, union_mode;
    float iou_thr;
    int keys_in_smem;      // key buffers of the sort held in shared memory: 2 (both), 1 (the first), 0
    int table_ints;        // 32 * nc
    int step_mode;         // 0: class offset step = max coordinate + 1 (torchvision batched_nms); 1: fixed_step, always shifted
    float fixed_step;      //    (Ultralytics: boxes + cls * max_wh, utils/general.py:698-699; 0 = agnostic)
    int max_cand;          // > 0: only the best max_cand candidates of an image enter NMS (max_nms, utils/general.py:694-695)
    int small_max;         // images with at most this many candidates are finished inside S1 (set by the launcher)
};

struct YppNmsParams {
    YppWs ws;
    int B, N, nc;
    float iou_thr;
    int union_mode;
    int sparse_min_len;    // per-class segments at least this long try the sweep-and-prune path first
    int big_smem_bytes;    // dynamic shared memory of the long-segment kernel (its sweep-and-prune path sizes itself by it)
    int* big_seen;         // page-locked host word the long-segment kernel sets when it finds work (ypp_adaptive_launch), or null
};

struct YppGatherParams {
    YppWs ws;
    int B, N;
    int* cand_count_out;
    int* keep_count_out;
    int* keep_idx;
    float* det_rows;
    int row_width;
    long long det_capacity;
    int* det_total;
    float img_idx_base;
    const float* src_rows;   // optional [B,N,6]: copy the class column verbatim (nms_rows entry)
    int max_det;             // > 0: at most this many rows per image (utils/general.py:701-702)
    int xyxy_out;            // rows carry x1,y1,x2,y2 instead of cx,cy,w,h
    const long long* cursor_in;   // optional: rows already in det_rows; this call's rows go after them
    long long* cursor_out;        // optional: *cursor_in + rows of this call (a different word than cursor_in)
    long long* host_probe;        // optional, pinned host memory (device-accessible): receives (*cursor_out << 24) | probe_seq
    unsigned probe_seq;
};

int ypp_launch_decode_compact(const YppDecodeParams& p, int B, cudaStream_t st);
int ypp_launch_head_decode(const YppDecodeParams& p, int B, const float* const* feature, const float* const* weight,
                           const int* cin, float* const* head_out, cudaStream_t st);
int ypp_launch_decode_dense(const YppDecodeParams& p, int B, float* out, cudaStream_t st);
int ypp_launch_rows_compact(const float* rows, int B, int N, float conf, const YppWs& ws, cudaStream_t st);
int ypp_launch_pred_compact(const float* pred, int B, int n_boxes, int nc, float conf, int multi_label,
                            const unsigned char* class_allow, const YppWs& ws, int n_rows, cudaStream_t st);
int ypp_launch_decode_yolov1(const float* input, int B, int S, int nc, int num_boxes, float input_size, float* out, cudaStream_t st);
int ypp_launch_organise(const YppOrgParams& p, cudaStream_t st);
int ypp_launch_nms(const YppNmsParams& p, cudaStream_t st);
int ypp_launch_gather(const YppGatherParams& p, cudaStream_t st);
int ypp_org_smem_bytes(int N, int nc, int* keys_in_smem, int* table_ints);

int ypp_launch_detection_ints(const float* rows, long long K, int row_width, int mode, const float* scale_xy, int n_img,
                               float img_base, int* out, cudaStream_t st);
int ypp_launch_confusion(const float* det, const int* det_off, const float* lab, const int* lab_off, int n_img, int nc,
                         float conf, float thr, int* bestg, float* bestiou, unsigned long long* matrix, cudaStream_t st);
int ypp_launch_target_rows(const float* annot, int B, int T, float input_size, float img_base,
                           float* out, int* gt_count, long long capacity, const long long* cursor_in,
                           long long* cursor_out, long long* host_probe, unsigned probe_seq, cudaStream_t st);
int ypp_launch_box_iou(const float* a, int N, const float* b, int M, float* out, cudaStream_t st);
int ypp_launch_bbox_iou(const float* a, int na, const float* b, int nb, int xywh, int mode, float eps, float one_eps,
                        float* out, cudaStream_t st);
size_t ypp_map_ws_bytes(long long G, long long K, int nc);
int ypp_launch_map(const float* gt, long long G, const float* det, long long K, int nc, float iou_thr,
                   void* ws, size_t ws_bytes, unsigned char* tp, float* per_class, float* map,
                   cudaStream_t st);

int ypp_launch_map_flags(const float* gt, long long G, const float* det, long long K, int nc, float iou_thr, void* ws,
                         size_t ws_bytes, unsigned char* tp, float* triples, long long* gt_per_class, cudaStream_t st);
int ypp_launch_map_ap_flags(const float* triples, long long n, long long seg_stride, const long long* seg_count,
                            const long long* gt_per_class, int nc, int c_lo, int c_hi, void* ws, size_t ws_bytes,
                            float* per_class, cudaStream_t st);
int ypp_adaptive_launch_set(int enable);      // nms.cu
int ypp_launch_map_mean(const float* per_class, int nc, float* map, cudaStream_t st);

// ---- programmatic dependent launch (PDL) --------------------------------------------------
// The chain D1 -> S1 -> N1 -> N1b -> G is five short kernels; with the attribute below the next kernel's
// CTAs are scheduled while the previous one drains and park in ypp_grid_wait() (griddepcontrol.wait:
// returns once the WHOLE previous grid has completed and flushed), so launch latency and CTA set-up
// leave the critical path.  Every kernel of the chain calls ypp_grid_wait() first thing in every CTA
// (completion then chains transitively) and releases its successor right after.  Measured on the C2
// chain: early launch pays for D1 (after the previous call's G), N1b and G (-2 us each group) and
// costs ~2.5 us each for S1 and N1, whose CTAs then land on the SMs in a worse order; hence the default mask.
#ifdef __CUDACC__
__device__ __forceinline__ void ypp_grid_wait()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif

template <class... KArgs, class... Args>
inline cudaError_t ypp_launch_chain(int edge, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    static int pdl = -1;
    if (pdl < 0) { const char* e = getenv("YPP_PDL"); pdl = e ? atoi(e) : 25; }     // bit per kernel: D1, S1, N1, N1b, G
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = ((pdl >> edge) & 1) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}
