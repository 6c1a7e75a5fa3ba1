// extern "C" surface of libypp.so (declared in include/ypp.h): argument checking, parameter
// blocks, stream-ordered launches.  No allocation, no synchronisation, no global state.
#include "ypp_kernels.cuh"
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <new>

static thread_local char g_err[512] = "";
static thread_local long long g_launches = 0;

void ypp_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int ypp_check_launch(const char* what)
{
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        ypp_set_error("%s: %s", what, cudaGetErrorString(e));
        return YPP_E_CUDA;
    }
    return YPP_OK;
}

void ypp_count_launch(int n) { g_launches += n; }

namespace {


int check_cfg(const ypp_cfg* cfg)
{
    if (!cfg) { ypp_set_error("cfg is null"); return YPP_E_BADARG; }
    if (cfg->B <= 0 || cfg->nc <= 0) { ypp_set_error("B=%d nc=%d must be positive", cfg->B, cfg->nc); return YPP_E_BADARG; }
    if (cfg->nc > YPP_MAX_CLASSES) { ypp_set_error("nc=%d > %d", cfg->nc, YPP_MAX_CLASSES); return YPP_E_UNSUPPORTED; }
    if (cfg->offset_mode < 0 || cfg->offset_mode > 3 || cfg->union_mode < 0 || cfg->union_mode > 1 ||
        cfg->score_mode < 0 || cfg->score_mode > 1) { ypp_set_error("bad mode enum"); return YPP_E_BADARG; }
    if (cfg->B > 65535) { ypp_set_error("B=%d > 65535", cfg->B); return YPP_E_UNSUPPORTED; }
    return YPP_OK;
}

int build_decode_params(const ypp_cfg* cfg, const ypp_scale* sc, bool dense, YppDecodeParams* p, bool need_head = true)
{
    int rc = check_cfg(cfg);
    if (rc) return rc;
    if (!sc || cfg->n_scales < 1 || cfg->n_scales > YPP_MAX_SCALES) { ypp_set_error("n_scales=%d out of range", cfg->n_scales); return YPP_E_BADARG; }
    memset(p, 0, sizeof(*p));
    p->n_scales = cfg->n_scales;
    p->nc = cfg->nc;
    p->C = 5 + cfg->nc;
    p->conf = cfg->conf_thr;
    p->score_mode = cfg->score_mode;
    // rows of a plane whose size is not a multiple of 4 floats (13x13, 19x19) start at any 4-byte offset: their
    // copies fetch the enclosing 16-byte-aligned span, so a staged row needs up to 2 x 3 floats of slack
    p->pitch = YPP_TILE_BOXES;
    for (int i = 0; i < cfg->n_scales; ++i)
        if ((sc[i].H * sc[i].W) % 4 != 0) p->pitch = YPP_TILE_BOXES + 8;
    int box_base = 0, tile = 0;
    for (int i = 0; i < cfg->n_scales; ++i) {
        const ypp_scale& s = sc[i];
        if ((need_head && !s.head) || s.A < 1 || s.A > YPP_MAX_ANCHORS || s.H < 1 || s.W < 1) { ypp_set_error("scale %d: bad head/A/H/W", i); return YPP_E_BADARG; }
        YppScaleParams& d = p->sc[i];
        d.head = s.head; d.A = s.A; d.H = s.H; d.W = s.W; d.HW = s.H * s.W;
        d.box_base = box_base;
        d.tpp = (d.HW + YPP_TILE_BOXES - 1) / YPP_TILE_BOXES;
        d.elems = (long long)cfg->B * s.A * p->C * d.HW;
        // staged (cp.async.bulk) path: the tensor must start 16-byte aligned and the tile must fit shared memory
        d.tma = (!dense && (((uintptr_t)s.head & 15u) == 0) && p->C * p->pitch * 4 <= 100 * 1024) ? 1 : 0;
        if (d.tma) p->stage_bytes = p->C * p->pitch * 4;
        for (int a = 0; a < s.A; ++a) { d.aw[a] = s.anchor_w[a]; d.ah[a] = s.anchor_h[a]; }
        d.sw = s.stride_w; d.sh = s.stride_h;
        d.inv_w = 1.0f / (float)s.W;
        p->tile_base[i] = tile;
        tile += s.A * d.tpp;
        box_base += s.A * d.HW;
        if (box_base > YPP_MAX_BOXES) { ypp_set_error("more than %d boxes per image", YPP_MAX_BOXES); return YPP_E_UNSUPPORTED; }
    }
    for (int i = cfg->n_scales; i <= YPP_MAX_SCALES; ++i) p->tile_base[i] = tile;
    p->N = box_base;
    return YPP_OK;
}

int bind_ws(int B, int N, int nc, void* ws, size_t ws_bytes, YppWs* out)
{
    if (!ws) { ypp_set_error("workspace is null"); return YPP_E_WORKSPACE; }
    if (((uintptr_t)ws & 255u) != 0) { ypp_set_error("workspace must be 256-byte aligned"); return YPP_E_WORKSPACE; }
    *out = ypp_make_ws(ws, B, N, nc);
    if (out->total_bytes > ws_bytes) { ypp_set_error("workspace %zu bytes < required %zu", ws_bytes, out->total_bytes); return YPP_E_WORKSPACE; }
    return YPP_OK;
}

int run_sort_nms(const ypp_cfg* cfg, int N, const YppWs& w, int* cand_count, int* keep_count, int* keep_idx,
                 float* det_rows, int row_width, int64_t det_capacity, int* det_total, float img_idx_base,
                 const float* src_rows, bool do_sort, bool do_nms, cudaStream_t st,
                 const long long* cursor_in = nullptr, long long* cursor_out = nullptr, bool do_rows = true,
                 const ypp_probe* probe = nullptr)
{
    int rc = YPP_OK;
    if (do_sort) {
        YppOrgParams o;
        memset(&o, 0, sizeof(o));
        o.ws = w; o.B = cfg->B; o.N = N; o.nc = cfg->nc;
        o.offset_mode = cfg->offset_mode; o.union_mode = cfg->union_mode; o.iou_thr = cfg->iou_thr;
        rc = ypp_launch_organise(o, st);
        if (rc) return rc;
    }
    if (do_nms) {
        YppNmsParams n;
        memset(&n, 0, sizeof(n));
        n.ws = w; n.B = cfg->B; n.N = N; n.nc = cfg->nc; n.iou_thr = cfg->iou_thr; n.union_mode = cfg->union_mode;
        rc = ypp_launch_nms(n, st);
        if (rc) return rc;
    }
    if (do_rows) {
        if (det_rows && row_width != 6 && row_width != 7) { ypp_set_error("row_width must be 6 or 7"); return YPP_E_BADARG; }
        YppGatherParams g;
        memset(&g, 0, sizeof(g));
        g.ws = w; g.B = cfg->B; g.N = N;
        g.cand_count_out = cand_count; g.keep_count_out = keep_count; g.keep_idx = keep_idx;
        g.det_rows = det_rows; g.row_width = row_width; g.det_capacity = det_capacity; g.det_total = det_total;
        g.img_idx_base = img_idx_base; g.src_rows = src_rows;
        g.cursor_in = cursor_in; g.cursor_out = cursor_out;
        if (probe && probe->host_word && cursor_out) { g.host_probe = (long long*)probe->host_word; g.probe_seq = probe->seq; }
        rc = ypp_launch_gather(g, st);
    }
    return rc;
}

}  // namespace

extern "C" {

size_t ypp_workspace_bytes(int B, int n_boxes, int nc)
{
    if (B <= 0 || n_boxes <= 0 || nc <= 0) return 0;
    return ypp_make_ws(nullptr, B, n_boxes, nc).total_bytes;
}

float ypp_effective_iou_thr(double iou_thr, int compare_in_double)
{
    float tf = (float)iou_thr;
    if (compare_in_double && (double)tf > iou_thr) tf = nextafterf(tf, -INFINITY);
    return tf;
}

int ypp_decode_dense(const ypp_cfg* cfg, const ypp_scale* scales, float* out_BN6, void* stream)
{
    YppDecodeParams p;
    int rc = build_decode_params(cfg, scales, true, &p);
    if (rc) return rc;
    if (!out_BN6) { ypp_set_error("out is null"); return YPP_E_BADARG; }
    return ypp_launch_decode_dense(p, cfg->B, out_BN6, (cudaStream_t)stream);
}

int ypp_decode_compact(const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes, void* stream)
{
    YppDecodeParams p;
    int rc = build_decode_params(cfg, scales, false, &p);
    if (rc) return rc;
    rc = bind_ws(cfg->B, p.N, cfg->nc, ws, ws_bytes, &p.ws);
    if (rc) return rc;
    return ypp_launch_decode_compact(p, cfg->B, (cudaStream_t)stream);
}

int ypp_rows_compact(const ypp_cfg* cfg, const float* rows_BN6, int n_boxes, void* ws, size_t ws_bytes, void* stream)
{
    int rc = check_cfg(cfg);
    if (rc) return rc;
    if (!rows_BN6 || n_boxes <= 0 || n_boxes > YPP_MAX_BOXES) { ypp_set_error("rows null or n_boxes=%d out of range", n_boxes); return YPP_E_BADARG; }
    YppWs w;
    rc = bind_ws(cfg->B, n_boxes, cfg->nc, ws, ws_bytes, &w);
    if (rc) return rc;
    return ypp_launch_rows_compact(rows_BN6, cfg->B, n_boxes, cfg->conf_thr, w, (cudaStream_t)stream);
}

int ypp_sort_segments(const ypp_cfg* cfg, int n_boxes, void* ws, size_t ws_bytes, void* stream)
{
    int rc = check_cfg(cfg);
    if (rc) return rc;
    YppWs w;
    rc = bind_ws(cfg->B, n_boxes, cfg->nc, ws, ws_bytes, &w);
    if (rc) return rc;
    return run_sort_nms(cfg, n_boxes, w, nullptr, nullptr, nullptr, nullptr, 7, 0, nullptr, 0.f, nullptr, true, false,
                        (cudaStream_t)stream, nullptr, nullptr, false);
}

int ypp_nms(const ypp_cfg* cfg, int n_boxes, void* ws, size_t ws_bytes, int* cand_count, int* keep_count, int* keep_idx,
            float* det_rows, int row_width, int64_t det_capacity, int* det_total, float img_idx_base, void* stream)
{
    int rc = check_cfg(cfg);
    if (rc) return rc;
    YppWs w;
    rc = bind_ws(cfg->B, n_boxes, cfg->nc, ws, ws_bytes, &w);
    if (rc) return rc;
    return run_sort_nms(cfg, n_boxes, w, cand_count, keep_count, keep_idx, det_rows, row_width, det_capacity, det_total,
                        img_idx_base, nullptr, false, true, (cudaStream_t)stream);
}

int ypp_nms_append(const ypp_cfg* cfg, int n_boxes, void* ws, size_t ws_bytes, int* cand_count, int* keep_count,
                   float* det_rows, int row_width, int64_t det_capacity, const int64_t* cursor_in, int64_t* cursor_out,
                   float img_idx_base, void* stream, const ypp_probe* probe)
{
    int rc = check_cfg(cfg);
    if (rc) return rc;
    if (!det_rows || !cursor_in || !cursor_out || cursor_in == cursor_out) {
        ypp_set_error("nms_append: det_rows and two different cursor words are required"); return YPP_E_BADARG;
    }
    YppWs w;
    rc = bind_ws(cfg->B, n_boxes, cfg->nc, ws, ws_bytes, &w);
    if (rc) return rc;
    return run_sort_nms(cfg, n_boxes, w, cand_count, keep_count, nullptr, det_rows, row_width, det_capacity, nullptr,
                        img_idx_base, nullptr, false, true, (cudaStream_t)stream, (const long long*)cursor_in,
                        (long long*)cursor_out, true, probe);
}

int ypp_postprocess(const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes, int* cand_count,
                    int* keep_count, int* keep_idx, float* det_rows, int row_width, int64_t det_capacity,
                    int* det_total, float img_idx_base, void* stream)
{
    YppDecodeParams p;
    int rc = build_decode_params(cfg, scales, false, &p);
    if (rc) return rc;
    rc = bind_ws(cfg->B, p.N, cfg->nc, ws, ws_bytes, &p.ws);
    if (rc) return rc;
    rc = ypp_launch_decode_compact(p, cfg->B, (cudaStream_t)stream);
    if (rc) return rc;
    return run_sort_nms(cfg, p.N, p.ws, cand_count, keep_count, keep_idx, det_rows, row_width, det_capacity, det_total,
                        img_idx_base, nullptr, true, true, (cudaStream_t)stream);
}

int ypp_postprocess_append(const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes, int* cand_count,
                           int* keep_count, float* det_rows, int row_width, int64_t det_capacity,
                           const int64_t* cursor_in, int64_t* cursor_out, float img_idx_base, void* stream,
                           const ypp_probe* probe)
{
    YppDecodeParams p;
    int rc = build_decode_params(cfg, scales, false, &p);
    if (rc) return rc;
    if (!det_rows || !cursor_in || !cursor_out || cursor_in == cursor_out) {
        ypp_set_error("postprocess_append: det_rows and two different cursor words are required"); return YPP_E_BADARG;
    }
    rc = bind_ws(cfg->B, p.N, cfg->nc, ws, ws_bytes, &p.ws);
    if (rc) return rc;
    rc = ypp_launch_decode_compact(p, cfg->B, (cudaStream_t)stream);
    if (rc) return rc;
    return run_sort_nms(cfg, p.N, p.ws, cand_count, keep_count, nullptr, det_rows, row_width, det_capacity, nullptr,
                        img_idx_base, nullptr, true, true, (cudaStream_t)stream, (const long long*)cursor_in,
                        (long long*)cursor_out, true, probe);
}

int ypp_run_stages(const ypp_cfg* cfg, const ypp_scale* scales, int n_boxes, void* ws, size_t ws_bytes, int stages,
                   int* cand_count, int* keep_count, int* keep_idx, float* det_rows, int row_width, int64_t det_capacity,
                   int* det_total, const int64_t* cursor_in, int64_t* cursor_out, float img_idx_base, void* stream,
                   const ypp_probe* probe)
{
    if (stages <= 0 || stages > 15) { ypp_set_error("run_stages: stages=%d is not a mask of YPP_STAGE_*", stages); return YPP_E_BADARG; }
    if ((cursor_in == nullptr) != (cursor_out == nullptr) || (cursor_in && (const int64_t*)cursor_out == cursor_in)) {
        ypp_set_error("run_stages: append mode needs two different cursor words"); return YPP_E_BADARG;
    }
    if (cursor_in && !det_rows) { ypp_set_error("run_stages: append mode needs det_rows"); return YPP_E_BADARG; }
    YppWs w;
    int rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (stages & YPP_STAGE_DECODE) {
        YppDecodeParams p;
        rc = build_decode_params(cfg, scales, false, &p);
        if (rc) return rc;
        if (n_boxes > 0 && n_boxes != p.N) { ypp_set_error("run_stages: n_boxes=%d but the scales hold %d boxes", n_boxes, p.N); return YPP_E_BADARG; }
        n_boxes = p.N;
        rc = bind_ws(cfg->B, p.N, cfg->nc, ws, ws_bytes, &p.ws);
        if (rc) return rc;
        rc = ypp_launch_decode_compact(p, cfg->B, st);
        if (rc) return rc;
        w = p.ws;
    } else {
        rc = check_cfg(cfg);
        if (rc) return rc;
        if (n_boxes <= 0 || n_boxes > YPP_MAX_BOXES) { ypp_set_error("run_stages: n_boxes=%d out of range", n_boxes); return YPP_E_BADARG; }
        rc = bind_ws(cfg->B, n_boxes, cfg->nc, ws, ws_bytes, &w);
        if (rc) return rc;
    }
    return run_sort_nms(cfg, n_boxes, w, cand_count, keep_count, keep_idx, det_rows, row_width, det_capacity, det_total,
                        img_idx_base, nullptr, (stages & YPP_STAGE_SORT) != 0, (stages & YPP_STAGE_SUPPRESS) != 0, st,
                        (const long long*)cursor_in, (long long*)cursor_out, (stages & YPP_STAGE_ROWS) != 0, probe);
}

// ---- ypp_postprocess without caller scratch --------------------------------------------------------------
// The chain's whole workspace in a file-scope device array (YPP_STATIC_WS_BYTES; every GPU that loads the library
// holds it until exit).  In HOST code `g_static_ws` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) unsigned char g_static_ws[YPP_STATIC_WS_BYTES];

static void* static_ws_ptr(int* rc)
{
    static void* ptr_d[YPP_MAX_DEVICES];              // per device ordinal; a racing first call stores the same value twice
    int dev = 0;
    *rc = YPP_OK;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= YPP_MAX_DEVICES) {
        ypp_set_error("postprocess_static: device ordinal %d out of range", dev); *rc = YPP_E_UNSUPPORTED; return nullptr;
    }
    if (!ptr_d[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_ws);
        if (e != cudaSuccess || !p) {
            ypp_set_error("postprocess_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); *rc = YPP_E_CUDA; return nullptr;
        }
        ptr_d[dev] = p;
    }
    return ptr_d[dev];
}

int ypp_postprocess_static(const ypp_cfg* cfg, const ypp_scale* scales, int* cand_count, int* keep_count, int* keep_idx,
                           float* det_rows, int row_width, int64_t det_capacity, int* det_total, float img_idx_base,
                           void* stream)
{
    YppDecodeParams p;
    int rc = build_decode_params(cfg, scales, false, &p);
    if (rc) return rc;
    const size_t need = ypp_make_ws(nullptr, cfg->B, p.N, cfg->nc).total_bytes;
    if (need > (size_t)YPP_STATIC_WS_BYTES) {
        ypp_set_error("postprocess_static: B=%d x %d boxes x %d classes needs %zu bytes of scratch, the built-in array holds %d "
                      "(use ypp_postprocess with a caller workspace)", cfg->B, p.N, cfg->nc, need, YPP_STATIC_WS_BYTES);
        return YPP_E_WORKSPACE;
    }
    void* ws = static_ws_ptr(&rc);
    if (rc) return rc;
    rc = bind_ws(cfg->B, p.N, cfg->nc, ws, YPP_STATIC_WS_BYTES, &p.ws);
    if (rc) return rc;
    rc = ypp_launch_decode_compact(p, cfg->B, (cudaStream_t)stream);
    if (rc) return rc;
    return run_sort_nms(cfg, p.N, p.ws, cand_count, keep_count, keep_idx, det_rows, row_width, det_capacity, det_total,
                        img_idx_base, nullptr, true, true, (cudaStream_t)stream);
}

// the heads' 1x1 convolutions fused with D1: one tcgen05 kernel over every scale writes the slot records
static int head_decode_all(const ypp_cfg* cfg, const ypp_scale* scales, const ypp_head_src* src, void* ws, size_t ws_bytes,
                           YppDecodeParams* p, cudaStream_t st)
{
    int rc = build_decode_params(cfg, scales, false, p, false);
    if (rc) return rc;
    if (!src) { ypp_set_error("head sources are null"); return YPP_E_BADARG; }
    for (int i = 0; i < cfg->n_scales; ++i)
        if (!src[i].feature || !src[i].weight || src[i].cin < 1) { ypp_set_error("scale %d: feature / weight / cin missing", i); return YPP_E_BADARG; }
    rc = bind_ws(cfg->B, p->N, cfg->nc, ws, ws_bytes, &p->ws);
    if (rc) return rc;
    const float* feat[YPP_MAX_SCALES];
    const float* wgt[YPP_MAX_SCALES];
    float* hout[YPP_MAX_SCALES];
    int cin[YPP_MAX_SCALES];
    for (int i = 0; i < cfg->n_scales; ++i) { feat[i] = src[i].feature; wgt[i] = src[i].weight; hout[i] = src[i].head_out; cin[i] = src[i].cin; }
    return ypp_launch_head_decode(*p, cfg->B, feat, wgt, cin, hout, st);
}

int ypp_head_decode_compact(const ypp_cfg* cfg, const ypp_scale* scales, const ypp_head_src* src, void* ws, size_t ws_bytes,
                            void* stream)
{
    YppDecodeParams p;
    return head_decode_all(cfg, scales, src, ws, ws_bytes, &p, (cudaStream_t)stream);
}

int ypp_head_postprocess(const ypp_cfg* cfg, const ypp_scale* scales, const ypp_head_src* src, void* ws, size_t ws_bytes,
                         int* cand_count, int* keep_count, int* keep_idx, float* det_rows, int row_width,
                         int64_t det_capacity, int* det_total, float img_idx_base, void* stream)
{
    YppDecodeParams p;
    int rc = head_decode_all(cfg, scales, src, ws, ws_bytes, &p, (cudaStream_t)stream);
    if (rc) return rc;
    return run_sort_nms(cfg, p.N, p.ws, cand_count, keep_count, keep_idx, det_rows, row_width, det_capacity, det_total,
                        img_idx_base, nullptr, true, true, (cudaStream_t)stream);
}

int ypp_nms_rows(const ypp_cfg* cfg, const float* rows_BN6, int n_boxes, void* ws, size_t ws_bytes, int* cand_count,
                 int* keep_count, int* keep_idx, float* det_rows, int row_width, int64_t det_capacity, int* det_total,
                 float img_idx_base, void* stream)
{
    int rc = ypp_rows_compact(cfg, rows_BN6, n_boxes, ws, ws_bytes, stream);
    if (rc) return rc;
    YppWs w = ypp_make_ws(ws, cfg->B, n_boxes, cfg->nc);
    return run_sort_nms(cfg, n_boxes, w, cand_count, keep_count, keep_idx, det_rows, row_width, det_capacity, det_total,
                        img_idx_base, rows_BN6, true, true, (cudaStream_t)stream);
}

int64_t ypp_ultra_rows(int n_boxes, int nc, int multi_label)
{
    return (multi_label && nc > 1) ? (int64_t)n_boxes * nc : (int64_t)n_boxes;
}

int ypp_nms_ultralytics(const ypp_ultra_cfg* cfg, const float* pred_BNC, int n_boxes, void* ws, size_t ws_bytes,
                        int* cand_count, int* keep_count, int* keep_idx, float* det_rows, int row_width,
                        int64_t det_capacity, int* det_total, float img_idx_base, void* stream)
{
    if (!cfg || !pred_BNC) { ypp_set_error("nms_ultralytics: null pointer"); return YPP_E_BADARG; }
    if (cfg->nc < 1 || cfg->nc > YPP_MAX_CLASSES || cfg->B < 1 || cfg->B > 65535 || n_boxes < 1) {
        ypp_set_error("nms_ultralytics: nc=%d B=%d n_boxes=%d out of range", cfg->nc, cfg->B, n_boxes); return YPP_E_BADARG;
    }
    if (cfg->union_mode < 0 || cfg->union_mode > 1) { ypp_set_error("bad mode enum"); return YPP_E_BADARG; }
    if (!(cfg->max_wh >= 0.0f)) { ypp_set_error("nms_ultralytics: max_wh must be >= 0"); return YPP_E_BADARG; }
    if (det_rows && row_width != 6 && row_width != 7) { ypp_set_error("row_width must be 6 or 7"); return YPP_E_BADARG; }
    const int multi = (cfg->multi_label && cfg->nc > 1) ? 1 : 0;                 // multi_label &= nc > 1
    const int64_t rows = ypp_ultra_rows(n_boxes, cfg->nc, multi);
    // the sort key holds 20 bits of row number; the long-segment kernel keeps one removed-bit per row in shared memory
    if (rows > YPP_MAX_BOXES) { ypp_set_error("nms_ultralytics: %lld rows per image > %d", (long long)rows, YPP_MAX_BOXES); return YPP_E_UNSUPPORTED; }
    const int N = (int)rows;
    YppWs w;
    int rc = bind_ws(cfg->B, N, cfg->nc, ws, ws_bytes, &w);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    rc = ypp_launch_pred_compact(pred_BNC, cfg->B, n_boxes, cfg->nc, cfg->conf_thr, multi, cfg->class_allow, w, N, st);
    if (rc) return rc;
    YppOrgParams o;
    memset(&o, 0, sizeof(o));
    o.ws = w; o.B = cfg->B; o.N = N; o.nc = cfg->nc;
    o.offset_mode = YPP_OFFSET_TRICK; o.union_mode = cfg->union_mode; o.iou_thr = cfg->iou_thr;
    o.step_mode = 1;
    o.fixed_step = cfg->agnostic ? 0.0f : cfg->max_wh;       // c = x[:, 5:6] * (0 if agnostic else max_wh)
    o.max_cand = cfg->max_nms > 0 ? cfg->max_nms : 0;
    rc = ypp_launch_organise(o, st);
    if (rc) return rc;
    YppNmsParams n;
    memset(&n, 0, sizeof(n));
    n.ws = w; n.B = cfg->B; n.N = N; n.nc = cfg->nc; n.iou_thr = cfg->iou_thr; n.union_mode = cfg->union_mode;
    rc = ypp_launch_nms(n, st);
    if (rc) return rc;
    YppGatherParams g;
    memset(&g, 0, sizeof(g));
    g.ws = w; g.B = cfg->B; g.N = N;
    g.cand_count_out = cand_count; g.keep_count_out = keep_count; g.keep_idx = keep_idx;
    g.det_rows = det_rows; g.row_width = row_width; g.det_capacity = det_capacity; g.det_total = det_total;
    g.img_idx_base = img_idx_base;
    g.max_det = cfg->max_det > 0 ? cfg->max_det : 0;
    g.xyxy_out = 1;
    return ypp_launch_gather(g, st);
}

int ypp_decode_yolov1(const float* input, int B, int S, int nc, int num_boxes, float input_size, float* out_BS6, void* stream)
{
    if (!input || !out_BS6) { ypp_set_error("decode_yolov1: null pointer"); return YPP_E_BADARG; }
    if (B < 1 || S < 1 || nc < 1 || num_boxes < 1 || (int64_t)B * S * S > 0x7fffffffLL) {
        ypp_set_error("decode_yolov1: B=%d S=%d nc=%d num_boxes=%d out of range", B, S, nc, num_boxes); return YPP_E_BADARG;
    }
    return ypp_launch_decode_yolov1(input, B, S, nc, num_boxes, input_size, out_BS6, (cudaStream_t)stream);
}

int ypp_target_rows(const float* annot_BT5, int B, int T, float input_size, float img_idx_base, float* out_G7,
                    int* gt_count, void* stream)
{
    if (!annot_BT5 || !out_G7 || !gt_count || B <= 0 || T <= 0) { ypp_set_error("target_rows: bad argument"); return YPP_E_BADARG; }
    return ypp_launch_target_rows(annot_BT5, B, T, input_size, img_idx_base, out_G7, gt_count, (long long)B * T, nullptr,
                                  nullptr, nullptr, 0u, (cudaStream_t)stream);
}

int ypp_target_rows_append(const float* annot_BT5, int B, int T, float input_size, float img_idx_base, float* out_G7,
                           int64_t capacity, int64_t* cursor, void* stream, const ypp_probe* probe)
{
    if (!annot_BT5 || !out_G7 || !cursor || B <= 0 || T <= 0 || capacity < 0) { ypp_set_error("target_rows_append: bad argument"); return YPP_E_BADARG; }
    return ypp_launch_target_rows(annot_BT5, B, T, input_size, img_idx_base, out_G7, nullptr, capacity, (const long long*)cursor,
                                  (long long*)cursor, probe ? (long long*)probe->host_word : nullptr, probe ? probe->seq : 0u,
                                  (cudaStream_t)stream);
}

int ypp_detection_ints(const float* rows, int64_t K, int row_width, int mode, const float* scale_xy, int n_img,
                       float img_idx_base, int* out_K4, void* stream)
{
    if (K < 0 || (row_width != 6 && row_width != 7)) { ypp_set_error("detection_ints: K >= 0 and row_width 6 or 7"); return YPP_E_BADARG; }
    if (mode != YPP_WRITER_PRED_FILE && mode != YPP_WRITER_COCO) { ypp_set_error("detection_ints: unknown mode %d", mode); return YPP_E_BADARG; }
    if (K == 0) return YPP_OK;
    if (!rows || !out_K4) { ypp_set_error("detection_ints: null pointer"); return YPP_E_BADARG; }
    if (((uintptr_t)out_K4) & 15u) { ypp_set_error("detection_ints: out must be 16-byte aligned"); return YPP_E_BADARG; }
    if (mode == YPP_WRITER_COCO && (!scale_xy || n_img <= 0)) { ypp_set_error("detection_ints: COCO mode needs scale_xy[n_img,2]"); return YPP_E_BADARG; }
    if ((K + 255) / 256 > 0x7fffffffLL) { ypp_set_error("detection_ints: K too large"); return YPP_E_UNSUPPORTED; }
    return ypp_launch_detection_ints(rows, (long long)K, row_width, mode, scale_xy, n_img, img_idx_base, out_K4, (cudaStream_t)stream);
}

int ypp_confusion_update(const float* det_K6, const int* det_off, const float* lab_G5, const int* lab_off, int n_img,
                         int64_t K, int nc, float conf_thr, float iou_thr, void* scratch, size_t scratch_bytes,
                         int64_t* matrix, void* stream)
{
    if (n_img < 0 || K < 0 || nc < 1) { ypp_set_error("confusion_update: bad size"); return YPP_E_BADARG; }
    if (n_img == 0) return YPP_OK;
    if (!det_off || !lab_off || !matrix) { ypp_set_error("confusion_update: null pointer"); return YPP_E_BADARG; }
    if (K > 0 && (!det_K6 || !scratch || scratch_bytes < (size_t)K * 8)) {
        ypp_set_error("confusion_update: scratch %zu bytes < %lld", scratch_bytes, (long long)K * 8); return YPP_E_WORKSPACE;
    }
    return ypp_launch_confusion(det_K6, det_off, lab_G5, lab_off, n_img, nc, conf_thr, iou_thr, (int*)scratch,
                                (float*)scratch + K, (unsigned long long*)matrix, (cudaStream_t)stream);
}

int ypp_box_iou(const float* a_N4, int N, const float* b_M4, int M, float* out_NM, void* stream)
{
    if (N < 0 || M < 0) { ypp_set_error("box_iou: negative size"); return YPP_E_BADARG; }
    if (N == 0 || M == 0) return YPP_OK;
    if (!a_N4 || !b_M4 || !out_NM) { ypp_set_error("box_iou: null pointer"); return YPP_E_BADARG; }
    if ((((uintptr_t)a_N4) | ((uintptr_t)b_M4)) & 15u) { ypp_set_error("box_iou: inputs must be 16-byte aligned"); return YPP_E_BADARG; }
    if ((N + 7) / 8 > 65535) { ypp_set_error("box_iou: N too large"); return YPP_E_UNSUPPORTED; }
    return ypp_launch_box_iou(a_N4, N, b_M4, M, out_NM, (cudaStream_t)stream);
}

int ypp_bbox_iou(const float* box1, int n1, const float* box2, int n2, int xywh, int mode, double eps, float* out, void* stream)
{
    if (n1 < 0 || n2 < 0 || mode < 0 || mode > 3) { ypp_set_error("bbox_iou: bad size or mode"); return YPP_E_BADARG; }
    if (n1 == 0 || n2 == 0) return YPP_OK;
    if (n1 != n2 && n1 != 1 && n2 != 1) { ypp_set_error("bbox_iou: %d and %d boxes do not broadcast", n1, n2); return YPP_E_BADARG; }
    if (!box1 || !box2 || !out) { ypp_set_error("bbox_iou: null pointer"); return YPP_E_BADARG; }
    if ((((uintptr_t)box1) | ((uintptr_t)box2)) & 15u) { ypp_set_error("bbox_iou: inputs must be 16-byte aligned"); return YPP_E_BADARG; }
    // eps and 1 + eps are Python floats in the reference: rounded to fp32 once when they meet an fp32 tensor
    return ypp_launch_bbox_iou(box1, n1, box2, n2, xywh ? 1 : 0, mode, (float)eps, (float)(1.0 + eps), out, (cudaStream_t)stream);
}

size_t ypp_map_workspace_bytes(int64_t G, int64_t K, int nc) { return ypp_map_ws_bytes(G, K, nc); }

int ypp_map_match(const float* gt_G7, int64_t G, const float* det_K7, int64_t K, int nc, float iou_thr, void* ws,
                  size_t ws_bytes, unsigned char* tp, float* per_class, float* map, void* stream)
{
    if (G < 0 || K < 0 || nc <= 0 || !map || !ws) { ypp_set_error("map_match: bad argument"); return YPP_E_BADARG; }
    if ((G > 0 && !gt_G7) || (K > 0 && !det_K7)) { ypp_set_error("map_match: null rows"); return YPP_E_BADARG; }
    if (nc > 65535 || K >= (1LL << 31) || G >= (1LL << 31)) { ypp_set_error("map_match: size out of range"); return YPP_E_UNSUPPORTED; }
    return ypp_launch_map(gt_G7, G, det_K7, K, nc, iou_thr, ws, ws_bytes, tp, per_class, map, (cudaStream_t)stream);
}

int ypp_map_flags(const float* gt_G7, int64_t G, const float* det_K7, int64_t K, int nc, float iou_thr, void* ws,
                  size_t ws_bytes, unsigned char* tp, float* triples, int64_t* gt_per_class, void* stream)
{
    if (G < 0 || K < 0 || nc <= 0 || !ws) { ypp_set_error("map_flags: bad argument"); return YPP_E_BADARG; }
    if ((G > 0 && !gt_G7) || (K > 0 && !det_K7)) { ypp_set_error("map_flags: null rows"); return YPP_E_BADARG; }
    if (nc > 65535 || K >= (1LL << 31) || G >= (1LL << 31)) { ypp_set_error("map_flags: size out of range"); return YPP_E_UNSUPPORTED; }
    return ypp_launch_map_flags(gt_G7, G, det_K7, K, nc, iou_thr, ws, ws_bytes, tp, triples, (long long*)gt_per_class,
                                (cudaStream_t)stream);
}

int ypp_map_ap_from_flags(const float* triples, int64_t n_seg, int64_t seg_stride, const int64_t* seg_count,
                          const int64_t* gt_per_class, int nc, int c_lo, int c_hi, void* ws, size_t ws_bytes,
                          float* per_class, void* stream)
{
    if (n_seg < 0 || seg_stride < 0 || nc <= 0 || c_lo < 0 || c_hi > nc || c_lo > c_hi || !ws || !gt_per_class || !per_class) {
        ypp_set_error("map_ap_from_flags: bad argument"); return YPP_E_BADARG;
    }
    const int64_t n = n_seg * seg_stride;
    if (n > 0 && !triples) { ypp_set_error("map_ap_from_flags: null rows"); return YPP_E_BADARG; }
    if (nc > 65535 || n >= (1LL << 31)) { ypp_set_error("map_ap_from_flags: size out of range"); return YPP_E_UNSUPPORTED; }
    if (c_lo == c_hi) return YPP_OK;
    return ypp_launch_map_ap_flags(triples, n, seg_stride > 0 ? seg_stride : 1, (const long long*)seg_count,
                                   (const long long*)gt_per_class, nc, c_lo, c_hi, ws, ws_bytes, per_class, (cudaStream_t)stream);
}

int ypp_map_mean(const float* per_class_nc4, int nc, float* map, void* stream)
{
    if (!per_class_nc4 || !map || nc <= 0) { ypp_set_error("map_mean: bad argument"); return YPP_E_BADARG; }
    return ypp_launch_map_mean(per_class_nc4, nc, map, (cudaStream_t)stream);
}

// ---- batches in flight: the stream choreography of a validation loop, in one call per batch ----------------------
struct ypp_pipeline {
    int depth, device;
    cudaStream_t stream[YPP_PIPELINE_MAX_DEPTH];
    cudaEvent_t ready[YPP_PIPELINE_MAX_DEPTH], done[YPP_PIPELINE_MAX_DEPTH];
    long long submitted;                      // tickets 0 .. submitted-1 exist; ticket t ran in slot t % depth
};

int ypp_pipeline_create(int depth, ypp_pipeline** out)
{
    if (!out || depth < 1 || depth > YPP_PIPELINE_MAX_DEPTH) { ypp_set_error("pipeline_create: depth must be 1..%d", YPP_PIPELINE_MAX_DEPTH); return YPP_E_BADARG; }
    ypp_pipeline* p = new (std::nothrow) ypp_pipeline();
    if (!p) { ypp_set_error("pipeline_create: out of host memory"); return YPP_E_CUDA; }
    p->depth = depth; p->submitted = 0;
    cudaError_t e = cudaGetDevice(&p->device);
    for (int i = 0; i < depth && e == cudaSuccess; ++i) {
        e = cudaStreamCreateWithFlags(&p->stream[i], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ready[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->done[i], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) { ypp_set_error("pipeline_create: %s", cudaGetErrorString(e)); delete p; return YPP_E_CUDA; }
    *out = p;
    return YPP_OK;
}

int ypp_pipeline_destroy(ypp_pipeline* p)
{
    if (!p) return YPP_OK;
    for (int i = 0; i < p->depth; ++i) {      // destruction of a busy stream / event is deferred by the runtime
        cudaStreamDestroy(p->stream[i]); cudaEventDestroy(p->ready[i]); cudaEventDestroy(p->done[i]);
    }
    delete p;
    return YPP_OK;
}

int ypp_pipeline_next_slot(const ypp_pipeline* p) { return p ? (int)(p->submitted % p->depth) : -1; }

void* ypp_pipeline_stream(const ypp_pipeline* p, int slot) { return (p && slot >= 0 && slot < p->depth) ? (void*)p->stream[slot] : nullptr; }

int ypp_pipeline_submit(ypp_pipeline* pl, const ypp_cfg* cfg, const ypp_scale* scales, void* ws, size_t ws_bytes,
                        int* cand_count, int* keep_count, float* det_rows, int row_width, int64_t det_capacity,
                        const int64_t* cursor_in, int64_t* cursor_out, const ypp_probe* probe, float img_idx_base,
                        void* caller_stream, int64_t* ticket)
{
    if (!pl) { ypp_set_error("pipeline_submit: null pipeline"); return YPP_E_BADARG; }
    YppDecodeParams p;
    int rc = build_decode_params(cfg, scales, false, &p);
    if (rc) return rc;
    if (!det_rows || !cursor_in || !cursor_out || cursor_in == cursor_out) {
        ypp_set_error("pipeline_submit: det_rows and two different cursor words are required"); return YPP_E_BADARG;
    }
    if (row_width != 6 && row_width != 7) { ypp_set_error("row_width must be 6 or 7"); return YPP_E_BADARG; }
    rc = bind_ws(cfg->B, p.N, cfg->nc, ws, ws_bytes, &p.ws);
    if (rc) return rc;
    const int slot = (int)(pl->submitted % pl->depth);
    cudaStream_t st = pl->stream[slot];
    // the side stream starts behind everything the caller's stream has queued: the heads, the accumulator's state
    cudaError_t e = cudaEventRecord(pl->ready[slot], (cudaStream_t)caller_stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(st, pl->ready[slot], 0);
    if (e != cudaSuccess) { ypp_set_error("pipeline_submit: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
    rc = ypp_launch_decode_compact(p, cfg->B, st);
    if (rc) return rc;
    rc = run_sort_nms(cfg, p.N, p.ws, nullptr, nullptr, nullptr, nullptr, row_width, 0, nullptr, 0.f, nullptr, true, true, st,
                      nullptr, nullptr, false);
    if (rc) return rc;
    // the rows go behind the previous batch's rows: its gather wrote the cursor word this one reads
    if (pl->submitted > 0 && pl->depth > 1) {
        e = cudaStreamWaitEvent(st, pl->done[(int)((pl->submitted - 1) % pl->depth)], 0);
        if (e != cudaSuccess) { ypp_set_error("pipeline_submit: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
    }
    rc = run_sort_nms(cfg, p.N, p.ws, cand_count, keep_count, nullptr, det_rows, row_width, det_capacity, nullptr, img_idx_base,
                      nullptr, false, false, st, (const long long*)cursor_in, (long long*)cursor_out, true, probe);
    if (rc) return rc;
    e = cudaEventRecord(pl->done[slot], st);
    if (e != cudaSuccess) { ypp_set_error("pipeline_submit: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
    if (ticket) *ticket = pl->submitted;
    pl->submitted += 1;
    return YPP_OK;
}

// 1: batch `ticket` has completed (rows, counts and cursor probe included); 0: not known to have
int ypp_pipeline_query(const ypp_pipeline* p, int64_t ticket)
{
    if (!p || ticket < 0 || ticket >= p->submitted) return 0;
    // the slot's event belongs to the newest ticket of the slot; that one completes after `ticket` (same stream)
    return cudaEventQuery(p->done[(int)(ticket % p->depth)]) == cudaSuccess ? 1 : 0;
}

// the HOST waits for batch `ticket` (newer batches stay in flight)
int ypp_pipeline_wait_ticket(const ypp_pipeline* p, int64_t ticket)
{
    if (!p || ticket < 0 || ticket >= p->submitted) { ypp_set_error("pipeline_wait_ticket: no such ticket"); return YPP_E_BADARG; }
    // a reused slot's event belongs to a newer batch on the same stream: waiting for that one covers `ticket` too
    const cudaError_t e = cudaEventSynchronize(p->done[(int)(ticket % p->depth)]);
    if (e != cudaSuccess) { ypp_set_error("pipeline_wait_ticket: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
    return YPP_OK;
}

// `stream` waits ON THE DEVICE for every batch submitted so far (batches complete in order: the last one's event)
int ypp_pipeline_join(const ypp_pipeline* p, void* stream)
{
    if (!p) { ypp_set_error("pipeline_join: null pipeline"); return YPP_E_BADARG; }
    if (p->submitted == 0) return YPP_OK;
    const cudaError_t e = cudaStreamWaitEvent((cudaStream_t)stream, p->done[(int)((p->submitted - 1) % p->depth)], 0);
    if (e != cudaSuccess) { ypp_set_error("pipeline_join: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
    return YPP_OK;
}

int ypp_peek_async(const void* device_src, void* host_dst, size_t bytes, void* stream)
{
    if (!device_src || !host_dst || bytes == 0 || bytes > 4096) { ypp_set_error("peek_async: bad argument"); return YPP_E_BADARG; }
    const cudaError_t e = cudaMemcpyAsync(host_dst, device_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
    if (e != cudaSuccess) { ypp_set_error("peek_async: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
    return YPP_OK;
}

int ypp_adaptive_launch(int enable) { return ypp_adaptive_launch_set(enable); }

const char* ypp_last_error(void) { return g_err; }
int ypp_version(void) { return 100; }
int64_t ypp_launch_count(int reset)
{
    const long long v = g_launches;
    if (reset) g_launches = 0;
    return v;
}

}  // extern "C"
