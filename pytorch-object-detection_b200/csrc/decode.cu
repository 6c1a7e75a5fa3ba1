// D1: fused head decode + confidence threshold + stream compaction (and the dense decode).
//
// Reference semantics: decode_predictions, utils/yolov3_utils.py:12-59 and
// utils/yolov2_utils.py:12-58; the threshold is the torch.where of nms_v3,
// utils/yolo_utils.py:334.
//
// One pass over the NCHW head by persistent CTAs.  A tile = up to 256 consecutive cells of one anchor
// plane group: its C = 5+nc channel rows (256 floats each, contiguous in the NCHW tensor) are
// pulled into shared memory by the async-copy engine (cp.async.bulk, one bulk copy per channel
// row, completion on an mbarrier), so the whole tile is in flight at once without holding
// registers, and DRAM sees only large sequential requests.  Thread t then decodes cell t from
// shared memory (conflict-free columns).  Heads whose plane size is not a multiple of 4 floats
// (13x13, 19x19: channel rows are only 4-byte aligned) are staged the same way: each row copy fetches
// the enclosing span of whole 16-byte words and the consumers add the row's 0..3 float shift.
//
// Compaction is deterministic: the candidates of every 32-box slot are written in box-index order
// at keys[b][slot_first ..] with their count in tile_cnt[b][slot]; S1 concatenates the slots, so
// the sort input is already in box-index order and a STABLE sort by score alone gives the
// reference's tie order.  No atomics, nothing to clear between calls.
#include "decode_common.cuh"
#include <cstdlib>

namespace {

// Warp-specialised persistent CTAs: 8 consumer warps + 1 producer warp, no block barrier in the
// loop.  The producer walks the CTA's (image, tile) work items and keeps YPP_D1_STAGES tiles in
// flight (cp.async.bulk rows -> stage, completion on full[stage]); consumer warp w decodes cells
// 32w..32w+31 of each tile straight out of the stage, releases it on empty[stage] and compacts its
// own slot, so warps never wait for each other -- only for data.
#ifndef YPP_D1_STAGES
#define YPP_D1_STAGES 2
#endif
#define YPP_D1_CONSUMERS (YPP_TILE_BOXES / 32)
#define YPP_SHIFT_PITCH (YPP_TILE_BOXES + 8)   /* row pitch of a staged tile with shifted rows (compile-time: LDS offsets fold) */
#define YPP_D1_MAX_PRODUCERS 4            /* UBLKCP issue is serialised per warp: spread the rows of a tile */
#define YPP_D1_THREADS (32 * (YPP_D1_CONSUMERS + YPP_D1_MAX_PRODUCERS))


__global__ void __launch_bounds__(YPP_D1_THREADS, 3)
ypp_decode_compact_kernel(const __grid_constant__ YppDecodeParams p, const int B)
{
    extern __shared__ __align__(128) float stage_mem[];      // [STAGES][C][256] when some scale is staged
    __shared__ __align__(8) u64 full[YPP_D1_STAGES], empty[YPP_D1_STAGES];
    ypp_grid_wait();       // the previous call's G may still be reading the workspace
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tiles_per_img = p.tile_base[p.n_scales];
    const long long total = (long long)tiles_per_img * B;
    const int stage_floats = p.C * p.pitch;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < YPP_D1_STAGES; ++i) { mbar_init(&full[i], (int)(blockDim.x >> 5) - YPP_D1_CONSUMERS); mbar_init(&empty[i], YPP_D1_CONSUMERS); }
    }
    __syncthreads();
    int b = (int)(blockIdx.x / (unsigned)tiles_per_img), tile = (int)(blockIdx.x - (unsigned)b * (unsigned)tiles_per_img);
    const int gb = (int)(gridDim.x / (unsigned)tiles_per_img), gt = (int)(gridDim.x - (unsigned)gb * (unsigned)tiles_per_img);
    int k = 0;
    if (warp >= YPP_D1_CONSUMERS) {
        // ---------------- producers: warp pw copies channel rows [c_lo, c_hi) of every tile
        const int P = (int)(blockDim.x >> 5) - YPP_D1_CONSUMERS, pw = warp - YPP_D1_CONSUMERS;
        const int per = (p.C + P - 1) / P;
        const int c_lo = min(p.C, pw * per), c_hi = min(p.C, c_lo + per);
        for (long long w = blockIdx.x; w < total; w += gridDim.x, ++k, b += gb, tile += gt) {
            if (tile >= tiles_per_img) { tile -= tiles_per_img; ++b; }
            const int stg = k % YPP_D1_STAGES, use = k / YPP_D1_STAGES;
            if (use > 0) mbar_wait(&empty[stg], (u32)(use - 1) & 1u);        // consumers are done with this stage
            const TileRef t = tile_ref(p, tile);
            const YppScaleParams& s = p.sc[t.si];
            if (s.tma && c_hi > c_lo && (s.HW & 3) == 0) {
                // 16-byte-aligned rows: one copy of ncells floats per row
                const float* src = s.head + ((size_t)(b * s.A + t.a) * p.C) * s.HW + t.cell0;
                const u32 row_bytes = (u32)t.ncells * 4u;                    // multiple of 16: HW % 4 == 0
                float* stage = stage_mem + stg * stage_floats;
                if (lane == 0) mbar_expect_tx(&full[stg], row_bytes * (u32)(c_hi - c_lo));
                __syncwarp();
                for (int c = c_lo + lane; c < c_hi; c += 32)
                    bulk_g2s(stage + c * YPP_TILE_BOXES, src + (size_t)c * s.HW, row_bytes, &full[stg]);
            } else if (s.tma && c_hi > c_lo) {
                // row c of the tile = floats [e0, e0 + ncells) of the head; the copy fetches the enclosing span of
                // whole 16-byte words [a0, a0 + n_al), so cell ct of row c lands at stage[c * 264 + (e0 & 3) + ct]
                const long long e_tile = ((long long)(b * s.A + t.a) * p.C) * s.HW + t.cell0;
                float* stage = stage_mem + stg * stage_floats;
                u32 bytes = 0;
                for (int c = c_lo + lane; c < c_hi; c += 32) {
                    const long long e0 = e_tile + (long long)c * s.HW;
                    const int sh = (int)(e0 & 3);
                    const long long a0 = e0 - sh;
                    const int n_el = t.ncells + sh;
                    int n_al = (n_el + 3) & ~3;
                    if (a0 + n_al > s.elems) {                               // last words of the tensor: never read past it
                        n_al = (int)((s.elems - a0) & ~3LL);
                        for (int j = n_al; j < n_el; ++j) stage[c * YPP_SHIFT_PITCH + j] = __ldg(s.head + a0 + j);
                    }
                    bytes += (u32)n_al * 4u;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) bytes += __shfl_xor_sync(YPP_FULL_MASK, bytes, o);
                __syncwarp();                                                // the tail stores above are ordered before ..
                if (lane == 0) mbar_expect_tx(&full[stg], bytes);            // .. this arrive (release) publishes them
                __syncwarp();
                for (int c = c_lo + lane; c < c_hi; c += 32) {
                    const long long e0 = e_tile + (long long)c * s.HW;
                    const long long a0 = e0 & ~3LL;
                    int n_al = (t.ncells + (int)(e0 & 3) + 3) & ~3;
                    if (a0 + n_al > s.elems) n_al = (int)((s.elems - a0) & ~3LL);
                    if (n_al > 0) bulk_g2s(stage + c * YPP_SHIFT_PITCH, s.head + a0, (u32)n_al * 4u, &full[stg]);
                }
            } else if (lane == 0) {
                mbar_arrive(&full[stg]);                                     // nothing to stage: consumers read the head directly
            }
        }
        return;
    }
    // ---------------- consumers
    const int nslots = tiles_per_img * YPP_D1_CONSUMERS;
    for (long long w = blockIdx.x; w < total; w += gridDim.x, ++k, b += gb, tile += gt) {
        if (tile >= tiles_per_img) { tile -= tiles_per_img; ++b; }
        const int stg = k % YPP_D1_STAGES, use = k / YPP_D1_STAGES;
        const TileRef t = tile_ref(p, tile);
        const YppScaleParams& s = p.sc[t.si];
        const int ct = tid;                                                  // cell inside the tile (warp w: 32w..32w+31)
        Decoded d;
        d.box = make_float4(0.f, 0.f, 0.f, 0.f); d.score = 0.f; d.cls = 0;
        bool pass;
        mbar_wait(&full[stg], (u32)use & 1u);
        const int hwm = s.HW & 3;
        if (s.tma && hwm == 0) {
            pass = decode_one<1>(p, s, t.a, t.cell0 + ct, ct < t.ncells, stage_mem + stg * stage_floats + ct, YPP_TILE_BOXES, 0, 0, d);
        } else if (s.tma) {
            // (first float of row 0 of this tile) mod 4; unsigned wrap-around keeps the low two bits
            const int q = (int)((((unsigned)(b * s.A + t.a) * (unsigned)p.C) * (unsigned)s.HW + (unsigned)t.cell0) & 3u);
            pass = decode_one<2>(p, s, t.a, t.cell0 + ct, ct < t.ncells, stage_mem + stg * stage_floats + ct, YPP_SHIFT_PITCH, q, hwm, d);
        } else {
            const float* src = s.head + ((size_t)(b * s.A + t.a) * p.C) * s.HW + t.cell0;
            pass = decode_one<0>(p, s, t.a, t.cell0 + ct, ct < t.ncells, src + ct, s.HW, 0, 0, d);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[stg]);                             // this warp no longer reads the stage
        slot_store(p.ws, p.N, b, tile * YPP_D1_CONSUMERS + warp, nslots, t.first + 32 * warp, pass,
                   d.box, d.score, d.cls, t.first + ct);
    }
}

// ---- dense decode: every box, rows [B, N, 6] (= decode_predictions + cat) ---------------
__global__ void __launch_bounds__(YPP_DECODE_THREADS)
ypp_decode_dense_kernel(const __grid_constant__ YppDecodeParams p, float* __restrict__ out)
{
    const int b = blockIdx.y, tid = threadIdx.x;
    const TileRef t = tile_ref(p, (int)blockIdx.x);
    if (tid >= t.ncells) return;
    const YppScaleParams& s = p.sc[t.si];
    const size_t HW = s.HW;
    const float* col = s.head + ((size_t)(b * s.A + t.a) * p.C) * HW + t.cell0 + tid;
    const float obj = __ldg(col + 4 * HW);
    const float t0 = __ldg(col), t1 = __ldg(col + HW), t2 = __ldg(col + 2 * HW), t3 = __ldg(col + 3 * HW);
    ClsTrack tr;
    if (p.nc > 1 || p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) class_sweep([&](int c) { return __ldg(col + (5 + c) * HW); }, p.nc, tr);
    else cls_init(tr);
    const Decoded d = decode_cell(p, s, t.a, t.cell0 + tid, t0, t1, t2, t3, ypp_sigmoid(obj), tr, col + 5 * HW, (int)HW);
    float2* o = reinterpret_cast<float2*>(out + ((size_t)b * p.N + t.first + tid) * 6);
    o[0] = make_float2(d.box.x, d.box.y);
    o[1] = make_float2(d.box.z, d.box.w);
    o[2] = make_float2(d.score, (float)d.cls);
}

// ---- compaction from decoded rows (nms_v3 entry, utils/yolo_utils.py:321-337) -----------
__global__ void __launch_bounds__(YPP_DECODE_THREADS)
ypp_rows_compact_kernel(const float* __restrict__ rows, int N, float conf, YppWs ws)
{
    ypp_grid_wait();
    const int b = blockIdx.y;
    const int n = blockIdx.x * YPP_DECODE_THREADS + threadIdx.x;
    const int slot = n >> 5;
    const int nslots = (int)gridDim.x * (YPP_DECODE_THREADS / 32);
    bool pass = false;
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
    float score = 0.f;
    int cls = 0;
    if (n < N) {
        const float2* r = reinterpret_cast<const float2*>(rows + ((size_t)b * N + n) * 6);
        const float2 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
        box = make_float4(r0.x, r0.y, r1.x, r1.y);
        score = r2.x;
        int c = (int)r2.y;
        // class ids that the 12-bit key field cannot hold are clamped (documented: ids must be
        // integers in [0, nc)); the row keeps its own class column verbatim in the output
        if (!(r2.y >= 0.0f && r2.y <= (float)YPP_KEY_CLS_MASK)) c = min(max(c, 0), (int)YPP_KEY_CLS_MASK);
        cls = c;
        pass = score > conf;
    }
    slot_store(ws, N, b, slot, nslots, slot << 5, pass, box, score, cls, n);
}

// ---- compaction from Ultralytics-style predictions [B, n, 5+nc] (utils/general.py:640-684) ----
// One thread per row of the workspace: a box (best class only) or a (box, class) pair (multi_label;
// row = box * nc + j, which is the order nonzero() lists the pairs in).  A row's box is stored under
// its own row number, so the sort / NMS / gather kernels run unchanged on n * nc "boxes".
template <bool kMulti>
__global__ void __launch_bounds__(YPP_DECODE_THREADS)
ypp_pred_compact_kernel(const float* __restrict__ pred, int n_boxes, int nc, float conf,
                        const unsigned char* __restrict__ allow, YppWs ws, int n_rows)
{
    ypp_grid_wait();
    const int b = blockIdx.y;
    const int n = blockIdx.x * YPP_DECODE_THREADS + threadIdx.x;
    const int slot = n >> 5;
    const int nslots = (int)gridDim.x * (YPP_DECODE_THREADS / 32);
    const int C = 5 + nc;
    bool pass = false;
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
    float score = 0.f;
    int cls = 0;
    if (n < n_rows) {
        const int i = kMulti ? n / nc : n;
        const float* r = pred + ((size_t)b * n_boxes + i) * C;
        const float obj = __ldg(r + 4);
        if (obj > conf) {                                        // xc = prediction[..., 4] > conf_thres
            if (kMulti) {
                cls = n - i * nc;
                score = __fmul_rn(__ldg(r + 5 + cls), obj);      // x[:, 5:] *= x[:, 4:5]
            } else {
                // x[:, 5:].max(1): the first maximum; a NaN is maximal (and then fails the threshold)
                float best = __fmul_rn(__ldg(r + 5), obj);
                for (int c = 1; c < nc; ++c) {
                    const float v = __fmul_rn(__ldg(r + 5 + c), obj);
                    if (v > best || (v != v && best == best)) { best = v; cls = c; }
                }
                score = best;
            }
            pass = score > conf && (allow == nullptr || allow[cls] != 0);
            if (pass) box = make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
        }
    }
    slot_store(ws, n_rows, b, slot, nslots, slot << 5, pass, box, score, cls, n);
}

// ---- YOLO v1 decode (utils/yolov1_utils.py:12-66): one thread per grid cell ------------------
__global__ void __launch_bounds__(128)
ypp_decode_yolov1_kernel(const float* __restrict__ in, int cells_total, int S, int nc, int nb, float stride_w, float stride_h,
                         float* __restrict__ out)
{
    const int t = blockIdx.x * 128 + threadIdx.x;
    if (t >= cells_total) return;
    const int cell = t % (S * S), gy = cell / S, gx = cell - gy * S;
    const float* r = in + (size_t)t * (nc + 5 * nb);
    // best predictor: torch.argmax over the confidences (first maximum, NaN maximal)
    int k = 0;
    float best = ypp_sigmoid(__ldg(r + nc));
    for (int q = 1; q < nb; ++q) {
        const float v = ypp_sigmoid(__ldg(r + nc + 5 * q));
        if (v > best || (v != v && best == best)) { best = v; k = q; }
    }
    // sum over predictors of one_hot * value (:43-48): the selected value plus zeros, NaN where any is NaN
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    for (int q = 0; q < nb; ++q) {
        const float sel = q == k ? 1.0f : 0.0f;
#pragma unroll
        for (int e = 0; e < 5; ++e) acc[e] = __fadd_rn(acc[e], __fmul_rn(sel, ypp_sigmoid(__ldg(r + nc + 5 * q + e))));
    }
    // class: argmax of the sigmoids (saturation can tie values whose logits differ)
    int cls = 0;
    float cb = ypp_sigmoid(__ldg(r));
    for (int c = 1; c < nc; ++c) {
        const float v = ypp_sigmoid(__ldg(r + c));
        if (v > cb || (v != v && cb == cb)) { cb = v; cls = c; }
    }
    float* o = out + (size_t)t * 6;
    o[0] = __fmul_rn(__fadd_rn(acc[1], (float)gx), stride_w);
    o[1] = __fmul_rn(__fadd_rn(acc[2], (float)gy), stride_h);
    o[2] = __fmul_rn(__fmul_rn(acc[3], (float)S), stride_w);
    o[3] = __fmul_rn(__fmul_rn(acc[4], (float)S), stride_h);
    o[4] = acc[0];
    o[5] = (float)cls;
}

}  // namespace

// ---------------------------------------------------------------- host launchers

int ypp_launch_decode_compact(const YppDecodeParams& p, int B, cudaStream_t st)
{
    // function attributes and occupancy are per device: cache them per device (one process may drive several)
    static int configured_d[YPP_MAX_DEVICES], ctas_d[YPP_MAX_DEVICES], sms_d[YPP_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= YPP_MAX_DEVICES) { ypp_set_error("decode_compact: device ordinal %d out of range", dev); return YPP_E_UNSUPPORTED; }
    int& configured = configured_d[dev];
    int& ctas_per_sm = ctas_d[dev];
    int& sms = sms_d[dev];
    const int smem = YPP_D1_STAGES * p.stage_bytes;
    // one producer warp per ~20 channel rows of a tile (env YPP_D1_PRODUCERS overrides, for tuning)
    static int producers_env = -1;
    if (producers_env < 0) { const char* e = getenv("YPP_D1_PRODUCERS"); producers_env = e ? atoi(e) : 0; }
    int P = producers_env > 0 ? producers_env : (p.C + 19) / 20;
    P = P < 1 ? 1 : (P > YPP_D1_MAX_PRODUCERS ? YPP_D1_MAX_PRODUCERS : P);
    const int threads = 32 * (YPP_D1_CONSUMERS + P);
    const int cfg_key = smem * 8 + P + 1;                 // never 0 (the zero-initialised cache)
    if (cfg_key != configured) {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(ypp_decode_compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) { ypp_set_error("decode_compact: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
        }
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, ypp_decode_compact_kernel, threads, smem);
        if (e != cudaSuccess || ctas_per_sm < 1) { ypp_set_error("decode_compact: occupancy query failed"); return YPP_E_CUDA; }
        configured = cfg_key;
    }
    const long long total = (long long)p.tile_base[p.n_scales] * B;
    long long grid = (long long)sms * ctas_per_sm;
    if (grid > total) grid = total;
    ypp_launch_chain(0, ypp_decode_compact_kernel, dim3((unsigned)grid), dim3(threads), smem, st, p, B);
    ypp_count_launch(1);
    return ypp_check_launch("decode_compact");
}

int ypp_launch_decode_dense(const YppDecodeParams& p, int B, float* out, cudaStream_t st)
{
    dim3 grid(p.tile_base[p.n_scales], B);
    ypp_decode_dense_kernel<<<grid, YPP_DECODE_THREADS, 0, st>>>(p, out);
    ypp_count_launch(1);
    return ypp_check_launch("decode_dense");
}

int ypp_launch_rows_compact(const float* rows, int B, int N, float conf, const YppWs& ws, cudaStream_t st)
{
    dim3 grid((N + YPP_TILE_BOXES - 1) / YPP_TILE_BOXES, B);
    ypp_launch_chain(0, ypp_rows_compact_kernel, grid, dim3(YPP_DECODE_THREADS), 0, st, rows, N, conf, ws);
    ypp_count_launch(1);
    return ypp_check_launch("rows_compact");
}

int ypp_launch_pred_compact(const float* pred, int B, int n_boxes, int nc, float conf, int multi_label,
                            const unsigned char* class_allow, const YppWs& ws, int n_rows, cudaStream_t st)
{
    dim3 grid((n_rows + YPP_TILE_BOXES - 1) / YPP_TILE_BOXES, B);
    if (multi_label) ypp_launch_chain(0, ypp_pred_compact_kernel<true>, grid, dim3(YPP_DECODE_THREADS), 0, st, pred, n_boxes, nc, conf, class_allow, ws, n_rows);
    else             ypp_launch_chain(0, ypp_pred_compact_kernel<false>, grid, dim3(YPP_DECODE_THREADS), 0, st, pred, n_boxes, nc, conf, class_allow, ws, n_rows);
    ypp_count_launch(1);
    return ypp_check_launch("pred_compact");
}

int ypp_launch_decode_yolov1(const float* input, int B, int S, int nc, int num_boxes, float input_size, float* out, cudaStream_t st)
{
    const int cells = B * S * S;
    // stride = FloatTensor([input_size / layer_w, input_size / layer_h] * 2): Python doubles rounded to fp32
    const float sw = (float)((double)input_size / (double)S), sh = sw;
    ypp_decode_yolov1_kernel<<<(cells + 127) / 128, 128, 0, st>>>(input, cells, S, nc, num_boxes, sw, sh, out);
    ypp_count_launch(1);
    return ypp_check_launch("decode_yolov1");
}
