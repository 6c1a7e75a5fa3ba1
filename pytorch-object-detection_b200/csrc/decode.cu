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
#include "ypp_kernels.cuh"
#include <cstdlib>

namespace {

// ---- class arg-max over sigmoid(logit) without evaluating nc sigmoids -------------------
// The reference takes argmax AFTER the sigmoid (yolov3_utils.py:51,57), so fp32 collisions
// (saturation, rounding) tie towards the lowest index.  The computed sigmoid is monotone,
// hence the winner is among the logits within `tol` of the largest logit; only when a
// second logit falls inside that band (rare) are sigmoids evaluated.
struct ClsTrack {
    float m1, m2;   // largest / second largest logit seen (m2 counts duplicates of m1)
    int i1;         // first index attaining m1
};

__device__ __forceinline__ void cls_init(ClsTrack& t) { t.m1 = -INFINITY; t.m2 = -INFINITY; t.i1 = 0; }

// branch-free; a NaN logit never wins a > compare and is dropped by fmaxf (documented: NaN
// logits are not supported, torch.argmax would treat them as maximal)
__device__ __forceinline__ void cls_push(ClsTrack& t, float v, int c)
{
    const bool gt = v > t.m1;
    t.m2 = fmaxf(t.m2, gt ? t.m1 : v);
    t.i1 = gt ? c : t.i1;
    t.m1 = gt ? v : t.m1;
}

__device__ __forceinline__ float cls_tol(float m1)
{
    // |d sigmoid| < 1 ulp(sigmoid)  <=>  d logit < 2^-23 (1 + e^m); 8x margin
    return 9.6e-7f * (1.0f + __expf(fmaxf(m1, 0.0f)));
}

// slow path: exact argmax over sigmoid for logits inside the band (first max wins);
// cls_base/stride address either the staged tile or the head itself
__device__ __noinline__ int cls_resolve(const float* cls_base, int stride, int nc, float m1, float tol, int q, int hwm)
{
    int best = 0;
    float best_v = -1.0f;
    for (int c = 0; c < nc; ++c) {
        const float v = cls_base[(size_t)c * stride + ((q + c * hwm) & 3)];      // q = hwm = 0: rows without a shift
        if (v >= m1 - tol) {
            const float s = ypp_sigmoid(v);
            if (s > best_v) { best_v = s; best = c; }
        }
    }
    return best;
}

template <class Load>
__device__ __forceinline__ void class_sweep(Load ld, int nc, ClsTrack& tr)      // ld(c) = logit of class c
{
    cls_init(tr);
    int c = 0;
    for (; c + 4 <= nc; c += 4) {
        const float v0 = ld(c), v1 = ld(c + 1), v2 = ld(c + 2), v3 = ld(c + 3);
        cls_push(tr, v0, c); cls_push(tr, v1, c + 1); cls_push(tr, v2, c + 2); cls_push(tr, v3, c + 3);
    }
    for (; c < nc; ++c) cls_push(tr, ld(c), c);
}

// decode one cell from its five box/objectness logits and the class track
struct Decoded { float4 box; float score; int cls; };

__device__ __forceinline__ Decoded decode_cell(const YppDecodeParams& p, const YppScaleParams& s, int a, int cell,
                                               float t0, float t1, float t2, float t3, float sobj,
                                               const ClsTrack& tr, const float* cls_base, int cls_stride,
                                               int cls_q = 0, int cls_hwm = 0)
{
    // y = cell / W without an integer division: float estimate, then one step of correction
    int y = __float2int_rz(__fmul_rn((float)cell, s.inv_w));
    int x = cell - y * s.W;
    if (x < 0) { --y; x += s.W; }
    if (x >= s.W) { ++y; x -= s.W; }
    Decoded d;
    // pxy = sigmoid(t) + grid, then *= stride (yolov3_utils.py:47,54)
    const float px = __fmul_rn(__fadd_rn(ypp_sigmoid(t0), (float)x), s.sw);
    const float py = __fmul_rn(__fadd_rn(ypp_sigmoid(t1), (float)y), s.sh);
    // pwh = exp(t) * scaled_anchor, then *= stride: two roundings (:48,54)
    const float pw = __fmul_rn(__fmul_rn(expf(t2), s.aw[a]), s.sw);
    const float ph = __fmul_rn(__fmul_rn(expf(t3), s.ah[a]), s.sh);
    d.box = make_float4(px, py, pw, ph);
    int ci = 0;
    if (p.nc > 1) {
        ci = tr.i1;
        const float tol = cls_tol(tr.m1);
        if (!(tr.m1 - tr.m2 > tol)) ci = cls_resolve(cls_base, cls_stride, p.nc, tr.m1, tol, cls_q, cls_hwm);
    }
    d.cls = ci;
    d.score = sobj;
    if (p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) d.score = __fmul_rn(sobj, ypp_sigmoid(tr.m1));
    return d;
}

__device__ __forceinline__ float max_nan(float a, float b)
{
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

// ---- warp-wide compaction in box-index order ------------------------------------------------
// The compaction unit is a SLOT of 32 consecutive boxes handled by one warp (one ballot): its
// candidates go to keys[b][slot_first ..] in lane order, its count and extrema to the slot record.
// S1 concatenates the slots.  No block barrier, no atomics, nothing to clear between calls.
__device__ __forceinline__ void slot_store(const YppWs& ws, int N, int b, int slot, int nslots, int first,
                                           bool pass, const float4 box, const float score, const int cls, int idx)
{
    const int lane = threadIdx.x & 31;
    const u32 m = __ballot_sync(YPP_FULL_MASK, pass);
    u32 emax = 0, enx = 0, eny = 0;
    if (m) {                                       // warp-uniform
        // extrema for the coordinate trick / cross-class detector
        if (pass) {
            const float4 q = ypp_xyxy(box);
            emax = ypp_enc(max_nan(max_nan(q.x, q.y), max_nan(q.z, q.w)));   // a NaN coordinate stays on top, like boxes.max()
            enx = ypp_enc(-q.x);
            eny = ypp_enc(-q.y);
        }
        emax = __reduce_max_sync(YPP_FULL_MASK, emax);
        enx = __reduce_max_sync(YPP_FULL_MASK, enx);
        eny = __reduce_max_sync(YPP_FULL_MASK, eny);
    }
    if (lane == 0) {
        ws.tile_cnt[(size_t)b * ws.tiles_max + slot] = __popc(m);
        *reinterpret_cast<uint4*>(ws.tile_ext + 4 * ((size_t)b * ws.tiles_max + slot)) = make_uint4(emax, enx, eny, 0u);
        if (b == 0) {
            ws.tile_first[slot] = first;
            if (slot == 0) *ws.ntiles = nslots;
        }
    }
    if (pass) {
        const int pos = first + __popc(m & ((1u << lane) - 1u));
        ws.keys[(size_t)b * N + pos] = ypp_make_key(score, idx, cls);
        ws.box[(size_t)b * N + idx] = box;
    }
}

// ---- async bulk copy + mbarrier (sm_90+ PTX) ----------------------------------------------
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(u64* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, u32 bytes, u64* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "YPP_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra YPP_DONE;\n"
        "bra YPP_WAIT;\n"
        "YPP_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// tile -> (scale, anchor, first cell, cells)
struct TileRef { int si, a, cell0, ncells, first; };

__device__ __forceinline__ TileRef tile_ref(const YppDecodeParams& p, int bx)
{
    TileRef t;
    t.si = 0;
#pragma unroll
    for (int k = 1; k < YPP_MAX_SCALES; ++k)
        if (k < p.n_scales && bx >= p.tile_base[k]) t.si = k;
    const YppScaleParams& s = p.sc[t.si];
    int tl = bx - p.tile_base[t.si];
    t.a = 0;
    while (tl >= s.tpp) { tl -= s.tpp; ++t.a; }            // A <= 8 uniform steps instead of a division
    t.cell0 = tl * YPP_TILE_BOXES;
    t.ncells = min(YPP_TILE_BOXES, s.HW - t.cell0);
    t.first = s.box_base + t.a * s.HW + t.cell0;
    return t;
}

// one cell of the head itself (kMode 0: global loads, stride HW), of a staged tile (1: shared-memory column
// `col`, rows `stride` floats apart) or of a staged tile whose rows carry their own 0..3 float shift (2: row c
// starts ((q + c * hwm) & 3) floats into its slot -- planes whose size is not a multiple of 4 floats)
template <int kMode>
__device__ __forceinline__ bool decode_one(const YppDecodeParams& p, const YppScaleParams& s, int a, int cell, bool valid,
                                           const float* col, int stride, int q, int hwm, Decoded& d)
{
    constexpr bool kStaged = kMode != 0, kShift = kMode == 2;
    auto ld = [&](int c) -> float {
        const float* ptr = col + (size_t)c * stride + (kShift ? ((q + c * hwm) & 3) : 0);
        return kStaged ? *ptr : __ldg(ptr);
    };
    bool pass = false;
    float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f, sobj = 0.f;
    if (valid) {
        const float obj = ld(4);
        if (!kStaged) { t0 = ld(0); t1 = ld(1); t2 = ld(2); t3 = ld(3); }
        sobj = ypp_sigmoid(obj);                                         // yolov3_utils.py:50
        pass = sobj > p.conf;                                            // yolo_utils.py:334 (strict)
    }
#ifdef YPP_D1_NOCOMPUTE                                                  /* tuning experiment: copy pipeline only */
    pass = false;
#endif
    // the staged path sweeps the classes warp-wide (shared memory reads are cheap, no divergence);
    // the direct path only in lanes that hold a candidate (every read is a global load)
    if (__any_sync(YPP_FULL_MASK, pass) && (kStaged || pass)) {
        ClsTrack tr;
        if (p.nc > 1 || p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) class_sweep([&](int c) { return ld(5 + c); }, p.nc, tr);
        else cls_init(tr);
        if (pass) {
            if (kStaged) { t0 = ld(0); t1 = ld(1); t2 = ld(2); t3 = ld(3); }
            d = decode_cell(p, s, a, cell, t0, t1, t2, t3, sobj, tr, col + 5 * (size_t)stride, stride,
                            kShift ? ((q + 5 * hwm) & 3) : 0, kShift ? hwm : 0);
            if (p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) pass = d.score > p.conf;
        }
    }
    return pass;
}

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

__device__ __forceinline__ void mbar_arrive(u64* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

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
