// Device helpers shared by the decode kernels (decode.cu) and the fused head kernel (head.cu):
// class arg-max tracking, the decode of one cell, the per-slot compaction, mbarrier / bulk-copy PTX.
#pragma once
#include "ypp_kernels.cuh"

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

// YPP_SCORE_OBJ_TIMES_CLS: the class is the first maximum of the PRODUCTS conf_j = sigmoid(cls_j) * sigmoid(obj)
// (utils/general.py:669 `x[:, 5:] *= x[:, 4:5]`, :679 `conf, j = x[:, 5:].max(1)`).  The fp32 product is monotone in
// sigmoid(cls_j), so its maximum P belongs to the largest sigmoid, but a slightly smaller sigmoid at a LOWER index
// may round to the same product and then wins the arg-max.  Only logits inside `band` of the largest can do that.
__device__ __noinline__ int cls_resolve_product(const float* cls_base, int stride, int nc, float m1, float band, float sobj,
                                                float P, int fallback, int q, int hwm)
{
    for (int c = 0; c < nc; ++c) {
        const float v = cls_base[(size_t)c * stride + ((q + c * hwm) & 3)];
        if (v >= m1 - band && __fmul_rn(ypp_sigmoid(v), sobj) == P) return c;
    }
    return fallback;
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
    d.score = sobj;
    if (p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) {
        d.score = __fmul_rn(ypp_sigmoid(tr.m1), sobj);
        // a sigmoid within 2^-21 relative of the largest is the most that can still round to the same product
        const float band = 8.0f * cls_tol(tr.m1);
        if (p.nc > 1 && !(tr.m1 - tr.m2 > band))
            ci = cls_resolve_product(cls_base, cls_stride, p.nc, tr.m1, band, sobj, d.score, ci, cls_q, cls_hwm);
    }
    d.cls = ci;
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
#ifndef YPP_MBAR_SUSPEND_NS
#define YPP_MBAR_SUSPEND_NS 20000u      /* upper bound of one suspension; the barrier's completion ends it at once */
#endif
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "YPP_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@p bra YPP_DONE;\n"
        "bra YPP_WAIT;\n"
        "YPP_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity), "r"(YPP_MBAR_SUSPEND_NS) : "memory");
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

__device__ __forceinline__ void mbar_arrive(u64* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

}  // namespace
