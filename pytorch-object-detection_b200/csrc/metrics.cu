// mAP@.5 side of the path: ground-truth ingest, N x M box_iou, and the TP/FP matching + AP.
//
// Reference semantics:
//   get_target_boxes_for_map      utils/yolo_utils.py:682-710
//   box_iou                       utils/metrics.py:233-255
//   mean_average_precision        utils/yolo_utils.py:351-456  (metrics_per_class :458-574)
//   bbox_iou (IoU branch)         utils/yolo_utils.py:75-123,143-144
//
// The reference matches detections one by one in Python.  The greedy rule is order-free once
// restated: the best ground truth g*(d) of a detection depends only on d and the ground truths
// (strict >, first maximum wins, utils/yolo_utils.py:412-417), and d is a true positive iff
// IoU(d, g*) > thr and d is the FIRST detection in descending-score order that selects g*
// (:419-430).  So: one stable radix sort of the detections by (class, score desc), one parallel
// pass computing g*(d) and an atomicMin of the sort rank per ground truth, one pass comparing.
#include "ypp_kernels.cuh"
#include <limits.h>

namespace {

// ------------------------------------------------------------------ GT ingest
__global__ void __launch_bounds__(1024)
ypp_target_rows_kernel(const float* __restrict__ annot, int B, int T, float input_size, float img_base,
                       float* __restrict__ out, int* __restrict__ gt_count, long long capacity,
                       const long long* cursor_in, long long* cursor_out, long long* host_probe, unsigned probe_seq)
{
    __shared__ int warp_tot[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int total = B * T;
    // In append mode the kernel sits between two chains of a validation loop and is launched early (PDL): its CTA is
    // scheduled while the kernel ahead drains and it lets the next chain's D1 be scheduled at once.  Every global
    // read and write - the annotations, the cursor word the previous ground-truth kernel wrote, the rows - comes
    // after griddepcontrol.wait, i.e. after the whole grid ahead has completed and flushed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const long long base = cursor_in ? *cursor_in : 0;      // append mode: rows already in `out`
    int running = 0;
    for (int i0 = 0; i0 < total; i0 += 1024) {
        const int i = i0 + tid;
        bool ok = false;
        float r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
        if (i < total) {
            const float* r = annot + (size_t)i * 5;
            r0 = r[0]; r1 = r[1]; r2 = r[2]; r3 = r[3]; r4 = r[4];
            float s = __fadd_rn(r0, r1); s = __fadd_rn(s, r2); s = __fadd_rn(s, r3); s = __fadd_rn(s, r4);
            ok = !(s <= 0.0f);                       // `if target[b, t].sum() <= 0: continue` (:697)
        }
        const u32 m = __ballot_sync(YPP_FULL_MASK, ok);
        if (lane == 0) warp_tot[warp] = __popc(m);
        __syncthreads();
        int before = 0, tot = 0;
        for (int w = 0; w < 32; ++w) { const int t = warp_tot[w]; if (w < warp) before += t; tot += t; }
        const long long row = base + running + before + __popc(m & ((1u << lane) - 1u));
        if (ok && row < capacity) {
            float* o = out + (size_t)row * 7;
            o[0] = img_base + (float)(i / T);
            o[1] = __fmul_rn(r0, input_size); o[2] = __fmul_rn(r1, input_size);
            o[3] = __fmul_rn(r2, input_size); o[4] = __fmul_rn(r3, input_size);
            o[5] = 1.0f; o[6] = r4;
        }
        running += tot;
        __syncthreads();
    }
    if (tid == 0) {
        if (gt_count) *gt_count = running;
        if (cursor_out) *cursor_out = base + running;       // one block: may be the word cursor_in points at
        if (host_probe) *(volatile long long*)host_probe = ((base + running) << 24) | (long long)(probe_seq & 0xffffffu);
    }
}

// ------------------------------------------------------------------ box_iou (N x M)
__global__ void __launch_bounds__(256)
ypp_box_iou_kernel(const float4* __restrict__ a, int N, const float4* __restrict__ b, int M, float* __restrict__ out)
{
    const int j = blockIdx.x * 32 + (threadIdx.x & 31);
    const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= N || j >= M) return;
    const float4 p = __ldg(a + i), q = __ldg(b + j);
    const float s1 = __fmul_rn(__fsub_rn(p.z, p.x), __fsub_rn(p.w, p.y));
    const float s2 = __fmul_rn(__fsub_rn(q.z, q.x), __fsub_rn(q.w, q.y));
    const float w = fmaxf(__fsub_rn(fminf(p.z, q.z), fmaxf(p.x, q.x)), 0.0f);
    const float h = fmaxf(__fsub_rn(fminf(p.w, q.w), fmaxf(p.y, q.y)), 0.0f);
    const float inter = __fmul_rn(w, h);
    out[(size_t)i * M + j] = __fdiv_rn(inter, __fsub_rn(__fadd_rn(s1, s2), inter));
}

// ------------------------------------------------------------------ bbox_iou of utils/metrics.py:192-230
// The Ultralytics flavour (not the one of utils/yolo_utils.py that the mAP matching uses: no abs, eps 1e-7 on the
// union, an xywh flag, and the GIoU / DIoU / CIoU returns): box1 [1|n,4] against box2 [n,4], element by element,
// one fp32 rounding per torch operation in the reference's operand order.  mode: 0 IoU, 1 GIoU, 2 DIoU, 3 CIoU.
__global__ void __launch_bounds__(256)
ypp_bbox_iou_kernel(const float4* __restrict__ a, int na, const float4* __restrict__ b, int nb, int n, int xywh, int mode,
                    float eps, float one_eps, float* __restrict__ out)
{
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= n) return;
    const float4 p = __ldg(a + (na > 1 ? k : 0)), q = __ldg(b + (nb > 1 ? k : 0));
    float ax1, ay1, ax2, ay2, bx1, by1, bx2, by2, w1, h1, w2, h2;
    if (xywh) {
        w1 = p.z; h1 = p.w; w2 = q.z; h2 = q.w;
        const float w1h = __fmul_rn(w1, 0.5f), h1h = __fmul_rn(h1, 0.5f), w2h = __fmul_rn(w2, 0.5f), h2h = __fmul_rn(h2, 0.5f);   // x / 2, exact
        ax1 = __fsub_rn(p.x, w1h); ax2 = __fadd_rn(p.x, w1h); ay1 = __fsub_rn(p.y, h1h); ay2 = __fadd_rn(p.y, h1h);
        bx1 = __fsub_rn(q.x, w2h); bx2 = __fadd_rn(q.x, w2h); by1 = __fsub_rn(q.y, h2h); by2 = __fadd_rn(q.y, h2h);
    } else {
        ax1 = p.x; ay1 = p.y; ax2 = p.z; ay2 = p.w;
        bx1 = q.x; by1 = q.y; bx2 = q.z; by2 = q.w;
        w1 = __fsub_rn(ax2, ax1); h1 = __fadd_rn(__fsub_rn(ay2, ay1), eps);
        w2 = __fsub_rn(bx2, bx1); h2 = __fadd_rn(__fsub_rn(by2, by1), eps);
    }
    const float iw = fmaxf(__fsub_rn(fminf(ax2, bx2), fmaxf(ax1, bx1)), 0.0f);
    const float ih = fmaxf(__fsub_rn(fminf(ay2, by2), fmaxf(ay1, by1)), 0.0f);
    const float inter = __fmul_rn(iw, ih);
    const float uni = __fadd_rn(__fsub_rn(__fadd_rn(__fmul_rn(w1, h1), __fmul_rn(w2, h2)), inter), eps);
    const float iou = __fdiv_rn(inter, uni);
    float res = iou;
    if (mode != 0) {
        const float cw = __fsub_rn(fmaxf(ax2, bx2), fminf(ax1, bx1));
        const float ch = __fsub_rn(fmaxf(ay2, by2), fminf(ay1, by1));
        if (mode == 1) {
            const float ca = __fadd_rn(__fmul_rn(cw, ch), eps);
            res = __fsub_rn(iou, __fdiv_rn(__fsub_rn(ca, uni), ca));
        } else {
            const float c2 = __fadd_rn(__fadd_rn(__fmul_rn(cw, cw), __fmul_rn(ch, ch)), eps);
            const float dx = __fsub_rn(__fsub_rn(__fadd_rn(bx1, bx2), ax1), ax2);
            const float dy = __fsub_rn(__fsub_rn(__fadd_rn(by1, by2), ay1), ay2);
            const float rho2 = __fmul_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), 0.25f);
            if (mode == 2) {
                res = __fsub_rn(iou, __fdiv_rn(rho2, c2));
            } else {
                const float da = __fsub_rn(atanf(__fdiv_rn(w2, h2)), atanf(__fdiv_rn(w1, h1)));
                const float v = __fmul_rn(0.40528473456935109f, __fmul_rn(da, da));            // 4 / pi^2
                const float alpha = __fdiv_rn(v, __fadd_rn(__fsub_rn(v, iou), one_eps));
                res = __fsub_rn(iou, __fadd_rn(__fdiv_rn(rho2, c2), __fmul_rn(v, alpha)));
            }
        }
    }
    out[k] = res;
}

// ------------------------------------------------------------------ matching

// class ids that are not integers in [0, nc) can never equal a loop class c (:370-375): marker nc
__device__ __forceinline__ u32 map_cls(float cf, int nc)
{
    return (cf >= 0.0f && cf < (float)nc && (float)(int)cf == cf) ? (u32)cf : (u32)nc;
}

// bbox_iou IoU branch on cxcywh (abs areas, eps on the union): utils/yolo_utils.py:103-123
__device__ __forceinline__ float map_iou(const float* a, const float* b)
{
    const float ahw = __fmul_rn(a[2], 0.5f), ahh = __fmul_rn(a[3], 0.5f);
    const float bhw = __fmul_rn(b[2], 0.5f), bhh = __fmul_rn(b[3], 0.5f);
    const float ax1 = __fsub_rn(a[0], ahw), ay1 = __fsub_rn(a[1], ahh), ax2 = __fadd_rn(a[0], ahw), ay2 = __fadd_rn(a[1], ahh);
    const float bx1 = __fsub_rn(b[0], bhw), by1 = __fsub_rn(b[1], bhh), bx2 = __fadd_rn(b[0], bhw), by2 = __fadd_rn(b[1], bhh);
    const float iw = fmaxf(__fsub_rn(fminf(ax2, bx2), fmaxf(ax1, bx1)), 0.0f);
    const float ih = fmaxf(__fsub_rn(fminf(ay2, by2), fmaxf(ay1, by1)), 0.0f);
    const float inter = __fmul_rn(iw, ih);
    const float sa = fabsf(__fmul_rn(__fsub_rn(ax2, ax1), __fsub_rn(ay2, ay1)));
    const float sb = fabsf(__fmul_rn(__fsub_rn(bx2, bx1), __fsub_rn(by2, by1)));
    const float uni = __fadd_rn(__fsub_rn(__fadd_rn(sa, sb), inter), 1e-6f);
    return __fdiv_rn(inter, uni);
}

// lanes of the warp that hold the same 8-bit digit (eight ballots; lanes with ok == false belong to no group)
__device__ __forceinline__ u32 same_digit_lanes(int v, bool ok)
{
    u32 peers = __ballot_sync(YPP_FULL_MASK, ok);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool bit = (v >> i) & 1;
        const u32 bal = __ballot_sync(YPP_FULL_MASK, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

// One stable LSD pass (8 bits at `shift`) over items [0, len) of a class, by the whole CTA: warp w owns the contiguous
// ranks [w L, (w+1) L) and walks them 32 at a time, so "warp, then step, then lane" is the input order; counters per
// (digit, warp) make the scatter stable.  tab: 256 x 33 ints of shared memory.
#define AP_THREADS 1024
#define AP_TAB_PITCH 33
__device__ void cta_radix_pass(const u64* __restrict__ src, u64* __restrict__ dst, long long len, long long L, int shift,
                               int* tab, int* warp_tot)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 256 * AP_TAB_PITCH; i += AP_THREADS) tab[i] = 0;
    __syncthreads();
    const long long r_begin = min(len, (long long)warp * L), r_end = min(len, r_begin + L);
    for (long long r = r_begin + lane; r < r_end; r += 32)
        atomicAdd(&tab[(int)((src[r] >> shift) & 255ull) * AP_TAB_PITCH + warp], 1);       // only this warp touches column `warp`
    __syncthreads();
    {   // exclusive scan in (digit, warp) order: thread t owns the 8 entries digit = t / 4, warps 8 (t % 4) ..
        const int d = tid >> 2, w0 = (tid & 3) * 8;
        int v[8], sum = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { v[k] = tab[d * AP_TAB_PITCH + w0 + k]; sum += v[k]; }
        int inc = sum;
#pragma unroll
        for (int q = 1; q < 32; q <<= 1) { const int o = __shfl_up_sync(YPP_FULL_MASK, inc, q); if (lane >= q) inc += o; }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        int run = inc - sum;
        for (int w = 0; w < warp; ++w) run += warp_tot[w];
#pragma unroll
        for (int k = 0; k < 8; ++k) { tab[d * AP_TAB_PITCH + w0 + k] = run; run += v[k]; }
    }
    __syncthreads();
    u64 kn = (r_begin + lane < r_end) ? src[r_begin + lane] : 0ull;
    for (long long r0 = r_begin; r0 < r_end; r0 += 32) {
        const long long r = r0 + lane;
        const bool ok = r < r_end;
        const u64 k = kn;
        kn = (r + 32 < r_end) ? src[r + 32] : 0ull;          // the next step's key is in flight during this step's ballots
        const int d = ok ? (int)((k >> shift) & 255ull) : 0;
        const u32 peers = same_digit_lanes(d, ok);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        int base = 0;
        if (ok) base = tab[d * AP_TAB_PITCH + warp];
        __syncwarp();
        if (ok) {
            if (rank == 0) tab[d * AP_TAB_PITCH + warp] = base + __popc(peers);
            dst[base + rank] = k;
        }
        __syncwarp();
    }
    __syncthreads();                                   // the block's global writes are visible to the block after it
}

// one CTA per class c_lo + blockIdx.x.  Its items (u64: (~enc(score)) << 1 | tp, in input order) lie at
// [table[class row, block 0], start of the next class) of `items`; they are sorted here by the 32 score bits (four stable
// passes, ending in `items` again), then: cumulative precision/recall, envelope, AP.  The class's ground-truth count comes
// from gt_per_class.  term[] / prec[] (one float per item) are scratch at the class's own positions.
__global__ void __launch_bounds__(AP_THREADS)
map_ap_kernel(const int* __restrict__ table, int nblocks, int nc, int c_lo, int c_hi,
              u64* __restrict__ items, u64* __restrict__ items2, const long long* __restrict__ gt_per_class,
              float* __restrict__ prec, float* __restrict__ term, float* __restrict__ class_ap, float* __restrict__ per_class)
{
    __shared__ int tab[256 * AP_TAB_PITCH];
    __shared__ int warp_i[32];
    __shared__ float warp_f[32];
    __shared__ int carry_i;
    __shared__ float carry_f;
    const int c = c_lo + blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long s_lo = table[(size_t)(c - c_lo) * nblocks], s_hi = table[(size_t)(c - c_lo + 1) * nblocks];   // the table has one entry more: the total
    const long long s_ng = gt_per_class[c];
    if (tid == 0) carry_i = 0;
    if (s_ng != 0 && s_hi > s_lo) {
        const long long len = s_hi - s_lo;
        const long long L = ((len + 31) / 32 + 31) / 32 * 32;
        u64* a = items + s_lo;
        u64* b = items2 + s_lo;
        for (int pass = 0; pass < 4; ++pass) {
            cta_radix_pass(a, b, len, L, 1 + 8 * pass, tab, warp_i);
            u64* t = a; a = b; b = t;
        }
    }
    const u64* __restrict__ sorted = items;            // four passes: back where they started
    __syncthreads();
    const long long lo = s_lo, hi = s_hi;
    const long long ng = s_ng;
    float* pc = per_class ? per_class + 4 * (size_t)c : nullptr;
    if (ng == 0) {                                           // class skipped, AP 0 (:379-381)
        if (pc && tid < 4) pc[tid] = 0.0f;
        if (tid == 0) class_ap[c] = 0.0f;
        return;
    }
    const float rec_den = (float)((double)ng + 1e-6);         // python float64 (G + eps) -> fp32
    constexpr int IT = 4, CH = AP_THREADS * IT;              // ranks per thread and per round
    // forward: cumulative TP -> precision at every rank (thread = IT consecutive ranks)
    for (long long p0 = lo; p0 < hi; p0 += CH) {
        const long long pb = p0 + (long long)tid * IT;
        int t[IT], tsum = 0;
#pragma unroll
        for (int q = 0; q < IT; ++q) { t[q] = (pb + q < hi) ? (int)(sorted[pb + q] & 1ull) : 0; tsum += t[q]; }
        int inc = tsum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(YPP_FULL_MASK, inc, d); if (lane >= d) inc += o; }
        if (lane == 31) warp_i[warp] = inc;
        __syncthreads();
        const int wv = warp_i[lane];
        int winc = wv;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(YPP_FULL_MASK, winc, d); if (lane >= d) winc += o; }
        const int tot = __shfl_sync(YPP_FULL_MASK, winc, 31);
        const int before = __shfl_sync(YPP_FULL_MASK, winc - wv, warp);
        const int cbase = carry_i;
        int tpc = cbase + before + inc - tsum;
#pragma unroll
        for (int q = 0; q < IT; ++q) {
            const long long p = pb + q;
            if (p < hi) {
                tpc += t[q];
                const float ftp = (float)tpc, ffp = (float)((int)(p - lo + 1) - tpc);
                prec[p] = __fdiv_rn(ftp, __fadd_rn(__fadd_rn(ftp, ffp), 1e-6f));             // :435
                if (t[q]) term[lo + tpc - 1] = __uint_as_float((u32)(p - lo));               // rank of the tpc-th TP (int bits)
            }
        }
        __syncthreads();
        if (tid == 0) carry_i = cbase + tot;
        __syncthreads();
    }
    const int n_tp = carry_i;
    // backward: precision envelope = running max from the right (:442-443); thread 0 owns the right-most IT ranks
    if (tid == 0) carry_f = -INFINITY;
    __syncthreads();
    const long long len = hi - lo;
    for (long long r0 = 0; r0 < len; r0 += CH) {
        const long long pr = hi - 1 - r0 - (long long)tid * IT;      // my right-most rank; I walk pr, pr-1, ..
        float v[IT], m = -INFINITY;
#pragma unroll
        for (int q = 0; q < IT; ++q) { v[q] = (pr - q >= lo) ? prec[pr - q] : -INFINITY; m = fmaxf(m, v[q]); }
        float inc = m;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const float o = __shfl_up_sync(YPP_FULL_MASK, inc, d); if (lane >= d) inc = fmaxf(inc, o); }
        if (lane == 31) warp_f[warp] = inc;
        __syncthreads();
        float before = carry_f, tot = carry_f;
        for (int w = 0; w < 32; ++w) { const float x = warp_f[w]; if (w < warp) before = fmaxf(before, x); tot = fmaxf(tot, x); }
        float run = fmaxf(before, __shfl_up_sync(YPP_FULL_MASK, inc, 1));     // everything to my right
        if (lane == 0) run = before;
#pragma unroll
        for (int q = 0; q < IT; ++q) {
            if (pr - q >= lo) { run = fmaxf(run, v[q]); prec[pr - q] = run; }
        }
        __syncthreads();
        if (tid == 0) carry_f = tot;
        __syncthreads();
    }
    // AP = sum over the recall change points (every true positive) of (recall step) x (envelope there), :445-452.
    // The terms are independent: every thread computes some (same roundings as the reference's one-by-one
    // evaluation); only the fp32 ADDITIONS stay sequential, in rank order, so the sum rounds as the reference's:
    // the terms are staged in shared memory CH at a time and one thread adds them up.
    static_assert(CH * sizeof(float) <= sizeof(tab), "the staged terms reuse the radix table");
    float* stage = reinterpret_cast<float*>(tab);                  // the sort is long over
    float ap = 0.0f;
    for (int t0 = 0; t0 < n_tp; t0 += CH) {
        for (int t = t0 + tid; t < min(n_tp, t0 + CH); t += AP_THREADS) {
            const long long p = lo + (long long)__float_as_uint(term[lo + t]);
            const float r1 = __fdiv_rn((float)(t + 1), rec_den), r0 = __fdiv_rn((float)t, rec_den);
            stage[t - t0] = __fmul_rn(__fsub_rn(r1, r0), prec[p]);
        }
        __syncthreads();
        if (tid == 0) {
            const int n = min(CH, n_tp - t0);
            for (int k = 0; k < n; ++k) ap = __fadd_rn(ap, stage[k]);
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (pc) { pc[0] = ap; pc[1] = (float)n_tp; pc[2] = (float)((int)len - n_tp); pc[3] = (float)((int)ng - n_tp); }
        class_ap[c] = ap;
    }
}

// mean over ALL nc classes (torch.mean of the stacked per-class APs, :456), fp32, in class order
__global__ void map_mean_kernel(int nc, const float* __restrict__ class_ap, int stride, float* __restrict__ map_out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float s = 0.0f;
    for (int c = 0; c < nc; ++c) s = __fadd_rn(s, class_ap[(size_t)c * stride]);
    *map_out = __fdiv_rn(s, (float)nc);
}

// ---- matching (M1): no sort at all ----------------------------------------------------------------------------
// A detection's best ground truth g*(d) depends on d and its image's ground truths only (same image id - float
// equality, utils/yolo_utils.py:407 -, same class, strict >, first maximum in row order wins, :412-417); d is a true
// positive iff IoU > thr and d is the first in (score descending, row order) among the detections that chose g*(d)
// (:419-430).  "First" is a 64-bit minimum: claim[g] = min over those detections of (~enc(score)) << 32 | d.
// The ground truths of a validation run arrive in image order (the accumulators append them batch after batch), so
// the rows of an image are found by a binary search on the image column and the ground truths are never sorted.
// Any other row order is still handled - `unsorted` is raised on the device by the prepare kernel and every
// detection then scans all G rows: correct, O(K G).
__global__ void map_gt_prepare_kernel(const float* __restrict__ gt, long long G, int nc, u64* __restrict__ claim,
                                      unsigned long long* __restrict__ per_class, int* __restrict__ unsorted)
{
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    claim[g] = ~0ull;
    const u32 cls = map_cls(gt[g * 7 + 6], nc);
    if (cls < (u32)nc) atomicAdd(per_class + cls, 1ull);
    if (g > 0 && !(gt[g * 7] >= gt[(g - 1) * 7])) *unsorted = 1;        // every writer writes the same value
}

__global__ void map_claim_kernel(const float* __restrict__ det, long long K, const float* __restrict__ gt, long long G, int nc,
                                 const int* __restrict__ unsorted, float iou_thr, int* __restrict__ best_g, u64* __restrict__ claim)
{
    const long long d = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= K) return;
    const float* dr = det + (size_t)d * 7;
    const u32 cls = map_cls(dr[6], nc);
    const float img = dr[0];
    const bool sorted = *unsorted == 0;
    long long s = 0;
    if (sorted) {                                             // first row whose image id is not below mine
        long long hi = G;
        while (s < hi) { const long long mid = (s + hi) >> 1; if (gt[mid * 7] < img) s = mid + 1; else hi = mid; }
    }
    float best = 0.0f;
    int bg = -1;
    if (cls < (u32)nc) {
        for (; s < G; ++s) {
            const float* gr = gt + (size_t)s * 7;
            const float gi = gr[0];
            if (gi != img) { if (sorted) break; continue; }
            if (map_cls(gr[6], nc) != cls) continue;
            const float iou = map_iou(dr + 1, gr + 1);
            if (iou > best) { best = iou; bg = (int)s; }
        }
    }
    if (!(best > iou_thr)) bg = -1;
    best_g[d] = bg;
    if (bg >= 0) atomicMin(claim + bg, ((u64)(~ypp_enc(dr[5])) << 32) | (u64)(u32)d);
}

__global__ void map_flags_kernel(const float* __restrict__ det, long long K, int nc, const int* __restrict__ best_g,
                                 const u64* __restrict__ claim, const long long* __restrict__ gt_per_class,
                                 unsigned char* __restrict__ tp_out, float* __restrict__ triples)
{
    const long long d = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= K) return;
    const float score = det[d * 7 + 5], cf = det[d * 7 + 6];
    const u32 cls = map_cls(cf, nc);
    const int bg = best_g[d];
    const bool tp = bg >= 0 && claim[bg] == (((u64)(~ypp_enc(score)) << 32) | (u64)(u32)d);
    const bool counted = cls < (u32)nc && gt_per_class[cls] > 0;
    if (tp_out) tp_out[d] = counted ? (tp ? 1 : 0) : 255;
    if (triples) { triples[d * 3] = score; triples[d * 3 + 1] = cf; triples[d * 3 + 2] = tp ? 1.0f : 0.0f; }
}

// ---- AP from flagged detections gathered from every rank (M2), for the classes [c_lo, c_hi) ----------------------
// Input: n_seg segments of seg_stride triples (score, cls, tp) each, of which the first seg_count[s] are valid
// (the padded all-gather buffer as it is).  Three steps, four launches:
//   select   the triples of my classes, GROUPED BY CLASS and in input order inside a class (per-(class, block) counts ->
//            one scan in class-major order -> scatter): item = (~enc(score)) << 1 | tp, one u64 each;
//   sort+AP  one CTA per class: stable LSD radix sort of its own items by the 32 score bits (ties keep the input order:
//            segment after segment, row after row - the order one process would have seen), then the precision /
//            recall scans and the AP, all in the same kernel.
#define SEL_THREADS 256
#define SEL_ITEMS 8
#define SEL_TILE (SEL_THREADS * SEL_ITEMS)
#define SEL_MAX_CLASSES 128            /* classes per call the shared-memory counters hold; more: the call is split */
__device__ __forceinline__ bool sel_take(const float* __restrict__ tr, long long i, long long seg_stride,
                                         const long long* __restrict__ seg_count, int nc, int c_lo, int c_hi, u32& cls)
{
    const long long sgm = i / seg_stride, r = i - sgm * seg_stride;
    if (seg_count && r >= seg_count[sgm]) return false;
    cls = map_cls(tr[i * 3 + 1], nc);
    return cls >= (u32)c_lo && cls < (u32)c_hi;
}

// table[(c - c_lo) * nblocks + block] = items of class c in this block's tile
__global__ void __launch_bounds__(SEL_THREADS)
map_select_count_kernel(const float* __restrict__ tr, long long n, long long seg_stride, const long long* __restrict__ seg_count,
                        int nc, int c_lo, int c_hi, int* __restrict__ table, int nblocks)
{
    __shared__ int cnt[SEL_MAX_CLASSES];
    const int ncls = c_hi - c_lo;
    for (int c = threadIdx.x; c < ncls; c += SEL_THREADS) cnt[c] = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * SEL_TILE;
    for (int it = 0; it < SEL_ITEMS; ++it) {
        const long long i = base + (long long)it * SEL_THREADS + threadIdx.x;
        u32 cls;
        if (i < n && sel_take(tr, i, seg_stride, seg_count, nc, c_lo, c_hi, cls)) atomicAdd(&cnt[cls - c_lo], 1);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < ncls; c += SEL_THREADS) table[(size_t)c * nblocks + blockIdx.x] = cnt[c];
}

// exclusive scan of the per-block counts of the selection, one CTA
__global__ void __launch_bounds__(1024)
block_offsets_kernel(int* __restrict__ counts, int n)
{
    __shared__ int warp_tot[32];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += 1024) {
        const int i = i0 + tid;
        const int v = i < n ? counts[i] : 0;
        int inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(YPP_FULL_MASK, inc, d); if (lane >= d) inc += o; }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        int before = 0, tot = 0;
        for (int w = 0; w < 32; ++w) { const int t = warp_tot[w]; if (w < warp) before += t; tot += t; }
        const int c = carry;
        if (i < n) counts[i] = c + before + inc - v;
        __syncthreads();
        if (tid == 0) carry = c + tot;
        __syncthreads();
    }
}

// the table holds exclusive offsets now (block_offsets_kernel over the class-major table): entry (c, block) is where
// this block's items of class c go.  Stable inside the block: 256 items at a time, rank among the lanes of the same
// class by MATCH.ANY, per-(warp, class) counts turned into offsets by the thread that owns the class.
__global__ void __launch_bounds__(SEL_THREADS)
map_select_scatter_kernel(const float* __restrict__ tr, long long n, long long seg_stride, const long long* __restrict__ seg_count,
                          int nc, int c_lo, int c_hi, const int* __restrict__ table, int nblocks, u64* __restrict__ items)
{
    __shared__ int run[SEL_MAX_CLASSES];
    __shared__ int wc[SEL_THREADS / 32][SEL_MAX_CLASSES];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncls = c_hi - c_lo;
    for (int c = tid; c < ncls; c += SEL_THREADS) run[c] = table[(size_t)c * nblocks + blockIdx.x];
    const long long base = (long long)blockIdx.x * SEL_TILE;
    for (int it = 0; it < SEL_ITEMS; ++it) {
        for (int c = tid; c < ncls; c += SEL_THREADS)
            for (int w = 0; w < SEL_THREADS / 32; ++w) wc[w][c] = 0;
        __syncthreads();
        const long long i = base + (long long)it * SEL_THREADS + tid;
        u32 cls = 0;
        const bool take = i < n && sel_take(tr, i, seg_stride, seg_count, nc, c_lo, c_hi, cls);
        const int key = take ? (int)(cls - c_lo) : -1 - lane;                 // lanes without an item match nobody
        const unsigned peers = __match_any_sync(YPP_FULL_MASK, key);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        if (take && rank == 0) wc[warp][key] = __popc(peers);
        __syncthreads();
        for (int c = tid; c < ncls; c += SEL_THREADS) {                       // per-warp counts -> destinations, in warp order
            int r = run[c];
            for (int w = 0; w < SEL_THREADS / 32; ++w) { const int t = wc[w][c]; wc[w][c] = r; r += t; }
            run[c] = r;
        }
        __syncthreads();
        if (take) items[wc[warp][key] + rank] = ((u64)(~ypp_enc(tr[i * 3])) << 1) | (tr[i * 3 + 2] != 0.0f ? 1ull : 0ull);
        __syncthreads();
    }
}

struct MapWs {
    u64* items[2];
    int* best_g; float* prec; float* term; float* class_ap; float* triples;
    u64* claim; unsigned long long* gt_per_class; int* table; int* unsorted;
    size_t total;
};

MapWs make_map_ws(void* base, long long G, long long K, int nc)
{
    MapWs w;
    char* p = (char*)base;
    size_t off = 0;
    auto take = [&](size_t bytes) { char* r = p + off; off += ypp_align_up(bytes ? bytes : 8, 256); return r; };
    const long long Kp = K > 0 ? K : 1, Gp = G > 0 ? G : 1;
    const size_t nb = (size_t)((Kp + SEL_TILE - 1) / SEL_TILE);
    const size_t ncl = (size_t)(nc < SEL_MAX_CLASSES ? (nc > 0 ? nc : 1) : SEL_MAX_CLASSES);
    w.items[0] = (u64*)take(8 * Kp); w.items[1] = (u64*)take(8 * Kp);
    w.best_g = (int*)take(4 * Kp);
    w.prec = (float*)take(4 * Kp);
    w.term = (float*)take(4 * Kp);
    w.triples = (float*)take(12 * Kp);
    w.class_ap = (float*)take(4 * (size_t)(nc > 0 ? nc : 1));
    w.claim = (u64*)take(8 * Gp);
    w.gt_per_class = (unsigned long long*)take(8 * (size_t)(nc > 0 ? nc : 1) + 8);     // + the `unsorted` word: one memset
    w.unsorted = (int*)(w.gt_per_class + (nc > 0 ? nc : 1));
    w.table = (int*)take(4 * (ncl * nb + 1));
    w.total = off;
    return w;
}


// ---- detection writers: the integer columns of the pred-file / COCO rows -------------------
// make_pred_file_yolov3.py:73-86, yolo2coco_pred_file_yolov3.py:83-100.  The reference works on
// 0-dim fp32 tensors: `w / 2`, `cx - ...` are fp32 operations, `bbox[0] * (width / input_size)`
// multiplies by the Python-float ratio rounded to fp32, int() truncates toward zero.
__global__ void __launch_bounds__(256)
ypp_detection_ints_kernel(const float* __restrict__ rows, long long K, int row_width, int mode,
                          const float* __restrict__ scale_xy, int n_img, float img_base, int* __restrict__ out)
{
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const float* r = rows + k * row_width;
    const int o = row_width == 7 ? 1 : 0;
    float cx = r[o], cy = r[o + 1], w = r[o + 2], h = r[o + 3];
    int4 v;
    if (mode == YPP_WRITER_COCO) {
        int img = o ? (int)(r[0] - img_base) : 0;
        img = min(max(img, 0), n_img - 1);
        const float sx = scale_xy[2 * img], sy = scale_xy[2 * img + 1];
        cx = __fmul_rn(cx, sx); cy = __fmul_rn(cy, sy); w = __fmul_rn(w, sx); h = __fmul_rn(h, sy);
        v.x = __float2int_rz(__fsub_rn(cx, __fmul_rn(w, 0.5f)));
        v.y = __float2int_rz(__fsub_rn(cy, __fmul_rn(h, 0.5f)));
        v.z = __float2int_rz(w);
        v.w = __float2int_rz(h);
    } else {
        const float hw = __fmul_rn(w, 0.5f), hh = __fmul_rn(h, 0.5f);        // x / 2 is exact either way
        v.x = __float2int_rz(__fsub_rn(cx, hw));
        v.y = __float2int_rz(__fsub_rn(cy, hh));
        v.z = __float2int_rz(__fadd_rn(cx, hw));
        v.w = __float2int_rz(__fadd_rn(cy, hh));
    }
    reinterpret_cast<int4*>(out)[k] = v;
}

// ------------------------------------------------------------------ ConfusionMatrix.process_batch
// utils/metrics.py:122-160, one CTA per image.  The reference builds the (gt, det, iou) triples with
// iou > thr, sorts them by iou, keeps per detection its best triple, sorts again and keeps per ground
// truth its best remaining triple.  Restated without the sorts: g*(d) = the ground truth of highest
// iou with d (above thr), d*(g) = the detection of highest iou among those with g*(d) = g.  Ties (the
// reference's argsort is unstable there) go to the lower index.  Then, as the reference does:
// matched gt -> matrix[cls(d*)][cls(g)], unmatched gt -> matrix[nc][cls(g)], and ONLY IF the image has
// any triple at all, every unmatched detection -> matrix[cls(d)][nc]  (:150-157).
#define YPP_CM_THREADS 256
__global__ void __launch_bounds__(YPP_CM_THREADS)
ypp_confusion_kernel(const float* __restrict__ det, const int* __restrict__ det_off, const float* __restrict__ lab,
                     const int* __restrict__ lab_off, int nc, float conf, float thr, int* __restrict__ bestg,
                     float* __restrict__ bestiou, unsigned long long* __restrict__ matrix)
{
    __shared__ int any_match;
    const int img = blockIdx.x, tid = threadIdx.x;
    const int d0 = det_off[img], nd = det_off[img + 1] - d0;
    const int l0 = lab_off[img], nl = lab_off[img + 1] - l0;
    if (tid == 0) any_match = 0;
    __syncthreads();
    const int W = nc + 1;
    for (int d = tid; d < nd; d += YPP_CM_THREADS) {
        const float* r = det + (size_t)(d0 + d) * 6;
        int g_best = -2;                                        // -2: below the confidence threshold (:134)
        float i_best = 0.0f;
        if (r[4] > conf) {
            g_best = -1;
            const float4 q = make_float4(r[0], r[1], r[2], r[3]);
            const float s2 = __fmul_rn(__fsub_rn(q.z, q.x), __fsub_rn(q.w, q.y));
            for (int g = 0; g < nl; ++g) {
                const float* l = lab + (size_t)(l0 + g) * 5;
                const float4 p = make_float4(l[1], l[2], l[3], l[4]);
                const float s1 = __fmul_rn(__fsub_rn(p.z, p.x), __fsub_rn(p.w, p.y));
                const float w = fmaxf(__fsub_rn(fminf(p.z, q.z), fmaxf(p.x, q.x)), 0.0f);
                const float h = fmaxf(__fsub_rn(fminf(p.w, q.w), fmaxf(p.y, q.y)), 0.0f);
                const float inter = __fmul_rn(w, h);
                const float iou = __fdiv_rn(inter, __fsub_rn(__fadd_rn(s1, s2), inter));      // box_iou(labels, detections)
                if (iou > thr && (g_best < 0 || iou > i_best)) { g_best = g; i_best = iou; }
            }
            if (g_best >= 0) any_match = 1;
        }
        bestg[d0 + d] = g_best;
        bestiou[d0 + d] = i_best;
    }
    __syncthreads();
    for (int g = tid; g < nl; g += YPP_CM_THREADS) {
        int dm = -1;
        float im = 0.0f;
        for (int d = 0; d < nd; ++d)
            if (bestg[d0 + d] == g) { const float v = bestiou[d0 + d]; if (dm < 0 || v > im) { dm = d; im = v; } }
        const int gc = (int)lab[(size_t)(l0 + g) * 5];
        if (gc < 0 || gc >= nc) continue;
        if (dm >= 0) {
            bestg[d0 + dm] = -3;                                 // matched: no longer equal to any g
            const int dc = (int)det[(size_t)(d0 + dm) * 6 + 5];
            if (dc >= 0 && dc < nc) atomicAdd(matrix + (size_t)dc * W + gc, 1ull);
        } else {
            atomicAdd(matrix + (size_t)nc * W + gc, 1ull);
        }
    }
    __syncthreads();
    if (!any_match) return;
    for (int d = tid; d < nd; d += YPP_CM_THREADS) {
        if (bestg[d0 + d] < -1) continue;                        // filtered out, or matched
        const int dc = (int)det[(size_t)(d0 + d) * 6 + 5];
        if (dc >= 0 && dc < nc) atomicAdd(matrix + (size_t)dc * W + nc, 1ull);
    }
}

}  // namespace

int ypp_launch_target_rows(const float* annot, int B, int T, float input_size, float img_base,
                           float* out, int* gt_count, long long capacity, const long long* cursor_in,
                           long long* cursor_out, long long* host_probe, unsigned probe_seq, cudaStream_t st)
{
    if (cursor_in)      // append mode: the buffers are the accumulator's own, nothing of the kernel before is read
        ypp_launch_chain(0, ypp_target_rows_kernel, dim3(1), dim3(1024), 0, st, annot, B, T, input_size, img_base, out, gt_count,
                         capacity, cursor_in, cursor_out, host_probe, probe_seq);
    else
        ypp_target_rows_kernel<<<1, 1024, 0, st>>>(annot, B, T, input_size, img_base, out, gt_count, capacity, cursor_in,
                                                   cursor_out, host_probe, probe_seq);
    ypp_count_launch(1);
    return ypp_check_launch("target_rows");
}

int ypp_launch_box_iou(const float* a, int N, const float* b, int M, float* out, cudaStream_t st)
{
    dim3 grid((M + 31) / 32, (N + 7) / 8);
    ypp_box_iou_kernel<<<grid, 256, 0, st>>>((const float4*)a, N, (const float4*)b, M, out);
    ypp_count_launch(1);
    return ypp_check_launch("box_iou");
}

int ypp_launch_bbox_iou(const float* a, int na, const float* b, int nb, int xywh, int mode, float eps, float one_eps,
                        float* out, cudaStream_t st)
{
    const int n = na > nb ? na : nb;
    ypp_bbox_iou_kernel<<<(n + 255) / 256, 256, 0, st>>>((const float4*)a, na, (const float4*)b, nb, n, xywh, mode, eps, one_eps, out);
    ypp_count_launch(1);
    return ypp_check_launch("bbox_iou");
}

size_t ypp_map_ws_bytes(long long G, long long K, int nc)
{
    return make_map_ws(nullptr, G, K, nc).total;
}

// M1: per-class ground-truth counts, claims, flags.  gpc_out: caller's [nc] int64 or NULL (then the workspace's).
static int launch_m1(const MapWs& w, const float* gt, long long G, const float* det, long long K, int nc, float iou_thr,
                     unsigned char* tp, float* triples, unsigned long long* gpc, cudaStream_t st)
{
    if (gpc == w.gt_per_class) cudaMemsetAsync(w.gt_per_class, 0, 8 * (size_t)nc + 8, st);
    else { cudaMemsetAsync(gpc, 0, 8 * (size_t)nc, st); cudaMemsetAsync(w.unsorted, 0, 4, st); }
    if (G > 0) {
        map_gt_prepare_kernel<<<(unsigned)((G + 255) / 256), 256, 0, st>>>(gt, G, nc, w.claim, gpc, w.unsorted);
        ypp_count_launch(1);
    }
    if (K > 0) {
        map_claim_kernel<<<(unsigned)((K + 255) / 256), 256, 0, st>>>(det, K, gt, G, nc, w.unsorted, iou_thr, w.best_g, w.claim);
        map_flags_kernel<<<(unsigned)((K + 255) / 256), 256, 0, st>>>(det, K, nc, w.best_g, w.claim, (const long long*)gpc, tp, triples);
        ypp_count_launch(2);
    }
    return YPP_OK;
}

// M2 for the classes [c_lo, c_hi) of n triples: select grouped by class -> per-class sort + AP (<= SEL_MAX_CLASSES per round)
static void launch_m2(const MapWs& w, const float* triples, long long n, long long seg_stride, const long long* seg_count,
                      const long long* gpc, int nc, int c_lo, int c_hi, float* per_class, cudaStream_t st)
{
    const int nb = (int)((max(n, 1LL) + SEL_TILE - 1) / SEL_TILE);
    for (int c0 = c_lo; c0 < c_hi; c0 += SEL_MAX_CLASSES) {
        const int c1 = min(c_hi, c0 + SEL_MAX_CLASSES), ncls = c1 - c0;
        if (n > 0) {
            cudaMemsetAsync(w.table + (size_t)ncls * nb, 0, 4, st);           // the extra entry: after the scan it holds the total
            map_select_count_kernel<<<nb, SEL_THREADS, 0, st>>>(triples, n, seg_stride, seg_count, nc, c0, c1, w.table, nb);
            block_offsets_kernel<<<1, 1024, 0, st>>>(w.table, ncls * nb + 1);
            map_select_scatter_kernel<<<nb, SEL_THREADS, 0, st>>>(triples, n, seg_stride, seg_count, nc, c0, c1, w.table, nb, w.items[0]);
            ypp_count_launch(3);
        } else {
            cudaMemsetAsync(w.table, 0, 4 * ((size_t)ncls * nb + 1), st);
        }
        map_ap_kernel<<<ncls, AP_THREADS, 0, st>>>(w.table, nb, nc, c0, c1, w.items[0], w.items[1], gpc, w.prec, w.term, w.class_ap, per_class);
        ypp_count_launch(1);
    }
}

// mean_average_precision in one process = M1, then M2 over every class of the one buffer
int ypp_launch_map(const float* gt, long long G, const float* det, long long K, int nc, float iou_thr,
                   void* ws, size_t ws_bytes, unsigned char* tp, float* per_class, float* map,
                   cudaStream_t st)
{
    MapWs w = make_map_ws(ws, G, K, nc);
    if (ws_bytes < w.total) { ypp_set_error("map_match: workspace %zu < %zu", ws_bytes, w.total); return YPP_E_WORKSPACE; }
    launch_m1(w, gt, G, det, K, nc, iou_thr, tp, w.triples, w.gt_per_class, st);
    launch_m2(w, w.triples, K, K > 0 ? K : 1, nullptr, (const long long*)w.gt_per_class, nc, 0, nc, per_class, st);
    map_mean_kernel<<<1, 32, 0, st>>>(nc, w.class_ap, 1, map);
    ypp_count_launch(1);
    return ypp_check_launch("map_match");
}

int ypp_launch_map_flags(const float* gt, long long G, const float* det, long long K, int nc, float iou_thr, void* ws,
                         size_t ws_bytes, unsigned char* tp, float* triples, long long* gt_per_class, cudaStream_t st)
{
    MapWs w = make_map_ws(ws, G, K, nc);
    if (ws_bytes < w.total) { ypp_set_error("map_flags: workspace %zu < %zu", ws_bytes, w.total); return YPP_E_WORKSPACE; }
    launch_m1(w, gt, G, det, K, nc, iou_thr, tp, triples, gt_per_class ? (unsigned long long*)gt_per_class : w.gt_per_class, st);
    return ypp_check_launch("map_flags");
}

int ypp_launch_map_ap_flags(const float* triples, long long n, long long seg_stride, const long long* seg_count,
                            const long long* gt_per_class, int nc, int c_lo, int c_hi, void* ws, size_t ws_bytes,
                            float* per_class, cudaStream_t st)
{
    MapWs w = make_map_ws(ws, 0, n, nc);
    if (ws_bytes < w.total) { ypp_set_error("map_ap_from_flags: workspace %zu < %zu", ws_bytes, w.total); return YPP_E_WORKSPACE; }
    launch_m2(w, triples, n, seg_stride, seg_count, gt_per_class, nc, c_lo, c_hi, per_class, st);
    return ypp_check_launch("map_ap_from_flags");
}

int ypp_launch_map_mean(const float* per_class, int nc, float* map, cudaStream_t st)
{
    map_mean_kernel<<<1, 32, 0, st>>>(nc, per_class, 4, map);
    ypp_count_launch(1);
    return ypp_check_launch("map_mean");
}

int ypp_launch_detection_ints(const float* rows, long long K, int row_width, int mode, const float* scale_xy, int n_img,
                              float img_base, int* out, cudaStream_t st)
{
    const unsigned grid = (unsigned)((K + 255) / 256);
    ypp_detection_ints_kernel<<<grid, 256, 0, st>>>(rows, K, row_width, mode, scale_xy, n_img, img_base, out);
    ypp_count_launch(1);
    return ypp_check_launch("detection_ints");
}

int ypp_launch_confusion(const float* det, const int* det_off, const float* lab, const int* lab_off, int n_img, int nc,
                         float conf, float thr, int* bestg, float* bestiou, unsigned long long* matrix, cudaStream_t st)
{
    ypp_confusion_kernel<<<n_img, YPP_CM_THREADS, 0, st>>>(det, det_off, lab, lab_off, nc, conf, thr, bestg, bestiou, matrix);
    ypp_count_launch(1);
    return ypp_check_launch("confusion");
}
