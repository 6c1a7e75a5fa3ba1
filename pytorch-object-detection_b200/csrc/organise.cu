// S1: per image ordering.  One CTA per image:
//   1. sort the compacted 64-bit keys (~score | box index | class) ascending in shared memory:
//      the result is the stable descending-score order of torchvision::nms
//      (scores.sort(stable=True, descending=True); ties -> lower box index) for the whole image;
//   2. stable counting partition by class -> per (image, class) segments in the same order;
//   3. emit xyxy boxes (utils/yolo_utils.py:339-343) in segment order, shifted by
//      cls * (max_coordinate + 1) when torchvision would take the coordinate-trick path
//      (_batched_nms_coordinate_trick), and the segment -> global-rank map;
//   4. coordinate trick only: detect cross-class suppression edges (possible when a box with
//      negative x1,y1 is shifted onto the previous class band); such an image is re-emitted as
//      ONE class-agnostic segment so the greedy scan sees exactly what torchvision's does.
#include "ypp_kernels.cuh"

namespace {

#define YPP_DET_LIST 64

struct DetEntry { float4 box; int cls; int rank; };

#define YPP_DET_BINS 64

struct OrgSmemHeader {
    DetEntry lowl[YPP_DET_LIST];
    DetEntry highl[YPP_DET_LIST];
    u32 bin_hi[YPP_DET_BINS];     // A role: per x-bin max of enc(y2 - step); after the scan: suffix max
    u32 bin_lo[YPP_DET_BINS];     // B role: per x-bin max of enc(-y1);      after the scan: prefix max
    int n_low, n_high, flagged, pad;
};

// Cross-class suppression under the coordinate trick needs a pair (A of class c, B of class
// c' > c) whose SHIFTED boxes intersect: x1_b + k*step < x2_a and y1_b + k*step < y2_a with
// k >= 1, step = max_coord + 1.  With u_a = x2_a - step, v_a = y2_a - step (both <= -1) that is a
// 2-D dominance  x1_b < u_a  and  y1_b < v_a.  A 64-bin staircase over x (suffix max of v for
// the A role, prefix min of y1 for the B role) leaves only boxes near the Pareto frontier as
// candidates; those few are then tested pairwise with the exact shifted arithmetic.
struct DetGeom {
    float step, x1min, lim, slack, inv_w;
    bool on;
};

__device__ __forceinline__ int det_bin(const DetGeom& g, float x)
{
    const float t = (x - g.x1min) * g.inv_w;
    return min(max((int)t, 0), YPP_DET_BINS - 1);
}

__device__ __forceinline__ void ce(u64* a, int i, int l)
{
    const u64 x = a[i], y = a[l];
    if (x > y) { a[i] = y; a[l] = x; }
}

__device__ __forceinline__ u64 shfl_xor_u64(u64 v, int m)
{
    return ((u64)__shfl_xor_sync(YPP_FULL_MASK, (u32)(v >> 32), m) << 32) | __shfl_xor_sync(YPP_FULL_MASK, (u32)v, m);
}

// compare-exchange across lanes: the lane whose `upper` is false keeps the minimum
__device__ __forceinline__ void ce_lane(u64& x, int m, bool upper)
{
    const u64 o = shfl_xor_u64(x, m);
    x = (upper == (x > o)) ? x : o;       // upper keeps max, lower keeps min
}

// last six steps (strides 32,16,..,1) of an ascending bitonic merge on a 64-element window held
// two per lane: x0 = w[lane], x1 = w[lane + 32]
__device__ __forceinline__ void merge64_tail(u64& x0, u64& x1, int lane)
{
    if (x0 > x1) { const u64 t = x0; x0 = x1; x1 = t; }
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        const bool up = (lane & m) != 0;
        ce_lane(x0, m, up);
        ce_lane(x1, m, up);
    }
}

// full ascending bitonic sort of a 64-element window in registers (merge sizes 2..64, every
// merge = mirror step + halving steps, all compare-exchanges ascending)
__device__ __forceinline__ void sort64(u64& x0, u64& x1, int lane)
{
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
        // mirror step inside each lane half: partner index = i ^ (k-1)  (lane ^ (k-1))
        const bool up = (lane & (k >> 1)) != 0;
        ce_lane(x0, k - 1, up);
        ce_lane(x1, k - 1, up);
#pragma unroll
        for (int m = k >> 2; m > 0; m >>= 1) {
            const bool u2 = (lane & m) != 0;
            ce_lane(x0, m, u2);
            ce_lane(x1, m, u2);
        }
    }
    // k = 64 mirror: element lane of the low half pairs with element (31 - lane) of the high half
    {
        const u64 o1 = shfl_xor_u64(x1, 31);   // value x1 of lane 31-lane
        const u64 o0 = shfl_xor_u64(x0, 31);   // value x0 of lane 31-lane
        // my x0 (index lane) vs partner index 63-lane = x1 of lane (31-lane): keep min
        const u64 n0 = x0 < o1 ? x0 : o1;
        // my x1 (index 32+lane) vs partner index 31-lane = x0 of lane (31-lane): keep max
        const u64 n1 = x1 > o0 ? x1 : o0;
        x0 = n0; x1 = n1;
    }
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        const bool up = (lane & m) != 0;
        ce_lane(x0, m, up);
        ce_lane(x1, m, up);
    }
}

// Bitonic network with every compare-exchange ascending (mirror first step), so a virtual
// +inf padding beyond n never moves: pairs whose upper index is >= n are simply skipped in
// shared memory, and windows are padded with ~0 in registers.  Strides >= 64 go through
// shared memory with a block barrier per step; strides <= 32 run in registers on 64-element
// windows (one warp per window) with shuffles.
__device__ void bitonic_sort(u64* a, int n)
{
    int n_pad = 64, lg = 6;
    while (n_pad < n) { n_pad <<= 1; ++lg; }
    const int half = n_pad >> 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_win = n_pad >> 6;
    const u64 INF = ~0ull;
    // phase 0: sort every 64-window
    for (int w = warp; w < n_win; w += YPP_ORG_THREADS / 32) {
        const int i0 = (w << 6) + lane, i1 = i0 + 32;
        u64 x0 = i0 < n ? a[i0] : INF, x1 = i1 < n ? a[i1] : INF;
        sort64(x0, x1, lane);
        if (i0 < n) a[i0] = x0;
        if (i1 < n) a[i1] = x1;
    }
    __syncthreads();
    for (int lk = 7; lk <= lg; ++lk) {
        const int k = 1 << lk, hk = k >> 1;
        // mirror step: pair r of block blk: (blk*k + r, blk*k + k-1-r)
        for (int t = tid; t < half; t += YPP_ORG_THREADS) {
            const int blk = t >> (lk - 1), r = t & (hk - 1);
            const int i = (blk << lk) + r, l = (blk << lk) + (k - 1 - r);
            if (l < n) ce(a, i, l);
        }
        __syncthreads();
        for (int lj = lk - 2; lj >= 6; --lj) {
            const int j = 1 << lj;
            for (int t = tid; t < half; t += YPP_ORG_THREADS) {
                const int blk = t >> lj, r = t & (j - 1);
                const int i = (blk << (lj + 1)) + r, l = i + j;
                if (l < n) ce(a, i, l);
            }
            __syncthreads();
        }
        // strides 32..1 in registers
        for (int w = warp; w < n_win; w += YPP_ORG_THREADS / 32) {
            const int i0 = (w << 6) + lane, i1 = i0 + 32;
            if ((w << 6) >= n) continue;
            u64 x0 = i0 < n ? a[i0] : INF, x1 = i1 < n ? a[i1] : INF;
            merge64_tail(x0, x1, lane);
            if (i0 < n) a[i0] = x0;
            if (i1 < n) a[i1] = x1;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(YPP_ORG_THREADS, 1)
ypp_organise_kernel(const __grid_constant__ YppOrgParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    OrgSmemHeader* hd = reinterpret_cast<OrgSmemHeader*>(smem_raw);
    int* table = reinterpret_cast<int*>(smem_raw + sizeof(OrgSmemHeader));   // [32][nc]
    int* seg = table + 32 * p.nc;                                            // [nc+1]
    size_t koff = sizeof(OrgSmemHeader) + sizeof(int) * (size_t)(33 * p.nc + 1);
    koff = (koff + 15) / 16 * 16;

    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nc = p.nc, N = p.N;
    const int n = min(p.ws.cand_count[b], N);
    u64* gkeys = p.ws.keys + (size_t)b * N;
    u64* keys = p.keys_in_smem ? reinterpret_cast<u64*>(smem_raw + koff) : gkeys;
    int* gseg = p.ws.seg_start + (size_t)b * (nc + 1);

    if (n == 0) {
        for (int c = tid; c <= nc; c += YPP_ORG_THREADS) gseg[c] = 0;
        return;
    }
    if (p.keys_in_smem) {
        for (int i = tid; i < n; i += YPP_ORG_THREADS) keys[i] = gkeys[i];
    }
    if (tid == 0) { hd->n_low = 0; hd->n_high = 0; hd->flagged = 0; }
    if (tid < YPP_DET_BINS) { hd->bin_hi[tid] = 0; hd->bin_lo[tid] = 0; }
    for (int i = tid; i < 32 * nc; i += YPP_ORG_THREADS) table[i] = 0;
    unsigned char* kf = p.ws.keepflag + (size_t)b * N;
    for (int i = tid; i < n; i += YPP_ORG_THREADS) kf[i] = 0;
    __syncthreads();

    bitonic_sort(keys, n);       // ends with __syncthreads()

    if (p.keys_in_smem) {
        for (int i = tid; i < n; i += YPP_ORG_THREADS) gkeys[i] = keys[i];
    }

    // ---- which torchvision path would run for this image
    bool trick = p.offset_mode == YPP_OFFSET_TRICK;
    if (p.offset_mode == YPP_OFFSET_AUTO_CPU) trick = 4LL * n <= 4000;
    if (p.offset_mode == YPP_OFFSET_AUTO_CUDA) trick = 4LL * n <= 100000;
    const float M = ypp_dec(p.ws.maxc[b]);
    const float step = __fadd_rn(M, 1.0f);                 // max_coordinate + 1
    DetGeom dg;
    dg.step = step;
    dg.x1min = -ypp_dec(p.ws.nminx1[b]);
    dg.slack = fmaf(fabsf(step) * (float)nc, 1e-6f, 1e-3f);
    dg.lim = -1.0f + dg.slack;
    // no detector when offsets are not finite (every shifted IoU is NaN on both paths) or when no
    // box reaches x1 < -1 (nothing can be shifted onto the previous class band)
    dg.on = trick && nc > 1 && (fabsf(M) <= 3.0e38f) && dg.x1min < dg.lim;
    {
        const float range = dg.lim - dg.x1min;
        dg.inv_w = (range > 64.0f * 4.0f * dg.slack) ? (float)YPP_DET_BINS / range : 0.0f;
    }
    const bool detect = dg.on;
    if (trick && nc > 1 && !(step > 0.0f) && tid == 0) hd->flagged = 1;   // degenerate offsets: be exact

    const float4* gbox = p.ws.box + (size_t)b * N;

    // ---- stable partition by class: warp w owns ranks [w*L, (w+1)*L)
    const int L = ((n + 31) / 32 + 31) / 32 * 32;
    const int r_begin = warp * L, r_end = min(n, r_begin + L);
    int* my = table + warp * nc;
    for (int r0 = r_begin; r0 < r_end; r0 += 32) {
        const int r = r0 + lane;
        const bool ok = r < r_end;
        u64 k = 0;
        if (ok) k = keys[r];
        const int c = ok ? ypp_key_cls(k) : -1 - lane;
        const unsigned peers = __match_any_sync(YPP_FULL_MASK, c);
        if (ok && (peers & ((1u << lane) - 1u)) == 0) my[c] += __popc(peers);
        if (detect && ok) {
            const float4 xy = ypp_xyxy(gbox[ypp_key_idx(k)]);
            const float u = xy.z - step, v = xy.w - step;
            if (u > dg.x1min - dg.slack) atomicMax(&hd->bin_hi[det_bin(dg, u)], ypp_enc(v));
            if (xy.x < dg.lim && xy.y < dg.lim) atomicMax(&hd->bin_lo[det_bin(dg, xy.x)], ypp_enc(-xy.y));
        }
        __syncwarp();
    }
    __syncthreads();
    if (detect && warp == 1) {
        // suffix max over bin_hi (A's in bins >= j can dominate a B in bin j),
        // prefix max over bin_lo (B's in bins <= i can be dominated by an A in bin i); 2 bins per lane
        u32 h0 = hd->bin_hi[lane], h1 = hd->bin_hi[lane + 32];
        u32 l0 = hd->bin_lo[lane], l1 = hd->bin_lo[lane + 32];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 a1 = __shfl_down_sync(YPP_FULL_MASK, h1, d), a0 = __shfl_down_sync(YPP_FULL_MASK, h0, d);
            if (lane + d < 32) { h1 = max(h1, a1); h0 = max(h0, a0); }
            const u32 b0 = __shfl_up_sync(YPP_FULL_MASK, l0, d), b1 = __shfl_up_sync(YPP_FULL_MASK, l1, d);
            if (lane >= d) { l0 = max(l0, b0); l1 = max(l1, b1); }
        }
        h0 = max(h0, __shfl_sync(YPP_FULL_MASK, h1, 0));        // bins 0..31 also see all of 32..63
        l1 = max(l1, __shfl_sync(YPP_FULL_MASK, l0, 31));       // bins 32..63 also see all of 0..31
        hd->bin_hi[lane] = h0; hd->bin_hi[lane + 32] = h1;
        hd->bin_lo[lane] = l0; hd->bin_lo[lane + 32] = l1;
    }
    for (int c = tid; c < nc; c += YPP_ORG_THREADS) {
        int t = 0;
        for (int w = 0; w < 32; ++w) t += table[w * nc + c];
        seg[c] = t;
    }
    __syncthreads();
    if (warp == 0) {   // exclusive scan of seg[0..nc) -> seg[0..nc]
        int run = 0;
        for (int c0 = 0; c0 < nc; c0 += 32) {
            const int c = c0 + lane;
            const int v = c < nc ? seg[c] : 0;
            int inc = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(YPP_FULL_MASK, inc, d);
                if (lane >= d) inc += o;
            }
            if (c < nc) seg[c] = run + inc - v;
            run += __shfl_sync(YPP_FULL_MASK, inc, 31);
        }
        if (lane == 0) seg[nc] = run;
    }
    __syncthreads();
    for (int c = tid; c < nc; c += YPP_ORG_THREADS) {
        int run = seg[c];
        for (int w = 0; w < 32; ++w) { const int t = table[w * nc + c]; table[w * nc + c] = run; run += t; }
    }
    __syncthreads();

    float4* sbox = p.ws.sbox + (size_t)b * N;
    int* grank = p.ws.grank + (size_t)b * N;

    for (int r0 = r_begin; r0 < r_end; r0 += 32) {
        const int r = r0 + lane;
        const bool ok = r < r_end;
        u64 k = 0;
        int c = -1 - lane;
        if (ok) { k = keys[r]; c = ypp_key_cls(k); }
        const unsigned peers = __match_any_sync(YPP_FULL_MASK, c);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        int q = 0;
        if (ok) q = my[c] + rank;
        __syncwarp();
        if (ok && rank == 0) my[c] += __popc(peers);
        __syncwarp();
        if (ok) {
            const float4 cb = gbox[ypp_key_idx(k)];
            float4 xy = ypp_xyxy(cb);
            if (detect) {
                // B role: some A in an x-bin >= mine has v_a > y1 ; A role: some B in an x-bin <= mine has y1 < v_a
                bool is_low = false, is_high = false;
                if (xy.x < dg.lim && xy.y < dg.lim) {
                    const int j = max(det_bin(dg, xy.x) - 1, 0);
                    const u32 e = hd->bin_hi[j];
                    is_low = e != 0 && xy.y < ypp_dec(e) + dg.slack;
                }
                const float u = xy.z - step, v = xy.w - step;
                if (u > dg.x1min - dg.slack) {
                    const int i = min(det_bin(dg, u) + 1, YPP_DET_BINS - 1);
                    const u32 e = hd->bin_lo[i];
                    is_high = e != 0 && v > -ypp_dec(e) - dg.slack;
                }
                if (is_low || is_high) {
                    const float off = __fmul_rn((float)c, step);
                    DetEntry e;
                    e.box = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
                    e.cls = c; e.rank = r;
                    if (is_low) { const int s = atomicAdd(&hd->n_low, 1); if (s < YPP_DET_LIST) hd->lowl[s] = e; }
                    if (is_high) { const int s = atomicAdd(&hd->n_high, 1); if (s < YPP_DET_LIST) hd->highl[s] = e; }
                }
            }
            if (trick) {
                const float off = __fmul_rn((float)c, step);    // idxs.to(boxes) * (max_coordinate + 1)
                xy = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
            }
            sbox[q] = xy;
            grank[q] = r;
        }
    }
    __syncthreads();

    if (trick && nc > 1) {
        const int nl = hd->n_low, nh = hd->n_high;
        if (nl > 0 && nh > 0) {
            if (nl > YPP_DET_LIST || nh > YPP_DET_LIST) {
                if (tid == 0) hd->flagged = 1;
            } else {
                const bool neg_thr = !(p.iou_thr >= 0.0f);
                for (int t = tid; t < nl * nh; t += YPP_ORG_THREADS) {
                    const DetEntry lo = hd->lowl[t / nh], hi = hd->highl[t % nh];
                    if (lo.cls == hi.cls) continue;
                    const DetEntry& row = lo.rank < hi.rank ? lo : hi;
                    const DetEntry& col = lo.rank < hi.rank ? hi : lo;
                    const float sa = ypp_area(row.box);
                    const bool s = neg_thr ? ypp_suppresses<false>(row.box, sa, col.box, p.iou_thr, p.union_mode)
                                           : ypp_suppresses<true>(row.box, sa, col.box, p.iou_thr, p.union_mode);
                    if (s) hd->flagged = 1;
                }
            }
        }
        __syncthreads();
        if (hd->flagged) {
            // class-agnostic single segment in global score order, shifted coordinates
            for (int r = tid; r < n; r += YPP_ORG_THREADS) {
                const u64 k = keys[r];
                const float off = __fmul_rn((float)ypp_key_cls(k), step);
                const float4 xy = ypp_xyxy(gbox[ypp_key_idx(k)]);
                sbox[r] = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
                grank[r] = r;
            }
            for (int c = tid; c <= nc; c += YPP_ORG_THREADS) seg[c] = c == 0 ? 0 : n;
            if (tid == 0) p.ws.imgflag[b] = 1;
            __syncthreads();
        }
    }

    for (int c = tid; c <= nc; c += YPP_ORG_THREADS) gseg[c] = seg[c];
    for (int c = tid; c < nc; c += YPP_ORG_THREADS) {
        if (seg[c + 1] - seg[c] > YPP_WARP_SEG) {
            const int s = atomicAdd(p.ws.q, 1);
            p.ws.bigq[s] = b * nc + c;
        }
    }
}

}  // namespace

int ypp_org_smem_bytes(int N, int nc, int* keys_in_smem)
{
    size_t fixed = sizeof(OrgSmemHeader) + sizeof(int) * (size_t)(33 * nc + 1);
    fixed = (fixed + 15) / 16 * 16;
    const size_t cap = 227 * 1024;
    if (fixed > cap) return -1;
    if (fixed + (size_t)N * 8 <= cap) { *keys_in_smem = 1; return (int)(fixed + (size_t)N * 8); }
    *keys_in_smem = 0;
    return (int)fixed;
}

int ypp_launch_organise(const YppOrgParams& p, cudaStream_t st)
{
    int kin = 0;
    const int smem = ypp_org_smem_bytes(p.N, p.nc, &kin);
    if (smem < 0) { ypp_set_error("organise: nc=%d needs more shared memory than one CTA has", p.nc); return YPP_E_UNSUPPORTED; }
    YppOrgParams q = p;
    q.keys_in_smem = kin;
    static int configured_smem = 0;
    if (smem > configured_smem) {
        cudaError_t e = cudaFuncSetAttribute(ypp_organise_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) { ypp_set_error("organise: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
        configured_smem = 227 * 1024;
    }
    ypp_organise_kernel<<<p.B, YPP_ORG_THREADS, smem, st>>>(q);
    ypp_count_launch(1);
    return ypp_check_launch("organise");
}
