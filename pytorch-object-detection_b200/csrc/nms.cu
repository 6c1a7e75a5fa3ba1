// N1 + N2: bitmask IoU suppression and greedy keep-scan per (image, class) segment, and the
// final row gather.
//
// Reference semantics: torchvision.ops.batched_nms as called by nms_v3
// (utils/yolo_utils.py:346) -> torch.ops.torchvision.nms: boxes visited in descending score
// (stable), box j suppressed iff an earlier KEPT box i has IoU(i,j) > thr (strict).
//
// torchvision's CUDA kernel writes an n x n/64 bit matrix to HBM and scans it with one CTA.
// Here nothing but the keep flags reaches HBM.  Two ways through a segment, same result:
//   - sweep-and-prune (segments of ordinary boxes, see nms_sparse / nms_big_sparse): greedy NMS only
//     needs the edges (i before j, IoU > thr), so the boxes are counting-sorted by x1, every box
//     meets only the later boxes inside the x1 range an IoU above thr allows, the edges go to a
//     list in shared memory and the greedy order is a fix-point over that list;
//   - tile sweeps (anything else, or when the edge list overflows): tiles in score order, the
//     diagonal tile resolved by a ballot fix-point, and only the KEPT rows broadcast against the
//     later, still-alive columns, so suppressed rows never do any pair work.
// Segments of up to YPP_WARP_SEG boxes (the per-class case) are owned by one 4-warp CTA; longer
// ones (single-class heads, classes a cross-class edge merged) go to a 1024-thread CTA kernel
// walking the per-image lists of long segments S1 left.
//
// Pair test.  torchvision evaluates  inter / union > thr  with an IEEE division.  The same
// decision is reached without dividing whenever  d = fma(-thr, union, inter)  is clearly away
// from zero: |d| > 2^-20 * union  leaves 8x more room than the rounding of the quotient, the
// half-ulp of thr and the rounding of d together can consume (valid for 2^-10 <= thr <= 1).
// Pairs inside that band (practically none) are re-evaluated with the exact division, so keep
// sets stay bit-exact.
#include "ypp_kernels.cuh"
#include <cstdlib>

namespace {

#define YPP_PAIR_BAND 9.5367431640625e-07f      /* 2^-20 */

// column-side constants of the pair test: kFma: (w_b, h_b); else (S_b, unused)
template <bool kFma>
__device__ __forceinline__ void pair_col(const float4 b, float& p0, float& p1)
{
    const float wb = __fsub_rn(b.z, b.x), hb = __fsub_rn(b.w, b.y);
    if (kFma) { p0 = wb; p1 = hb; }
    else      { p0 = __fmul_rn(wb, hb); p1 = 0.0f; }
}

// a = row (earlier in score order, area sa), b = column.  dead |= sure suppression,
// amb |= the decision needs the exact division.
template <bool kFma>
__device__ __forceinline__ void pair_fast(const float4 a, const float sa, const float4 b, const float p0,
                                          const float p1, const float thr, bool& dead, bool& amb)
{
    const float l = fmaxf(a.x, b.x), t = fmaxf(a.y, b.y);
    const float r = fminf(a.z, b.z), bt = fminf(a.w, b.w);
    const float w = fmaxf(__fsub_rn(r, l), 0.0f), h = fmaxf(__fsub_rn(bt, t), 0.0f);
    const float inter = __fmul_rn(w, h);
    const float uni = kFma ? __fsub_rn(__fmaf_rn(p0, p1, sa), inter) : __fsub_rn(__fadd_rn(sa, p0), inter);
    const float au = fabsf(uni);                       // inter > 0 implies uni > 0; |.| keeps inter == 0 false
    const float d = __fmaf_rn(-thr, au, inter);
    const float band = __fmul_rn(au, YPP_PAIR_BAND);
    dead |= d > band;
    amb |= !(fabsf(d) > band);                         // also catches NaN / inf
}

__device__ __forceinline__ bool pair_exact(const float4 a, const float sa, const float4 b, const float thr,
                                           const int union_fma)
{
    return ypp_suppresses<false>(a, sa, b, thr, union_fma);
}

// Lean form used by the sweeps of both kernels (the accumulating case: one column against many kept
// rows; NaN values are dropped by fmaxf, which is right: every NaN case has an exact answer of "not
// suppressed").  With c1 = 1 + thr, ta = -thr*S_a, tb = -thr*S_b:
//   inter - thr*(S_a + S_b - inter) = c1*inter + ta + tb,
// so a column only accumulates g = max over rows of fma(c1, inter, ta) and adds its own tb at the end.
// The clamp of the y extent rides on the multiplier: hs = sat(h * 2^-k) (k per segment such that
// every extent * 2^-k <= 1; power-of-two scaling is exact) and c1k = c1 * 2^k undoes it.  inter is
// torchvision's own fp32 value; the union differs from torchvision's by a few roundings of an area,
// |error of g + tb| <= 20 * 2^-24 * (largest |area|), inside the segment's decision band
// 2 * 2^-20 * (largest |area|), so "beyond the band" still decides and "inside" still goes to the
// exact division.  ta, tb use |area|: an inverted box never overlaps anything (inter <= 0), and with
// |area| its pairs can only come out "alive" or "ask the exact form", never "suppressed".
// 5 fma-pipe + 4 (+ half a 3-input max) alu-pipe instructions per pair.
__device__ __forceinline__ float pair_g(const float4 a, const float ta, const float4 b, const float c1k, const float scale)
{
    const float l = fmaxf(a.x, b.x), t = fmaxf(a.y, b.y);
    const float r = fminf(a.z, b.z), bt = fminf(a.w, b.w);
    const float w = __fsub_rn(r, l);
    const float hs = __saturatef(__fmul_rn(__fsub_rn(bt, t), scale));
    return __fmaf_rn(c1k, __fmul_rn(w, hs), ta);
}

// box of one segment position: S1's word for it holds the box index and the class.  cxcywh -> xyxy exactly as nms_v3
// does, plus the coordinate trick's class offset when torchvision would apply it (S1 decided per image)
__device__ __forceinline__ float4 pos_box(const YppWs& ws, size_t ibase, u64 g, bool trick, float step)
{
    float4 xy = ypp_xyxy(ws.box[ibase + ypp_key_idx(g)]);
    if (trick) {
        const float off = __fmul_rn((float)ypp_key_cls(g), step);    // idxs.to(boxes) * (max_coordinate + 1)
        xy = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
    }
    return xy;
}

// a row that can never raise dmax (padding of the kept-row lists)
#define YPP_DUMMY_BOX make_float4(1e30f, 1e30f, -1e30f, -1e30f)
#define YPP_DUMMY_TA (-1e30f)

// ------------------------------------------------------------------ per-class segments: one 4-warp CTA each
#define YPP_NMSQ_WARPS 4
#define YPP_NMSQ_THREADS (32 * YPP_NMSQ_WARPS)

// sweep-and-prune path of the per-class kernel (see nms_sparse below)
#define YPP_SP_BINS 256
#define YPP_SP_MIN_LEN 192
#define YPP_SP_EDGES 1024
#define YPP_SP_ROUNDS 24
#define YPP_SP_LONG 32              /* ranges this long are swept by a whole warp (even) */

struct SegDense {
    float tb[YPP_WARP_SEG];             // -thr * |area| of every box (its share of thr * union)
    float4 kbox[YPP_NMSQ_WARPS][8];     // per warp: its kept rows of the current tile, padded to a multiple of 2
    float kta[YPP_NMSQ_WARPS][8];       // -thr * area of those rows
    u32 in_words[32];          // byte w of word j: rows 8w..8w+7 of the tile that suppress column j
    uint4 dpart[32];           // per lane: the four warps' partial dead-column bits
};

struct SegSparse {
    u32 hist[YPP_SP_BINS];              // boxes per x1 bin, then the scatter cursors
    u32 edges[YPP_SP_EDGES];            // (earlier rank << 16) | later rank of every pair with IoU > thr
    float tb[YPP_WARP_SEG];             // -thr * |area| by position
    u32 rowlist[YPP_WARP_SEG];          // position | end of its range << 16, longest range first
    unsigned short bstart[YPP_SP_BINS + 2];     // first position of every bin
    unsigned short rk[YPP_WARP_SEG];    // position (x1 order) -> rank inside the segment (score order)
    unsigned char state[YPP_WARP_SEG];  // by rank: 0 undecided, 1 kept, 2 suppressed
    unsigned char pend[YPP_WARP_SEG];   // some undecided earlier box could still suppress me
    unsigned char kill[YPP_WARP_SEG];   // a kept earlier box suppresses me
    int wtot[YPP_NMSQ_WARPS];
    u32 xlo[YPP_NMSQ_WARPS], xhi[YPP_NMSQ_WARPS];
};

struct SegSmem {
    float4 box[YPP_WARP_SEG];           // score order (tile sweeps) or x1 order (sweep-and-prune)
    union {
        SegDense d;
        SegSparse s;
    };
    u32 amax[YPP_NMSQ_WARPS];
    u32 cmax[YPP_NMSQ_WARPS];
    int nedges;
};

// One warp's kept rows (kb/ka, npad of them, nk real) against G groups of 32 columns starting at
// group k (lane = column mod 32).  Bit g of the lane's `dead >> k` says its column of group k+g needs no
// work (suppressed, or past the end of the segment: those are marked dead up front and hold dummies).
// Returns the bits of the groups whose column in this lane died.
template <int G, bool kFast>
__device__ __forceinline__ u32 sweep_groups(const float4* __restrict__ sbox, const float* __restrict__ stb,
                                            const float4* __restrict__ kb, const float* __restrict__ ka, int npad, int nk,
                                            int k, u32 dead, float c1k, float scale, float Bseg, float thr, int union_fma)
{
    const int lane = threadIdx.x & 31;
    const u32 vm = ~(dead >> k) & ((1u << G) - 1u);
    if (!__any_sync(YPP_FULL_MASK, vm != 0u)) return 0u;
    const int j = (k << 5) + lane;
    float4 c[G];
#pragma unroll
    for (int g = 0; g < G; ++g) c[g] = sbox[j + 32 * g];
    u32 dbits = 0;
    bool exact = !kFast;
    if (kFast) {
        float dm[G];
#pragma unroll
        for (int g = 0; g < G; ++g) dm[g] = -INFINITY;
        for (int r = 0; r < npad; r += 2) {
            const float4 a = kb[r], a2 = kb[r + 1];
            const float ta = ka[r], ta2 = ka[r + 1];
#pragma unroll
            for (int g = 0; g < G; ++g)
                dm[g] = fmaxf(fmaxf(dm[g], pair_g(a, ta, c[g], c1k, scale)), pair_g(a2, ta2, c[g], c1k, scale));
        }
        u32 amb = 0;
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float D = __fadd_rn(dm[g], stb[j + 32 * g]);
            if (D > Bseg) dbits |= 1u << g;
            if (!(fabsf(D) > Bseg)) amb |= 1u << g;               // inside the band, or NaN
        }
        exact = (amb & vm) != 0u;
    }
    if (exact) {                                    // a pair inside the band (or thr out of range)
        dbits = 0;
        for (int r = 0; r < nk; ++r) {
            const float4 a = kb[r];
            const float sa = ypp_area(a);
#pragma unroll
            for (int g = 0; g < G; ++g)
                if (pair_exact(a, sa, c[g], thr, union_fma)) dbits |= 1u << g;
        }
    }
    return (dbits & vm) << k;
}

// ---- sweep-and-prune path -------------------------------------------------------------------
// The tile sweeps below test every (kept row, later column) pair: len^2/2 pair tests per segment, and
// on detector output nearly all of them are between boxes that do not even touch.  Greedy NMS only
// needs the EDGES (i earlier than j, IoU(i, j) > thr); which pairs are looked at, and in what order,
// is free as long as no edge is missed.  So for segments of ordinary boxes (finite, x2 > x1, y2 > y1):
//   1. counting sort of the boxes by x1 into YPP_SP_BINS bins over the segment's x1 range;
//   2. every box p only meets the boxes after it in that order whose bin starts before
//        xe = x1 + (1 - 0.9 thr) * w_p   (every step rounded up).
//      For x1_c >= x1_p:  IoU <= overlap_x / max(w_p, w_c) <= 1 - (x1_c - x1_p) / w_p, so a column
//      beyond xe has IoU <= 0.9 thr in exact arithmetic, which no rounding of the fp32 quotient (a few
//      2^-24 relative; the union never cancels) lifts above thr.  Boxes of p's own bin after p are
//      always met, whatever their x1;
//   3. one thread per box, boxes handed out in order of decreasing range length (a second counting
//      sort), so the lanes of a warp run loops of nearly the same length.  The pair test is the lean
//      form of the tile sweeps (pair_g + the segment's decision band); the rare pair inside the band
//      is decided by the exact division with the EARLIER box as torchvision's row (the fma form of
//      the union is not symmetric);
//   4. the greedy order over the edge list as a fix-point: a box is kept once every earlier box with
//      an edge to it is known suppressed, suppressed once one of them is known kept.
// Anything unusual - a box that is not ordinary, coordinates or areas outside the lean test's range,
// more than YPP_SP_EDGES edges (crowds of overlapping boxes), a dependency chain deeper than
// YPP_SP_ROUNDS - returns false with nothing written, and the segment goes through the tile sweeps.
// On the C2 dense workload (409 boxes per segment, 18 edges) this path does 11 % of the pair tests.
#ifdef YPP_SP_STATS
// tuning build only: 0 attempts, 1 done, 2 not ordinary, 3 edge overflow, 4 too deep, 5 pair tests, 6 edges, 7 rounds,
// 8.. clock cycles of CTA phases (load, sorts, sweep, resolve), 12 sweep iterations of warp 0
__device__ unsigned long long ypp_sp_stats[16];
#define SP_STAT(i, v) atomicAdd(&ypp_sp_stats[i], (unsigned long long)(v))
#define SP_CLOCK(var) const long long var = clock64()
#else
#define SP_STAT(i, v)
#define SP_CLOCK(var)
#endif

template <int kBins>
__device__ __forceinline__ int sp_bin(const float x, const float xmin, const float inv)
{
    const float v = __fmul_rn(__fsub_rn(x, xmin), inv);          // monotone in x; NaN (0 * inf) -> bin 0
    return (int)fminf(fmaxf(v, 0.0f), (float)(kBins - 1));
}
template <int kBins>
__device__ __forceinline__ int sp_bin_end(const float x, const float xmin, const float inv)
{
    const float v = __fmul_rn(__fsub_rn(x, xmin), inv);          // >= sp_bin(x); NaN -> last bin
    return !(v < (float)kBins) ? kBins - 1 : max((int)v, 0);
}

// hist[256] -> exclusive offsets in place (the scatter cursors); optionally kept in bstart as well
__device__ __forceinline__ void sp_scan(SegSparse& s, const int tid, const bool keep_starts, const int len)
{
    const int lane = tid & 31, w = tid >> 5;
    const int h0 = (int)s.hist[2 * tid], h1 = (int)s.hist[2 * tid + 1];
    int incl = h0 + h1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(YPP_FULL_MASK, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s.wtot[w] = incl;
    __syncthreads();
    int off = incl - (h0 + h1);
    for (int k = 0; k < w; ++k) off += s.wtot[k];
    if (keep_starts) {
        s.bstart[2 * tid] = (unsigned short)off;
        s.bstart[2 * tid + 1] = (unsigned short)(off + h0);
        if (tid == 0) s.bstart[YPP_SP_BINS] = (unsigned short)len;
    }
    s.hist[2 * tid] = (u32)off;
    s.hist[2 * tid + 1] = (u32)(off + h0);
    __syncthreads();
}

// one pair of the sweep: box a (position order, rank ra) against the box at position c
template <bool kFma>
__device__ __forceinline__ void sp_pair(SegSmem& sm, const float4 a, const float ta, const int ra, const int c,
                                        const float c1k, const float scale, const float Bseg, const float thr, int& ntests)
{
    const float4 bq = sm.box[c];
#ifdef YPP_SP_STATS
    ++ntests;
#endif
    const float D = __fadd_rn(pair_g(a, ta, bq, c1k, scale), sm.s.tb[c]);
    if (D < -Bseg) return;                              // surely not suppressed (no NaN: everything is finite)
    const int rb = sm.s.rk[c];
    bool dead = D > Bseg;
    if (!dead) {                                        // inside the band: the exact division decides
        const bool a_first = ra < rb;
        const float4 hi = a_first ? a : bq, lo = a_first ? bq : a;
        dead = pair_exact(hi, ypp_area(hi), lo, thr, kFma ? 1 : 0);
    }
    if (dead) {
        const int e = atomicAdd(&sm.nedges, 1);
        if (e < YPP_SP_EDGES) sm.s.edges[e] = ((u32)min(ra, rb) << 16) | (u32)max(ra, rb);
    }
}

template <bool kFma>
__device__ __noinline__ bool nms_sparse(SegSmem& sm, const YppWs& ws, const size_t ibase, const u64* __restrict__ gpk,
                                        unsigned char* __restrict__ keepflag, const int len, const int b,
                                        const bool trick, const float step, const float thr)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    SegSparse& S = sm.s;
    SP_CLOCK(t0);
    if (tid == 0) SP_STAT(0, 1);
    // ---- boxes in score order -> registers; ordinary? x1 range, largest |area| and |coordinate|
    float4 q[4];
    bool odd = false;
    u32 xlo = 0xffffffffu, xhi = 0u, amax_bits = 0u, cmax_bits = 0u;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int i = tid + u * YPP_NMSQ_THREADS;
        q[u] = make_float4(0.f, 0.f, 1.f, 1.f);
        if (i < len) {
            q[u] = pos_box(ws, ibase, gpk[i], trick, step);
            const float bw = __fsub_rn(q[u].z, q[u].x), bh = __fsub_rn(q[u].w, q[u].y);
            const bool ok = bw > 0.0f && bh > 0.0f && bw < INFINITY && bh < INFINITY &&
                            fabsf(q[u].x) < INFINITY && fabsf(q[u].y) < INFINITY;
            odd |= !ok;
            const u32 e = ypp_enc(q[u].x);
            xlo = min(xlo, e); xhi = max(xhi, e);
            amax_bits = max(amax_bits, __float_as_uint(fabsf(__fmul_rn(bw, bh))));
            cmax_bits = max(max(cmax_bits, __float_as_uint(fabsf(q[u].x))), max(__float_as_uint(fabsf(q[u].y)),
                            max(__float_as_uint(fabsf(q[u].z)), __float_as_uint(fabsf(q[u].w)))));
        }
    }
    xlo = __reduce_min_sync(YPP_FULL_MASK, xlo);
    xhi = __reduce_max_sync(YPP_FULL_MASK, xhi);
    amax_bits = __reduce_max_sync(YPP_FULL_MASK, amax_bits);
    cmax_bits = __reduce_max_sync(YPP_FULL_MASK, cmax_bits);
    if (lane == 0) { S.xlo[w] = xlo; S.xhi[w] = xhi; sm.amax[w] = amax_bits; sm.cmax[w] = cmax_bits; }
    for (int i = tid; i < YPP_SP_BINS; i += YPP_NMSQ_THREADS) S.hist[i] = 0;
    if (tid == 0) sm.nedges = 0;
    if (__syncthreads_or(odd)) { if (tid == 0) SP_STAT(2, 1); return false; }
    // the lean pair test's constants, exactly as the tile sweeps derive them
    const float Bseg = 2.0f * YPP_PAIR_BAND *
        __uint_as_float(max(max(sm.amax[0], sm.amax[1]), max(sm.amax[2], sm.amax[3])));
    const int cexp = (int)(max(max(sm.cmax[0], sm.cmax[1]), max(sm.cmax[2], sm.cmax[3])) >> 23);
    if (cexp < 67 || cexp > 187 || !(Bseg < INFINITY)) { if (tid == 0) SP_STAT(2, 1); return false; }    // uniform
    const float scale = __uint_as_float((u32)(252 - cexp) << 23);
    const float c1k = __fmul_rn(__fadd_rn(1.0f, thr), __uint_as_float((u32)(cexp + 2) << 23));
    const float xmin = ypp_dec(min(min(S.xlo[0], S.xlo[1]), min(S.xlo[2], S.xlo[3])));
    const float xmax = ypp_dec(max(max(S.xhi[0], S.xhi[1]), max(S.xhi[2], S.xhi[3])));
    const float inv = __fdiv_rn((float)YPP_SP_BINS, __fsub_rn(xmax, xmin));      // inf when every x1 is the same: one bin
    SP_CLOCK(t1);
    // ---- counting sort by x1 bin
    int mybin[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        mybin[u] = sp_bin<YPP_SP_BINS>(q[u].x, xmin, inv);
        if (tid + u * YPP_NMSQ_THREADS < len) atomicAdd(&S.hist[mybin[u]], 1u);
    }
    __syncthreads();
    sp_scan(S, tid, true, len);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int i = tid + u * YPP_NMSQ_THREADS;
        if (i < len) {
            const int pos = (int)atomicAdd(&S.hist[mybin[u]], 1u);
            sm.box[pos] = q[u];
            S.tb[pos] = __fmul_rn(-thr, fabsf(ypp_area(q[u])));
            S.rk[pos] = (unsigned short)i;
        }
        if (i < YPP_WARP_SEG) { S.state[i] = 0; S.pend[i] = 0; S.kill[i] = 0; }
    }
    __syncthreads();
    // ---- range of every box; boxes in order of decreasing range length
    for (int i = tid; i < YPP_SP_BINS; i += YPP_NMSQ_THREADS) S.hist[i] = 0;
    __syncthreads();
    const float ct = __fsub_ru(1.0f, __fmul_rd(0.9f, thr));
    u32 myrow[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int pr = tid + u * YPP_NMSQ_THREADS;
        mybin[u] = YPP_SP_BINS - 1;
        myrow[u] = 0;
        if (pr < len) {
            const float4 a = sm.box[pr];
            const float xe = __fadd_ru(a.x, __fmul_ru(ct, __fsub_ru(a.z, a.x)));
            const int cend = S.bstart[sp_bin_end<YPP_SP_BINS>(xe, xmin, inv) + 1];
            const int cnt = max(cend - pr - 1, 0);
            mybin[u] = YPP_SP_BINS - 1 - min(cnt >> 1, YPP_SP_BINS - 1);
            myrow[u] = (u32)pr | ((u32)cend << 16);
            atomicAdd(&S.hist[mybin[u]], 1u);
        }
    }
    __syncthreads();
    sp_scan(S, tid, false, len);
#pragma unroll
    for (int u = 0; u < 4; ++u)
        if (tid + u * YPP_NMSQ_THREADS < len) S.rowlist[atomicAdd(&S.hist[mybin[u]], 1u)] = myrow[u];
    __syncthreads();
    SP_CLOCK(t2);
    // ---- sweep.  Ranges of YPP_SP_LONG boxes and more: one warp per box, lanes over its range (the few
    // very wide boxes meet most of the segment; handed to single threads they would be the whole run time);
    // the others: one thread per box
    {
        int ntests = 0;
        const int nlong = (int)S.hist[YPP_SP_BINS - 1 - YPP_SP_LONG / 2];      // cursor of that bin after the scatter = its end
        for (int k = w; k < nlong; k += YPP_NMSQ_WARPS) {
            const u32 row = S.rowlist[k];
            const int pr = (int)(row & 0xffffu), cend = (int)(row >> 16);
            const float4 a = sm.box[pr];
            const float ta = S.tb[pr];
            const int ra = S.rk[pr];
            for (int c = pr + 1 + lane; c < cend; c += 32)
                sp_pair<kFma>(sm, a, ta, ra, c, c1k, scale, Bseg, thr, ntests);
        }
        for (int k = nlong + tid; k < len; k += YPP_NMSQ_THREADS) {
            const u32 row = S.rowlist[k];
            const int pr = (int)(row & 0xffffu), cend = (int)(row >> 16);
            if (pr + 1 >= cend) continue;
            const float4 a = sm.box[pr];
            const float ta = S.tb[pr];
            const int ra = S.rk[pr];
            for (int c = pr + 1; c < cend; ++c)
                sp_pair<kFma>(sm, a, ta, ra, c, c1k, scale, Bseg, thr, ntests);
            if (*(volatile int*)&sm.nedges >= YPP_SP_EDGES) break;
        }
#ifdef YPP_SP_STATS
        SP_STAT(5, ntests);
#endif
    }
    __syncthreads();
    SP_CLOCK(t3);
    const int ne = sm.nedges;
    if (ne > YPP_SP_EDGES) { if (tid == 0) SP_STAT(3, 1); return false; }
    if (tid == 0) SP_STAT(6, ne);
    // ---- greedy order over the edges
    bool settled = ne == 0;
    for (int round = 0; !settled && round < YPP_SP_ROUNDS; ++round) {
        for (int e = tid; e < ne; e += YPP_NMSQ_THREADS) {
            const u32 ed = S.edges[e];
            const int st = S.state[ed >> 16];
            if (st == 1) S.kill[ed & 0xffffu] = 1;
            else if (st == 0) S.pend[ed & 0xffffu] = 1;
        }
        __syncthreads();
        bool open = false;
        for (int i = tid; i < len; i += YPP_NMSQ_THREADS) {
            if (S.state[i] == 0) {
                if (S.kill[i]) S.state[i] = 2;
                else if (!S.pend[i]) S.state[i] = 1;
                else open = true;
            }
            S.pend[i] = 0;
        }
        settled = !__syncthreads_or(open);
        if (tid == 0) SP_STAT(7, 1);
    }
    if (!settled) { if (tid == 0) SP_STAT(4, 1); return false; }
#ifdef YPP_SP_STATS
    if (tid == 0) { SP_STAT(1, 1); SP_STAT(8, t1 - t0); SP_STAT(9, t2 - t1); SP_STAT(10, t3 - t2); SP_STAT(11, clock64() - t3); }
#endif
    int kept = 0;
    for (int i = tid; i < len; i += YPP_NMSQ_THREADS)
        if (S.state[i] != 2) { keepflag[ypp_gpk_rank(gpk[i])] = 1; ++kept; }      // ne == 0: every state is still 0
    kept = __reduce_add_sync(YPP_FULL_MASK, kept);
    if (lane == 0 && kept) atomicAdd(ws.keep_count + b, kept);
    return true;
}

// 32-row tiles in score order.  Warp w owns rows 8w..8w+7 of every tile: it tests them against the
// tile's own columns (diagonal), the four byte-wide partial masks are exchanged through shared
// memory, every warp resolves the greedy order redundantly (ballot fix-point: a row is kept once
// every earlier row that could suppress it is known dead, dead once one of them is known kept),
// and then broadcasts ITS kept rows against all later columns (lane = column mod 32).
#ifndef YPP_NMSQ_MIN_BLOCKS
#define YPP_NMSQ_MIN_BLOCKS 9       /* resident CTAs per SM the register allocation is capped for */
#endif
template <bool kFast, bool kFma>
__global__ void __launch_bounds__(YPP_NMSQ_THREADS, YPP_NMSQ_MIN_BLOCKS)
ypp_nms_seg_kernel(const __grid_constant__ YppNmsParams p)
{
    __shared__ SegSmem sm;
    ypp_grid_wait();
    const int c = blockIdx.x, b = blockIdx.y;
    const int* seg_start = p.ws.seg_start + (size_t)b * (p.nc + 1);
    const int s0 = seg_start[c], len = seg_start[c + 1] - s0;
    if (len <= 0 || len > YPP_WARP_SEG) return;
    const size_t ibase = (size_t)b * p.N;
    const u64* __restrict__ gpk = p.ws.gpk + ibase + s0;
    unsigned char* __restrict__ keepflag = p.ws.keepflag + ibase;
    const float thr = p.iou_thr;
    const int union_fma = p.union_mode;
    const bool trick = (p.ws.imgflag[b] & 2) != 0;
    const float step = p.ws.img_step[b];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (kFast && thr >= 0.0f && len >= p.sparse_min_len) {        // uniform over the CTA
        if (nms_sparse<kFma>(sm, p.ws, ibase, gpk, keepflag, len, b, trick, step, thr)) return;
        __syncthreads();
    }
    u32 amax_bits = 0, cmax_bits = 0;   // uint order == float order for non-negative floats; NaN sorts on top
    for (int i = tid; i < len; i += YPP_NMSQ_THREADS) {
        const float4 q = pos_box(p.ws, ibase, gpk[i], trick, step);
        const float ar = fabsf(ypp_area(q));
        sm.box[i] = q;
        sm.d.tb[i] = __fmul_rn(-thr, ar);
        amax_bits = max(amax_bits, __float_as_uint(ar));
        cmax_bits = max(max(cmax_bits, __float_as_uint(fabsf(q.x))), max(__float_as_uint(fabsf(q.y)),
                        max(__float_as_uint(fabsf(q.z)), __float_as_uint(fabsf(q.w)))));
    }
    amax_bits = __reduce_max_sync(YPP_FULL_MASK, amax_bits);
    cmax_bits = __reduce_max_sync(YPP_FULL_MASK, cmax_bits);
    if (lane == 0) { sm.amax[w] = amax_bits; sm.cmax[w] = cmax_bits; }
    // the last tile is filled up with dummies (never a valid column: marked dead below; never a live row)
    for (int i = len + tid; i < ((len + 31) & ~31); i += YPP_NMSQ_THREADS) { sm.box[i] = YPP_DUMMY_BOX; sm.d.tb[i] = YPP_DUMMY_TA; }
    __syncthreads();
    // one decision band for the whole segment (NaN / inf areas make every compare fail -> exact path)
    float Bseg = 2.0f * YPP_PAIR_BAND *
        __uint_as_float(max(max(sm.amax[0], sm.amax[1]), max(sm.amax[2], sm.amax[3])));
    // extents are below 2^(E-125) when the largest |coordinate| has biased exponent E: scale = 2^-(E-125)
    const int cexp = (int)(max(max(sm.cmax[0], sm.cmax[1]), max(sm.cmax[2], sm.cmax[3])) >> 23);
    if (cexp < 67 || cexp > 187) Bseg = INFINITY;        // coordinates beyond 2^+-60 (or not finite): exact path only
    const int cex = min(max(cexp, 67), 187);
    const float scale = __uint_as_float((u32)(252 - cex) << 23);
    const float c1k = __fmul_rn(__fadd_rn(1.0f, thr), __uint_as_float((u32)(cex + 2) << 23));
    const int ntiles = (len + 31) >> 5;
    const u32 wmask = 0xffu << (8 * w);
    // bit k: column lane + 32k is suppressed (merged over the four warps) or lies past the end of the segment
    u32 dead = ((ntiles - 1) << 5) + lane >= len ? 1u << (ntiles - 1) : 0u;
    int total = 0;
    for (int T = 0; T < ntiles; ++T) {
        const int base = T << 5;
        const int j = base + lane;
        const bool alive_l = !((dead >> T) & 1u);
        const u32 alive = __ballot_sync(YPP_FULL_MASK, alive_l);       // identical in all four warps
        if (alive == 0) continue;
        // ---- diagonal tile, column view: which of MY rows (i < lane, alive) suppress column lane
        const float4 cb = sm.box[j];
        const float my_ta = sm.d.tb[j];
        u32 in = 0;
        if (kFast) {
            // my 8 rows, unconditionally (dead rows and the dummies past the end are masked out below)
            float lo = INFINITY;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int i = base + 8 * w + q;
                const float d = __fadd_rn(pair_g(sm.box[i], sm.d.tb[i], cb, c1k, scale), my_ta);
                if (d > Bseg) in |= 1u << (8 * w + q);
                lo = fminf(lo, fabsf(d));             // a NaN d is never a suppression: dropped
            }
            if (!(lo > Bseg)) {                       // some pair inside the band (or a NaN): exact pass
                in = 0;
                for (int q = 0; q < 8; ++q) {
                    const int i = base + 8 * w + q;
                    if (pair_exact(sm.box[i], ypp_area(sm.box[i]), cb, thr, union_fma)) in |= 1u << (8 * w + q);
                }
            }
        } else {
            for (int q = 0; q < 8; ++q) {
                const int i = base + 8 * w + q;
                if (pair_exact(sm.box[i], ypp_area(sm.box[i]), cb, thr, union_fma)) in |= 1u << (8 * w + q);
            }
        }
        in &= alive & ((1u << lane) - 1u);            // only alive rows before me count
        reinterpret_cast<unsigned char*>(sm.d.in_words)[lane * 4 + w] = (unsigned char)(in >> (8 * w));
        __syncthreads();
        in = sm.d.in_words[lane];
        // ---- greedy order as a fix-point
        u32 K = 0, D = ~alive;
        bool undecided = alive_l;
        while ((K | D) != 0xffffffffu) {
            const bool dead_now = undecided && (in & K) != 0u;
            const bool keep_now = undecided && !dead_now && (in & ~D) == 0u;
            K |= __ballot_sync(YPP_FULL_MASK, keep_now);
            D |= __ballot_sync(YPP_FULL_MASK, dead_now);
            undecided = undecided && !(dead_now || keep_now);
        }
        total += __popc(K);
        if (w == 0 && ((K >> lane) & 1u)) keepflag[ypp_gpk_rank(gpk[j])] = 1;
        if (T + 1 >= ntiles) break;
        // ---- broadcast my kept rows against the later columns (four column groups per sweep)
        const u32 Kw = K & wmask;
        const int nkw = __popc(Kw);
        u32 part = 0;
        if (nkw) {
            const int npad = (nkw + 1) & ~1;
            if ((Kw >> lane) & 1u) {
                const int s = __popc(Kw & ((1u << lane) - 1u));
                sm.d.kbox[w][s] = cb;
                sm.d.kta[w][s] = my_ta;
            }
            if (lane >= nkw && lane < npad) { sm.d.kbox[w][lane] = YPP_DUMMY_BOX; sm.d.kta[w][lane] = YPP_DUMMY_TA; }
            __syncwarp();
            const float4* kb = sm.d.kbox[w];
            const float* ka = sm.d.kta[w];
            int k = T + 1;
            for (; k + 4 <= ntiles; k += 4)
                part |= sweep_groups<4, kFast>(sm.box, sm.d.tb, kb, ka, npad, nkw, k, dead, c1k, scale, Bseg, thr, union_fma);
            if (k + 2 <= ntiles) {
                part |= sweep_groups<2, kFast>(sm.box, sm.d.tb, kb, ka, npad, nkw, k, dead, c1k, scale, Bseg, thr, union_fma);
                k += 2;
            }
            if (k < ntiles)
                part |= sweep_groups<1, kFast>(sm.box, sm.d.tb, kb, ka, npad, nkw, k, dead, c1k, scale, Bseg, thr, union_fma);
        }
        reinterpret_cast<u32*>(&sm.d.dpart[lane])[w] = part;
        __syncthreads();
        const uint4 dp = sm.d.dpart[lane];
        dead |= dp.x | dp.y | dp.z | dp.w;
    }
    if (tid == 0 && total) atomicAdd(p.ws.keep_count + b, total);
}

// ------------------------------------------------------------------ CTA-owned (long) segments
struct TileSmem {
    u32 diag[64][2];          // diag[i] = columns j (tile-local) suppressed by row i, j > i
    float4 kbox[64];          // kept rows of the current tile, compacted (padded to an even count)
    float karea[64];          // -thr * |area| of those rows
    int nkept;
    int total_kept;
    u32 amax, cmax;           // largest |area| / |coordinate| of the segment (float bits)
};

// Process one segment.  seg: boxes in score order (global memory), cache: shared staging for
// the first `cache_cap` boxes, removed: bitmap words (>= (len+31)/32).
template <int THREADS, bool kFast, bool kFma>
__device__ void nms_segment(const float4* __restrict__ seg, const u64* __restrict__ gpk,
                            unsigned char* __restrict__ keepflag, int len,
                            float4* cache, int cache_cap, u32* removed, TileSmem& ts,
                            float thr, int union_fma)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncache = min(len, cache_cap);
    if (tid == 0) { ts.total_kept = 0; ts.amax = 0; ts.cmax = 0; }
    __syncthreads();
    u32 amax_bits = 0, cmax_bits = 0;
    for (int i = tid; i < len; i += THREADS) {
        const float4 q = seg[i];
        if (i < ncache) cache[i] = q;
        amax_bits = max(amax_bits, __float_as_uint(fabsf(ypp_area(q))));
        cmax_bits = max(max(cmax_bits, __float_as_uint(fabsf(q.x))), max(__float_as_uint(fabsf(q.y)),
                        max(__float_as_uint(fabsf(q.z)), __float_as_uint(fabsf(q.w)))));
    }
    amax_bits = __reduce_max_sync(YPP_FULL_MASK, amax_bits);
    cmax_bits = __reduce_max_sync(YPP_FULL_MASK, cmax_bits);
    if (lane == 0) { atomicMax(&ts.amax, amax_bits); atomicMax(&ts.cmax, cmax_bits); }
    for (int i = tid; i < (len + 31) / 32 + 2; i += THREADS) removed[i] = 0;
    __syncthreads();
    // the sweeps use the lean pair test of the per-class kernel (see pair_g): one band per segment
    float Bseg = 2.0f * YPP_PAIR_BAND * __uint_as_float(ts.amax);
    const int cexp = (int)(ts.cmax >> 23);
    if (cexp < 67 || cexp > 187) Bseg = INFINITY;
    const int cex = min(max(cexp, 67), 187);
    const float scale = __uint_as_float((u32)(252 - cex) << 23);
    const float c1k = __fmul_rn(__fadd_rn(1.0f, thr), __uint_as_float((u32)(cex + 2) << 23));
    const float nthr = -thr;

    for (int t0 = 0; t0 < len; t0 += 64) {
        const int rows = min(64, len - t0);
        // ---- diagonal tile: pair (i, j), i = p / 64, j = p % 64; a warp covers half a row
        for (int p = tid; p < 64 * 64; p += THREADS) {
            const int i = p >> 6, j = p & 63;
            bool bit = false;
            if (j > i && j < rows) {     // j < rows implies i < rows
                const float4 a = (t0 + i) < ncache ? cache[t0 + i] : seg[t0 + i];
                const float4 bq = (t0 + j) < ncache ? cache[t0 + j] : seg[t0 + j];
                const float sa = ypp_area(a);
                if (kFast) {
                    float p0, p1;
                    pair_col<kFma>(bq, p0, p1);
                    bool amb = false;
                    pair_fast<kFma>(a, sa, bq, p0, p1, thr, bit, amb);
                    if (amb) bit = pair_exact(a, sa, bq, thr, union_fma);
                } else {
                    bit = pair_exact(a, sa, bq, thr, union_fma);
                }
            }
            const u32 m = __ballot_sync(YPP_FULL_MASK, bit);
            if (lane == 0) ts.diag[i][j >> 5] = m;
        }
        __syncthreads();
        // ---- greedy resolve of the tile by warp 0 (all lanes redundantly; result identical)
        if (warp == 0) {
            u64 alive = ~(((u64)removed[(t0 >> 5) + 1] << 32) | (u64)removed[t0 >> 5]);
            if (rows < 64) alive &= (1ull << rows) - 1ull;
            u64 keep = 0;
            const u64* dg = reinterpret_cast<const u64*>(&ts.diag[0][0]);
#pragma unroll 8
            for (int k = 0; k < 64; ++k) {
                const u64 d = dg[k];
                if ((alive >> k) & 1ull) { keep |= 1ull << k; alive &= ~d; }
            }
            // compact the kept rows; lanes own rows lane and lane+32
            const u32 klo = (u32)keep, khi = (u32)(keep >> 32);
            const int nlo = __popc(klo);
            if ((klo >> lane) & 1u) {
                const int s = __popc(klo & ((1u << lane) - 1u));
                const int r = t0 + lane;
                const float4 bx = r < ncache ? cache[r] : seg[r];
                ts.kbox[s] = bx; ts.karea[s] = __fmul_rn(nthr, fabsf(ypp_area(bx)));
                keepflag[ypp_gpk_rank(gpk[r])] = 1;
            }
            if ((khi >> lane) & 1u) {
                const int s = nlo + __popc(khi & ((1u << lane) - 1u));
                const int r = t0 + 32 + lane;
                const float4 bx = r < ncache ? cache[r] : seg[r];
                ts.kbox[s] = bx; ts.karea[s] = __fmul_rn(nthr, fabsf(ypp_area(bx)));
                keepflag[ypp_gpk_rank(gpk[r])] = 1;
            }
            if (lane == 0) {
                const int nk = nlo + __popc(khi);
                ts.nkept = nk; ts.total_kept += nk;
                if (nk < 64) { ts.kbox[nk] = YPP_DUMMY_BOX; ts.karea[nk] = YPP_DUMMY_TA; }     // rows go in pairs
            }
        }
        __syncthreads();
        // ---- broadcast kept rows against later, still-alive columns (rows in chunks of 8)
        const int nk = ts.nkept;
        if (nk > 0) {
            for (int j = t0 + 64 + tid; j < len; j += THREADS) {
                if ((removed[j >> 5] >> (j & 31)) & 1u) continue;
                const float4 bq = j < ncache ? cache[j] : seg[j];
                bool dead = false;
                bool exact = !kFast;
                if (kFast) {
                    float dm = -INFINITY;
                    for (int r = 0; r < nk; r += 2)
                        dm = fmaxf(fmaxf(dm, pair_g(ts.kbox[r], ts.karea[r], bq, c1k, scale)),
                                   pair_g(ts.kbox[r + 1], ts.karea[r + 1], bq, c1k, scale));
                    const float D = __fadd_rn(dm, __fmul_rn(nthr, fabsf(ypp_area(bq))));
                    dead = D > Bseg;
                    exact = !dead && !(D < -Bseg);
                }
                if (exact) {
                    for (int r = 0; r < nk && !dead; ++r) dead = pair_exact(ts.kbox[r], ypp_area(ts.kbox[r]), bq, thr, union_fma);
                }
                if (dead) atomicOr(&removed[j >> 5], 1u << (j & 31));
            }
        }
        __syncthreads();
    }
}

// ---- sweep-and-prune for long segments -----------------------------------------------------
// The same idea as nms_sparse for the segments the 1024-thread CTAs own (single-class heads, merged
// classes): x1 counting sort into YPP_BSP_BINS bins, per-box range from the IoU bound, lean pair test,
// edge list, fix-point.  At these lengths every range is long, so every box is swept by a warp (lanes
// over its range); the boxes live in the dynamic shared memory the tile sweeps use as their cache
// (18 + 3 bytes per box, up to ~9700 boxes; the edge list takes the rest).  Returns false with nothing written when the
// segment is too long, holds a box that is not ordinary, overflows the edge list or resolves too
// deep; nms_segment (tile sweeps) then takes it.
#define YPP_BSP_BINS 1024
#define YPP_BIG_SMEM_MAX (222 * 1024)  /* dynamic shared memory of the long-segment kernel */
#define YPP_BSP_MIN_EDGES 2048          /* the edge list takes whatever shared memory the boxes leave; at least this */
#define YPP_BSP_FIXED (4 * YPP_BSP_BINS + 2 * (YPP_BSP_BINS + 2) + 64)

#define YPP_BSP_SIDE 32                 /* the bins are a YPP_BSP_SIDE x YPP_BSP_SIDE grid over (x1, y1), x-major */
struct BigSparseHdr {
    u32 xlo, xhi, ylo, yhi, amax, cmax;
    int nedges, odd;
    int wtot[YPP_NMS_BIG_THREADS / 32];
};

template <bool kFma>
__device__ bool nms_big_sparse(const float4* __restrict__ seg, const u64* __restrict__ gpk,
                               unsigned char* __restrict__ keepflag, const int len, unsigned char* smem,
                               const int smem_bytes, BigSparseHdr& H, const float thr, int& kept_total)
{
    constexpr int T = YPP_NMS_BIG_THREADS;
    static_assert(T == YPP_BSP_BINS, "the bin scan gives every thread one bin");
    static_assert(YPP_BSP_SIDE * YPP_BSP_SIDE == YPP_BSP_BINS, "square grid of bins");
    constexpr int SD = YPP_BSP_SIDE;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (len > 65535) return false;                                                          // uniform
    const long long spare = (long long)smem_bytes - (long long)len * 21 - YPP_BSP_FIXED;
    if (spare < 4 * YPP_BSP_MIN_EDGES) return false;                                        // uniform
    const int ecap = (int)(spare / 4);
    // carve: boxes | hist | bstart | rk | state | pend | kill | edges
    float4* box = reinterpret_cast<float4*>(smem);
    u32* hist = reinterpret_cast<u32*>(smem + (size_t)len * 16);
    unsigned short* bstart = reinterpret_cast<unsigned short*>(hist + YPP_BSP_BINS);
    unsigned short* rk = bstart + YPP_BSP_BINS + 2;
    unsigned char* state = reinterpret_cast<unsigned char*>(rk + len + (len & 1));
    unsigned char* pend = state + len;
    unsigned char* kill = pend + len;
    u32* edges = reinterpret_cast<u32*>(smem + (((size_t)(kill + len - smem) + 15) & ~(size_t)15));
    if (tid == 0) { H.xlo = 0xffffffffu; H.xhi = 0u; H.ylo = 0xffffffffu; H.yhi = 0u; H.amax = 0u; H.cmax = 0u; H.nedges = 0; H.odd = 0; }
    for (int i = tid; i < YPP_BSP_BINS; i += T) hist[i] = 0;
    __syncthreads();
    // ---- pass 1: ordinary? x1 range, largest |area| and |coordinate|
    {
        bool odd = false;
        u32 xlo = 0xffffffffu, xhi = 0u, ylo = 0xffffffffu, yhi = 0u, amax_bits = 0u, cmax_bits = 0u;
        for (int i = tid; i < len; i += T) {
            const float4 q = seg[i];
            const float bw = __fsub_rn(q.z, q.x), bh = __fsub_rn(q.w, q.y);
            odd |= !(bw > 0.0f && bh > 0.0f && bw < INFINITY && bh < INFINITY && fabsf(q.x) < INFINITY && fabsf(q.y) < INFINITY);
            const u32 e = ypp_enc(q.x), ey = ypp_enc(q.y);
            xlo = min(xlo, e); xhi = max(xhi, e);
            ylo = min(ylo, ey); yhi = max(yhi, ey);
            amax_bits = max(amax_bits, __float_as_uint(fabsf(__fmul_rn(bw, bh))));
            cmax_bits = max(max(cmax_bits, __float_as_uint(fabsf(q.x))), max(__float_as_uint(fabsf(q.y)),
                            max(__float_as_uint(fabsf(q.z)), __float_as_uint(fabsf(q.w)))));
        }
        xlo = __reduce_min_sync(YPP_FULL_MASK, xlo);
        xhi = __reduce_max_sync(YPP_FULL_MASK, xhi);
        ylo = __reduce_min_sync(YPP_FULL_MASK, ylo);
        yhi = __reduce_max_sync(YPP_FULL_MASK, yhi);
        amax_bits = __reduce_max_sync(YPP_FULL_MASK, amax_bits);
        cmax_bits = __reduce_max_sync(YPP_FULL_MASK, cmax_bits);
        if (lane == 0) {
            atomicMin(&H.xlo, xlo); atomicMax(&H.xhi, xhi); atomicMin(&H.ylo, ylo); atomicMax(&H.yhi, yhi);
            atomicMax(&H.amax, amax_bits); atomicMax(&H.cmax, cmax_bits);
        }
        if (__syncthreads_or(odd)) return false;
    }
    const float Bseg = 2.0f * YPP_PAIR_BAND * __uint_as_float(H.amax);
    const int cexp = (int)(H.cmax >> 23);
    if (cexp < 67 || cexp > 187 || !(Bseg < INFINITY)) return false;                     // uniform
    const float scale = __uint_as_float((u32)(252 - cexp) << 23);
    const float c1k = __fmul_rn(__fadd_rn(1.0f, thr), __uint_as_float((u32)(cexp + 2) << 23));
    const float xmin = ypp_dec(H.xlo), xmax = ypp_dec(H.xhi), ymin = ypp_dec(H.ylo), ymax = ypp_dec(H.yhi);
    const float inv = __fdiv_rn((float)SD, __fsub_rn(xmax, xmin));           // inf when every x1 is the same: one column
    const float invy = __fdiv_rn((float)SD, __fsub_rn(ymax, ymin));
    auto bin_of = [&](const float4 q) { return sp_bin<SD>(q.x, xmin, inv) * SD + sp_bin<SD>(q.y, ymin, invy); };
    // ---- pass 2: histogram, scan (one bin per thread), pass 3: scatter
    for (int i = tid; i < len; i += T) atomicAdd(&hist[bin_of(seg[i])], 1u);
    __syncthreads();
    {
        const int h = (int)hist[tid];
        int incl = h;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(YPP_FULL_MASK, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) H.wtot[warp] = incl;
        __syncthreads();
        int off = incl - h;
        for (int k = 0; k < warp; ++k) off += H.wtot[k];
        bstart[tid] = (unsigned short)off;
        if (tid == 0) bstart[YPP_BSP_BINS] = (unsigned short)len;
        hist[tid] = (u32)off;
    }
    __syncthreads();
    for (int i = tid; i < len; i += T) {
        const float4 q = seg[i];
        const int pos = (int)atomicAdd(&hist[bin_of(q)], 1u);
        box[pos] = q;
        rk[pos] = (unsigned short)i;
        state[i] = 0; pend[i] = 0; kill[i] = 0;
    }
    __syncthreads();
    // ---- sweep: a warp per box, lanes over its windows.  Box p (bin column xb_p) meets the boxes AFTER it in the
    // (x-major) order that lie in columns xb_p .. column of xe and, inside every column, in the rows of its y window:
    //   x:  xe = x1 + (1 - 0.9 thr) w_p                       (as in nms_sparse: beyond it IoU <= 0.9 thr)
    //   y:  y1_c <= y1_p + (1 - 0.9 thr) h_p                  (y1_c >= y1_p: IoU <= overlap_y / h_p <= 1 - dy / h_p)
    //       y1_c >= y1_p - (1 - 0.9 thr) / (0.9 thr) h_p      (y1_c <  y1_p: IoU <= 1 - dy / h_c, and IoU > 0.9 thr
    //                                                          needs h_c <= h_p / (0.9 thr))
    // every step rounded outwards, windows widened to whole bins: a pair outside has IoU <= 0.9 thr in exact
    // arithmetic, which no rounding of the fp32 quotient lifts above thr.  Both windows follow from IoU > 0.9 thr alone,
    // so a real edge lies in the window of BOTH its boxes and is met from the side that comes first in the order.
    {
        const float ct = __fsub_ru(1.0f, __fmul_rd(0.9f, thr));
        const float cdown = __fdiv_ru(ct, __fmul_rd(0.9f, thr));
        const float nthr = -thr;
        for (int pr = warp; pr < len; pr += T / 32) {
            if (*(volatile int*)&H.nedges >= ecap) break;
            const float4 a = box[pr];
            const float hp = __fsub_ru(a.w, a.y);
            const float xe = __fadd_ru(a.x, __fmul_ru(ct, __fsub_ru(a.z, a.x)));
            const float yu = __fadd_ru(a.y, __fmul_ru(ct, hp));
            const float yd = __fsub_rd(a.y, __fmul_ru(cdown, hp));
            const int xb0 = sp_bin<SD>(a.x, xmin, inv), xb1 = sp_bin_end<SD>(xe, xmin, inv);
            const int yb0 = sp_bin<SD>(yd, ymin, invy), yb1 = sp_bin_end<SD>(yu, ymin, invy);
            const float ta = __fmul_rn(nthr, fabsf(ypp_area(a)));
            const int ra = rk[pr];
            for (int xb = xb0; xb <= xb1; ++xb) {
                int c0 = bstart[xb * SD + yb0];
                const int c1 = bstart[xb * SD + yb1 + 1];
                if (xb == xb0) c0 = max(c0, pr + 1);                 // my own column: only what comes after me
                for (int c = c0 + lane; c < c1; c += 32) {
                    const float4 bq = box[c];
                    const float D = __fadd_rn(pair_g(a, ta, bq, c1k, scale), __fmul_rn(nthr, fabsf(ypp_area(bq))));
                    if (D < -Bseg) continue;
                    const int rb = rk[c];
                    bool dead = D > Bseg;
                    if (!dead) {
                        const bool a_first = ra < rb;
                        const float4 hi = a_first ? a : bq, lo = a_first ? bq : a;
                        dead = pair_exact(hi, ypp_area(hi), lo, thr, kFma ? 1 : 0);
                    }
                    if (dead) {
                        const int e = atomicAdd(&H.nedges, 1);
                        if (e < ecap) edges[e] = ((u32)min(ra, rb) << 16) | (u32)max(ra, rb);
                    }
                }
            }
        }
    }
    __syncthreads();
    const int ne = H.nedges;
    if (ne > ecap) return false;
    // ---- greedy order over the edges
    bool settled = ne == 0;
    for (int round = 0; !settled && round < YPP_SP_ROUNDS; ++round) {
        for (int e = tid; e < ne; e += T) {
            const u32 ed = edges[e];
            const int st = state[ed >> 16];
            if (st == 1) kill[ed & 0xffffu] = 1;
            else if (st == 0) pend[ed & 0xffffu] = 1;
        }
        __syncthreads();
        bool open = false;
        for (int i = tid; i < len; i += T) {
            if (state[i] == 0) {
                if (kill[i]) state[i] = 2;
                else if (!pend[i]) state[i] = 1;
                else open = true;
            }
            pend[i] = 0;
        }
        settled = !__syncthreads_or(open);
    }
    if (!settled) return false;
    int kept = 0;
    for (int i = tid; i < len; i += T)
        if (state[i] != 2) { keepflag[ypp_gpk_rank(gpk[i])] = 1; ++kept; }
    kept = __reduce_add_sync(YPP_FULL_MASK, kept);
    if (lane == 0) H.wtot[warp] = kept;
    __syncthreads();
    kept_total = 0;
    for (int k = 0; k < T / 32; ++k) kept_total += H.wtot[k];
    __syncthreads();                              // the shared memory is reused by the CTA's next segment
    return true;
}

// persistent CTAs walk the per-image lists of long segments S1 left (bigq[b][0 .. bigcnt[b]))
template <bool kFast, bool kFma>
__global__ void __launch_bounds__(YPP_NMS_BIG_THREADS, 1)
ypp_nms_big_kernel(const __grid_constant__ YppNmsParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* cache = reinterpret_cast<float4*>(smem_raw);
    u32* removed = reinterpret_cast<u32*>(smem_raw + sizeof(float4) * YPP_BIG_CACHE);
    __shared__ TileSmem ts;
    __shared__ BigSparseHdr bsh;
    ypp_grid_wait();
    const int per = p.ws.bigq_per_image;
    {   // nearly always there is nothing to do: find out with one round of loads
        int any = 0;
        for (int b = threadIdx.x; b < p.B; b += YPP_NMS_BIG_THREADS) any |= p.ws.bigcnt[b];
        if (!__syncthreads_or(any)) return;
        // long segments exist on this device: the host sizes the next launches for them (ypp_adaptive_launch)
        if (p.big_seen && blockIdx.x == 0 && threadIdx.x == 0) *(volatile int*)p.big_seen = 1;
    }
    for (int it = blockIdx.x; it < p.B * per; it += gridDim.x) {
        const int b = it / per, k = it - b * per;
        if (k >= min(p.ws.bigcnt[b], per)) continue;          // uniform over the CTA
        const int c = p.ws.bigq[it];
        const int* seg_start = p.ws.seg_start + (size_t)b * (p.nc + 1);
        const int s0 = seg_start[c], len = seg_start[c + 1] - s0;
        const size_t base = (size_t)b * p.N;
        {   // materialise the segment's boxes (score order) for the sweeps below
            const bool trick = (p.ws.imgflag[b] & 2) != 0;
            const float step = p.ws.img_step[b];
            for (int i = threadIdx.x; i < len; i += YPP_NMS_BIG_THREADS)
                p.ws.sbox[base + s0 + i] = pos_box(p.ws, base, p.ws.gpk[base + s0 + i], trick, step);
            __syncthreads();
        }
        if (kFast && p.iou_thr >= 0.0f && p.sparse_min_len != 0x7fffffff) {
            int kept = 0;
            if (nms_big_sparse<kFma>(p.ws.sbox + base + s0, p.ws.gpk + base + s0, p.ws.keepflag + base, len, smem_raw,
                                     p.big_smem_bytes, bsh, p.iou_thr, kept)) {
                if (threadIdx.x == 0 && kept) atomicAdd(p.ws.keep_count + b, kept);
                continue;                                   // uniform over the CTA
            }
            __syncthreads();
        }
        nms_segment<YPP_NMS_BIG_THREADS, kFast, kFma>(p.ws.sbox + base + s0, p.ws.gpk + base + s0,
                                                      p.ws.keepflag + base, len, cache, YPP_BIG_CACHE,
                                                      removed, ts, p.iou_thr, p.union_mode);
        if (threadIdx.x == 0 && ts.total_kept) atomicAdd(p.ws.keep_count + b, ts.total_kept);
        __syncthreads();
    }
}

// ------------------------------------------------------------------ gather
// keep flags (global score order) -> packed rows + kept box indices.  grid (chunks, B): a CTA
// counts the flags of the chunks before its own (uint4 sweeps over the byte flags), then
// places its own YPP_GATHER_CHUNK ranks with a ballot prefix.
#define YPP_GATHER_THREADS 256
#define YPP_GATHER_ROUNDS 4
#define YPP_GATHER_CHUNK (YPP_GATHER_ROUNDS * YPP_GATHER_THREADS)

__device__ __forceinline__ int popc_bytes16(const uint4 v)
{
    return __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);     // flags are 0/1 bytes
}

__global__ void __launch_bounds__(YPP_GATHER_THREADS, 4)
ypp_gather_kernel(const __grid_constant__ YppGatherParams p)
{
    __shared__ int warp_tot[YPP_GATHER_ROUNDS][YPP_GATHER_THREADS / 32];
    __shared__ __align__(16) float stage[YPP_GATHER_CHUNK * 7 + 4];
    __shared__ long long red_img[YPP_GATHER_THREADS / 32];
    __shared__ int red_pre[YPP_GATHER_THREADS / 32];
    ypp_grid_wait();
    const int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = p.N;
    const int r_begin = blockIdx.x * YPP_GATHER_CHUNK;
    const unsigned char* kf = p.ws.keepflag + (size_t)b * N;
    const u64* keys = p.ws.skeys + (size_t)b * N;
    // the chunk's own flags and keys are asked for first, together with the counts: they need nothing the prefix
    // sums below produce, and the kernel is a chain of dependent load latencies (five before, three now).  Ranks
    // past the image's candidates read allocated but unwritten words; r < r_end masks them below
    unsigned char fl[YPP_GATHER_ROUNDS];
    u64 k[YPP_GATHER_ROUNDS];
#pragma unroll
    for (int u = 0; u < YPP_GATHER_ROUNDS; ++u) {
        const int r = r_begin + u * YPP_GATHER_THREADS + tid;
        fl[u] = r < N ? kf[r] : (unsigned char)0;
        k[u] = r < N ? keys[r] : 0ull;
    }
    const int cap = p.max_det > 0 ? p.max_det : 0x7fffffff;          // rows an image may contribute
    const bool writes_total = blockIdx.x == 0 && b == (int)gridDim.y - 1 && (p.det_total != nullptr || p.cursor_out != nullptr);
    // rows of the images before this one, and kept flags of this image before my chunk - also asked for before the
    // candidate count is known: a chunk that starts past the count leaves below and its sums are never used
    long long img_part = 0;
    for (int i = tid; i < b; i += YPP_GATHER_THREADS) img_part += min(p.ws.keep_count[i], cap);
    int pre_part = 0;
    {
        const int pre = r_begin;                                  // < N; == min(r_begin, n) for every chunk that stays
        int head = (int)((16 - ((uintptr_t)kf & 15)) & 15);       // b*N need not be a multiple of 16
        if (head > pre) head = pre;
        if (tid < head) pre_part += kf[tid];
        const int nvec = (pre - head) >> 4;
        const uint4* v = reinterpret_cast<const uint4*>(kf + head);
        for (int i = tid; i < nvec; i += YPP_GATHER_THREADS) pre_part += popc_bytes16(v[i]);
        const int tail0 = head + (nvec << 4);
        if (tail0 + tid < pre) pre_part += kf[tail0 + tid];
    }
    const int n = min(p.ws.cand_count[b], N);
    const long long base = p.cursor_in ? *p.cursor_in : 0;          // append mode: rows already in det_rows
    if (blockIdx.x == 0 && tid == 0) {
        if (p.cand_count_out) p.cand_count_out[b] = n;
        if (p.keep_count_out) p.keep_count_out[b] = min(p.ws.keep_count[b], cap);
    }
    if (r_begin >= n && !writes_total) return;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        img_part += __shfl_xor_sync(YPP_FULL_MASK, img_part, d);
        pre_part += __shfl_xor_sync(YPP_FULL_MASK, pre_part, d);
    }
    if (lane == 0) { red_img[warp] = img_part; red_pre[warp] = pre_part; }
    __syncthreads();
    long long img_rows = 0;
    int pre_rows = 0;
#pragma unroll
    for (int w = 0; w < YPP_GATHER_THREADS / 32; ++w) { img_rows += red_img[w]; pre_rows += red_pre[w]; }
    if (writes_total && tid == 0) {
        const long long rows = img_rows + min(p.ws.keep_count[b], cap);
        if (p.det_total) *p.det_total = (int)rows;
        if (p.cursor_out) *p.cursor_out = base + rows;               // never the word cursor_in points at
        // the host's view of the cursor: one 8-byte store into pinned memory, count and sequence number together
        if (p.host_probe) *(volatile long long*)p.host_probe = ((base + rows) << 24) | (long long)(p.probe_seq & 0xffffffu);
    }
    if (r_begin >= n || pre_rows >= cap) return;
    const float4* box = p.ws.box + (size_t)b * N;
    const float img = p.img_idx_base + (float)b;
    const int r_end = min(n, r_begin + YPP_GATHER_CHUNK);
    // the chunk's YPP_GATHER_ROUNDS rounds of 256 ranks at once: flags -> ballots -> one exchange of the
    // per-(round, warp) counts -> keys -> boxes -> rows, so every load of a round overlaps the others'
    bool f[YPP_GATHER_ROUNDS];
    u32 m[YPP_GATHER_ROUNDS];
#pragma unroll
    for (int u = 0; u < YPP_GATHER_ROUNDS; ++u) {
        const int r = r_begin + u * YPP_GATHER_THREADS + tid;
        f[u] = r < r_end && fl[u];
    }
#pragma unroll
    for (int u = 0; u < YPP_GATHER_ROUNDS; ++u) {
        m[u] = __ballot_sync(YPP_FULL_MASK, f[u]);
        if (lane == 0) warp_tot[u][warp] = __popc(m[u]);
    }
    __syncthreads();
    int pos[YPP_GATHER_ROUNDS];
    int running = pre_rows;
#pragma unroll
    for (int u = 0; u < YPP_GATHER_ROUNDS; ++u) {
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < YPP_GATHER_THREADS / 32; ++w) {
            const int t = warp_tot[u][w];
            if (w < warp) before += t;
            total += t;
        }
        pos[u] = running + before + __popc(m[u] & ((1u << lane) - 1u));      // position inside the image
        running += total;
        f[u] = f[u] && pos[u] < cap;                                         // max_det: later ranks are dropped
    }
    running = min(running, cap);
    float4 bx[YPP_GATHER_ROUNDS];
    float cls[YPP_GATHER_ROUNDS];
    const bool want_rows = p.det_rows != nullptr;
#pragma unroll
    for (int u = 0; u < YPP_GATHER_ROUNDS; ++u) {
        if (!f[u]) continue;
        const int idx = ypp_key_idx(k[u]);
        if (p.keep_idx) p.keep_idx[(size_t)b * N + pos[u]] = idx;
        if (want_rows) {
            bx[u] = box[idx];
            if (p.xyxy_out) bx[u] = ypp_xyxy(bx[u]);
            cls[u] = (float)ypp_key_cls(k[u]);
            if (p.src_rows) cls[u] = p.src_rows[((size_t)b * N + idx) * 6 + 5];
        }
    }
    if (!want_rows) return;
    // rows of the chunk are contiguous in the output: stage them in shared memory (stride 6 or 7 floats:
    // conflict-free) and write them out as full lines instead of 7 scattered 4-byte stores per row.  The stage starts
    // at the same offset inside a 16-byte word as the output does, so the body of the copy moves float4s
    const int width = p.row_width;
    const long long first = base + img_rows + pre_rows;
    float* out = p.det_rows + first * width;
    const int phase = (int)(((uintptr_t)out >> 2) & 3);
#pragma unroll
    for (int u = 0; u < YPP_GATHER_ROUNDS; ++u) {
        if (!f[u]) continue;
        float* o = stage + phase + (pos[u] - pre_rows) * width;
        if (width == 7) *o++ = img;
        o[0] = bx[u].x; o[1] = bx[u].y; o[2] = bx[u].z; o[3] = bx[u].w;
        o[4] = ypp_key_score(k[u]); o[5] = cls[u];
    }
    __syncthreads();
    long long nrows = running - pre_rows;
    if (first + nrows > p.det_capacity) nrows = p.det_capacity > first ? p.det_capacity - first : 0;
    const int nf = (int)nrows * width;
    const float* src = stage + phase;
    const int head = min((4 - phase) & 3, nf);                       // floats before the first aligned word
    if (tid < head) out[tid] = src[tid];
    const int nvec = (nf - head) >> 2;
    {
        float4* o4 = reinterpret_cast<float4*>(out + head);
        const float4* s4 = reinterpret_cast<const float4*>(src + head);
        for (int i = tid; i < nvec; i += YPP_GATHER_THREADS) o4[i] = s4[i];
    }
    const int tail0 = head + (nvec << 2);
    if (tail0 + tid < nf) out[tail0 + tid] = src[tail0 + tid];
}

// ---- how wide the long-segment kernel is launched (ypp_adaptive_launch) ------------------------------------
// N1b's CTAs take a whole SM each (1024 threads, up to 222 KB of shared memory).  Launched 148 wide "in case", the
// kernel that nearly always finds nothing to do still needs every SM fully vacated once per batch - with several
// batches in flight that is a device-wide hiccup per step (C2: 869 k -> 891 k images/s without it).  When the caller
// allowed it, the library keeps one page-locked host word per device that N1b sets the first time it finds a long
// segment; until then many-class configurations launch N1b YPP_BIG_IDLE_CTAS wide (the kernel walks its work list
// whatever the grid, so results do not depend on this), from then on at full width.  The word is host memory of the
// library itself, pinned with cudaHostRegister - nothing is allocated - and it is only ever read as a hint.
#define YPP_BIG_IDLE_CTAS 8
static int g_big_seen[YPP_MAX_DEVICES][16];                // one cache line per device
static int* g_big_seen_dev[YPP_MAX_DEVICES];               // the same words as the device sees them
static int g_adaptive = 0;                                 // 0 off, 1 on, -1 registration failed

// the device-visible address of this device's word (null: feature off), looked up once per device
static int* big_seen_word(int dev)
{
    if (g_adaptive <= 0) return nullptr;
    if (!g_big_seen_dev[dev]) {
        void* d = nullptr;
        if (cudaHostGetDevicePointer(&d, &g_big_seen[dev][0], 0) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        g_big_seen_dev[dev] = (int*)d;
    }
    return g_big_seen_dev[dev];
}

template <bool kFast, bool kFma>
int launch_nms_t(const YppNmsParams& p, cudaStream_t st)
{
    // N1 (segments <= 512) and N1b (longer ones) read what S1 left and write disjoint keep flags: their order is free.
    // Few classes (N / nc >= 1024: long segments are the rule, N1b is the step): N1b goes FIRST - launched early (PDL)
    // its <= 148 CTAs settle on the SMs S1's one-CTA-per-image grid leaves idle and start the moment S1 ends; with
    // three batches in flight C3 gains 17 % (217 k -> 253 k images/s).  Many classes (N1b nearly always finds nothing
    // to do): N1b stays LAST - first, its whole-SM CTAs have to wait for SMs the neighbouring batches occupy and hold
    // N1 back (C2: 869 k -> 829 k images/s).  YPP_BIG_FIRST=0/1 forces an order (tuning only).
    static int forced = -2;
    if (forced == -2) { const char* e = getenv("YPP_BIG_FIRST"); forced = e ? (atoi(e) != 0) : -1; }
    const bool big_first = forced >= 0 ? forced != 0 : (long long)p.N >= 2LL * YPP_WARP_SEG * p.nc;
    int rc = 0;
    auto launch_seg = [&]() {
        dim3 grid(p.nc, p.B);
        ypp_launch_chain(2, ypp_nms_seg_kernel<kFast, kFma>, grid, dim3(YPP_NMSQ_THREADS), 0, st, p);
        ypp_count_launch(1);
        return ypp_check_launch("nms_seg");
    };
    if (!big_first) { rc = launch_seg(); if (rc) return rc; }
    if (p.N > YPP_WARP_SEG) {
        int smem = (int)(sizeof(float4) * YPP_BIG_CACHE + sizeof(u32) * ((size_t)p.N / 32 + 4));
        const int want = (int)((long long)p.N * 21 + YPP_BSP_FIXED + 4 * 16384);        // the sweep-and-prune layout for a whole image
        if (smem < want) smem = want < YPP_BIG_SMEM_MAX ? want : YPP_BIG_SMEM_MAX;
        static int sms_d[YPP_MAX_DEVICES];                 // per device: 0 = the kernel's attribute is not set yet
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= YPP_MAX_DEVICES) { ypp_set_error("nms_big: device ordinal %d out of range", dev); return YPP_E_UNSUPPORTED; }
        if (!sms_d[dev]) {
            cudaError_t e = cudaFuncSetAttribute(ypp_nms_big_kernel<kFast, kFma>, cudaFuncAttributeMaxDynamicSharedMemorySize, YPP_BIG_SMEM_MAX);
            if (e != cudaSuccess) { ypp_set_error("nms_big: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
            int sms = 148;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            sms_d[dev] = sms > 0 ? sms : 148;
        }
        if (smem > YPP_BIG_SMEM_MAX) { ypp_set_error("nms_big: %d boxes per image exceed the removed-bitmap budget", p.N); return YPP_E_UNSUPPORTED; }
        int ctas = p.B * p.ws.bigq_per_image;
        if (ctas > sms_d[dev]) ctas = sms_d[dev];
        YppNmsParams pb = p;
        pb.big_smem_bytes = smem;
        pb.big_seen = big_first ? nullptr : big_seen_word(dev);      // few-class batches say nothing about many-class ones
        // many classes and no long segment seen on this device so far: a narrow launch (see above)
        if (!big_first && pb.big_seen && !*(volatile int*)&g_big_seen[dev][0] && ctas > YPP_BIG_IDLE_CTAS) ctas = YPP_BIG_IDLE_CTAS;
        ypp_launch_chain(3, ypp_nms_big_kernel<kFast, kFma>, dim3(ctas), dim3(YPP_NMS_BIG_THREADS), smem, st, pb);
        ypp_count_launch(1);
        rc = ypp_check_launch("nms_big");
        if (rc) return rc;
    }
    if (big_first) rc = launch_seg();
    return rc;
}

}  // namespace

int ypp_adaptive_launch_set(int enable)
{
    if (!enable) { if (g_adaptive > 0) g_adaptive = 0; return YPP_OK; }
    if (g_adaptive != 0) { if (g_adaptive < 0) return YPP_E_CUDA; g_adaptive = 1; return YPP_OK; }
    static bool registered = false;
    if (!registered) {
        cudaError_t e = cudaHostRegister(g_big_seen, sizeof(g_big_seen), cudaHostRegisterMapped | cudaHostRegisterPortable);
        if (e != cudaSuccess) {
            cudaGetLastError();
            g_adaptive = -1;
            ypp_set_error("adaptive launch: cudaHostRegister: %s", cudaGetErrorString(e));
            return YPP_E_CUDA;
        }
        registered = true;
    }
    g_adaptive = 1;
    return YPP_OK;
}

#ifdef YPP_SP_STATS
extern "C" int ypp_debug_sp_stats(unsigned long long* out16, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out16, ypp_sp_stats, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(ypp_sp_stats, z, sizeof(z)); }
    return 0;
}
#endif

int ypp_launch_nms(const YppNmsParams& p0, cudaStream_t st)
{
    // YPP_SPARSE_MIN: shortest per-class segment that tries the sweep-and-prune path (0 = never; tuning knob)
    static int sparse_min = -1;
    if (sparse_min < 0) { const char* e = getenv("YPP_SPARSE_MIN"); sparse_min = e ? atoi(e) : YPP_SP_MIN_LEN; if (sparse_min < 0) sparse_min = 0; }
    YppNmsParams p = p0;
    p.sparse_min_len = sparse_min > 0 ? sparse_min : 0x7fffffff;
    // the division-free pair test is proven for thresholds in [2^-10, 1]; anything else divides
    const bool fast = p.iou_thr >= 0.0009765625f && p.iou_thr <= 1.0f;
    if (!fast) return launch_nms_t<false, false>(p, st);
    return p.union_mode == YPP_UNION_FMA ? launch_nms_t<true, true>(p, st) : launch_nms_t<true, false>(p, st);
}

int ypp_launch_gather(const YppGatherParams& p, cudaStream_t st)
{
    dim3 grid((p.N + YPP_GATHER_CHUNK - 1) / YPP_GATHER_CHUNK, p.B);
    ypp_launch_chain(4, ypp_gather_kernel, grid, dim3(YPP_GATHER_THREADS), 0, st, p);
    ypp_count_launch(1);
    return ypp_check_launch("gather");
}
