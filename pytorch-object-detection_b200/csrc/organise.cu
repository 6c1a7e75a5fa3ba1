// S1: per image ordering.  One CTA per image:
//   1. concatenate the per-slot (32 boxes) candidate keys D1 left (box-index order) into shared memory;
//   2. sort -> the stable descending-score order of torchvision::nms (scores.sort(stable=True,
//      descending=True); ties -> lower box index): one bucketing pass on the top bits in which the
//      image's scores differ + ranking inside the (tiny) buckets; stable LSD radix passes on the
//      score word when many scores tie;
//   3. stable counting partition by class -> per (image, class) segments in the same order;
//   4. emit, per segment position, one word holding the global rank, the box index and the class
//      (gpk) and, per image, whether torchvision would take the coordinate-trick path
//      (_batched_nms_coordinate_trick) and its offset step max_coordinate + 1; N1 reads the word,
//      then the box, converts it to xyxy (utils/yolo_utils.py:339-343) and applies cls * step;
//   5. coordinate trick only: detect cross-class suppression edges (possible when a box with
//      negative x1,y1 is shifted onto the previous class band); the classes such an edge couples
//      share ONE segment, so the greedy scan sees exactly what torchvision's single pass does.
// Images with at most YPP_SMALL_MAX candidates skip 2..5: organise_small below finishes them here.
#include "ypp_kernels.cuh"

#ifdef YPP_PROFILE_S1
__device__ long long ypp_s1_clk[16];
#define S1_MARK(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) ypp_s1_clk[i] = clock64(); } while (0)
__device__ __forceinline__ long long s1_gtime() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
// [9] longest CTA (cycles), [10] earliest CTA start, [11] latest CTA end (globaltimer ns; [10] must be preset large)
#define S1_CTA_BEGIN() const long long s1_c0 = clock64(); if (threadIdx.x == 0) atomicMin(&ypp_s1_clk[10], s1_gtime())
#define S1_CTA_END() do { if (threadIdx.x == 0) { atomicMax(&ypp_s1_clk[9], clock64() - s1_c0); atomicMax(&ypp_s1_clk[11], s1_gtime()); } } while (0)
#else
#define S1_MARK(i) do { } while (0)
#define S1_CTA_BEGIN() do { } while (0)
#define S1_CTA_END() do { } while (0)
#endif

namespace {

#define YPP_DET_LIST 64
#define YPP_DET_BINS 64
#define YPP_DET_EDGES 64
#define YPP_ORG_WARPS (YPP_ORG_THREADS / 32)
#define YPP_RADIX_MAX 256
#define YPP_TABLE_PITCH 33              // table[digit * 33 + warp]: leaders of one warp hit distinct banks

struct DetEntry { float4 box; u64 key; int cls; int pad; };

struct OrgSmemHeader {
    DetEntry lowl[YPP_DET_LIST];
    DetEntry highl[YPP_DET_LIST];
    u32 bin_hi[YPP_DET_BINS];     // A role: per x-bin max of enc(y2 - step); after the scan: suffix max
    u32 bin_lo[YPP_DET_BINS];     // B role: per x-bin max of enc(-y1);      after the scan: prefix max
    int warp_sums[YPP_ORG_WARPS];
    int n_low, n_high, flagged, n;
    int n_edges, max_bucket, n_big, epad;
    u32 edges[YPP_DET_EDGES];     // (class of the earlier box << 16) | class of the box it suppresses
    u32 maxc, nminx1, nminy1, kmin, kmax, pad[3];
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

// block-wide exclusive scan of one int per thread (all YPP_ORG_THREADS threads call it)
__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int& total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(YPP_FULL_MASK, inc, d);
        if (lane >= d) inc += o;
    }
    __syncthreads();                       // previous users of warp_sums are done
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    const int ws = warp_sums[lane];        // YPP_ORG_WARPS == 32
    int winc = ws;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(YPP_FULL_MASK, winc, d);
        if (lane >= d) winc += o;
    }
    total = __shfl_sync(YPP_FULL_MASK, winc, 31);
    const int wbase = __shfl_sync(YPP_FULL_MASK, winc - ws, warp);
    return wbase + inc - v;
}

// lanes of the warp that hold the same nbits-bit value (MATCH.ANY is far slower than nbits ballots
// when every warp of the CTA is issuing it); lanes with ok == false belong to no group
__device__ __forceinline__ u32 same_value_lanes(int v, int nbits, bool ok)
{
    u32 peers = __ballot_sync(YPP_FULL_MASK, ok);
    for (int i = 0; i < nbits; ++i) {
        const bool bit = (v >> i) & 1;
        const u32 bal = __ballot_sync(YPP_FULL_MASK, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

// One stable LSD pass on digit ((key >> 32) - kmin) >> shift  (bw bits).  Warp w owns the contiguous
// ranks [w*L, (w+1)*L) and walks them 32 at a time, so "warp order, then chunk order, then lane
// order" is the input order; per (digit, warp) counters make the scatter stable.
__device__ void radix_pass(const u64* src, u64* dst, int n, int L, int shift, int bw,
                           u32 kmin, int* table, int* warp_sums)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int R = 1 << bw;
    const u32 dmask = (u32)R - 1u;
    for (int i = tid; i < R * YPP_TABLE_PITCH; i += YPP_ORG_THREADS) table[i] = 0;
    __syncthreads();
    const int r_begin = warp * L, r_end = min(n, r_begin + L);
    for (int r0 = r_begin; r0 < r_end; r0 += 32) {
        const int r = r0 + lane;
        const bool ok = r < r_end;
        // counting needs no order: shared-memory atomics (only this warp touches column `warp`)
        if (ok) atomicAdd(&table[(int)((((u32)(src[r] >> 32) - kmin) >> shift) & dmask) * YPP_TABLE_PITCH + warp], 1);
    }
    __syncthreads();
    // exclusive scan in (digit, warp) order; logical entry e = digit * 32 + warp
    {
        const int E = R * 32;
        const int ipt = max(1, E / YPP_ORG_THREADS);       // R/32 entries per thread (R >= 32) else 1
        const int e0 = tid * ipt;
        int vals[YPP_RADIX_MAX / 32];
        int sum = 0;
#pragma unroll
        for (int k = 0; k < YPP_RADIX_MAX / 32; ++k) {
            vals[k] = 0;
            const int e = e0 + k;
            if (k < ipt && e < E) { vals[k] = table[(e >> 5) * YPP_TABLE_PITCH + (e & 31)]; sum += vals[k]; }
        }
        int total;
        int run = block_exclusive_scan(sum, warp_sums, total);
#pragma unroll
        for (int k = 0; k < YPP_RADIX_MAX / 32; ++k) {
            const int e = e0 + k;
            if (k < ipt && e < E) { table[(e >> 5) * YPP_TABLE_PITCH + (e & 31)] = run; run += vals[k]; }
        }
    }
    __syncthreads();
    for (int r0 = r_begin; r0 < r_end; r0 += 32) {
        const int r = r0 + lane;
        const bool ok = r < r_end;
        u64 k = 0;
        int d = 0;
        if (ok) { k = src[r]; d = (int)((((u32)(k >> 32) - kmin) >> shift) & dmask); }
        const unsigned peers = same_value_lanes(d, bw, ok);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        int base = 0;
        if (ok) base = table[d * YPP_TABLE_PITCH + warp];
        __syncwarp();
        if (ok) {
            if (rank == 0) table[d * YPP_TABLE_PITCH + warp] = base + __popc(peers);
            dst[base + rank] = k;
        }
        __syncwarp();
    }
    __syncthreads();
}

// Sort by bucketing + ranking.  The keys are unique 64-bit words, so ANY sort of them is the stable
// score order; stability of the passes themselves is not needed.  One MSD pass on the top
// dbits <= 12 bits of the score span drops every key into its bucket with plain shared-memory
// atomics (order inside a bucket arbitrary); buckets hold a handful of keys, so a key's final
// place is bucket start + the number of smaller keys in its bucket.  Returns false (nothing written
// to dst) when some bucket is too crowded for that (many tied scores): the caller takes the LSD passes.
// Two bucket functions, both monotone in the key: kLinear = false, the top dbits of the score WORD (equal
// buckets per binade: right for scores spread over many binades, e.g. conf 0.001 on a trained net);
// kLinear = true, the score VALUE mapped linearly onto the buckets (right for scores spread evenly over
// their range: half of a uniform sample shares the top binade, 27 keys per bucket on C4 against 5.5).
// The caller tries the linear one first with a tight limit, then the word one, then the LSD passes.
#define YPP_BUCKET_BITS 12
#define YPP_BUCKET_MAX 96
#define YPP_BUCKET_MAX_LINEAR 32
template <bool kLinear>
__device__ __forceinline__ int bucket_of(const u64 k, const u32 kmin, const int shift, const float smax, const float scale, const int R)
{
    if (!kLinear) return (int)(((u32)(k >> 32) - kmin) >> shift);
    const float v = __fmul_rn(__fsub_rn(smax, ypp_key_score(k)), scale);      // 0 for the best score, grows with the key
    return min(max((int)v, 0), R - 1);
}

template <bool kLinear>
__device__ bool bucket_rank_sort(const u64* src, u64* tmp, u64* dst, int n, int bits, u32 kmin, u32 kmax,
                                 int* table, OrgSmemHeader* hd)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const int dbits = kLinear ? YPP_BUCKET_BITS : min(bits, YPP_BUCKET_BITS), shift = bits - dbits, R = 1 << dbits;
    const float smax = ypp_dec(~kmin), smin = ypp_dec(~kmax);               // smallest key word = largest score
    const float scale = __fdiv_rn((float)R - 0.5f, __fsub_rn(smax, smin));
    if (kLinear && !(scale > 0.0f && scale < INFINITY)) return false;       // uniform: no finite score range
    const int limit = kLinear ? YPP_BUCKET_MAX_LINEAR : YPP_BUCKET_MAX;
    for (int i = tid; i < R; i += YPP_ORG_THREADS) table[i] = 0;
    if (tid == 0) hd->max_bucket = 0;
    __syncthreads();
    for (int i = tid; i < n; i += YPP_ORG_THREADS) atomicAdd(&table[bucket_of<kLinear>(src[i], kmin, shift, smax, scale, R)], 1);
    __syncthreads();
    {
        const int ipt = (R + YPP_ORG_THREADS - 1) / YPP_ORG_THREADS;         // 1..4 entries per thread
        const int e0 = tid * ipt;
        int vals[(1 << YPP_BUCKET_BITS) / YPP_ORG_THREADS];
        int sum = 0, mx = 0;
#pragma unroll
        for (int k = 0; k < (1 << YPP_BUCKET_BITS) / YPP_ORG_THREADS; ++k) {
            vals[k] = (k < ipt && e0 + k < R) ? table[e0 + k] : 0;
            sum += vals[k];
            mx = max(mx, vals[k]);
        }
        mx = __reduce_max_sync(YPP_FULL_MASK, mx);
        if (lane == 0 && mx > 0) atomicMax(&hd->max_bucket, mx);
        int total;
        int run = block_exclusive_scan(sum, hd->warp_sums, total);           // syncs inside: max_bucket is complete after it
        if (hd->max_bucket > limit) return false;                            // uniform: every thread reads the same word
#pragma unroll
        for (int k = 0; k < (1 << YPP_BUCKET_BITS) / YPP_ORG_THREADS; ++k)
            if (k < ipt && e0 + k < R) { table[e0 + k] = run; run += vals[k]; }
    }
    __syncthreads();
    for (int i = tid; i < n; i += YPP_ORG_THREADS) {
        const u64 k = src[i];
        tmp[atomicAdd(&table[bucket_of<kLinear>(k, kmin, shift, smax, scale, R)], 1)] = k;
    }
    __syncthreads();                       // table[d] is now the END of bucket d
    for (int i = tid; i < n; i += YPP_ORG_THREADS) {
        const u64 k = tmp[i];
        const int d = bucket_of<kLinear>(k, kmin, shift, smax, scale, R);
        const int lo = d ? table[d - 1] : 0, hi = table[d];
        int rank = 0;
        for (int q = lo; q < hi; ++q) rank += tmp[q] < k;
        dst[lo + rank] = k;
    }
    __syncthreads();
    return true;
}

// ---- images with a handful of candidates (deployment: tens per image at conf 0.25) -----------------------
// The general path is a dozen block-wide phases whose cost is their latency, not their work: 10 us of S1 and as much
// again in N1 + N1b for 50 boxes.  Up to YPP_SMALL_MAX candidates the image's CTA does everything here, exactly as
// torchvision states it and without any of the machinery above:
//   rank      place of a key = number of smaller keys (unique 64-bit words: that IS the stable score order);
//   pairs     every (earlier i, later j) through the exact pair test; coordinate trick: ALL pairs on the shifted boxes
//             (torchvision's single pass over boxes + cls * step - cross-class suppression needs no detector here),
//             otherwise same-class pairs on the plain boxes (_batched_nms_vanilla);
//   greedy    fix-point over the bit rows: j is dropped once a kept earlier box suppresses it, kept once every earlier
//             box that suppresses it is known dropped.
// It leaves what G reads (sorted keys, keep flags by rank, counts) and NO segments, so N1 / N1b find nothing to do.
#define YPP_SMALL_MAX 256
#define YPP_SMALL_WORDS (YPP_SMALL_MAX / 32)
static_assert(sizeof(OrgSmemHeader) % 16 == 0, "the table behind the header holds float4s in the small path");
static_assert(YPP_SMALL_MAX * 4 == YPP_ORG_THREADS, "four threads rank one key");
static_assert((4 * 8 + 16 + 4 * YPP_SMALL_WORDS) * YPP_SMALL_MAX + 64 * YPP_SMALL_WORDS <= 4 * YPP_RADIX_MAX * YPP_TABLE_PITCH,
              "the small path lives in the radix table");

__device__ void organise_small(const YppOrgParams& p, const int b, const int n, const u64* __restrict__ src, int* table,
                               const bool trick, const float step)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = p.N, nc = p.nc;
    float4* sb = reinterpret_cast<float4*>(table);                          // 16-byte aligned: the header's size is a multiple of 16                             // boxes by rank (xyxy, shifted under the trick)
    u64* uk = reinterpret_cast<u64*>(sb + YPP_SMALL_MAX);                   // keys, box-index order
    u64* sk = uk + YPP_SMALL_MAX;                                           // keys by rank
    u32* msk = reinterpret_cast<u32*>(sk + YPP_SMALL_MAX);                  // [j][w]: earlier ranks 32w.. that suppress rank j
    u32* K = msk + YPP_SMALL_MAX * YPP_SMALL_WORDS;                         // kept ranks
    u32* D = K + YPP_SMALL_WORDS;                                           // dropped ranks
    const float4* gbox = p.ws.box + (size_t)b * N;
    u64* skeys = p.ws.skeys + (size_t)b * N;
    unsigned char* kf = p.ws.keepflag + (size_t)b * N;                      // zeroed for ranks < n by the caller

    S1_MARK(3);
    if (tid < n) uk[tid] = src[tid];
    if (tid < YPP_SMALL_WORDS) { K[tid] = 0u; D[tid] = 0u; }
    __syncthreads();
    // ---- rank: four threads per key, each counts a quarter of the others
    {
        const int i = tid >> 2, q = tid & 3;
        const u64 ki = i < n ? uk[i] : 0ull;
        int r = 0;
        if (i < n)
            for (int j = q; j < n; j += 4) r += uk[j] < ki;
        r += __shfl_xor_sync(YPP_FULL_MASK, r, 1);
        r += __shfl_xor_sync(YPP_FULL_MASK, r, 2);
        if (i < n && q == 0) {
            float4 xy = ypp_xyxy(gbox[ypp_key_idx(ki)]);                     // utils/yolo_utils.py:339-343
            if (trick) {
                const float off = __fmul_rn((float)ypp_key_cls(ki), step);  // idxs.to(boxes) * (max_coordinate + 1)
                xy = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
            }
            sb[r] = xy;
            sk[r] = ki;
            skeys[r] = ki;
        }
    }
    __syncthreads();
    S1_MARK(4);
    // ---- pairs: thread (j, w) tests rank j against the earlier ranks of word w
    const int n_in = p.max_cand > 0 ? min(n, p.max_cand) : n;               // ranks past max_cand: never kept, suppress nothing
    const bool neg_thr = !(p.iou_thr >= 0.0f);
    for (int t = tid; t < YPP_SMALL_MAX * YPP_SMALL_WORDS; t += YPP_ORG_THREADS) {
        const int j = t / YPP_SMALL_WORDS, w = t % YPP_SMALL_WORDS;
        u32 bits = 0u;
        if (j < n_in && 32 * w < j) {
            const float4 bj = sb[j];
            const int cj = ypp_key_cls(sk[j]);
            const int i_end = min(32 * w + 32, j);
            for (int i = 32 * w; i < i_end; ++i) {
                if (!trick && ypp_key_cls(sk[i]) != cj) continue;
                const float4 bi = sb[i];
                const float sa = ypp_area(bi);
                const bool sup = neg_thr ? ypp_suppresses<false>(bi, sa, bj, p.iou_thr, p.union_mode)
                                         : ypp_suppresses<true>(bi, sa, bj, p.iou_thr, p.union_mode);
                if (sup) bits |= 1u << (i - 32 * w);
            }
        }
        msk[t] = bits;
    }
    __syncthreads();
    S1_MARK(5);
    // ---- greedy order as a fix-point; thread j < 256 owns rank j, its bit lives in word `warp`
    {
        u32 m[YPP_SMALL_WORDS];
        const bool mine = tid < n_in;
#pragma unroll
        for (int w = 0; w < YPP_SMALL_WORDS; ++w) m[w] = mine ? msk[tid * YPP_SMALL_WORDS + w] : 0u;
        bool undecided = mine;
        for (;;) {
            bool hit = false, wait = false;
#pragma unroll
            for (int w = 0; w < YPP_SMALL_WORDS; ++w) {
                hit |= (m[w] & K[w]) != 0u;
                wait |= (m[w] & ~D[w]) != 0u;
            }
            const bool drop_now = undecided && hit, keep_now = undecided && !wait;      // !wait implies !hit
            undecided = undecided && !drop_now && !keep_now;
            __syncthreads();                                                // everyone has read K and D
            if (warp < YPP_SMALL_WORDS) {
                const u32 kb = __ballot_sync(YPP_FULL_MASK, keep_now), db = __ballot_sync(YPP_FULL_MASK, drop_now);
                if (lane == 0) { K[warp] |= kb; D[warp] |= db; }
            }
            if (!__syncthreads_or(undecided)) break;
        }
    }
    S1_MARK(6);
    if (tid < n_in && ((K[warp] >> lane) & 1u)) kf[tid] = 1;
    if (tid == 0) {
        int kept = 0;
#pragma unroll
        for (int w = 0; w < YPP_SMALL_WORDS; ++w) kept += __popc(K[w]);
        p.ws.keep_count[b] = kept;
        p.ws.bigcnt[b] = 0;
        p.ws.imgflag[b] = (trick ? 2 : 0) | 4;                              // bit 2: the image went this way (diagnostics)
        p.ws.img_step[b] = step;
    }
    int* gseg = p.ws.seg_start + (size_t)b * (nc + 1);
    for (int c = tid; c <= nc; c += YPP_ORG_THREADS) gseg[c] = 0;
    S1_MARK(7);
    S1_MARK(8);
}

__global__ void __launch_bounds__(YPP_ORG_THREADS, 1)
ypp_organise_kernel(const __grid_constant__ YppOrgParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    OrgSmemHeader* hd = reinterpret_cast<OrgSmemHeader*>(smem_raw);
    int* table = reinterpret_cast<int*>(smem_raw + sizeof(OrgSmemHeader));   // radix counters / tile offsets / class table
    int* seg = table + 32 * p.nc;                                            // [nc+1] (class-table phase)
    const size_t koff = (sizeof(OrgSmemHeader) + sizeof(int) * (size_t)p.table_ints + 15) / 16 * 16;

    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nc = p.nc, N = p.N;
    const u64* gkeys = p.ws.keys + (size_t)b * N;
    u64* skeys = p.ws.skeys + (size_t)b * N;
    float4* sbox = p.ws.sbox + (size_t)b * N;
    u64* gpk = p.ws.gpk + (size_t)b * N;
    int* gseg = p.ws.seg_start + (size_t)b * (nc + 1);
    // the two key buffers of the sort: both in shared memory when they fit, else the first one only (the
    // second in the not-yet-written sbox region), else both global (skeys / sbox)
    u64* bufA = p.keys_in_smem ? reinterpret_cast<u64*>(smem_raw + koff) : skeys;
    u64* bufB = p.keys_in_smem == 2 ? bufA + N : reinterpret_cast<u64*>(sbox);

    ypp_grid_wait();
    S1_CTA_BEGIN();
    S1_MARK(0);
    // ---- 1. tile offsets (exclusive scan of the D1 tile counts) and per-image extrema.  Everything this step reads is
    // asked for at once, BEFORE the slot count arrives (count -> counts -> first index -> keys used to be four dependent
    // loads deep): words of slots past the count are allocated but unwritten, and masked once the count is here
    const int tiles_max = p.ws.tiles_max;
    const int* tcnt = p.ws.tile_cnt + (size_t)b * tiles_max;
    const uint4* text = reinterpret_cast<const uint4*>(p.ws.tile_ext) + (size_t)b * tiles_max;
    int* tile_off = table;                                   // [ntiles + 1], dead before the first radix pass
    int* tile_first_s = table + tiles_max + 1;               // [ntiles] first box index of each slot, likewise
    int v0 = 0, tf0 = 0;
    uint4 e0 = make_uint4(0u, 0u, 0u, 0u);
    if (tid < tiles_max) { v0 = tcnt[tid]; e0 = text[tid]; tf0 = p.ws.tile_first[tid]; }
    const int ntiles = min(*p.ws.ntiles, tiles_max);
    if (tid == 0) {
        hd->maxc = 0; hd->nminx1 = 0; hd->nminy1 = 0;
        hd->n_low = 0; hd->n_high = 0; hd->flagged = 0; hd->n_edges = 0; hd->n_big = 0;
        hd->kmin = 0xffffffffu; hd->kmax = 0u;
    }
    if (tid < YPP_DET_BINS) { hd->bin_hi[tid] = 0; hd->bin_lo[tid] = 0; }
    {
        int run = 0;
        u32 M = 0, NX = 0, NY = 0;
        for (int t0 = 0; t0 < ntiles; t0 += YPP_ORG_THREADS) {
            const int t = t0 + tid;
            int v = v0, tf = tf0;
            uint4 e = e0;
            if (t0 > 0 && t < ntiles) { v = tcnt[t]; e = text[t]; tf = p.ws.tile_first[t]; }
            if (t >= ntiles) v = 0;
            if (v > 0) { M = max(M, e.x); NX = max(NX, e.y); NY = max(NY, e.z); }
            int total;
            const int ex = block_exclusive_scan(v, hd->warp_sums, total);   // syncs: the header init above is visible
            if (t < ntiles) { tile_off[t] = run + ex; tile_first_s[t] = tf; }
            run += total;
        }
        M = __reduce_max_sync(YPP_FULL_MASK, M);
        NX = __reduce_max_sync(YPP_FULL_MASK, NX);
        NY = __reduce_max_sync(YPP_FULL_MASK, NY);
        if (lane == 0 && M) { atomicMax(&hd->maxc, M); atomicMax(&hd->nminx1, NX); atomicMax(&hd->nminy1, NY); }
        if (tid == 0) {
            tile_off[ntiles] = run;
            hd->n = min(run, N);
            p.ws.cand_count[b] = min(run, N);
            p.ws.keep_count[b] = 0;
        }
    }
    __syncthreads();
    const int n = hd->n;
    if (n == 0) {
        for (int c = tid; c <= nc; c += YPP_ORG_THREADS) gseg[c] = 0;
        if (tid == 0) p.ws.bigcnt[b] = 0;
        return;
    }
    S1_MARK(1);
    // ---- 2. gather the slots, track the range of the score words
    {
        u32 kmin = 0xffffffffu, kmax = 0u;
        // slots hold <= 32 keys: one lane each; four slots per thread in flight
        for (int i0 = tid; i0 < ntiles * 32; i0 += 4 * YPP_ORG_THREADS) {
            u64 k[4];
            int dst[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = (i0 + u * YPP_ORG_THREADS) >> 5;
                dst[u] = -1;
                if (t < ntiles) {
                    const int off = tile_off[t], cnt = tile_off[t + 1] - off;
                    if (lane < cnt) { dst[u] = off + lane; k[u] = gkeys[tile_first_s[t] + lane]; }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (dst[u] >= 0) {
                    bufA[dst[u]] = k[u];
                    const u32 hi = (u32)(k[u] >> 32);
                    kmin = min(kmin, hi); kmax = max(kmax, hi);
                }
            }
        }
        kmin = __reduce_min_sync(YPP_FULL_MASK, kmin);
        kmax = __reduce_max_sync(YPP_FULL_MASK, kmax);
        if (lane == 0) { atomicMin(&hd->kmin, kmin); atomicMax(&hd->kmax, kmax); }
        unsigned char* kf = p.ws.keepflag + (size_t)b * N;
        for (int i = tid; i < n; i += YPP_ORG_THREADS) kf[i] = 0;
    }
    __syncthreads();
    S1_MARK(2);
    // ---- which torchvision path would run for this image, and the geometry of the cross-class detector
    bool trick = p.offset_mode == YPP_OFFSET_TRICK;
    if (p.offset_mode == YPP_OFFSET_AUTO_CPU) trick = 4LL * n <= 4000;
    if (p.offset_mode == YPP_OFFSET_AUTO_CUDA) trick = 4LL * n <= 100000;
    const float M = ypp_dec(hd->maxc);
    float step = __fadd_rn(M, 1.0f);                       // max_coordinate + 1
    // fixed class offsets (boxes + cls * max_wh, one torchvision.ops.nms over everything): the class bands
    // are disjoint as long as no coordinate exceeds step - 1, exactly the situation the detector below
    // was built for; otherwise (or step 0 = class-agnostic, or NaN) the image is one segment
    bool one_segment = false;
    if (p.step_mode) {
        trick = true;
        step = p.fixed_step;
        one_segment = !(M <= __fadd_rn(step, -1.0f)) || !(step > 0.0f);
    }
    if (n <= p.small_max) {                                  // uniform over the CTA
        organise_small(p, b, n, bufA, table, trick, step);
        return;
    }
    DetGeom dg;
    dg.step = step;
    dg.x1min = -ypp_dec(hd->nminx1);
    dg.slack = fmaf(fabsf(step) * (float)nc, 1e-6f, 1e-3f);
    dg.lim = -1.0f + dg.slack;
    // no detector when offsets are not finite (every shifted IoU is NaN on both paths) or when no
    // box reaches x1 < -1 (nothing can be shifted onto the previous class band)
    dg.on = trick && nc > 1 && !one_segment && (fabsf(M) <= 3.0e38f) && dg.x1min < dg.lim;
    {
        const float range = dg.lim - dg.x1min;
        dg.inv_w = (range > 64.0f * 4.0f * dg.slack) ? (float)YPP_DET_BINS / range : 0.0f;
    }
    const bool detect = dg.on;
    const float4* gbox = p.ws.box + (size_t)b * N;
    if (trick && nc > 1 && (one_segment || !(step > 0.0f)) && tid == 0) hd->flagged = 1;   // degenerate offsets: be exact

    // ---- 3. detector sweeps, BEFORE the sort: the keys are still in box-index order, so the box reads
    // are (nearly) sequential instead of a random gather
    if (detect) {
        for (int i0 = tid; i0 < n; i0 += 4 * YPP_ORG_THREADS) {       // four box reads in flight per thread
            float4 bx[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int i = i0 + q * YPP_ORG_THREADS;
                if (i < n) bx[q] = gbox[ypp_key_idx(bufA[i])];
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (i0 + q * YPP_ORG_THREADS >= n) break;
                const float4 xy = ypp_xyxy(bx[q]);
                const float u = xy.z - step, v = xy.w - step;
                // read first: after a few updates nearly every candidate loses to the bin's current value
                if (u > dg.x1min - dg.slack) {
                    const int j = det_bin(dg, u);
                    const u32 e = ypp_enc(v);
                    if (e > hd->bin_hi[j]) atomicMax(&hd->bin_hi[j], e);
                }
                if (xy.x < dg.lim && xy.y < dg.lim) {
                    const int j = det_bin(dg, xy.x);
                    const u32 e = ypp_enc(-xy.y);
                    if (e > hd->bin_lo[j]) atomicMax(&hd->bin_lo[j], e);
                }
            }
        }
        __syncthreads();
        S1_MARK(12);
        if (warp == 1) {
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
        __syncthreads();
        S1_MARK(13);
        for (int i0 = tid; i0 < n; i0 += 4 * YPP_ORG_THREADS) {
          float4 bx[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
              const int i = i0 + q * YPP_ORG_THREADS;
              if (i < n) bx[q] = gbox[ypp_key_idx(bufA[i])];
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int i = i0 + q * YPP_ORG_THREADS;
            if (i >= n) break;
            const u64 k = bufA[i];
            const float4 xy = ypp_xyxy(bx[q]);
            // B role: some A in an x-bin >= mine has v_a > y1 ; A role: some B in an x-bin <= mine has y1 < v_a
            bool is_low = false, is_high = false;
            if (xy.x < dg.lim && xy.y < dg.lim) {
                const int j = max(det_bin(dg, xy.x) - 1, 0);
                const u32 e = hd->bin_hi[j];
                is_low = e != 0 && xy.y < ypp_dec(e) + dg.slack;
            }
            const float u = xy.z - step, v = xy.w - step;
            if (u > dg.x1min - dg.slack) {
                const int j = min(det_bin(dg, u) + 1, YPP_DET_BINS - 1);
                const u32 e = hd->bin_lo[j];
                is_high = e != 0 && v > -ypp_dec(e) - dg.slack;
            }
            if (is_low || is_high) {
                const int c = ypp_key_cls(k);
                const float off = __fmul_rn((float)c, step);
                DetEntry e;
                e.box = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
                e.cls = c; e.key = k;
                if (is_low) { const int s = atomicAdd(&hd->n_low, 1); if (s < YPP_DET_LIST) hd->lowl[s] = e; }
                if (is_high) { const int s = atomicAdd(&hd->n_high, 1); if (s < YPP_DET_LIST) hd->highl[s] = e; }
            }
          }
        }
        __syncthreads();
        S1_MARK(14);
        // exact test of the candidate pairs: a cross-class pair whose SHIFTED boxes suppress is an edge
        {
            const int nl = hd->n_low, nh = hd->n_high;
            if (nl > 0 && nh > 0) {
                const bool neg_thr = !(p.iou_thr >= 0.0f);
                auto test_pair = [&](const DetEntry& x, const DetEntry& y) {
                    if (x.cls == y.cls) return;
                    // boxes that do not intersect have IoU 0 (or NaN): no edge unless thr is negative
                    if (!neg_thr && !(fmaxf(x.box.x, y.box.x) < fminf(x.box.z, y.box.z) &&
                                      fmaxf(x.box.y, y.box.y) < fminf(x.box.w, y.box.w))) return;
                    const DetEntry& row = x.key < y.key ? x : y;               // smaller key = earlier in score order
                    const DetEntry& col = x.key < y.key ? y : x;
                    const float sa = ypp_area(row.box);
                    const bool sup = neg_thr ? ypp_suppresses<false>(row.box, sa, col.box, p.iou_thr, p.union_mode)
                                             : ypp_suppresses<true>(row.box, sa, col.box, p.iou_thr, p.union_mode);
                    if (sup) {
                        const int e = atomicAdd(&hd->n_edges, 1);
                        if (e < YPP_DET_EDGES) hd->edges[e] = ((u32)row.cls << 16) | (u32)col.cls;
                    }
                };
                const bool lo_fits = nl <= YPP_DET_LIST, hi_fits = nh <= YPP_DET_LIST;
                if (!lo_fits && !hi_fits) {
                    if (tid == 0) hd->flagged = 1;
                } else if (lo_fits && hi_fits) {
                    for (int t = tid; t < nl * nh; t += YPP_ORG_THREADS) test_pair(hd->lowl[t / nh], hd->highl[t % nh]);
                } else {
                    // one role has too many candidates to list: test EVERY box that meets the role's necessary condition
                    // against the list that fits (keys are still in box-index order: sequential box reads)
                    const float y1min = -ypp_dec(hd->nminy1);
                    const DetEntry* lst = lo_fits ? hd->lowl : hd->highl;
                    const int nlst = lo_fits ? nl : nh;
                    for (int i = tid; i < n; i += YPP_ORG_THREADS) {
                        DetEntry e;
                        e.key = bufA[i];
                        e.cls = ypp_key_cls(e.key);
                        const float4 xy = ypp_xyxy(gbox[ypp_key_idx(e.key)]);
                        // the role's necessary condition first (same tests as the detector sweep)
                        if (lo_fits ? !(xy.z - step > dg.x1min - dg.slack && xy.w - step > y1min - dg.slack)
                                    : !(xy.x < dg.lim && xy.y < dg.lim)) continue;
                        const float off = __fmul_rn((float)e.cls, step);
                        e.box = make_float4(__fadd_rn(xy.x, off), __fadd_rn(xy.y, off), __fadd_rn(xy.z, off), __fadd_rn(xy.w, off));
                        for (int j = 0; j < nlst; ++j) test_pair(e, lst[j]);
                    }
                }
            }
        }
        __syncthreads();
    }
    S1_MARK(3);
    // ---- 4. stable sort by score word
    const int L = ((n + 31) / 32 + 31) / 32 * 32;            // ranks per warp, multiple of 32
    u64* keys = bufA;                                        // where the sorted keys end up
    {
        const u32 kmin = hd->kmin, span = hd->kmax - kmin;
        const int bits = 32 - __clz(span);                   // 0 when every score is equal: already in order
        // bucket + rank (sorted keys land back in bufA); LSD passes only when scores tie in crowds
        if (bits > 0 && !bucket_rank_sort<true>(bufA, bufB, bufA, n, bits, kmin, hd->kmax, table, hd) &&
            !bucket_rank_sort<false>(bufA, bufB, bufA, n, bits, kmin, hd->kmax, table, hd)) {
            const int passes = (bits + 7) / 8;
            const int bw = (bits + passes - 1) / passes;
            u64* other = bufB;
            for (int ps = 0; ps < passes; ++ps) {
                radix_pass(keys, other, n, L, ps * bw, bw, kmin, table, hd->warp_sums);
                u64* t = keys; keys = other; other = t;
            }
        }
    }
    S1_MARK(4);
    if (keys != skeys) {
        for (int i = tid; i < n; i += YPP_ORG_THREADS) skeys[i] = keys[i];
        if (!p.keys_in_smem) { __syncthreads(); keys = skeys; }
    }
    S1_MARK(5);
    // ---- 5. coordinate trick: an edge couples two classes: they are merged into one segment (greedy NMS
    // over the union in score order is then exactly what torchvision's single pass does; classes that
    // share no edge stay independent).  Too many candidates or edges -> one class-agnostic segment.
    __syncthreads();                                         // table is free again
    int* cmap = seg + nc + 1;                                // [nc] class -> segment that holds it
    const int ne = hd->n_edges;
    const bool was_flagged = hd->flagged != 0;               // read by every thread before anyone sets it below
    const bool merge_all = trick && nc > 1 && (was_flagged || ne > YPP_DET_EDGES);
    for (int c = tid; c < nc; c += YPP_ORG_THREADS) cmap[c] = merge_all ? 0 : c;
    for (int i = tid; i < 32 * nc; i += YPP_ORG_THREADS) table[i] = 0;       // the per-(warp, class) counters of step 6
    __syncthreads();
    if (merge_all) {
        if (tid == 0) hd->flagged = 1;
    } else if (trick && nc > 1 && ne > 0) {                  // uniform; rare
        if (tid == 0) {                                      // union by smallest class id; a handful of edges
            for (int e = 0; e < ne; ++e) {
                int a = (int)(hd->edges[e] >> 16), c2 = (int)(hd->edges[e] & 0xffffu);
                while (cmap[a] != a) a = cmap[a];
                while (cmap[c2] != c2) c2 = cmap[c2];
                if (a != c2) cmap[max(a, c2)] = min(a, c2);
            }
            hd->flagged = 1;
        }
        __syncthreads();
        for (int c0 = 0; c0 < nc; c0 += YPP_ORG_THREADS) {
            const int c = c0 + tid;
            int r = c < nc ? c : 0;
            while (cmap[r] != r) r = cmap[r];
            __syncthreads();
            if (c < nc) cmap[c] = r;
        }
        __syncthreads();
    }
    // ---- 6. class counting (warp w owns ranks [w*L, (w+1)*L)), segment starts.  table[w*nc + c] ends up as the number
    // of class-c ranks in the warps before w, i.e. warp w's first position RELATIVE to the segment start seg[c]
    const int n_in = p.max_cand > 0 ? min(n, p.max_cand) : n;        // ranks past max_cand join no segment: never kept
    const int r_begin = warp * L, r_end = min(n_in, r_begin + L);
    const int cbits = 32 - __clz(max(nc - 1, 1));            // bits of a class id
    int* my = table + warp * nc;
    for (int r0 = r_begin; r0 < r_end; r0 += 32) {
        const int r = r0 + lane;
        if (r < r_end) atomicAdd(&my[cmap[ypp_key_cls(keys[r])]], 1);
    }
    __syncthreads();
    for (int c = warp; c < nc; c += YPP_ORG_THREADS / 32) {  // one warp per class: lane = the warp whose count it holds
        const int v = table[lane * nc + c];
        int inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(YPP_FULL_MASK, inc, d);
            if (lane >= d) inc += o;
        }
        table[lane * nc + c] = inc - v;
        if (lane == 31) seg[c] = inc;
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
    S1_MARK(6);
    // ---- 7. stable ranking: segment position -> global rank, box index and class in one word (N1 gathers the boxes
    // through it: one coalesced load, then the box - not rank, then key, then box)
    // two 32-rank chunks (A, then B) per step: B's positions also count A's lanes of the same class,
    // found with the ballots A needs anyway
    const u32 lt = (1u << lane) - 1u;
    for (int r0 = r_begin; r0 < r_end; r0 += 64) {
        const int rA = r0 + lane, rB = rA + 32;
        const bool okA = rA < r_end, okB = rB < r_end;
        const u64 kA = okA ? keys[rA] : 0ull, kB = okB ? keys[rB] : 0ull;
        const int cA = okA ? cmap[ypp_key_cls(kA)] : 0;
        const int cB = okB ? cmap[ypp_key_cls(kB)] : 0;
        u32 pA = __ballot_sync(YPP_FULL_MASK, okA), pB = __ballot_sync(YPP_FULL_MASK, okB);
        u32 xB = pA, xA = pB;        // xB: A lanes of my B class; xA: B lanes of my A class
        for (int i = 0; i < cbits; ++i) {
            const bool bitA = (cA >> i) & 1, bitB = (cB >> i) & 1;
            const u32 balA = __ballot_sync(YPP_FULL_MASK, bitA), balB = __ballot_sync(YPP_FULL_MASK, bitB);
            pA &= bitA ? balA : ~balA;
            xA &= bitA ? balB : ~balB;
            pB &= bitB ? balB : ~balB;
            xB &= bitB ? balA : ~balA;
        }
        const int rankA = __popc(pA & lt), inB = __popc(pB & lt);
        const int baseA = okA ? my[cA] : 0, baseB = okB ? my[cB] : 0;
        const int segA = seg[cA], segB = seg[cB];
        __syncwarp();
        if (okA) {
            if (rankA == 0) my[cA] = baseA + __popc(pA) + __popc(xA);
            gpk[segA + baseA + rankA] = ypp_make_gpk(rA, kA);
        }
        if (okB) {
            if (inB == 0 && xB == 0u) my[cB] = baseB + __popc(pB);     // class absent from A: B's leader counts it
            gpk[segB + baseB + __popc(xB) + inB] = ypp_make_gpk(rB, kB);
        }
        __syncwarp();
    }
    S1_MARK(7);
    __syncthreads();
    if (tid == 0) {
        // bits 8.. are diagnostics: detector list lengths and suppressing cross-class pairs (each clamped to 255)
        p.ws.imgflag[b] = (hd->flagged ? 1 : 0) | (trick ? 2 : 0) | (min(hd->n_low, 255) << 8) | (min(hd->n_high, 255) << 16) |
                          (min(hd->n_edges, 255) << 24);
        p.ws.img_step[b] = step;
    }
    for (int c = tid; c <= nc; c += YPP_ORG_THREADS) gseg[c] = seg[c];
    for (int c = tid; c < nc; c += YPP_ORG_THREADS) {
        if (seg[c + 1] - seg[c] > YPP_WARP_SEG) {
            const int s = atomicAdd(&hd->n_big, 1);
            if (s < p.ws.bigq_per_image) p.ws.bigq[b * p.ws.bigq_per_image + s] = c;
        }
    }
    __syncthreads();
    if (tid == 0) p.ws.bigcnt[b] = hd->n_big;
    S1_MARK(8);
    S1_CTA_END();
}

}  // namespace

// shared memory: header + max(radix table, class table, tile offsets) + (two key buffers when they fit)
int ypp_org_smem_bytes(int N, int nc, int* keys_in_smem, int* table_ints)
{
    int tints = YPP_RADIX_MAX * YPP_TABLE_PITCH;
    if (34 * nc + 1 > tints) tints = 34 * nc + 1;
    const int tiles_max = (N / YPP_TILE_BOXES + YPP_MAX_SCALES * YPP_MAX_ANCHORS + 8) * (YPP_TILE_BOXES / 32);
    if (2 * tiles_max + 1 > tints) tints = 2 * tiles_max + 1;     // tile offsets + first indices
    *table_ints = tints;
    size_t fixed = sizeof(OrgSmemHeader) + sizeof(int) * (size_t)tints;
    fixed = (fixed + 15) / 16 * 16;
    const size_t cap = 227 * 1024;
    if (fixed > cap) return -1;
    if (fixed + (size_t)N * 16 <= cap) { *keys_in_smem = 2; return (int)(fixed + (size_t)N * 16); }
    if (fixed + (size_t)N * 8 <= cap) { *keys_in_smem = 1; return (int)(fixed + (size_t)N * 8); }
    *keys_in_smem = 0;
    return (int)fixed;
}

#ifdef YPP_PROFILE_S1
extern "C" int ypp_debug_s1_clocks(long long* out16)
{
    const int rc = (int)cudaMemcpyFromSymbol(out16, ypp_s1_clk, sizeof(long long) * 16);
    long long reset[16] = {0};
    reset[10] = 0x7fffffffffffffffLL;
    cudaMemcpyToSymbol(ypp_s1_clk, reset, sizeof(reset));
    return rc;
}
#endif

int ypp_launch_organise(const YppOrgParams& p, cudaStream_t st)
{
    int kin = 0, tints = 0;
    const int smem = ypp_org_smem_bytes(p.N, p.nc, &kin, &tints);
    if (smem < 0) { ypp_set_error("organise: nc=%d needs more shared memory than one CTA has", p.nc); return YPP_E_UNSUPPORTED; }
    YppOrgParams q = p;
    q.keys_in_smem = kin;
    q.table_ints = tints;
    static int small_max = -1;                             // YPP_SMALL: tuning / tests only (0 = every image takes the general path)
    if (small_max < 0) { const char* e = getenv("YPP_SMALL"); small_max = e ? atoi(e) : YPP_SMALL_MAX; if (small_max > YPP_SMALL_MAX) small_max = YPP_SMALL_MAX; if (small_max < 0) small_max = 0; }
    q.small_max = small_max;
    static int configured_d[YPP_MAX_DEVICES];              // per device, like the attribute itself
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= YPP_MAX_DEVICES) { ypp_set_error("organise: device ordinal %d out of range", dev); return YPP_E_UNSUPPORTED; }
    int& configured_smem = configured_d[dev];
    if (smem > configured_smem) {
        cudaError_t e = cudaFuncSetAttribute(ypp_organise_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) { ypp_set_error("organise: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
        configured_smem = 227 * 1024;
    }
    ypp_launch_chain(1, ypp_organise_kernel, dim3(p.B), dim3(YPP_ORG_THREADS), smem, st, q);
    ypp_count_launch(1);
    return ypp_check_launch("organise");
}
