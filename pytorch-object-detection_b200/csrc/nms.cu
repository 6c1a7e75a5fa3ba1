// N1 + N2: bitmask IoU suppression and greedy keep-scan per (image, class) segment, and the
// final row gather.
//
// Reference semantics: torchvision.ops.batched_nms as called by nms_v3
// (utils/yolo_utils.py:346) -> torch.ops.torchvision.nms: boxes visited in descending score
// (stable), box j suppressed iff an earlier KEPT box i has IoU(i,j) > thr (strict).
//
// torchvision's CUDA kernel writes an n x n/64 bit matrix to HBM and scans it with one CTA.
// Here a segment is processed by one CTA in 64-row tiles, entirely on chip:
//   - the 64x64 diagonal tile becomes a bitmask in shared memory (warp ballots);
//   - one warp resolves the tile greedily (64 dependent steps on a register bitmask);
//   - only the KEPT rows of the tile are then broadcast against the later columns that are
//     still alive, setting bits in a shared-memory `removed` bitmap.
// No mask ever reaches HBM and suppressed rows never do any pair work.
#include "ypp_kernels.cuh"

namespace {

struct TileSmem {
    u32 diag[64][2];          // diag[i] = columns j (tile-local) suppressed by row i, j > i
    float4 kbox[64];          // kept rows of the current tile, compacted
    float karea[64];
    int nkept;
    int total_kept;
};

// Process one segment.  seg: boxes in score order (global memory), cache: shared staging for
// the first `cache_cap` boxes, removed: bitmap words (>= (len+31)/32).
template <int THREADS, bool kEarlyOut>
__device__ void nms_segment(const float4* __restrict__ seg, const int* __restrict__ grank,
                            unsigned char* __restrict__ keepflag, int len,
                            float4* cache, int cache_cap, u32* removed, TileSmem& ts,
                            float thr, int union_fma)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncache = min(len, cache_cap);
    for (int i = tid; i < ncache; i += THREADS) cache[i] = seg[i];
    for (int i = tid; i < (len + 31) / 32; i += THREADS) removed[i] = 0;
    if (tid == 0) ts.total_kept = 0;
    __syncthreads();

    for (int t0 = 0; t0 < len; t0 += 64) {
        const int rows = min(64, len - t0);
        // ---- diagonal tile: pair (i, j), i = p / 64, j = p % 64; a warp covers half a row
        for (int p = tid; p < 64 * 64; p += THREADS) {
            const int i = p >> 6, j = p & 63;
            bool bit = false;
            if (j > i && j < rows) {     // j < rows implies i < rows
                const float4 a = (t0 + i) < ncache ? cache[t0 + i] : seg[t0 + i];
                const float4 bq = (t0 + j) < ncache ? cache[t0 + j] : seg[t0 + j];
                bit = ypp_suppresses<kEarlyOut>(a, ypp_area(a), bq, thr, union_fma);
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
                ts.kbox[s] = bx; ts.karea[s] = ypp_area(bx);
                keepflag[grank[r]] = 1;
            }
            if ((khi >> lane) & 1u) {
                const int s = nlo + __popc(khi & ((1u << lane) - 1u));
                const int r = t0 + 32 + lane;
                const float4 bx = r < ncache ? cache[r] : seg[r];
                ts.kbox[s] = bx; ts.karea[s] = ypp_area(bx);
                keepflag[grank[r]] = 1;
            }
            if (lane == 0) { ts.nkept = nlo + __popc(khi); ts.total_kept += nlo + __popc(khi); }
        }
        __syncthreads();
        // ---- broadcast kept rows against later, still-alive columns
        const int nk = ts.nkept;
        if (nk > 0) {
            for (int j = t0 + 64 + tid; j < len; j += THREADS) {
                if ((removed[j >> 5] >> (j & 31)) & 1u) continue;
                const float4 bq = j < ncache ? cache[j] : seg[j];
                bool dead = false;
                for (int r = 0; r < nk; ++r) {
                    if (ypp_suppresses<kEarlyOut>(ts.kbox[r], ts.karea[r], bq, thr, union_fma)) { dead = true; break; }
                }
                if (dead) atomicOr(&removed[j >> 5], 1u << (j & 31));
            }
        }
        __syncthreads();
    }
}

// ---- small segments: one CTA per (class, image)
template <bool kEarlyOut>
__global__ void __launch_bounds__(YPP_NMS_SMALL_THREADS)
ypp_nms_small_kernel(const __grid_constant__ YppNmsParams p)
{
    __shared__ float4 cache[YPP_SMALL_SEG];
    __shared__ u32 removed[YPP_SMALL_SEG / 32 + 2];
    __shared__ TileSmem ts;
    const int c = blockIdx.x, b = blockIdx.y;
    const int* seg_start = p.ws.seg_start + (size_t)b * (p.nc + 1);
    const int s0 = seg_start[c], len = seg_start[c + 1] - s0;
    if (len <= 0 || len > YPP_SMALL_SEG) return;
    const size_t base = (size_t)b * p.N;
    nms_segment<YPP_NMS_SMALL_THREADS, kEarlyOut>(p.ws.sbox + base + s0, p.ws.grank + base + s0,
                                                  p.ws.keepflag + base, len, cache, YPP_SMALL_SEG,
                                                  removed, ts, p.iou_thr, p.union_mode);
    if (threadIdx.x == 0 && ts.total_kept) atomicAdd(p.ws.keep_count + b, ts.total_kept);
}

// ---- big segments: persistent CTAs pull (image, class) pairs from the queue S1 filled
template <bool kEarlyOut>
__global__ void __launch_bounds__(YPP_NMS_BIG_THREADS, 1)
ypp_nms_big_kernel(const __grid_constant__ YppNmsParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* cache = reinterpret_cast<float4*>(smem_raw);
    u32* removed = reinterpret_cast<u32*>(smem_raw + sizeof(float4) * YPP_BIG_CACHE);
    __shared__ TileSmem ts;
    __shared__ int item;
    const int qn = p.ws.q[0];
    for (;;) {
        if (threadIdx.x == 0) item = atomicAdd(p.ws.q + 1, 1);
        __syncthreads();
        const int it = item;
        __syncthreads();
        if (it >= qn) return;
        const int e = p.ws.bigq[it];
        const int b = e / p.nc, c = e - b * p.nc;
        const int* seg_start = p.ws.seg_start + (size_t)b * (p.nc + 1);
        const int s0 = seg_start[c], len = seg_start[c + 1] - s0;
        const size_t base = (size_t)b * p.N;
        nms_segment<YPP_NMS_BIG_THREADS, kEarlyOut>(p.ws.sbox + base + s0, p.ws.grank + base + s0,
                                                    p.ws.keepflag + base, len, cache, YPP_BIG_CACHE,
                                                    removed, ts, p.iou_thr, p.union_mode);
        if (threadIdx.x == 0 && ts.total_kept) atomicAdd(p.ws.keep_count + b, ts.total_kept);
        __syncthreads();
    }
}

// ---- gather: keep flags (global score order) -> packed rows + kept box indices
#define YPP_GATHER_THREADS 512
__global__ void __launch_bounds__(YPP_GATHER_THREADS)
ypp_gather_kernel(const __grid_constant__ YppGatherParams p)
{
    __shared__ int warp_tot[YPP_GATHER_THREADS / 32];
    __shared__ long long base_s;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = p.N;
    const int n = min(p.ws.cand_count[b], N);
    // rows of the images before this one
    long long part = 0;
    for (int i = tid; i < b; i += YPP_GATHER_THREADS) part += p.ws.keep_count[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(YPP_FULL_MASK, part, d);
    if (tid == 0) base_s = 0;
    __syncthreads();
    if (lane == 0 && part) atomicAdd((unsigned long long*)&base_s, (unsigned long long)part);
    __syncthreads();
    const long long base = base_s;
    if (tid == 0) {
        if (p.cand_count_out) p.cand_count_out[b] = n;
        if (p.keep_count_out) p.keep_count_out[b] = p.ws.keep_count[b];
        if (b == (int)gridDim.x - 1 && p.det_total) *p.det_total = (int)(base + p.ws.keep_count[b]);
    }
    const unsigned char* kf = p.ws.keepflag + (size_t)b * N;
    const u64* keys = p.ws.keys + (size_t)b * N;
    const float4* box = p.ws.box + (size_t)b * N;
    const float img = p.img_idx_base + (float)b;
    int running = 0;
    for (int r0 = 0; r0 < n; r0 += YPP_GATHER_THREADS) {
        const int r = r0 + tid;
        const bool f = r < n && kf[r];
        const u32 m = __ballot_sync(YPP_FULL_MASK, f);
        if (lane == 0) warp_tot[warp] = __popc(m);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < YPP_GATHER_THREADS / 32; ++w) {
            const int t = warp_tot[w];
            if (w < warp) before += t;
            total += t;
        }
        if (f) {
            const int pos = running + before + __popc(m & ((1u << lane) - 1u));
            const u64 k = keys[r];
            const int idx = ypp_key_idx(k);
            if (p.keep_idx) p.keep_idx[(size_t)b * N + pos] = idx;
            const long long row = base + pos;
            if (p.det_rows && row < p.det_capacity) {
                const float4 bx = box[idx];
                float cls = (float)ypp_key_cls(k);
                if (p.src_rows) cls = p.src_rows[((size_t)b * N + idx) * 6 + 5];
                float* o = p.det_rows + row * p.row_width;
                if (p.row_width == 7) *o++ = img;
                o[0] = bx.x; o[1] = bx.y; o[2] = bx.z; o[3] = bx.w;
                o[4] = ypp_key_score(k); o[5] = cls;
            }
        }
        running += total;
        __syncthreads();
    }
}

}  // namespace

int ypp_launch_nms(const YppNmsParams& p, cudaStream_t st)
{
    const bool early = p.iou_thr >= 0.0f;
    dim3 grid(p.nc, p.B);
    if (early) ypp_nms_small_kernel<true><<<grid, YPP_NMS_SMALL_THREADS, 0, st>>>(p);
    else       ypp_nms_small_kernel<false><<<grid, YPP_NMS_SMALL_THREADS, 0, st>>>(p);
    ypp_count_launch(1);
    int rc = ypp_check_launch("nms_small");
    if (rc) return rc;
    if (p.N > YPP_SMALL_SEG) {
        const int smem = (int)(sizeof(float4) * YPP_BIG_CACHE + sizeof(u32) * ((size_t)p.N / 32 + 2));
        static bool configured = false;
        if (!configured) {
            cudaError_t e1 = cudaFuncSetAttribute(ypp_nms_big_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
            cudaError_t e2 = cudaFuncSetAttribute(ypp_nms_big_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
            if (e1 != cudaSuccess || e2 != cudaSuccess) { ypp_set_error("nms_big: cudaFuncSetAttribute failed"); return YPP_E_CUDA; }
            configured = true;
        }
        if (smem > 220 * 1024) { ypp_set_error("nms_big: %d boxes per image exceed the removed-bitmap budget", p.N); return YPP_E_UNSUPPORTED; }
        int ctas = p.B * p.ws.bigq_per_image;
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (ctas > sms) ctas = sms;
        if (early) ypp_nms_big_kernel<true><<<ctas, YPP_NMS_BIG_THREADS, smem, st>>>(p);
        else       ypp_nms_big_kernel<false><<<ctas, YPP_NMS_BIG_THREADS, smem, st>>>(p);
        ypp_count_launch(1);
        rc = ypp_check_launch("nms_big");
    }
    return rc;
}

int ypp_launch_gather(const YppGatherParams& p, cudaStream_t st)
{
    ypp_gather_kernel<<<p.B, YPP_GATHER_THREADS, 0, st>>>(p);
    ypp_count_launch(1);
    return ypp_check_launch("gather");
}
