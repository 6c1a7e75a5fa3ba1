// D1: fused head decode + confidence threshold + stream compaction (and the dense decode).
//
// Reference semantics: decode_predictions, utils/yolov3_utils.py:12-59 and
// utils/yolov2_utils.py:12-58; the threshold is the torch.where of nms_v3,
// utils/yolo_utils.py:334.  One pass over the NCHW head: lanes walk the cell dimension so
// every channel-plane read is a coalesced (128-bit per lane when H*W % 4 == 0) transaction.
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

__device__ __forceinline__ void cls_push(ClsTrack& t, float v, int c)
{
    // NaN logits: sigmoid(NaN)=NaN; torch.argmax treats NaN as maximal.  Not supported here
    // (documented): a NaN never wins a > compare and is ignored.
    if (v > t.m1) { t.m2 = t.m1; t.m1 = v; t.i1 = c; }
    else if (v > t.m2) { t.m2 = v; }
}

__device__ __forceinline__ float cls_tol(float m1)
{
    // |d sigmoid| < 1 ulp(sigmoid)  <=>  d logit < 2^-23 (1 + e^m); 8x margin
    return 9.6e-7f * (1.0f + __expf(fmaxf(m1, 0.0f)));
}

// slow path: exact argmax over sigmoid for logits inside the band (first max wins)
__device__ __noinline__ int cls_resolve(const float* __restrict__ cls_base, int HW, int nc, float m1, float tol)
{
    int best = 0;
    float best_v = -1.0f;
    for (int c = 0; c < nc; ++c) {
        const float v = __ldg(cls_base + (size_t)c * HW);
        if (v >= m1 - tol) {
            const float s = ypp_sigmoid(v);
            if (s > best_v) { best_v = s; best = c; }
        }
    }
    return best;
}

template <int V> struct VecLoad;
template <> struct VecLoad<1> {
    static __device__ __forceinline__ void ld(const float* p, float (&o)[1]) { o[0] = __ldg(p); }
};
template <> struct VecLoad<4> {
    static __device__ __forceinline__ void ld(const float* p, float (&o)[4])
    {
        const float4 v = __ldg(reinterpret_cast<const float4*>(p));
        o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
    }
};

// class-plane sweep for V consecutive cells: software-pipelined, 4 planes (4 x 16 B per lane
// when V = 4) in flight while the previous 4 are folded into the running arg-max.
template <int V>
__device__ __forceinline__ void class_sweep(const float* __restrict__ cp, int HW, int nc, ClsTrack (&tr)[V])
{
#pragma unroll
    for (int v = 0; v < V; ++v) cls_init(tr[v]);
    float cur[4][V], nxt[4][V];
    int c = 0;
    if (nc >= 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) VecLoad<V>::ld(cp + (size_t)k * HW, cur[k]);
        for (; c + 8 <= nc; c += 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k) VecLoad<V>::ld(cp + (size_t)(c + 4 + k) * HW, nxt[k]);
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int v = 0; v < V; ++v) cls_push(tr[v], cur[k][v], c + k);
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int v = 0; v < V; ++v) cur[k][v] = nxt[k][v];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int v = 0; v < V; ++v) cls_push(tr[v], cur[k][v], c + k);
        c += 4;
    }
    for (; c < nc; ++c) {
        float l0[V];
        VecLoad<V>::ld(cp + (size_t)c * HW, l0);
#pragma unroll
        for (int v = 0; v < V; ++v) cls_push(tr[v], l0[v], c);
    }
}

// decode V consecutive cells of one anchor from the already loaded t0..t3 planes and the class
// tracks.  Fills box/score/cls for the lanes in `need`.
template <int V>
__device__ __forceinline__ void decode_cells(const YppDecodeParams& p, const YppScaleParams& s,
                                             const float* __restrict__ base /* channel 0, cell0 */,
                                             int a, int cell0, unsigned need,
                                             const float (&t0)[V], const float (&t1)[V], const float (&t2)[V],
                                             const float (&t3)[V], const ClsTrack (&tr)[V],
                                             const float (&score_obj)[V],
                                             float4 (&box)[V], float (&score)[V], int (&cls)[V])
{
    const int HW = s.HW;
    const int nc = p.nc;
    const float aw = s.aw[a], ah = s.ah[a];
    int y = cell0 / s.W;
    int x = cell0 - y * s.W;
#pragma unroll
    for (int v = 0; v < V; ++v) {
        if (need & (1u << v)) {
            // pxy = sigmoid(t) + grid, then *= stride (yolov3_utils.py:47,54)
            const float px = __fmul_rn(__fadd_rn(ypp_sigmoid(t0[v]), (float)x), s.sw);
            const float py = __fmul_rn(__fadd_rn(ypp_sigmoid(t1[v]), (float)y), s.sh);
            // pwh = exp(t) * scaled_anchor, then *= stride: two roundings (:48,54)
            const float pw = __fmul_rn(__fmul_rn(expf(t2[v]), aw), s.sw);
            const float ph = __fmul_rn(__fmul_rn(expf(t3[v]), ah), s.sh);
            box[v] = make_float4(px, py, pw, ph);
            int ci = 0;
            if (nc > 1) {
                ci = tr[v].i1;
                const float tol = cls_tol(tr[v].m1);
                if (!(tr[v].m1 - tr[v].m2 > tol))
                    ci = cls_resolve(base + 5 * (size_t)HW + v, HW, nc, tr[v].m1, tol);
            }
            cls[v] = ci;
            score[v] = score_obj[v];
            if (p.score_mode == YPP_SCORE_OBJ_TIMES_CLS)
                score[v] = __fmul_rn(score_obj[v], ypp_sigmoid(tr[v].m1));
        }
        if (++x == s.W) { x = 0; ++y; }
    }
}

// ---- block-wide compaction of up to V candidates per thread ------------------------------
struct CompactSmem {
    int warp_cnt[YPP_DECODE_THREADS / 32];
    u32 warp_max[YPP_DECODE_THREADS / 32], warp_nx[YPP_DECODE_THREADS / 32], warp_ny[YPP_DECODE_THREADS / 32];
    int block_base;
};

template <int V>
__device__ __forceinline__ void compact_store(const YppWs& ws, int N, int b, CompactSmem& sm,
                                              unsigned pass, const float4 (&box)[V], const float (&score)[V],
                                              const int (&cls)[V], int idx0)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cnt = __popc(pass);
    // warp inclusive scan of cnt
    int inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(YPP_FULL_MASK, inc, d);
        if (lane >= d) inc += o;
    }
    // per-image extrema for the coordinate trick / cross-class detector
    u32 emax = 0, enx = 0, eny = 0;
#pragma unroll
    for (int v = 0; v < V; ++v) {
        if (pass & (1u << v)) {
            const float4 q = ypp_xyxy(box[v]);
            emax = max(emax, max(max(ypp_enc(q.x), ypp_enc(q.y)), max(ypp_enc(q.z), ypp_enc(q.w))));
            enx = max(enx, ypp_enc(-q.x));
            eny = max(eny, ypp_enc(-q.y));
        }
    }
    emax = __reduce_max_sync(YPP_FULL_MASK, emax);
    enx = __reduce_max_sync(YPP_FULL_MASK, enx);
    eny = __reduce_max_sync(YPP_FULL_MASK, eny);
    if (lane == 31) { sm.warp_cnt[warp] = inc; sm.warp_max[warp] = emax; sm.warp_nx[warp] = enx; sm.warp_ny[warp] = eny; }
    __syncthreads();
    if (warp == 0) {
        constexpr int NW = YPP_DECODE_THREADS / 32;
        int c = lane < NW ? sm.warp_cnt[lane] : 0;
        u32 m = lane < NW ? sm.warp_max[lane] : 0, nx = lane < NW ? sm.warp_nx[lane] : 0, ny = lane < NW ? sm.warp_ny[lane] : 0;
        int ex = c;
#pragma unroll
        for (int d = 1; d < NW; d <<= 1) {
            const int o = __shfl_up_sync(YPP_FULL_MASK, ex, d);
            if (lane >= d) ex += o;
        }
        m = __reduce_max_sync(YPP_FULL_MASK, m);
        nx = __reduce_max_sync(YPP_FULL_MASK, nx);
        ny = __reduce_max_sync(YPP_FULL_MASK, ny);
        if (lane < NW) sm.warp_cnt[lane] = ex - c;   // exclusive warp offsets
        const int total = __shfl_sync(YPP_FULL_MASK, ex, NW - 1);
        if (lane == 0) {
            int base = 0;
            if (total > 0) {
                base = atomicAdd(ws.cand_count + b, total);
                atomicMax(ws.maxc + b, m);
                atomicMax(ws.nminx1 + b, nx);
                atomicMax(ws.nminy1 + b, ny);
            }
            sm.block_base = base;
        }
    }
    __syncthreads();
    if (cnt) {
        int pos = sm.block_base + sm.warp_cnt[warp] + inc - cnt;
        u64* keys = ws.keys + (size_t)b * N;
        float4* bx = ws.box + (size_t)b * N;
#pragma unroll
        for (int v = 0; v < V; ++v) {
            if (pass & (1u << v)) {
                keys[pos++] = ypp_make_key(score[v], idx0 + v, cls[v]);
                bx[idx0 + v] = box[v];
            }
        }
    }
}

template <int V>
__device__ __forceinline__ void decode_compact_tile(const YppDecodeParams& p, const YppScaleParams& s,
                                                    int tile, int b, CompactSmem& sm)
{
    const int upp = s.HW / V;                       // units per anchor plane
    const int u = tile * YPP_DECODE_THREADS + threadIdx.x;
    const bool valid = u < s.units;
    const int a = valid ? u / upp : 0;
    const int cell0 = valid ? (u - a * upp) * V : 0;
    const float* base = s.head + ((size_t)(b * s.A + a) * p.C) * s.HW + cell0;

    // objectness and the four box planes are fetched together (one round trip); the nc class
    // planes only when some lane of the warp holds a candidate
    float obj[V], sobj[V], t0[V], t1[V], t2[V], t3[V];
    unsigned pass = 0;
    if (valid) {
        const size_t HW = s.HW;
        VecLoad<V>::ld(base + 4 * HW, obj);
        VecLoad<V>::ld(base, t0);
        VecLoad<V>::ld(base + HW, t1);
        VecLoad<V>::ld(base + 2 * HW, t2);
        VecLoad<V>::ld(base + 3 * HW, t3);
#pragma unroll
        for (int v = 0; v < V; ++v) {
            sobj[v] = ypp_sigmoid(obj[v]);                       // yolov3_utils.py:50
            if (sobj[v] > p.conf) pass |= 1u << v;               // yolo_utils.py:334 (strict)
        }
    }
    float4 box[V];
    float score[V];
    int cls[V];
    if (__any_sync(YPP_FULL_MASK, pass != 0)) {
        if (pass) {
            ClsTrack tr[V];
            if (p.nc > 1 || p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) class_sweep<V>(base + 5 * (size_t)s.HW, s.HW, p.nc, tr);
            else {
#pragma unroll
                for (int v = 0; v < V; ++v) cls_init(tr[v]);
            }
            decode_cells<V>(p, s, base, a, cell0, pass, t0, t1, t2, t3, tr, sobj, box, score, cls);
            if (p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) {
#pragma unroll
                for (int v = 0; v < V; ++v)
                    if ((pass & (1u << v)) && !(score[v] > p.conf)) pass &= ~(1u << v);
            }
        }
    }
    compact_store<V>(p.ws, p.N, b, sm, pass, box, score, cls, s.box_base + a * s.HW + cell0);
}

__global__ void __launch_bounds__(YPP_DECODE_THREADS, YPP_DECODE_MINB)
ypp_decode_compact_kernel(const __grid_constant__ YppDecodeParams p)
{
    __shared__ CompactSmem sm;
    const int b = blockIdx.y;
    int si = 0;
#pragma unroll
    for (int k = 1; k < YPP_MAX_SCALES; ++k)
        if (k < p.n_scales && (int)blockIdx.x >= p.tile_base[k]) si = k;
    const YppScaleParams& s = p.sc[si];
    const int tile = blockIdx.x - p.tile_base[si];
    if (s.vec == 4) decode_compact_tile<4>(p, s, tile, b, sm);
    else            decode_compact_tile<1>(p, s, tile, b, sm);
}

// ---- dense decode: every box, rows [B, N, 6] (= decode_predictions + cat) ---------------
__global__ void __launch_bounds__(YPP_DECODE_THREADS)
ypp_decode_dense_kernel(const __grid_constant__ YppDecodeParams p, float* __restrict__ out)
{
    const int b = blockIdx.y;
    int si = 0;
#pragma unroll
    for (int k = 1; k < YPP_MAX_SCALES; ++k)
        if (k < p.n_scales && (int)blockIdx.x >= p.tile_base[k]) si = k;
    const YppScaleParams& s = p.sc[si];
    const int tile = blockIdx.x - p.tile_base[si];
    const int u = tile * YPP_DECODE_THREADS + threadIdx.x;   // dense path: units are single boxes
    if (u >= s.A * s.HW) return;
    const int a = u / s.HW;
    const int cell = u - a * s.HW;
    const float* base = s.head + ((size_t)(b * s.A + a) * p.C) * s.HW + cell;
    float obj[1], sobj[1], score[1], t0[1], t1[1], t2[1], t3[1];
    float4 box[1];
    int cls[1];
    const size_t HW = s.HW;
    VecLoad<1>::ld(base + 4 * HW, obj);
    VecLoad<1>::ld(base, t0);
    VecLoad<1>::ld(base + HW, t1);
    VecLoad<1>::ld(base + 2 * HW, t2);
    VecLoad<1>::ld(base + 3 * HW, t3);
    ClsTrack tr[1];
    if (p.nc > 1 || p.score_mode == YPP_SCORE_OBJ_TIMES_CLS) class_sweep<1>(base + 5 * HW, s.HW, p.nc, tr);
    else cls_init(tr[0]);
    sobj[0] = ypp_sigmoid(obj[0]);
    decode_cells<1>(p, s, base, a, cell, 1u, t0, t1, t2, t3, tr, sobj, box, score, cls);
    float2* o = reinterpret_cast<float2*>(out + ((size_t)b * p.N + s.box_base + u) * 6);
    o[0] = make_float2(box[0].x, box[0].y);
    o[1] = make_float2(box[0].z, box[0].w);
    o[2] = make_float2(score[0], (float)cls[0]);
}

// ---- compaction from decoded rows (nms_v3 entry, utils/yolo_utils.py:321-337) -----------
__global__ void __launch_bounds__(YPP_DECODE_THREADS)
ypp_rows_compact_kernel(const float* __restrict__ rows, int N, float conf, YppWs ws)
{
    __shared__ CompactSmem sm;
    const int b = blockIdx.y;
    const int n = blockIdx.x * YPP_DECODE_THREADS + threadIdx.x;
    unsigned pass = 0;
    float4 box[1];
    float score[1];
    int cls[1];
    if (n < N) {
        const float2* r = reinterpret_cast<const float2*>(rows + ((size_t)b * N + n) * 6);
        const float2 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
        box[0] = make_float4(r0.x, r0.y, r1.x, r1.y);
        score[0] = r2.x;
        int c = (int)r2.y;
        if (!(r2.y >= 0.0f && r2.y <= (float)YPP_KEY_CLS_MASK && (float)c == r2.y)) {
            atomicOr(ws.q + 2, 1);          // class id not representable: flagged, clamped
            c = min(max(c, 0), (int)YPP_KEY_CLS_MASK);
        }
        cls[0] = c;
        if (score[0] > conf) pass = 1u;
    }
    compact_store<1>(ws, N, b, sm, pass, box, score, cls, n);
}

}  // namespace

// ---------------------------------------------------------------- host launchers

int ypp_launch_decode_compact(const YppDecodeParams& p, int B, cudaStream_t st)
{
    dim3 grid(p.tile_base[p.n_scales], B);
    ypp_decode_compact_kernel<<<grid, YPP_DECODE_THREADS, 0, st>>>(p);
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
    dim3 grid((N + YPP_DECODE_THREADS - 1) / YPP_DECODE_THREADS, B);
    ypp_rows_compact_kernel<<<grid, YPP_DECODE_THREADS, 0, st>>>(rows, N, conf, ws);
    ypp_count_launch(1);
    return ypp_check_launch("rows_compact");
}
