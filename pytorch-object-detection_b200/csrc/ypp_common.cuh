// Shared device/host helpers of libypp (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/ypp.h"

typedef unsigned long long u64;
typedef unsigned int u32;

#define YPP_FULL_MASK 0xffffffffu
#define YPP_KEY_CLS_BITS 12
#define YPP_KEY_IDX_BITS 20
#define YPP_KEY_CLS_MASK ((1u << YPP_KEY_CLS_BITS) - 1u)
#define YPP_KEY_IDX_MASK ((1u << YPP_KEY_IDX_BITS) - 1u)

// segments up to this long are owned by one warp; longer ones go to the persistent CTA kernel
#define YPP_WARP_SEG 512
// boxes the big kernel keeps staged in shared memory (x 16 B)
#define YPP_BIG_CACHE 11264
#define YPP_NMS_BIG_THREADS 1024
#define YPP_ORG_THREADS 1024
#define YPP_DECODE_THREADS 256     // == YPP_TILE_BOXES

// ---------------------------------------------------------------- workspace layout
// All per-image arrays are [B][N].  Nothing needs a memset between calls: D1 owns the slot
// records and S1 (re)initialises every per-image counter and list.
#define YPP_MAX_DEVICES 64         // launcher-side per-device caches
#define YPP_TILE_BOXES 256        // boxes per D1 tile (one per thread)
struct YppWs {
    int* cand_count;      // [B] written by S1
    int* keep_count;      // [B] zeroed by S1, accumulated by N1
    int* imgflag;         // [B] bit0: some classes share a segment (coordinate-trick cross-class edge), bit1: coordinate trick
    float* img_step;      // [B] max_coordinate + 1 of the image's candidates (the trick's class offset step)
    int* bigcnt;          // [B] segments longer than YPP_WARP_SEG in the image
    int* ntiles;          // [1] D1 slots (32 boxes, one warp) per image
    int* tile_first;      // [T] first box index of each slot (the same for every image)
    int* tile_cnt;        // [B][T] candidates each slot produced (compacted at keys[b][tile_first ..])
    u32* tile_ext;        // [B][T][4] order-encoded max(x1,y1,x2,y2), max(-x1), max(-y1) over the slot's candidates
    int* bigq;            // [B][bigq_per_image] their classes
    int* seg_start;       // [B][nc+1] segment starts in class order
    u64* keys;            // [B][N] (~enc(score))<<32 | idx<<12 | cls, compacted per tile by D1 (box-index order)
    u64* skeys;           // [B][N] the same keys in stable descending-score order (S1)
    float4* box;          // [B][N] cx,cy,w,h indexed by box index n
    float4* sbox;         // [B][N] x1,y1,x2,y2 (+ class offset) in class-segment order
    u64* gpk;             // [B][N] class-order position -> global score rank << 32 | box index << 12 | class (the key's low word)
    unsigned char* keepflag;  // [B][N] by global score rank
    int tiles_max;
    int bigq_per_image;
    size_t total_bytes;
};

static inline size_t ypp_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static inline YppWs ypp_make_ws(void* base, int B, int N, int nc)
{
    YppWs w;
    char* p = (char*)base;
    size_t off = 0;
    auto take = [&](size_t bytes) { char* r = p + off; off += ypp_align_up(bytes, 256); return r; };
    w.cand_count = (int*)take(sizeof(int) * B);
    w.keep_count = (int*)take(sizeof(int) * B);
    w.imgflag = (int*)take(sizeof(int) * B);
    w.img_step = (float*)take(sizeof(float) * B);
    w.bigcnt = (int*)take(sizeof(int) * B);
    w.ntiles = (int*)take(sizeof(int));
    // a tile never spans two anchor planes, so a scale adds at most A partial tiles
    w.tiles_max = (N / YPP_TILE_BOXES + YPP_MAX_SCALES * YPP_MAX_ANCHORS + 8) * (YPP_TILE_BOXES / 32);   // 32-box slots
    w.tile_first = (int*)take(sizeof(int) * w.tiles_max);
    w.tile_cnt = (int*)take(sizeof(int) * (size_t)B * w.tiles_max);
    w.tile_ext = (u32*)take(sizeof(u32) * 4 * (size_t)B * w.tiles_max);
    int per = N / YPP_WARP_SEG + 1;
    if (per > nc) per = nc;
    w.bigq_per_image = per;
    w.bigq = (int*)take(sizeof(int) * (size_t)B * per);
    w.seg_start = (int*)take(sizeof(int) * (size_t)B * (nc + 1));
    w.keys = (u64*)take(sizeof(u64) * (size_t)B * N);
    w.skeys = (u64*)take(sizeof(u64) * (size_t)B * N);
    w.box = (float4*)take(sizeof(float4) * (size_t)B * N);
    w.sbox = (float4*)take(sizeof(float4) * (size_t)B * N);
    w.gpk = (u64*)take(sizeof(u64) * (size_t)B * N);
    w.keepflag = (unsigned char*)take((size_t)B * N);
    w.total_bytes = off;
    return w;
}

// ---------------------------------------------------------------- error plumbing (api.cu)
void ypp_set_error(const char* fmt, ...);
int ypp_check_launch(const char* what);
void ypp_count_launch(int n);

#ifdef __CUDACC__
// ---------------------------------------------------------------- device helpers

// monotone float -> uint map (works for negatives): a < b  <=>  enc(a) < enc(b)
__device__ __forceinline__ u32 ypp_enc(float f)
{
    u32 b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ypp_dec(u32 e)
{
    u32 b = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
    return __uint_as_float(b);
}

// 1/(1+exp(-x)) with IEEE add/div and the accurate expf: the formula ATen's CUDA sigmoid uses
__device__ __forceinline__ float ypp_sigmoid(float x)
{
    return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)));
}

__device__ __forceinline__ u64 ypp_make_key(float score, int idx, int cls)
{
    return ((u64)(~ypp_enc(score)) << 32) | ((u64)(u32)idx << YPP_KEY_CLS_BITS) | (u64)(u32)cls;
}
__device__ __forceinline__ float ypp_key_score(u64 k) { return ypp_dec(~(u32)(k >> 32)); }
__device__ __forceinline__ int ypp_key_idx(u64 k) { return (int)(((u32)k >> YPP_KEY_CLS_BITS) & YPP_KEY_IDX_MASK); }
__device__ __forceinline__ int ypp_key_cls(u64 k) { return (int)((u32)k & YPP_KEY_CLS_MASK); }
// S1's per-position word: rank in the image's score order on top of the key's low word (index, class)
__device__ __forceinline__ u64 ypp_make_gpk(int rank, u64 key) { return ((u64)(u32)rank << 32) | (u64)(u32)key; }
__device__ __forceinline__ int ypp_gpk_rank(u64 g) { return (int)(g >> 32); }

// cxcywh -> xyxy exactly as nms_v3 does (utils/yolo_utils.py:339-343): c -/+ (wh / 2)
__device__ __forceinline__ float4 ypp_xyxy(float4 c)
{
    const float hw = __fmul_rn(c.z, 0.5f), hh = __fmul_rn(c.w, 0.5f);   // x/2 == x*0.5 exactly
    return make_float4(__fsub_rn(c.x, hw), __fsub_rn(c.y, hh), __fadd_rn(c.x, hw), __fadd_rn(c.y, hh));
}

__device__ __forceinline__ float ypp_area(float4 b)
{
    return __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
}

// torchvision::nms pair test. a = row (earlier in score order), b = column.
// union_fma: union = fma(w_b, h_b, S_a) - inter ; else (S_a + S_b) - inter.
template <bool kEarlyOut>
__device__ __forceinline__ bool ypp_suppresses(const float4 a, const float sa, const float4 b,
                                               const float thr, const int union_fma)
{
    const float l = fmaxf(a.x, b.x), t = fmaxf(a.y, b.y);
    const float r = fminf(a.z, b.z), bt = fminf(a.w, b.w);
    const float w = fmaxf(__fsub_rn(r, l), 0.0f), h = fmaxf(__fsub_rn(bt, t), 0.0f);
    const float inter = __fmul_rn(w, h);
    if (kEarlyOut && !(inter > 0.0f)) return false;   // iou is 0 or NaN: never > thr when thr >= 0
    const float wb = __fsub_rn(b.z, b.x), hb = __fsub_rn(b.w, b.y);
    float uni;
    if (union_fma) uni = __fsub_rn(__fmaf_rn(wb, hb, sa), inter);
    else           uni = __fsub_rn(__fadd_rn(sa, __fmul_rn(wb, hb)), inter);
    return __fdiv_rn(inter, uni) > thr;
}
#endif  // __CUDACC__
