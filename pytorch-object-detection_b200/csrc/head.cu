// F1: the detection head's 1x1 convolution fused with D1.
//
// Reference: every scale ends in  nn.Conv2d(Cin, A*(5+nc), 1, 1, bias=False)
// (models/detector/yolov3.py:53-69, yolov4_tiny.py, yolov2.py) whose output [B, A(5+nc), H, W] is what
// decode_predictions reads (utils/yolov3_utils.py:12-59).  A 1x1 convolution is the GEMM
//     head[b, hw, c] = sum_k feature[b, hw, k] * weight[c, k]
// so here the head tensor never exists in HBM: the accumulator tile of 128 cells x A(5+nc) channels is
// decoded, thresholded and compacted straight out of tensor memory into the same slot records D1
// writes, and S1 -> N1 -> G run unchanged.
//
// sm_100a design: one launch over every scale, one CTA per (image, 128 cells), widest scale first.
// A = feature tile [128 cells x 32 channels] and B = weight tile [N x 32 channels] (N = A(5+nc) rounded
// up to 16), both K-major, arrive by TMA (cp.async.bulk.tensor, 128-byte swizzle) into a small ring of
// stages; one thread issues tcgen05.mma.kind::tf32 (M 128, N, K 8) four times per stage into a TMEM
// accumulator and commits the stage back to the producer; after the last commit eight warps read the
// TMEM lanes (tcgen05.ld 32x32b, two warps per lane quarter), park the tile channel-major in the (now
// idle) ring and run D1's own decode_one / slot_store on it, so everything downstream of the
// accumulator is bit-identical to D1 fed the same head values.  Features are channels-last
// ([B, H, W, Cin], torch.channels_last): K-major rows of any H*W, 16-byte aligned for Cin % 4 == 0.
// Arithmetic: TF32 products, fp32 accumulation - what cuDNN does for this convolution under
// torch.backends.cudnn.allow_tf32 (the default); parity is stated against that, see tests.
#include "decode_common.cuh"
#include <cuda.h>
#include <cstdlib>
#include <cstring>

#define YPP_HEAD_STAGES 4            /* most stages of the TMA ring */
#define YPP_HEAD_DEFAULT_STAGES 2    /* 4 CTAs per SM with two stages each beat 2 CTAs with four (72 vs 77 us on C2) */
#define YPP_HEAD_BM 128            /* cells per CTA = UMMA M = TMEM lanes */
#define YPP_HEAD_BK 32             /* fp32 channels per stage row: 128 bytes = one swizzle span */
#define YPP_HEAD_EPI_WARPS 8       /* epilogue warps: warp w reads TMEM lanes 32 (w % 4) .., two warps per lane quarter */
#define YPP_HEAD_PRODUCER YPP_HEAD_EPI_WARPS          /* the TMA producer warp */
#define YPP_HEAD_MMA (YPP_HEAD_EPI_WARPS + 1)        /* the MMA issuer, owner of the TMEM allocation */
#define YPP_HEAD_THREADS (32 * (YPP_HEAD_EPI_WARPS + 2))
#define YPP_HEAD_A_BYTES (YPP_HEAD_BM * YPP_HEAD_BK * 4)

struct YppHeadParams {
    YppDecodeParams d;
    int order[YPP_MAX_SCALES];          // scales in the order their tiles are handed out (most input channels first)
    int tile_start[YPP_MAX_SCALES + 1]; // blockIdx.x ranges of order[0], order[1], ...
    int cin[YPP_MAX_SCALES];            // input channels (K) per scale
    float* head_out[YPP_MAX_SCALES];    // optional [B, ac, H, W]: the head tensor itself
    int ac;                 // A * (5 + nc): real output channels
    int n_pad;              // UMMA N: ac rounded up to a multiple of 16
    int tmem_cols;          // power of two >= max(32, n_pad)
    int stages;             // depth of the TMA ring (<= YPP_HEAD_STAGES)
    u32 idesc;              // tcgen05 instruction descriptor (tf32 x tf32 -> f32, K-major A and B, M 128, N n_pad)
};

struct YppHeadMaps {
    CUtensorMap a[YPP_MAX_SCALES];      // features [B][HW][Cin], box 1 x 128 x 32
    CUtensorMap b[YPP_MAX_SCALES];      // weight [ac][Cin], box n_pad x 32
};

namespace {

__device__ __forceinline__ void tma_load_3d(u32 dst, const CUtensorMap* tm, int c0, int c1, int c2, u64* bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"((u64)tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(u32 dst, const CUtensorMap* tm, int c0, int c1, u64* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"((u64)tm), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// shared-memory matrix descriptor, K-major, 128-byte swizzle: 8-row groups 1024 bytes apart
__device__ __forceinline__ u64 umma_desc(u32 saddr)
{
    return (u64)((saddr & 0x3ffffu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_tf32(u32 tmem_d, u64 da, u64 db, u32 idesc, u32 accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(u64* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(u32 taddr, u32 (&v)[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(YPP_HEAD_THREADS, 4)
ypp_head_decode_kernel(const __grid_constant__ YppHeadMaps tm, const __grid_constant__ YppHeadParams hp)
{
    extern __shared__ unsigned char ring_raw[];
    __shared__ __align__(8) u64 full[YPP_HEAD_STAGES], empty[YPP_HEAD_STAGES], accum;
    __shared__ u32 tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const YppDecodeParams& p = hp.d;
    int oi = 0;
#pragma unroll
    for (int k = 1; k < YPP_MAX_SCALES; ++k)
        if (k < p.n_scales && (int)blockIdx.x >= hp.tile_start[k]) oi = k;
    const int si = hp.order[oi];
    const YppScaleParams& s = p.sc[si];
    const CUtensorMap* tm_a = &tm.a[si];
    const CUtensorMap* tm_b = &tm.b[si];
    const int b = blockIdx.y, cell0 = ((int)blockIdx.x - hp.tile_start[oi]) * YPP_HEAD_BM;
    const bool has_cells = cell0 < s.HW;                     // 2 * tpp tiles per plane cover every D1 slot, some past the end
    const int nk = (hp.cin[si] + YPP_HEAD_BK - 1) / YPP_HEAD_BK;
    const int nst = hp.stages;
    const u32 stage_bytes = YPP_HEAD_A_BYTES + (u32)hp.n_pad * (YPP_HEAD_BK * 4);
    unsigned char* ring = ring_raw + ((1024u - (smem_u32(ring_raw) & 1023u)) & 1023u);      // swizzle atoms are 1024-byte aligned
    const u32 ring_addr = smem_u32(ring);
    int kc0 = 0;
    if (warp == YPP_HEAD_PRODUCER && lane == 0) {
#pragma unroll
        for (int i = 0; i < YPP_HEAD_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(&accum, 1);
        // the first loads leave before the TMEM allocation and the block barrier below: nobody else touches the
        // mbarriers until then, and the features are read-only (no need for the grid dependency either)
        if (has_cells)
            for (; kc0 < nk && kc0 < nst; ++kc0) {
                mbar_expect_tx(&full[kc0], stage_bytes);
                const u32 dst = ring_addr + (u32)kc0 * stage_bytes;
                tma_load_3d(dst, tm_a, kc0 * YPP_HEAD_BK, cell0, b, &full[kc0]);
                tma_load_2d(dst + YPP_HEAD_A_BYTES, tm_b, kc0 * YPP_HEAD_BK, 0, &full[kc0]);
            }
    }
    if (warp == YPP_HEAD_MMA) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(hp.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const u32 tmem = tmem_slot;
    ypp_grid_wait();                                          // the previous call's G may still be reading the workspace

    if (has_cells && warp == YPP_HEAD_PRODUCER && lane == 0) {
        // ---------------- TMA producer (the first ring-full is already in flight)
        for (int kc = kc0; kc < nk; ++kc) {
            const int st = kc % nst, use = kc / nst;
            if (use > 0) mbar_wait(&empty[st], (u32)(use - 1) & 1u);
            mbar_expect_tx(&full[st], stage_bytes);
            const u32 dst = ring_addr + (u32)st * stage_bytes;
            tma_load_3d(dst, tm_a, kc * YPP_HEAD_BK, cell0, b, &full[st]);              // cells past H*W arrive as zeros
            tma_load_2d(dst + YPP_HEAD_A_BYTES, tm_b, kc * YPP_HEAD_BK, 0, &full[st]);   // rows past ac arrive as zeros
        }
    } else if (has_cells && warp == YPP_HEAD_MMA && lane == 0) {
        // ---------------- MMA issuer
        for (int kc = 0; kc < nk; ++kc) {
            const int st = kc % nst, use = kc / nst;
            mbar_wait(&full[st], (u32)use & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const u32 a_addr = ring_addr + (u32)st * stage_bytes, b_addr = a_addr + YPP_HEAD_A_BYTES;
#pragma unroll
            for (int k = 0; k < YPP_HEAD_BK / 8; ++k)          // UMMA K = 8 tf32 = 32 bytes along the swizzled row
                umma_tf32(tmem, umma_desc(a_addr + 32u * k), umma_desc(b_addr + 32u * k), hp.idesc, (u32)(kc | k));
            umma_commit(&empty[st]);                          // the stage is free once these four have read it
        }
        umma_commit(&accum);
    } else if (warp < YPP_HEAD_EPI_WARPS) {
        // ---------------- epilogue.  Warps q and q + 4 share the TMEM lane quarter q (cells 32q .. 32q+31 of the
        // tile) and split its column chunks; the tile is parked channel-major in the ring (idle by then), after
        // which any warp can decode any 32-cell slot: the 4 A slots of the tile are dealt round-robin to the warps.
        float* stg = reinterpret_cast<float*>(ring);          // [ac][128]
        const int q = warp & 3, g = warp >> 2;
        if (has_cells) {
            const int ct = q * 32 + lane, cell = cell0 + ct;
            mbar_wait(&accum, 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int j = g; j < hp.n_pad / 16; j += YPP_HEAD_EPI_WARPS / 4) {
                u32 v[16];
                tmem_ld16(tmem + ((u32)(q * 32) << 16) + (u32)(j * 16), v);
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (j * 16 + i < hp.ac) stg[(j * 16 + i) * YPP_HEAD_BM + ct] = __uint_as_float(v[i]);
            }
            asm volatile("bar.sync 1, %0;" ::"n"(32 * YPP_HEAD_EPI_WARPS) : "memory");     // the epilogue warps only
            float* ho = hp.head_out[si];
            if (ho != nullptr && cell < s.HW)
                for (int c = g; c < hp.ac; c += YPP_HEAD_EPI_WARPS / 4)
                    ho[((size_t)b * hp.ac + c) * s.HW + cell] = stg[c * YPP_HEAD_BM + ct];
        }
        const int nslots = p.tile_base[p.n_scales] * (YPP_TILE_BOXES / 32);
        for (int sl = warp; sl < 4 * s.A; sl += YPP_HEAD_EPI_WARPS) {
            const int a = sl >> 2, qq = sl & 3;
            const int ct = qq * 32 + lane, cell = cell0 + ct;
            Decoded d;
            d.box = make_float4(0.f, 0.f, 0.f, 0.f); d.score = 0.f; d.cls = 0;
            const bool pass = decode_one<1>(p, s, a, cell, cell < s.HW, stg + a * p.C * YPP_HEAD_BM + ct, YPP_HEAD_BM, 0, 0, d);
            const int tile = p.tile_base[si] + a * s.tpp + cell0 / YPP_TILE_BOXES;
            const int wslot = ((cell0 & (YPP_TILE_BOXES - 1)) >> 5) + qq;
            const int first = s.box_base + a * s.HW + cell0;
            slot_store(p.ws, p.N, b, tile * (YPP_TILE_BOXES / 32) + wslot, nslots, first + 32 * qq, pass,
                       d.box, d.score, d.cls, first + ct);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == YPP_HEAD_MMA)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(hp.tmem_cols) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)ptr;
    }
    return fn;
}

}  // namespace

int ypp_launch_head_decode(const YppDecodeParams& p, int B, const float* const* feature, const float* const* weight,
                           const int* cin, float* const* head_out, cudaStream_t st)
{
    const int A = p.sc[0].A;
    const int ac = A * p.C;
    const int n_pad = (ac + 15) & ~15;
    if (n_pad > 256) { ypp_set_error("head_decode: A*(5+nc) = %d exceeds one UMMA tile (256)", ac); return YPP_E_UNSUPPORTED; }
    EncodeTiledFn enc = encode_tiled();
    if (!enc) { ypp_set_error("head_decode: cuTensorMapEncodeTiled is not available from this driver"); return YPP_E_CUDA; }
    YppHeadMaps tm;
    memset(&tm, 0, sizeof(tm));
    YppHeadParams hp;
    memset(&hp, 0, sizeof(hp));
    hp.d = p;
    for (int i = 0; i < p.n_scales; ++i) {
        const YppScaleParams& s = p.sc[i];
        if (s.A != A) { ypp_set_error("head_decode: every scale must have the same number of anchors"); return YPP_E_UNSUPPORTED; }
        if (cin[i] < 4 || (cin[i] & 3) != 0) { ypp_set_error("head_decode: Cin = %d must be a positive multiple of 4", cin[i]); return YPP_E_UNSUPPORTED; }
        if (((uintptr_t)feature[i] & 15u) != 0 || ((uintptr_t)weight[i] & 15u) != 0) { ypp_set_error("head_decode: feature / weight must be 16-byte aligned"); return YPP_E_BADARG; }
        {   // features [B][HW][Cin], box 1 x 128 x 32
            const cuuint64_t dims[3] = {(cuuint64_t)cin[i], (cuuint64_t)s.HW, (cuuint64_t)B};
            const cuuint64_t strides[2] = {(cuuint64_t)cin[i] * 4, (cuuint64_t)s.HW * cin[i] * 4};
            const cuuint32_t box[3] = {YPP_HEAD_BK, YPP_HEAD_BM, 1}, es[3] = {1, 1, 1};
            const CUresult r = enc(&tm.a[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)feature[i], dims, strides, box, es,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { ypp_set_error("head_decode: tensor map of the features failed (%d)", (int)r); return YPP_E_CUDA; }
        }
        {   // weight [ac][Cin], box n_pad x 32
            const cuuint64_t dims[2] = {(cuuint64_t)cin[i], (cuuint64_t)ac};
            const cuuint64_t strides[1] = {(cuuint64_t)cin[i] * 4};
            const cuuint32_t box[2] = {YPP_HEAD_BK, (cuuint32_t)n_pad}, es[2] = {1, 1};
            const CUresult r = enc(&tm.b[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)weight[i], dims, strides, box, es,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { ypp_set_error("head_decode: tensor map of the weight failed (%d)", (int)r); return YPP_E_CUDA; }
        }
        hp.cin[i] = cin[i];
        hp.head_out[i] = head_out ? head_out[i] : nullptr;
        hp.order[i] = i;
    }
    // the longest K loops first: their CTAs run longest, the short ones fill the tail
    for (int i = 1; i < p.n_scales; ++i)
        for (int j = i; j > 0 && hp.cin[hp.order[j]] > hp.cin[hp.order[j - 1]]; --j) { const int t = hp.order[j]; hp.order[j] = hp.order[j - 1]; hp.order[j - 1] = t; }
    int tiles = 0;
    for (int i = 0; i < p.n_scales; ++i) { hp.tile_start[i] = tiles; tiles += 2 * p.sc[hp.order[i]].tpp; }
    for (int i = p.n_scales; i <= YPP_MAX_SCALES; ++i) hp.tile_start[i] = tiles;
    hp.ac = ac; hp.n_pad = n_pad;
    hp.tmem_cols = 32;
    while (hp.tmem_cols < n_pad) hp.tmem_cols <<= 1;
    static int stages_env = -1;
    if (stages_env < 0) { const char* e = getenv("YPP_HEAD_STAGES"); stages_env = e ? atoi(e) : 0; }
    hp.stages = stages_env >= 1 && stages_env <= YPP_HEAD_STAGES ? stages_env : YPP_HEAD_DEFAULT_STAGES;
    // c_format f32 (1 << 4), a/b format tf32 (2 << 7, 2 << 10), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
    hp.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((u32)(n_pad >> 3) << 17) | ((u32)(YPP_HEAD_BM >> 4) << 24);
    int smem = hp.stages * (YPP_HEAD_A_BYTES + n_pad * YPP_HEAD_BK * 4);
    if (smem < ac * YPP_HEAD_BM * 4) smem = ac * YPP_HEAD_BM * 4;        // the epilogue parks the whole tile in the ring
    smem += 1024;
    static int configured_d[YPP_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= YPP_MAX_DEVICES) { ypp_set_error("head_decode: device ordinal %d out of range", dev); return YPP_E_UNSUPPORTED; }
    if (configured_d[dev] < smem) {
        cudaError_t e = cudaFuncSetAttribute(ypp_head_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { ypp_set_error("head_decode: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return YPP_E_CUDA; }
        configured_d[dev] = smem;
    }
    dim3 grid(tiles, B);
    ypp_head_decode_kernel<<<grid, YPP_HEAD_THREADS, smem, st>>>(tm, hp);
    ypp_count_launch(1);
    return ypp_check_launch("head_decode");
}
