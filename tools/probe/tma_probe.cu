// Micro-benchmark (tuning aid for D1; not part of the library): what HBM read bandwidth do
//   mode 0: a persistent, mbarrier-pipelined cp.async.bulk row-copy pattern over the real head layout
//           (slabs of ROWS planes of PLANEB bytes; a tile = ROWS rows x ROWB bytes, one per plane)
//   mode 1: a plain grid-stride LDG.128 read
// reach when every launch reads LAUNCH_MB and launches run back to back over a 272 MB ring?
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;
__device__ __forceinline__ u32 s32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_tx(u64* b, u32 n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void g2s(void* d, const void* s, u32 n, u64* b) { asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(d)), "l"(s), "r"(n), "r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(u64* b, u32 ph) { asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(s32(b)), "r"(ph) : "memory"); }

template <int STAGES>
__global__ void probe(const char* src, int nslabs, int tps, int rows, int rowb, int planeb, int issuers, float* out)
{
    extern __shared__ __align__(128) char smem[];
    __shared__ __align__(8) u64 bar[STAGES];
    const int tid = threadIdx.x;
    const int stage_bytes = rows * rowb;
    const long long ntiles = (long long)nslabs * tps;
    if (tid == 0) for (int i = 0; i < STAGES; ++i) mbar_init(&bar[i], 1);
    __syncthreads();
    auto issue = [&](long long t, int st) {
        const int slab = (int)(t / tps), j = (int)(t - (long long)slab * tps);
        const int nb = min(rowb, planeb - j * rowb);
        if (tid == 0) mbar_tx(&bar[st], (u32)nb * rows);
        __syncthreads();
        if (tid < issuers) {
            const char* base = src + (size_t)slab * rows * planeb + (size_t)j * rowb;
            for (int r = tid; r < rows; r += issuers) g2s(smem + (size_t)st * stage_bytes + (size_t)r * rowb, base + (size_t)r * planeb, (u32)nb, &bar[st]);
        }
    };
    long long t = blockIdx.x;
    for (int i = 0; i < STAGES; ++i) if (t + (long long)i * gridDim.x < ntiles) issue(t + (long long)i * gridDim.x, i);
    float acc = 0.f;
    int k = 0;
    for (; t < ntiles; t += gridDim.x, ++k) {
        const int st = k % STAGES;
        mbar_wait(&bar[st], (k / STAGES) & 1);
        acc += reinterpret_cast<const float*>(smem + (size_t)st * stage_bytes)[tid];
        __syncthreads();
        if (t + (long long)STAGES * gridDim.x < ntiles) issue(t + (long long)STAGES * gridDim.x, st);
    }
    if (acc == 123.456f) out[0] = acc;
}

__global__ void plain(const uint4* __restrict__ src, long long n16, float* out)
{
    u32 acc = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        const uint4 a = __ldg(src + i), b = __ldg(src + i + stride), c = __ldg(src + i + 2 * stride), d = __ldg(src + i + 3 * stride);
        acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
    }
    for (; i < n16; i += stride) { const uint4 a = __ldg(src + i); acc ^= a.x ^ a.y ^ a.z ^ a.w; }
    if (acc == 0x12345u) out[0] = 1.f;
}

int main(int argc, char** argv)
{
    const int mode = argc > 1 ? atoi(argv[1]) : 0;
    const int launch_mb = argc > 2 ? atoi(argv[2]) : 68;
    const int rows = argc > 3 ? atoi(argv[3]) : 75, rowb = argc > 4 ? atoi(argv[4]) : 1024, stages = argc > 5 ? atoi(argv[5]) : 2;
    const int threads = argc > 6 ? atoi(argv[6]) : 256, cps = argc > 7 ? atoi(argv[7]) : 1, issuers = argc > 8 ? atoi(argv[8]) : 32;
    const int planeb = argc > 9 ? atoi(argv[9]) : 10816;   // 52*52*4
    const size_t total = 272ull << 20;    // > L2
    char* d; float* o;
    cudaMalloc(&d, total + (1 << 20)); cudaMalloc(&o, 4);
    cudaMemset(d, 1, total);
    const size_t slab = (size_t)planeb * rows;
    const int nslabs = (int)(((size_t)launch_mb << 20) / slab);
    const size_t launch_bytes = nslabs * slab;
    const int nring = (int)(total / launch_bytes);
    const int tps = (planeb + rowb - 1) / rowb;
    const int smem = rows * rowb * stages;
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * cps;
    cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(probe<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(probe<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9;
    const int reps = 8;
    for (int it = 0; it < 5; ++it) {
        cudaEventRecord(e0);
        for (int r = 0; r < reps; ++r)
            for (int q = 0; q < nring; ++q) {
                const char* base = d + (size_t)q * launch_bytes;
                if (mode == 1) plain<<<grid, threads>>>(reinterpret_cast<const uint4*>(base), (long long)(launch_bytes / 16), o);
                else if (stages == 2) probe<2><<<grid, threads, smem>>>(base, nslabs, tps, rows, rowb, planeb, issuers, o);
                else if (stages == 3) probe<3><<<grid, threads, smem>>>(base, nslabs, tps, rows, rowb, planeb, issuers, o);
                else probe<4><<<grid, threads, smem>>>(base, nslabs, tps, rows, rowb, planeb, issuers, o);
            }
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        ms /= reps * nring;
        if (it && ms < best) best = ms;
    }
    printf("mode=%d launch=%.1fMB rows=%d rowb=%d stages=%d threads=%d ctas/sm=%d issuers=%d planeb=%d: %.2f us/launch -> %.0f GB/s  (%s)\n", mode, launch_bytes / 1e6, rows, rowb,
           stages, threads, cps, issuers, planeb, best * 1e3, launch_bytes / best / 1e6, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
