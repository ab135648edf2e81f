// store_patterns.cu — microbenchmark behind DESIGN §4.4 (round 2): how fast can the DecGridRL observation blocks
// (E blocks of 5808 B = [4,3,11,11] f32, env-major, 761 MB at C5) leave the SMs, pattern by pattern?
// Every variant writes the same bytes (0/1 floats expanded from a per-group bit stream through the 16-entry LUT, as
// dec_env_kernel does) and is timed alone with CUDA events; the L2 (126 MB) is far smaller than the buffer.
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o store_patterns store_patterns.cu
//   run:   ./store_patterns [E=131072] [reps=20]   -> one JSON line per variant
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

constexpr int BLK_F4 = 363;                  // float4 per observation block (5808 B)
constexpr int BITS_WORDS = 46;               // 1452 bits

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_s2g(void* g, const void* s, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(g), "r"(smem_u32(s)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ float4 lut4(uint32_t n) {
    return make_float4((n & 1u) ? 1.f : 0.f, (n & 2u) ? 1.f : 0.f, (n & 4u) ? 1.f : 0.f, (n & 8u) ? 1.f : 0.f);
}

// ---- V0/V1: G lanes per block, lane writes float4 q = lg + G*t (what dec_env_kernel does; G = 8 / 16 / 32) --------
template <int G, int MODE>   // MODE 0 .cs, 1 default
__global__ void __launch_bounds__(256, 4) k_group(float* out, int E, long long pitch_f4) {
    __shared__ float4 lut[16];
    __shared__ uint32_t bits[256 / G][BITS_WORDS + 2];
    if (threadIdx.x < 16) lut[threadIdx.x] = lut4(threadIdx.x);
    const int lg = threadIdx.x & (G - 1), gi = threadIdx.x / G, gpc = 256 / G;
    for (int k = lg; k < BITS_WORDS + 2; k += G) bits[gi][k] = 0x9e3779b9u * (k + 1 + gi);
    __syncthreads();
    for (int e = blockIdx.x * gpc + gi; e < E; e += gridDim.x * gpc) {
        float4* d = reinterpret_cast<float4*>(out) + (size_t)e * pitch_f4 + lg;
        const int iters = (BLK_F4 - lg + G - 1) / G;
#pragma unroll 4
        for (int t = 0; t < iters; ++t) {
            const int q = lg + t * G;
            const float4 v = lut[(bits[gi][q >> 3] >> ((q & 7) * 4)) & 15u];
            if (MODE == 0) __stcs(d + t * G, v); else d[t * G] = v;
        }
    }
}

// ---- V2: whole CTA writes the CTA's 32 adjacent blocks as one contiguous burst (bits of 32 blocks in smem) ---------
__global__ void __launch_bounds__(256, 4) k_cta_burst(float* out, int E) {
    __shared__ float4 lut[16];
    __shared__ uint32_t bits[32 * 363 / 8 + 8];          // 32 blocks x 363 nibbles-of-4-floats
    if (threadIdx.x < 16) lut[threadIdx.x] = lut4(threadIdx.x);
    for (int k = threadIdx.x; k < 32 * 363 / 8 + 8; k += 256) bits[k] = 0x9e3779b9u * (k + 1);
    __syncthreads();
    for (int e0 = blockIdx.x * 32; e0 < E; e0 += gridDim.x * 32) {
        float4* d = reinterpret_cast<float4*>(out) + (size_t)e0 * BLK_F4;
        const int n = min(32, E - e0) * BLK_F4;
#pragma unroll 4
        for (int q = threadIdx.x; q < n; q += 256)
            __stcs(d + q, lut[(bits[q >> 3] >> ((q & 7) * 4)) & 15u]);
    }
}

// ---- V3: plain fill (ceiling): grid-stride float4 over the whole buffer ------------------------------------------
__global__ void __launch_bounds__(256) k_fill(float4* out, long long n) {
    const float4 v = make_float4(1.f, 0.f, 1.f, 0.f);
    for (long long q = blockIdx.x * 256ll + threadIdx.x; q < n; q += (long long)gridDim.x * 256) __stcs(out + q, v);
}

// ---- V4: TMA bulk store: G lanes expand CH float4 of the block into a shared-memory chunk, one lane issues
//      cp.async.bulk.global.shared::cta; NBUF chunk buffers per group (wait_group.read NBUF-1 before a refill) -------
template <int G, int CH, int NBUF>
__global__ void __launch_bounds__(256) k_bulk(float* out, int E, int groups_per_cta) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ float4 lut[16];
    __shared__ uint32_t bits[32][BITS_WORDS + 2];
    if (threadIdx.x < 16) lut[threadIdx.x] = lut4(threadIdx.x);
    const int lg = threadIdx.x & (G - 1), gi = threadIdx.x / G;
    if (gi < groups_per_cta) for (int k = lg; k < BITS_WORDS + 2; k += G) bits[gi][k] = 0x9e3779b9u * (k + 1 + gi);
    __syncthreads();
    if (gi >= groups_per_cta) return;
    constexpr int NCH = BLK_F4 / CH;                      // chunks per block (CH divides 363)
    float4* stage = reinterpret_cast<float4*>(smem) + (size_t)gi * NBUF * CH;
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1)));
    int buf = 0;
    for (int e = blockIdx.x * groups_per_cta + gi; e < E; e += gridDim.x * groups_per_cta) {
        float4* d = reinterpret_cast<float4*>(out) + (size_t)e * BLK_F4;
        for (int c = 0; c < NCH; ++c) {
            if (lg == 0) bulk_wait_read<NBUF - 1>();      // the buffer about to be refilled has been read out
            __syncwarp(gmask);
            float4* s = stage + buf * CH;
            for (int q = lg; q < CH; q += G) {
                const int qq = c * CH + q;
                s[q] = lut[(bits[gi][qq >> 3] >> ((qq & 7) * 4)) & 15u];
            }
            fence_proxy_async();
            __syncwarp(gmask);
            if (lg == 0) bulk_s2g(d + c * CH, s, CH * 16);
            buf = (buf + 1) % NBUF;
        }
    }
    if (lg == 0) bulk_wait_all();
}

template <class F> float time_it(F f, int reps) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i) f();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; ++i) f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, a, b));
    CK(cudaGetLastError());
    return ms / reps;
}

void report(const char* name, float ms, double bytes, const char* note) {
    printf("{\"bench\": \"store_patterns\", \"variant\": \"%s\", \"ms\": %.5f, \"GBps\": %.1f, \"note\": \"%s\"}\n", name, ms,
           bytes / (ms * 1e-3) / 1e9, note);
    fflush(stdout);
}

int main(int argc, char** argv) {
    const int E = argc > 1 ? atoi(argv[1]) : 131072;
    const int reps = argc > 2 ? atoi(argv[2]) : 20;
    int sms = 148;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const double bytes = (double)E * 5808.0;
    float* out;
    CK(cudaMalloc(&out, (size_t)E * 5888 + 4096));        // room for the 128-byte padded pitch
    CK(cudaMemset(out, 0, (size_t)E * 5888));
    const int grid4 = sms * 4;

    report("fill_cs_float4", time_it([&] { k_fill<<<sms * 8, 256>>>(reinterpret_cast<float4*>(out), (long long)E * BLK_F4); }, reps), bytes,
           "ceiling: contiguous grid-stride float4 .cs stores");
    report("group8_cs", time_it([&] { k_group<8, 0><<<grid4, 256>>>(out, E, BLK_F4); }, reps), bytes, "dec_env_kernel today: 8 lanes per block, 4 CTAs/SM");
    report("group8_default", time_it([&] { k_group<8, 1><<<grid4, 256>>>(out, E, BLK_F4); }, reps), bytes, "same, default store policy");
    report("group16_cs", time_it([&] { k_group<16, 0><<<grid4, 256>>>(out, E, BLK_F4); }, reps), bytes, "16 lanes per block");
    report("group32_cs", time_it([&] { k_group<32, 0><<<grid4, 256>>>(out, E, BLK_F4); }, reps), bytes, "warp per block");
    report("group8_cs_pitch5888", time_it([&] { k_group<8, 0><<<grid4, 256>>>(out, E, 368); }, reps), bytes,
           "blocks padded to 46 whole 128-byte lines (what alignment alone would give; not the API layout)");
    report("group32_cs_pitch5888", time_it([&] { k_group<32, 0><<<grid4, 256>>>(out, E, 368); }, reps), bytes, "warp per block, padded pitch");
    report("cta_burst_cs", time_it([&] { k_cta_burst<<<grid4, 256>>>(out, E); }, reps), bytes, "CTA writes its 32 adjacent blocks as one 186 kB burst");
    {
        auto run = [&](auto kern, int G, int CH, int NBUF, int gpc, int ctas_per_sm, const char* name, const char* note) {
            const size_t sm = (size_t)gpc * NBUF * CH * 16;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
            report(name, time_it([&] { kern<<<sms * ctas_per_sm, 256, sm>>>(out, E, gpc); }, reps), bytes, note);
        };
        run(k_bulk<8, 33, 2>, 8, 33, 2, 32, 4, "bulk_g8_528Bx2_4cta", "TMA bulk store, 528 B chunks double-buffered, 128 groups/SM");
        run(k_bulk<8, 33, 3>, 8, 33, 3, 32, 4, "bulk_g8_528Bx3_4cta", "528 B chunks triple-buffered, 128 groups/SM");
        run(k_bulk<8, 121, 2>, 8, 121, 2, 32, 1, "bulk_g8_1936Bx2_1cta", "1936 B chunks double-buffered, 32 groups/SM");
        run(k_bulk<8, 121, 2>, 8, 121, 2, 16, 3, "bulk_g8_1936Bx2_3cta", "1936 B chunks double-buffered, 48 groups/SM");
        run(k_bulk<8, 363, 1>, 8, 363, 1, 32, 1, "bulk_g8_5808Bx1_1cta", "whole block staged, one bulk store per env, 32 groups/SM");
        run(k_bulk<32, 363, 2>, 32, 363, 2, 8, 2, "bulk_g32_5808Bx2_2cta", "warp per block, whole block double-buffered, 16 warps/SM");
        run(k_bulk<32, 121, 2>, 32, 121, 2, 8, 4, "bulk_g32_1936Bx2_4cta", "warp per block, 1936 B chunks, 32 warps/SM");
    }
    CK(cudaFree(out));
    return 0;
}
