// Denominators for the gather roofline (SURVEY.md 8d asks for a measured L2 figure next to the HBM copy peak):
//   (1) HBM streaming read        : 8 GB buffer, LDG.64 per lane, read once per pass
//   (2) L2-resident read          : 48 MB buffer (fits the 126 MB L2), same access, many passes
//   (3) L2-resident row gather    : every warp reads 256 contiguous bytes at a pseudo-random 256 B-aligned offset
//                                   of the 48 MB buffer (the generic search kernel's access shape)
//   (4) shared-memory LDS.64      : 32 lanes x 8 B = 256 B per warp instruction (the tiled kernel's access shape)
// Build + run on the box:  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/peaks scripts/peaks_micro.cu && /tmp/peaks
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__global__ void stream_read(const float2* __restrict__ p, size_t n, int passes, float* out) {
    float a = 0.f, b = 0.f;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int k = 0; k < passes; ++k)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            float2 v = __ldg(p + i);
            a += v.x; b += v.y;
        }
    if (a + b == 123.456f) out[0] = a;
}

__global__ void row_gather(const float2* __restrict__ p, uint32_t n_rows, int iters, float* out) {
    // n_rows = number of 256 B rows (power of two); one row = 32 float2
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    uint32_t h = warp * 2654435761u + 12345u;
    float a = 0.f, b = 0.f;
#pragma unroll 4
    for (int k = 0; k < iters; ++k) {
        h = h * 1664525u + 1013904223u;
        uint32_t row = (h >> 8) & (n_rows - 1);
        float2 v = __ldg(p + (size_t)row * 32 + lane);
        a += v.x; b += v.y;
    }
    if (a + b == 123.456f) out[0] = a;
}

__global__ void __launch_bounds__(512, 1) smem_lds64(int iters, float* out) {
    extern __shared__ float2 sm[];
    const int n = 96 * 1024 / 8;   // 96 KB of pairs
    for (int i = threadIdx.x; i < n; i += blockDim.x) sm[i] = make_float2((float)i, 1.f);
    __syncthreads();
    float a0 = 0, a1 = 0, a2 = 0, a3 = 0, b0 = 0, b1 = 0, b2 = 0, b3 = 0;
    int base = (threadIdx.x >> 5) * 37 * 32 + (threadIdx.x & 31);
    for (int k = 0; k < iters; ++k) {
        int o = (base + k * 32 * 4) % (n - 4 * 32 * 8);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            float2 v0 = sm[o + (u * 4 + 0) * 32], v1 = sm[o + (u * 4 + 1) * 32];
            float2 v2 = sm[o + (u * 4 + 2) * 32], v3 = sm[o + (u * 4 + 3) * 32];
            a0 += v0.x; b0 += v0.y; a1 += v1.x; b1 += v1.y; a2 += v2.x; b2 += v2.y; a3 += v3.x; b3 += v3.y;
        }
    }
    float s = a0 + a1 + a2 + a3 + b0 + b1 + b2 + b3;
    if (s == 123.456f) out[0] = s;
}

template <class F> static float time_ms(F f, int reps) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    float* out; CK(cudaMalloc(&out, 4));
    const size_t big = (size_t)8 << 30, small = (size_t)48 << 20;
    float2* p; CK(cudaMalloc(&p, big)); CK(cudaMemset(p, 0, big));
    {   // (1) HBM
        size_t n = big / 8;
        float ms = time_ms([&] { stream_read<<<sms * 8, 512>>>(p, n, 1, out); }, 5);
        printf("{\"what\": \"hbm_stream_read\", \"bytes\": %zu, \"ms\": %.3f, \"GBps\": %.1f}\n", big, ms, big / ms / 1e6);
    }
    {   // (2) L2 resident streaming
        size_t n = small / 8; int passes = 200;
        float ms = time_ms([&] { stream_read<<<sms * 8, 512>>>(p, n, passes, out); }, 5);
        printf("{\"what\": \"l2_resident_read\", \"bytes\": %zu, \"passes\": %d, \"ms\": %.3f, \"GBps\": %.1f}\n", small, passes, ms,
               (double)small * passes / ms / 1e6);
    }
    {   // (3) L2 resident 256 B row gather
        uint32_t rows = 1u << 17;  // 32 MB
        int iters = 4096; int blocks = sms * 8, threads = 512;
        float ms = time_ms([&] { row_gather<<<blocks, threads>>>(p, rows, iters, out); }, 5);
        double bytes = (double)blocks * threads * iters * 8;
        printf("{\"what\": \"l2_resident_row_gather_256B\", \"working_set\": %u, \"ms\": %.3f, \"GBps\": %.1f}\n", rows * 256u, ms, bytes / ms / 1e6);
    }
    {   // (4) shared memory LDS.64
        CK(cudaFuncSetAttribute(smem_lds64, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        int iters = 20000;
        float ms = time_ms([&] { smem_lds64<<<sms, 512, 96 * 1024>>>(iters, out); }, 5);
        double bytes = (double)sms * 512 * iters * 32.0 * 8;
        double gbps = bytes / ms / 1e6;
        printf("{\"what\": \"smem_lds64\", \"ms\": %.3f, \"GBps\": %.1f, \"bytes_per_clk_per_sm_at_max_clock\": %.1f, \"sms\": %d, \"max_clock_mhz\": %.0f}\n",
               ms, gbps, gbps * 1e9 / sms / (clk_khz * 1e3), sms, clk_khz / 1e3);
    }
    return 0;
}
