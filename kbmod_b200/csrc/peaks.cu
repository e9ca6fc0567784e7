// peaks.cu -- kb_measure_peaks: the denominators of the gather roofline, measured on the device the caller is
// about to benchmark, in the same process (SURVEY.md 8d asks for a measured L2 figure next to the HBM copy peak).
//   [0] HBM streaming read       : 2 GB buffer, one LDG.64 per lane, read once per pass
//   [1] L2-resident read         : 48 MB buffer (fits the 126 MB L2), same access, many passes
//   [2] L2-resident row gather   : every warp reads 256 contiguous bytes at a pseudo-random 256 B-aligned offset of a
//                                  32 MB buffer (the access shape of a warp gathering 32 x-adjacent float pairs)
//   [3] shared-memory LDS.64     : 32 lanes x 8 B = 256 B per warp instruction (the tiled kernel's access shape)
//   [4] shared-memory LDS.32     : 128 B per warp instruction (one uint16 pair per lane)
//   [5] shared-memory LDS.U16    : 64 B per warp instruction (one uint8 pair per lane)
// All in GB/s (1e9 bytes), best of a few repetitions, CUDA events on the default stream.  No reference counterpart.
#include "common.cuh"

namespace kb {

__global__ void pk_stream_read(const float2 *__restrict__ p, size_t n, int passes, float *out) {
    float a = 0.f, b = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int k = 0; k < passes; ++k)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const float2 v = __ldg(p + i);
            a += v.x;
            b += v.y;
        }
    if (a + b == 123.456f) out[0] = a;
}

__global__ void pk_row_gather(const float2 *__restrict__ p, uint32_t n_rows, int iters, float *out) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    uint32_t h = warp * 2654435761u + 12345u;
    float a = 0.f, b = 0.f;
#pragma unroll 4
    for (int k = 0; k < iters; ++k) {
        h = h * 1664525u + 1013904223u;
        const uint32_t row = (h >> 8) & (n_rows - 1);
        const float2 v = __ldg(p + (size_t)row * 32 + lane);
        a += v.x;
        b += v.y;
    }
    if (a + b == 123.456f) out[0] = a;
}

// W = bytes per lane per load: 8 (LDS.64), 4 (LDS.32), 2 (LDS.U16)
template <int W>
__global__ void __launch_bounds__(512, 1) pk_smem(int iters, float *out) {
    extern __shared__ __align__(16) unsigned char pk_sm[];
    constexpr int kBytes = 96 * 1024;
    for (int i = threadIdx.x; i < kBytes / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(pk_sm)[i] = (uint32_t)i;
    __syncthreads();
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(pk_sm);
    uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int n = kBytes / W;                   // elements
    const int base = (threadIdx.x >> 5) * 37 * 32 + (threadIdx.x & 31);
    for (int k = 0; k < iters; ++k) {
        const int o = (base + k * 32 * 4) % (n - 4 * 32 * 8);
#pragma unroll
        for (int u = 0; u < 32; ++u) {
            const uint32_t addr = sbase + (uint32_t)(o + u * 32) * W;
            if (W == 8) {
                uint32_t x, y;
                asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(addr));
                acc[u & 7] += x ^ y;
            } else if (W == 4) {
                uint32_t x;
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x) : "r"(addr));
                acc[u & 7] += x;
            } else {
                uint16_t x;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(x) : "r"(addr));
                acc[u & 7] += x;
            }
        }
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; ++i) s += acc[i];
    if (s == 0x12345678u) out[0] = (float)s;
}

template <class F>
static float best_ms(F f, int reps) {
    cudaEvent_t e0, e1;
    KB_CUDA(cudaEventCreate(&e0));
    KB_CUDA(cudaEventCreate(&e1));
    f();
    KB_CUDA(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(e0);
        f();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    KB_CUDA(cudaGetLastError());
    return best;
}

}  // namespace kb

using namespace kb;

extern "C" int kb_measure_peaks(double *gbps, int n) {
    KB_API_BEGIN
    if (!gbps || n < 6) throw Error("kb_measure_peaks: need room for 6 values");
    int dev = 0, sms = 0;
    KB_CUDA(cudaGetDevice(&dev));
    KB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    float *out = nullptr;
    float2 *p = nullptr;
    const size_t big = (size_t)2 << 30, small = (size_t)48 << 20;
    KB_CUDA(cudaMalloc(&out, 4));
    if (cudaMalloc(&p, big) != cudaSuccess) {
        cudaFree(out);
        throw Error("kb_measure_peaks: cannot allocate the 2 GB probe buffer");
    }
    try {
        KB_CUDA(cudaMemset(p, 0, big));
        {
            const size_t nel = big / 8;
            const float ms = best_ms([&] { pk_stream_read<<<sms * 8, 512>>>(p, nel, 1, out); }, 5);
            gbps[0] = (double)big / ms / 1e6;
        }
        {
            const size_t nel = small / 8;
            const int passes = 100;
            const float ms = best_ms([&] { pk_stream_read<<<sms * 8, 512>>>(p, nel, passes, out); }, 5);
            gbps[1] = (double)small * passes / ms / 1e6;
        }
        {
            const uint32_t rows = 1u << 17;  // 32 MB
            const int iters = 2048, blocks = sms * 8, threads = 512;
            const float ms = best_ms([&] { pk_row_gather<<<blocks, threads>>>(p, rows, iters, out); }, 5);
            gbps[2] = (double)blocks * threads * iters * 8 / ms / 1e6;
        }
        const int iters = 2500;
        KB_CUDA(cudaFuncSetAttribute(pk_smem<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        KB_CUDA(cudaFuncSetAttribute(pk_smem<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        KB_CUDA(cudaFuncSetAttribute(pk_smem<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        const double loads = (double)sms * 512 * iters * 32.0;
        gbps[3] = loads * 8 / best_ms([&] { pk_smem<8><<<sms, 512, 96 * 1024>>>(iters, out); }, 5) / 1e6;
        gbps[4] = loads * 4 / best_ms([&] { pk_smem<4><<<sms, 512, 96 * 1024>>>(iters, out); }, 5) / 1e6;
        gbps[5] = loads * 2 / best_ms([&] { pk_smem<2><<<sms, 512, 96 * 1024>>>(iters, out); }, 5) / 1e6;
    } catch (...) {
        cudaFree(p);
        cudaFree(out);
        throw;
    }
    cudaFree(p);
    cudaFree(out);
    KB_API_END
}
