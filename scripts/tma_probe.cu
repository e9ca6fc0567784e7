// Debug probe: one TMA 3-D box load with NaN OOB fill + one bulk copy, checked on the host.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("ERR %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); return 1; } } while (0)
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap tmap, const int *hdr_src, float *out, int *hdr_out, int bw, int rows, int c0, int c1, int c2, int mode) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t *bar = (uint64_t *)(smem + 32768);
    int *hdr = (int *)(smem + 32768 + 64);
    uint32_t b = s32(bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t bytes = (mode & 1 ? (uint32_t)(2 * bw * rows * 4) : 0) + (mode & 2 ? 64u : 0u);
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(b), "r"(bytes) : "memory");
        if (mode & 1)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(s32(smem)), "l"(&tmap), "r"(c0), "r"(c1), "r"(c2), "r"(b) : "memory");
        if (mode & 2)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(hdr)), "l"(hdr_src), "r"(64), "r"(b) : "memory");
    }
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_LOOP:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE;\n\tbra WAIT_LOOP;\n\tDONE:\n\t}" ::"r"(b), "r"(0) : "memory");
    for (int i = threadIdx.x; i < 2 * bw * rows; i += blockDim.x) out[i] = ((float *)smem)[i];
    for (int i = threadIdx.x; i < 16; i += blockDim.x) hdr_out[i] = hdr[i];
}
typedef CUresult (*EncFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
    int mode = argc > 1 ? atoi(argv[1]) : 3;
    const int T = 3, H = 24, W = 32, Wp = 32, ROWS = 4; const int BW = argc > 2 ? atoi(argv[2]) : 40; const int fill = argc > 3 ? atoi(argv[3]) : 1; const int cx = argc > 4 ? atoi(argv[4]) : -3; const int cy = argc > 5 ? atoi(argv[5]) : 22; const int l2p = argc > 6 ? atoi(argv[6]) : 3;
    std::vector<float> h((size_t)T * H * Wp * 2);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)i;
    float *d; CK(cudaMalloc(&d, h.size() * 4)); CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    void *sym = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q));
    CUtensorMap tm;
    cuuint64_t gdim[3] = {2 * W, H, T}, gstr[2] = {(cuuint64_t)Wp * 8, (cuuint64_t)Wp * 8 * H};
    cuuint32_t box[3] = {(cuuint32_t)(2 * BW), ROWS, 1}, es[3] = {1, 1, 1};
    CUresult r = ((EncFn)sym)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)l2p, fill ? CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA : CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    int hh[16]; for (int i = 0; i < 16; ++i) hh[i] = 100 + i;
    int *dh, *dho; float *dout; CK(cudaMalloc(&dh, 64)); CK(cudaMalloc(&dho, 64)); CK(cudaMalloc(&dout, 2 * BW * ROWS * 4));
    CK(cudaMemcpy(dh, hh, 64, cudaMemcpyHostToDevice));
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000));
    probe<<<1, 64, 40000>>>(tm, dh, dout, dho, BW, ROWS, 2 * cx, cy, 1, mode);
    CK(cudaDeviceSynchronize());
    std::vector<float> o(2 * BW * ROWS); int ho[16];
    CK(cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(ho, dho, 64, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int r0 = 0; r0 < ROWS; ++r0) for (int c = 0; c < BW; ++c) {
        int y = cy + r0, x = cx + c; float a = o[(r0 * BW + c) * 2];
        bool oob = y < 0 || y >= H || x < 0 || x >= W;
        float want = oob ? NAN : h[((size_t)(1 * H + y) * Wp + x) * 2];
        if (oob ? !(fill ? std::isnan(a) : a == 0.0f) : a != want) { if (bad < 5) printf("mismatch r%d c%d got %f want %f\n", r0, c, a, want); ++bad; }
    }
    unsigned nanbits; memcpy(&nanbits, &o[0], 4);
    printf("BW %d fill %d cx %d cy %d l2p %d: ", BW, fill, cx, cy, l2p); printf("mode %d bad=%d hdr0=%d hdr15=%d oob bits=%08x\n", mode, bad, ho[0], ho[15], nanbits);
    return 0;
}
