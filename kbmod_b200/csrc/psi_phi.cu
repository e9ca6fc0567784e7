// psi_phi.cu -- psi/phi generation (PSF correlation with shared-memory tiles + halos), uint8/uint16
// quantisation and the layout transforms of the psi/phi stack.
//
// Arithmetic contract: bit-identical to the reference's CPU routines (image_utils_cpp.cpp:20-68,110-177 and
// psi_phi_array.cpp:219-291): taps visited j-outer / i-inner, invalid and out-of-image taps skipped (not
// zero-weighted), every float op rounded separately (no FMA contraction), IEEE division.
#include <float.h>
#include <math_constants.h>

#include "common.cuh"

namespace kb {

static constexpr int CT_X = 32, CT_Y = 16;  // output tile per CTA (one thread per output pixel)

// MODE 0: out = correlate(img, psf).  MODE 1: fused psi/phi: the pre-pass p_psi = sci/var, p_phi = 1/var is
// applied while the tile (with its halo) is staged into shared memory, then both correlations run from the
// same staged tile.
template <int MODE>
__global__ void __launch_bounds__(CT_X *CT_Y)
        conv_kernel(const float *__restrict__ in0, const float *__restrict__ in1, int H, int W,
                    const float *__restrict__ psf, const float *__restrict__ psf2, int dim, float psf_total,
                    float psf2_total, float2 *__restrict__ out_pairs, float *__restrict__ out0,
                    float *__restrict__ out1, int y_write_lo, int y_write_hi) {
    extern __shared__ float smem[];
    const int rad = (dim - 1) / 2;
    const int tw = CT_X + 2 * rad, th = CT_Y + 2 * rad;
    float *s_a = smem;                                   // staged p_psi (or img)
    float *s_b = s_a + (MODE == 1 ? tw * th : 0);        // staged p_phi
    float *s_k = s_b + tw * th;                          // psf
    float *s_k2 = s_k + (MODE == 1 ? dim * dim : 0);     // psf^2
    const int tid = threadIdx.y * CT_X + threadIdx.x;
    const int x0 = blockIdx.x * CT_X - rad, y0 = blockIdx.y * CT_Y - rad;

    for (int i = tid; i < dim * dim; i += CT_X * CT_Y) {
        s_k[i] = psf[i];
        if (MODE == 1) s_k2[i] = psf2[i];
    }
    for (int i = tid; i < tw * th; i += CT_X * CT_Y) {
        const int yy = y0 + i / tw, xx = x0 + i % tw;
        float a = KB_NAN_F, b = KB_NAN_F;  // out-of-image taps are skipped == treated as invalid
        if (xx >= 0 && xx < W && yy >= 0 && yy < H) {
            const size_t g = (size_t)yy * W + xx;
            if (MODE == 0) {
                a = in0[g];
            } else {
                const float s = in0[g], v = in1[g];
                const bool vok = isfinite(v) && v != 0.0f;
                a = (vok && isfinite(s)) ? __fdiv_rn(s, v) : KB_NAN_F;   // image_utils_cpp.cpp:142-149
                b = vok ? __fdiv_rn(1.0f, v) : KB_NAN_F;                // :166-171 (double div == f32 div)
            }
        }
        s_a[i] = a;
        if (MODE == 1) s_b[i] = b;
    }
    __syncthreads();

    const int x = blockIdx.x * CT_X + threadIdx.x, y = blockIdx.y * CT_Y + threadIdx.y;
    if (x >= W || y >= H || y < y_write_lo || y >= y_write_hi) return;  // rows outside the wanted band: halo only
    const int cx = threadIdx.x + rad, cy = threadIdx.y + rad;
    const float ca = s_a[cy * tw + cx];
    float ra, rb = 0.0f;
    {
        float sum = 0.0f, portion = 0.0f;
        if (!isfinite(ca)) {
            ra = ca;  // masked pixels pass through unchanged (:41-44)
        } else {
            for (int j = 0; j < dim; ++j) {
                const float *row = s_a + (threadIdx.y + j) * tw + threadIdx.x;
                const float *krow = s_k + j * dim;
                for (int i = 0; i < dim; ++i) {
                    const float pix = row[i];
                    if (isfinite(pix)) {
                        const float w = krow[i];
                        portion = __fadd_rn(portion, w);
                        sum = __fadd_rn(sum, __fmul_rn(pix, w));
                    }
                }
            }
            ra = (portion == 0.0f) ? KB_NAN_F : __fdiv_rn(__fmul_rn(sum, psf_total), portion);
        }
    }
    if (MODE == 1) {
        const float cb = s_b[cy * tw + cx];
        float sum = 0.0f, portion = 0.0f;
        if (!isfinite(cb)) {
            rb = cb;
        } else {
            for (int j = 0; j < dim; ++j) {
                const float *row = s_b + (threadIdx.y + j) * tw + threadIdx.x;
                const float *krow = s_k2 + j * dim;
                for (int i = 0; i < dim; ++i) {
                    const float pix = row[i];
                    if (isfinite(pix)) {
                        const float w = krow[i];
                        portion = __fadd_rn(portion, w);
                        sum = __fadd_rn(sum, __fmul_rn(pix, w));
                    }
                }
            }
            rb = (portion == 0.0f) ? KB_NAN_F : __fdiv_rn(__fmul_rn(sum, psf2_total), portion);
        }
    }
    const size_t g = (size_t)y * W + x;
    if (MODE == 0) {
        out0[g] = ra;
    } else {
        if (out_pairs) out_pairs[g] = make_float2(ra, rb);
        if (out0) out0[g] = ra;
        if (out1) out1[g] = rb;
    }
}

static size_t conv_smem(int mode, int dim) {
    const int rad = (dim - 1) / 2;
    const size_t tile = (size_t)(CT_X + 2 * rad) * (CT_Y + 2 * rad);
    return sizeof(float) * ((mode == 1 ? 2 : 1) * tile + (mode == 1 ? 2 : 1) * (size_t)dim * dim);
}

int launch_convolve(const float *d_img, int H, int W, const float *d_psf, int dim, float psf_total, float *d_out,
                    cudaStream_t st) {
    dim3 grid((W + CT_X - 1) / CT_X, (H + CT_Y - 1) / CT_Y), block(CT_X, CT_Y);
    const size_t smem = conv_smem(0, dim);
    if (smem > 220 * 1024) throw Error("PSF kernel too large for the convolution tile.");
    if (smem > 48 * 1024)
        KB_CUDA(cudaFuncSetAttribute(conv_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_kernel<0><<<grid, block, smem, st>>>(d_img, nullptr, H, W, d_psf, nullptr, dim, psf_total, 0.0f, nullptr,
                                              d_out, nullptr, 0, H);
    KB_CUDA(cudaGetLastError());
    return 1;
}

int launch_psi_phi(const float *d_sci, const float *d_var, int H, int W, const float *d_psf, const float *d_psf2,
                   int dim, float psf_total, float psf2_total, float2 *d_out_pairs, float *d_psi_plane,
                   float *d_phi_plane, cudaStream_t st, int y_write_lo, int y_write_hi) {
    if (y_write_hi < 0) y_write_hi = H;
    dim3 grid((W + CT_X - 1) / CT_X, (H + CT_Y - 1) / CT_Y), block(CT_X, CT_Y);
    const size_t smem = conv_smem(1, dim);
    if (smem > 220 * 1024) throw Error("PSF kernel too large for the convolution tile.");
    if (smem > 48 * 1024)
        KB_CUDA(cudaFuncSetAttribute(conv_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_kernel<1><<<grid, block, smem, st>>>(d_sci, d_var, H, W, d_psf, d_psf2, dim, psf_total, psf2_total,
                                              d_out_pairs, d_psi_plane, d_phi_plane, y_write_lo, y_write_hi);
    KB_CUDA(cudaGetLastError());
    return 1;
}

__global__ void interleave_kernel(const float *__restrict__ psi, const float *__restrict__ phi, uint64_t n,
                                  float2 *__restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        out[i] = make_float2(psi[i], phi[i]);
}

int launch_interleave(const float *d_psi, const float *d_phi, uint64_t n, float2 *d_out, cudaStream_t st) {
    interleave_kernel<<<148 * 8, 256, 0, st>>>(d_psi, d_phi, n, d_out);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- min / max over finite values (psi_phi_array.cpp:219-232) -----------------------------------------
// Floats are mapped to unsigned ints whose order matches the float order, so atomicMin/atomicMax apply.
__device__ __host__ inline uint32_t float_to_ordered(float f) {
    uint32_t u;
#ifdef __CUDA_ARCH__
    u = __float_as_uint(f);
#else
    memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
float ordered_to_float(uint32_t u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

__global__ void minmax_kernel(const float2 *__restrict__ pairs, uint64_t n, uint32_t *mm) {
    uint32_t mn0 = 0xffffffffu, mx0 = 0u, mn1 = 0xffffffffu, mx1 = 0u;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const float2 v = pairs[i];
        if (isfinite(v.x)) {
            const uint32_t o = float_to_ordered(v.x);
            mn0 = min(mn0, o);
            mx0 = max(mx0, o);
        }
        if (isfinite(v.y)) {
            const uint32_t o = float_to_ordered(v.y);
            mn1 = min(mn1, o);
            mx1 = max(mx1, o);
        }
    }
    for (int d = 16; d > 0; d >>= 1) {
        mn0 = min(mn0, __shfl_xor_sync(0xffffffffu, mn0, d));
        mx0 = max(mx0, __shfl_xor_sync(0xffffffffu, mx0, d));
        mn1 = min(mn1, __shfl_xor_sync(0xffffffffu, mn1, d));
        mx1 = max(mx1, __shfl_xor_sync(0xffffffffu, mx1, d));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(mm + 0, mn0);
        atomicMax(mm + 1, mx0);
        atomicMin(mm + 2, mn1);
        atomicMax(mm + 3, mx1);
    }
}

int launch_minmax(const float2 *d_pairs, uint64_t n, uint32_t *d_mm, cudaStream_t st) {
    const uint32_t init[4] = {0xffffffffu, 0u, 0xffffffffu, 0u};
    KB_CUDA(cudaMemcpyAsync(d_mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
    minmax_kernel<<<148 * 8, 256, 0, st>>>(d_pairs, n, d_mm);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// The kernels below visit rows [r0, r0 + nr) of each of the T images of H rows (the whole stack: r0 = 0, nr = H).
__device__ __forceinline__ uint64_t stack_row(uint64_t r, int H, int r0, int nr) {
    return (r / (uint64_t)nr) * (uint64_t)H + (uint64_t)r0 + r % (uint64_t)nr;
}

// One pass over rows [r0, r0 + nr) of every image of a freshly built float stack (reference order, unpitched):
// out[0] = 1 when some pair with both values finite is exactly (0, 0) (then invalid pairs cannot be stored as zeros),
// out[1..4] = {psi_min, psi_max, phi_min, phi_max} over the finite values of each channel, as ordered uints,
// out[5] = 1 when some pair is invalid (either value not finite).
__global__ void scan_raw_rows_kernel(const float2 *__restrict__ pairs, int T, int H, int r0, int nr, int W,
                                     uint32_t *out) {
    const uint64_t total = (uint64_t)T * nr * (uint64_t)W;
    uint32_t mn0 = 0xffffffffu, mx0 = 0u, mn1 = 0xffffffffu, mx1 = 0u;
    bool hit = false, bad = false;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const float2 v = pairs[stack_row(i / W, H, r0, nr) * (uint64_t)W + i % W];
        const bool fa = isfinite(v.x), fb = isfinite(v.y);
        if (fa) {
            const uint32_t o = float_to_ordered(v.x);
            mn0 = min(mn0, o);
            mx0 = max(mx0, o);
        }
        if (fb) {
            const uint32_t o = float_to_ordered(v.y);
            mn1 = min(mn1, o);
            mx1 = max(mx1, o);
        }
        hit |= fa && fb && v.x == 0.0f && v.y == 0.0f;
        bad |= !(fa && fb);
    }
    for (int d = 16; d > 0; d >>= 1) {
        mn0 = min(mn0, __shfl_xor_sync(0xffffffffu, mn0, d));
        mx0 = max(mx0, __shfl_xor_sync(0xffffffffu, mx0, d));
        mn1 = min(mn1, __shfl_xor_sync(0xffffffffu, mn1, d));
        mx1 = max(mx1, __shfl_xor_sync(0xffffffffu, mx1, d));
    }
    const bool any_hit = __any_sync(0xffffffffu, hit), any_bad = __any_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0) {
        if (any_hit) atomicExch(out, 1u);
        if (any_bad) atomicExch(out + 5, 1u);   // some pair is invalid
        atomicMin(out + 1, mn0);
        atomicMax(out + 2, mx0);
        atomicMin(out + 3, mn1);
        atomicMax(out + 4, mx1);
    }
}

int launch_scan_raw_rows(const float2 *d_pairs, int T, int H, int r0, int nr, int W, uint32_t *d_out, cudaStream_t st) {
    const uint32_t init[6] = {0u, 0xffffffffu, 0u, 0xffffffffu, 0u, 0u};
    KB_CUDA(cudaMemcpyAsync(d_out, init, sizeof(init), cudaMemcpyHostToDevice, st));
    if (nr > 0) scan_raw_rows_kernel<<<148 * 8, 256, 0, st>>>(d_pairs, T, H, r0, nr, W, d_out);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- quantiser (psi_phi_array_ds.h:40-43, psi_phi_array.cpp:264-291) --------------------------------
__device__ __forceinline__ unsigned encode_value(float v, float mn, float safe_mx, float scale) {
    if (!isfinite(v)) return 0u;
    float c = (v < safe_mx) ? v : safe_mx;  // std::min(value, max_val)
    c = (c < mn) ? mn : c;                  // std::max(., min_val)
    const float q = __fdiv_rn(__fsub_rn(c, mn), scale);
    const float code = (float)__dadd_rn((double)q, 1.0);
    return (unsigned)(int)code;  // static_cast<T>: truncate, then wrap to the storage width
}

template <typename T2, typename T1>
__global__ void encode_kernel(const float2 *__restrict__ pairs, uint64_t n, float pmn, float psafe, float pscale,
                              float fmn, float fsafe, float fscale, T2 *__restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const float2 v = pairs[i];
        T2 o;
        o.x = (T1)encode_value(v.x, pmn, psafe, pscale);
        o.y = (T1)encode_value(v.y, fmn, fsafe, fscale);
        out[i] = o;
    }
}

int launch_encode(const float2 *d_pairs, uint64_t n, int nb, const float psi_p[3], const float phi_p[3], void *d_out,
                  cudaStream_t st) {
    const float psafe = (float)((double)psi_p[1] - (double)psi_p[2] / 100.0);
    const float fsafe = (float)((double)phi_p[1] - (double)phi_p[2] / 100.0);
    if (nb == 1)
        encode_kernel<uchar2, unsigned char><<<148 * 8, 256, 0, st>>>(d_pairs, n, psi_p[0], psafe, psi_p[2],
                                                                      phi_p[0], fsafe, phi_p[2], (uchar2 *)d_out);
    else
        encode_kernel<ushort2, unsigned short><<<148 * 8, 256, 0, st>>>(d_pairs, n, psi_p[0], psafe, psi_p[2],
                                                                        phi_p[0], fsafe, phi_p[2], (ushort2 *)d_out);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- reference layout -> canonical pitched search copy ------------------------------------------------
template <typename T2>
__device__ __forceinline__ T2 canon(T2 v);
template <>
__device__ __forceinline__ float2 canon(float2 v) {
    return (isfinite(v.x) && isfinite(v.y)) ? v : make_float2(KB_NAN_F, KB_NAN_F);
}
template <>
__device__ __forceinline__ ushort2 canon(ushort2 v) {
    return (v.x != 0 && v.y != 0) ? v : make_ushort2(0, 0);
}
template <>
__device__ __forceinline__ uchar2 canon(uchar2 v) {
    return (v.x != 0 && v.y != 0) ? v : make_uchar2(0, 0);
}

template <typename T2>
__device__ __forceinline__ T2 invalid_pair();
template <>
__device__ __forceinline__ float2 invalid_pair<float2>() { return make_float2(KB_NAN_F, KB_NAN_F); }
template <>
__device__ __forceinline__ ushort2 invalid_pair<ushort2>() { return make_ushort2(0, 0); }
template <>
__device__ __forceinline__ uchar2 invalid_pair<uchar2>() { return make_uchar2(0, 0); }

template <typename T2>
__global__ void canonicalise_kernel(const T2 *src, int src_pitch, T2 *dst, int T, int H, int r0, int nr, int W, int Wp) {
    const uint64_t total = (uint64_t)T * nr * (uint64_t)Wp;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t g = stack_row(i / Wp, H, r0, nr);
        const int c = (int)(i % Wp);
        dst[g * Wp + c] = (c < W) ? canon(src[g * src_pitch + c]) : invalid_pair<T2>();  // padding columns: invalid
    }
}

// float stacks, zero-marker form: invalid -> (+0, +0)
__global__ void canonicalise_zero_kernel(const float2 *src, int src_pitch, float2 *dst, int T, int H, int r0, int nr,
                                         int W, int Wp) {
    const uint64_t total = (uint64_t)T * nr * (uint64_t)Wp;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t g = stack_row(i / Wp, H, r0, nr);
        const int c = (int)(i % Wp);
        float2 v = make_float2(0.0f, 0.0f);
        if (c < W) {
            const float2 s = src[g * src_pitch + c];
            if (isfinite(s.x) && isfinite(s.y)) v = s;
        }
        dst[g * Wp + c] = v;
    }
}

__global__ void detect_valid_zero_kernel(const float2 *src, int src_pitch, uint64_t rows, int W, uint32_t *flag) {
    const uint64_t total = rows * (uint64_t)W;
    bool hit = false;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const float2 s = src[(i / W) * src_pitch + (i % W)];
        hit |= (s.x == 0.0f && s.y == 0.0f);  // finite by construction
    }
    if (__any_sync(0xffffffffu, hit) && (threadIdx.x & 31) == 0) atomicExch(flag, 1u);
}

int launch_detect_valid_zero_pairs(const void *d_src, int src_pitch, int T, int H, int W, uint32_t *d_flag,
                                   cudaStream_t st) {
    KB_CUDA(cudaMemsetAsync(d_flag, 0, sizeof(uint32_t), st));
    detect_valid_zero_kernel<<<148 * 8, 256, 0, st>>>((const float2 *)d_src, src_pitch, (uint64_t)T * H, W, d_flag);
    KB_CUDA(cudaGetLastError());
    return 1;
}
int launch_canonicalise_rows(const void *d_src, int src_pitch, void *d_dst, int T, int H, int W, int Wp, int nb,
                             int zero_marker, int r0, int nr, cudaStream_t st) {
    if (d_src == d_dst && src_pitch != Wp) throw Error("in-place canonicalise needs equal pitches");
    if (nr <= 0) return 0;
    if (nb == 4 && zero_marker)
        canonicalise_zero_kernel<<<148 * 8, 256, 0, st>>>((const float2 *)d_src, src_pitch, (float2 *)d_dst, T, H, r0, nr, W, Wp);
    else if (nb == 4)
        canonicalise_kernel<float2><<<148 * 8, 256, 0, st>>>((const float2 *)d_src, src_pitch, (float2 *)d_dst, T, H, r0, nr, W, Wp);
    else if (nb == 2)
        canonicalise_kernel<ushort2><<<148 * 8, 256, 0, st>>>((const ushort2 *)d_src, src_pitch, (ushort2 *)d_dst, T, H, r0, nr, W, Wp);
    else
        canonicalise_kernel<uchar2><<<148 * 8, 256, 0, st>>>((const uchar2 *)d_src, src_pitch, (uchar2 *)d_dst, T, H, r0, nr, W, Wp);
    KB_CUDA(cudaGetLastError());
    return 1;
}
int launch_canonicalise(const void *d_src, int src_pitch, void *d_dst, int T, int H, int W, int Wp, int nb,
                        int zero_marker, cudaStream_t st) {
    return launch_canonicalise_rows(d_src, src_pitch, d_dst, T, H, W, Wp, nb, zero_marker, 0, H, st);
}

// ---- the screening copy: every pair of the canonical search copy as two signed 16-bit fixed-point numbers ----------
// word = (uint16)rint(psi / step_psi) | (uint16)rint(phi / step_phi) << 16, an invalid pair = 0.  The search kernel
// adds these words up exactly, in integers, to bound a candidate's likelihood from above; candidates that cannot be
// ruled out are then evaluated exactly from the search copy.  |value - q * step| <= 0.502 * step (the quotient's own
// rounding included); a valid pair whose two numbers both round to 0 gets phi = 1 instead (one more half step), so
// that "word == 0" means "no data" -- which is what the kernel's error bound (search_tiled.cu, screen_bounds) and
// its observation counter assume.
__device__ __forceinline__ int quantise_value(float v, float step) {
    const int q = __float2int_rn(__fdiv_rn(v, step));
    return max(-32767, min(32767, q));
}

template <int NB>
__global__ void quantise_kernel(const void *__restrict__ src, uint32_t *__restrict__ dst, int T, int H, int r0, int nr,
                                int Wp, int zero_marker, float psi_scale, float psi_min, float phi_scale,
                                float phi_min, float step_psi, float step_phi) {
    const uint64_t total = (uint64_t)T * nr * (uint64_t)Wp;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t g = stack_row(i / Wp, H, r0, nr) * (uint64_t)Wp + i % Wp;
        float a, b;
        bool ok;
        if (NB == 4) {
            const float2 v = reinterpret_cast<const float2 *>(src)[g];
            a = v.x;
            b = v.y;
            ok = zero_marker ? (((__float_as_uint(a) | __float_as_uint(b)) << 1) != 0u) : (b == b);
        } else if (NB == 2) {
            const ushort2 v = reinterpret_cast<const ushort2 *>(src)[g];
            ok = v.x != 0;  // canonical: both or none
            a = ok ? (float)__dadd_rn(__dmul_rn((double)v.x - 1.0, (double)psi_scale), (double)psi_min) : 0.0f;
            b = ok ? (float)__dadd_rn(__dmul_rn((double)v.y - 1.0, (double)phi_scale), (double)phi_min) : 0.0f;
        } else {
            const uchar2 v = reinterpret_cast<const uchar2 *>(src)[g];
            ok = v.x != 0;
            a = ok ? (float)__dadd_rn(__dmul_rn((double)v.x - 1.0, (double)psi_scale), (double)psi_min) : 0.0f;
            b = ok ? (float)__dadd_rn(__dmul_rn((double)v.y - 1.0, (double)phi_scale), (double)phi_min) : 0.0f;
        }
        uint32_t w = 0u;
        if (ok) {
            const int qa = quantise_value(a, step_psi);
            int qb = quantise_value(b, step_phi);
            if (qa == 0 && qb == 0) qb = 1;
            w = ((uint32_t)qa & 0xffffu) | ((uint32_t)qb << 16);
        }
        dst[g] = w;
    }
}

int launch_quantise_rows(const void *d_search, uint32_t *d_screen, int T, int H, int Wp, int nb, int zero_marker,
                         float psi_scale, float psi_min, float phi_scale, float phi_min, float step_psi,
                         float step_phi, int r0, int nr, cudaStream_t st) {
    if (nr <= 0) return 0;
    if (nb == 4)
        quantise_kernel<4><<<148 * 8, 256, 0, st>>>(d_search, d_screen, T, H, r0, nr, Wp, zero_marker, psi_scale, psi_min,
                                                    phi_scale, phi_min, step_psi, step_phi);
    else if (nb == 2)
        quantise_kernel<2><<<148 * 8, 256, 0, st>>>(d_search, d_screen, T, H, r0, nr, Wp, zero_marker, psi_scale, psi_min,
                                                    phi_scale, phi_min, step_psi, step_phi);
    else
        quantise_kernel<1><<<148 * 8, 256, 0, st>>>(d_search, d_screen, T, H, r0, nr, Wp, zero_marker, psi_scale, psi_min,
                                                    phi_scale, phi_min, step_psi, step_phi);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- run planes of the screening copy -----------------------------------------------------------------------------------
// Exposures taken minutes apart (several per night) put a slow mover on the same pixel, so a candidate's gathers for
// such a run of images read one pixel position in each of them.  For every run the screening copy carries one more
// plane: per pixel the sum of the run's valid values, quantised with m times the step (m = images in the run), so
// that  m * q_run * step  stands for the m single words with the same error bound (<= 0.502 m step; a pixel with data
// in the run never maps to the word 0).  The search kernel adds  m * q_run  with one IDP.2A whose byte multiplier
// is m, wherever all candidates of a group keep their pixel offsets through the run.
template <int NB>
__device__ __forceinline__ bool decoded_pair(const void *src, uint64_t g, int zero_marker, float psi_scale, float psi_min,
                                             float phi_scale, float phi_min, float &a, float &b) {
    if (NB == 4) {
        const float2 v = reinterpret_cast<const float2 *>(src)[g];
        a = v.x;
        b = v.y;
        return zero_marker ? (((__float_as_uint(a) | __float_as_uint(b)) << 1) != 0u) : (b == b);
    } else if (NB == 2) {
        const ushort2 v = reinterpret_cast<const ushort2 *>(src)[g];
        if (v.x == 0) return false;  // canonical: both or none
        a = (float)__dadd_rn(__dmul_rn((double)v.x - 1.0, (double)psi_scale), (double)psi_min);
        b = (float)__dadd_rn(__dmul_rn((double)v.y - 1.0, (double)phi_scale), (double)phi_min);
        return true;
    } else {
        const uchar2 v = reinterpret_cast<const uchar2 *>(src)[g];
        if (v.x == 0) return false;
        a = (float)__dadd_rn(__dmul_rn((double)v.x - 1.0, (double)psi_scale), (double)psi_min);
        b = (float)__dadd_rn(__dmul_rn((double)v.y - 1.0, (double)phi_scale), (double)phi_min);
        return true;
    }
}

template <int NB>
__global__ void run_planes_kernel(const void *__restrict__ src, uint32_t *__restrict__ screen, int T, int H, int r0, int nr,
                                  int Wp, int zero_marker, float psi_scale, float psi_min, float phi_scale, float phi_min,
                                  float step_psi, float step_phi, const int2 *__restrict__ runs, int n_runs) {
    const uint64_t per_run = (uint64_t)nr * (uint64_t)Wp, total = per_run * (uint64_t)n_runs;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i / per_run);
        const uint64_t in_img = (uint64_t)r0 * Wp + i % per_run;   // (row, column) offset inside an image
        const int2 run = runs[c];                                   // {first image, number of images}
        double sa = 0.0, sb = 0.0;
        bool any = false;
        for (int k = 0; k < run.y; ++k) {
            float a, b;
            if (decoded_pair<NB>(src, (uint64_t)(run.x + k) * H * Wp + in_img, zero_marker, psi_scale, psi_min, phi_scale,
                                 phi_min, a, b)) {
                sa += (double)a;
                sb += (double)b;
                any = true;
            }
        }
        uint32_t w = 0u;
        if (any) {
            const double m = (double)run.y;
            int qa = __double2int_rn(sa / (m * (double)step_psi)), qb = __double2int_rn(sb / (m * (double)step_phi));
            qa = max(-32767, min(32767, qa));
            qb = max(-32767, min(32767, qb));
            if (qa == 0 && qb == 0) qb = 1;
            w = ((uint32_t)qa & 0xffffu) | ((uint32_t)qb << 16);
        }
        screen[(uint64_t)(T + c) * H * Wp + in_img] = w;
    }
}

int launch_run_planes_rows(const void *d_search, uint32_t *d_screen, int T, int H, int Wp, int nb, int zero_marker,
                           float psi_scale, float psi_min, float phi_scale, float phi_min, float step_psi, float step_phi,
                           const int2 *d_runs, int n_runs, int r0, int nr, cudaStream_t st) {
    if (nr <= 0 || n_runs <= 0) return 0;
#define KB_RUNS(NBV)                                                                                                   \
    run_planes_kernel<NBV><<<148 * 8, 256, 0, st>>>(d_search, d_screen, T, H, r0, nr, Wp, zero_marker, psi_scale, psi_min, \
                                                    phi_scale, phi_min, step_psi, step_phi, d_runs, n_runs)
    if (nb == 4) KB_RUNS(4);
    else if (nb == 2) KB_RUNS(2);
    else KB_RUNS(1);
#undef KB_RUNS
    KB_CUDA(cudaGetLastError());
    return 1;
}

// max |psi| and max |phi| over the valid pairs of a canonical float search copy (padding and invalid pairs are NaN
// or zero and do not matter): d_mm = {psi_min, psi_max, phi_min, phi_max} as ordered uints, like launch_minmax.
int launch_minmax_pitched(const void *d_search, uint64_t n_pairs, uint32_t *d_mm, cudaStream_t st) {
    return launch_minmax(reinterpret_cast<const float2 *>(d_search), n_pairs, d_mm, st);
}

}  // namespace kb
