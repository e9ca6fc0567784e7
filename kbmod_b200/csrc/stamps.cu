// stamps.cu -- coadded stamps of the search results, on the device.
//
//   reference (Python: numpy, numba, torch)                                       here
//   predict_pixel_locations(centered, as_int)   trajectory_utils.py:64-73        stamp_center()
//   extract_stamp / extract_stamp_stack         core/stamp_utils.py:352-439      gathered on the fly, never stored
//   coadd_sum / coadd_mean / coadd_median / coadd_weighted   stamp_utils.py:241-349   coadd_kernel
//   append_coadds (loop over results)           filters/stamp_filters.py:72-160  one CTA per (result, 128 stamp pixels)
//
// The reference's stamp stacks are float64 (np.full(..., np.nan)), so sums, means and the weighted mean are carried
// in double, image by image in time order (numpy reduces axis 0 of a T x S x S array slice by slice), and only the
// final value is narrowed to float32 -- bit-identical to the reference.  The median is torch.nanmedian's: the lower of
// the two middle values for an even count.  Pixels with no data at any time give 0 in every coadd.
#include <math.h>
#include <math_constants.h>

#include "common.cuh"

namespace kb {

static constexpr int kStampThreads = 128;

__device__ __forceinline__ int stamp_center(double v, double t, double p0) {
    // pos = v * t + p0 (two roundings, no FMA), + 0.5, astype(int): truncation toward zero
    return (int)__dadd_rn(__dadd_rn(__dmul_rn(v, t), p0), 0.5);
}

struct CoaddOut {
    float *sum, *mean, *median, *weighted;   // each n * S * S or nullptr
};

__global__ void __launch_bounds__(kStampThreads)
        coadd_kernel(const float *__restrict__ sci, const float *__restrict__ var, int T, int H, int W,
                     const double *__restrict__ times, const double *__restrict__ x0, const double *__restrict__ y0,
                     const double *__restrict__ vx, const double *__restrict__ vy, int radius,
                     const uint8_t *__restrict__ obs_valid, CoaddOut out, int med_stride) {
    extern __shared__ unsigned char s_raw[];
    int *s_px = reinterpret_cast<int *>(s_raw);            // T
    int *s_py = s_px + T;                                  // T
    float *s_val = reinterpret_cast<float *>(s_py + T);    // T * med_stride (median only)
    const uint64_t r = blockIdx.x;
    const int S = 2 * radius + 1, npix = S * S;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const bool use = obs_valid ? obs_valid[r * T + t] != 0 : true;
        s_px[t] = use ? stamp_center(vx[r], times[t], x0[r]) : INT_MIN;   // INT_MIN = image not selected
        s_py[t] = use ? stamp_center(vy[r], times[t], y0[r]) : INT_MIN;
    }
    __syncthreads();
    const int p = blockIdx.y * blockDim.x + threadIdx.x;
    const bool active = p < npix;
    const int dy = active ? p / S - radius : 0, dx = active ? p % S - radius : 0;
    const size_t img = (size_t)H * W;
    double s = 0.0, ws = 0.0, wsum = 0.0;
    int cnt = 0, n_sel = 0;
    if (active) {
        for (int t = 0; t < T; ++t) {
            const int cx = s_px[t], cy = s_py[t];
            float v = KB_NAN_F, vv = KB_NAN_F;
            if (cx != INT_MIN) {
                const long long x = (long long)cx + dx, y = (long long)cy + dy;
                if (x >= 0 && x < W && y >= 0 && y < H) {
                    v = sci[(size_t)t * img + (size_t)y * W + x];
                    if (var) vv = var[(size_t)t * img + (size_t)y * W + x];
                }
                ++n_sel;
                const bool is_nan = v != v;
                s = __dadd_rn(s, is_nan ? 0.0 : (double)v);                    // np.nansum
                cnt += is_nan ? 0 : 1;
                if (out.weighted && !is_nan && !(vv != vv) && vv != 0.0f) {   // stamp_utils.py:329-339
                    const double w = __ddiv_rn(1.0, (double)vv);
                    ws = __dadd_rn(ws, __dmul_rn((double)v, w));
                    wsum = __dadd_rn(wsum, w);
                }
            }
            if (out.median) s_val[(size_t)t * med_stride + threadIdx.x] = v;   // NaN where not selected / no data
        }
        if (out.weighted && cnt == 0 && n_sel > 0) {
            // _mask_all_nans turned this pixel's science values into 0.0 at every selected time, so every time with
            // a usable variance contributes 0.0 * w to the numerator and w to the denominator
            ws = 0.0;
            wsum = 0.0;
            for (int t = 0; t < T; ++t) {
                const int cx = s_px[t], cy = s_py[t];
                if (cx == INT_MIN) continue;
                const long long x = (long long)cx + dx, y = (long long)cy + dy;
                if (!(x >= 0 && x < W && y >= 0 && y < H)) continue;
                const float vv = var[(size_t)t * img + (size_t)y * W + x];
                if (!(vv != vv) && vv != 0.0f) {
                    const double w = __ddiv_rn(1.0, (double)vv);
                    ws = __dadd_rn(ws, __dmul_rn(0.0, w));
                    wsum = __dadd_rn(wsum, w);
                }
            }
        }
        const size_t o = (size_t)r * npix + p;
        if (out.sum) out.sum[o] = __double2float_rn(s);
        if (out.mean) out.mean[o] = cnt ? __double2float_rn(__ddiv_rn(s, (double)cnt)) : 0.0f;
        if (out.weighted) {
            if (wsum == 0.0) wsum = 1e24;
            out.weighted[o] = n_sel ? __double2float_rn(__ddiv_rn(ws, wsum)) : 0.0f;
        }
        if (out.median) {
            float med = 0.0f;
            if (cnt > 0) {
                const int want = (cnt - 1) / 2;
                for (int i = 0; i < T; ++i) {
                    const float a = s_val[(size_t)i * med_stride + threadIdx.x];
                    if (a != a) continue;
                    int less = 0;
                    for (int j = 0; j < T; ++j) {
                        const float b = s_val[(size_t)j * med_stride + threadIdx.x];
                        less += (b < a || (b == a && j < i)) ? 1 : 0;
                    }
                    if (less == want) {
                        med = a;
                        break;
                    }
                }
            }
            out.median[o] = med;
        }
    }
}

// append_all_stamps (filters/stamp_filters.py:163-211): every stamp of every result, T x S x S per result, NaN where
// there is no image.  One thread per output pixel.
__global__ void all_stamps_kernel(const float *__restrict__ sci, int T, int H, int W, const double *__restrict__ times,
                                  const double *__restrict__ x0, const double *__restrict__ y0,
                                  const double *__restrict__ vx, const double *__restrict__ vy, uint64_t n, int radius,
                                  float *__restrict__ out) {
    const int S = 2 * radius + 1;
    const uint64_t per = (uint64_t)T * S * S, total = n * per;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / per;
        const int rem = (int)(i % per), t = rem / (S * S), p = rem % (S * S);
        const long long x = (long long)stamp_center(vx[r], times[t], x0[r]) + (p % S - radius);
        const long long y = (long long)stamp_center(vy[r], times[t], y0[r]) + (p / S - radius);
        float v = KB_NAN_F;
        if (x >= 0 && x < W && y >= 0 && y < H) v = sci[(size_t)t * H * W + (size_t)y * W + x];
        out[i] = v;
    }
}

namespace {
struct DBuf {
    void *p = nullptr;
    explicit DBuf(size_t bytes) { KB_CUDA(cudaMalloc(&p, bytes ? bytes : 1)); }
    ~DBuf() { cudaFree(p); }
    template <class T> T *as() { return static_cast<T *>(p); }
};
}  // namespace
}  // namespace kb

using namespace kb;

extern "C" KB_EXPORT int kb_coadd_stamps(const float *const *sci, const float *const *var, int T, int H, int W,
                                         const double *times, const double *x0, const double *y0, const double *vx,
                                         const double *vy, uint64_t n, int radius, const uint8_t *obs_valid,
                                         float *out_sum, float *out_mean, float *out_median, float *out_weighted) {
    KB_API_BEGIN
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0)
        throw Error("no CUDA device: kbmod_b200 has no CPU path");
    if (radius <= 0) throw Error("Invalid stamp radius " + std::to_string(radius));
    if (radius > 255) throw Error("stamp radius above 255 is not supported");
    if (T <= 0 || H <= 0 || W <= 0) throw Error("empty image stack");
    if (!sci || !times || (n && (!x0 || !y0 || !vx || !vy))) throw Error("null argument");
    if (out_weighted && !var) throw Error("the weighted coadd needs the variance images");
    if (n == 0) return 0;
    if (n > 0x7fffffffull) throw Error("too many trajectories");
    const int S = 2 * radius + 1, npix = S * S;
    const size_t img = (size_t)H * W;
    // the median keeps every thread's T values in shared memory: narrow the CTA until they fit
    int threads = kStampThreads;
    if (out_median)
        while (threads > 32 && (size_t)T * threads * 4 + 8 * (size_t)T > 200 * 1024) threads /= 2;
    const size_t smem = 8 * (size_t)T + (out_median ? (size_t)T * threads * 4 : 0);
    if (smem > 220 * 1024) throw Error("too many images for the median coadd (shared memory)");
    DBuf d_sci(img * T * 4), d_var(out_weighted ? img * T * 4 : 0), d_t(8 * (size_t)T), d_par(4 * 8 * n),
            d_ov(obs_valid ? n * T : 0);
    for (int t = 0; t < T; ++t) {
        KB_CUDA(cudaMemcpy(d_sci.as<float>() + img * t, sci[t], img * 4, cudaMemcpyHostToDevice));
        if (out_weighted) KB_CUDA(cudaMemcpy(d_var.as<float>() + img * t, var[t], img * 4, cudaMemcpyHostToDevice));
    }
    KB_CUDA(cudaMemcpy(d_t.p, times, 8 * (size_t)T, cudaMemcpyHostToDevice));
    double *par = d_par.as<double>();
    KB_CUDA(cudaMemcpy(par, x0, 8 * n, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(par + n, y0, 8 * n, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(par + 2 * n, vx, 8 * n, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(par + 3 * n, vy, 8 * n, cudaMemcpyHostToDevice));
    if (obs_valid) KB_CUDA(cudaMemcpy(d_ov.p, obs_valid, n * T, cudaMemcpyHostToDevice));
    const size_t ob = (size_t)n * npix * 4;
    DBuf d_sum(out_sum ? ob : 0), d_mean(out_mean ? ob : 0), d_med(out_median ? ob : 0), d_w(out_weighted ? ob : 0);
    CoaddOut o{out_sum ? d_sum.as<float>() : nullptr, out_mean ? d_mean.as<float>() : nullptr,
               out_median ? d_med.as<float>() : nullptr, out_weighted ? d_w.as<float>() : nullptr};
    KB_CUDA(cudaFuncSetAttribute(coadd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)n, (unsigned)((npix + threads - 1) / threads));
    coadd_kernel<<<grid, threads, smem>>>(d_sci.as<float>(), out_weighted ? d_var.as<float>() : nullptr, T, H, W,
                                          d_t.as<double>(), par, par + n, par + 2 * n, par + 3 * n, radius,
                                          obs_valid ? d_ov.as<uint8_t>() : nullptr, o, threads);
    KB_CUDA(cudaGetLastError());
    if (out_sum) KB_CUDA(cudaMemcpy(out_sum, d_sum.p, ob, cudaMemcpyDeviceToHost));
    if (out_mean) KB_CUDA(cudaMemcpy(out_mean, d_mean.p, ob, cudaMemcpyDeviceToHost));
    if (out_median) KB_CUDA(cudaMemcpy(out_median, d_med.p, ob, cudaMemcpyDeviceToHost));
    if (out_weighted) KB_CUDA(cudaMemcpy(out_weighted, d_w.p, ob, cudaMemcpyDeviceToHost));
    KB_API_END
}

extern "C" KB_EXPORT int kb_extract_stamps(const float *const *sci, int T, int H, int W, const double *times,
                                           const double *x0, const double *y0, const double *vx, const double *vy,
                                           uint64_t n, int radius, float *out) {
    KB_API_BEGIN
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0)
        throw Error("no CUDA device: kbmod_b200 has no CPU path");
    if (radius < 1) throw Error("Invalid stamp radius: " + std::to_string(radius));
    if (radius > 255) throw Error("stamp radius above 255 is not supported");
    if (T <= 0 || H <= 0 || W <= 0) throw Error("empty image stack");
    if (n == 0) return 0;
    if (!sci || !times || !x0 || !y0 || !vx || !vy || !out) throw Error("null argument");
    const int S = 2 * radius + 1;
    const size_t img = (size_t)H * W;
    const uint64_t total = n * (uint64_t)T * S * S;
    if (total >= (1ull << 36)) throw Error("too many stamp pixels for one call");
    DBuf d_sci(img * T * 4), d_t(8 * (size_t)T), d_par(4 * 8 * n), d_out(total * 4);
    for (int t = 0; t < T; ++t)
        KB_CUDA(cudaMemcpy(d_sci.as<float>() + img * t, sci[t], img * 4, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_t.p, times, 8 * (size_t)T, cudaMemcpyHostToDevice));
    double *par = d_par.as<double>();
    KB_CUDA(cudaMemcpy(par, x0, 8 * n, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(par + n, y0, 8 * n, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(par + 2 * n, vx, 8 * n, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(par + 3 * n, vy, 8 * n, cudaMemcpyHostToDevice));
    const unsigned grid = (unsigned)std::min<uint64_t>((total + 255) / 256, 148ull * 32);
    all_stamps_kernel<<<grid, 256>>>(d_sci.as<float>(), T, H, W, d_t.as<double>(), par, par + n, par + 2 * n, par + 3 * n,
                                     n, radius, d_out.as<float>());
    KB_CUDA(cudaGetLastError());
    KB_CUDA(cudaMemcpy(out, d_out.p, total * 4, cudaMemcpyDeviceToHost));
    KB_API_END
}

// extract_curve_values (core/stamp_utils.py:87-141 over _extract_curve_values, :477-512): values[r][t] =
// imgs[t][y[r][t]][x[r][t]] where that pixel exists, NaN elsewhere; float64 like the reference's np.full(..., nan).
namespace kb {
__global__ void curve_values_kernel(const float *__restrict__ img, int H, int W, const int64_t *__restrict__ x,
                                    const int64_t *__restrict__ y, uint64_t R, int T, int t, double *__restrict__ out) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const int64_t xi = x[r * T + t], yi = y[r * T + t];
    double v = CUDART_NAN;
    if (xi >= 0 && xi < W && yi >= 0 && yi < H) v = (double)img[(size_t)yi * W + (size_t)xi];
    out[r * T + t] = v;
}
}  // namespace kb

extern "C" KB_EXPORT int kb_extract_curve_values(const float *const *imgs, int T, int H, int W, const int64_t *x_vals,
                                                 const int64_t *y_vals, uint64_t n, double *out) {
    KB_API_BEGIN
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0)
        throw Error("no CUDA device: kbmod_b200 has no CPU path");
    if (T <= 0 || H <= 0 || W <= 0) throw Error("empty image stack");
    if (n == 0) return 0;
    if (!imgs || !x_vals || !y_vals || !out) throw Error("null argument");
    const size_t img = (size_t)H * W, cells = (size_t)n * T;
    DBuf d_img[2] = {DBuf(img * 4), DBuf(img * 4)}, d_x(cells * 8), d_y(cells * 8), d_out(cells * 8);
    KB_CUDA(cudaMemcpy(d_x.p, x_vals, cells * 8, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_y.p, y_vals, cells * 8, cudaMemcpyHostToDevice));
    for (int t = 0; t < T; ++t) {   // one image at a time: nothing but the current image has to be on the device
        KB_CUDA(cudaMemcpy(d_img[t & 1].p, imgs[t], img * 4, cudaMemcpyHostToDevice));
        kb::curve_values_kernel<<<(unsigned)((n + 255) / 256), 256>>>(d_img[t & 1].as<float>(), H, W, d_x.as<int64_t>(),
                                                                      d_y.as<int64_t>(), n, T, t, d_out.as<double>());
    }
    KB_CUDA(cudaGetLastError());
    KB_CUDA(cudaMemcpy(out, d_out.p, cells * 8, cudaMemcpyDeviceToHost));
    KB_API_END
}
