// curve_filters.cu -- the step that follows the search in the reference pipeline, on the
// device: sigma-G clipping of per-result likelihood curves and the likelihood update.
//
//   reference (Python, numpy + torch)                                         here
//   Results.compute_likelihood_curves      results.py:568-606                 lh_curve_value()
//   SigmaGClipping.compute_clipped_sigma_g_matrix  sigma_g_filter.py:114-168  sigmag_matrix_kernel (one warp per curve)
//   Results._update_likelihood             results.py:608-645                 update_likelihood_kernel
//
// Everything is float32 with IEEE operations in the reference's order, so the keep matrix is bit-identical:
//   * torch.nanquantile: sort the non-NaN values, rank = fl32(q * (n - 1)), value = lerp(v[floor], v[ceil], frac) with
//     torch's lerp (one fused multiply-add from the nearer end);
//   * bounds median -/+ fl32(n_sigma * coeff) * max(upper - lower, 1e-5), strict comparisons on the unmasked curve;
//   * the sums of _update_likelihood follow numpy's float32 pairwise summation (blocks of 128, 8 partial sums).
#include <math.h>

#include <math_constants.h>

#include <vector>

#include "common.cuh"

namespace kb {

static constexpr int kCurveWarps = 4;          // curves per CTA
static constexpr unsigned kMaxCurveT = 4096;   // 4 warps x 4096 floats = 64 KB of shared memory

__device__ __forceinline__ float lh_curve_value(float psi, float phi, bool obs_ok) {
    // results.py:600-605: valid = (phi != 0) & isfinite(psi) & isfinite(phi) [& obs_valid]; lh = psi / sqrt(phi)
    const bool valid = (phi != 0.0f) && isfinite(psi) && isfinite(phi) && obs_ok;
    return valid ? __fdiv_rn(psi, __fsqrt_rn(phi)) : KB_NAN_F;
}

__device__ __forceinline__ float torch_lerp(float a, float b, float w) {
    const float d = __fsub_rn(b, a);
    return (fabsf(w) < 0.5f) ? __fmaf_rn(w, d, a) : __fmaf_rn(__fsub_rn(w, 1.0f), d, b);
}

// One warp per curve.  `lh` (R x T) is either given, or built from psi/phi[/obs_valid] when lh == nullptr.
// valid_out (R x T bytes) receives the keep matrix.
__global__ void __launch_bounds__(kCurveWarps * 32)
        sigmag_matrix_kernel(const float *__restrict__ lh, const float *__restrict__ psi,
                             const float *__restrict__ phi, const uint8_t *__restrict__ obs_valid, uint64_t R,
                             unsigned T, float q_low, float q_high, float mult, int clip_negative,
                             uint8_t *__restrict__ valid_out) {
    extern __shared__ float s_all[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *v = s_all + (size_t)warp * T;       // masked curve
    __shared__ float s_stat[kCurveWarps][6];   // order statistics below/above for the three quantiles
    const unsigned FULL = 0xffffffffu;
    for (uint64_t r = (uint64_t)blockIdx.x * kCurveWarps + warp; r < R; r += (uint64_t)gridDim.x * kCurveWarps) {
        const size_t base = (size_t)r * T;
        int n = 0;
        for (unsigned t = lane; t < T; t += 32) {
            float x = lh ? lh[base + t] : lh_curve_value(psi[base + t], phi[base + t], obs_valid ? obs_valid[base + t] != 0 : true);
            if (clip_negative && !(x > 0.0f)) x = KB_NAN_F;   // torch.where(lh > 0, lh, nan)
            v[t] = x;
            n += (x == x) ? 1 : 0;
        }
        n = __reduce_add_sync(FULL, n);
        __syncwarp();
        float lo = KB_NAN_F, med = KB_NAN_F, hi = KB_NAN_F;
        if (n > 0) {
            // ranks (float32, like the q tensor times the integer count) and their floor / ceil positions
            const float nm1 = (float)(n - 1);
            float rk[3] = {__fmul_rn(q_low, nm1), __fmul_rn(0.5f, nm1), __fmul_rn(q_high, nm1)};
            int want[6];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                if (rk[i] < 0.0f) rk[i] = 0.0f;
                want[2 * i] = (int)floorf(rk[i]);
                want[2 * i + 1] = min((int)ceilf(rk[i]), n - 1);
            }
            // rank by counting: position of v[t] among the non-NaN values, ties by index (a stable sort)
            for (unsigned t = lane; t < T; t += 32) {
                const float x = v[t];
                if (x != x) continue;
                int pos = 0;
                for (unsigned j = 0; j < T; ++j) {
                    const float y = v[j];   // broadcast read
                    pos += (y < x || (y == x && j < t)) ? 1 : 0;
                }
#pragma unroll
                for (int i = 0; i < 6; ++i)
                    if (pos == want[i]) s_stat[warp][i] = x;
            }
            __syncwarp();
            lo = torch_lerp(s_stat[warp][0], s_stat[warp][1], __fsub_rn(rk[0], floorf(rk[0])));
            med = torch_lerp(s_stat[warp][2], s_stat[warp][3], __fsub_rn(rk[1], floorf(rk[1])));
            hi = torch_lerp(s_stat[warp][4], s_stat[warp][5], __fsub_rn(rk[2], floorf(rk[2])));
        }
        float delta = __fsub_rn(hi, lo);
        if (delta < 1e-5f) delta = 1e-5f;          // delta[delta < 1e-5] = 1e-5 (a NaN delta stays NaN)
        const float ns = __fmul_rn(mult, delta);
        const float lower = __fsub_rn(med, ns), upper = __fadd_rn(med, ns);
        for (unsigned t = lane; t < T; t += 32) {
            // the comparison uses the curve itself, not its clip_negative-masked copy (sigma_g_filter.py:165)
            const float x = lh ? lh[base + t] : lh_curve_value(psi[base + t], phi[base + t], obs_valid ? obs_valid[base + t] != 0 : true);
            valid_out[base + t] = (isfinite(x) && x < upper && x > lower) ? 1 : 0;
        }
        __syncwarp();
    }
}

// numpy's float32 pairwise sum of one contiguous row of (curve * obs_valid) (loops_utils.h.src, pairwise_sum):
// n < 8: sequential from -0.0; n <= 128: eight partial sums, combined as a tree, tail added in order; else split at
// n/2 rounded down to a multiple of 8.
__device__ float row_term(const float *c, const uint8_t *m, unsigned i) {
    return m ? __fmul_rn(c[i], m[i] ? 1.0f : 0.0f) : c[i];
}
__device__ __forceinline__ float pairwise_leaf(const float *c, const uint8_t *m, unsigned lo, unsigned n) {
    if (n < 8) {
        float res = -0.0f;
        for (unsigned i = 0; i < n; ++i) res = __fadd_rn(res, row_term(c, m, lo + i));
        return res;
    }
    float r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = row_term(c, m, lo + k);
    unsigned i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = __fadd_rn(r[k], row_term(c, m, lo + i + k));
    }
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __fadd_rn(res, row_term(c, m, lo + i));
    return res;
}
// The recursion depth is bounded by the row length (kMaxCurveT = 4096 -> 5 splits), so it is unrolled at compile
// time and the call graph stays static.
template <int DEPTH>
__device__ __noinline__ float pairwise_sum_row(const float *c, const uint8_t *m, unsigned lo, unsigned n) {
    if (n <= 128) return pairwise_leaf(c, m, lo, n);
    unsigned n2 = n / 2;
    n2 -= n2 % 8;
    if constexpr (DEPTH > 0) {
        return __fadd_rn(pairwise_sum_row<DEPTH - 1>(c, m, lo, n2), pairwise_sum_row<DEPTH - 1>(c, m, lo + n2, n - n2));
    } else {
        return KB_NAN_F;  // unreachable for n <= kMaxCurveT
    }
}

__global__ void update_likelihood_kernel(const float *__restrict__ psi, const float *__restrict__ phi,
                                         const uint8_t *__restrict__ obs_valid, uint64_t R, unsigned T,
                                         double *__restrict__ lh_out, double *__restrict__ flux_out,
                                         long long *__restrict__ obs_out) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float *ps = psi + (size_t)r * T, *ph = phi + (size_t)r * T;
    const uint8_t *m = obs_valid ? obs_valid + (size_t)r * T : nullptr;
    const float psi_sum = pairwise_sum_row<6>(ps, m, 0, T), phi_sum = pairwise_sum_row<6>(ph, m, 0, T);
    long long n = T;
    if (m) {
        n = 0;
        for (unsigned t = 0; t < T; ++t) n += m[t] ? 1 : 0;
    }
    const bool nz = phi_sum != 0.0f;   // results.py:640 (a NaN sum is "non zero" there too)
    lh_out[r] = nz ? (double)__fdiv_rn(psi_sum, __fsqrt_rn(phi_sum)) : 0.0;
    flux_out[r] = nz ? (double)__fdiv_rn(psi_sum, phi_sum) : 0.0;
    obs_out[r] = n;
}

// ---- host side -----------------------------------------------------------------------------------------------
namespace {
struct DevBuf {
    void *p = nullptr;
    explicit DevBuf(size_t bytes) { KB_CUDA(cudaMalloc(&p, bytes ? bytes : 1)); }
    ~DevBuf() { cudaFree(p); }
    template <class T> T *as() { return static_cast<T *>(p); }
};

void check_shape(uint64_t R, unsigned T) {
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0)
        throw Error("no CUDA device: kbmod_b200 has no CPU path");
    if (T == 0) throw Error("curves need at least one time step");
    if (T > kMaxCurveT) throw Error("curves longer than " + std::to_string(kMaxCurveT) + " time steps are not supported");
    if (R * (uint64_t)T >= (1ull << 40)) throw Error("curve matrix too large");
}

// One curve, one CTA: rank the selected values by counting (ties in index order), scatter them into sorted order,
// then thread 0 takes numpy's three percentiles and every thread tests its own points.
__device__ __forceinline__ double np_lerp(double a, double b, double t) {   // numpy/lib/function_base.py _lerp
    const double d = __dsub_rn(b, a);
    return (t >= 0.5) ? __dsub_rn(b, __dmul_rn(d, __dsub_rn(1.0, t))) : __dadd_rn(a, __dmul_rn(d, t));
}
__device__ __forceinline__ double np_percentile_sorted(const double *v, unsigned m, double pct) {
    const double q = pct / 100.0;
    const double virt = __dmul_rn((double)(m - 1), q);
    if (virt >= (double)(m - 1)) return v[m - 1];
    const double fl = floor(virt);
    const unsigned lo = (unsigned)fl;
    return np_lerp(v[lo], v[lo + 1], __dsub_rn(virt, fl));
}
__global__ void __launch_bounds__(1024)
        sigmag_curve_kernel(const double *__restrict__ lh, unsigned n, double low_pct, double high_pct, double n_sigma,
                            double coeff, int clip_negative, double *__restrict__ sorted, uint8_t *__restrict__ flags) {
    __shared__ unsigned s_m;
    __shared__ int s_nan;
    __shared__ double s_lo, s_hi;
    if (threadIdx.x == 0) {
        s_m = 0;
        s_nan = 0;
    }
    __syncthreads();
    // selected = all points, or the positive ones (clip_negative); rank by counting among the selected
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
        const double vi = lh[i];
        const bool sel = clip_negative ? (vi > 0.0) : true;
        if (!sel) continue;
        if (vi != vi) s_nan = 1;   // numpy's percentile of data with a NaN is NaN: nothing passes
        unsigned rank = 0;
        for (unsigned j = 0; j < n; ++j) {
            const double vj = lh[j];
            const bool selj = clip_negative ? (vj > 0.0) : true;
            if (selj && (vj < vi || (vj == vi && j < i))) ++rank;
        }
        sorted[rank] = vi;
        atomicAdd(&s_m, 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned m = s_m;
        if (m == 0 || s_nan) {
            s_lo = CUDART_NAN;
            s_hi = CUDART_NAN;
        } else {
            const double lower = np_percentile_sorted(sorted, m, low_pct);
            const double median = np_percentile_sorted(sorted, m, 50.0);
            const double upper = np_percentile_sorted(sorted, m, high_pct);
            const double delta = fmax(__dsub_rn(upper, lower), 1e-8);
            const double ns = __dmul_rn(n_sigma, __dmul_rn(coeff, delta));
            s_lo = __dsub_rn(median, ns);
            s_hi = __dadd_rn(median, ns);
        }
    }
    __syncthreads();
    const double lo = s_lo, hi = s_hi;
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) flags[i] = (lh[i] > lo && lh[i] < hi) ? 1 : 0;
}

void launch_clip(const float *d_lh, const float *d_psi, const float *d_phi, const uint8_t *d_obs, uint64_t R,
                 unsigned T, float q_low, float q_high, float mult, int clip_negative, uint8_t *d_valid) {
    const size_t smem = sizeof(float) * (size_t)T * kCurveWarps;
    KB_CUDA(cudaFuncSetAttribute(sigmag_matrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const uint64_t ctas = (R + kCurveWarps - 1) / kCurveWarps;
    const unsigned grid = (unsigned)std::min<uint64_t>(std::max<uint64_t>(ctas, 1), 148ull * 16);
    sigmag_matrix_kernel<<<grid, kCurveWarps * 32, smem>>>(d_lh, d_psi, d_phi, d_obs, R, T, q_low, q_high, mult,
                                                           clip_negative, d_valid);
    KB_CUDA(cudaGetLastError());
}
}  // namespace
}  // namespace kb

using namespace kb;

extern "C" {

KB_EXPORT int kb_sigmag_clip_matrix(const float *lh, uint64_t n_rows, uint32_t n_times, float q_low, float q_high,
                                    float n_sigma_coeff, int clip_negative, uint8_t *valid_out) {
    KB_API_BEGIN
    if (n_rows == 0) return 0;
    if (!lh || !valid_out) throw Error("null argument");
    check_shape(n_rows, n_times);
    const size_t n = (size_t)n_rows * n_times;
    DevBuf d_lh(n * 4), d_valid(n);
    KB_CUDA(cudaMemcpy(d_lh.p, lh, n * 4, cudaMemcpyHostToDevice));
    launch_clip(d_lh.as<float>(), nullptr, nullptr, nullptr, n_rows, n_times, q_low, q_high, n_sigma_coeff,
                clip_negative, d_valid.as<uint8_t>());
    KB_CUDA(cudaMemcpy(valid_out, d_valid.p, n, cudaMemcpyDeviceToHost));
    KB_API_END
}

KB_EXPORT int kb_update_likelihood(const float *psi, const float *phi, const uint8_t *obs_valid, uint64_t n_rows,
                                   uint32_t n_times, double *lh_out, double *flux_out, int64_t *obs_count_out) {
    KB_API_BEGIN
    if (n_rows == 0) return 0;
    if (!psi || !phi || !lh_out || !flux_out || !obs_count_out) throw Error("null argument");
    check_shape(n_rows, n_times);
    const size_t n = (size_t)n_rows * n_times;
    DevBuf d_psi(n * 4), d_phi(n * 4), d_obs(obs_valid ? n : 0), d_lh(n_rows * 8), d_flux(n_rows * 8), d_cnt(n_rows * 8);
    KB_CUDA(cudaMemcpy(d_psi.p, psi, n * 4, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_phi.p, phi, n * 4, cudaMemcpyHostToDevice));
    if (obs_valid) KB_CUDA(cudaMemcpy(d_obs.p, obs_valid, n, cudaMemcpyHostToDevice));
    update_likelihood_kernel<<<(unsigned)((n_rows + 127) / 128), 128>>>(
            d_psi.as<float>(), d_phi.as<float>(), obs_valid ? d_obs.as<uint8_t>() : nullptr, n_rows, n_times,
            d_lh.as<double>(), d_flux.as<double>(), d_cnt.as<long long>());
    KB_CUDA(cudaGetLastError());
    KB_CUDA(cudaMemcpy(lh_out, d_lh.p, n_rows * 8, cudaMemcpyDeviceToHost));
    KB_CUDA(cudaMemcpy(flux_out, d_flux.p, n_rows * 8, cudaMemcpyDeviceToHost));
    KB_CUDA(cudaMemcpy(obs_count_out, d_cnt.p, n_rows * 8, cudaMemcpyDeviceToHost));
    KB_API_END
}

KB_EXPORT int kb_sigmag_filter_curves(const float *psi, const float *phi, uint64_t n_rows, uint32_t n_times,
                                      float q_low, float q_high, float n_sigma_coeff, int clip_negative,
                                      uint8_t *obs_valid_io, int has_obs_valid, double *lh_out, double *flux_out,
                                      int64_t *obs_count_out) {
    KB_API_BEGIN
    if (n_rows == 0) return 0;
    if (!psi || !phi || !obs_valid_io || !lh_out || !flux_out || !obs_count_out) throw Error("null argument");
    check_shape(n_rows, n_times);
    const size_t n = (size_t)n_rows * n_times;
    DevBuf d_psi(n * 4), d_phi(n * 4), d_obs(n), d_new(n), d_lh(n_rows * 8), d_flux(n_rows * 8), d_cnt(n_rows * 8);
    KB_CUDA(cudaMemcpy(d_psi.p, psi, n * 4, cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_phi.p, phi, n * 4, cudaMemcpyHostToDevice));
    if (has_obs_valid) KB_CUDA(cudaMemcpy(d_obs.p, obs_valid_io, n, cudaMemcpyHostToDevice));
    // apply_clipped_sigma_g (sigma_g_filter.py:171-190): curves with filter_obs -> clip -> update_obs_valid (which
    // replaces the mask and recomputes likelihood, flux and obs_count from it)
    launch_clip(nullptr, d_psi.as<float>(), d_phi.as<float>(), has_obs_valid ? d_obs.as<uint8_t>() : nullptr, n_rows,
                n_times, q_low, q_high, n_sigma_coeff, clip_negative, d_new.as<uint8_t>());
    update_likelihood_kernel<<<(unsigned)((n_rows + 127) / 128), 128>>>(
            d_psi.as<float>(), d_phi.as<float>(), d_new.as<uint8_t>(), n_rows, n_times, d_lh.as<double>(),
            d_flux.as<double>(), d_cnt.as<long long>());
    KB_CUDA(cudaGetLastError());
    KB_CUDA(cudaMemcpy(obs_valid_io, d_new.p, n, cudaMemcpyDeviceToHost));
    KB_CUDA(cudaMemcpy(lh_out, d_lh.p, n_rows * 8, cudaMemcpyDeviceToHost));
    KB_CUDA(cudaMemcpy(flux_out, d_flux.p, n_rows * 8, cudaMemcpyDeviceToHost));
    KB_CUDA(cudaMemcpy(obs_count_out, d_cnt.p, n_rows * 8, cudaMemcpyDeviceToHost));
    KB_API_END
}


// SigmaGClipping.compute_clipped_sigma_g (filters/sigma_g_filter.py:86-112): ONE likelihood curve, float64, numpy's
// percentile (method "linear": virtual index q/100 * (n - 1), value = lerp of the two neighbours, taken from the
// nearer end exactly as numpy's _lerp does), delta floored at 1e-8, strict bounds on the whole curve.  good_idx (room
// for n entries) receives the indices kept, in ascending order.
KB_EXPORT int kb_sigmag_clip_curve(const double *lh, uint32_t n, double low_pct, double high_pct, double n_sigma,
                                   double coeff, int clip_negative, int64_t *good_idx, uint32_t *n_good) {
    KB_API_BEGIN
    if (!n_good) throw Error("null argument");
    *n_good = 0;
    if (n == 0) return 0;
    if (!lh || !good_idx) throw Error("null argument");
    if (n > 65536) throw Error("likelihood curve too long (at most 65536 points)");
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0)
        throw Error("no CUDA device: kbmod_b200 has no CPU path");
    DevBuf d_lh((size_t)n * 8), d_sorted((size_t)n * 8), d_flags(n), d_cnt(4);
    KB_CUDA(cudaMemcpy(d_lh.p, lh, (size_t)n * 8, cudaMemcpyHostToDevice));
    sigmag_curve_kernel<<<1, 1024>>>(d_lh.as<double>(), n, low_pct, high_pct, n_sigma, coeff, clip_negative,
                                     d_sorted.as<double>(), d_flags.as<uint8_t>());
    KB_CUDA(cudaGetLastError());
    std::vector<uint8_t> flags(n);
    KB_CUDA(cudaMemcpy(flags.data(), d_flags.p, n, cudaMemcpyDeviceToHost));
    uint32_t k = 0;
    for (uint32_t i = 0; i < n; ++i)
        if (flags[i]) good_idx[k++] = (int64_t)i;
    *n_good = k;
    KB_API_END
}

}  // extern "C"
