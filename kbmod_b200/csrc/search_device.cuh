// search_device.cuh -- device routines shared by the generic and the tiled search kernels.
//
// Everything here follows the reference's GPU-path semantics (kernels/kernels.cu) and is written so the
// results are bit-identical to the host-compiled reference: fp64 index arithmetic without contraction,
// IEEE float division / square root, sums taken in the same order.
#pragma once
#include <float.h>
#include <math_constants.h>

#include "common.cuh"

namespace kb {

// kernels.cu:33-35 predict_index: (int)floor(pos0 + vel0*time + 0.5f), evaluated in fp64.
// The host build of the reference does not contract the multiply-add; neither do we.
__device__ __forceinline__ int predict_index(int pos0, float vel0, double time) {
    const double prod = __dmul_rn((double)vel0, time);
    const double s = __dadd_rn((double)(float)pos0, prod);
    return __double2int_rd(__dadd_rn(s, 0.5));
}

// kernels.cu:58-68 decode of one stored value: code 0 -> NaN, else (code-1)*scale+min in fp64 -> f32.
__device__ __forceinline__ float decode_code(unsigned code, float scale, float mn) {
    if (code == 0u) return KB_NAN_F;
    const double v = __dadd_rn(__dmul_rn((double)code - 1.0, (double)scale), (double)mn);
    return (float)v;
}

// kernels.cu:37-71 read_encoded_psi_phi on the pitched search copy. Returns true when both values are finite.
template <int NB>
__device__ __forceinline__ bool load_pair(const StackView &sv, int t, int iy, int ix, float &psi, float &phi) {
    if ((unsigned)ix >= (unsigned)sv.W || (unsigned)iy >= (unsigned)sv.H) return false;
    const size_t idx = ((size_t)t * sv.H + iy) * sv.Wp + ix;
    if (NB == 4) {
        const float2 v = __ldg(reinterpret_cast<const float2 *>(sv.data) + idx);
        psi = v.x;
        phi = v.y;
    } else if (NB == 2) {
        const ushort2 v = __ldg(reinterpret_cast<const ushort2 *>(sv.data) + idx);
        psi = decode_code(v.x, sv.psi_scale, sv.psi_min);
        phi = decode_code(v.y, sv.phi_scale, sv.phi_min);
    } else {
        const uchar2 v = __ldg(reinterpret_cast<const uchar2 *>(sv.data) + idx);
        psi = decode_code(v.x, sv.psi_scale, sv.psi_min);
        phi = decode_code(v.y, sv.phi_scale, sv.phi_min);
    }
    return isfinite(psi) && isfinite(phi);
}

// kernels.cu:195-197: lh = psi_sum / sqrt(phi_sum), flux = psi_sum / phi_sum, -1 when phi_sum <= 0.
__device__ __forceinline__ void finish_sums(float psi_sum, float phi_sum, float &lh, float &flux) {
    if (phi_sum > 0.0f) {
        lh = __fdiv_rn(psi_sum, __fsqrt_rn(phi_sum));
        flux = __fdiv_rn(psi_sum, phi_sum);
    } else {
        lh = -1.0f;
        flux = -1.0f;
    }
}

// kernels.cu:77-147 SigmaGFilteredIndicesCU: sort an index array by value with the same exchange sort (so equal
// values end in the same order), pick the three percentile ranks, return the kept range of sorted positions.
__device__ inline void sigmag_indices(const float *values, int n, float sgl0, float sgl1, float coeff, float width,
                                      int *idx, int &min_keep, int &max_keep) {
    if (n == 0) {
        min_keep = 0;
        max_keep = -1;
        return;
    }
    if ((double)sgl0 < 0.0001) sgl0 = (float)0.0001;
    if ((double)sgl1 > 0.9999) sgl1 = (float)0.9999;
    for (int j = 0; j < n; ++j) idx[j] = j;
    for (int j = 0; j < n; ++j) {
        int ij = idx[j];
        float vj = values[ij];
        for (int k = j + 1; k < n; ++k) {
            const int ik = idx[k];
            const float vk = values[ik];
            if (vj > vk) {
                idx[k] = ij;
                ij = ik;
                vj = vk;
            }
        }
        idx[j] = ij;
    }
    int pL = (int)((double)ceilf(__fmul_rn((float)n, sgl0)) + 0.001) - 1;
    pL = pL < 0 ? 0 : pL;
    pL = pL >= n ? n - 1 : pL;
    int pH = (int)((double)ceilf(__fmul_rn((float)n, sgl1)) + 0.001) - 1;
    pH = pH < 0 ? 0 : pH;
    pH = pH >= n ? n - 1 : pH;
    int med = (int)(ceil((double)n * 0.5) + 0.001) - 1;
    med = med < 0 ? 0 : med;
    med = med >= n ? n - 1 : med;
    const float sigma_g = __fmul_rn(coeff, __fsub_rn(values[idx[pH]], values[idx[pL]]));
    const float w = __fmul_rn(width, sigma_g);
    const float lo = __fsub_rn(values[idx[med]], w);
    const float hi = __fadd_rn(values[idx[med]], w);
    int start = 0;
    while (start < med && values[idx[start]] < lo) ++start;
    min_keep = start;
    int end = med + 1;
    while (end < n && values[idx[end]] <= hi) ++end;
    max_keep = end - 1;
}

// The sigma-G branch of evaluateTrajectory (kernels.cu:213-241) for one (start pixel, candidate): gather the
// valid observations again, clip, and re-sum the kept ones in ascending-lc order.  Rarely taken (only when the
// unclipped lh already passes min_lh), so it lives out of line with its arrays in local memory.
template <int NB, int MAXT>
__device__ __noinline__ void sigmag_recompute(const StackView &sv, const double *times, int x, int y, float vx,
                                              float vy, float sgl_L, float sgl_H, float coeff, float &lh,
                                              float &flux, uint8_t *mask) {
    float psi_a[MAXT], phi_a[MAXT], lc[MAXT];
    int idx[MAXT];
    int seen = 0;
    for (int t = 0; t < sv.T; ++t) {
        const double tm = times[t];
        float a, b;
        if (load_pair<NB>(sv, t, predict_index(y, vy, tm), predict_index(x, vx, tm), a, b)) {
            if (seen < MAXT) {
                psi_a[seen] = a;
                phi_a[seen] = b;
                lc[seen] = (b != 0.0f) ? __fdiv_rn(a, b) : 0.0f;
                if (mask) idx[seen] = t;  // borrow idx[] for the time index until the sort
            }
            ++seen;
        }
    }
    // Remember the time index of each observation when a keep-mask is wanted.
    uint16_t *tix = nullptr;
    uint16_t tix_store[MAXT];
    if (mask) {
        for (int i = 0; i < seen; ++i) tix_store[i] = (uint16_t)idx[i];
        tix = tix_store;
        for (int t = 0; t < sv.T; ++t) mask[t] = 0;
    }
    int lo = 0, hi = seen - 1;
    sigmag_indices(lc, seen, sgl_L, sgl_H, coeff, 2.0f, idx, lo, hi);
    if (lo < 0) lo = 0;
    if (hi >= seen) hi = seen - 1;
    float ps = 0.0f, ph = 0.0f;
    for (int i = lo; i <= hi; ++i) {
        ps = __fadd_rn(ps, psi_a[idx[i]]);
        ph = __fadd_rn(ph, phi_a[idx[i]]);
        if (tix) mask[tix[idx[i]]] = 1;
    }
    finish_sums(ps, ph, lh, flux);
}

// evaluateTrajectory (kernels.cu:154-242), complete: used by the epilogue of the tiled kernel, by the
// single-trajectory entry point and by the generic kernel.  `times` may point to shared or global memory.
template <int NB, bool SIGMAG, int MAXT>
__device__ __forceinline__ void evaluate_exact(const StackView &sv, const double *times, const kb_params &p, int x,
                                               int y, float vx, float vy, float &lh, float &flux, int &obs,
                                               uint8_t *mask) {
    float psi_sum = 0.0f, phi_sum = 0.0f;
    int seen = 0;
    for (int t = 0; t < sv.T; ++t) {
        const double tm = times[t];
        float a, b;
        const bool ok = load_pair<NB>(sv, t, predict_index(y, vy, tm), predict_index(x, vx, tm), a, b);
        if (ok) {
            psi_sum = __fadd_rn(psi_sum, a);
            phi_sum = __fadd_rn(phi_sum, b);
            ++seen;
        }
        if (mask) mask[t] = ok ? 1 : 0;
    }
    obs = seen;
    finish_sums(psi_sum, phi_sum, lh, flux);
    if (SIGMAG) {
        if (!(seen < p.min_observations || seen == 0 || lh < p.min_lh))
            sigmag_recompute<NB, MAXT>(sv, times, x, y, vx, vy, p.sgl_L, p.sgl_H, p.sigmag_coeff, lh, flux, mask);
    }
}

// Ordering used for the per-pixel list when candidates are not visited in list order: higher lh first; equal lh
// -> the candidate that comes first in the caller's list (what the reference's strict '>' insertion in list
// order produces, kernels.cu:324-330).  Empty slots carry cand = -1 and therefore win ties, as they do in the
// reference where they are present before any candidate is visited.  NaN never beats anything.
__device__ __forceinline__ bool beats(float lh_a, int cand_a, float lh_b, int cand_b) {
    return lh_a > lh_b || (lh_a == lh_b && cand_a < cand_b);
}

template <int K>
struct TopK {
    float lh[K];
    int cand[K];
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int r = 0; r < K; ++r) {
            lh[r] = -FLT_MAX;
            cand[r] = -1;
        }
    }
    __device__ __forceinline__ float worst() const { return lh[K - 1]; }
    __device__ __forceinline__ void insert(float nlh, int ncand) {
#pragma unroll
        for (int r = 0; r < K; ++r) {
            if (beats(nlh, ncand, lh[r], cand[r])) {
                const float tl = lh[r];
                const int tc = cand[r];
                lh[r] = nlh;
                cand[r] = ncand;
                nlh = tl;
                ncand = tc;
            }
        }
    }
};

}  // namespace kb
