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
        if (sv.zero_marker) return ((__float_as_uint(v.x) | __float_as_uint(v.y)) << 1) != 0u;  // (0,0) == invalid
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
                                              float &flux, uint8_t *mask, int *n_valid = nullptr) {
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
    if (n_valid) *n_valid = seen;
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

// Warp-cooperative form of the same sigma-G branch: the 32 lanes of a warp work on ONE (start pixel, candidate)
// whose parameters are warp-uniform.  Lane l holds observations t = l, l+32, ... (E per lane), so the re-gather
// is one parallel round of loads; the order statistics come from rank-by-counting over warp shuffles (no sort,
// no local memory); the kept values are then summed in ascending-lc order exactly like the reference does.
// The reference's exchange sort leaves equal lc values in a history-dependent order; that order only matters
// when two tied observations carry different (psi, phi).  In that (rare) case *needs_serial is set and the
// caller falls back to sigmag_recompute for bit-exactness.  Result is warp-uniform.  T <= 32*E.
// What comes back is exactly what evaluateTrajectory leaves in the trajectory (kernels.cu:154-242): the unclipped
// lh / flux (sums in time order; -1 when phi_sum <= 0) if there are too few observations or that lh is below min_lh
// (kernels.cu:201-203: the clip is skipped), the clipped values otherwise.
// The re-gather: lane l takes the observations t = l, l + 32, ...  Returns the number of valid ones.
template <int NB, int E>
__device__ __forceinline__ int warp_sigmag_load(const StackView &sv, const double *times, int x, int y, float vx,
                                                float vy, float (&a)[E], float (&b)[E], float (&lc)[E], bool (&ok)[E],
                                                unsigned (&okm)[E]) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int n = 0;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const int t = e * 32 + lane;
        a[e] = 0.0f;
        b[e] = 0.0f;
        ok[e] = false;
        if (t < sv.T) {
            const double tm = times[t];
            ok[e] = load_pair<NB>(sv, t, predict_index(y, vy, tm), predict_index(x, vx, tm), a[e], b[e]);
        }
        lc[e] = (b[e] != 0.0f) ? __fdiv_rn(a[e], b[e]) : 0.0f;
        okm[e] = __ballot_sync(FULL, ok[e]);
        n += __popc(okm[e]);
    }
    return n;
}

// Everything after the re-gather (see above).
template <int NB, int E>
__device__ __forceinline__ void warp_sigmag_eval(const StackView &sv, const float (&a)[E], const float (&b)[E],
                                                 const float (&lc)[E], const bool (&ok)[E], const unsigned (&okm)[E],
                                                 int n, float sgl_L, float sgl_H, float coeff, int min_obs,
                                                 float min_lh, float &lh, float &flux, int &n_valid,
                                                 bool &needs_serial, float2 *ws = nullptr) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    needs_serial = false;
    n_valid = n;
    {   // kernels.cu:175-197: the unclipped sums, valid observations in time order (every lane runs the same chain)
        float ups = 0.0f, uph = 0.0f;
        if (ws) {
            // through the warp's shared-memory scratch: every observation at its time index, invalid ones as (+0, +0)
            // -- adding +0 to a sum that started at +0 is an exact no-op -- then one broadcast read per image
#pragma unroll
            for (int e = 0; e < E; ++e) {
                const int t = e * 32 + lane;
                if (t < sv.T) ws[t] = ok[e] ? make_float2(a[e], b[e]) : make_float2(0.0f, 0.0f);
            }
            __syncwarp();
            for (int t = 0; t < sv.T; ++t) {
                const float2 v = ws[t];
                ups = __fadd_rn(ups, v.x);
                uph = __fadd_rn(uph, v.y);
            }
            __syncwarp();   // ws is written again below
        } else
#pragma unroll
        for (int es = 0; es < E; ++es) {
            const int nsrc = min(32, sv.T - es * 32);  // warp-uniform
            for (int ls = 0; ls < nsrc; ++ls) {
                const float va = __shfl_sync(FULL, a[es], ls), vb = __shfl_sync(FULL, b[es], ls);
                if ((okm[es] >> ls) & 1u) {
                    ups = __fadd_rn(ups, va);
                    uph = __fadd_rn(uph, vb);
                }
            }
        }
        finish_sums(ups, uph, lh, flux);
    }
    // kernels.cu:201-203: too few observations (or none), or an unclipped lh below min_lh -> the clip is skipped
    if (n < min_obs || n == 0 || lh < min_lh) return;
    // rank[e] = position of my element in ascending order (ties: earlier image first).
    // Fast pass: count the strictly smaller values (invalid observations travel as NaN, which compares false).
    // Without equal values -- the usual case -- that count is the rank, and the ranks of the n valid observations
    // are 0 .. n-1, which add up to n (n - 1) / 2; equal values share a (smaller) rank, so any tie lowers the sum.
    int rank[E];
    bool has_tie = false;
    {
        float lcs[E];
#pragma unroll
        for (int e = 0; e < E; ++e) {
            lcs[e] = ok[e] ? lc[e] : KB_NAN_F;
            rank[e] = 0;
        }
#pragma unroll
        for (int es = 0; es < E; ++es) {
            const int nsrc = min(32, sv.T - es * 32);  // warp-uniform
#pragma unroll 4
            for (int ls = 0; ls < nsrc; ++ls) {
                const float v = __shfl_sync(FULL, lcs[es], ls);
#pragma unroll
                for (int e = 0; e < E; ++e) rank[e] += (v < lcs[e]) ? 1 : 0;
            }
        }
        int rsum = 0;
#pragma unroll
        for (int e = 0; e < E; ++e) rsum += ok[e] ? rank[e] : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) rsum += __shfl_xor_sync(FULL, rsum, d);
        has_tie = rsum != n * (n - 1) / 2;
    }
    if (__any_sync(FULL, has_tie)) {
        // Slow pass: ranks with the tie rule, then the payload check below.
#pragma unroll
        for (int e = 0; e < E; ++e) rank[e] = 0;
#pragma unroll
        for (int es = 0; es < E; ++es) {
            unsigned m = okm[es];
            while (m) {
                const int ls = __ffs(m) - 1;
                m &= m - 1;
                const float v = __shfl_sync(FULL, lc[es], ls);
                const int ts = es * 32 + ls;
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const int t = e * 32 + lane;
                    if (v < lc[e] || (v == lc[e] && ts < t)) rank[e] += 1;
                }
            }
        }
        // Equal lc values: harmless when the tied observations carry identical (psi, phi); otherwise the
        // reference's tie order matters for the rounding of the sums and the serial routine must decide.
        bool tie_bad = false;
#pragma unroll
        for (int es = 0; es < E; ++es) {
            unsigned m = okm[es];
            while (m) {
                const int ls = __ffs(m) - 1;
                m &= m - 1;
                const float v = __shfl_sync(FULL, lc[es], ls);
                const float va = __shfl_sync(FULL, a[es], ls);
                const float vb = __shfl_sync(FULL, b[es], ls);
                const int ts = es * 32 + ls;
#pragma unroll
                for (int e = 0; e < E; ++e)
                    if (ok[e] && v == lc[e] && ts != e * 32 + lane && (va != a[e] || vb != b[e])) tie_bad = true;
            }
        }
        if (__any_sync(FULL, tie_bad)) {
            needs_serial = true;
            return;
        }
    }
    // kernels.cu:94-127 percentile ranks
    if ((double)sgl_L < 0.0001) sgl_L = (float)0.0001;
    if ((double)sgl_H > 0.9999) sgl_H = (float)0.9999;
    int pL = (int)((double)ceilf(__fmul_rn((float)n, sgl_L)) + 0.001) - 1;
    pL = pL < 0 ? 0 : pL;
    pL = pL >= n ? n - 1 : pL;
    int pH = (int)((double)ceilf(__fmul_rn((float)n, sgl_H)) + 0.001) - 1;
    pH = pH < 0 ? 0 : pH;
    pH = pH >= n ? n - 1 : pH;
    int med = (int)(ceil((double)n * 0.5) + 0.001) - 1;
    med = med < 0 ? 0 : med;
    med = med >= n ? n - 1 : med;
    float vL = 0.0f, vH = 0.0f, vM = 0.0f;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        unsigned m = __ballot_sync(FULL, ok[e] && rank[e] == pL);
        if (m) vL = __shfl_sync(FULL, lc[e], __ffs(m) - 1);
        m = __ballot_sync(FULL, ok[e] && rank[e] == pH);
        if (m) vH = __shfl_sync(FULL, lc[e], __ffs(m) - 1);
        m = __ballot_sync(FULL, ok[e] && rank[e] == med);
        if (m) vM = __shfl_sync(FULL, lc[e], __ffs(m) - 1);
    }
    const float sigma_g = __fmul_rn(coeff, __fsub_rn(vH, vL));
    const float w = __fmul_rn(2.0f, sigma_g);
    const float lo = __fsub_rn(vM, w), hi = __fadd_rn(vM, w);
    // kernels.cu:134-146: first sorted position (below the median) not < lo; last (above the median) still <= hi
    int n_lo = 0, n_hi = 0;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        n_lo += __popc(__ballot_sync(FULL, ok[e] && rank[e] < med && lc[e] < lo));
        n_hi += __popc(__ballot_sync(FULL, ok[e] && rank[e] > med && lc[e] <= hi));
    }
    const int first = n_lo, last = med + n_hi;
    float ps = 0.0f, ph = 0.0f;
    if (ws) {
        // `ws`: 32*E pairs of shared memory owned by this warp.  Scatter every observation to its sorted position,
        // then add the kept range in ascending-lc order (every lane runs the same chain on broadcast reads).
#pragma unroll
        for (int e = 0; e < E; ++e)
            if (ok[e]) ws[rank[e]] = make_float2(a[e], b[e]);
        __syncwarp();
        for (int r = first; r <= last; ++r) {
            const float2 v = ws[r];
            ps = __fadd_rn(ps, v.x);
            ph = __fadd_rn(ph, v.y);
        }
        __syncwarp();  // the next call overwrites ws
        finish_sums(ps, ph, lh, flux);
        return;
    }
    for (int r = first; r <= last; ++r) {
#pragma unroll
        for (int e = 0; e < E; ++e) {
            const unsigned m = __ballot_sync(FULL, ok[e] && rank[e] == r);
            if (m) {
                const int src = __ffs(m) - 1;
                ps = __fadd_rn(ps, __shfl_sync(FULL, a[e], src));
                ph = __fadd_rn(ph, __shfl_sync(FULL, b[e], src));
            }
        }
    }
    finish_sums(ps, ph, lh, flux);
}

// ---- a cheap, certain "no" for the deferred clips -------------------------------------------------------------------
// Of the (start pixel, candidate) pairs whose screening bound reaches min_lh only about one in a hundred survives the
// clip (a trajectory that crosses one bright source has a high unclipped lh and a noise-level clipped one).  Whether
// a pair is rejected does not need the reference's ordered float sums, only bounds of them:
//   * the order statistics and the kept set are functions of the lc values alone (exact here);
//   * two float sums of the same n numbers taken in different orders differ by at most 2 gamma_(n-1) sum|v_i|
//     (gamma_k = k u / (1 - k u), u = 2^-24), so a sum taken by a warp reduction brackets the reference's.
// warp_sigmag_rejects returns true only when the reference certainly rejects the pair (too few observations, or the
// unclipped lh below min_lh, or the clipped lh below min_lh: kernels.cu:201-203,318-320); false means "evaluate".
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
}
// The largest of the lanes' values (finite floats, or "none" = -FLT_MAX..): one REDUX on order-preserving integers.
__device__ __forceinline__ float warp_max_finite(float v, bool have) {
    const unsigned bits = __float_as_uint(v);
    const unsigned key = have ? (bits ^ ((bits >> 31) ? 0xffffffffu : 0x80000000u)) : 0u;
    const unsigned m = __reduce_max_sync(0xffffffffu, key);
    return __uint_as_float(m ^ ((m >> 31) ? 0x80000000u : 0xffffffffu));
}
// [lo, hi] certainly contains fl(psi_sum / fl(sqrt(phi_sum))) for the reference's sequential float sums of the same n
// numbers that (sp, sf) are float sums of in some other order; sa = float sum of |psi_i|, all phi_i >= 0.
// False: no statement (phi_sum may not be positive, overflow).
__device__ __forceinline__ bool lh_interval(float sp, float sa, float sf, int n, float &lo, float &hi) {
    const float g = __fmul_rn((float)n, 1.5e-7f);   // > 2 gamma_(n-1) (1 + gamma_n) = 1.2e-7 n for n <= 256
    const float dp = __fmul_rn(g, sa) + 1e-30f, df = __fmul_rn(g, sf) + 1e-30f;
    const float p_hi = sp + dp, p_lo = sp - dp, f_hi = sf + df, f_lo = sf - df;
    if (!(f_lo > 0.0f) || !(f_hi < FLT_MAX) || !(fabsf(p_hi) < FLT_MAX) || !(fabsf(p_lo) < FLT_MAX)) return false;
    const float s_lo = __fsqrt_rn(f_lo), s_hi = __fsqrt_rn(f_hi);
    hi = p_hi >= 0.0f ? __fdiv_rn(p_hi, s_lo) : __fdiv_rn(p_hi, s_hi);
    lo = p_lo >= 0.0f ? __fdiv_rn(p_lo, s_hi) : __fdiv_rn(p_lo, s_lo);
    // the roundings of these few operations and of the reference's own square root and division: < 1e-6 relative
    hi = hi + (fabsf(hi) * 4e-6f + 1e-30f);
    lo = lo - (fabsf(lo) * 4e-6f + 1e-30f);
    return true;
}
template <int E>
__device__ __forceinline__ bool warp_sigmag_rejects(int T, const float (&a)[E], const float (&b)[E],
                                                    const float (&lc)[E], const bool (&ok)[E], int n, float sgl_L,
                                                    float sgl_H, float coeff, int min_obs, float min_lh) {
    const unsigned FULL = 0xffffffffu;
    if (n == 0) return false;        // lh = flux = -1 there: left to the exact routine
    if (n < min_obs) return true;    // kernels.cu:318: obs_count < min_observations never enters the list
    bool odd = false;                // a negative phi or a non-finite lc: no bounds
    float sp = 0.0f, sa = 0.0f, sf = 0.0f;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        odd |= ok[e] && (b[e] < 0.0f || !(fabsf(lc[e]) < FLT_MAX));
        if (ok[e]) {
            sp = __fadd_rn(sp, a[e]);
            sa = __fadd_rn(sa, fabsf(a[e]));
            sf = __fadd_rn(sf, b[e]);
        }
    }
    if (__any_sync(FULL, odd)) return false;
    float lo, hi;
    if (!lh_interval(warp_sum(sp), warp_sum(sa), warp_sum(sf), n, lo, hi)) return false;
    if (hi < min_lh) return true;    // kernels.cu:201-203: the clip is skipped, and :318-320 drops lh < min_lh
    if (!(lo >= min_lh)) return false;
    // The clip happens.  cnt[e] = number of strictly smaller values; the value at sorted position p is the largest
    // one with cnt <= p (equal values included, whatever their order).
    float lcs[E];
    int cnt[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
        lcs[e] = ok[e] ? lc[e] : KB_NAN_F;
        cnt[e] = 0;
    }
#pragma unroll
    for (int es = 0; es < E; ++es) {
        const int nsrc = min(32, T - es * 32);  // warp-uniform
#pragma unroll 4
        for (int ls = 0; ls < nsrc; ++ls) {
            const float v = __shfl_sync(FULL, lcs[es], ls);
#pragma unroll
            for (int e = 0; e < E; ++e) cnt[e] += (v < lcs[e]) ? 1 : 0;
        }
    }
    // kernels.cu:94-127 percentile positions
    if ((double)sgl_L < 0.0001) sgl_L = (float)0.0001;
    if ((double)sgl_H > 0.9999) sgl_H = (float)0.9999;
    int pL = (int)((double)ceilf(__fmul_rn((float)n, sgl_L)) + 0.001) - 1;
    pL = pL < 0 ? 0 : pL;
    pL = pL >= n ? n - 1 : pL;
    int pH = (int)((double)ceilf(__fmul_rn((float)n, sgl_H)) + 0.001) - 1;
    pH = pH < 0 ? 0 : pH;
    pH = pH >= n ? n - 1 : pH;
    int med = (int)(ceil((double)n * 0.5) + 0.001) - 1;
    med = med < 0 ? 0 : med;
    med = med >= n ? n - 1 : med;
    float vL = -FLT_MAX, vH = -FLT_MAX, vM = -FLT_MAX;
    bool hL = false, hH = false, hM = false;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        if (ok[e] && cnt[e] <= pL) { vL = hL ? fmaxf(vL, lc[e]) : lc[e]; hL = true; }
        if (ok[e] && cnt[e] <= pH) { vH = hH ? fmaxf(vH, lc[e]) : lc[e]; hH = true; }
        if (ok[e] && cnt[e] <= med) { vM = hM ? fmaxf(vM, lc[e]) : lc[e]; hM = true; }
    }
    vL = warp_max_finite(vL, hL);
    vH = warp_max_finite(vH, hH);
    vM = warp_max_finite(vM, hM);
    // kernels.cu:130-146: the kept range of sorted positions == the values inside [lo_c, hi_c] (sorted, sigma_g >= 0)
    const float sigma_g = __fmul_rn(coeff, __fsub_rn(vH, vL));
    const float w = __fmul_rn(2.0f, sigma_g);
    const float lo_c = __fsub_rn(vM, w), hi_c = __fadd_rn(vM, w);
    if (!(fabsf(lo_c) < FLT_MAX) || !(fabsf(hi_c) < FLT_MAX) || !(lo_c <= vM) || !(vM <= hi_c)) return false;
    sp = 0.0f;
    sa = 0.0f;
    sf = 0.0f;
    int nk = 0;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const bool keep = ok[e] && lc[e] >= lo_c && lc[e] <= hi_c;
        nk += __popc(__ballot_sync(FULL, keep));
        if (keep) {
            sp = __fadd_rn(sp, a[e]);
            sa = __fadd_rn(sa, fabsf(a[e]));
            sf = __fadd_rn(sf, b[e]);
        }
    }
    if (!lh_interval(warp_sum(sp), warp_sum(sa), warp_sum(sf), nk, lo, hi)) return false;
    return hi < min_lh;              // kernels.cu:318-320: the clipped lh is below min_lh
}

template <int NB, int E>
__device__ __forceinline__ void warp_sigmag_body(const StackView &sv, const double *times, int x, int y, float vx,
                                                 float vy, float sgl_L, float sgl_H, float coeff, int min_obs,
                                                 float min_lh, float &lh, float &flux, int &n_valid,
                                                 bool &needs_serial, float2 *ws = nullptr) {
    float a[E], b[E], lc[E];
    bool ok[E];
    unsigned okm[E];
    const int n = warp_sigmag_load<NB, E>(sv, times, x, y, vx, vy, a, b, lc, ok, okm);
    warp_sigmag_eval<NB, E>(sv, a, b, lc, ok, okm, n, sgl_L, sgl_H, coeff, min_obs, min_lh, lh, flux, n_valid,
                            needs_serial, ws);
}

// Out of line for the search kernels (whose register budget belongs to the sweep); clip_kernel inlines the parts.
template <int NB, int E>
__device__ __noinline__ void warp_sigmag(const StackView &sv, const double *times, int x, int y, float vx,
                                            float vy, float sgl_L, float sgl_H, float coeff, int min_obs, float min_lh,
                                            float &lh, float &flux, int &n_valid, bool &needs_serial,
                                            float2 *ws = nullptr) {
    warp_sigmag_body<NB, E>(sv, times, x, y, vx, vy, sgl_L, sgl_H, coeff, min_obs, min_lh, lh, flux, n_valid,
                            needs_serial, ws);
}

template <int NB, bool SIGMAG, int MAXT>
__device__ __forceinline__ void evaluate_exact(const StackView &sv, const double *times, const kb_params &p, int x,
                                               int y, float vx, float vy, float &lh, float &flux, int &obs,
                                               uint8_t *mask);

// Evaluates, for every lane of the warp that `need`s it, the whole of evaluateTrajectory with sigma-G on
// (kernels.cu:154-242), one lane at a time, cooperatively.  (x, y, vx, vy) are the caller's per-lane values.  For
// the lanes that needed it: n_valid receives the number of valid observations; lh/flux receive the unclipped values
// when the reference skips the clip (n_valid < min_obs, n_valid == 0 or unclipped lh < min_lh; kernels.cu:201-203),
// the clipped ones otherwise.
// MAXT <= 256 selects the cooperative routine (E = MAXT/32); larger stacks use the serial routine.
template <int NB, int MAXT>
__device__ __noinline__ void sigmag_for_lanes(bool need, const StackView &sv, const double *times, int x, int y,
                                                 float vx, float vy, const kb_params &p, float &lh, float &flux,
                                                 int &n_valid, float2 *ws = nullptr) {
    if constexpr (MAXT > 256) {
        if (need) evaluate_exact<NB, true, MAXT>(sv, times, p, x, y, vx, vy, lh, flux, n_valid, nullptr);
        return;
    } else {
    constexpr int E = (MAXT + 31) / 32;
    const int lane = threadIdx.x & 31;
    unsigned m = __ballot_sync(0xffffffffu, need);
    while (m) {
        const int L = __ffs(m) - 1;
        m &= m - 1;
        const int bx = __shfl_sync(0xffffffffu, x, L), by = __shfl_sync(0xffffffffu, y, L);
        const float bvx = __shfl_sync(0xffffffffu, vx, L), bvy = __shfl_sync(0xffffffffu, vy, L);
        float rl = 0.0f, rf = 0.0f;
        int rn = 0;
        bool serial;
        warp_sigmag<NB, E>(sv, times, bx, by, bvx, bvy, p.sgl_L, p.sgl_H, p.sigmag_coeff, p.min_observations, p.min_lh,
                           rl, rf, rn, serial, ws);
        if (lane == L) {
            n_valid = rn;
            lh = rl;
            flux = rf;
            if (serial)   // tied lc values with different payloads: the reference's exchange-sort order decides
                sigmag_recompute<NB, MAXT>(sv, times, x, y, vx, vy, p.sgl_L, p.sgl_H, p.sigmag_coeff, lh, flux,
                                           nullptr);
        }
    }
    }
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
