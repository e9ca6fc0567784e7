// search_pairs.cu -- the candidate-parallel search for small start windows and long per-pixel lists.
//
// KBMOD's TrajectoryExplorer searches a handful of start pixels (11 x 11) with thousands of candidates and keeps ALL
// of them per pixel (results_per_pixel up to 10 000; trajectory_explorer.py:221-255 -> stack_search.cpp:221-284).
// One thread per start pixel leaves a 148-SM device with 121 busy threads, each inserting into a list of thousands
// kept in global memory.  Here the work is spread over (start pixel, candidate) PAIRS instead:
//   1. one thread per pair evaluates it exactly (evaluateTrajectory, kernels.cu:154-242) and applies the rejection
//      rules of the per-pixel loop (kernels.cu:318-320: obs_count < min_obs; with sigma-G on, lh < min_lh; a NaN lh
//      never passes the strict '>' of the insertion);
//   2. two stable radix sorts (by lh descending, then by start pixel) turn the pair array into per-pixel lists ordered
//      by (lh descending, candidate index ascending) -- the order the reference's in-order strict-'>' insertion
//      produces, exact lh ties aside -- with the rejected pairs at the end of every list;
//   3. the first results_per_pixel entries of every list become that pixel's slots.
// Used when results_per_pixel > 8 (the tiled kernel keeps at most 8 in registers) and the pair array is small enough.
#include <float.h>

#include "search_device.cuh"

namespace kb {

__device__ __forceinline__ uint32_t pair_lh_key(float lh) {   // ascending unsigned order == descending lh
    const uint32_t u = __float_as_uint(lh == 0.0f ? 0.0f : lh);
    const uint32_t ordered = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ~ordered;
}

template <int NB, bool SIGMAG, int MAXT>
__global__ void __launch_bounds__(128)
        pair_eval_kernel(StackView sv, kb_params p, const float *__restrict__ cvx, const float *__restrict__ cvy,
                         uint32_t n_cand, uint64_t n_pairs, Slot *__restrict__ pairs, uint32_t *__restrict__ keys,
                         uint32_t *__restrict__ vals) {
    extern __shared__ double s_times[];
    for (int i = threadIdx.x; i < sv.T; i += blockDim.x) s_times[i] = sv.times[i];
    __syncthreads();
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    const int sw = p.x_start_max - p.x_start_min;
    const uint32_t pix = (uint32_t)(i / n_cand), c = (uint32_t)(i % n_cand);
    const int x = p.x_start_min + (int)(pix % (uint32_t)sw), y = p.y_start_min + (int)(pix / (uint32_t)sw);
    float lh, flux;
    int obs;
    evaluate_exact<NB, SIGMAG, MAXT>(sv, s_times, p, x, y, __ldg(cvx + c), __ldg(cvy + c), lh, flux, obs, nullptr);
    const bool rejected = obs < p.min_observations || (SIGMAG && lh < p.min_lh) || !(lh == lh);
    pairs[i] = rejected ? Slot{-FLT_MAX, 0.0f, 0, -1} : Slot{lh, flux, obs, (int)c};
    keys[i] = rejected ? 0xffffffffu : pair_lh_key(lh);
    vals[i] = (uint32_t)i;
}

__global__ void pair_pixel_keys_kernel(const uint32_t *__restrict__ vals, uint64_t n, uint32_t n_cand,
                                       uint32_t *__restrict__ keys) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = vals[i] / n_cand;
}

// sorted position j holds pair vals[j]; within its pixel's run (n_cand entries) its rank is j - pixel * n_cand
__global__ void pair_scatter_kernel(const Slot *__restrict__ pairs, const uint32_t *__restrict__ vals, uint64_t n,
                                    uint32_t n_cand, uint32_t K, Slot *__restrict__ slots) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t pi = vals[j];
    const uint64_t pix = pi / n_cand, rank = j - pix * n_cand;
    if (rank < K) slots[pix * K + rank] = pairs[pi];
}

__global__ void pair_fill_empty_kernel(Slot *slots, uint64_t n_pixels, uint32_t K, uint32_t first) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t span = K - first;
    if (i >= n_pixels * span) return;
    slots[(i / span) * K + first + i % span] = Slot{-FLT_MAX, 0.0f, 0, -1};
}

template <int NB>
static void launch_pair_eval(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                             uint32_t n_cand, uint64_t n_pairs, Slot *d_pairs, uint32_t *keys, uint32_t *vals,
                             cudaStream_t st) {
    const unsigned gs = (unsigned)((n_pairs + 127) / 128);
    const size_t sm = sizeof(double) * (size_t)sv.T;
    if (sm > 48 * 1024) throw Error("Too many images for the candidate-parallel search.");
#define KB_PAIR(SG, MT) pair_eval_kernel<NB, SG, MT><<<gs, 128, sm, st>>>(sv, p, d_vx, d_vy, n_cand, n_pairs, d_pairs, keys, vals)
    if (!p.do_sigmag_filter) KB_PAIR(false, 1);
    else if (sv.T <= 64) KB_PAIR(true, 64);
    else if (sv.T <= 256) KB_PAIR(true, 256);
    else if (sv.T <= kMaxTSigma) KB_PAIR(true, kMaxTSigma);
    else throw Error("The on-device sigma-G filter supports at most " + std::to_string(kMaxTSigma) + " images. Got " + std::to_string(sv.T));
#undef KB_PAIR
    KB_CUDA(cudaGetLastError());
}

bool search_pairs_usable(const kb_params &p, uint64_t n_cand, std::string *why_not) {
    const uint64_t sw = (uint64_t)(p.x_start_max - p.x_start_min), sh = (uint64_t)(p.y_start_max - p.y_start_min);
    if (n_cand == 0) {
        if (why_not) *why_not = "no candidates";
        return false;
    }
    if (sw * sh * n_cand > kMaxSearchPairs || n_cand >= (1ull << 31)) {
        if (why_not) *why_not = "start pixels x candidates exceeds the pair buffer (" + std::to_string(kMaxSearchPairs) + ")";
        return false;
    }
    return true;
}

int launch_search_pairs(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, uint64_t n_cand,
                        Slot *d_slots, Slot *d_pairs, ResultScratch &scr, cudaStream_t st) {
    const uint64_t sw = (uint64_t)(p.x_start_max - p.x_start_min), sh = (uint64_t)(p.y_start_max - p.y_start_min);
    const uint64_t n_pix = sw * sh, n = n_pix * n_cand;
    const uint32_t K = p.results_per_pixel;
    result_scratch_reserve(scr, n);
    int launches = 0;
    switch (sv.nb) {
        case 4: launch_pair_eval<4>(sv, p, d_vx, d_vy, (uint32_t)n_cand, n, d_pairs, scr.keys[0], scr.vals[0], st); break;
        case 2: launch_pair_eval<2>(sv, p, d_vx, d_vy, (uint32_t)n_cand, n, d_pairs, scr.keys[0], scr.vals[0], st); break;
        case 1: launch_pair_eval<1>(sv, p, d_vx, d_vy, (uint32_t)n_cand, n, d_pairs, scr.keys[0], scr.vals[0], st); break;
        default: throw Error("bad num_bytes");
    }
    launches += 1 + radix_sort_u32_pairs(scr, n, st);
    const unsigned gs = (unsigned)((n + 255) / 256);
    pair_pixel_keys_kernel<<<gs, 256, 0, st>>>(scr.vals[0], n, (uint32_t)n_cand, scr.keys[0]);
    launches += 1 + radix_sort_u32_pairs(scr, n, st);
    if (K > n_cand) {   // more slots than candidates: the tail of every list stays empty
        const uint64_t m = n_pix * (K - (uint32_t)n_cand);
        pair_fill_empty_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(d_slots, n_pix, K, (uint32_t)n_cand);
        ++launches;
    }
    pair_scatter_kernel<<<gs, 256, 0, st>>>(d_pairs, scr.vals[0], n, (uint32_t)n_cand, K, d_slots);
    KB_CUDA(cudaGetLastError());
    return launches + 1;
}

}  // namespace kb
