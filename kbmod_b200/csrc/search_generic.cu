// search_generic.cu -- the general search kernel and the small per-trajectory kernels.
//
// One thread per start pixel, candidates visited in list order, exact fp64 index arithmetic per observation,
// gathers straight from the L2/HBM-resident search copy.  It accepts every input the API allows (any bounds,
// any results_per_pixel, any candidate list) and is the fallback of the TMA-staged tiled kernel
// (search_tiled.cu), which covers the throughput-critical shapes.
// Reference: kernels/kernels.cu:252-332 searchFilterImages, :154-242 evaluateTrajectory.
#include "search_device.cuh"

namespace kb {

static constexpr int GX = 32, GY = 8;

template <int NB, bool SIGMAG, int MAXT>
__global__ void __launch_bounds__(GX *GY)
        search_generic_kernel(StackView sv, kb_params p, const float *__restrict__ cvx,
                              const float *__restrict__ cvy, uint64_t n_cand, Slot *__restrict__ slots) {
    extern __shared__ double s_times[];
    for (int i = threadIdx.y * GX + threadIdx.x; i < sv.T; i += GX * GY) s_times[i] = sv.times[i];
    __syncthreads();

    const int sw = p.x_start_max - p.x_start_min, sh = p.y_start_max - p.y_start_min;
    const int xi = blockIdx.x * GX + threadIdx.x, yi = blockIdx.y * GY + threadIdx.y;
    // Threads past the search bounds stay in the loop (the sigma-G clip is warp-cooperative and needs whole
    // warps); they just never write.
    const bool active = xi < sw && yi < sh;
    const int x = xi + p.x_start_min, y = yi + p.y_start_min;
    const int K = (int)p.results_per_pixel;
    Slot *my = slots + (active ? ((size_t)yi * sw + xi) * K : 0);
    const bool in_regs = K <= kMaxRegK;

    // Per-pixel sorted list: registers when K <= 8, else the output slots themselves (kernels.cu:293-301).
    float r_lh[kMaxRegK], r_flux[kMaxRegK];
    int r_obs[kMaxRegK], r_cand[kMaxRegK];
#pragma unroll
    for (int r = 0; r < kMaxRegK; ++r) {
        r_lh[r] = -FLT_MAX;
        r_flux[r] = 0.0f;
        r_obs[r] = 0;
        r_cand[r] = -1;
    }
    if (!in_regs && active)
        for (int r = 0; r < K; ++r) my[r] = Slot{-FLT_MAX, 0.0f, 0, -1};

    for (uint64_t c = 0; c < n_cand; ++c) {
        const float vx = __ldg(cvx + c), vy = __ldg(cvy + c);
        float psi_sum = 0.0f, phi_sum = 0.0f;
        int seen = 0;
        for (int t = 0; t < sv.T; ++t) {
            const double tm = s_times[t];
            float a, b;
            if (load_pair<NB>(sv, t, predict_index(y, vy, tm), predict_index(x, vx, tm), a, b)) {
                psi_sum = __fadd_rn(psi_sum, a);
                phi_sum = __fadd_rn(phi_sum, b);
                ++seen;
            }
        }
        float lh, flux;
        finish_sums(psi_sum, phi_sum, lh, flux);
        if (SIGMAG) {
            const bool need = active && !(seen < p.min_observations || seen == 0 || lh < p.min_lh);
            int nv = seen;
            sigmag_for_lanes<NB, MAXT>(need, sv, s_times, x, y, vx, vy, p, lh, flux, nv);  // all lanes take part
        }
        if (!active || seen < p.min_observations || (SIGMAG && lh < p.min_lh)) continue;  // kernels.cu:318-320
        int cc = (int)c;
        if (in_regs) {
            if (!(lh > r_lh[kMaxRegK - 1])) continue;
#pragma unroll
            for (int r = 0; r < kMaxRegK; ++r) {
                if (lh > r_lh[r]) {  // strict '>' in list order (kernels.cu:324-330)
                    float tf = r_lh[r]; r_lh[r] = lh; lh = tf;
                    tf = r_flux[r]; r_flux[r] = flux; flux = tf;
                    int ti = r_obs[r]; r_obs[r] = seen; seen = ti;
                    ti = r_cand[r]; r_cand[r] = cc; cc = ti;
                }
            }
        } else {
            if (!(lh > my[K - 1].lh)) continue;
            Slot cur{lh, flux, seen, cc};
            for (int r = 0; r < K; ++r) {
                const Slot s = my[r];
                if (cur.lh > s.lh) {
                    my[r] = cur;
                    cur = s;
                }
            }
        }
    }
    if (in_regs && active) {
#pragma unroll
        for (int r = 0; r < kMaxRegK; ++r)
            if (r < K) my[r] = Slot{r_lh[r], r_flux[r], r_obs[r], r_cand[r]};
    }
}

template <int NB>
static int launch_generic_nb(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                             uint64_t n_cand, Slot *d_slots, cudaStream_t st) {
    const int sw = p.x_start_max - p.x_start_min, sh = p.y_start_max - p.y_start_min;
    dim3 grid((sw + GX - 1) / GX, (sh + GY - 1) / GY), block(GX, GY);
    const size_t smem = sizeof(double) * (size_t)sv.T;
    if (smem > 200 * 1024) throw Error("Too many images for the search kernel (times do not fit shared memory).");
#define KB_LAUNCH_G(SG, MT)                                                                               \
    do {                                                                                                  \
        auto kfn = search_generic_kernel<NB, SG, MT>;                                                     \
        if (smem > 48 * 1024)                                                                             \
            KB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
        kfn<<<grid, block, smem, st>>>(sv, p, d_vx, d_vy, n_cand, d_slots);                               \
    } while (0)
    if (!p.do_sigmag_filter) {
        KB_LAUNCH_G(false, 1);
    } else if (sv.T <= 32) {
        KB_LAUNCH_G(true, 32);
    } else if (sv.T <= 64) {
        KB_LAUNCH_G(true, 64);
    } else if (sv.T <= 256) {
        KB_LAUNCH_G(true, 256);
    } else if (sv.T <= kMaxTSigma) {
        KB_LAUNCH_G(true, kMaxTSigma);
    } else {
        throw Error("The on-device sigma-G filter supports at most " + std::to_string(kMaxTSigma) +
                    " images. Got " + std::to_string(sv.T));
    }
#undef KB_LAUNCH_G
    KB_CUDA(cudaGetLastError());
    return 1;
}

int launch_search_generic(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                          uint64_t n_cand, Slot *d_slots, cudaStream_t st) {
    switch (sv.nb) {
        case 4: return launch_generic_nb<4>(sv, p, d_vx, d_vy, n_cand, d_slots, st);
        case 2: return launch_generic_nb<2>(sv, p, d_vx, d_vy, n_cand, d_slots, st);
        case 1: return launch_generic_nb<1>(sv, p, d_vx, d_vy, n_cand, d_slots, st);
    }
    throw Error("bad num_bytes");
}

// ---- evaluate_single_trajectory(use_kernel=true): one thread per trajectory ---------------------------
template <int NB, int MAXT>
__global__ void evaluate_kernel(StackView sv, kb_params p, kb_trajectory *trjs, uint64_t n, uint8_t *mask) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    kb_trajectory t = trjs[i];
    float lh, flux;
    int obs;
    uint8_t *m = mask ? mask + i * (uint64_t)sv.T : nullptr;
    if (p.do_sigmag_filter)
        evaluate_exact<NB, true, MAXT>(sv, sv.times, p, t.x, t.y, t.vx, t.vy, lh, flux, obs, m);
    else
        evaluate_exact<NB, false, 1>(sv, sv.times, p, t.x, t.y, t.vx, t.vy, lh, flux, obs, m);
    t.lh = lh;
    t.flux = flux;
    t.obs_count = obs;
    trjs[i] = t;
}

template <int NB>
static int launch_evaluate_nb(const StackView &sv, const kb_params &p, kb_trajectory *d, uint64_t n, uint8_t *mask,
                              cudaStream_t st) {
    const int bs = 64;
    const unsigned gs = (unsigned)((n + bs - 1) / bs);
    if (p.do_sigmag_filter && sv.T > kMaxTSigma)
        throw Error("The on-device sigma-G filter supports at most " + std::to_string(kMaxTSigma) + " images.");
    if (sv.T <= 64 || !p.do_sigmag_filter)
        evaluate_kernel<NB, 64><<<gs, bs, 0, st>>>(sv, p, d, n, mask);
    else if (sv.T <= 256)
        evaluate_kernel<NB, 256><<<gs, bs, 0, st>>>(sv, p, d, n, mask);
    else
        evaluate_kernel<NB, kMaxTSigma><<<gs, bs, 0, st>>>(sv, p, d, n, mask);
    KB_CUDA(cudaGetLastError());
    return 1;
}

int launch_evaluate(const StackView &sv, const kb_params &p, kb_trajectory *d_trjs, uint64_t n, uint8_t *d_mask,
                    cudaStream_t st) {
    if (n == 0) return 0;
    switch (sv.nb) {
        case 4: return launch_evaluate_nb<4>(sv, p, d_trjs, n, d_mask, st);
        case 2: return launch_evaluate_nb<2>(sv, p, d_trjs, n, d_mask, st);
        case 1: return launch_evaluate_nb<1>(sv, p, d_trjs, n, d_mask, st);
    }
    throw Error("bad num_bytes");
}

// ---- sigmag_filtered_indices test hook: the device routine on one vector ----------------------------
__global__ void sigmag_indices_kernel(const float *vals, int n, float sgl0, float sgl1, float coeff, float width,
                                      int *idx, int *range) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        int lo, hi;
        sigmag_indices(vals, n, sgl0, sgl1, coeff, width, idx, lo, hi);
        range[0] = lo;
        range[1] = hi;
    }
}

int launch_sigmag_indices(const float *d_vals, int n, float sgl0, float sgl1, float coeff, float width, int *d_idx,
                          int *d_range, cudaStream_t st) {
    sigmag_indices_kernel<<<1, 32, 0, st>>>(d_vals, n, sgl0, sgl1, coeff, width, d_idx, d_range);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- get_all_psi_phi_curves (stack_search.cpp:14-31): reads the *reference-layout* array (not canonicalised;
// psi and phi validity are independent here) with Trajectory::get_x_index rounding (common.h:71-79: the
// position is narrowed to float before the floor, unlike predict_index). ---------------------------------
__global__ void curves_kernel(const void *data, const double *times, int T, int H, int W, int nb, float psi_min,
                              float psi_scale, float phi_min, float phi_scale, const kb_trajectory *trjs,
                              uint64_t n, float *out) {
    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= n * (uint64_t)T) return;
    const uint64_t r = gid / T;
    const int t = (int)(gid % T);
    const kb_trajectory tr = trjs[r];
    const double tm = times[t];
    const float xp = (float)__dadd_rn(__dadd_rn((double)tr.x, __dmul_rn(tm, (double)tr.vx)), 0.5);
    const float yp = (float)__dadd_rn(__dadd_rn((double)tr.y, __dmul_rn(tm, (double)tr.vy)), 0.5);
    const int col = (int)floorf(xp), row = (int)floorf(yp);
    float psi = KB_NAN_F, phi = KB_NAN_F;
    if (row >= 0 && col >= 0 && row < H && col < W) {
        const size_t idx = 2 * ((size_t)t * H * W + (size_t)row * W + col);
        if (nb == 4) {
            psi = reinterpret_cast<const float *>(data)[idx];
            phi = reinterpret_cast<const float *>(data)[idx + 1];
        } else if (nb == 2) {
            psi = decode_code(reinterpret_cast<const uint16_t *>(data)[idx], psi_scale, psi_min);
            phi = decode_code(reinterpret_cast<const uint16_t *>(data)[idx + 1], phi_scale, phi_min);
        } else {
            psi = decode_code(reinterpret_cast<const uint8_t *>(data)[idx], psi_scale, psi_min);
            phi = decode_code(reinterpret_cast<const uint8_t *>(data)[idx + 1], phi_scale, phi_min);
        }
    }
    out[r * 2 * T + t] = isfinite(psi) ? psi : 0.0f;
    out[r * 2 * T + T + t] = isfinite(phi) ? phi : 0.0f;
}

int launch_curves(const void *d_ref_layout, const double *d_times, int T, int H, int W, int nb, float psi_min,
                  float psi_scale, float phi_min, float phi_scale, const kb_trajectory *d_trjs, uint64_t n,
                  float *d_out, cudaStream_t st) {
    if (n == 0) return 0;
    const uint64_t total = n * (uint64_t)T;
    curves_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(d_ref_layout, d_times, T, H, W, nb, psi_min,
                                                                    psi_scale, phi_min, phi_scale, d_trjs, n, d_out);
    KB_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace kb
