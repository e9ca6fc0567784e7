// search_tiled_kernel.cuh -- device code of the TMA-staged tiled search kernel and its per-encoding launchers.
// Included by search_tiled.cu (host plan + dispatcher) and by search_tiled_nb{4,2,1}.cu, each of which
// instantiates the kernels of one stack encoding so the three compile in parallel.  See search_tiled.cu for the
// description of the algorithm.
#pragma once
#include <cuda.h>
#include <float.h>
#include <stddef.h>

#include "search_device.cuh"

namespace kb {

static constexpr int TBX = 32;         // tile width  (lanes)
static constexpr int TBY = 15;         // tile height (consumer warps); +1 producer warp = 512 threads
static constexpr int TNT = (TBY + 1) * 32;
static constexpr int TG = 16;          // candidates per group
static constexpr int TS = 4;           // pipeline stages
static constexpr int TCH = 4;          // rows per TMA box
static constexpr int TMAXTC = 8;       // max images per stage
static constexpr int STAGE_BYTES = 48 * 1024;
static constexpr int NMAPS = 16;       // tensor maps with box widths base, base+step, ...

// One pipeline stage as the host lays it out (and as it lands in shared memory by one bulk copy):
//   meta   {nt | last << 8, group, t0, tx_bytes}
//   box[k] {ox, oy, nrows | widx << 16, slot_off}   for the nt images of the stage (k < TMAXTC)
//   offs[k][j] byte offset of candidate j's pixel inside image k's slot, relative to the thread's own pixel
struct StageRec {
    int4 meta;
    int4 box[TMAXTC];
    int pad[4];
    int offs[TMAXTC][TG];
};
static_assert(sizeof(StageRec) % 16 == 0 && offsetof(StageRec, offs) % 16 == 0, "bulk copies need 16-byte units");
static constexpr int HDR_BYTES = (sizeof(StageRec) + 127) / 128 * 128;

struct TmapSet {
    CUtensorMap m[NMAPS];
};

struct PlanView {
    const StageRec *stages;  // [n_stages]
    const int *cand;         // [n_groups][TG]  candidate index or -1
    const int *exact;        // candidates evaluated by the exact routine
    int n_exact, n_groups, n_stages, w_base, w_step, max_ch, T;
    uint2 *clipq;                   // deferred sigma-G requests {start-pixel index, candidate}
    unsigned long long *clip_cnt;  // [0] requests, [1] survivors, [2] clipped in line (queue full)
    unsigned long long clip_cap;
    unsigned pixel_base;            // start-pixel index of this launch's first pixel within the whole search area
};

struct TiledPlan {
    TmapSet maps;
    PlanView v{};
    void *d_block = nullptr;  // one allocation holding all tables
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the gather kernel
    size_t smem = 0;
    int reach_hi = 0;  // a tile at row y0 reads stack rows below y0 + reach_hi (largest box bottom, padded)
    int nb = 4, K = 8;
    bool sigmag = false;
};

// ---- PTX wrappers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_LOOP:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
            "@p bra DONE;\n\t"
            "bra WAIT_LOOP;\n\t"
            "DONE:\n\t}" ::"r"(bar),
            "r"(parity)
            : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2,
                                            uint32_t bar) {
    asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
            : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

// ---- pair decode from shared memory ---------------------------------------------------------------------
template <int NB>
struct PairLoad;
template <>
struct PairLoad<4> {
    static constexpr int kPairBytes = 8;
    template <bool ZM>
    __device__ static __forceinline__ bool load(uint32_t addr, const StackView &sv, float &psi, float &phi) {
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(psi), "=f"(phi) : "r"(addr));
        // canonical copy: an invalid pair is (+0, +0) in zero-marker form, else (NaN, NaN)
        if (ZM) return ((__float_as_uint(psi) | __float_as_uint(phi)) << 1) != 0u;
        return phi == phi;
    }
};
// Encoded stacks: the host verified (plan creation) that fmaf((float)(code-1), scale, min) reproduces the
// reference's fp64 decode for every code, so the gather decodes in fp32.
template <>
struct PairLoad<2> {
    static constexpr int kPairBytes = 4;
    template <bool ZM>
    __device__ static __forceinline__ bool load(uint32_t addr, const StackView &sv, float &psi, float &phi) {
        uint32_t w;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(addr));
        const float a = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7410)) - 8388609.0f;  // code - 1, exact
        const float b = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7432)) - 8388609.0f;
        psi = fmaf(a, sv.psi_scale, sv.psi_min);
        phi = fmaf(b, sv.phi_scale, sv.phi_min);
        return w != 0u;  // canonical copy: an invalid pair is (0, 0)
    }
};
template <>
struct PairLoad<1> {
    static constexpr int kPairBytes = 2;
    template <bool ZM>
    __device__ static __forceinline__ bool load(uint32_t addr, const StackView &sv, float &psi, float &phi) {
        uint16_t h;
        asm volatile("ld.shared.u16 %0, [%1];" : "=h"(h) : "r"(addr));
        const uint32_t w = h;
        const float a = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7440)) - 8388609.0f;
        const float b = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7441)) - 8388609.0f;
        psi = fmaf(a, sv.psi_scale, sv.psi_min);
        phi = fmaf(b, sv.phi_scale, sv.phi_min);
        return w != 0u;
    }
};

// ---- the kernel -----------------------------------------------------------------------------------------
// ZM (float stacks in zero-marker form, StackView::zero_marker): invalid pairs and out-of-image pixels read as
// (+0, +0), so the inner loop adds unconditionally.
// COUNT: keep a per-candidate observation count in registers.  Only needed when sigma-G is off and min_obs > 0;
// with sigma-G on, the count of the few candidates that reach min_lh falls out of the clip's own re-gather.
template <int NB, bool ZM, bool COUNT, bool SIGMAG, int MAXT, int K>
__global__ void __launch_bounds__(TNT, 1)
        search_tiled_kernel(const __grid_constant__ TmapSet maps, const PlanView pv, const StackView sv,
                            const kb_params p, const float *__restrict__ cvx, const float *__restrict__ cvy,
                            Slot *__restrict__ slots) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *s_stage = smem;                                              // TS * STAGE_BYTES
    uint8_t *s_hdr = smem + TS * STAGE_BYTES;                             // TS * HDR_BYTES
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(smem + TS * (STAGE_BYTES + HDR_BYTES));  // full[TS], empty[TS]
    double *s_times = reinterpret_cast<double *>(s_bar + 2 * TS);         // T doubles (exact routines)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = smem_u32(s_bar);
    if (threadIdx.x == 0) {
        for (int i = 0; i < TS; ++i) {
            mbar_init(bar0 + 8 * i, 1);           // full: one arrive (producer) + tx bytes
            mbar_init(bar0 + 8 * (TS + i), TBY);  // empty: one arrive per consumer warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < sv.T; i += TNT) s_times[i] = sv.times[i];
    __syncthreads();

    const int sw = p.x_start_max - p.x_start_min, sh = p.y_start_max - p.y_start_min;
    const int tile_x0 = p.x_start_min + blockIdx.x * TBX, tile_y0 = p.y_start_min + blockIdx.y * TBY;
    const int n_stages = pv.n_stages;
    constexpr int PB = PairLoad<NB>::kPairBytes;
    constexpr int XE = (NB == 4) ? 2 : 1;  // tensor-map elements per pixel along x

    if (warp == TBY) {
        // ===================== producer warp: TMA box loads + the stage record =================================
        // Lane l < 9 keeps int4 #l of the *next* stage record (meta + 8 boxes) in a register, loaded one stage ahead
        // so the global-memory latency overlaps the wait for a free slot; fields are handed out by shuffles.
        const unsigned FULLM = 0xffffffffu;
        auto fetch = [&](int s) -> int4 {
            if (lane < 1 + TMAXTC && s < n_stages) return __ldg(reinterpret_cast<const int4 *>(pv.stages + s) + lane);
            return make_int4(0, 0, 0, 0);
        };
        int4 nxt = fetch(0);
        for (int s = 0; s < n_stages; ++s) {
            const int slot = s % TS;
            const uint32_t full = bar0 + 8 * slot, empty = bar0 + 8 * (TS + slot);
            const int4 cur = nxt;
            nxt = fetch(s + 1);
            if (s >= TS) mbar_wait(empty, ((s / TS) - 1) & 1);
            const int nt = __shfl_sync(FULLM, cur.x, 0) & 0xff, t0 = __shfl_sync(FULLM, cur.z, 0);
            const int tx = __shfl_sync(FULLM, cur.w, 0);
            if (lane == 0) mbar_arrive_expect_tx(full, (uint32_t)tx);
            __syncwarp();
            const uint32_t stage_base = smem_u32(s_stage + (size_t)slot * STAGE_BYTES);
            const int total = nt * pv.max_ch;
            for (int i0 = 0; i0 < total; i0 += 32) {
                const int i = i0 + lane;
                const int k = min(i / pv.max_ch, TMAXTC - 1), ch = i % pv.max_ch;
                const int bxx = __shfl_sync(FULLM, cur.x, k + 1), bxy = __shfl_sync(FULLM, cur.y, k + 1);
                const int bxz = __shfl_sync(FULLM, cur.z, k + 1), bxw = __shfl_sync(FULLM, cur.w, k + 1);
                const int nrows = bxz & 0xffff, widx = bxz >> 16;
                if (i < total && ch * TCH < nrows) {
                    const int pitch = pv.w_base + pv.w_step * widx;
                    tma_load_3d(stage_base + bxw + ch * TCH * pitch * PB, &maps.m[widx], XE * (tile_x0 + bxx),
                                tile_y0 + bxy + ch * TCH, t0 + k, full);
                }
            }
            if (lane == 0)
                bulk_load(smem_u32(s_hdr + (size_t)slot * HDR_BYTES), pv.stages + s, (uint32_t)sizeof(StageRec), full);
        }
    } else {
        // ===================== consumer warps: gather, reduce, rank ========================================
        const int px = tile_x0 + lane, py = tile_y0 + warp;
        const bool in_bounds = (px - p.x_start_min) < sw && (py - p.y_start_min) < sh;
        const unsigned pixel_index = pv.pixel_base + (unsigned)((py - p.y_start_min) * sw + (px - p.x_start_min));
        float psi[TG], phi[TG];
        int cnt[COUNT ? TG : 1];
        TopK<K> top;
        top.init();
#pragma unroll
        for (int j = 0; j < TG; ++j) {
            psi[j] = 0.0f;
            phi[j] = 0.0f;
            if (COUNT) cnt[j] = 0;
        }
        for (int s = 0; s < n_stages; ++s) {
            const int slot = s % TS;
            const uint32_t full = bar0 + 8 * slot, empty = bar0 + 8 * (TS + slot);
            mbar_wait(full, (s / TS) & 1);
            const StageRec *hdr = reinterpret_cast<const StageRec *>(s_hdr + (size_t)slot * HDR_BYTES);
            const int4 meta = hdr->meta;
            const int nt = meta.x & 0xff;
            const uint32_t base = smem_u32(s_stage + (size_t)slot * STAGE_BYTES);
            for (int k = 0; k < nt; ++k) {
                const int4 bx = hdr->box[k];
                const int pitch = pv.w_base + pv.w_step * (bx.z >> 16);
                const uint32_t bk = base + bx.w + (uint32_t)((warp * pitch + lane) * PB);
                const int4 *ofs = reinterpret_cast<const int4 *>(hdr->offs[k]);
                // Neighbouring candidates of a group often land on the same pixel of an image (always for t ~ 0,
                // mostly for dense velocity grids); the host orders a group as a chain of nearest neighbours, and
                // inside each run of four a candidate whose offset equals its predecessor's reuses the pair already
                // in registers.  The comparison is warp-uniform (offsets are per (candidate, image), not per pixel);
                // the four runs stay independent so their loads overlap.  Only where it compiles to predicated
                // loads (float pairs, no observation counter); the other variants load every time.
                constexpr bool DEDUP = (NB == 4) && !COUNT;
#pragma unroll
                for (int j4 = 0; j4 < TG / 4; ++j4) {
                    const int4 o = ofs[j4];
                    const int off[4] = {o.x, o.y, o.z, o.w};
                    float a = 0.0f, b = 0.0f;
                    bool ok = false;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int j = j4 * 4 + q;
                        if (!DEDUP || q == 0 || off[q] != off[q - 1])
                            ok = PairLoad<NB>::template load<ZM>(bk + off[q], sv, a, b);
                        if (ZM) {
                            psi[j] = __fadd_rn(psi[j], a);
                            phi[j] = __fadd_rn(phi[j], b);
                            if (COUNT && NB == 4) {
                                // valid <=> the pair is not (+-0, +-0): one LOP3 with a predicate result and one
                                // predicated add in place (the compiler's select form costs two more per gather)
                                asm("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
                                    "lop3.b32 t, %1, 0x7fffffff, %2, 0xc8;\n\t"
                                    "setp.ne.u32 p, t, 0;\n\t"
                                    "@p add.s32 %0, %0, 1;\n\t}"
                                    : "+r"(cnt[j])
                                    : "r"(__float_as_uint(a)), "r"(__float_as_uint(b)));
                            } else if (COUNT && ok) {
                                cnt[j] += 1;
                            }
                        } else if (ok) {
                            psi[j] = __fadd_rn(psi[j], a);
                            phi[j] = __fadd_rn(phi[j], b);
                            if (COUNT) cnt[j] += 1;
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty);
            if (meta.x >> 8) {
                // ---- all images of this group are in: turn sums into ranked candidates -------------------
                const int g = meta.y;
#pragma unroll
                for (int j = 0; j < TG; ++j) {
                    const int c = __ldg(pv.cand + g * TG + j);  // warp-uniform
                    const float ps = psi[j], ph = phi[j];
                    psi[j] = 0.0f;
                    phi[j] = 0.0f;
                    int seen = 0;
                    if (COUNT) {
                        seen = cnt[j];
                        cnt[j] = 0;
                    }
                    if (c < 0) continue;
                    bool alive = !COUNT || seen >= p.min_observations;  // kernels.cu:318 obs_count < min_obs
                    // Conservative reject: est approximates lh to ~1e-6 relative; anything clearly below the
                    // bar (min_lh when sigma-G is on, else the current K-th best) cannot matter.
                    const float bar_lh = SIGMAG ? p.min_lh : top.worst();
                    const float est = (ph > 0.0f) ? ps * rsqrtf(ph) : -1.0f;
                    alive = alive && !(est < bar_lh - (fabsf(est) + fabsf(bar_lh)) * 2e-6f - 1e-30f);
                    if (!SIGMAG) {
                        if (!alive) continue;
                        float lh, flux;
                        finish_sums(ps, ph, lh, flux);
                        if (beats(lh, c, top.worst(), top.cand[K - 1])) top.insert(lh, c);
                    } else {
                        if (!__any_sync(0xffffffffu, alive)) continue;
                        float lh = -1.0f, flux = -1.0f;
                        if (alive) finish_sums(ps, ph, lh, flux);
                        alive = alive && !(lh < p.min_lh);  // kernels.cu:201-203 then :319
                        // The clip itself is deferred: append (pixel, candidate) to the global queue; clip_kernel
                        // works it off afterwards with every SM (the requests cluster around bright sources).
                        alive = alive && in_bounds;
                        const unsigned am = __ballot_sync(0xffffffffu, alive);
                        if (am) {
                            unsigned long long base_slot = 0;
                            if (lane == 0) base_slot = atomicAdd(pv.clip_cnt, (unsigned long long)__popc(am));
                            base_slot = __shfl_sync(0xffffffffu, base_slot, 0);
                            const unsigned long long slot_q = base_slot + __popc(am & ((1u << lane) - 1u));
                            bool inline_need = false;
                            if (alive) {
                                if (slot_q < pv.clip_cap) pv.clipq[slot_q] = make_uint2(pixel_index, (unsigned)c);
                                else inline_need = true;
                            }
                            if (__any_sync(0xffffffffu, inline_need)) {  // queue full: clip it here and now
                                int nv = 0;
                                sigmag_for_lanes<NB, MAXT>(inline_need, sv, s_times, px, py, __ldg(cvx + c),
                                                           __ldg(cvy + c), p, lh, flux, nv);
                                if (inline_need) {
                                    atomicAdd(pv.clip_cnt + 2, 1ull);
                                    if (!(nv < p.min_observations) && !(lh < p.min_lh) &&
                                        beats(lh, c, top.worst(), top.cand[K - 1]))
                                        top.insert(lh, c);
                                }
                            }
                        }
                    }
                }
            }
        }
        // ---- candidates whose offsets could not be tabulated exactly: full per-pixel arithmetic ----------
        for (int e = 0; e < pv.n_exact; ++e) {
            const int c = __ldg(pv.exact + e);
            float lh, flux;
            int obs;
            evaluate_exact<NB, SIGMAG, MAXT>(sv, s_times, p, px, py, __ldg(cvx + c), __ldg(cvy + c), lh, flux, obs,
                                             nullptr);
            if (obs < p.min_observations || (SIGMAG && lh < p.min_lh)) continue;
            if (beats(lh, c, top.worst(), top.cand[K - 1])) top.insert(lh, c);
        }
        // ---- epilogue: flux and obs_count of the winners, written in the reference's slot order ------------
        const int xi = px - p.x_start_min, yi = py - p.y_start_min;
#pragma unroll
        for (int r = 0; r < K; ++r) {
            if (r >= (int)p.results_per_pixel) break;  // warp-uniform
            Slot out{-FLT_MAX, 0.0f, 0, -1};
            const int c = top.cand[r];
            const float vx = (c >= 0) ? __ldg(cvx + c) : 0.0f, vy = (c >= 0) ? __ldg(cvy + c) : 0.0f;
            float lh = 0.0f, flux = 0.0f;
            int obs = 0;
            if (c >= 0) evaluate_exact<NB, false, 1>(sv, s_times, p, px, py, vx, vy, lh, flux, obs, nullptr);
            if (SIGMAG) {
                const bool need = c >= 0 && !(obs < p.min_observations || obs == 0 || lh < p.min_lh);
                int nv = obs;
                sigmag_for_lanes<NB, MAXT>(need, sv, s_times, px, py, vx, vy, p, lh, flux, nv);
            }
            if (c >= 0) out = Slot{lh, flux, obs, c};
            if (xi < sw && yi < sh) slots[((size_t)yi * sw + xi) * p.results_per_pixel + r] = out;
        }
    }
}

// ---- the deferred sigma-G clips: one warp per queued (pixel, candidate), all SMs ------------------------------
template <int NB, int MAXT>
__global__ void __launch_bounds__(256)
        clip_kernel(const StackView sv, const kb_params p, const float *__restrict__ cvx, const float *__restrict__ cvy,
                    const uint2 *__restrict__ queue, unsigned long long *cnt, unsigned long long cap,
                    ClipSurvivor *__restrict__ surv, unsigned long long surv_cap) {
    extern __shared__ double s_times[];
    for (int i = threadIdx.x; i < sv.T; i += blockDim.x) s_times[i] = sv.times[i];
    __syncthreads();
    // per-warp scratch for the ordered sum of the cooperative clip (MAXT pairs each), after the time table
    float2 *ws = (MAXT <= 256) ? reinterpret_cast<float2 *>(s_times + ((sv.T + 1) & ~1)) + (threadIdx.x >> 5) * MAXT
                               : nullptr;
    const unsigned long long n = min(cnt[0], cap);
    const int sw = p.x_start_max - p.x_start_min;
    const int lane = threadIdx.x & 31;
    const unsigned long long nwarps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    const unsigned long long i0 = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    uint2 nxt = (i0 < n) ? queue[i0] : make_uint2(0u, 0u);
    for (unsigned long long i = i0; i < n; i += nwarps) {
        const uint2 it = nxt;
        if (i + nwarps < n) nxt = queue[i + nwarps];  // in flight while this request is clipped
        const int x = p.x_start_min + (int)(it.x % (unsigned)sw), y = p.y_start_min + (int)(it.x / (unsigned)sw);
        const int c = (int)it.y;
        float lh = -1.0f, flux = -1.0f;  // what the reference keeps when there is no valid observation at all
        int nv = 0;
        sigmag_for_lanes<NB, MAXT>(lane == 0, sv, s_times, x, y, __ldg(cvx + c), __ldg(cvy + c), p, lh, flux, nv, ws);
        if (lane == 0) {
            // Queued candidates had an unclipped lh >= min_lh.  The clip replaces lh/flux when it ran
            // (nv >= max(min_obs, 1)); nv < min_obs is rejected (kernels.cu:318); nv == 0 keeps lh = flux = -1.
            if (!(nv < p.min_observations) && !(lh < p.min_lh)) {
                const unsigned long long k = atomicAdd(cnt + 1, 1ull);  // counts past surv_cap too: host reruns
                if (k < surv_cap) surv[k] = ClipSurvivor{it.x, c, lh, flux, nv};
            }
        }
    }
}

template <int NB>
static void launch_clip(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                        ClipScratch &clip, cudaStream_t st) {
    const size_t tbytes = sizeof(double) * (size_t)((sv.T + 1) & ~1);
    const int grid = 148 * 8;
#define KB_CLIP(MT)                                                                                              \
    clip_kernel<NB, MT><<<grid, 256, tbytes + ((MT) <= 256 ? sizeof(float2) * 8 * (MT) : 0), st>>>(              \
            sv, p, d_vx, d_vy, clip.queue, clip.counters, clip.cap, clip.surv, clip.surv_cap)
    if (sv.T <= 32) KB_CLIP(32);
    else if (sv.T <= 64) KB_CLIP(64);
    else if (sv.T <= 128) KB_CLIP(128);
    else if (sv.T <= 256) KB_CLIP(256);
    else KB_CLIP(kMaxTSigma);
#undef KB_CLIP
    KB_CUDA(cudaGetLastError());
}

template <int NB, bool ZM, bool COUNT, bool SIGMAG, int MAXT>
static void launch_k(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                     Slot *d_slots, dim3 grid, cudaStream_t st) {
#define KB_LAUNCH_T(KK)                                                                                         \
    do {                                                                                                        \
        auto kfn = search_tiled_kernel<NB, ZM, COUNT, SIGMAG, MAXT, KK>;                                                   \
        KB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->smem));       \
        kfn<<<grid, TNT, plan->smem, st>>>(plan->maps, plan->v, sv, p, d_vx, d_vy, d_slots);                    \
    } while (0)
    // One list length: the first K entries of a sorted top-8 list are the top-K list (K <= 8).
    KB_LAUNCH_T(8);
#undef KB_LAUNCH_T
}

template <int NB, bool ZM>
static void launch_nb(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st) {
    if (!plan->sigmag) {
        if (p.min_observations > 0) launch_k<NB, ZM, true, false, 1>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
        else launch_k<NB, ZM, false, false, 1>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    } else if (sv.T <= 32) launch_k<NB, ZM, false, true, 32>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    else if (sv.T <= 64) launch_k<NB, ZM, false, true, 64>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    // (a MAXT = 256 instantiation makes ptxas spill the accumulators of the gather loop; the serial sigma-G routine
    // of the large variant only runs for the K winners of a pixel and when the clip queue overflows)
    else launch_k<NB, ZM, false, true, kMaxTSigma>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
}

// One entry per stack encoding (defined in search_tiled_nb*.cu).
void launch_tiled_nb4(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st);
void launch_tiled_nb2(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st);
void launch_tiled_nb1(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st);
void launch_clip_nb4(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                     cudaStream_t st);
void launch_clip_nb2(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                     cudaStream_t st);
void launch_clip_nb1(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                     cudaStream_t st);

}  // namespace kb
