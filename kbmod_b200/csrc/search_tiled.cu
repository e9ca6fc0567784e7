// search_tiled.cu -- the throughput path of the search: a TMA-staged, time-tiled gather-reduce kernel.
//
// Decomposition (B200-first, not the reference's one-thread-per-pixel serial loop):
//   * A CTA owns a tile of 32 x BY start pixels (one consumer warp per row, lanes along x) and visits the
//     candidate list in *groups* of G candidates that are neighbours in velocity space.  Every consumer thread
//     keeps the G running (psi_sum, phi_sum, count) triples in registers.
//   * For a group g and an image t, all G trajectories of all tile pixels land inside one small box of the
//     image: origin (tile + min offset), size (32 + spread_x) x (BY + spread_y).  A dedicated producer warp
//     fetches those boxes with TMA (cp.async.bulk.tensor, 3-D map over [t][y][x-pairs]) into a ring of shared
//     memory stages, several images per stage, signalling mbarriers.  Out-of-image parts of a box are filled by
//     the TMA unit with NaN (float) / 0 (codes) -- exactly the reference's "outside the image = no data".
//   * Consumers wait on the stage's mbarrier and gather from shared memory: lane l of row r reads the pair at
//     (r + dy_j(t), l + dx_j(t)) -- x-adjacent lanes read adjacent pairs, so every 64-bit LDS is conflict-free.
//     The per-(candidate, image) integer offsets dx, dy come from a table built on the host with the
//     reference's fp64 arithmetic; the table arrives in the same stage by a 1-D bulk copy.
//   * After the last image of a group the G sums are turned into likelihoods.  A cheap conservative test
//     (MUFU.RSQ estimate against the current threshold) rejects almost everything; survivors take the exact
//     IEEE path, the rare sigma-G candidates re-gather their observations (search_device.cuh), and winners
//     enter the per-pixel top-K list held in registers (ordered by lh, then candidate index == the order the
//     reference's in-order strict-'>' insertion produces).
//   * Epilogue: the K winners per pixel are re-evaluated exactly to obtain flux and obs_count and written out.
//
// Exactness of the offset table: the reference computes ix = floor((x + fl(vx*t)) + 0.5) in fp64 per pixel.
// With |x|, |vx*t| < 2^22 the two roundings perturb the sum by < 2^-29, so ix = x + floor(fl(vx*t) + 0.5)
// unless fl(vx*t)+0.5 lies within 2^-27 of an integer without being exactly one.  Candidates with such an
// image are not tabulated; every thread evaluates them with the exact per-pixel routine instead.
#include <cuda.h>
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <limits.h>
#include <stddef.h>
#include <string.h>

#include <algorithm>
#include <memory>
#include <numeric>

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
};

struct TiledPlan {
    TmapSet maps;
    PlanView v{};
    void *d_block = nullptr;  // one allocation holding all tables
    size_t smem = 0;
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
    __device__ static __forceinline__ bool load(uint32_t addr, const StackView &sv, float &psi, float &phi) {
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(psi), "=f"(phi) : "r"(addr));
        // canonical copy: an invalid pair is (+0, +0) in zero-marker form, else (NaN, NaN)
        if (sv.zero_marker) return ((__float_as_uint(psi) | __float_as_uint(phi)) << 1) != 0u;
        return phi == phi;
    }
};
// Encoded stacks: the host verified (plan creation) that fmaf((float)(code-1), scale, min) reproduces the
// reference's fp64 decode for every code, so the gather decodes in fp32.
template <>
struct PairLoad<2> {
    static constexpr int kPairBytes = 4;
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
        const unsigned pixel_index = (unsigned)((py - p.y_start_min) * sw + (px - p.x_start_min));
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
#pragma unroll
                for (int j4 = 0; j4 < TG / 4; ++j4) {
                    const int4 o = ofs[j4];
                    const int off[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int j = j4 * 4 + q;
                        float a, b;
                        const bool ok = PairLoad<NB>::load(bk + off[q], sv, a, b);
                        if (ZM) {
                            psi[j] = __fadd_rn(psi[j], a);
                            phi[j] = __fadd_rn(phi[j], b);
                            if (COUNT && ok) cnt[j] += 1;
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

// ---- host: plan construction -------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        KB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres));
        if (!sym || qres != cudaDriverEntryPointSuccess) throw Error("cuTensorMapEncodeTiled is not available");
        fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

// floor(fl(fl(v*t) + 0.5)) plus the "cannot be tabulated" test described in the file header.
static inline bool tab_offset(float v, double t, int &off) {
    const double prod = (double)v * t;
    if (!(fabs(prod) < 4194304.0)) {
        off = 0;
        return false;
    }
    const double q = prod + 0.5;
    const double fl = floor(q);
    off = (int)fl;
    const double r = q - fl;
    const double dist = (r < 0.5) ? r : 1.0 - r;
    if (dist > 7.450580596923828e-09) return true;          // 2^-27
    return r == 0.0 && (q - 0.5) == prod;                   // exactly on a boundary: no rounding anywhere
}

static bool decode_is_exact_in_fp32(int nb, float scale, float mn) {
    const unsigned n = (nb == 1) ? 255u : 65535u;
    for (unsigned code = 1; code <= n; ++code) {
        const float ref = (float)(((double)(float)code - 1.0) * (double)scale + (double)mn);
        const float fast = fmaf((float)(code - 1), scale, mn);
        if (memcmp(&ref, &fast, 4) != 0) return false;
    }
    return true;
}

TiledPlan *tiled_plan_create(const StackView &sv, const kb_params &p, const float *h_vx, const float *h_vy,
                             uint64_t n_cand, cudaStream_t st, std::string *why_not) {
    auto no = [&](const char *why) -> TiledPlan * {
        if (why_not) *why_not = why;
        return nullptr;
    };
    const int T = sv.T, K = (int)p.results_per_pixel;
    if (K > kMaxRegK) return no("results_per_pixel > 8");
    if (n_cand == 0) return no("no candidates");
    if (p.do_sigmag_filter && T > kMaxTSigma) return no("too many images for sigma-G");
    if (sizeof(double) * (size_t)T > 16 * 1024) return no("too many images for the tiled kernel's time table");
    if (sv.nb != 4) {
        if (!decode_is_exact_in_fp32(sv.nb, sv.psi_scale, sv.psi_min) ||
            !decode_is_exact_in_fp32(sv.nb, sv.phi_scale, sv.phi_min))
            return no("fp32 decode would not be bit-exact for this scale/offset");
    }
    std::vector<double> times(T);
    KB_CUDA(cudaMemcpyAsync(times.data(), sv.times, sizeof(double) * T, cudaMemcpyDeviceToHost, st));
    KB_CUDA(cudaStreamSynchronize(st));
    double tabs = 0.0;
    for (double t : times) {
        if (!std::isfinite(t)) return no("non-finite time");
        tabs = std::max(tabs, fabs(t));
    }
    const int PB = 2 * sv.nb;
    const int bw_limit = (sv.nb == 4) ? 128 : 256;  // TMA box: at most 256 elements along x

    // ---- candidates whose offsets cannot be tabulated exactly go to the exact list ---------------------------------
    const int N = (int)n_cand;
    std::vector<int> exact, tabulated;
    tabulated.reserve(N);
    for (int c = 0; c < N; ++c) {
        bool ok = true;
        for (int t = 0; t < T && ok; ++t) {
            int o;
            ok = tab_offset(h_vx[c], times[t], o) && tab_offset(h_vy[c], times[t], o);
        }
        (ok ? tabulated : exact).push_back(c);
    }
    const int NTab = (int)tabulated.size();

    // ---- group the rest: k-d style recursive bisection of velocity space into leaves of <= TG candidates ---------
    // A group's box is (TBX + sx) x (TBY + sy) pixels with sx, sy ~ velocity extent * max|t|; each cut goes
    // across the direction that shrinks that box most, at a multiple of TG so the leaves come out full.
    const int XA = 16 / PB;                          // TMA: box origin must be 16-byte aligned in global memory
    const int bw_round = (sv.nb == 1) ? 16 : 8;      // keeps every TMA destination 128-byte aligned
    const int sy_lim = 40;
    const int rows_lim = (TBY + sy_lim + TCH - 1) / TCH * TCH;
    const int w_cap = std::min(TBX + bw_round + (NMAPS - 1) * bw_round, bw_limit);  // widest tensor-map box
    const int sx_lim = std::max(8, std::min(STAGE_BYTES / PB / rows_lim, w_cap) - TBX - XA - bw_round);
    std::vector<int> group_start;
    int max_sx = 0, max_sy = 0;
    {
        struct Seg { int lo, hi; };
        std::vector<Seg> stack_;
        stack_.push_back({0, NTab});
        std::vector<Seg> leaves;
        while (!stack_.empty()) {
            const Seg sg = stack_.back();
            stack_.pop_back();
            const int n = sg.hi - sg.lo;
            if (n <= 0) continue;
            float x0 = FLT_MAX, x1 = -FLT_MAX, y0 = FLT_MAX, y1 = -FLT_MAX;
            for (int i = sg.lo; i < sg.hi; ++i) {
                const int c = tabulated[i];
                x0 = std::min(x0, h_vx[c]); x1 = std::max(x1, h_vx[c]);
                y0 = std::min(y0, h_vy[c]); y1 = std::max(y1, h_vy[c]);
            }
            const int sx = (int)ceil(((double)x1 - x0) * tabs) + 1, sy = (int)ceil(((double)y1 - y0) * tabs) + 1;
            if (n <= TG && sx <= sx_lim && sy <= sy_lim) {
                leaves.push_back(sg);
                max_sx = std::max(max_sx, sx);
                max_sy = std::max(max_sy, sy);
                continue;
            }
            if (n == 1) return no("a single candidate does not fit a stage");
            // cut across x if that leaves the smaller box, else across y
            const double ax = (TBX + 0.5 * sx) * (double)(TBY + sy), ay = (TBX + (double)sx) * (TBY + 0.5 * sy);
            const bool cut_x = (sx > sx_lim) || (!(sy > sy_lim) && ax <= ay);
            int left = (n > TG) ? ((n / 2 + TG - 1) / TG) * TG : n / 2;
            if (left >= n) left = n / 2;
            auto first = tabulated.begin() + sg.lo;
            if (cut_x)
                std::nth_element(first, first + left, tabulated.begin() + sg.hi, [&](int a, int b) {
                    return h_vx[a] != h_vx[b] ? h_vx[a] < h_vx[b] : (h_vy[a] != h_vy[b] ? h_vy[a] < h_vy[b] : a < b);
                });
            else
                std::nth_element(first, first + left, tabulated.begin() + sg.hi, [&](int a, int b) {
                    return h_vy[a] != h_vy[b] ? h_vy[a] < h_vy[b] : (h_vx[a] != h_vx[b] ? h_vx[a] < h_vx[b] : a < b);
                });
            stack_.push_back({sg.lo + left, sg.hi});  // popped second: leaves come out in in-order traversal
            stack_.push_back({sg.lo, sg.lo + left});
        }
        for (const Seg &l : leaves) group_start.push_back(l.lo);
        group_start.push_back(NTab);
    }
    const int n_groups = (int)group_start.size() - 1;
    if (n_groups == 0 && exact.empty()) return no("no candidates");
    // Box widths come in steps (one tensor map each); a box is as wide as its own (group, image) needs.
    const int w_step = bw_round, w_base = TBX + w_step;
    const int w_max = std::min(w_base + (NMAPS - 1) * w_step, bw_limit);
    if (TBX + max_sx + XA - 1 > w_max) return no("velocity spread too large for the TMA box");
    const int rmax = (TBY + max_sy + TCH - 1) / TCH * TCH;
    const int max_ch = rmax / TCH;

    // ---- stage records, group by group (greedy packing of consecutive images into <= STAGE_BYTES) ---------------
    std::vector<std::vector<StageRec>> per_group(std::max(n_groups, 1));
    std::vector<int> cand((size_t)std::max(n_groups, 1) * TG, -1);
    bool overflow = false;
#pragma omp parallel for schedule(dynamic, 8)
    for (int g = 0; g < n_groups; ++g) {
        const int b = group_start[g], n = group_start[g + 1] - b;
        int dx[TG], dy[TG];
        for (int j = 0; j < n; ++j) cand[(size_t)g * TG + j] = tabulated[b + j];
        std::vector<StageRec> &out = per_group[g];
        StageRec cur;
        memset(&cur, 0, sizeof(cur));
        int nt = 0, used = 0, tx = 0, t0 = 0;
        auto flush = [&](bool last) {
            cur.meta = make_int4(nt | ((last ? 1 : 0) << 8), g, t0, tx + (int)sizeof(StageRec));
            out.push_back(cur);
            memset(&cur, 0, sizeof(cur));
            nt = used = tx = 0;
        };
        for (int t = 0; t < T; ++t) {
            int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
            for (int j = 0; j < n; ++j) {
                const int c = tabulated[b + j];
                tab_offset(h_vx[c], times[t], dx[j]);
                tab_offset(h_vy[c], times[t], dy[j]);
                mnx = std::min(mnx, dx[j]); mxx = std::max(mxx, dx[j]);
                mny = std::min(mny, dy[j]); mxy = std::max(mxy, dy[j]);
            }
            mnx -= ((p.x_start_min + mnx) % XA + XA) % XA;  // tile origins are multiples of 32 past x_start_min
            const int need_w = TBX + (mxx - mnx), nrows = TBY + (mxy - mny);
            const int widx = std::max(0, (need_w - w_base + w_step - 1) / w_step);
            const int pitch = w_base + w_step * widx;
            const int rows_pad = (nrows + TCH - 1) / TCH * TCH;
            const int bytes = rows_pad * pitch * PB;
            if (pitch > w_max || widx >= NMAPS || nrows > rmax || bytes > STAGE_BYTES) {
                overflow = true;
                break;
            }
            if (nt == TMAXTC || used + bytes > STAGE_BYTES) flush(false);
            if (nt == 0) t0 = t;
            cur.box[nt] = make_int4(mnx, mny, nrows | (widx << 16), used);
            for (int j = 0; j < TG; ++j)
                cur.offs[nt][j] = (j < n) ? ((dy[j] - mny) * pitch + (dx[j] - mnx)) * PB : 0;
            used += bytes;
            tx += bytes;
            ++nt;
        }
        if (nt > 0) flush(true);
    }
    if (overflow) return no("velocity spread too large for a shared-memory stage");
    std::vector<StageRec> stages;
    {
        size_t total = 0;
        for (auto &v : per_group) total += v.size();
        if (total >= (1ull << 31)) return no("too many stages");
        stages.reserve(std::max<size_t>(total, 1));
        for (auto &v : per_group) {
            stages.insert(stages.end(), v.begin(), v.end());
            std::vector<StageRec>().swap(v);
        }
    }
    const int n_stages = (int)stages.size();
    if (getenv("KB_DEBUG"))
        fprintf(stderr, "[kbmod_b200] tiled plan: N=%d tab=%d exact=%zu groups=%d stages=%d max_sx=%d max_sy=%d "
                "w_base=%d w_step=%d rmax=%d sx_lim=%d sy_lim=%d\n", N, NTab, exact.size(), n_groups, n_stages,
                max_sx, max_sy, w_base, w_step, rmax, sx_lim, sy_lim);

    std::unique_ptr<TiledPlan> plan(new TiledPlan());
    plan->nb = sv.nb;
    plan->K = K;
    plan->sigmag = p.do_sigmag_filter != 0;
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t b_st = up(std::max<size_t>(stages.size(), 1) * sizeof(StageRec)), b_cand = up(cand.size() * 4),
                 b_ex = up(std::max<size_t>(exact.size(), 1) * 4);
    KB_CUDA(cudaMalloc(&plan->d_block, b_st + b_cand + b_ex));
    uint8_t *d = (uint8_t *)plan->d_block;
    if (!stages.empty())
        KB_CUDA(cudaMemcpyAsync(d, stages.data(), stages.size() * sizeof(StageRec), cudaMemcpyHostToDevice, st));
    KB_CUDA(cudaMemcpyAsync(d + b_st, cand.data(), cand.size() * 4, cudaMemcpyHostToDevice, st));
    if (!exact.empty())
        KB_CUDA(cudaMemcpyAsync(d + b_st + b_cand, exact.data(), exact.size() * 4, cudaMemcpyHostToDevice, st));
    KB_CUDA(cudaStreamSynchronize(st));  // the host vectors die with this scope
    PlanView &v = plan->v;
    v.stages = (const StageRec *)d;
    v.cand = (const int *)(d + b_st);
    v.exact = (const int *)(d + b_st + b_cand);
    v.n_exact = (int)exact.size();
    v.n_groups = n_groups;
    v.n_stages = n_stages;
    v.w_base = w_base;
    v.w_step = w_step;
    v.max_ch = max_ch;
    v.T = T;
    v.clipq = nullptr;
    v.clip_cnt = nullptr;
    v.clip_cap = 0;
    plan->smem = (size_t)TS * (STAGE_BYTES + HDR_BYTES) + 2 * TS * 8 + sizeof(double) * (size_t)T;

    // ---- tensor maps over the search copy: dims (x elements, rows, images), one per box width ------------------
    const bool f32 = sv.nb == 4;
    const CUtensorMapDataType dt = f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                       : (sv.nb == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT16);
    const cuuint64_t gdim[3] = {(cuuint64_t)(f32 ? 2 * sv.W : sv.W), (cuuint64_t)sv.H, (cuuint64_t)sv.T};
    const cuuint64_t gstr[2] = {(cuuint64_t)sv.Wp * PB, (cuuint64_t)sv.Wp * PB * sv.H};
    const cuuint32_t estr[3] = {1, 1, 1};
    for (int i = 0; i < NMAPS; ++i) {
        const int bw = std::min(w_base + i * w_step, w_max);
        const cuuint32_t box[3] = {(cuuint32_t)(f32 ? 2 * bw : bw), (cuuint32_t)TCH, 1};
        const CUresult r = encode_fn()(&plan->maps.m[i], dt, 3, const_cast<void *>(sv.data), gdim, gstr, box, estr,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                       (f32 && !sv.zero_marker) ? CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA
                                                                : CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            cudaFree(plan->d_block);
            plan->d_block = nullptr;
            return no("cuTensorMapEncodeTiled failed");
        }
    }
    return plan.release();
}

void tiled_plan_destroy(TiledPlan *plan) {
    if (!plan) return;
    if (plan->d_block) cudaFree(plan->d_block);
    delete plan;
}

// ---- the deferred sigma-G clips: one warp per queued (pixel, candidate), all SMs ------------------------------
template <int NB, int MAXT>
__global__ void __launch_bounds__(256)
        clip_kernel(const StackView sv, const kb_params p, const float *__restrict__ cvx, const float *__restrict__ cvy,
                    const uint2 *__restrict__ queue, unsigned long long *cnt, unsigned long long cap,
                    ClipSurvivor *__restrict__ surv) {
    extern __shared__ double s_times[];
    for (int i = threadIdx.x; i < sv.T; i += blockDim.x) s_times[i] = sv.times[i];
    __syncthreads();
    const unsigned long long n = min(cnt[0], cap);
    const int sw = p.x_start_max - p.x_start_min;
    const int lane = threadIdx.x & 31;
    const unsigned long long nwarps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    for (unsigned long long i = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n;
         i += nwarps) {
        const uint2 it = queue[i];
        const int x = p.x_start_min + (int)(it.x % (unsigned)sw), y = p.y_start_min + (int)(it.x / (unsigned)sw);
        const int c = (int)it.y;
        float lh = -1.0f, flux = -1.0f;  // what the reference keeps when there is no valid observation at all
        int nv = 0;
        sigmag_for_lanes<NB, MAXT>(lane == 0, sv, s_times, x, y, __ldg(cvx + c), __ldg(cvy + c), p, lh, flux, nv);
        if (lane == 0) {
            // Queued candidates had an unclipped lh >= min_lh.  The clip replaces lh/flux when it ran
            // (nv >= max(min_obs, 1)); nv < min_obs is rejected (kernels.cu:318); nv == 0 keeps lh = flux = -1.
            if (!(nv < p.min_observations) && !(lh < p.min_lh)) {
                const unsigned long long k = atomicAdd(cnt + 1, 1ull);
                surv[k] = ClipSurvivor{it.x, c, lh, flux, nv};
            }
        }
    }
}

template <int NB>
static void launch_clip(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                        ClipScratch &clip, cudaStream_t st) {
    const size_t smem = sizeof(double) * (size_t)sv.T;
    const int grid = 148 * 8;
#define KB_CLIP(MT) clip_kernel<NB, MT><<<grid, 256, smem, st>>>(sv, p, d_vx, d_vy, clip.queue, clip.counters, clip.cap, clip.surv)
    if (sv.T <= 32) KB_CLIP(32);
    else if (sv.T <= 64) KB_CLIP(64);
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
    const int K = plan->K;
    if (K <= 1) KB_LAUNCH_T(1);
    else if (K <= 2) KB_LAUNCH_T(2);
    else if (K <= 4) KB_LAUNCH_T(4);
    else KB_LAUNCH_T(8);
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
    else if (sv.T <= 256) launch_k<NB, ZM, false, true, 256>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    else launch_k<NB, ZM, false, true, kMaxTSigma>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
}

int launch_search_tiled(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx,
                        const float *d_vy, Slot *d_slots, ClipScratch &clip, ResultScratch &scr, cudaStream_t st) {
    const int sw = p.x_start_max - p.x_start_min, sh = p.y_start_max - p.y_start_min;
    dim3 grid((sw + TBX - 1) / TBX, (sh + TBY - 1) / TBY);
    if (grid.y > 65535) throw Error("search area too tall for the tiled kernel grid");
    if (plan->sigmag) {
        // Queue for the deferred clips: every (pixel, candidate) pair if that is small, else 32M requests (256 MB);
        // the kernel clips in line once the queue is full, so the capacity only affects speed.
        const uint64_t want = std::min<uint64_t>((uint64_t)sw * sh * (uint64_t)std::max(plan->v.n_groups * TG, 1),
                                                 32ull << 20);
        if (!clip.counters) KB_CUDA(cudaMalloc(&clip.counters, 4 * sizeof(unsigned long long)));
        if (clip.cap < want) {
            if (clip.queue) cudaFree(clip.queue);
            if (clip.surv) cudaFree(clip.surv);
            clip.queue = nullptr;
            clip.surv = nullptr;
            clip.cap = 0;
            KB_CUDA(cudaMalloc(&clip.queue, sizeof(uint2) * want));
            KB_CUDA(cudaMalloc(&clip.surv, sizeof(ClipSurvivor) * want));
            clip.cap = want;
        }
        KB_CUDA(cudaMemsetAsync(clip.counters, 0, 4 * sizeof(unsigned long long), st));
        plan->v.clipq = clip.queue;
        plan->v.clip_cnt = clip.counters;
        plan->v.clip_cap = clip.cap;
    }
    switch (sv.nb) {
        case 4:
            if (sv.zero_marker) launch_nb<4, true>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
            else launch_nb<4, false>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
            break;
        case 2: launch_nb<2, false>(plan, sv, p, d_vx, d_vy, d_slots, grid, st); break;
        case 1: launch_nb<1, false>(plan, sv, p, d_vx, d_vy, d_slots, grid, st); break;
        default: throw Error("bad num_bytes");
    }
    KB_CUDA(cudaGetLastError());
    int launches = 1;
    if (plan->sigmag) {
        switch (sv.nb) {
            case 4: launch_clip<4>(sv, p, d_vx, d_vy, clip, st); break;
            case 2: launch_clip<2>(sv, p, d_vx, d_vy, clip, st); break;
            default: launch_clip<1>(sv, p, d_vx, d_vy, clip, st); break;
        }
        ++launches;
        unsigned long long cnt[4];
        KB_CUDA(cudaMemcpyAsync(cnt, clip.counters, sizeof(cnt), cudaMemcpyDeviceToHost, st));
        KB_CUDA(cudaStreamSynchronize(st));
        if (getenv("KB_DEBUG"))
            fprintf(stderr, "[kbmod_b200] sigma-G clips: requests=%llu (queue cap %llu) survivors=%llu in-line=%llu\n",
                    cnt[0], (unsigned long long)clip.cap, cnt[1], cnt[2]);
        launches += merge_clip_survivors(clip.surv, cnt[1], d_slots, (int)p.results_per_pixel, scr, st);
    }
    return launches;
}

}  // namespace kb
