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

#ifndef KB_TBY
#define KB_TBY 15
#endif
#ifndef KB_TCH
#define KB_TCH 8
#endif
#ifndef KB_TPX
#define KB_TPX 2
#endif
static constexpr int TPX = KB_TPX;     // start pixels per lane along x: x, x + 32, ...
static constexpr int TBX = 32 * TPX;   // tile width
static constexpr int TBY = KB_TBY;     // tile height (consumer warps); +1 producer warp
static constexpr int TNT = (TBY + 1) * 32;
static constexpr int TGMAX = 16;       // most candidates per group (record layouts are sized for it)
// Candidates per group.  Sixteen halve the per-image and per-stage bookkeeping of the sweep; the sigma-G-off variants
// also keep eight lower bounds (and, with min_obs > 0, sixteen observation counters) per pixel in registers and stay
// at eight.
__host__ __device__ constexpr int group_size(bool sigmag) { return sigmag ? 16 : 8; }
#ifndef KB_TS
#define KB_TS 4
#endif
#ifndef KB_STAGE_KB
#define KB_STAGE_KB 48
#endif
static constexpr int TS = KB_TS;       // pipeline stages
static constexpr int TCH = KB_TCH;     // rows per TMA box
static constexpr int TMAXTC = 8;       // max images per stage
static constexpr int STAGE_BYTES = KB_STAGE_KB * 1024;
static constexpr int NMAPS = 16;       // tensor maps with box widths base, base+step, ...
static constexpr int kScreenPairBytes = 4;  // one 32-bit word of the screening copy per pixel pair

// One pipeline stage as the host lays it out (and as it lands in shared memory by one bulk copy):
//   meta   {nt | last << 8, group, t0, tx_bytes}
//   box[k] {ox, oy, nrows | widx << 8 | plane << 12, slot_off}   for the nt records of the stage (k < TMAXTC): the
//          producer's view; plane = image index, or T + run index for a run's summed plane
//   img[k] what a consumer needs for image k: where its box starts in the stage, the box's row pitch in bytes, and
//          offs[j], the byte offset of candidate j's pixel inside the box relative to the thread's own pixel
struct ImageRec {
    int slot_off, pitch_bytes;
    int mult_lo, mult_hi;   // IDP.2A multipliers of the record: m and m << 8, m = images the record stands for
    int offs[TGMAX];
};
struct StageRec {
    int4 meta;
    int4 box[TMAXTC];
    ImageRec img[TMAXTC];
};
static_assert(sizeof(StageRec) % 16 == 0 && offsetof(StageRec, img) % 16 == 0 && sizeof(ImageRec) % 16 == 0,
              "bulk copies and 128-bit shared loads need 16-byte units");
static constexpr int HDR_BYTES = (sizeof(StageRec) + 127) / 128 * 128;

struct TmapSet {
    CUtensorMap m[NMAPS];
};

struct PlanView {
    const StageRec *stages;  // [n_stages]
    const StageRec *stages_plain;  // schedule of single images (== stages unless the stack has run planes and the search
    int n_stages_plain;            // counts observations): used by the tiles that have to count
    const int *cand;         // [n_groups][group size]  candidate index or -1
    const int *exact;        // candidates evaluated by the exact routine
    int n_exact, n_groups, n_stages, w_base, w_step, max_ch, T;
    uint2 *clipq;                   // deferred sigma-G requests {start-pixel index, candidate}
    unsigned long long *clip_cnt;  // [0] requests, [1] survivors, [2] clipped in line (queue full)
    unsigned long long clip_cap;
    unsigned pixel_base;            // start-pixel index of this launch's first pixel within the whole search area
    // screening: value of one quantisation step and the bound of |reference float sum - step * integer sum|
    float step_psi, step_phi, e_psi, e_phi;
    // sigma-G searches: the bar is min_lh for every pixel; bar_c3 = step_phi * bar^2, bar_c4 = -e_phi * bar^2 (both
    // scaled by 0.99999), bar_ok = 0 when min_lh is not a positive squarable number (then nothing is screened out)
    float bar_c3, bar_c4;
    int bar_ok;
    // sigma-G off: requests for the deferred exact evaluation {pixel, candidate, upper bound of lh} (the counters and
    // the capacity are clip_cnt / clip_cap), and every pixel's final bar (its 8th largest lower bound)
    ExactReq *exq;
    float *bar_out;
    // extent of the schedule's boxes relative to a tile's origin: [ext_x0, ext_x1) x [ext_y0, ext_y1) pixels
    int ext_x0, ext_x1, ext_y0, ext_y1;
    // optional launch order of the tiles (1-D grid): tile_order[blockIdx.x] = {tile column, tile row}.  Tiles that
    // count observations follow the longer single-image schedule and are started first.
    const uint2 *tile_order;
};

struct TiledPlan {
    TmapSet maps;
    PlanView v{};
    void *d_block = nullptr;  // one allocation holding all tables
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the gather kernel
    float ms_accum = 0.0f;    // gather-kernel time of the last search (summed over its chunks)
    unsigned long long requests_accum = 0;  // deferred exact evaluations the last search queued
    uint2 *d_order = nullptr; // tile launch order of the current launch (counting searches over run planes)
    size_t order_cap = 0;
    bool order_tiles = false; // the schedule has a longer single-image variant for the tiles that count
    size_t smem = 0;
    int reach_hi = 0;  // a tile at row y0 reads stack rows below y0 + reach_hi (largest box bottom, padded)
    int nb = 4, K = 8;
    bool sigmag = false;
    int tg = 8;   // candidates per group of this schedule (group_size(sigmag))
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
// The wait suspends the warp in hardware for up to the hinted time before it reports "not yet" (a spinning warp
// takes issue slots from the warps that have work).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_LOOP:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
            "@p bra DONE;\n\t"
            "bra WAIT_LOOP;\n\t"
            "DONE:\n\t}" ::"r"(bar),
            "r"(parity), "r"(20000u)
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

template <int OFF>
__device__ __forceinline__ int4 lds_v4(uint32_t addr) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4+%5];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr), "n"(OFF));
    return v;
}
template <int OFF>
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t w;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(w) : "r"(addr), "n"(OFF));
    return w;
}

// bounds of the reference's lh = psi_sum / sqrt(phi_sum) (-1 for phi_sum <= 0) from the bounds of the two sums; the
// 4e-6 covers MUFU.RSQ and the reference's own division and square-root roundings
__device__ __forceinline__ void lh_bounds(float ps, float ph, float e_psi, float e_phi, float &lb, float &ub) {
    const float psi_hi = ps + e_psi, psi_lo = ps - e_psi, phi_hi = ph + e_phi, phi_lo = ph - e_phi;
    if (!(phi_lo > 0.0f)) {   // phi_sum may be <= 0 (lh = -1) or tiny (lh huge): no bounds
        lb = -CUDART_INF_F;
        ub = CUDART_INF_F;
        return;
    }
    const float rl = rsqrtf(phi_lo), rh = rsqrtf(phi_hi);
    ub = (psi_hi >= 0.0f) ? psi_hi * rl * 1.000004f : psi_hi * rh * 0.999996f;
    lb = (psi_lo >= 0.0f) ? psi_lo * rh * 0.999996f : psi_lo * rl * 1.000004f;
}

// The gather of one stage: for each of its nt records, the words of the lane's TPX pixels for the TG candidates.
// DO_COUNT also counts the valid observations (word != 0: the quantiser never maps a valid pair to 0) with one compare
// and one predicated add in place (the compiler's select form costs a move more).  A run record counts as its m
// images: it is only counted in stacks without invalid pixels, where a word is 0 exactly when the (common) pixel of
// the run's images lies outside the image.
template <int TG, bool DO_COUNT, int CT>
__device__ __forceinline__ void gather_images(int nt, uint32_t rec, uint32_t lane_base, int warp, int (&P)[TPX][TG],
                                              int (&F)[TPX][TG], int (&cnt)[CT][CT == 1 ? 1 : TG]) {
    constexpr int PB = kScreenPairBytes;
    for (int k = 0; k < nt; ++k, rec += (uint32_t)sizeof(ImageRec)) {
        const int4 ir = lds_v4<0>(rec);   // {slot_off, pitch_bytes, multiplier for psi, multiplier for phi}
        const uint32_t bk = lane_base + (uint32_t)ir.x + (uint32_t)(warp * ir.y);
#pragma unroll
        for (int j4 = 0; j4 < TG / 4; ++j4) {
            const int4 o = (j4 == 0) ? lds_v4<16>(rec) : ((j4 == 1) ? lds_v4<32>(rec) : ((j4 == 2) ? lds_v4<48>(rec) : lds_v4<64>(rec)));
            static_assert(TG <= TGMAX && TGMAX == 16, "offset loads above cover 16 candidates");
            const int off[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int j = j4 * 4 + q;
                const uint32_t addr = bk + off[q];
                uint32_t w[TPX];
                w[0] = lds_u32<0>(addr);
                if (TPX > 1) w[TPX - 1] = lds_u32<32 * PB>(addr);
#pragma unroll
                for (int h = 0; h < TPX; ++h) {
                    P[h][j] = __dp2a_lo((int)w[h], ir.z, P[h][j]);   // += m * low half  (psi, signed 16 bit)
                    F[h][j] = __dp2a_lo((int)w[h], ir.w, F[h][j]);   // += m * high half (phi)
                    if (DO_COUNT) {   // + the number of images the record stands for
                        asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p add.s32 %0, %0, %2;\n\t}"
                            : "+r"(cnt[h][j])
                            : "r"(w[h]), "r"(ir.z));
                    }
                }
            }
        }
    }
}

// ---- the kernel -----------------------------------------------------------------------------------------
// The gather loop works on the SCREENING copy of the stack (StackView::screen): one 32-bit word per pixel pair, psi
// and phi as signed 16-bit multiples of sv.q_step_psi / sv.q_step_phi, 0 for an invalid pair (and for everything
// outside the image: the TMA unit fills with zeros).  Per (start pixel, candidate, image) it costs one LDS.32 and
// two IDP.2A (dp2a picks one 16-bit half and adds it to a 32-bit sum), and the sums are EXACT integers, so
//     | psi_sum_ref - step_psi * P |  <=  pv.e_psi,      | phi_sum_ref - step_phi * F |  <=  pv.e_phi
// hold rigorously for the reference's float sums (quantisation <= 0.502 / 1.002 steps per observation, float
// summation error <= T^2 * 2^-24 * max|value|; screen_bounds() in search_tiled.cu).  From them an upper bound of the
// candidate's likelihood follows; a candidate whose bound is below the bar (min_lh with sigma-G on, else the pixel's
// current K-th best) cannot appear in the results and is dropped.  Everything else -- a tiny fraction -- is
// evaluated EXACTLY from the search copy with the reference's own arithmetic: through the deferred clip queue when
// sigma-G is on (clip_kernel re-gathers, checks the unclipped likelihood, clips), in line otherwise.  The result is
// therefore identical to evaluating every candidate exactly; the screening only decides how much work that takes.
// With sigma-G off the bar is the pixel's K-th best lh, which only exact evaluation knows; the sweep uses bounds
// instead: the 8th largest LOWER bound seen so far is a valid bar (eight candidates are known to reach it), and a
// candidate whose UPPER bound reaches the bar is queued with that bound.  exact_kernel afterwards evaluates the
// queued candidates whose bound still reaches the pixel's final bar -- about ten per pixel -- and the survivors are
// merged into the slots by (lh, candidate index).
// NB is the encoding of the exact search copy (4 = float, 2 = uint16, 1 = uint8), used by the exact routines only.
// COUNT: keep a per-candidate observation count (sigma-G off and min_obs > 0), so candidates with too few
// observations are dropped before the exact evaluation.
template <int NB, bool COUNT, bool SIGMAG, int MAXT, int K>
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
    const uint2 tile = pv.tile_order ? pv.tile_order[blockIdx.x] : make_uint2(blockIdx.x, blockIdx.y);
    const int tile_x0 = p.x_start_min + (int)tile.x * TBX, tile_y0 = p.y_start_min + (int)tile.y * TBY;
    constexpr int PB = kScreenPairBytes;
    constexpr int TG = group_size(SIGMAG);
    // Observations only have to be counted where some can be invalid: a stack with invalid pixels, or a tile from
    // which some box of the schedule reaches over the image's edge (CTA-uniform).  In a stack with invalid pixels the
    // counting follows the schedule of single images (a run's summed word cannot say how many of its images had
    // data there); the host only builds that second schedule for such stacks.
    const bool count_here = COUNT && !(sv.all_valid && tile_x0 + pv.ext_x0 >= 0 && tile_x0 + pv.ext_x1 <= sv.W &&
                                       tile_y0 + pv.ext_y0 >= 0 && tile_y0 + pv.ext_y1 <= sv.H);
    const StageRec *const stage_list = count_here ? pv.stages_plain : pv.stages;
    const int n_stages = count_here ? pv.n_stages_plain : pv.n_stages;

    if (warp == TBY) {
        // ===================== producer warp: TMA box loads + the stage record =================================
        // Lane l < 9 keeps int4 #l of the *next* stage record (meta + 8 boxes) in a register, loaded one stage ahead
        // so the global-memory latency overlaps the wait for a free slot; fields are handed out by shuffles.
        const unsigned FULLM = 0xffffffffu;
        auto fetch = [&](int s) -> int4 {
            if (lane < 1 + TMAXTC && s < n_stages) return __ldg(reinterpret_cast<const int4 *>(stage_list + s) + lane);
            return make_int4(0, 0, 0, 0);
        };
        int4 nxt = fetch(0);
        for (int s = 0; s < n_stages; ++s) {
            const int slot = s % TS;
            const uint32_t full = bar0 + 8 * slot, empty = bar0 + 8 * (TS + slot);
            const int4 cur = nxt;
            nxt = fetch(s + 1);
            if (s >= TS) mbar_wait(empty, ((s / TS) - 1) & 1);
            const int nt = __shfl_sync(FULLM, cur.x, 0) & 0xff;
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
                const int nrows = bxz & 0xff, widx = (bxz >> 8) & 0xf, plane = bxz >> 12;
                if (i < total && ch * TCH < nrows) {
                    const int pitch = pv.w_base + pv.w_step * widx;
                    tma_load_3d(stage_base + bxw + ch * TCH * pitch * PB, &maps.m[widx], tile_x0 + bxx,
                                tile_y0 + bxy + ch * TCH, plane, full);
                }
            }
            if (lane == 0)
                bulk_load(smem_u32(s_hdr + (size_t)slot * HDR_BYTES), stage_list + s, (uint32_t)sizeof(StageRec), full);
        }
    } else {
        // ===================== consumer warps: gather, reduce, rank ========================================
        // A lane owns TPX start pixels of its row: x, x + 32, ...  Their words for one (candidate, image) lie 128
        // bytes apart in the staged box, so one address serves all of them (immediate offsets).
        const int py = tile_y0 + warp;
        int px[TPX];
        bool in_bounds[TPX];
        unsigned pixel_index[TPX];
#pragma unroll
        for (int h = 0; h < TPX; ++h) {
            px[h] = tile_x0 + lane + 32 * h;
            in_bounds[h] = (px[h] - p.x_start_min) < sw && (py - p.y_start_min) < sh;
            pixel_index[h] = pv.pixel_base + (unsigned)((py - p.y_start_min) * sw + (px[h] - p.x_start_min));
        }
        const uint32_t hdr0 = smem_u32(s_hdr), stage0 = smem_u32(s_stage);
        int P[TPX][TG], F[TPX][TG];
        int cnt[COUNT ? TPX : 1][COUNT ? TG : 1];
        TopK<K> top[TPX];
        float lb8[SIGMAG ? 1 : TPX][8];   // sigma-G off: the 8 largest lower bounds of lh so far, descending
#pragma unroll
        for (int h = 0; h < TPX; ++h) {
            if (!SIGMAG) {
#pragma unroll
                for (int i = 0; i < 8; ++i) lb8[h][i] = -CUDART_INF_F;
            }
#pragma unroll
            for (int j = 0; j < TG; ++j) {
                P[h][j] = 0;
                F[h][j] = 0;
                if (COUNT) cnt[h][j] = 0;
            }
        }
        for (int s = 0; s < n_stages; ++s) {
            const int slot = s % TS;
            const uint32_t full = bar0 + 8 * slot, empty = bar0 + 8 * (TS + slot);
            mbar_wait(full, (s / TS) & 1);
            const uint32_t hdr_a = hdr0 + (uint32_t)slot * HDR_BYTES;
            const int4 meta = lds_v4<0>(hdr_a);
            const int nt = meta.x & 0xff;
            const uint32_t lane_base = stage0 + (uint32_t)slot * STAGE_BYTES + (uint32_t)lane * PB;
            uint32_t rec = hdr_a + (uint32_t)offsetof(StageRec, img);
            if (count_here) gather_images<TG, true, COUNT ? TPX : 1>(nt, rec, lane_base, warp, P, F, cnt);
            else gather_images<TG, false, COUNT ? TPX : 1>(nt, rec, lane_base, warp, P, F, cnt);
            __syncwarp();
            if (lane == 0) mbar_arrive(empty);
            if (meta.x >> 8) {
                // ---- all images of this group are in: which candidates can still matter? ------------------
                // The reference's lh = psi_sum / sqrt(phi_sum) can reach the bar (> 0) only if
                //     psi_hi > 0  and  psi_hi^2 >= bar^2 * phi_lo,    psi_hi = step_psi P + e_psi,
                //                                                    phi_lo = step_phi F - e_phi
                // (phi_lo <= 0 puts no bound on the quotient; the right-hand side is then <= 0 and only the sign
                // of psi_hi decides -- with psi_sum <= 0 the lh is <= 0 or, for phi_sum <= 0, -1, below any bar
                // > 0).  The 1e-5 taken off bar^2 covers the roundings of this test and of the reference's own
                // division and square root.  A bar that is not a positive, squarable number: no screening.
                const int g = meta.y;
                unsigned passmask = 0u;   // bit h * TG + j: candidate j of the group needs the exact evaluation
#pragma unroll
                for (int h = 0; h < TPX; ++h) {
                    float c3, c4;
                    bool screen;
                    if (SIGMAG) {
                        c3 = pv.bar_c3;
                        c4 = pv.bar_c4;
                        screen = pv.bar_ok != 0;
                    } else {
                        const float bar = lb8[h][7];
                        screen = bar > 0.0f && bar < 1.0e15f;
                        const float b2 = bar * bar * 0.99999f;
                        c3 = pv.step_phi * b2;
                        c4 = -pv.e_phi * b2;
                    }
#pragma unroll
                    for (int j = 0; j < TG; ++j) {
                        const float Pf = __int2float_rn(P[h][j]), Ff = __int2float_rn(F[h][j]);
                        const float t = fmaf(Pf, pv.step_psi, pv.e_psi);
                        const float r = fmaf(Ff, c3, c4);
                        P[h][j] = 0;
                        F[h][j] = 0;
                        bool alive = in_bounds[h] && (!screen || (t > 0.0f && t * t >= r));
                        if (COUNT) {   // kernels.cu:318 obs_count < min_obs (every observation valid where nothing is counted)
                            alive = alive && (count_here ? cnt[h][j] : sv.T) >= p.min_observations;
                            cnt[h][j] = 0;
                        }
                        if (SIGMAG) {
                            passmask |= alive ? (1u << (h * TG + j)) : 0u;
                        } else {
                            // sigma-G off: raise the pixel's bar with the candidate's lower bound and queue the
                            // candidate with its upper bound for the exact evaluation after the sweep
                            const unsigned am = __ballot_sync(0xffffffffu, alive);
                            if (am) {
                                const int c = __ldg(pv.cand + g * TG + j);   // warp-uniform; < 0: padding
                                float lb, ub;
                                lh_bounds(Pf * pv.step_psi, Ff * pv.step_phi, pv.e_psi, pv.e_phi, lb, ub);
                                const bool on = alive && c >= 0;
                                if (on && lb > lb8[h][7]) {
#pragma unroll
                                    for (int i = 0; i < 8; ++i) {
                                        if (lb > lb8[h][i]) {
                                            const float tmp = lb8[h][i];
                                            lb8[h][i] = lb;
                                            lb = tmp;
                                        }
                                    }
                                    const float bar = lb8[h][7];
                                    screen = bar > 0.0f && bar < 1.0e15f;
                                    const float b2 = bar * bar * 0.99999f;
                                    c3 = pv.step_phi * b2;
                                    c4 = -pv.e_phi * b2;
                                }
                                const unsigned qm = __ballot_sync(0xffffffffu, on && ub >= lb8[h][7]);
                                if (qm) {
                                    unsigned long long base_slot = 0;
                                    if (lane == 0) base_slot = atomicAdd(pv.clip_cnt, (unsigned long long)__popc(qm));
                                    base_slot = __shfl_sync(0xffffffffu, base_slot, 0);
                                    const unsigned long long slot_q = base_slot + __popc(qm & ((1u << lane) - 1u));
                                    if ((qm >> lane) & 1u) {
                                        // past a full queue the request is dropped: the host sees the count exceed
                                        // the capacity and searches the area again in smaller pieces
                                        if (slot_q < pv.clip_cap) pv.exq[slot_q] = ExactReq{pixel_index[h], c, ub, 0u};
                                    }
                                }
                            }
                        }
                    }
                }
                unsigned wm = SIGMAG ? passmask : 0u;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) wm |= __shfl_xor_sync(0xffffffffu, wm, d);
                while (SIGMAG && wm) {   // warp-uniform loop over the (pixel slot, candidate) pairs some lane needs
                    const int b = __ffs(wm) - 1;
                    wm &= wm - 1;
                    const int h = b / TG, j = b % TG;
                    const int c = __ldg(pv.cand + g * TG + j);
                    if (c < 0) continue;   // padding of an incomplete group
                    const bool alive = (passmask >> b) & 1u;
                    static_assert(TPX <= 2, "pixel selection below handles two pixels per lane");
                    {
                        // The exact evaluation (unclipped lh against min_lh, then the clip) is deferred: append
                        // (pixel, candidate) to the global queue; clip_kernel works it off afterwards with every SM
                        // (the requests cluster around bright sources).
                        const unsigned am = __ballot_sync(0xffffffffu, alive);
                        unsigned long long base_slot = 0;
                        if (lane == 0) base_slot = atomicAdd(pv.clip_cnt, (unsigned long long)__popc(am));
                        base_slot = __shfl_sync(0xffffffffu, base_slot, 0);
                        const unsigned long long slot_q = base_slot + __popc(am & ((1u << lane) - 1u));
                        // past a full queue the request is dropped; the host sees the count exceed the capacity and
                        // searches the area again in smaller pieces
                        if (alive && slot_q < pv.clip_cap)
                            pv.clipq[slot_q] = make_uint2((h == 0) ? pixel_index[0] : pixel_index[TPX - 1], (unsigned)c);
                    }
                }
            }
        }
#pragma unroll
        for (int h = 0; h < TPX; ++h) {
            top[h].init();   // the per-pixel lists are not touched before the sweep is over
            // ---- candidates whose offsets could not be tabulated exactly: full per-pixel arithmetic ----------
            for (int e = 0; e < pv.n_exact; ++e) {
                const int c = __ldg(pv.exact + e);
                float lh, flux;
                int obs;
                evaluate_exact<NB, SIGMAG, MAXT>(sv, s_times, p, px[h], py, __ldg(cvx + c), __ldg(cvy + c), lh, flux, obs,
                                                 nullptr);
                if (obs < p.min_observations || (SIGMAG && lh < p.min_lh)) continue;
                if (beats(lh, c, top[h].worst(), top[h].cand[K - 1])) top[h].insert(lh, c);
            }
            // ---- epilogue: flux and obs_count of the winners, written in the reference's slot order ------------
            const int xi = px[h] - p.x_start_min, yi = py - p.y_start_min;
            if (!SIGMAG && in_bounds[h]) pv.bar_out[pixel_index[h]] = lb8[h][7];
#pragma unroll
            for (int r = 0; r < K; ++r) {
                if (r >= (int)p.results_per_pixel) break;  // warp-uniform
                Slot out{-FLT_MAX, 0.0f, 0, -1};
                const int c = top[h].cand[r];
                const float vx = (c >= 0) ? __ldg(cvx + c) : 0.0f, vy = (c >= 0) ? __ldg(cvy + c) : 0.0f;
                float lh = 0.0f, flux = 0.0f;
                int obs = 0;
                if (c >= 0) evaluate_exact<NB, false, 1>(sv, s_times, p, px[h], py, vx, vy, lh, flux, obs, nullptr);
                if (SIGMAG) {
                    const bool need = c >= 0 && !(obs < p.min_observations || obs == 0 || lh < p.min_lh);
                    int nv = obs;
                    sigmag_for_lanes<NB, MAXT>(need, sv, s_times, px[h], py, vx, vy, p, lh, flux, nv);
                }
                if (c >= 0) out = Slot{lh, flux, obs, c};
                if (xi < sw && yi < sh) slots[((size_t)yi * sw + xi) * p.results_per_pixel + r] = out;
            }
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
        const float vx = __ldg(cvx + c), vy = __ldg(cvy + c);
        bool rejected = false;
        if constexpr (MAXT <= 256) {
            // the request is warp-uniform: the cooperative routine in line, no hand-over between lanes
            constexpr int E = (MAXT + 31) / 32;
            float oa[E], ob[E], olc[E];
            bool ook[E];
            unsigned ookm[E];
            const int n_obs = warp_sigmag_load<NB, E>(sv, s_times, x, y, vx, vy, oa, ob, olc, ook, ookm);
            // most requests are certainly rejected (bounds of the float sums): no ordered sums for those
            rejected = warp_sigmag_rejects<E>(sv.T, oa, ob, olc, ook, n_obs, p.sgl_L, p.sgl_H, p.sigmag_coeff,
                                              p.min_observations, p.min_lh);
#ifndef KB_CLIP_VERIFY
            if (rejected) continue;
#endif
            bool serial;
            warp_sigmag_eval<NB, E>(sv, oa, ob, olc, ook, ookm, n_obs, p.sgl_L, p.sgl_H, p.sigmag_coeff,
                                    p.min_observations, p.min_lh, lh, flux, nv, serial, ws);
            if (serial && lane == 0)   // tied lc values with different payloads: the reference's exchange-sort order
                sigmag_recompute<NB, MAXT>(sv, s_times, x, y, vx, vy, p.sgl_L, p.sgl_H, p.sigmag_coeff, lh, flux,
                                           nullptr);
        } else {
            sigmag_for_lanes<NB, MAXT>(lane == 0, sv, s_times, x, y, vx, vy, p, lh, flux, nv, ws);
        }
#ifdef KB_CLIP_VERIFY
        if constexpr (MAXT <= 256) {   // test build: every "certain" rejection against the exact routine
            if (lane == 0 && rejected) atomicAdd(cnt + 3, 1ull);
            if (lane == 0 && rejected && !(nv < p.min_observations) && !(lh < p.min_lh)) {
                printf("kbmod_b200: warp_sigmag_rejects dropped pixel %u cand %d with lh %.9g nv %d\n", it.x, c, lh, nv);
                __trap();
            }
        }
#endif
        if (lane == 0) {
            // Queued candidates only passed the screening bound.  sigmag_for_lanes returns what evaluateTrajectory
            // does (kernels.cu:154-242): the unclipped lh when it is below min_lh or there are too few
            // observations (lh = flux = -1 when there is none), else the clipped one; kernels.cu:318-320 then
            // rejects nv < min_obs and lh < min_lh.
            if (!(nv < p.min_observations) && !(lh < p.min_lh)) {
                const unsigned long long k = atomicAdd(cnt + 1, 1ull);  // counts past surv_cap too: host reruns
                if (k < surv_cap) surv[k] = ClipSurvivor{it.x, c, lh, flux, nv};
            }
        }
    }
}

// ---- sigma-G off: exact evaluation of the queued candidates, one warp per request, all SMs -----------------------
// Lanes take consecutive images (32 gathers in flight); the sums are then taken in time order, which is the
// reference's order (kernels.cu:175-197), every lane running the same chain on shuffled values.
template <int NB>
__global__ void __launch_bounds__(256)
        exact_kernel(const StackView sv, const kb_params p, const float *__restrict__ cvx, const float *__restrict__ cvy,
                     const ExactReq *__restrict__ queue, unsigned long long *cnt, unsigned long long cap,
                     const float *__restrict__ bar, ClipSurvivor *__restrict__ surv, unsigned long long surv_cap) {
    extern __shared__ double s_times[];
    for (int i = threadIdx.x; i < sv.T; i += blockDim.x) s_times[i] = sv.times[i];
    __syncthreads();
    const unsigned FULL = 0xffffffffu;
    const unsigned long long n = min(cnt[0], cap);
    const int sw = p.x_start_max - p.x_start_min;
    const int lane = threadIdx.x & 31;
    const unsigned long long nwarps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    for (unsigned long long i = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += nwarps) {
        const ExactReq rq = queue[i];
        const float b = __ldg(bar + rq.pixel);
        if (!(rq.ub >= b)) continue;   // the pixel's bar rose above this candidate's bound after it was queued
        const int x = p.x_start_min + (int)(rq.pixel % (unsigned)sw), y = p.y_start_min + (int)(rq.pixel / (unsigned)sw);
        const float vx = __ldg(cvx + rq.cand), vy = __ldg(cvy + rq.cand);
        float ps = 0.0f, ph = 0.0f;
        int seen = 0;
        for (int t0 = 0; t0 < sv.T; t0 += 32) {
            const int t = t0 + lane;
            float a = 0.0f, c = 0.0f;
            bool ok = false;
            if (t < sv.T) {
                const double tm = s_times[t];
                ok = load_pair<NB>(sv, t, predict_index(y, vy, tm), predict_index(x, vx, tm), a, c);
            }
            const unsigned okm = __ballot_sync(FULL, ok);
            seen += __popc(okm);
            const int nsrc = min(32, sv.T - t0);
            for (int ls = 0; ls < nsrc; ++ls) {
                const float va = __shfl_sync(FULL, a, ls), vc = __shfl_sync(FULL, c, ls);
                if ((okm >> ls) & 1u) {
                    ps = __fadd_rn(ps, va);
                    ph = __fadd_rn(ph, vc);
                }
            }
        }
        float lh, flux;
        finish_sums(ps, ph, lh, flux);
        // kernels.cu:318: obs_count < min_obs never enters the list; a NaN lh never passes the insertion's '>'
        if (lane == 0 && !(seen < p.min_observations) && lh >= b) {
            const unsigned long long k = atomicAdd(cnt + 1, 1ull);   // counts past surv_cap too: host reruns
            if (k < surv_cap) surv[k] = ClipSurvivor{rq.pixel, rq.cand, lh, flux, seen};
        }
    }
}

template <int NB>
static void launch_exact(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                         ClipScratch &clip, cudaStream_t st) {
    const size_t tbytes = sizeof(double) * (size_t)sv.T;
    exact_kernel<NB><<<148 * 8, 256, tbytes, st>>>(sv, p, d_vx, d_vy, static_cast<const ExactReq *>(clip.queue), clip.counters,
                                                   clip.cap, clip.bar, clip.surv, clip.surv_cap);
    KB_CUDA(cudaGetLastError());
}

template <int NB>
static void launch_clip(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                        ClipScratch &clip, cudaStream_t st) {
    const size_t tbytes = sizeof(double) * (size_t)((sv.T + 1) & ~1);
    const int grid = 148 * 8;
#define KB_CLIP(MT)                                                                                              \
    clip_kernel<NB, MT><<<grid, 256, tbytes + ((MT) <= 256 ? sizeof(float2) * 8 * (MT) : 0), st>>>(              \
            sv, p, d_vx, d_vy, static_cast<const uint2 *>(clip.queue), clip.counters, clip.cap, clip.surv, clip.surv_cap)
    if (sv.T <= 32) KB_CLIP(32);
    else if (sv.T <= 64) KB_CLIP(64);
    else if (sv.T <= 128) KB_CLIP(128);
    else if (sv.T <= 256) KB_CLIP(256);
    else KB_CLIP(kMaxTSigma);
#undef KB_CLIP
    KB_CUDA(cudaGetLastError());
}

template <int NB, bool COUNT, bool SIGMAG, int MAXT>
static void launch_k(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                     Slot *d_slots, dim3 grid, cudaStream_t st) {
#define KB_LAUNCH_T(KK)                                                                                         \
    do {                                                                                                        \
        auto kfn = search_tiled_kernel<NB, COUNT, SIGMAG, MAXT, KK>;                                            \
        KB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->smem));       \
        kfn<<<grid, TNT, plan->smem, st>>>(plan->maps, plan->v, sv, p, d_vx, d_vy, d_slots);                    \
    } while (0)
    // One list length: the first K entries of a sorted top-8 list are the top-K list (K <= 8).
    KB_LAUNCH_T(8);
#undef KB_LAUNCH_T
}

template <int NB>
static void launch_nb(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st) {
    if (!plan->sigmag) {
        if (p.min_observations > 0) launch_k<NB, true, false, 1>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
        else launch_k<NB, false, false, 1>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    } else if (sv.T <= 32) launch_k<NB, false, true, 32>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    else if (sv.T <= 64) launch_k<NB, false, true, 64>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
    // (the serial sigma-G routine of the large variant only runs for the K winners of a pixel and when the clip
    // queue overflows)
    else launch_k<NB, false, true, kMaxTSigma>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
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
void launch_exact_nb4(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                      cudaStream_t st);
void launch_exact_nb2(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                      cudaStream_t st);
void launch_exact_nb1(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                      cudaStream_t st);

}  // namespace kb
