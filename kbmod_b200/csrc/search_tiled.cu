// search_tiled.cu -- the throughput path of the search: a TMA-staged, time-tiled gather-reduce kernel that screens
// every (start pixel, candidate) with exact integer sums over a 16-bit fixed-point copy of the stack and evaluates
// exactly only what the screening cannot rule out (see search_tiled_kernel.cuh).
//
// Decomposition (B200-first, not the reference's one-thread-per-pixel serial loop):
//   * A CTA owns a tile of 64 x 15 start pixels (one consumer warp per row, every lane two pixels: x and x + 32) and
//     visits the candidate list in *groups* of G = 16 (sigma-G on) or 8 (off) candidates that are neighbours in
//     velocity space.  Every consumer thread keeps the G running integer (psi, phi) sums of both its pixels in
//     registers.
//   * For a group and an image, all G trajectories of all tile pixels land inside one small box of the image:
//     origin (tile + min offset), size (64 + spread_x) x (15 + spread_y).  A dedicated producer warp fetches those
//     boxes of the SCREENING copy (one 32-bit word per pixel: two int16 in fixed point, 0 = no data) with TMA
//     (cp.async.bulk.tensor, 3-D map over [plane][y][x]) into a ring of shared-memory stages, several images per
//     stage, signalling mbarriers.  Out-of-image parts of a box are zero-filled by the TMA unit -- exactly the
//     reference's "outside the image = no data".  The stage's schedule record arrives by a 1-D bulk copy.
//   * Consumers wait on the stage's mbarrier and gather: lane l of row r reads the word at
//     (r + dy_j(t), l + dx_j(t)) -- x-adjacent lanes read adjacent words, so every LDS.32 is conflict-free -- and adds
//     its halves to the two sums with one IDP.2A each.  The per-(candidate, image) integer offsets dx, dy come from
//     a table built on the host with the reference's fp64 arithmetic.
//   * Run planes: exposures taken minutes apart (a run) mostly keep a candidate's pixel offset.  The screening copy
//     carries one summed plane per run (and per half/quarter/pair of one); a group whose candidates all keep their
//     offsets through a run gathers the run's plane once, with the run length as the IDP.2A multiplier.
//   * After the last image of a group the integer sums give rigorous bounds of the exact float likelihood (see
//     screen_bounds).  Sigma-G on: a pair whose upper bound cannot reach min_lh is dropped, the rest is queued for
//     clip_kernel (warp-cooperative sigma-G with the reference's exact arithmetic).  Sigma-G off: the eight largest
//     lower bounds per pixel are a rising bar; pairs whose upper bound reaches it are queued for exact_kernel.
//     Survivors of either queue enter the per-pixel top-K lists (ordered by lh, then candidate index == the order
//     the reference's in-order strict-'>' insertion produces) in surv_merge_kernel.  Only exactly evaluated values
//     are ever written, so results are bit-identical to the reference's.
//   * Observation counts (min_obs with sigma-G off) are only swept in tiles where an observation can be invalid (a
//     stack with invalid pixels, or boxes that leave the image); elsewhere the count is T.
//
// Exactness of the offset table: the reference computes ix = floor((x + fl(vx*t)) + 0.5) in fp64 per pixel.
// With |x|, |vx*t| < 2^22 the two roundings perturb the sum by < 2^-29, so ix = x + floor(fl(vx*t) + 0.5)
// unless fl(vx*t)+0.5 lies within 2^-27 of an integer without being exactly one.  Candidates with such an
// image are not tabulated; every thread evaluates them with the exact per-pixel routine instead.
#include <cuda.h>
#include <chrono>
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

#include "search_tiled_kernel.cuh"

namespace kb {

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
    double fl = (double)(long long)q;    // floor without the libm call: |q| < 2^22 + 1, truncate then step down
    if (fl > q) fl -= 1.0;
    off = (int)fl;
    const double r = q - fl;
    const double dist = (r < 0.5) ? r : 1.0 - r;
    if (dist > 7.450580596923828e-09) return true;          // 2^-27
    return r == 0.0 && (q - 0.5) == prod;                   // exactly on a boundary: no rounding anywhere
}

// Bounds of |reference float sum - step * exact integer sum| over T observations (see the kernel's header):
//   quantisation: <= 0.502 step per observation for psi; phi may be moved by one more half step so that a valid pair
//                 never maps to the word 0 (1.002 step); 0.52 / 1.02 leave room for the kernel's own float products
//   the reference's sequential float sum: <= T^2 * 2^-24 * max|value| = T^2 * step / 512 (max|value| <= 32767.5 step)
static void screen_bounds(int T, float step_psi, float step_phi, float &e_psi, float &e_phi) {
    const double t = (double)T;
    e_psi = (float)((double)step_psi * (0.52 * t + t * t / 500.0 + 2.0));
    e_phi = (float)((double)step_phi * (1.02 * t + t * t / 500.0 + 2.0));
}

TiledPlan *tiled_plan_create(const StackView &sv, const kb_params &p, const float *h_vx, const float *h_vy,
                             uint64_t n_cand, cudaStream_t st, std::string *why_not, const double *h_times,
                             const int2 *h_runs) {
    auto no = [&](const char *why) -> TiledPlan * {
        if (why_not) *why_not = why;
        return nullptr;
    };
    const int T = sv.T, K = (int)p.results_per_pixel;
    const int TG = group_size(p.do_sigmag_filter != 0);   // candidates per group of this schedule
    const bool dbg = getenv("KB_DEBUG") != nullptr;
    auto t_phase = std::chrono::steady_clock::now();
    auto phase = [&](const char *what) {
        if (!dbg) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[kbmod_b200] tiled plan: %-28s %.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_phase).count());
        t_phase = now;
    };
    if (K > kMaxRegK) return no("results_per_pixel > 8");
    if (n_cand == 0) return no("no candidates");
    if (p.do_sigmag_filter && T > kMaxTSigma) return no("too many images for sigma-G");
    if (sizeof(double) * (size_t)T > 16 * 1024) return no("too many images for the tiled kernel's time table");
    if (!sv.screen) return no("no screening copy of the stack");
    if ((double)T * 32767.0 >= 2147483647.0) return no("too many images for 32-bit screening sums");
    std::vector<double> times(T);
    if (h_times) {
        std::copy(h_times, h_times + T, times.begin());
    } else {
        KB_CUDA(cudaMemcpyAsync(times.data(), sv.times, sizeof(double) * T, cudaMemcpyDeviceToHost, st));
        KB_CUDA(cudaStreamSynchronize(st));
    }
    double tabs = 0.0;
    for (double t : times) {
        if (!std::isfinite(t)) return no("non-finite time");
        tabs = std::max(tabs, fabs(t));
    }
    const int PB = kScreenPairBytes;  // the TMA boxes come from the screening copy: one 32-bit word per pixel
    const int bw_limit = 256;         // TMA box: at most 256 elements along x

    // ---- candidates whose offsets cannot be tabulated exactly go to the exact list ---------------------------------
    const int N = (int)n_cand;
    std::vector<int> exact, tabulated;
    tabulated.reserve(N);
    {
        std::vector<uint8_t> tab_ok((size_t)N);
#pragma omp parallel for schedule(static)
        for (int c = 0; c < N; ++c) {
            bool ok = true;
            for (int t = 0; t < T && ok; ++t) {
                int o;
                ok = tab_offset(h_vx[c], times[t], o) && tab_offset(h_vy[c], times[t], o);
            }
            tab_ok[c] = ok ? 1 : 0;
        }
        for (int c = 0; c < N; ++c) (tab_ok[c] ? tabulated : exact).push_back(c);
    }
    const int NTab = (int)tabulated.size();
    phase("tabulation test");

    // ---- group the rest: k-d style recursive bisection of velocity space into leaves of <= TG candidates ---------
    // A group's box is (TBX + sx) x (TBY + sy) pixels with sx, sy ~ velocity extent * max|t|; each cut goes
    // across the direction that shrinks that box most, at a multiple of TG so the leaves come out full.
    const int XA = 16 / PB;                          // TMA: box origin must be 16-byte aligned in global memory
    const int bw_round = 8;                          // keeps every TMA destination 128-byte aligned
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
        // Order every leaf as a greedy chain of nearest neighbours in velocity space: consecutive candidates then
        // share their pixel offset on many images, and the kernel skips the repeated shared-memory load.
        for (const Seg &l : leaves) {
            int *c = tabulated.data() + l.lo;
            const int n = l.hi - l.lo;
            int first = 0;
            for (int i = 1; i < n; ++i)
                if (h_vy[c[i]] < h_vy[c[first]] || (h_vy[c[i]] == h_vy[c[first]] && h_vx[c[i]] < h_vx[c[first]])) first = i;
            std::swap(c[0], c[first]);
            for (int i = 1; i < n; ++i) {
                int best = i;
                double bd = DBL_MAX;
                for (int k = i; k < n; ++k) {
                    const double ddx = (double)h_vx[c[k]] - h_vx[c[i - 1]], ddy = (double)h_vy[c[k]] - h_vy[c[i - 1]];
                    const double d2 = ddx * ddx + ddy * ddy;
                    if (d2 < bd || (d2 == bd && c[k] < c[best])) {
                        bd = d2;
                        best = k;
                    }
                }
                std::swap(c[i], c[best]);
            }
        }
        for (const Seg &l : leaves) group_start.push_back(l.lo);
        group_start.push_back(NTab);
    }
    const int n_groups = (int)group_start.size() - 1;
    if (n_groups == 0 && exact.empty()) return no("no candidates");
    phase("grouping");
    // Box widths come in steps (one tensor map each); a box is as wide as its own (group, image) needs.
    const int w_step = bw_round, w_base = TBX + w_step;
    const int w_max = std::min(w_base + (NMAPS - 1) * w_step, bw_limit);
    if (TBX + max_sx + XA - 1 > w_max) return no("velocity spread too large for the TMA box");
    const int rmax = (TBY + max_sy + TCH - 1) / TCH * TCH;
    const int max_ch = rmax / TCH;

    // ---- stage records, group by group (greedy packing of consecutive records into <= STAGE_BYTES) ---------------
    // A record is one image -- or, with `use_runs`, a whole run of closely spaced exposures through which every
    // candidate of the group keeps its pixel offsets: the gather then reads the run's summed plane of the screening
    // copy once, with the run's length as the multiplier of its IDP.2A.
    std::vector<int> group_reach(std::max(n_groups, 1), 0);
    std::vector<int4> group_ext(std::max(n_groups, 1), make_int4(INT_MAX, INT_MIN, INT_MAX, INT_MIN));
    std::vector<int> cand((size_t)std::max(n_groups, 1) * TG, -1);
    // runs that start at image t, longest first (the handle lists a maximal run before its halves, quarters, ...)
    std::vector<std::vector<int>> run_at(T);
    for (int c = 0; c < sv.n_runs && h_runs; ++c) run_at[h_runs[c].x].push_back(c);
    for (auto &v : run_at)
        std::stable_sort(v.begin(), v.end(), [&](int a, int b2) { return h_runs[a].y > h_runs[b2].y; });
    long long n_records[2] = {0, 0};
    // Stage records of a schedule in one plain buffer (no value-initialisation of 100+ MB).
    struct StageBuf {
        StageRec *p = nullptr;
        size_t n = 0;
        StageBuf() = default;
        StageBuf(const StageBuf &) = delete;
        StageBuf &operator=(const StageBuf &) = delete;
        ~StageBuf() { free(p); }
        bool empty() const { return n == 0; }
        size_t size() const { return n; }
        const StageRec *data() const { return p; }
    };
    // The records of group g, packed greedily into stages of <= STAGE_BYTES.
    auto pack_group = [&](int g, bool use_runs, std::vector<StageRec> &out, long long &records, bool &overflow) {
        const int b = group_start[g], n = group_start[g + 1] - b;
        for (int j = 0; j < n; ++j) cand[(size_t)g * TG + j] = tabulated[b + j];
        // pixel offsets of the group's candidates at every image, once
        std::vector<int> off_x((size_t)T * TGMAX), off_y((size_t)T * TGMAX);
        for (int j = 0; j < n; ++j) {
            const int c = tabulated[b + j];
            for (int t = 0; t < T; ++t) {
                tab_offset(h_vx[c], times[t], off_x[(size_t)t * TGMAX + j]);
                tab_offset(h_vy[c], times[t], off_y[(size_t)t * TGMAX + j]);
            }
        }
        out.reserve((size_t)T / 6 + 2);
        StageRec cur;
        memset(&cur, 0, sizeof(cur));
        int nt = 0, used = 0, tx = 0, t0 = 0;
        auto flush = [&](bool last) {
            cur.meta = make_int4(nt | ((last ? 1 : 0) << 8), g, t0, tx + (int)sizeof(StageRec));
            out.push_back(cur);
            memset(&cur, 0, sizeof(cur));
            nt = used = tx = 0;
        };
        for (int t = 0; t < T;) {
            int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
            const int *dx = &off_x[(size_t)t * TGMAX], *dy = &off_y[(size_t)t * TGMAX];
            for (int j = 0; j < n; ++j) {
                mnx = std::min(mnx, dx[j]); mxx = std::max(mxx, dx[j]);
                mny = std::min(mny, dy[j]); mxy = std::max(mxy, dy[j]);
            }
            int plane = t, mult = 1;
            if (use_runs && !run_at[t].empty()) {
                // how many images from t on do all candidates stay on their pixels?  the longest run that
                // starts here and fits is read as one record
                const int longest = h_runs[run_at[t][0]].y;
                int hold = 1;
                for (; hold < longest; ++hold) {
                    const int *hx = &off_x[(size_t)(t + hold) * TGMAX], *hy = &off_y[(size_t)(t + hold) * TGMAX];
                    bool same = true;
                    for (int j = 0; j < n && same; ++j) same = hx[j] == dx[j] && hy[j] == dy[j];
                    if (!same) break;
                }
                for (int r : run_at[t])
                    if (h_runs[r].y <= hold) {
                        plane = T + r;
                        mult = h_runs[r].y;
                        break;
                    }
            }
            mnx -= ((p.x_start_min + mnx) % XA + XA) % XA;  // tile origins are multiples of 32 past x_start_min
            const int need_w = TBX + (mxx - mnx), nrows = TBY + (mxy - mny);
            const int widx = std::max(0, (need_w - w_base + w_step - 1) / w_step);
            const int pitch = w_base + w_step * widx;
            const int rows_pad = (nrows + TCH - 1) / TCH * TCH;
            const int bytes = rows_pad * pitch * PB;
            if (pitch > w_max || widx >= NMAPS || nrows > rmax || nrows > 255 || bytes > STAGE_BYTES) {
                overflow = true;
                break;
            }
            if (nt == TMAXTC || used + bytes > STAGE_BYTES) flush(false);
            if (nt == 0) t0 = t;
            cur.box[nt] = make_int4(mnx, mny, nrows | (widx << 8) | (plane << 12), used);
            group_reach[g] = std::max(group_reach[g], mny + rows_pad);
            group_ext[g].x = std::min(group_ext[g].x, mnx);
            group_ext[g].y = std::max(group_ext[g].y, mnx + pitch);
            group_ext[g].z = std::min(group_ext[g].z, mny);
            group_ext[g].w = std::max(group_ext[g].w, mny + rows_pad);
            cur.img[nt].slot_off = used;
            cur.img[nt].pitch_bytes = pitch * PB;
            cur.img[nt].mult_lo = mult;         // byte 0: multiplies the low half of the word (psi)
            cur.img[nt].mult_hi = mult << 8;    // byte 1: multiplies the high half (phi)
            for (int j = 0; j < TGMAX; ++j)
                cur.img[nt].offs[j] = (j < n) ? ((dy[j] - mny) * pitch + (dx[j] - mnx)) * PB : 0;
            used += bytes;
            tx += bytes;
            ++nt;
            ++records;
            t += mult;
        }
        if (nt > 0) flush(true);
    };
    auto build = [&](bool use_runs, StageBuf &stages) -> const char * {
        std::vector<std::vector<StageRec>> per_group(std::max(n_groups, 1));
        bool overflow = false;
        long long records = 0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : records) reduction(|| : overflow)
        for (int g = 0; g < n_groups; ++g) {
            bool ov = false;
            pack_group(g, use_runs, per_group[g], records, ov);
            overflow = overflow || ov;
        }
        n_records[use_runs ? 1 : 0] = records;
        if (overflow) return "velocity spread too large for a shared-memory stage";
        std::vector<size_t> first((size_t)n_groups + 1, 0);
        for (int g = 0; g < n_groups; ++g) first[(size_t)g + 1] = first[g] + per_group[g].size();
        const size_t total = first[n_groups];
        if (total >= (1ull << 31)) return "too many stages";
        stages.p = static_cast<StageRec *>(malloc(std::max<size_t>(total, 1) * sizeof(StageRec)));
        if (!stages.p) return "out of host memory for the schedule";
        stages.n = total;
#pragma omp parallel for schedule(dynamic, 64)
        for (int g = 0; g < n_groups; ++g) {
            if (!per_group[g].empty())
                memcpy(stages.p + first[g], per_group[g].data(), per_group[g].size() * sizeof(StageRec));
            std::vector<StageRec>().swap(per_group[g]);
        }
        return nullptr;
    };
    // The schedule every tile uses -- with run records where the stack has runs -- and, for searches that count
    // observations in a stack with invalid pixels, a second one of single images for the tiles that count (a run's
    // summed word cannot say how many of its images had data; without invalid pixels it is all or none of them).
    const bool have_runs = sv.n_runs > 0 && h_runs != nullptr;
    const bool counting = !p.do_sigmag_filter && p.min_observations > 0;
    StageBuf stages, stages_plain;
    if (const char *err = build(have_runs, stages)) return no(err);
    if (have_runs && counting && !sv.all_valid)
        if (const char *err = build(false, stages_plain)) return no(err);
    const int n_stages = (int)stages.size();
    phase("stage records");
    if (dbg && getenv("KB_DEBUG_PLAN_HASH")) {
        unsigned long long hsh = 1469598103934665603ull;
        const unsigned char *q = reinterpret_cast<const unsigned char *>(stages.data());
        for (size_t i = 0; i < stages.size() * sizeof(StageRec); ++i) hsh = (hsh ^ q[i]) * 1099511628211ull;
        for (int c2 : cand) hsh = (hsh ^ (unsigned)c2) * 1099511628211ull;
        fprintf(stderr, "[kbmod_b200] tiled plan: schedule hash %016llx\n", hsh);
    }
    if (dbg)
        fprintf(stderr, "[kbmod_b200] tiled plan: N=%d tab=%d exact=%zu groups=%d stages=%d (+%zu single-image) "
                "records=%lld of %lld max_sx=%d max_sy=%d w_base=%d w_step=%d rmax=%d sx_lim=%d sy_lim=%d\n", N, NTab,
                exact.size(), n_groups, n_stages, stages_plain.size(), have_runs ? n_records[1] : n_records[0],
                (long long)n_groups * T, max_sx, max_sy, w_base, w_step, rmax, sx_lim, sy_lim);

    std::unique_ptr<TiledPlan> plan(new TiledPlan());
    plan->nb = sv.nb;
    plan->K = K;
    plan->sigmag = p.do_sigmag_filter != 0;
    plan->tg = TG;
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t b_st = up(std::max<size_t>(stages.size(), 1) * sizeof(StageRec)), b_cand = up(cand.size() * 4),
                 b_ex = up(std::max<size_t>(exact.size(), 1) * 4),
                 b_pl = stages_plain.empty() ? 0 : up(stages_plain.size() * sizeof(StageRec));
    KB_CUDA(cudaMalloc(&plan->d_block, b_st + b_cand + b_ex + b_pl));
    uint8_t *d = (uint8_t *)plan->d_block;
    if (!stages.empty())
        KB_CUDA(cudaMemcpyAsync(d, stages.data(), stages.size() * sizeof(StageRec), cudaMemcpyHostToDevice, st));
    if (!stages_plain.empty())
        KB_CUDA(cudaMemcpyAsync(d + b_st + b_cand + b_ex, stages_plain.data(), stages_plain.size() * sizeof(StageRec),
                                cudaMemcpyHostToDevice, st));
    KB_CUDA(cudaMemcpyAsync(d + b_st, cand.data(), cand.size() * 4, cudaMemcpyHostToDevice, st));
    if (!exact.empty())
        KB_CUDA(cudaMemcpyAsync(d + b_st + b_cand, exact.data(), exact.size() * 4, cudaMemcpyHostToDevice, st));
    KB_CUDA(cudaStreamSynchronize(st));  // the host vectors die with this scope
    phase("upload");
    PlanView &v = plan->v;
    v.stages = (const StageRec *)d;
    v.cand = (const int *)(d + b_st);
    v.exact = (const int *)(d + b_st + b_cand);
    v.n_exact = (int)exact.size();
    v.n_groups = n_groups;
    v.n_stages = n_stages;
    v.stages_plain = stages_plain.empty() ? v.stages : (const StageRec *)(d + b_st + b_cand + b_ex);
    v.n_stages_plain = stages_plain.empty() ? n_stages : (int)stages_plain.size();
    v.tile_order = nullptr;
    plan->order_tiles = counting;   // tiles that count are slower: launched first (launch_tiled_any)
    v.w_base = w_base;
    v.w_step = w_step;
    v.max_ch = max_ch;
    v.T = T;
    v.clipq = nullptr;
    v.clip_cnt = nullptr;
    v.clip_cap = 0;
    v.step_psi = sv.q_step_psi;
    v.step_phi = sv.q_step_phi;
    screen_bounds(T, sv.q_step_psi, sv.q_step_phi, v.e_psi, v.e_phi);
    {
        const float bar = p.min_lh, b2 = bar * bar * 0.99999f;
        v.bar_ok = (p.do_sigmag_filter && bar > 0.0f && bar < 1.0e15f) ? 1 : 0;
        v.bar_c3 = v.bar_ok ? v.step_phi * b2 : 0.0f;
        v.bar_c4 = v.bar_ok ? -v.e_phi * b2 : 0.0f;
    }
    plan->smem = (size_t)TS * (STAGE_BYTES + HDR_BYTES) + 2 * TS * 8 + sizeof(double) * (size_t)T;
    // Rows a tile can touch below its origin: the deepest box bottom of any group; candidates evaluated per pixel
    // (exact list, winners in the epilogue, clips) go as far as their own last offset.
    int reach_hi = TBY;
    for (int r : group_reach) reach_hi = std::max(reach_hi, r);
    for (int c : exact) reach_hi = std::max(reach_hi, TBY + (int)ceil(fabs((double)h_vy[c]) * tabs) + 2);
    plan->reach_hi = reach_hi;
    v.ext_x0 = v.ext_y0 = 0;
    v.ext_x1 = TBX;
    v.ext_y1 = TBY;
    for (int g = 0; g < n_groups; ++g) {
        v.ext_x0 = std::min(v.ext_x0, group_ext[g].x);
        v.ext_x1 = std::max(v.ext_x1, group_ext[g].y);
        v.ext_y0 = std::min(v.ext_y0, group_ext[g].z);
        v.ext_y1 = std::max(v.ext_y1, group_ext[g].w);
    }
    v.pixel_base = 0;

    // ---- tensor maps over the screening copy: dims (pixels, rows, images) of 32-bit words, one per box width; what
    // lies outside the image is filled with zeros by the TMA unit == the screening copy's "no data" word
    const cuuint64_t gdim[3] = {(cuuint64_t)sv.W, (cuuint64_t)sv.H, (cuuint64_t)(sv.T + sv.n_runs)};
    const cuuint64_t gstr[2] = {(cuuint64_t)sv.Wp * PB, (cuuint64_t)sv.Wp * PB * sv.H};
    const cuuint32_t estr[3] = {1, 1, 1};
    for (int i = 0; i < NMAPS; ++i) {
        const int bw = std::min(w_base + i * w_step, w_max);
        const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)TCH, 1};
        const CUresult r = encode_fn()(&plan->maps.m[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3,
                                       const_cast<uint32_t *>(sv.screen), gdim, gstr, box, estr,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            cudaFree(plan->d_block);
            plan->d_block = nullptr;
            return no("cuTensorMapEncodeTiled failed");
        }
    }
    return plan.release();
}

float tiled_plan_kernel_ms(TiledPlan *plan) { return plan ? plan->ms_accum : 0.0f; }
unsigned long long tiled_plan_requests(TiledPlan *plan) { return plan ? plan->requests_accum : 0ull; }

void tiled_plan_destroy(TiledPlan *plan) {
    if (!plan) return;
    if (plan->d_block) cudaFree(plan->d_block);
    if (plan->d_order) cudaFree(plan->d_order);
    if (plan->ev0) cudaEventDestroy(plan->ev0);
    if (plan->ev1) cudaEventDestroy(plan->ev1);
    delete plan;
}

static void launch_tiled_any(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx,
                             const float *d_vy, Slot *d_slots, cudaStream_t st, bool may_order) {
    const int sw = p.x_start_max - p.x_start_min, sh = p.y_start_max - p.y_start_min;
    dim3 grid((sw + TBX - 1) / TBX, (sh + TBY - 1) / TBY);
    plan->v.tile_order = nullptr;
    if (may_order && plan->order_tiles && sv.all_valid) {   // (with invalid pixels every tile counts)
        // Tiles from which a box reaches over the image's edge count observations and follow the single-image
        // schedule, several times longer than the run schedule of the interior tiles: start them first, so the
        // short tiles fill in behind them instead of the long ones forming the tail of the launch.
        const PlanView &v = plan->v;
        const size_t n_tiles = (size_t)grid.x * grid.y;
        std::vector<uint2> slow, fast;
        for (unsigned ty = 0; ty < grid.y; ++ty)
            for (unsigned tx = 0; tx < grid.x; ++tx) {
                const int x0 = p.x_start_min + (int)tx * TBX, y0 = p.y_start_min + (int)ty * TBY;
                const bool inside = x0 + v.ext_x0 >= 0 && x0 + v.ext_x1 <= sv.W && y0 + v.ext_y0 >= 0 && y0 + v.ext_y1 <= sv.H;
                (inside ? fast : slow).push_back(make_uint2(tx, ty));
            }
        if (!slow.empty() && !fast.empty()) {
            slow.insert(slow.end(), fast.begin(), fast.end());
            if (plan->order_cap < n_tiles) {
                if (plan->d_order) cudaFree(plan->d_order);
                plan->d_order = nullptr;
                plan->order_cap = 0;
                KB_CUDA(cudaMalloc(&plan->d_order, sizeof(uint2) * n_tiles));
                plan->order_cap = n_tiles;
            }
            KB_CUDA(cudaMemcpyAsync(plan->d_order, slow.data(), sizeof(uint2) * n_tiles, cudaMemcpyHostToDevice, st));
            plan->v.tile_order = plan->d_order;
            grid = dim3((unsigned)n_tiles, 1);
        }
    }
    switch (sv.nb) {
        case 4: launch_tiled_nb4(plan, sv, p, d_vx, d_vy, d_slots, grid, st); break;
        case 2: launch_tiled_nb2(plan, sv, p, d_vx, d_vy, d_slots, grid, st); break;
        case 1: launch_tiled_nb1(plan, sv, p, d_vx, d_vy, d_slots, grid, st); break;
        default: throw Error("bad num_bytes");
    }
    KB_CUDA(cudaGetLastError());
}

static constexpr uint64_t kExactQueueCap = 768ull << 20;   // sigma-G off: requests per launch (12 GB of the 180)

int launch_search_tiled(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx,
                        const float *d_vy, Slot *d_slots, ClipScratch &clip, ResultScratch &scr, cudaStream_t st,
                        const UploadProgress *up, bool *overflowed, bool is_chunk) {
    const int sw = p.x_start_max - p.x_start_min, sh = p.y_start_max - p.y_start_min;
    if (!is_chunk) {
        plan->ms_accum = 0.0f;
        plan->requests_accum = 0;
        if (overflowed) *overflowed = false;
    }
    if (!plan->sigmag && !up && !is_chunk) {
        // Sigma-G off: every pixel queues the candidates whose bound reached its rising bar -- a few hundred per pixel,
        // more for longer candidate lists.  A search area whose requests would not fit the queue is searched in
        // chunks of rows, each with its own exact evaluation and merge.
        const double n_cand = (double)std::max(plan->v.n_groups * plan->tg, 2);
        const double per_pixel = 64.0 + 24.0 * log2(n_cand);
        int rows = (int)std::min<double>((double)sh, std::max(1.0, (double)kExactQueueCap / (per_pixel * (double)sw)));
        rows = std::max(TBY, rows / TBY * TBY);
        if (rows < sh) {
            int launches = 0;
            for (int ya = p.y_start_min; ya < p.y_start_max; ya += rows) {
                kb_params pc = p;
                pc.y_start_min = ya;
                pc.y_start_max = std::min(p.y_start_max, ya + rows);
                launches += launch_search_tiled(plan, sv, pc, d_vx, d_vy,
                                                d_slots + (size_t)(ya - p.y_start_min) * sw * p.results_per_pixel, clip, scr,
                                                st, nullptr, overflowed, true);
            }
            return launches;
        }
    }
    dim3 grid((sw + TBX - 1) / TBX, (sh + TBY - 1) / TBY);
    if (grid.y > 65535) throw Error("search area too tall for the tiled kernel grid");
    {
        // Queue of the deferred exact evaluations.  Sigma-G on: {pixel, candidate} of everything whose bound reaches
        // min_lh (C4 with 20 bright movers raises 108M requests); off: every candidate whose upper bound reached the
        // pixel's rising bar, with that bound (a few hundred per pixel).  Every (pixel, candidate) pair if that is
        // small, else a fixed share of the 180 GB; past a full queue the kernel evaluates in line, so the capacity
        // only affects speed.  Survivors are a small fraction: their buffer starts at 1/8 (it grows on demand).
        const uint64_t all_pairs = (uint64_t)sw * sh * (uint64_t)std::max(plan->v.n_groups * plan->tg, 1);
        const size_t elem = plan->sigmag ? sizeof(uint2) : sizeof(ExactReq);
        const uint64_t want = std::min<uint64_t>(all_pairs, plan->sigmag ? (256ull << 20) : kExactQueueCap);
        if (!clip.counters) KB_CUDA(cudaMalloc(&clip.counters, 4 * sizeof(unsigned long long)));
        if (clip.queue_bytes < want * elem) {
            if (clip.queue) cudaFree(clip.queue);
            clip.queue = nullptr;
            clip.queue_bytes = 0;
            KB_CUDA(cudaMalloc(&clip.queue, want * elem));
            clip.queue_bytes = want * elem;
        }
        clip.cap = want;
        const uint64_t want_surv = std::min<uint64_t>(want, std::max<uint64_t>(want / 8, 4ull << 20));
        if (clip.surv_cap < want_surv) {
            if (clip.surv) cudaFree(clip.surv);
            clip.surv = nullptr;
            clip.surv_cap = 0;
            KB_CUDA(cudaMalloc(&clip.surv, sizeof(ClipSurvivor) * want_surv));
            clip.surv_cap = want_surv;
        }
        if (!plan->sigmag && clip.bar_cap < (uint64_t)sw * sh) {
            if (clip.bar) cudaFree(clip.bar);
            clip.bar = nullptr;
            clip.bar_cap = 0;
            KB_CUDA(cudaMalloc(&clip.bar, sizeof(float) * (uint64_t)sw * sh));
            clip.bar_cap = (uint64_t)sw * sh;
        }
        KB_CUDA(cudaMemsetAsync(clip.counters, 0, 4 * sizeof(unsigned long long), st));
        plan->v.clipq = plan->sigmag ? static_cast<uint2 *>(clip.queue) : nullptr;
        plan->v.exq = plan->sigmag ? nullptr : static_cast<ExactReq *>(clip.queue);
        plan->v.bar_out = clip.bar;
        plan->v.clip_cnt = clip.counters;
        plan->v.clip_cap = clip.cap;
    }
    if (!plan->ev0) {
        KB_CUDA(cudaEventCreate(&plan->ev0));
        KB_CUDA(cudaEventCreate(&plan->ev1));
    }
    KB_CUDA(cudaEventRecord(plan->ev0, st));
    int launches = 0;
    if (!up || up->n_chunks <= 1) {
        if (up && up->n_chunks == 1) KB_CUDA(cudaStreamWaitEvent(st, up->ev[0], 0));
        plan->v.pixel_base = 0;
        launch_tiled_any(plan, sv, p, d_vx, d_vy, d_slots, st, !up);
        launches = 1;
    } else {
        // The stack is still arriving: search row bands as their rows land.  Band b ends where a tile's reach stays
        // inside the rows of chunk b; bands alternate between two streams so one band's tail overlaps the next.
        KB_CUDA(cudaEventRecord(up->ready, st));  // candidates, counters, slots are ordered before the bands
        bool used[2] = {false, false};
        int ya = p.y_start_min, nb_launched = 0;
        for (int b = 0; b < up->n_chunks; ++b) {
            const bool last = (b == up->n_chunks - 1);
            int yb = last ? p.y_start_max : std::min(p.y_start_max, up->row_end[b] - plan->reach_hi);
            if (!last && yb > ya) yb = ya + (yb - ya) / TBY * TBY;  // whole tiles, so the tiling equals one launch's
            if (yb <= ya) continue;
            kb_params pb = p;
            pb.y_start_min = ya;
            pb.y_start_max = yb;
            const int k = nb_launched & 1;
            cudaStream_t bs = up->band_stream[k];
            KB_CUDA(cudaStreamWaitEvent(bs, up->ready, 0));
            KB_CUDA(cudaStreamWaitEvent(bs, up->ev[b], 0));
            const size_t row_off = (size_t)(ya - p.y_start_min) * sw;
            plan->v.pixel_base = (unsigned)row_off;
            launch_tiled_any(plan, sv, pb, d_vx, d_vy, d_slots + row_off * p.results_per_pixel, bs, false);
            used[k] = true;
            ++nb_launched;
            ++launches;
            ya = yb;
        }
        plan->v.pixel_base = 0;
        for (int k = 0; k < 2; ++k)
            if (used[k]) {
                KB_CUDA(cudaEventRecord(up->band_done[k], up->band_stream[k]));
                KB_CUDA(cudaStreamWaitEvent(st, up->band_done[k], 0));
            }
        KB_CUDA(cudaStreamWaitEvent(st, up->ev[up->n_chunks - 1], 0));  // clips and the epilogue read any row
    }
    KB_CUDA(cudaEventRecord(plan->ev1, st));
    {
        unsigned long long cnt[4];
        for (int attempt = 0; attempt < 2; ++attempt) {
            if (plan->sigmag) {
                switch (sv.nb) {
                    case 4: launch_clip_nb4(sv, p, d_vx, d_vy, clip, st); break;
                    case 2: launch_clip_nb2(sv, p, d_vx, d_vy, clip, st); break;
                    default: launch_clip_nb1(sv, p, d_vx, d_vy, clip, st); break;
                }
            } else {
                switch (sv.nb) {
                    case 4: launch_exact_nb4(sv, p, d_vx, d_vy, clip, st); break;
                    case 2: launch_exact_nb2(sv, p, d_vx, d_vy, clip, st); break;
                    default: launch_exact_nb1(sv, p, d_vx, d_vy, clip, st); break;
                }
            }
            ++launches;
            KB_CUDA(cudaMemcpyAsync(cnt, clip.counters, sizeof(cnt), cudaMemcpyDeviceToHost, st));
            KB_CUDA(cudaStreamSynchronize(st));
            if (attempt == 0) {   // the stream is idle: the gather kernel's events have fired
                float ms = 0.0f;
                if (cudaEventElapsedTime(&ms, plan->ev0, plan->ev1) == cudaSuccess) plan->ms_accum += ms;
            }
            if (cnt[1] <= clip.surv_cap) break;
            // more survivors than the buffer holds: enlarge it and work the (intact) queue off again
            cudaFree(clip.surv);
            clip.surv = nullptr;
            clip.surv_cap = 0;
            KB_CUDA(cudaMalloc(&clip.surv, sizeof(ClipSurvivor) * cnt[1]));
            clip.surv_cap = cnt[1];
            KB_CUDA(cudaMemsetAsync(clip.counters + 1, 0, sizeof(unsigned long long), st));
        }
        if (getenv("KB_DEBUG"))
            fprintf(stderr, "[kbmod_b200] deferred exact evaluations (%s): requests=%llu (queue cap %llu) survivors=%llu in-line=%llu"
                            " certain rejections (counted by a KB_CLIP_VERIFY build only)=%llu\n",
                    plan->sigmag ? "sigma-G clips" : "bounds reached the bar", cnt[0], (unsigned long long)clip.cap, cnt[1], cnt[2], cnt[3]);
        if (cnt[0] <= clip.cap) plan->requests_accum += cnt[0];   // (an overflowed area is counted by its halves)
        if (cnt[0] > clip.cap) {
            // The queue was too small and requests were dropped.  Search the two halves of the area again, each with
            // the whole queue to itself (the slots are simply overwritten); a single row of tiles that still does not
            // fit -- or an upload in flight, whose bands are already sized -- is left to the caller's fallback.
            if (!up && sh > TBY) {
                const int half = std::max(TBY, (sh / 2 + TBY - 1) / TBY * TBY);
                kb_params pa = p, pb = p;
                pa.y_start_max = p.y_start_min + half;
                pb.y_start_min = p.y_start_min + half;
                launches += launch_search_tiled(plan, sv, pa, d_vx, d_vy, d_slots, clip, scr, st, nullptr, overflowed, true);
                launches += launch_search_tiled(plan, sv, pb, d_vx, d_vy,
                                                d_slots + (size_t)half * sw * p.results_per_pixel, clip, scr, st, nullptr,
                                                overflowed, true);
                return launches;
            }
            if (overflowed) *overflowed = true;
            return launches;
        }
        launches += merge_clip_survivors(clip.surv, cnt[1], d_slots, (int)p.results_per_pixel, scr, st);
    }
    return launches;
}

}  // namespace kb
