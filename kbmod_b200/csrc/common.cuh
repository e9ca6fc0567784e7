// common.cuh -- internal declarations shared by the translation units of libkbmod_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/kbmod_b200.h"

namespace kb {

// The reference's NO_DATA is C's NAN (common.h:34): quiet NaN 0x7fc00000.  Use the same bit pattern so float
// stacks are memcmp-equal (CUDART_NAN_F is 0x7fffffff).
#define KB_NAN_F __int_as_float(0x7fc00000)

// ---- error plumbing -------------------------------------------------------------------------------
void set_error(const std::string &msg);

struct Error : public std::runtime_error {
    explicit Error(const std::string &m) : std::runtime_error(m) {}
};

#define KB_CUDA(expr)                                                                                  \
    do {                                                                                               \
        cudaError_t _e = (expr);                                                                       \
        if (_e != cudaSuccess)                                                                         \
            throw kb::Error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + __FILE__ + \
                            ":" + std::to_string(__LINE__) + " (" #expr ")");                          \
    } while (0)

#define KB_API_BEGIN try {
#define KB_API_END                          \
    }                                       \
    catch (const std::exception &e) {       \
        kb::set_error(e.what());            \
        return 1;                           \
    }                                       \
    catch (...) {                           \
        kb::set_error("unknown error");     \
        return 2;                           \
    }                                       \
    return 0;

// ---- device view of the psi/phi stack ----------------------------------------------------------------
// Search copy in HBM: [t][row][Wp] pixel pairs (psi, phi); pair = float2 / ushort2 / uchar2.  Rows are padded
// to Wp pixels so the row pitch is a multiple of 16 bytes (TMA global-stride rule).  Pairs are canonical:
// if either member is invalid (non-finite float, code 0) both are (NaN,NaN) / (0,0), so one test decides
// whether the observation counts (kernels.cu:184: both must be finite).
struct StackView {
    const void *data;
    const double *times;
    int T, H, W, Wp;
    int nb;  // bytes per value: 4, 2, 1
    float psi_min, psi_scale, phi_min, phi_scale;
    // float stacks only: 1 = invalid pairs are stored as (+0, +0) instead of (NaN, NaN).  Adding +0 to a running
    // float sum is an exact no-op, so the search kernel can accumulate without testing validity.  Chosen at upload
    // time when no *valid* pair is (0, 0) (upload_stack checks), which keeps "pair is zero" == "pair is invalid".
    int zero_marker;
    // Screening copy [t][row][Wp] of 32-bit words (psi_phi.cu quantise_kernel): psi and phi as signed 16-bit
    // multiples of q_step_psi / q_step_phi, invalid pair = 0.  Only the tiled kernel's gather loop reads it.
    const uint32_t *screen;
    float q_step_psi, q_step_phi;
    int all_valid;   // 1 = no invalid pair anywhere in the rows this stack holds (known from the constructor's scan)
    int n_runs;      // run planes that follow the T image planes of `screen` (psi_phi.cu run_planes_kernel)
};

// Per-pixel result slot written by the search kernels: 16 bytes, [pixel][K].
struct __align__(16) Slot {
    float lh;
    float flux;
    int obs;
    int cand;  // candidate index, -1 = empty slot
};

static constexpr int kMaxRegK = 8;      // results_per_pixel kept in registers
static constexpr int kMaxTSigma = 1024; // largest T for which the in-kernel sigma-G path is instantiated

// ---- launchers (each returns the number of kernels it launched) ------------------------------------
// search_generic.cu
int launch_search_generic(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                          uint64_t n_cand, Slot *d_slots, cudaStream_t st);
int launch_evaluate(const StackView &sv, const kb_params &p, kb_trajectory *d_trjs, uint64_t n, uint8_t *d_mask,
                    cudaStream_t st);
int launch_sigmag_indices(const float *d_vals, int n, float sgl0, float sgl1, float coeff, float width, int *d_idx,
                          int *d_range, cudaStream_t st);
int launch_curves(const void *d_ref_layout, const double *d_times, int T, int H, int W, int nb, float psi_min,
                  float psi_scale, float phi_min, float phi_scale, const kb_trajectory *d_trjs, uint64_t n,
                  float *d_out, cudaStream_t st);

// Deferred sigma-G clips (search_tiled.cu -> results.cu).  The tiled kernel does not clip in line: a candidate whose
// unclipped lh reaches min_lh is appended to a global queue; a second kernel clips the queue with every SM (the
// requests cluster around bright sources, so clipping inside the tile's CTA leaves a few SMs with all the work);
// the few survivors are then merged into the per-pixel slots.
struct ClipSurvivor {
    uint32_t pixel;  // start-pixel index yi * search_width + xi
    int cand;
    float lh, flux;
    int obs;
};
// sigma-G off: a candidate whose upper bound of lh reached the pixel's bar during the sweep, for exact_kernel
struct __align__(16) ExactReq {
    uint32_t pixel;  // start-pixel index yi * search_width + xi
    int cand;
    float ub;        // upper bound of the candidate's lh from the screening sums
    uint32_t pad;
};
struct ClipScratch {
    void *queue = nullptr;   // sigma-G on: uint2 {pixel, candidate}; off: ExactReq
    size_t queue_bytes = 0;
    ClipSurvivor *surv = nullptr;
    unsigned long long *counters = nullptr;  // [0] requests (may exceed cap), [1] survivors, [2] evaluated in line
    uint64_t cap = 0;       // queue capacity (requests) of the current search
    uint64_t surv_cap = 0;  // survivor capacity; the clip kernel is rerun with a larger buffer if it overflows
    float *bar = nullptr;   // sigma-G off: final bar of every start pixel (its 8th largest lower bound of lh)
    uint64_t bar_cap = 0;
};
void clip_scratch_free(ClipScratch &c);
struct ResultScratch;
int merge_clip_survivors(const ClipSurvivor *d_surv, uint64_t n, Slot *d_slots, int K, ResultScratch &scr,
                         cudaStream_t st);

// search_tiled.cu
struct TiledPlan;  // opaque: schedule for the TMA-staged kernel
// Returns nullptr (and a reason) when the tiled kernel cannot be used for this search.
// h_times: host copy of the T image times (avoids a device read-back while an upload is in flight).
// h_runs: the runs of closely spaced exposures the screening copy has planes for ({first image, count}, sv.n_runs).
TiledPlan *tiled_plan_create(const StackView &sv, const kb_params &p, const float *h_vx, const float *h_vy,
                             uint64_t n_cand, cudaStream_t st, std::string *why_not, const double *h_times = nullptr,
                             const int2 *h_runs = nullptr);
// A stack upload in flight on another stream: rows [0, row_end[c]) of every image are on the device (and
// canonical) once ev[c] has fired.  With it, launch_search_tiled starts on the rows that have arrived: the frame is
// searched in row bands, each waiting only for the chunk that completes its reach.
struct UploadProgress {
    static constexpr int kMaxChunks = 8;
    int n_chunks = 0;
    int row_end[kMaxChunks] = {0};
    cudaEvent_t ev[kMaxChunks] = {nullptr};
    cudaStream_t band_stream[2] = {nullptr, nullptr};
    cudaEvent_t band_done[2] = {nullptr, nullptr}, ready = nullptr;
};
int launch_search_tiled(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx,
                        const float *d_vy, Slot *d_slots, ClipScratch &clip, ResultScratch &scr, cudaStream_t st,
                        const UploadProgress *up = nullptr, bool *overflowed = nullptr, bool is_chunk = false);
void tiled_plan_destroy(TiledPlan *plan);
unsigned long long tiled_plan_requests(TiledPlan *plan);  // deferred exact evaluations queued by the last launch
float tiled_plan_kernel_ms(TiledPlan *plan);  // duration of the gather kernel of the last launch (CUDA events)

// psi_phi.cu
int launch_convolve(const float *d_img, int H, int W, const float *d_psf, int dim, float psf_total, float *d_out,
                    cudaStream_t st);
// psi/phi for one image, written interleaved (float2) at d_out_pairs[H*W]; or to separate planes.
// Only output rows [y_write_lo, y_write_hi) are written (band-limited construction: the other rows of the input are
// the convolution's halo); y_write_hi < 0 = all rows.
int launch_psi_phi(const float *d_sci, const float *d_var, int H, int W, const float *d_psf, const float *d_psf2,
                   int dim, float psf_total, float psf2_total, float2 *d_out_pairs, float *d_psi_plane,
                   float *d_phi_plane, cudaStream_t st, int y_write_lo = 0, int y_write_hi = -1);
// zero-pair flag + per-channel min/max + number of invalid pairs (saturating) of rows [r0, r0 + nr) of a float stack
// in reference order: d_out[6]
int launch_scan_raw_rows(const float2 *d_pairs, int T, int H, int r0, int nr, int W, uint32_t *d_out, cudaStream_t st);
int launch_interleave(const float *d_psi, const float *d_phi, uint64_t n, float2 *d_out, cudaStream_t st);
// min/max over finite values of both channels: d_mm = {psi_min, psi_max, phi_min, phi_max} as ordered uints.
int launch_minmax(const float2 *d_pairs, uint64_t n, uint32_t *d_mm, cudaStream_t st);
int launch_encode(const float2 *d_pairs, uint64_t n, int nb, const float psi_p[3], const float phi_p[3], void *d_out,
                  cudaStream_t st);
// rows of `src_pitch` pixel pairs -> canonical search copy with pitch Wp (in place when src == dst and pitches match).
int launch_canonicalise(const void *d_src, int src_pitch, void *d_dst, int T, int H, int W, int Wp, int nb,
                        int zero_marker, cudaStream_t st);
int launch_canonicalise_rows(const void *d_src, int src_pitch, void *d_dst, int T, int H, int W, int Wp, int nb,
                             int zero_marker, int r0, int nr, cudaStream_t st);
// The screening copy (one 32-bit word per pair: two signed 16-bit fixed-point numbers, 0 = invalid) of rows
// [r0, r0 + nr) of every image, from the canonical search copy; and the min/max reduction over a float search copy.
int launch_quantise_rows(const void *d_search, uint32_t *d_screen, int T, int H, int Wp, int nb, int zero_marker,
                         float psi_scale, float psi_min, float phi_scale, float phi_min, float step_psi,
                         float step_phi, int r0, int nr, cudaStream_t st);
int launch_minmax_pitched(const void *d_search, uint64_t n_pairs, uint32_t *d_mm, cudaStream_t st);
// Run planes of the screening copy (planes T .. T + n_runs - 1): per pixel the quantised sum over the images of a run
// of closely spaced exposures; d_runs[c] = {first image, number of images}.
int launch_run_planes_rows(const void *d_search, uint32_t *d_screen, int T, int H, int Wp, int nb, int zero_marker,
                           float psi_scale, float psi_min, float phi_scale, float phi_min, float step_psi, float step_phi,
                           const int2 *d_runs, int n_runs, int r0, int nr, cudaStream_t st);
// float stacks: sets *d_flag to 1 when some valid pair is exactly (0, 0) (then the zero marker cannot be used).
int launch_detect_valid_zero_pairs(const void *d_src, int src_pitch, int T, int H, int W, uint32_t *d_flag,
                                   cudaStream_t st);
float ordered_to_float(uint32_t u);

// results.cu
struct ResultScratch {
    uint32_t *keys[2] = {nullptr, nullptr};
    uint32_t *vals[2] = {nullptr, nullptr};
    uint32_t *block_counts = nullptr;  // compaction
    uint32_t *hist = nullptr;          // radix histograms
    uint32_t *scan_tmp = nullptr;      // tile sums of the multi-CTA scan
    uint64_t cap_items = 0, cap_blocks = 0, cap_hist = 0, cap_scan_tmp = 0;
    uint32_t *d_count = nullptr;  // number of survivors
    uint32_t *d_flag = nullptr;   // invalid-trajectory flag
};
void result_scratch_free(ResultScratch &s);
void result_scratch_reserve(ResultScratch &s, uint64_t items);   // room for `items` key/index pairs
int radix_sort_u32_pairs(ResultScratch &scr, uint64_t n, cudaStream_t st);  // scr.keys[0]/vals[0], stable, ascending
// search_pairs.cu: candidate-parallel search (one thread per (start pixel, candidate) pair + two stable sorts)
static constexpr uint64_t kMaxSearchPairs = 1ull << 27;
bool search_pairs_usable(const kb_params &p, uint64_t n_cand, std::string *why_not);
int launch_search_pairs(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, uint64_t n_cand,
                        Slot *d_slots, Slot *d_pairs, ResultScratch &scr, cudaStream_t st);
// clustering_grid.py: indices of the best trajectory of every spatial bin, bins in first-seen order.
int grid_filter_indices(const kb_trajectory *d_list, uint64_t n, double bin_width, double max_time,
                        uint32_t *d_idx_out, uint64_t *n_out, ResultScratch &scr, cudaStream_t st);
// slots[P*K] -> filtered, sorted kb_trajectory list. Two phases so the caller can size the output:
// filter_count() returns the number of survivors (syncs the stream); filter_finish() writes them to d_out.
int filter_count(const Slot *d_slots, uint64_t n_slots, float min_lh, int min_obs, ResultScratch &scr,
                 uint64_t *n_out, cudaStream_t st);
int filter_finish(const Slot *d_slots, uint64_t n_slots, uint64_t n_keep, int K, int sw, int x0, int y0, float min_lh,
                  int min_obs, const float *d_vx, const float *d_vy, ResultScratch &scr, kb_trajectory *d_out,
                  bool *invalid, cudaStream_t st);
int slots_to_trajectories(const Slot *d_slots, uint64_t n_slots, int K, int sw, int x0, int y0, const float *d_vx,
                          const float *d_vy, kb_trajectory *d_out, cudaStream_t st);
// Multi-GPU merge (nccl_merge.cu): blocks of {32-byte header with the record count, sorted records}; every record
// computes its own position in the merged list (stable with respect to block order).
int merge_header_bytes();
int merge_pack_block(void *d_block, const kb_trajectory *d_list, uint64_t n, cudaStream_t st);
int merge_sorted_runs(const void *d_blocks, size_t block_bytes, int world, uint64_t cap_in, kb_trajectory *d_out,
                      uint64_t keep, uint64_t *d_total, cudaStream_t st);
// Sort (key desc by lh) an existing device list of trajectories in place.
int sort_trajectories_by_lh(kb_trajectory *d_list, uint64_t n, ResultScratch &scr, kb_trajectory *d_tmp,
                            cudaStream_t st);

}  // namespace kb
