/*
 * kbmod_b200.h -- C-ABI of libkbmod_b200.so: KBMOD's shift-and-stack trajectory search for NVIDIA B200
 * (sm_100a).  Plain pointers and sizes only; every function returns 0 on success and a non-zero code on
 * failure, with the message available from kb_last_error() (the Python shim raises RuntimeError with it,
 * matching the reference, whose errors are std::runtime_error -> RuntimeError).
 *
 * Each entry point names the reference interface it replaces (paths under src/kbmod/search/ of
 * dirac-institute/kbmod @ 617cd33).  Host pointers are borrowed for the duration of the call only.
 * A handle is bound to the CUDA device that was current when it was created and is not thread-safe
 * (the reference object is not re-entrant either: stack_search.h:68-82 mutable results/params).
 */
#ifndef KBMOD_B200_H_
#define KBMOD_B200_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define KB_EXPORT __attribute__((visibility("default")))
#else
#define KB_EXPORT
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* common.h:55-67  struct Trajectory (28 bytes, tests/test_trajectory_list.py:28,47) */
typedef struct {
    float vx, vy, lh, flux;
    int32_t x, y, obs_count;
} kb_trajectory;

/* common.h:119-161  struct SearchParameters */
typedef struct {
    int32_t min_observations;
    float min_lh;
    int32_t do_sigmag_filter;
    float sgl_L, sgl_H, sigmag_coeff;
    int32_t encode_num_bytes; /* -1, 1 or 2; informational, the stack knows its own encoding */
    int32_t x_start_min, x_start_max, y_start_min, y_start_max;
    uint32_t results_per_pixel;
    uint64_t total_results;
} kb_params;

/* psi_phi_array_ds.h:50-69  struct PsiPhiArrayMeta */
typedef struct {
    uint64_t num_times, width, height, pixels_per_image, num_entries, block_size, total_array_size;
    int32_t num_bytes; /* 4 (float), 2 (uint16), 1 (uint8) */
    float psi_min_val, psi_max_val, psi_scale;
    float phi_min_val, phi_max_val, phi_scale;
} kb_meta;

/* Device-side timings of the last kb_search on a handle (CUDA events on the handle's stream). */
typedef struct {
    float ms_h2d;          /* stack upload when not preloaded (psi_phi_array.cpp:74-111 move_to_gpu) */
    float ms_prepare;      /* candidate upload + schedule construction */
    float ms_search;       /* the search kernel(s) (kernels.cu:252-332 searchFilterImages) */
    float ms_filter_sort;  /* filter + compaction + sort + gather (stack_search.cpp:268-277) */
    float ms_total;        /* everything between entry and the last kernel, on the device */
    uint64_t evals;        /* start pixels * candidates * times of that search */
    uint32_t search_launches; /* number of search-kernel launches */
    uint32_t total_launches;  /* all kernels launched by the call */
    uint32_t used_tiled;      /* 1 when the TMA-staged tiled kernel ran, 2 the candidate-parallel one, 0 the generic one */
    float ms_main_kernel;     /* the single gather kernel launch alone (search_tiled_kernel / search_generic_kernel) */
    uint64_t h2d_stack_bytes; /* psi/phi bytes copied host -> device inside the call (0 when the stack was resident) */
} kb_timing;

typedef struct kb_stack_s *kb_handle;

/* ---- library ---------------------------------------------------------------------------------- */
KB_EXPORT const char *kb_last_error(void);
KB_EXPORT const char *kb_version(void);
/* kernel_helpers.cpp:23-29 has_gpu() / kernels/kernel_memory.cu:15-60 */
KB_EXPORT int kb_device_count(int *count);
KB_EXPORT int kb_gpu_memory(size_t *free_bytes, size_t *total_bytes);
/* Select the CUDA device for handles created afterwards by this thread (the reference always uses the current
 * device and never sets it: kernels/kernel_memory.cu; one process per GPU needs this). */
KB_EXPORT int kb_set_device(int device);
/* "domain:bus:device.function" of the current device (>= 16 bytes).  No reference counterpart: a multi-GPU host
 * process uses it to find the CPUs of the GPU's own socket (/sys/bus/pci/devices/<id>/local_cpulist) so that the
 * host copy of its stack and the pinned staging buffers are placed in that socket's memory. */
KB_EXPORT int kb_device_pci_bus_id(char *buf, int len);

/* ---- image operations (image_utils_cpp.cpp) ------------------------------------------------------ */
/* convolve_image_gpu (:70-101) with the arithmetic of convolve_image_cpu (:20-68): mask-aware correlation,
 * renormalised by the PSF portion used; NaN when no tap contributes. img/out: H*W float32; psf: dim*dim. */
KB_EXPORT int kb_convolve(const float *img, int H, int W, const float *psf, int psf_dim, float *out);
/* generate_psi (:126-153) and generate_phi (:155-177), fused: psi = conv(sci/var, psf), phi = conv(1/var, psf^2).
 * Either output pointer may be NULL. */
KB_EXPORT int kb_generate_psi_phi(const float *sci, const float *var, int H, int W, const float *psf, int psf_dim,
                        float *psi_out, float *phi_out);

/* ---- psi/phi stack (psi_phi_array.cpp, PsiPhiArray) ---------------------------------------------- */
/* StackSearch ctor (stack_search.cpp:37-75) -> fill_psi_phi_array_from_image_arrays (psi_phi_array.cpp:374-410).
 * sci/var: T pointers to H*W float32 images; psf: T pointers to psf_dim[t]^2 float32 kernels;
 * num_bytes in {-1, 4, 2, 1}. Builds psi/phi and the encoded array on the device and leaves the search copy there
 * (kb_on_gpu still reports 0 until kb_preload, like the reference; kb_unload frees it); the host copy in the
 * reference layout [t][row][col][psi,phi] is filled by a background thread that every reader joins. */
KB_EXPORT int kb_create_from_images(const float *const *sci, const float *const *var, const float *const *psf,
                          const int *psf_dim, int T, int H, int W, const double *zeroed_times, int num_bytes,
                          kb_handle *out);
/* The same for rows [row_lo, row_hi) of every image only (no reference counterpart: the reference is single-GPU and
 * always builds whole images).  A rank of the multi-GPU search needs psi/phi only for its band of start rows plus
 * the candidates' reach; the convolution is local, so it uploads and convolves just those rows (plus the PSF radius)
 * of the caller's full-size images.  float32 stacks only.  Searches on such a handle must stay inside its rows
 * (kb_search checks), and it cannot be re-uploaded after kb_unload. */
KB_EXPORT int kb_create_from_images_rows(const float *const *sci, const float *const *var, const float *const *psf,
                               const int *psf_dim, int T, int H, int W, const double *zeroed_times, int num_bytes,
                               int row_lo, int row_hi, kb_handle *out);
/* fill_psi_phi_array (psi_phi_array.cpp:321-372): from precomputed psi/phi images. */
KB_EXPORT int kb_create_from_psi_phi(const float *const *psi, const float *const *phi, int T, int H, int W,
                           const double *zeroed_times, int num_bytes, kb_handle *out);
KB_EXPORT int kb_destroy(kb_handle h);
KB_EXPORT int kb_get_meta(kb_handle h, kb_meta *meta);
/* PsiPhiArray::get_cpu_array_ptr (psi_phi_array_ds.h:120): borrowed pointer to the host copy. */
KB_EXPORT int kb_host_array(kb_handle h, const void **ptr, uint64_t *nbytes);
/* PsiPhiArray::read_psi_phi (psi_phi_array.cpp:172-205), read_time (:207-212). Out of bounds -> NaN pair. */
KB_EXPORT int kb_read_psi_phi(kb_handle h, uint64_t t, int row, int col, float out[2]);
KB_EXPORT int kb_read_time(kb_handle h, uint64_t t, double *out);
/* StackSearch::preload_psi_phi_array / unload_psi_phi_array / psi_phi_array_on_gpu (stack_search.cpp:174-186). */
KB_EXPORT int kb_preload(kb_handle h);
KB_EXPORT int kb_unload(kb_handle h);
KB_EXPORT int kb_on_gpu(kb_handle h, int *flag);

/* ---- search (stack_search.cpp:221-284 search_all with on_gpu=true; kernels.cu:252-397) ----------- */
/* Evaluates every (start pixel in bounds, candidate) pair with the reference's GPU-path semantics, keeps the
 * best results_per_pixel per start pixel, drops lh < min_lh and obs_count < min_observations, and sorts by
 * lh descending on the device.  n_results receives the number of surviving results, which stay on the
 * device until read with kb_get_results.  Unlike the reference there is no 200-image limit. */
KB_EXPORT int kb_search(kb_handle h, const kb_params *params, const float *vx, const float *vy, uint64_t n_candidates,
              uint64_t *n_results);
/* TrajectoryList::get_batch (trajectory_list.cpp:85-94): copy [start, start+count) to out; n_copied may be
 * smaller at the end of the list. */
KB_EXPORT int kb_get_results(kb_handle h, uint64_t start, uint64_t count, kb_trajectory *out, uint64_t *n_copied);
KB_EXPORT int kb_num_results(kb_handle h, uint64_t *n);
/* StackSearch::set_results / clear_results (stack_search.cpp:329-338). */
KB_EXPORT int kb_set_results(kb_handle h, const kb_trajectory *list, uint64_t n);
KB_EXPORT int kb_clear_results(kb_handle h);
KB_EXPORT int kb_get_timing(kb_handle h, kb_timing *t);
/* The raw K*P per-pixel slots of the last search in the reference's order ((y_i*sw + x_i)*K + r), i.e. what
 * searchFilterImages leaves in `results` before the host filter (kernels.cu:285-301,322-330). Debug/tests. */
KB_EXPORT int kb_get_slots(kb_handle h, kb_trajectory *out, uint64_t capacity, uint64_t *n_slots);
/* 0 = choose automatically, 1 = force the generic kernel, 2 = force the tiled TMA kernel (error if unusable),
 * 3 = force the candidate-parallel kernel (small start windows with long per-pixel lists, as TrajectoryExplorer
 * drives StackSearch: trajectory_explorer.py:221-255). */
KB_EXPORT int kb_set_kernel_mode(kb_handle h, int mode);
/* Why the last kb_search on this handle did not run on the tiled kernel, or ran on it without its screening paying
 * off (a sigma-G search that had to evaluate more than 5 % of its pairs exactly); "" otherwise.  Borrowed string,
 * valid until the next search.  The Python layer logs it (the reference has one kernel and nothing to report). */
KB_EXPORT int kb_search_note(kb_handle h, const char **note);

/* StackSearch::evaluate_single_trajectory(trj, use_kernel=true) (stack_search.cpp:193-209 -> evaluateTrajectory
 * kernels.cu:154-242): fills lh, flux, obs_count of each trajectory in place, on the device.  keep_mask
 * (optional, n*T bytes) receives 1 where an observation contributed to the final lh/flux. */
KB_EXPORT int kb_evaluate_trajectories(kb_handle h, const kb_params *params, kb_trajectory *trjs, uint64_t n,
                             uint8_t *keep_mask);
/* StackSearch::get_all_psi_phi_curves (stack_search.cpp:14-31,302-318): out is n rows of 2*T float32
 * (psi values then phi values; invalid -> 0). */
KB_EXPORT int kb_psi_phi_curves(kb_handle h, const kb_trajectory *trjs, uint64_t n, float *out);
/* sigmag_filtered_indices test wrapper (kernel_helpers.cpp:86-106) evaluated by the device routine the search
 * kernel uses: returns the indices kept, in ascending-value order. idx_out has room for n ints. */
KB_EXPORT int kb_sigmag_filtered_indices(const float *values, int n, float sgl0, float sgl1, float coeff, float width,
                               int *idx_out, int *n_kept);

/* ---- the filter that follows the search in the reference pipeline ----------------------- */
/* SigmaGClipping.compute_clipped_sigma_g_matrix (filters/sigma_g_filter.py:114-168): lh is n_rows x n_times
 * float32 (NaN = no data); valid_out (n_rows x n_times bytes) receives 1 where the point lies strictly inside
 * median -/+ n_sigma_coeff * max(q_high percentile - q_low percentile, 1e-5), percentiles as torch.nanquantile
 * computes them (linear interpolation, float32).  q_low/q_high are fractions in (0,1); n_sigma_coeff is the
 * product n_sigma * coeff narrowed to float32; clip_negative != 0 ignores values <= 0 in the percentiles only.
 * Host pointers; n_times <= 4096. */
KB_EXPORT int kb_sigmag_clip_matrix(const float *lh, uint64_t n_rows, uint32_t n_times, float q_low, float q_high,
                                    float n_sigma_coeff, int clip_negative, uint8_t *valid_out);
/* Results._update_likelihood (results.py:608-645): per row, float32 sums of psi*obs_valid and phi*obs_valid in
 * numpy's pairwise order; likelihood = psi_sum/sqrt(phi_sum), flux = psi_sum/phi_sum (0 where phi_sum == 0),
 * obs_count = number of valid points (n_times when obs_valid is NULL). */
KB_EXPORT int kb_update_likelihood(const float *psi, const float *phi, const uint8_t *obs_valid, uint64_t n_rows,
                                   uint32_t n_times, double *lh_out, double *flux_out, int64_t *obs_count_out);
/* apply_clipped_sigma_g (filters/sigma_g_filter.py:171-190) in one call: likelihood curves from psi/phi and the
 * current obs_valid (results.py:568-606), the clip above, then the likelihood update with the new mask.
 * obs_valid_io is read when has_obs_valid != 0 and always receives the new mask. */
KB_EXPORT int kb_sigmag_filter_curves(const float *psi, const float *phi, uint64_t n_rows, uint32_t n_times,
                                      float q_low, float q_high, float n_sigma_coeff, int clip_negative,
                                      uint8_t *obs_valid_io, int has_obs_valid, double *lh_out, double *flux_out,
                                      int64_t *obs_count_out);

/* SigmaGClipping.compute_clipped_sigma_g (filters/sigma_g_filter.py:86-112): one float64 likelihood curve of n points.
 * Percentiles as numpy.percentile computes them (method "linear", float64) over all points, or over the positive
 * ones when clip_negative != 0 (none positive: nothing is kept); kept are the points strictly inside
 * median -/+ n_sigma * coeff * max(upper - lower, 1e-8).  good_idx (room for n entries) receives their indices in
 * ascending order, *n_good their number.  Host pointers; n <= 65536. */
KB_EXPORT int kb_sigmag_clip_curve(const double *lh, uint32_t n, double low_pct, double high_pct, double n_sigma,
                                   double coeff, int clip_negative, int64_t *good_idx, uint32_t *n_good);

/* TrajectoryClusterGrid / apply_trajectory_grid_filter (filters/clustering_grid.py:13-176): every trajectory falls
 * into the spatial bin (int(x/w), int(y/w), int((x + max_time*vx)/w), int((y + max_time*vy)/w)); idx_out receives,
 * for every occupied bin in the order the bins are first seen, the index of its best trajectory (largest lh; the
 * earlier one wins a tie); *n_out = number of bins.  max_time*v is a double product, in add_trajectory (:72-75)
 * and in the vectorised add_trajectory_list (:109-113) alike (the velocity columns reach numpy as Python floats).
 * Host pointers; idx_out has room for n entries. */
KB_EXPORT int kb_grid_filter(const kb_trajectory *trjs, uint64_t n, double bin_width, double max_time,
                             uint32_t *idx_out, uint64_t *n_out);

/* ---- coadded stamps of the results ----------------------------------------------------- */
/* append_coadds (filters/stamp_filters.py:72-160, nightly = False) over coadd_sum / coadd_mean / coadd_median /
 * coadd_weighted (core/stamp_utils.py:241-349): for each of the n trajectories (x0, y0, vx, vy as the float64
 * columns of the reference's result table) the (2*radius+1)^2 stamp centred on int(v*t + p0 + 0.5) is cut from
 * every image selected by obs_valid (n x T bytes, NULL = all), NaN where there is no image, and coadded over time in
 * float64 exactly as the reference does; each output is n x S x S float32 (NULL = not wanted).  var may be NULL
 * unless out_weighted is given.  Host pointers; sci/var are T pointers to H x W float32 images. */
KB_EXPORT int kb_coadd_stamps(const float *const *sci, const float *const *var, int T, int H, int W,
                              const double *times, const double *x0, const double *y0, const double *vx,
                              const double *vy, uint64_t n, int radius, const uint8_t *obs_valid, float *out_sum,
                              float *out_mean, float *out_median, float *out_weighted);

/* append_all_stamps (filters/stamp_filters.py:163-211) over extract_stamp_stack (core/stamp_utils.py:16-84,352-439):
 * out is n x T x S x S float32, the stamp of every trajectory at every image, NaN where there is no image. */
KB_EXPORT int kb_extract_stamps(const float *const *sci, int T, int H, int W, const double *times, const double *x0,
                                const double *y0, const double *vx, const double *vy, uint64_t n, int radius,
                                float *out);

/* extract_curve_values (core/stamp_utils.py:87-141): out[r][t] = imgs[t][y_vals[r][t]][x_vals[r][t]] as float64 where
 * that pixel exists, NaN elsewhere.  x_vals / y_vals / out are n x T, row-major; imgs are T pointers to H x W float32
 * images (any stack: science, variance, psi, phi).  Host pointers. */
KB_EXPORT int kb_extract_curve_values(const float *const *imgs, int T, int H, int W, const int64_t *x_vals,
                                      const int64_t *y_vals, uint64_t n, double *out);

/* ---- measurement hooks (no reference counterpart) ---------------------------------------------------- */
/* CUDA events recorded on the handle's own stream, so a caller can time a region of calls on the device:
 * kb_timer_start records the start event; kb_timer_stop records the stop event, waits for it and returns the
 * elapsed milliseconds between the two. */
KB_EXPORT int kb_timer_start(kb_handle h);
KB_EXPORT int kb_timer_stop(kb_handle h, float *ms);
/* Roofline denominators measured on the current device, in this process (no reference counterpart; SURVEY.md 8d
 * asks for a measured L2 figure beside the HBM one).  gbps[0..5] receive, in GB/s: HBM streaming read, L2-resident
 * streaming read, L2-resident 256-byte row gather, and the shared-memory bandwidth of 64-, 32- and 16-bit loads
 * per lane (the access shapes of the float32, uint16 and uint8 gathers of the tiled search kernel).  Takes about
 * 0.2 s and 2 GB of device memory for its duration.  n = capacity of gbps (at least 6). */
KB_EXPORT int kb_measure_peaks(double *gbps, int n);

/* ---- for callers that cannot allocate device memory --------------------------------------------------------
 * No counterpart in the reference (its sort is a host-side __gnu_parallel::sort, trajectory_list.cpp:96-107).
 *
 * Sorts the `n` trajectories at DEVICE address `d_list` in place by lh, highest first, ties keeping their input
 * order -- the same stable radix sort kb_search uses for its result list -- without a single cudaMalloc/cudaFree.
 * It exists for two kinds of caller.  Services that give every last byte of HBM to their own memory pool at
 * start-up and hand this library buffers out of it: a late cudaMalloc simply fails there.  And latency-critical
 * loops that keep several streams busy: cudaMalloc and cudaFree synchronise the whole device, which stalls
 * every other stream each time a result list is sorted.
 *
 * The sort's working memory is a fixed array that is part of the library image: every GPU that loads
 * libkbmod_b200.so carries it (about 12 MB) until the process exits.  Consequences:
 *   - n may be at most KB_STATIC_SORT_MAX; larger lists are refused with an error (nothing is sorted).
 *   - All calls on a device share that one array.  Calls whose GPU work can overlap (different streams, or
 *     different host threads) race on it: issue them on one stream, or order them with events.
 *   - The first call on each device looks up where that device placed the array (cudaGetSymbolAddress).  Make
 *     that first call during set-up, before any CUDA stream capture begins on the device.  The CUDA runtime loads
 *     the library's device code lazily, so the array's 12 MB are claimed on a device at that first call (or at the
 *     first other kernel of this library's sorting code there), not when the library is dlopen'ed: a caller that
 *     hands all remaining HBM to its own pool must make the first call before it does so.
 *   - The looked-up address is remembered per device ordinal for the life of the process; after cudaDeviceReset
 *     it is stale, and this entry point must not be used again in that process.
 * The work is enqueued on `cuda_stream` (a cudaStream_t passed as void*; NULL = the default stream) and the call
 * returns without waiting for it. */
#define KB_STATIC_SORT_MAX (1u << 18)
KB_EXPORT int kb_sort_trajectories_static(kb_trajectory *d_list, uint64_t n, void *cuda_stream);

/* ---- multi-GPU merge (no reference counterpart: the reference is single-GPU) ---------------------- */
/* One process per GPU.  Rank 0 calls kb_nccl_unique_id and ships the 128 bytes to the other ranks by any
 * means; every rank then calls kb_nccl_init.  kb_merge_results_allgather all-gathers every rank's sorted results
 * over NCCL and leaves the global best `top_n` (0 = all) as this handle's result list on every rank.  With
 * top_n > 0 the host does not wait for anything before the gather (the counts travel in the blocks). */
KB_EXPORT int kb_nccl_unique_id(uint8_t id[128]);
KB_EXPORT int kb_nccl_init(const uint8_t id[128], int rank, int world_size);
KB_EXPORT int kb_nccl_finalize(void);
KB_EXPORT int kb_merge_results_allgather(kb_handle h, uint64_t top_n, uint64_t *n_results);
/* The merge kernel of kb_merge_results_allgather on one device, fed from host lists (each sorted by lh descending):
 * out receives the first top_n (0 = all) records of the merged list, equal lh values in (list, position) order;
 * *n_out their number.  out has room for min(top_n, sum of counts) records. */
KB_EXPORT int kb_merge_sorted_lists(const kb_trajectory *const *lists, const uint64_t *counts, int n_lists,
                                    uint64_t top_n, kb_trajectory *out, uint64_t *n_out);

#ifdef __cplusplus
}
#endif
#endif /* KBMOD_B200_H_ */
