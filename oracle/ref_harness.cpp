/*
 * ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never imported by the product path).
 *
 * Builds the reference's own single translation unit *in place* (bindings.cpp is found through
 * -I/root/reference/src/kbmod/search; no reference source is copied into this repository) and adds a
 * second, tiny Python module `ref_harness` to the same shared object.  The extra module lets the tests
 *   (1) read the raw bytes of a reference PsiPhiArray (psi_phi_array_ds.h:120 get_cpu_array_ptr), and
 *   (2) run the reference's *GPU-path semantics* on the host: the reference's host-compiled
 *       evaluateTrajectory (kernels/kernels.cu:154-242, it is __host__ __device__) called once per
 *       (start pixel, candidate), followed by the per-pixel sorted insertion of kernels.cu:293-330,
 *       which is restated below in our own words because it lives inside a __global__ function.
 * The output is the raw K*P slot array, i.e. exactly what searchFilterImages leaves in `results`
 * before StackSearch::search_all filters and sorts it (stack_search.cpp:268-277).
 */
#include "bindings.cpp"

#include <cfloat>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace refh {
using search::PsiPhiArray;
using search::SearchParameters;
using search::Trajectory;

// Raw bytes of the host psi/phi array in the reference layout [t][row][col][psi,phi].
py::array_t<uint8_t> psi_phi_bytes(PsiPhiArray &ppa) {
    const size_t n = ppa.get_total_array_size();
    py::array_t<uint8_t> out(n);
    if (ppa.get_cpu_array_ptr() == nullptr) throw std::runtime_error("psi/phi not allocated on host");
    std::memcpy(out.mutable_data(), ppa.get_cpu_array_ptr(), n);
    return out;
}

// One trajectory through the host-compiled device function (no has_gpu() gate, unlike
// StackSearch::evaluate_single_trajectory, stack_search.cpp:193-209).
Trajectory eval_host(PsiPhiArray &ppa, SearchParameters params, Trajectory trj) {
#ifdef HAVE_CUDA
    search::evaluateTrajectory(ppa.get_meta_data(), ppa.get_cpu_array_ptr(), ppa.get_cpu_time_array_ptr(),
                               params, &trj);
#else
    throw std::runtime_error("harness needs the CUDA build of the reference");
#endif
    return trj;
}

// GPU-path semantics for every start pixel in the bounds of `params`; returns the K*P slots as
// seven parallel numpy arrays (x, y, vx, vy, lh, flux, obs_count) in slot order
// ((y_i * search_width + x_i) * K + r), kernels.cu:285.
py::tuple search_slots(PsiPhiArray &ppa, SearchParameters params, std::vector<Trajectory> &cands) {
    const int sw = params.x_start_max - params.x_start_min;
    const int sh = params.y_start_max - params.y_start_min;
    if (sw <= 0 || sh <= 0) throw std::runtime_error("Invalid search bounds");
    const int K = params.results_per_pixel;
    const size_t total = (size_t)K * sw * sh;
    std::vector<Trajectory> slots(total);
    const long long n = (long long)cands.size();
    search::PsiPhiArrayMeta meta = ppa.get_meta_data();
    void *vect = ppa.get_cpu_array_ptr();
    double *times = ppa.get_cpu_time_array_ptr();

#pragma omp parallel for collapse(2) schedule(dynamic, 8)
    for (int yi = 0; yi < sh; ++yi) {
        for (int xi = 0; xi < sw; ++xi) {
            const int x = xi + params.x_start_min, y = yi + params.y_start_min;
            Trajectory *best = &slots[((size_t)yi * sw + xi) * K];
            for (int r = 0; r < K; ++r) {
                best[r].x = x; best[r].y = y; best[r].vx = 0.0f; best[r].vy = 0.0f;
                best[r].lh = -FLT_MAX; best[r].flux = 0.0f; best[r].obs_count = 0;
            }
            for (long long c = 0; c < n; ++c) {
                Trajectory cur;
                cur.x = x; cur.y = y; cur.vx = cands[c].vx; cur.vy = cands[c].vy; cur.obs_count = 0;
#ifdef HAVE_CUDA
                search::evaluateTrajectory(meta, vect, times, params, &cur);
#endif
                if (cur.obs_count < params.min_observations ||
                    (params.do_sigmag_filter && cur.lh < params.min_lh))
                    continue;
                for (int r = 0; r < K; ++r) {
                    if (cur.lh > best[r].lh) std::swap(cur, best[r]);
                }
            }
        }
    }
    py::array_t<int> ox(total), oy(total), oo(total);
    py::array_t<float> ovx(total), ovy(total), olh(total), ofl(total);
    for (size_t i = 0; i < total; ++i) {
        ox.mutable_data()[i] = slots[i].x; oy.mutable_data()[i] = slots[i].y;
        oo.mutable_data()[i] = slots[i].obs_count;
        ovx.mutable_data()[i] = slots[i].vx; ovy.mutable_data()[i] = slots[i].vy;
        olh.mutable_data()[i] = slots[i].lh; ofl.mutable_data()[i] = slots[i].flux;
    }
    return py::make_tuple(ox, oy, ovx, ovy, olh, ofl, oo);
}

SearchParameters make_params(int min_obs, float min_lh, bool sigmag, float sgl_L, float sgl_H, float coeff,
                             int num_bytes, int x0, int x1, int y0, int y1, int K) {
    SearchParameters p;
    p.min_observations = min_obs; p.min_lh = min_lh; p.do_sigmag_filter = sigmag;
    p.sgl_L = sgl_L; p.sgl_H = sgl_H; p.sigmag_coeff = coeff; p.encode_num_bytes = num_bytes;
    p.x_start_min = x0; p.x_start_max = x1; p.y_start_min = y0; p.y_start_max = y1;
    p.results_per_pixel = K; p.total_results = 0;
    return p;
}

// OpenMP threads the reference's CPU loops will use (search_cpu_only, cpu_search_algorithms.cpp:96-124, is an
// `omp parallel for`).  Launchers such as torchrun export OMP_NUM_THREADS=1; the benchmark's reference arm sets
// and reports the real number through these two.
int omp_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void set_omp_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
}  // namespace refh

PYBIND11_MODULE(ref_harness, m) {
    m.def("psi_phi_bytes", &refh::psi_phi_bytes);
    m.def("eval_host", &refh::eval_host);
    m.def("search_slots", &refh::search_slots);
    m.def("make_params", &refh::make_params);
    m.def("omp_threads", &refh::omp_threads);
    m.def("set_omp_threads", &refh::set_omp_threads);
}
