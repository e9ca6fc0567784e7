// api.cu -- the C-ABI of libkbmod_b200.so (include/kbmod_b200.h): handle management, host orchestration of
// StackSearch (stack_search.cpp) and PsiPhiArray (psi_phi_array.cpp) on top of the kernels in this directory.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <omp.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <mutex>
#include <thread>

#include "common.cuh"

namespace kb {

static thread_local std::string g_last_error;
void set_error(const std::string &msg) { g_last_error = msg; }

template <typename T>
struct DevBuf {  // RAII device buffer that only grows
    T *p = nullptr;
    uint64_t cap = 0;
    void ensure(uint64_t n) {
        if (n <= cap) return;
        release();
        KB_CUDA(cudaMalloc(&p, sizeof(T) * n));
        cap = n;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    ~DevBuf() { release(); }
};

struct Timer {
    cudaEvent_t a = nullptr, b = nullptr;
    Timer() {
        cudaEventCreate(&a);
        cudaEventCreate(&b);
    }
    ~Timer() {
        cudaEventDestroy(a);
        cudaEventDestroy(b);
    }
    void start(cudaStream_t st) { cudaEventRecord(a, st); }
    void stop(cudaStream_t st) { cudaEventRecord(b, st); }
    float ms() {
        float m = 0.f;
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&m, a, b);
        return m;
    }
};

}  // namespace kb

using namespace kb;

struct kb_stack_s {
    int device = 0;
    cudaStream_t stream = nullptr;
    int T = 0, H = 0, W = 0, Wp = 0, nb = 4;
    kb_meta meta{};
    std::vector<double> times;
    void *host_array = nullptr;  // reference layout [t][row][col][psi,phi]
    bool host_pinned = false;
    // The constructor leaves the stack on the device and fills the host copy from a background thread; everything
    // that reads the host copy joins it first (ensure_host).
    std::thread host_thread;
    std::atomic<bool> host_done{false};
    std::string host_error;
    void *d_raw = nullptr;        // reference-layout bytes on the device while the background copy drains them
    cudaStream_t host_stream = nullptr;
    bool dev_cache = false;       // d_search / d_screen hold the whole canonical stack (built by the constructor)
    int row_lo = 0, row_hi = 0;   // rows of every image this handle holds (the whole image unless band-limited)
    bool all_valid = false;       // no invalid pair in the rows this handle holds (constructor's scan)
    bool keep_raw = false;        // enough HBM to keep d_raw for good (psi/phi curves are gathered from it)
    void *d_search = nullptr;  // canonical pitched search copy
    uint32_t *d_screen = nullptr;  // 16-bit fixed-point screening copy of the same pixels (StackView::screen)
    // Runs of closely spaced exposures (several per night): {first image, count}.  The screening copy carries one
    // summed plane per run after its T image planes (psi_phi.cu run_planes_kernel).
    std::vector<int2> runs;
    int2 *d_runs = nullptr;
    float q_step_psi = 1.0f, q_step_phi = 1.0f;  // its quantisation steps; known once the stack's value range is
    bool q_known = false;
    double *d_times = nullptr;
    bool on_gpu = false;
    int zero_marker = 0;  // form of the search copy (StackView::zero_marker)
    int kernel_mode = 0;
    // search state
    DevBuf<float> d_vx, d_vy;
    DevBuf<Slot> d_slots, d_pairs;
    std::string search_note;   // why the last search did not run on the tiled kernel (empty when it did)
    // The schedule of the tiled kernel depends on the candidate list, not on the start bounds' extent: repeated
    // searches with the same candidates (row bands, repeated steps) reuse it instead of rebuilding and re-uploading it.
    TiledPlan *plan_cache = nullptr;
    std::vector<float> plan_vx, plan_vy;
    kb_params plan_params{};
    const void *plan_screen = nullptr;
    float plan_steps[2] = {0.0f, 0.0f};
    uint64_t n_slots = 0;
    kb_params last_params{};
    bool have_slots = false;
    DevBuf<kb_trajectory> d_results, d_tmp;
    uint64_t n_results = 0;
    ResultScratch scratch;
    ClipScratch clip;
    kb_timing timing{};
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    // pipelined upload (search_all on a stack that is not preloaded): copy stream, per-chunk events, band streams
    int zm_known = -1;  // zero_marker decided by an earlier upload of this (immutable) stack; -1 = not yet
    cudaStream_t copy_stream = nullptr;
    UploadProgress up;
};

static size_t pair_bytes(int nb) { return 2 * (size_t)nb; }
static size_t screen_bytes(kb_handle h) {   // T image planes + one plane per run
    return ((size_t)h->T + h->runs.size()) * h->H * h->Wp * sizeof(uint32_t);
}

// Consecutive images closer than this (days) form a run: 72 minutes, several exposures of one field in a night.
static constexpr double kRunGapDays = 0.05;
static void find_runs(kb_handle h) {
    h->runs.clear();
    const char *off = getenv("KB_NO_RUN_PLANES");
    if (off && off[0] == '1') return;
    // Every maximal run of closely spaced exposures, and its aligned halves, quarters, ... down to pairs: a group of
    // candidates that does not hold its pixels through a whole night often does through half of it.  Longest first
    // at every start image (the schedule takes the first that fits).
    int i = 0;
    while (i < h->T) {
        int j = i + 1;
        while (j < h->T && j - i < 64 && h->times[j] - h->times[j - 1] >= 0.0 && h->times[j] - h->times[j - 1] <= kRunGapDays) ++j;
        const int m = j - i;
        if (m >= 2) {
            h->runs.push_back(make_int2(i, m));
            int len = 1;
            while (len * 2 < m) len *= 2;           // largest power of two below m
            for (; len >= 2; len /= 2)
                for (int s0 = i; s0 + len <= j; s0 += len) h->runs.push_back(make_int2(s0, len));
        }
        i = j;
    }
    // keep the extra planes within a quarter of the image planes' footprint times four (they are a speed-up only)
    size_t free_b = 0, total_b = 0;
    const size_t plane = (size_t)h->H * h->Wp * sizeof(uint32_t);
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && h->runs.size() * plane > free_b / 8) {
        std::vector<int2> whole;
        for (const int2 &r : h->runs)   // drop the sub-runs, keep the maximal runs
            if (whole.empty() || r.x >= whole.back().x + whole.back().y) whole.push_back(r);
        h->runs.swap(whole);
        if (h->runs.size() * plane > free_b / 8) h->runs.clear();
    }
}
static void upload_runs(kb_handle h, cudaStream_t st) {
    if (h->runs.empty() || h->d_runs) return;
    KB_CUDA(cudaMalloc(&h->d_runs, sizeof(int2) * h->runs.size()));
    KB_CUDA(cudaMemcpyAsync(h->d_runs, h->runs.data(), sizeof(int2) * h->runs.size(), cudaMemcpyHostToDevice, st));
}
static void build_screen_rows(kb_handle h, int r0, int nr, cudaStream_t st) {
    launch_quantise_rows(h->d_search, h->d_screen, h->T, h->H, h->Wp, h->nb, h->zero_marker, h->meta.psi_scale,
                         h->meta.psi_min_val, h->meta.phi_scale, h->meta.phi_min_val, h->q_step_psi, h->q_step_phi, r0,
                         nr, st);
    launch_run_planes_rows(h->d_search, h->d_screen, h->T, h->H, h->Wp, h->nb, h->zero_marker, h->meta.psi_scale,
                           h->meta.psi_min_val, h->meta.phi_scale, h->meta.phi_min_val, h->q_step_psi, h->q_step_phi,
                           h->d_runs, (int)h->runs.size(), r0, nr, st);
}

static void use_device(kb_handle h) { KB_CUDA(cudaSetDevice(h->device)); }

static StackView make_view(kb_handle h) {
    StackView sv;
    sv.data = h->d_search;
    sv.times = h->d_times;
    sv.T = h->T;
    sv.H = h->H;
    sv.W = h->W;
    sv.Wp = h->Wp;
    sv.nb = h->nb;
    sv.psi_min = h->meta.psi_min_val;
    sv.psi_scale = h->meta.psi_scale;
    sv.phi_min = h->meta.phi_min_val;
    sv.phi_scale = h->meta.phi_scale;
    sv.zero_marker = h->zero_marker;
    sv.screen = h->d_screen;
    sv.q_step_psi = h->q_step_psi;
    sv.q_step_phi = h->q_step_phi;
    sv.all_valid = h->all_valid ? 1 : 0;
    sv.n_runs = (int)h->runs.size();
    return sv;
}

// ---- pinned staging rings ------------------------------------------------------------------------------------
// Caller-owned images and the host copy of the stack are ordinary pageable memory.  The driver moves such memory
// through one internal bounce buffer with a single-threaded memcpy (10 GB/s here); a small ring of page-locked slots
// filled (or drained) by all host threads keeps the PCIe link busy instead.  One ring per direction and device.
struct PinnedRing {
    static constexpr int kSlots = 4;
    static constexpr size_t kSlotBytes = 8u << 20;
    void *slot[kSlots] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev[kSlots] = {nullptr, nullptr, nullptr, nullptr};
    bool busy[kSlots] = {false, false, false, false};
    int next = 0;
    std::mutex mu;
    void init() {
        if (slot[0]) return;
        for (int i = 0; i < kSlots; ++i) {
            KB_CUDA(cudaHostAlloc(&slot[i], kSlotBytes, cudaHostAllocPortable));
            KB_CUDA(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
        }
    }
};
static PinnedRing g_ring_up[64], g_ring_down[64];

// `share`: fraction of the host threads to use (the background fill of the host copy takes a quarter, so the
// caller's own thread -- which is launching the first search meanwhile -- is not crowded out).
static void parallel_memcpy(void *dst, const void *src, size_t n, int share_div = 1) {
    const size_t piece = 256u << 10;
    const long long np = (long long)((n + piece - 1) / piece);
    if (np <= 2) {
        memcpy(dst, src, n);
        return;
    }
    const int nthreads = std::max(1, omp_get_max_threads() / std::max(share_div, 1));
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long long i = 0; i < np; ++i) {
        const size_t o = (size_t)i * piece;
        memcpy((char *)dst + o, (const char *)src + o, std::min(piece, n - o));
    }
}

// pageable host -> device, asynchronous with respect to `st` (returns when the last slot has been filled)
static void ring_h2d(int device, void *d_dst, const void *h_src, size_t bytes, cudaStream_t st) {
    PinnedRing &r = g_ring_up[device & 63];
    std::lock_guard<std::mutex> lock(r.mu);
    r.init();
    for (size_t off = 0; off < bytes; off += PinnedRing::kSlotBytes) {
        const size_t n = std::min(PinnedRing::kSlotBytes, bytes - off);
        const int i = r.next;
        r.next = (i + 1) % PinnedRing::kSlots;
        if (r.busy[i]) KB_CUDA(cudaEventSynchronize(r.ev[i]));
        parallel_memcpy(r.slot[i], (const char *)h_src + off, n);
        KB_CUDA(cudaMemcpyAsync((char *)d_dst + off, r.slot[i], n, cudaMemcpyHostToDevice, st));
        KB_CUDA(cudaEventRecord(r.ev[i], st));
        r.busy[i] = true;
    }
}

// device -> pageable host through the ring; synchronous (the data is in h_dst on return)
static void ring_d2h(int device, void *h_dst, const void *d_src, size_t bytes, cudaStream_t st) {
    PinnedRing &r = g_ring_down[device & 63];
    std::lock_guard<std::mutex> lock(r.mu);
    r.init();
    const size_t S = PinnedRing::kSlotBytes;
    const size_t nchunks = (bytes + S - 1) / S;
    auto issue = [&](size_t c) {
        const int i = (int)(c % PinnedRing::kSlots);
        KB_CUDA(cudaMemcpyAsync(r.slot[i], (const char *)d_src + c * S, std::min(S, bytes - c * S), cudaMemcpyDeviceToHost, st));
        KB_CUDA(cudaEventRecord(r.ev[i], st));
    };
    const size_t depth = PinnedRing::kSlots - 1;
    for (size_t c = 0; c < std::min(depth, nchunks); ++c) issue(c);
    for (size_t c = 0; c < nchunks; ++c) {
        const int i = (int)(c % PinnedRing::kSlots);
        KB_CUDA(cudaEventSynchronize(r.ev[i]));
        parallel_memcpy((char *)h_dst + c * S, r.slot[i], std::min(S, bytes - c * S), 4);
        if (c + depth < nchunks) issue(c + depth);
    }
}

// The host copy: ordinary memory, first touched by all host threads.  It is page-locked (cudaHostRegister) only when
// a search has to upload it (pin_host_array): registering costs about 0.25 s per GB and most stacks never need it,
// because the constructor leaves the stack on the device.
static void alloc_host_array(kb_handle h, size_t bytes) {
    void *p = nullptr;
    if (posix_memalign(&p, 4096, std::max<size_t>(bytes, 4096)) != 0 || !p)
        throw Error("Unable to allocate space for CPU PsiPhi array.");
    char *c = static_cast<char *>(p);
    const bool band = h->row_lo != 0 || h->row_hi != h->H;
    if (!band) {
        const long long npages = (long long)((bytes + 4095) / 4096);
        const int nthreads = std::max(1, omp_get_max_threads() / 4);   // runs beside the caller's first search
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (long long i = 0; i < npages; ++i) c[(size_t)i * 4096] = 0;  // first touch
    } else {
        // band-limited handle: only its rows are ever written; the other pages stay untouched (and uncommitted)
        const size_t row_bytes = (size_t)h->W * 2 * h->nb, img = row_bytes * h->H;
        for (int t = 0; t < h->T; ++t) {
            char *b0 = c + (size_t)t * img + (size_t)h->row_lo * row_bytes;
            const size_t nb = (size_t)(h->row_hi - h->row_lo) * row_bytes;
            const long long npages = (long long)((nb + 4095) / 4096);
#pragma omp parallel for schedule(static)
            for (long long i = 0; i < npages; ++i) b0[std::min((size_t)i * 4096, nb - 1)] = 0;
        }
    }
    h->host_array = p;
    h->host_pinned = false;
}

static void pin_host_array(kb_handle h) {
    if (h->host_pinned || !h->host_array) return;
    const char *nopin = getenv("KB_NO_PIN");
    if (nopin && nopin[0] == '1') return;
    if (cudaHostRegister(h->host_array, h->meta.total_array_size, cudaHostRegisterDefault) == cudaSuccess) h->host_pinned = true;
    else cudaGetLastError();
}

// Everything that reads (or frees) the host copy waits for the background thread that fills it.
static void ensure_host(kb_handle h) {
    if (h->host_thread.joinable()) h->host_thread.join();
    if (!h->host_error.empty()) {
        const std::string e = h->host_error;
        h->host_error.clear();
        throw Error("building the host copy of the psi/phi stack failed: " + e);
    }
}

// PsiPhiArray::set_meta_data (psi_phi_array.cpp:113-148)
static void set_meta(kb_handle h, int num_bytes, int T, int H, int W) {
    if (num_bytes != -1 && num_bytes != 1 && num_bytes != 2 && num_bytes != 4)
        throw Error("Invalid setting of num_bytes. Must be (-1 [use default], 1, 2, or 4). Got " +
                    std::to_string(num_bytes));
    if (T <= 0) throw Error("Invalid num_times passed to set_meta_data: " + std::to_string(T));
    if (W <= 0) throw Error("Invalid width passed to set_meta_data: " + std::to_string(W));
    if (H <= 0) throw Error("Invalid height passed to set_meta_data: " + std::to_string(H));
    h->nb = (num_bytes == 1 || num_bytes == 2) ? num_bytes : 4;
    h->T = T;
    h->H = H;
    h->W = W;
    h->Wp = (W + 15) / 16 * 16;
    kb_meta &m = h->meta;
    m.num_bytes = h->nb;
    m.block_size = h->nb;
    m.num_times = T;
    m.width = W;
    m.height = H;
    m.pixels_per_image = (uint64_t)W * H;
    m.num_entries = 2 * m.pixels_per_image * T;
    m.total_array_size = m.block_size * m.num_entries;
    m.psi_min_val = m.phi_min_val = 3.402823466e+38F;
    m.psi_max_val = m.phi_max_val = -3.402823466e+38F;
    m.psi_scale = m.phi_scale = 1.0f;
}

// compute_scale_params_from_image_vect (psi_phi_array.cpp:219-245), given min/max from the device reduction.
static void scale_from_minmax(float mn, float mx, int nb, float out[3]) {
    float scale = 1.0f;
    if (nb == 1 || nb == 2) {
        float width = mx - mn;
        if (width < 1e-6) width = 1e-6;
        const uint64_t num_values = (1u << (8 * nb)) - 1;
        scale = (float)((double)width / (double)num_values);
    }
    out[0] = mn;
    out[1] = mx;
    out[2] = scale;
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void set_screen_steps(kb_handle h, float amax_psi, float amax_phi);

// Given the float2 psi/phi stack on the device (reference order, unpitched; rows [row_lo, row_hi) of every image are
// valid), finish the handle (fill_psi_phi_array, psi_phi_array.cpp:321-372): scale parameters and codes for encoded
// stacks, the canonical search copy and the screening copy -- both stay on the device -- and, in the background, the
// host copy in the reference layout.  Takes ownership of d_pairs.
static void finish_stack(kb_handle h, float2 *d_pairs, const double *times) {
    const uint64_t npix = (uint64_t)h->T * h->H * h->W;
    const bool band = h->row_lo != 0 || h->row_hi != h->H;
    const int r0 = h->row_lo, nr = h->row_hi - h->row_lo;
    const size_t pb = 2 * (size_t)h->nb, rows = (size_t)h->T * h->H;
    cudaStream_t st = h->stream;
    h->times.assign(times, times + h->T);
    KB_CUDA(cudaMalloc(&h->d_search, rows * h->Wp * pb));
    find_runs(h);
    KB_CUDA(cudaMalloc(&h->d_screen, screen_bytes(h)));
    upload_runs(h, st);
    KB_CUDA(cudaMalloc(&h->d_times, sizeof(double) * h->T));
    KB_CUDA(cudaMemcpyAsync(h->d_times, h->times.data(), sizeof(double) * h->T, cudaMemcpyHostToDevice, st));
    const char *nozm = getenv("KB_NO_ZERO_MARKER");
    DevBuf<uint32_t> d_scan;
    d_scan.ensure(6);
    uint32_t scan[6];
    launch_scan_raw_rows(d_pairs, h->T, h->H, r0, nr, h->W, d_scan.p, st);
    KB_CUDA(cudaMemcpyAsync(scan, d_scan.p, sizeof(scan), cudaMemcpyDeviceToHost, st));
    KB_CUDA(cudaStreamSynchronize(st));
    // min / max over the finite values of each channel; none at all: the reference keeps FLT_MAX / -FLT_MAX (:223-232)
    float pmn = 3.402823466e+38F, pmx = -3.402823466e+38F, fmn = pmn, fmx = pmx;
    if (scan[1] <= scan[2]) { pmn = ordered_to_float(scan[1]); pmx = ordered_to_float(scan[2]); }
    if (scan[3] <= scan[4]) { fmn = ordered_to_float(scan[3]); fmx = ordered_to_float(scan[4]); }
    h->all_valid = scan[5] == 0;   // (an encoded stack has the same invalid pairs: only non-finite values map to code 0)
    void *d_ref = d_pairs;   // reference-layout bytes for the host copy
    if (h->nb == 4) {
        h->zm_known = (scan[0] || (nozm && nozm[0] == '1')) ? 0 : 1;
        set_screen_steps(h, pmn <= pmx ? std::max(fabsf(pmn), fabsf(pmx)) : 0.0f,
                         fmn <= fmx ? std::max(fabsf(fmn), fabsf(fmx)) : 0.0f);
    } else {
        if (band) throw Error("band-limited construction supports float32 stacks only (the scale of an encoded "
                              "stack depends on every row)");
        float pp[3], fp[3];
        scale_from_minmax(pmn, pmx, h->nb, pp);
        scale_from_minmax(fmn, fmx, h->nb, fp);
        // PsiPhiArray::set_psi_scaling / set_phi_scaling checks (psi_phi_array.cpp:150-168)
        if (pp[0] > pp[1]) throw Error("Min value needs to be < max value. Got " + std::to_string(pp[0]) + " and " + std::to_string(pp[1]));
        if (fp[0] > fp[1]) throw Error("Min value needs to be < max value. Got " + std::to_string(fp[0]) + " and " + std::to_string(fp[1]));
        h->meta.psi_min_val = pp[0]; h->meta.psi_max_val = pp[1]; h->meta.psi_scale = pp[2];
        h->meta.phi_min_val = fp[0]; h->meta.phi_max_val = fp[1]; h->meta.phi_scale = fp[2];
        void *d_enc = nullptr;
        KB_CUDA(cudaMalloc(&d_enc, h->meta.total_array_size));
        launch_encode(d_pairs, npix, h->nb, pp, fp, d_enc, st);
        d_ref = d_enc;
        h->zm_known = 0;
        set_screen_steps(h, std::max(fabsf(pp[0]), fabsf(pp[1])), std::max(fabsf(fp[0]), fabsf(fp[1])));
    }
    h->zero_marker = h->zm_known;
    launch_canonicalise_rows(d_ref, h->W, h->d_search, h->T, h->H, h->W, h->Wp, h->nb, h->zero_marker, r0, nr, st);
    build_screen_rows(h, r0, nr, st);
    KB_CUDA(cudaStreamSynchronize(st));
    if (d_ref != d_pairs) cudaFree(d_pairs);
    h->d_raw = d_ref;
    const char *nocache = getenv("KB_NO_DEVICE_CACHE");
    h->dev_cache = !(nocache && nocache[0] == '1');
    {   // the reference-layout bytes stay on the device too while HBM is plentiful (180 GB): get_all_psi_phi_curves
        // gathers from them (psi and phi validity are independent there, unlike in the canonical search copy)
        size_t free_b = 0, total_b = 0;
        const char *drop = getenv("KB_DROP_RAW");
        h->keep_raw = h->dev_cache && !(drop && drop[0] == '1') && cudaMemGetInfo(&free_b, &total_b) == cudaSuccess &&
                      free_b > 4 * (size_t)h->meta.total_array_size;
    }
    // ---- the host copy, off the caller's critical path ----
    KB_CUDA(cudaStreamCreateWithFlags(&h->host_stream, cudaStreamNonBlocking));
    const double t0 = now_s();
    h->host_thread = std::thread([h, t0, r0, nr, pb]() {
        try {
            KB_CUDA(cudaSetDevice(h->device));
            alloc_host_array(h, h->meta.total_array_size);
            const size_t img = (size_t)h->H * h->W * pb;
            if (nr == h->H) {
                ring_d2h(h->device, h->host_array, h->d_raw, (size_t)h->T * img, h->host_stream);
            } else {
                for (int t = 0; t < h->T; ++t) {
                    const size_t off = (size_t)t * img + (size_t)r0 * h->W * pb;
                    ring_d2h(h->device, (char *)h->host_array + off, (const char *)h->d_raw + off, (size_t)nr * h->W * pb, h->host_stream);
                }
            }
            if (getenv("KB_DEBUG"))
                fprintf(stderr, "[kbmod_b200] host copy (%.2f GB) filled in the background in %.3f s\n",
                        h->meta.total_array_size / 1e9, now_s() - t0);
        } catch (const std::exception &e) {
            h->host_error = e.what();
        }
        h->host_done.store(true);
    });
    if (!h->dev_cache) {   // no cache wanted: behave like the reference (host copy only)
        ensure_host(h);
        cudaFree(h->d_raw);
        h->d_raw = nullptr;
        cudaFree(h->d_search);
        cudaFree(h->d_screen);
        cudaFree(h->d_times);
        h->d_search = nullptr;
        h->d_screen = nullptr;
        h->d_times = nullptr;
    }
}

// The device-side source of the host copy is released as soon as the background copy is done and somebody looks.
static void release_raw_if_done(kb_handle h, bool wait, bool force = false) {
    if (wait) ensure_host(h);
    if (!h->d_raw || (h->keep_raw && !force)) return;
    if (!h->host_done.load()) return;   // still copying
    ensure_host(h);
    cudaFree(h->d_raw);
    h->d_raw = nullptr;
}

static kb_handle new_handle() {
    std::unique_ptr<kb_stack_s> h(new kb_stack_s());
    int cnt = 0;
    if (cudaGetDeviceCount(&cnt) != cudaSuccess || cnt == 0) {
        cudaGetLastError();
        throw Error("GPU is not available (kbmod_b200 has no CPU fallback).");
    }
    KB_CUDA(cudaGetDevice(&h->device));
    KB_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    return h.release();
}

// keep_alloc: search_all on a stack that is not preloaded uploads it for every search and drops it afterwards
// (the reference's semantics, stack_search.cpp:243,257).  The device allocation itself is kept for the next
// upload -- cudaMalloc/cudaFree of a multi-GB block every search costs milliseconds and synchronises the device.
static void free_device_stack(kb_handle h, bool keep_alloc = false) {
    if (!keep_alloc) {
        if (h->d_search) cudaFree(h->d_search);
        if (h->d_screen) cudaFree(h->d_screen);
        if (h->d_times) cudaFree(h->d_times);
        h->d_search = nullptr;
        h->d_screen = nullptr;
        h->d_times = nullptr;
        h->dev_cache = false;
    }
    h->on_gpu = false;
}

static void destroy_handle(kb_handle h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->host_thread.joinable()) h->host_thread.join();
    if (h->plan_cache) tiled_plan_destroy(h->plan_cache);
    if (h->d_runs) cudaFree(h->d_runs);
    if (h->d_raw) cudaFree(h->d_raw);
    if (h->host_stream) cudaStreamDestroy(h->host_stream);
    free_device_stack(h);
    if (h->host_array) {
        if (h->host_pinned) cudaHostUnregister(h->host_array);
        free(h->host_array);
    }
    result_scratch_free(h->scratch);
    clip_scratch_free(h->clip);
    if (h->ev_start) cudaEventDestroy(h->ev_start);
    if (h->ev_stop) cudaEventDestroy(h->ev_stop);
    for (int i = 0; i < UploadProgress::kMaxChunks; ++i)
        if (h->up.ev[i]) cudaEventDestroy(h->up.ev[i]);
    for (int i = 0; i < 2; ++i) {
        if (h->up.band_done[i]) cudaEventDestroy(h->up.band_done[i]);
        if (h->up.band_stream[i]) cudaStreamDestroy(h->up.band_stream[i]);
    }
    if (h->up.ready) cudaEventDestroy(h->up.ready);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

// Quantisation steps of the screening copy from the largest |psi| and |phi| of the stack: 32767 steps reach it.
static void set_screen_steps(kb_handle h, float amax_psi, float amax_phi) {
    auto step = [](float a) {
        if (!(a > 0.0f) || !std::isfinite(a)) return 1.0f;
        const float s = a / 32767.0f;
        return (s > 0.0f && std::isfinite(1.0f / s)) ? s : 1.0f;   // denormal range: screening is useless but safe
    };
    h->q_step_psi = step(amax_psi);
    h->q_step_phi = step(amax_phi);
    h->q_known = true;
}

// PsiPhiArray::move_to_gpu (psi_phi_array.cpp:74-111): upload + canonicalise into the pitched search copy.
static void upload_stack(kb_handle h) {
    if (h->on_gpu) return;
    if (h->dev_cache) {   // the copy the constructor built is still there: nothing to move
        h->zero_marker = h->zm_known;
        h->on_gpu = true;
        return;
    }
    ensure_host(h);
    pin_host_array(h);
    const size_t pb = pair_bytes(h->nb);
    const size_t rows = (size_t)h->T * h->H;
    const int r0 = h->row_lo, nr = h->row_hi - h->row_lo;   // the rows this handle holds (all of them as a rule)
    if (!h->d_search) KB_CUDA(cudaMalloc(&h->d_search, rows * h->Wp * pb));
    if (!h->d_screen) KB_CUDA(cudaMalloc(&h->d_screen, screen_bytes(h)));
    if (!h->d_times) KB_CUDA(cudaMalloc(&h->d_times, sizeof(double) * h->T));
    KB_CUDA(cudaMemcpyAsync(h->d_times, h->times.data(), sizeof(double) * h->T, cudaMemcpyHostToDevice, h->stream));
    upload_runs(h, h->stream);
    {
        cudaMemcpy3DParms q;
        memset(&q, 0, sizeof(q));
        q.srcPtr = make_cudaPitchedPtr(h->host_array, (size_t)h->W * pb, (size_t)h->W * pb, (size_t)h->H);
        q.dstPtr = make_cudaPitchedPtr(h->d_search, (size_t)h->Wp * pb, (size_t)h->Wp * pb, (size_t)h->H);
        q.srcPos = make_cudaPos(0, (size_t)r0, 0);
        q.dstPos = make_cudaPos(0, (size_t)r0, 0);
        q.extent = make_cudaExtent((size_t)h->W * pb, (size_t)nr, (size_t)h->T);
        q.kind = cudaMemcpyHostToDevice;
        KB_CUDA(cudaMemcpy3DAsync(&q, h->stream));
    }
    h->zero_marker = 0;
    const char *nozm = getenv("KB_NO_ZERO_MARKER");
    if (h->nb == 4 && h->zm_known < 0) {
        // One look at a float stack the first time it comes up (the host copy never changes afterwards): may
        // invalid pairs be stored as (0, 0)?
        DevBuf<uint32_t> d_flag;
        d_flag.ensure(1);
        launch_detect_valid_zero_pairs(h->d_search, h->Wp, h->T, h->H, h->W, d_flag.p, h->stream);
        uint32_t flag = 1;
        KB_CUDA(cudaMemcpyAsync(&flag, d_flag.p, sizeof(flag), cudaMemcpyDeviceToHost, h->stream));
        KB_CUDA(cudaStreamSynchronize(h->stream));
        h->zm_known = (flag || (nozm && nozm[0] == '1')) ? 0 : 1;
    }
    if (h->nb != 4) h->zm_known = 0;
    h->zero_marker = h->zm_known;
    launch_canonicalise_rows(h->d_search, h->Wp, h->d_search, h->T, h->H, h->W, h->Wp, h->nb, h->zero_marker, r0, nr,
                             h->stream);
    if (!h->q_known) {
        // ... and how large do the values get: the quantisation steps of the screening copy
        if (h->nb == 4) {
            DevBuf<uint32_t> d_mm;
            d_mm.ensure(4);
            launch_minmax_pitched(h->d_search, rows * h->Wp, d_mm.p, h->stream);  // canonical: padding is invalid
            uint32_t out[4];
            KB_CUDA(cudaMemcpyAsync(out, d_mm.p, sizeof(out), cudaMemcpyDeviceToHost, h->stream));
            KB_CUDA(cudaStreamSynchronize(h->stream));
            float amax[2] = {0.0f, 0.0f};
            for (int k = 0; k < 2; ++k)
                if (out[2 * k] <= out[2 * k + 1])
                    amax[k] = std::max(fabsf(ordered_to_float(out[2 * k])), fabsf(ordered_to_float(out[2 * k + 1])));
            set_screen_steps(h, amax[0], amax[1]);
        } else {
            set_screen_steps(h, std::max(fabsf(h->meta.psi_min_val), fabsf(h->meta.psi_max_val)),
                             std::max(fabsf(h->meta.phi_min_val), fabsf(h->meta.phi_max_val)));
        }
    }
    build_screen_rows(h, r0, nr, h->stream);
    h->on_gpu = true;
}

// The same upload in row chunks on a second stream, so a search can start on the rows that have arrived
// (launch_search_tiled with an UploadProgress).  Needs the zero-marker decision of an earlier upload: the host copy
// never changes after construction, so the decision holds.  Returns false when it cannot be used.
static bool upload_stack_pipelined(kb_handle h, int row_lo, int row_hi, uint64_t *bytes_copied) {
    if (getenv("KB_DEBUG"))
        fprintf(stderr, "[kbmod_b200] pipelined upload? on_gpu=%d zm_known=%d pinned=%d\n", (int)h->on_gpu, h->zm_known,
                (int)h->host_pinned);
    if (h->on_gpu || h->dev_cache || h->zm_known < 0 || !h->q_known) return false;
    ensure_host(h);
    pin_host_array(h);
    if (!h->host_pinned) return false;
    const char *off = getenv("KB_NO_PIPELINED_UPLOAD");
    if (off && off[0] == '1') return false;
    const size_t pb = pair_bytes(h->nb);
    const size_t rows = (size_t)h->T * h->H;
    const size_t bytes = rows * h->W * pb;
    size_t min_bytes = 64u << 20;  // small stacks: one copy is as fast
    if (const char *mb = getenv("KB_PIPELINED_UPLOAD_MIN_BYTES")) min_bytes = (size_t)strtoull(mb, nullptr, 10);
    if (getenv("KB_DEBUG"))
        fprintf(stderr, "[kbmod_b200] pipelined upload: %zu bytes (min %zu), H=%d\n", bytes, min_bytes, h->H);
    if (bytes < min_bytes || h->H < 2 * UploadProgress::kMaxChunks) return false;
    // Only the rows this search can touch (its start rows plus the candidates' reach) are brought up: the copy is
    // dropped again after the search, so the rest of the stack is never read.  Too thin a range: take everything.
    row_lo = std::max(h->row_lo, std::min(row_lo, h->row_hi));
    row_hi = std::max(h->row_lo, std::min(row_hi, h->row_hi));
    if (row_hi - row_lo < 2 * UploadProgress::kMaxChunks) {   // too thin a range: everything the handle holds
        row_lo = h->row_lo;
        row_hi = h->row_hi;
    }
    if (row_hi - row_lo < 2 * UploadProgress::kMaxChunks) return false;
    const int64_t span = row_hi - row_lo;
    if (!h->d_search) KB_CUDA(cudaMalloc(&h->d_search, rows * h->Wp * pb));
    if (!h->d_screen) KB_CUDA(cudaMalloc(&h->d_screen, screen_bytes(h)));
    if (!h->d_times) KB_CUDA(cudaMalloc(&h->d_times, sizeof(double) * h->T));
    UploadProgress &up = h->up;
    if (!h->copy_stream) {
        KB_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < UploadProgress::kMaxChunks; ++i) KB_CUDA(cudaEventCreateWithFlags(&up.ev[i], cudaEventDisableTiming));
        for (int i = 0; i < 2; ++i) {
            KB_CUDA(cudaStreamCreateWithFlags(&up.band_stream[i], cudaStreamNonBlocking));
            KB_CUDA(cudaEventCreateWithFlags(&up.band_done[i], cudaEventDisableTiming));
        }
        KB_CUDA(cudaEventCreateWithFlags(&up.ready, cudaEventDisableTiming));
    }
    // the copy stream starts after whatever the handle's stream did to the old copy
    KB_CUDA(cudaEventRecord(up.ready, h->stream));
    KB_CUDA(cudaStreamWaitEvent(h->copy_stream, up.ready, 0));
    KB_CUDA(cudaMemcpyAsync(h->d_times, h->times.data(), sizeof(double) * h->T, cudaMemcpyHostToDevice, h->copy_stream));
    upload_runs(h, h->copy_stream);
    if (const char *poison = getenv("KB_POISON_STACK"))  // tests: rows that are not uploaded must never be read
        if (poison[0] == '1') {
            KB_CUDA(cudaMemsetAsync(h->d_search, 0xA5, rows * h->Wp * pb, h->copy_stream));
            KB_CUDA(cudaMemsetAsync(h->d_screen, 0xA5, screen_bytes(h), h->copy_stream));
        }
    h->zero_marker = (h->nb == 4) ? h->zm_known : 0;
    up.n_chunks = UploadProgress::kMaxChunks;
    for (int c = 0; c < up.n_chunks; ++c) {
        const int r0 = row_lo + (int)(span * c / up.n_chunks), r1 = row_lo + (int)(span * (c + 1) / up.n_chunks);
        cudaMemcpy3DParms q;
        memset(&q, 0, sizeof(q));
        q.srcPtr = make_cudaPitchedPtr(h->host_array, (size_t)h->W * pb, (size_t)h->W * pb, (size_t)h->H);
        q.dstPtr = make_cudaPitchedPtr(h->d_search, (size_t)h->Wp * pb, (size_t)h->Wp * pb, (size_t)h->H);
        q.srcPos = make_cudaPos(0, (size_t)r0, 0);
        q.dstPos = make_cudaPos(0, (size_t)r0, 0);
        q.extent = make_cudaExtent((size_t)h->W * pb, (size_t)(r1 - r0), (size_t)h->T);
        q.kind = cudaMemcpyHostToDevice;
        KB_CUDA(cudaMemcpy3DAsync(&q, h->copy_stream));
        launch_canonicalise_rows(h->d_search, h->Wp, h->d_search, h->T, h->H, h->W, h->Wp, h->nb, h->zero_marker, r0,
                                 r1 - r0, h->copy_stream);
        build_screen_rows(h, r0, r1 - r0, h->copy_stream);
        up.row_end[c] = r1;
        KB_CUDA(cudaEventRecord(up.ev[c], h->copy_stream));
    }
    h->on_gpu = true;
    if (bytes_copied) *bytes_copied = (uint64_t)span * h->T * h->W * pb;
    return true;
}

static void check_params(kb_handle h, const kb_params *p) {
    if (p->x_start_min >= p->x_start_max)
        throw Error("Invalid search bounds for the x pixel [" + std::to_string(p->x_start_min) + ", " +
                    std::to_string(p->x_start_max) + "]");
    if (p->y_start_min >= p->y_start_max)
        throw Error("Invalid search bounds for the y pixel [" + std::to_string(p->y_start_min) + ", " +
                    std::to_string(p->y_start_max) + "]");
    if (p->results_per_pixel == 0) throw Error("Invalid results per pixel. Got 0");
    if (p->min_observations < 0) throw Error("min_obs must be >= 0. Got " + std::to_string(p->min_observations));
    if (p->do_sigmag_filter) {
        if (!(p->sgl_L < p->sgl_H) || p->sgl_L <= 0.0f || p->sgl_H >= 1.0f)
            throw Error("Invalid percentiles for sigma G filtering. Got [" + std::to_string(p->sgl_L) + ", " +
                        std::to_string(p->sgl_H) + "].");
        if (p->sigmag_coeff <= 0.0f)
            throw Error("Invalid coefficient for sigma G filtering. Got " + std::to_string(p->sigmag_coeff) + ".");
    }
    const int64_t lim = 1 << 22;  // index arithmetic is exact well inside this range
    if (p->x_start_min < -lim || p->x_start_max > lim || p->y_start_min < -lim || p->y_start_max > lim)
        throw Error("Search bounds are out of the supported range (+-4194304).");
    (void)h;
}

extern "C" {

const char *kb_last_error(void) { return g_last_error.c_str(); }
const char *kb_version(void) { return "kbmod_b200 0.2 (sm_100a)"; }

int kb_device_count(int *count) {
    KB_API_BEGIN
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
        cudaGetLastError();
        c = 0;
    }
    *count = c;
    KB_API_END
}

int kb_gpu_memory(size_t *free_bytes, size_t *total_bytes) {
    KB_API_BEGIN
    KB_CUDA(cudaMemGetInfo(free_bytes, total_bytes));
    KB_API_END
}

int kb_set_device(int device) {
    KB_API_BEGIN
    KB_CUDA(cudaSetDevice(device));
    KB_API_END
}

int kb_device_pci_bus_id(char *buf, int len) {
    KB_API_BEGIN
    if (!buf || len < 16) throw Error("kb_device_pci_bus_id: need a buffer of at least 16 bytes");
    int dev = 0;
    KB_CUDA(cudaGetDevice(&dev));
    KB_CUDA(cudaDeviceGetPCIBusId(buf, len, dev));
    KB_API_END
}

static float host_sum(const float *k, int n) {
    float s = 0.0f;
    for (int i = 0; i < n; ++i) s += k[i];  // image_utils_cpp.cpp:30-35 row-major float adds
    return s;
}

int kb_convolve(const float *img, int H, int W, const float *psf, int psf_dim, float *out) {
    KB_API_BEGIN
    if (H <= 0 || W <= 0 || psf_dim <= 0) throw Error("Invalid image or PSF dimensions.");
    const size_t n = (size_t)H * W;
    DevBuf<float> d_in, d_out, d_psf;
    d_in.ensure(n);
    d_out.ensure(n);
    d_psf.ensure((size_t)psf_dim * psf_dim);
    KB_CUDA(cudaMemcpy(d_in.p, img, n * sizeof(float), cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_psf.p, psf, sizeof(float) * psf_dim * psf_dim, cudaMemcpyHostToDevice));
    launch_convolve(d_in.p, H, W, d_psf.p, psf_dim, host_sum(psf, psf_dim * psf_dim), d_out.p, 0);
    KB_CUDA(cudaMemcpy(out, d_out.p, n * sizeof(float), cudaMemcpyDeviceToHost));
    KB_API_END
}

int kb_generate_psi_phi(const float *sci, const float *var, int H, int W, const float *psf, int psf_dim,
                        float *psi_out, float *phi_out) {
    KB_API_BEGIN
    if (H <= 0 || W <= 0 || psf_dim <= 0) throw Error("Invalid image or PSF dimensions.");
    const size_t n = (size_t)H * W, kn = (size_t)psf_dim * psf_dim;
    std::vector<float> k2(kn);
    for (size_t i = 0; i < kn; ++i) k2[i] = psf[i] * psf[i];  // square_psf_values (image_utils_cpp.cpp:110-120)
    DevBuf<float> d_sci, d_var, d_psi, d_phi, d_k, d_k2;
    d_sci.ensure(n); d_var.ensure(n); d_psi.ensure(n); d_phi.ensure(n); d_k.ensure(kn); d_k2.ensure(kn);
    // generate_phi takes no science image: feed zeros so the psi half is simply unused.
    if (sci) KB_CUDA(cudaMemcpy(d_sci.p, sci, n * sizeof(float), cudaMemcpyHostToDevice));
    else KB_CUDA(cudaMemset(d_sci.p, 0, n * sizeof(float)));
    KB_CUDA(cudaMemcpy(d_var.p, var, n * sizeof(float), cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_k.p, psf, kn * sizeof(float), cudaMemcpyHostToDevice));
    KB_CUDA(cudaMemcpy(d_k2.p, k2.data(), kn * sizeof(float), cudaMemcpyHostToDevice));
    launch_psi_phi(d_sci.p, d_var.p, H, W, d_k.p, d_k2.p, psf_dim, host_sum(psf, (int)kn), host_sum(k2.data(), (int)kn),
                   nullptr, d_psi.p, d_phi.p, 0);
    if (psi_out) KB_CUDA(cudaMemcpy(psi_out, d_psi.p, n * sizeof(float), cudaMemcpyDeviceToHost));
    if (phi_out) KB_CUDA(cudaMemcpy(phi_out, d_phi.p, n * sizeof(float), cudaMemcpyDeviceToHost));
    KB_API_END
}

static void create_from_images(const float *const *sci, const float *const *var, const float *const *psf,
                               const int *psf_dim, int T, int H, int W, const double *zeroed_times, int num_bytes,
                               int row_lo, int row_hi, kb_handle *out) {
    if (T <= 0) throw Error("No images in the to process.");
    kb_handle h = new_handle();
    float2 *d_pairs = nullptr;
    cudaStream_t copy_st = nullptr;
    cudaEvent_t copied[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr};
    auto cleanup = [&]() {
        for (int i = 0; i < 2; ++i) {
            if (copied[i]) cudaEventDestroy(copied[i]);
            if (consumed[i]) cudaEventDestroy(consumed[i]);
        }
        if (copy_st) cudaStreamDestroy(copy_st);
    };
    try {
        const double t_begin = now_s();
        set_meta(h, num_bytes, T, H, W);
        if (row_lo < 0 || row_hi > H || row_lo >= row_hi) throw Error("Invalid row range for the psi/phi stack.");
        h->row_lo = row_lo;
        h->row_hi = row_hi;
        const size_t n = (size_t)H * W;
        KB_CUDA(cudaMalloc(&d_pairs, sizeof(float2) * n * T));
        // every PSF and its square (square_psf_values, image_utils_cpp.cpp:110-120) in one small upload
        std::vector<size_t> koff(T + 1, 0);
        int rmax = 0;
        for (int t = 0; t < T; ++t) {
            const int dim = psf_dim[t];
            if (dim <= 0 || (dim & 1) == 0) throw Error("PSF kernels must be square with an odd dimension.");
            koff[t + 1] = koff[t] + (size_t)dim * dim;
            rmax = std::max(rmax, (dim - 1) / 2);
        }
        std::vector<float> kall(2 * koff[T]);
        std::vector<float> ksum(2 * (size_t)T);
        for (int t = 0; t < T; ++t) {
            const size_t kn = koff[t + 1] - koff[t];
            float *k1 = kall.data() + 2 * koff[t], *k2 = k1 + kn;
            for (size_t i = 0; i < kn; ++i) {
                k1[i] = psf[t][i];
                k2[i] = psf[t][i] * psf[t][i];
            }
            ksum[2 * t] = host_sum(k1, (int)kn);
            ksum[2 * t + 1] = host_sum(k2, (int)kn);
        }
        DevBuf<float> d_k;
        d_k.ensure(std::max<size_t>(kall.size(), 1));
        KB_CUDA(cudaMemcpyAsync(d_k.p, kall.data(), kall.size() * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        // Input rows of image t that rows [row_lo, row_hi) of its psi/phi depend on: the band plus the PSF radius.
        // Double-buffered staging: the upload of image t+1 (through the pinned ring) overlaps the convolution of t.
        const size_t max_rows = (size_t)std::min(H, row_hi - row_lo + 2 * rmax);
        DevBuf<float> d_sci[2], d_var[2];
        KB_CUDA(cudaStreamCreateWithFlags(&copy_st, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            d_sci[i].ensure(max_rows * W);
            d_var[i].ensure(max_rows * W);
            KB_CUDA(cudaEventCreateWithFlags(&copied[i], cudaEventDisableTiming));
            KB_CUDA(cudaEventCreateWithFlags(&consumed[i], cudaEventDisableTiming));
        }
        uint64_t h2d_bytes = 0;
        for (int t = 0; t < T; ++t) {
            const int b = t & 1;
            const int dim = psf_dim[t], rad = (dim - 1) / 2;
            const int in_lo = std::max(0, row_lo - rad), in_hi = std::min(H, row_hi + rad);
            const size_t nin = (size_t)(in_hi - in_lo) * W;
            if (t >= 2) KB_CUDA(cudaStreamWaitEvent(copy_st, consumed[b], 0));
            ring_h2d(h->device, d_sci[b].p, sci[t] + (size_t)in_lo * W, nin * sizeof(float), copy_st);
            ring_h2d(h->device, d_var[b].p, var[t] + (size_t)in_lo * W, nin * sizeof(float), copy_st);
            h2d_bytes += 2 * nin * sizeof(float);
            KB_CUDA(cudaEventRecord(copied[b], copy_st));
            KB_CUDA(cudaStreamWaitEvent(h->stream, copied[b], 0));
            const float *k1 = d_k.p + 2 * koff[t], *k2 = k1 + (koff[t + 1] - koff[t]);
            // the uploaded rows are convolved as an image of their own; its first row is row in_lo of the frame
            launch_psi_phi(d_sci[b].p, d_var[b].p, in_hi - in_lo, W, k1, k2, dim, ksum[2 * t], ksum[2 * t + 1],
                           d_pairs + (size_t)t * n + (size_t)in_lo * W, nullptr, nullptr, h->stream, row_lo - in_lo,
                           row_hi - in_lo);
            KB_CUDA(cudaEventRecord(consumed[b], h->stream));
        }
        KB_CUDA(cudaStreamSynchronize(h->stream));   // kall / ksum and the staging buffers die with this scope
        const double t_gen = now_s();
        if (getenv("KB_DEBUG"))
            fprintf(stderr, "[kbmod_b200] H2D (%.2f GB through the pinned ring) + psi/phi kernels: %.3f s\n", h2d_bytes / 1e9,
                    t_gen - t_begin);
        float2 *pairs = d_pairs;
        d_pairs = nullptr;   // finish_stack owns it from here
        finish_stack(h, pairs, zeroed_times);
        if (getenv("KB_DEBUG")) fprintf(stderr, "[kbmod_b200] scale / encode / search + screening copies: %.3f s\n", now_s() - t_gen);
    } catch (...) {
        cleanup();
        if (d_pairs) cudaFree(d_pairs);
        destroy_handle(h);
        throw;
    }
    cleanup();
    *out = h;
}

int kb_create_from_images(const float *const *sci, const float *const *var, const float *const *psf,
                          const int *psf_dim, int T, int H, int W, const double *zeroed_times, int num_bytes,
                          kb_handle *out) {
    KB_API_BEGIN
    create_from_images(sci, var, psf, psf_dim, T, H, W, zeroed_times, num_bytes, 0, H, out);
    KB_API_END
}

int kb_create_from_images_rows(const float *const *sci, const float *const *var, const float *const *psf,
                               const int *psf_dim, int T, int H, int W, const double *zeroed_times, int num_bytes,
                               int row_lo, int row_hi, kb_handle *out) {
    KB_API_BEGIN
    create_from_images(sci, var, psf, psf_dim, T, H, W, zeroed_times, num_bytes, row_lo, row_hi, out);
    KB_API_END
}

int kb_create_from_psi_phi(const float *const *psi, const float *const *phi, int T, int H, int W,
                           const double *zeroed_times, int num_bytes, kb_handle *out) {
    KB_API_BEGIN
    if (T <= 0) throw Error("Trying to fill PsiPhi from empty vectors.");
    kb_handle h = new_handle();
    float2 *d_pairs = nullptr;
    try {
        set_meta(h, num_bytes, T, H, W);
        h->row_lo = 0;
        h->row_hi = H;
        const size_t n = (size_t)H * W;
        KB_CUDA(cudaMalloc(&d_pairs, sizeof(float2) * n * T));
        DevBuf<float> d_a, d_b;
        d_a.ensure(n);
        d_b.ensure(n);
        for (int t = 0; t < T; ++t) {
            ring_h2d(h->device, d_a.p, psi[t], n * sizeof(float), h->stream);
            ring_h2d(h->device, d_b.p, phi[t], n * sizeof(float), h->stream);
            launch_interleave(d_a.p, d_b.p, n, d_pairs + (size_t)t * n, h->stream);
        }
        KB_CUDA(cudaStreamSynchronize(h->stream));
        float2 *pairs = d_pairs;
        d_pairs = nullptr;
        finish_stack(h, pairs, zeroed_times);
    } catch (...) {
        if (d_pairs) cudaFree(d_pairs);
        destroy_handle(h);
        throw;
    }
    *out = h;
    KB_API_END
}

int kb_destroy(kb_handle h) {
    KB_API_BEGIN
    destroy_handle(h);
    KB_API_END
}

int kb_get_meta(kb_handle h, kb_meta *meta) {
    KB_API_BEGIN
    *meta = h->meta;
    KB_API_END
}

int kb_host_array(kb_handle h, const void **ptr, uint64_t *nbytes) {
    KB_API_BEGIN
    use_device(h);
    release_raw_if_done(h, true);
    *ptr = h->host_array;
    *nbytes = h->meta.total_array_size;
    KB_API_END
}

int kb_read_psi_phi(kb_handle h, uint64_t t, int row, int col, float out[2]) {
    KB_API_BEGIN
    out[0] = out[1] = NAN;
    if (row >= 0 && col >= 0 && row < h->H && col < h->W && t < (uint64_t)h->T) {
        use_device(h);
        release_raw_if_done(h, true);
    }
    if (!h->host_array || row < 0 || col < 0 || row >= h->H || col >= h->W || t >= (uint64_t)h->T) return 0;
    const uint64_t idx = 2 * (h->meta.pixels_per_image * t + (uint64_t)row * h->W + (uint64_t)col);
    if (h->nb == 4) {
        out[0] = ((const float *)h->host_array)[idx];
        out[1] = ((const float *)h->host_array)[idx + 1];
    } else {
        const float a = (h->nb == 1) ? (float)((const uint8_t *)h->host_array)[idx] : (float)((const uint16_t *)h->host_array)[idx];
        const float b = (h->nb == 1) ? (float)((const uint8_t *)h->host_array)[idx + 1] : (float)((const uint16_t *)h->host_array)[idx + 1];
        // decode_uint_scalar (psi_phi_array_ds.h:45-47)
        out[0] = (a == 0.0f) ? NAN : (float)(((double)a - 1.0) * (double)h->meta.psi_scale + (double)h->meta.psi_min_val);
        out[1] = (b == 0.0f) ? NAN : (float)(((double)b - 1.0) * (double)h->meta.phi_scale + (double)h->meta.phi_min_val);
    }
    KB_API_END
}

int kb_read_time(kb_handle h, uint64_t t, double *out) {
    KB_API_BEGIN
    if (t >= (uint64_t)h->T) throw Error("Out of bounds read for time step. [" + std::to_string(t) + "]");
    *out = h->times[t];
    KB_API_END
}

int kb_preload(kb_handle h) {
    KB_API_BEGIN
    use_device(h);
    upload_stack(h);
    KB_CUDA(cudaStreamSynchronize(h->stream));
    KB_API_END
}

int kb_unload(kb_handle h) {
    KB_API_BEGIN
    use_device(h);
    KB_CUDA(cudaStreamSynchronize(h->stream));
    release_raw_if_done(h, true, true);   // from now on the host copy is the only one
    free_device_stack(h);
    KB_API_END
}

int kb_on_gpu(kb_handle h, int *flag) {
    KB_API_BEGIN
    *flag = h->on_gpu ? 1 : 0;
    KB_API_END
}

int kb_search_note(kb_handle h, const char **note) {
    KB_API_BEGIN
    *note = h->search_note.c_str();
    KB_API_END
}

int kb_set_kernel_mode(kb_handle h, int mode) {
    KB_API_BEGIN
    if (mode < 0 || mode > 3) throw Error("kernel mode must be 0, 1, 2 or 3");
    h->kernel_mode = mode;
    KB_API_END
}

int kb_search(kb_handle h, const kb_params *params, const float *vx, const float *vy, uint64_t n_cand,
              uint64_t *n_results) {
    KB_API_BEGIN
    use_device(h);
    check_params(h, params);
    if (params->min_observations > h->T)
        throw Error("min_obs cannot be greater than the number of images. min_obs = " +
                    std::to_string(params->min_observations) + ", num_imgs = " + std::to_string(h->T) + ".");
    for (uint64_t i = 0; i < n_cand; ++i)  // TrajectoryList::assert_valid (trajectory_list.cpp:155-164)
        if (!isfinite(vx[i]) || !isfinite(vy[i]))
            throw Error("Invalid trajectory detected at index " + std::to_string(i));
    if (n_cand >= (1ull << 31)) throw Error("Too many candidate trajectories.");
    kb_params p = *params;
    const uint64_t sw = (uint64_t)(p.x_start_max - p.x_start_min), sh = (uint64_t)(p.y_start_max - p.y_start_min);
    const uint64_t n_slots = sw * sh * p.results_per_pixel;
    // 32-bit slot and start-pixel indices downstream (compaction, sort, request queues): refuse before anything is
    // allocated or searched
    if (n_slots >= (1ull << 32)) throw Error("Too many result slots for 32-bit indices (start pixels x results_per_pixel).");
    p.total_results = n_slots;
    h->timing = kb_timing{};
    h->timing.evals = sw * sh * n_cand * (uint64_t)h->T;
    cudaStream_t st = h->stream;
    Timer t_total, t_h2d, t_prep, t_search, t_post;
    uint32_t launches = 0;

    release_raw_if_done(h, false);
    {   // band-limited stacks: every row the search can touch must be one the handle holds
        double vmax = 0.0, tabs = 0.0;
        for (uint64_t i = 0; i < n_cand; ++i) vmax = std::max(vmax, (double)fabsf(vy[i]));
        for (double t : h->times) tabs = std::max(tabs, fabs(t));
        const double R = std::min(ceil(vmax * tabs) + 1.0, 4.0e6);
        const int need_lo = std::max(0, (int)std::max(-4.0e6, (double)p.y_start_min - R));
        const int need_hi = std::min(h->H, (int)std::min(4.0e6, (double)p.y_start_max + R));
        if ((h->row_lo != 0 || h->row_hi != h->H) && need_lo < need_hi && (need_lo < h->row_lo || need_hi > h->row_hi))
            throw Error("this band-limited stack holds image rows [" + std::to_string(h->row_lo) + ", " +
                        std::to_string(h->row_hi) + ") but the search reaches rows [" + std::to_string(need_lo) + ", " +
                        std::to_string(need_hi) + ")");
    }
    t_total.start(st);
    const bool was_on_gpu = h->on_gpu;
    // the copy the constructor left on the device serves searches of a stack that is "not on the GPU" as well
    const bool cached = !was_on_gpu && h->dev_cache;
    if (cached) h->zero_marker = h->zm_known;
    // A stack brought up for this search only (possibly just the rows in reach) must not outlive the call, also
    // when the call fails half way.
    struct DropTransientStack {
        kb_handle h;
        bool active;
        ~DropTransientStack() {
            if (!active) return;
            cudaStreamSynchronize(h->stream);
            if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
            for (int i = 0; i < 2; ++i)
                if (h->up.band_stream[i]) cudaStreamSynchronize(h->up.band_stream[i]);
            free_device_stack(h, true);
        }
    } drop{h, !was_on_gpu && !cached};
    t_h2d.start(st);
    // Not preloaded: the stack comes up for this search only.  From the second search of a handle on, in row chunks
    // on a copy stream, and the tiled kernel follows the rows as they land.
    bool pipelined = false;
    if (!was_on_gpu && !cached) {
        uint64_t copied = h->meta.total_array_size;
        if (h->kernel_mode != 1) {
            // rows within reach of the start rows: |vy| * max|t| plus a tile, its padding and a margin
            double vmax = 0.0, tabs = 0.0;
            for (uint64_t i = 0; i < n_cand; ++i) vmax = std::max(vmax, (double)fabsf(vy[i]));
            for (double t : h->times) tabs = std::max(tabs, fabs(t));
            const double reach_d = std::min(vmax * tabs + 64.0, 4.0e6);
            const int reach = (int)reach_d;
            pipelined = upload_stack_pipelined(h, p.y_start_min - reach, p.y_start_max + reach, &copied);
        }
        if (!pipelined) upload_stack(h);
        h->timing.h2d_stack_bytes = copied;
        launches += pipelined ? 2 * UploadProgress::kMaxChunks : 2;  // canonicalise + quantise kernels
    }
    t_h2d.stop(st);

    t_prep.start(st);
    const double t_host0 = now_s();
    h->d_vx.ensure(std::max<uint64_t>(n_cand, 1));
    h->d_vy.ensure(std::max<uint64_t>(n_cand, 1));
    if (n_cand) {
        KB_CUDA(cudaMemcpyAsync(h->d_vx.p, vx, sizeof(float) * n_cand, cudaMemcpyHostToDevice, st));
        KB_CUDA(cudaMemcpyAsync(h->d_vy.p, vy, sizeof(float) * n_cand, cudaMemcpyHostToDevice, st));
    }
    h->d_slots.ensure(n_slots);
    const StackView sv = make_view(h);
    // Which kernel: the tiled one when it can take the search (results_per_pixel <= 8, ...); the candidate-parallel
    // one for longer per-pixel lists over few start pixels; the generic one for everything else.  Whenever the tiled
    // kernel is not the one that runs, the reason is kept for the caller (kb_search_note; the Python layer logs it).
    TiledPlan *plan = nullptr;
    std::string why;
    bool use_pairs = false;
    h->search_note.clear();
    if (h->kernel_mode == 0 || h->kernel_mode == 2) {
        const bool hit = h->plan_cache && h->plan_vx.size() == n_cand && h->plan_screen == sv.screen &&
                         h->plan_steps[0] == sv.q_step_psi && h->plan_steps[1] == sv.q_step_phi &&
                         h->plan_params.x_start_min == p.x_start_min &&
                         h->plan_params.results_per_pixel == p.results_per_pixel &&
                         h->plan_params.do_sigmag_filter == p.do_sigmag_filter &&
                         memcmp(&h->plan_params.min_lh, &p.min_lh, sizeof(float)) == 0 &&
                         (n_cand == 0 || (memcmp(h->plan_vx.data(), vx, sizeof(float) * n_cand) == 0 &&
                                          memcmp(h->plan_vy.data(), vy, sizeof(float) * n_cand) == 0));
        if (hit) {
            plan = h->plan_cache;
        } else {
            if (h->plan_cache) tiled_plan_destroy(h->plan_cache);
            h->plan_cache = nullptr;
            plan = tiled_plan_create(sv, p, vx, vy, n_cand, st, &why, h->times.data(), h->runs.data());
            if (plan) {
                h->plan_cache = plan;
                h->plan_vx.assign(vx, vx + n_cand);
                h->plan_vy.assign(vy, vy + n_cand);
                h->plan_params = p;
                h->plan_screen = sv.screen;
                h->plan_steps[0] = sv.q_step_psi;
                h->plan_steps[1] = sv.q_step_phi;
            }
        }
        if (!plan && h->kernel_mode == 2) throw Error("tiled kernel not usable: " + why);
    } else {
        why = h->kernel_mode == 1 ? "generic kernel requested" : "candidate-parallel kernel requested";
    }
    // automatic mode, a start window of a few tiles only (the tiled kernel gives every tile to one SM): spread the
    // (pixel, candidate) pairs over the device instead
    const uint64_t n_tiles = ((sw + 63) / 64) * ((sh + 14) / 15);
    if (plan && h->kernel_mode == 0 && n_tiles < 37 && search_pairs_usable(p, n_cand, nullptr)) {
        plan = nullptr;   // stays in the cache
        why = "the start window covers only " + std::to_string(n_tiles) + " tiles of the tiled kernel";
    }
    if (!plan && (h->kernel_mode == 3 || (h->kernel_mode == 0 && (p.results_per_pixel > (uint32_t)kMaxRegK || n_tiles < 37)))) {
        std::string why_pairs;
        use_pairs = search_pairs_usable(p, n_cand, &why_pairs);
        if (!use_pairs && h->kernel_mode == 3) throw Error("candidate-parallel kernel not usable: " + why_pairs);
        if (!use_pairs) why += "; " + why_pairs;
    }
    if (!plan) {
        h->search_note = std::string(use_pairs ? "candidate-parallel kernel used: " : "generic kernel used: ") + why;
        if (getenv("KB_DEBUG")) fprintf(stderr, "[kbmod_b200] %s\n", h->search_note.c_str());
    }
    if (use_pairs) h->d_pairs.ensure(sw * sh * n_cand);
    if (getenv("KB_DEBUG"))
        fprintf(stderr, "[kbmod_b200] search set-up on the host (buffers, schedule%s): %.2f ms\n",
                plan && plan == h->plan_cache ? "" : " n/a", 1e3 * (now_s() - t_host0));
    t_prep.stop(st);

    t_search.start(st);
    uint32_t sl = 0;
    try {
        if (pipelined && !plan)  // the generic kernel reads any row: wait for the whole stack
            KB_CUDA(cudaStreamWaitEvent(st, h->up.ev[h->up.n_chunks - 1], 0));
        if (plan) {
            bool overflowed = false;
            sl = launch_search_tiled(plan, sv, p, h->d_vx.p, h->d_vy.p, h->d_slots.p, h->clip, h->scratch, st,
                                     pipelined ? &h->up : nullptr, &overflowed);
            if (overflowed) {   // a single row of tiles overflowed the request queue: the generic kernel, slowly
                h->search_note = "generic kernel used: the request queue of the tiled kernel overflowed";
                if (pipelined) KB_CUDA(cudaStreamWaitEvent(st, h->up.ev[h->up.n_chunks - 1], 0));
                sl += launch_search_generic(sv, p, h->d_vx.p, h->d_vy.p, n_cand, h->d_slots.p, st);
                plan = nullptr;
            }
        }
        else if (use_pairs)
            sl = launch_search_pairs(sv, p, h->d_vx.p, h->d_vy.p, n_cand, h->d_slots.p, h->d_pairs.p, h->scratch, st);
        else sl = launch_search_generic(sv, p, h->d_vx.p, h->d_vy.p, n_cand, h->d_slots.p, st);
    } catch (...) {
        if (h->plan_cache) tiled_plan_destroy(h->plan_cache);   // do not trust a schedule whose launch failed
        h->plan_cache = nullptr;
        throw;
    }
    t_search.stop(st);
    launches += sl;
    h->timing.search_launches = sl;
    h->timing.used_tiled = plan ? 1 : (use_pairs ? 2 : 0);

    t_post.start(st);
    uint64_t n_keep = 0;
    launches += filter_count(h->d_slots.p, n_slots, p.min_lh, p.min_observations, h->scratch, &n_keep, st);
    const float main_ms = plan ? tiled_plan_kernel_ms(plan) : -1.0f;   // the stream is idle (filter_count synchronised)
    if (plan && p.do_sigmag_filter) {
        // Screening that lets a large share of the pairs through has stopped paying: say so (results are unaffected).
        const double pairs = (double)(p.x_start_max - p.x_start_min) * (double)(p.y_start_max - p.y_start_min) * (double)n_cand;
        const double req = (double)tiled_plan_requests(plan);
        if (pairs >= 1e8 && req > 0.05 * pairs) {
            char buf[256];
            snprintf(buf, sizeof(buf),
                     "screening passed %.0f %% of the (start pixel, candidate) pairs on to exact evaluation: the stack's "
                     "largest |psi| or phi is too far above its noise for the 16-bit screening copy (mask saturated "
                     "pixels), or min_lh is low", 100.0 * req / pairs);
            h->search_note = buf;
        }
    }
    h->d_results.ensure(std::max<uint64_t>(n_keep, 1));
    bool invalid = false;
    launches += filter_finish(h->d_slots.p, n_slots, n_keep, (int)p.results_per_pixel, (int)sw, p.x_start_min,
                              p.y_start_min, p.min_lh, p.min_observations, h->d_vx.p, h->d_vy.p, h->scratch,
                              h->d_results.p, &invalid, st);
    t_post.stop(st);
    t_total.stop(st);
    KB_CUDA(cudaStreamSynchronize(st));
    h->n_results = n_keep;
    h->n_slots = n_slots;
    h->last_params = p;
    h->have_slots = true;
    h->timing.ms_h2d = t_h2d.ms();
    h->timing.ms_prepare = t_prep.ms();
    h->timing.ms_search = t_search.ms();
    h->timing.ms_main_kernel = main_ms >= 0.0f ? main_ms : h->timing.ms_search;
    h->timing.ms_filter_sort = t_post.ms();
    h->timing.ms_total = t_total.ms();
    h->timing.total_launches = launches;
    // search_all drops the stack again unless preloaded (:257): `drop` does it on every way out of this function
    if (invalid) throw Error("Invalid trajectory detected in the search results.");
    if (n_results) *n_results = n_keep;
    KB_API_END
}

int kb_num_results(kb_handle h, uint64_t *n) {
    KB_API_BEGIN
    *n = h->n_results;
    KB_API_END
}

int kb_get_results(kb_handle h, uint64_t start, uint64_t count, kb_trajectory *out, uint64_t *n_copied) {
    KB_API_BEGIN
    use_device(h);
    if (count == 0) throw Error("count must be greater than 0");
    uint64_t n = 0;
    if (start < h->n_results) n = std::min(count, h->n_results - start);
    if (n) KB_CUDA(cudaMemcpy(out, h->d_results.p + start, sizeof(kb_trajectory) * n, cudaMemcpyDeviceToHost));
    if (n_copied) *n_copied = n;
    KB_API_END
}

int kb_set_results(kb_handle h, const kb_trajectory *list, uint64_t n) {
    KB_API_BEGIN
    use_device(h);
    for (uint64_t i = 0; i < n; ++i) {
        const kb_trajectory &t = list[i];
        if (!(isfinite(t.vx) && isfinite(t.vy) && isfinite(t.lh) && isfinite(t.flux) && t.obs_count >= 0))
            throw Error("Invalid trajectory detected at index " + std::to_string(i));
    }
    h->d_results.ensure(std::max<uint64_t>(n, 1));
    if (n) KB_CUDA(cudaMemcpy(h->d_results.p, list, sizeof(kb_trajectory) * n, cudaMemcpyHostToDevice));
    h->n_results = n;
    KB_API_END
}

int kb_clear_results(kb_handle h) {
    KB_API_BEGIN
    h->n_results = 0;
    KB_API_END
}

int kb_get_timing(kb_handle h, kb_timing *t) {
    KB_API_BEGIN
    *t = h->timing;
    KB_API_END
}

int kb_timer_start(kb_handle h) {
    KB_API_BEGIN
    use_device(h);
    if (!h->ev_start) {
        KB_CUDA(cudaEventCreate(&h->ev_start));
        KB_CUDA(cudaEventCreate(&h->ev_stop));
    }
    KB_CUDA(cudaStreamSynchronize(h->stream));
    KB_CUDA(cudaEventRecord(h->ev_start, h->stream));
    KB_API_END
}

int kb_timer_stop(kb_handle h, float *ms) {
    KB_API_BEGIN
    use_device(h);
    if (!h->ev_start) throw Error("kb_timer_start has not been called.");
    KB_CUDA(cudaEventRecord(h->ev_stop, h->stream));
    KB_CUDA(cudaEventSynchronize(h->ev_stop));
    KB_CUDA(cudaEventElapsedTime(ms, h->ev_start, h->ev_stop));
    KB_API_END
}

int kb_get_slots(kb_handle h, kb_trajectory *out, uint64_t capacity, uint64_t *n_slots) {
    KB_API_BEGIN
    use_device(h);
    if (!h->have_slots) throw Error("No search has been run on this handle.");
    if (n_slots) *n_slots = h->n_slots;
    if (!out) return 0;
    if (capacity < h->n_slots) throw Error("Output buffer too small for the result slots.");
    const kb_params &p = h->last_params;
    h->d_tmp.ensure(h->n_slots);
    slots_to_trajectories(h->d_slots.p, h->n_slots, (int)p.results_per_pixel, p.x_start_max - p.x_start_min,
                          p.x_start_min, p.y_start_min, h->d_vx.p, h->d_vy.p, h->d_tmp.p, h->stream);
    KB_CUDA(cudaMemcpyAsync(out, h->d_tmp.p, sizeof(kb_trajectory) * h->n_slots, cudaMemcpyDeviceToHost, h->stream));
    KB_CUDA(cudaStreamSynchronize(h->stream));
    KB_API_END
}

int kb_evaluate_trajectories(kb_handle h, const kb_params *params, kb_trajectory *trjs, uint64_t n,
                             uint8_t *keep_mask) {
    KB_API_BEGIN
    use_device(h);
    if (n == 0) return 0;
    const bool was_on_gpu = h->on_gpu;
    if (!was_on_gpu) upload_stack(h);
    DevBuf<kb_trajectory> d_t;
    DevBuf<uint8_t> d_m;
    d_t.ensure(n);
    if (keep_mask) d_m.ensure(n * (uint64_t)h->T);
    KB_CUDA(cudaMemcpyAsync(d_t.p, trjs, sizeof(kb_trajectory) * n, cudaMemcpyHostToDevice, h->stream));
    launch_evaluate(make_view(h), *params, d_t.p, n, keep_mask ? d_m.p : nullptr, h->stream);
    KB_CUDA(cudaMemcpyAsync(trjs, d_t.p, sizeof(kb_trajectory) * n, cudaMemcpyDeviceToHost, h->stream));
    if (keep_mask)
        KB_CUDA(cudaMemcpyAsync(keep_mask, d_m.p, n * (uint64_t)h->T, cudaMemcpyDeviceToHost, h->stream));
    KB_CUDA(cudaStreamSynchronize(h->stream));
    if (!was_on_gpu) free_device_stack(h, true);
    KB_API_END
}

int kb_psi_phi_curves(kb_handle h, const kb_trajectory *trjs, uint64_t n, float *out) {
    KB_API_BEGIN
    use_device(h);
    if (n == 0) return 0;
    // The curves are read from the reference-layout array (psi and phi validity are independent there).
    // Pinned host memory is directly addressable by the device (UVA); otherwise stage a temporary copy.
    // Preferably the reference-layout bytes the constructor kept on the device; else pinned host memory through UVA;
    // else a temporary upload.
    DevBuf<uint8_t> d_tmp_raw;
    const void *src = h->d_raw;
    if (!src) {
        ensure_host(h);
        src = h->host_array;
        if (!h->host_pinned) {
            d_tmp_raw.ensure(h->meta.total_array_size);
            KB_CUDA(cudaMemcpyAsync(d_tmp_raw.p, h->host_array, h->meta.total_array_size, cudaMemcpyHostToDevice, h->stream));
            src = d_tmp_raw.p;
        }
    }
    DevBuf<kb_trajectory> d_t;
    DevBuf<float> d_out;
    DevBuf<double> d_times;
    d_t.ensure(n);
    d_out.ensure(n * 2 * (uint64_t)h->T);
    d_times.ensure(h->T);
    KB_CUDA(cudaMemcpyAsync(d_times.p, h->times.data(), sizeof(double) * h->T, cudaMemcpyHostToDevice, h->stream));
    KB_CUDA(cudaMemcpyAsync(d_t.p, trjs, sizeof(kb_trajectory) * n, cudaMemcpyHostToDevice, h->stream));
    launch_curves(src, d_times.p, h->T, h->H, h->W, h->nb, h->meta.psi_min_val, h->meta.psi_scale,
                  h->meta.phi_min_val, h->meta.phi_scale, d_t.p, n, d_out.p, h->stream);
    KB_CUDA(cudaMemcpyAsync(out, d_out.p, sizeof(float) * n * 2 * (uint64_t)h->T, cudaMemcpyDeviceToHost, h->stream));
    KB_CUDA(cudaStreamSynchronize(h->stream));
    KB_API_END
}

int kb_sigmag_filtered_indices(const float *values, int n, float sgl0, float sgl1, float coeff, float width,
                               int *idx_out, int *n_kept) {
    KB_API_BEGIN
    *n_kept = 0;
    if (n <= 0) return 0;
    DevBuf<float> d_v;
    DevBuf<int> d_idx, d_range;
    d_v.ensure(n);
    d_idx.ensure(n);
    d_range.ensure(2);
    KB_CUDA(cudaMemcpy(d_v.p, values, sizeof(float) * n, cudaMemcpyHostToDevice));
    launch_sigmag_indices(d_v.p, n, sgl0, sgl1, coeff, width, d_idx.p, d_range.p, 0);
    int range[2];
    std::vector<int> idx(n);
    KB_CUDA(cudaMemcpy(range, d_range.p, sizeof(range), cudaMemcpyDeviceToHost));
    KB_CUDA(cudaMemcpy(idx.data(), d_idx.p, sizeof(int) * n, cudaMemcpyDeviceToHost));
    int k = 0;
    for (int i = range[0]; i <= range[1] && i < n; ++i)
        if (i >= 0) idx_out[k++] = idx[i];
    *n_kept = k;
    KB_API_END
}

int kb_grid_filter(const kb_trajectory *trjs, uint64_t n, double bin_width, double max_time, uint32_t *idx_out,
                   uint64_t *n_out) {
    KB_API_BEGIN
    if (!n_out) throw Error("null argument");
    *n_out = 0;
    if (!(bin_width >= 1.0) || !std::isfinite(bin_width))
        throw Error("Bin width must be at least 1. Got " + std::to_string(bin_width) + ".");
    if (!(max_time >= 0.0) || !std::isfinite(max_time))
        throw Error("Max time must be >= 0. Got " + std::to_string(max_time) + ".");
    if (n == 0) return 0;
    if (!trjs || !idx_out) throw Error("null argument");
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0)
        throw Error("no CUDA device: kbmod_b200 has no CPU path");
    DevBuf<kb_trajectory> d_list;
    DevBuf<uint32_t> d_idx;
    d_list.ensure(n);
    d_idx.ensure(n);
    KB_CUDA(cudaMemcpy(d_list.p, trjs, sizeof(kb_trajectory) * n, cudaMemcpyHostToDevice));
    ResultScratch scr;
    try {
        grid_filter_indices(d_list.p, n, bin_width, max_time, d_idx.p, n_out, scr, 0);
        KB_CUDA(cudaMemcpy(idx_out, d_idx.p, sizeof(uint32_t) * (*n_out), cudaMemcpyDeviceToHost));
    } catch (...) {
        result_scratch_free(scr);
        throw;
    }
    result_scratch_free(scr);
    KB_API_END
}

// The merge step of the multi-GPU search on one device, from host lists (what kb_merge_results_allgather does with
// the blocks ncclAllGather delivers): n_lists lists, each sorted by lh descending, merged into one; equal lh values
// keep (list, position) order.
int kb_merge_sorted_lists(const kb_trajectory *const *lists, const uint64_t *counts, int n_lists, uint64_t top_n,
                          kb_trajectory *out, uint64_t *n_out) {
    KB_API_BEGIN
    if (n_lists <= 0 || !counts || !lists || !n_out) throw Error("kb_merge_sorted_lists: bad arguments");
    uint64_t cap = 0, total = 0;
    for (int r = 0; r < n_lists; ++r) {
        cap = std::max(cap, counts[r]);
        total += counts[r];
    }
    const uint64_t keep = (top_n && top_n < total) ? top_n : total;
    *n_out = keep;
    if (keep == 0) return 0;
    const size_t hdr = (size_t)merge_header_bytes();
    const size_t blk = (hdr + sizeof(kb_trajectory) * cap + 15) / 16 * 16;
    DevBuf<uint8_t> d_blocks;
    DevBuf<kb_trajectory> d_out, d_in;
    DevBuf<uint64_t> d_total;
    d_blocks.ensure(blk * n_lists);
    d_out.ensure(keep);
    d_in.ensure(std::max<uint64_t>(cap, 1));
    d_total.ensure(1);
    for (int r = 0; r < n_lists; ++r) {
        if (counts[r]) KB_CUDA(cudaMemcpy(d_in.p, lists[r], sizeof(kb_trajectory) * counts[r], cudaMemcpyHostToDevice));
        merge_pack_block(d_blocks.p + (size_t)r * blk, d_in.p, counts[r], 0);
        KB_CUDA(cudaStreamSynchronize(0));
    }
    merge_sorted_runs(d_blocks.p, blk, n_lists, cap, d_out.p, keep, d_total.p, 0);
    KB_CUDA(cudaMemcpy(out, d_out.p, sizeof(kb_trajectory) * keep, cudaMemcpyDeviceToHost));
    KB_API_END
}

}  // extern "C"

// ---- accessors used by nccl_merge.cu ------------------------------------------------------------------
namespace kb {
cudaStream_t handle_stream(kb_handle h) { return h->stream; }
int handle_device(kb_handle h) { return h->device; }
kb_trajectory *handle_results(kb_handle h, uint64_t *n) {
    *n = h->n_results;
    return h->d_results.p;
}
// Room for `capacity` results in the handle's list (contents undefined afterwards); the caller fills it on the
// handle's stream and then sets the count.
kb_trajectory *handle_results_reserve(kb_handle h, uint64_t capacity) {
    h->d_results.ensure(std::max<uint64_t>(capacity, 1));
    return h->d_results.p;
}
void handle_set_result_count(kb_handle h, uint64_t n) { h->n_results = n; }
}  // namespace kb
