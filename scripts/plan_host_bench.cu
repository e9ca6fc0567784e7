// plan_host_bench.cu -- host-side timing and fingerprint of tiled_plan_create (the candidate schedule of the tiled
// search kernel) without a GPU.  The function is host-only until its first cudaMalloc, which fails on a machine
// without a device; KB_DEBUG=1 prints the phase timings before that, KB_DEBUG_PLAN_HASH=1 a hash of the schedule, so a
// change of the host code can be checked for bit-identical output and timed where there is no GPU.
//   cd kbmod_b200/csrc && make -j8 && nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -I. -Xcompiler -fopenmp \
//       ../../scripts/plan_host_bench.cu build/*.o -lcudart_static -lgomp -ldl -lpthread -lrt -o /tmp/plan_host_bench
//   KB_DEBUG=1 KB_DEBUG_PLAN_HASH=1 /tmp/plan_host_bench 400 400 250 0     # T, velocity grid nx x ny, sigma-G on/off
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "common.cuh"
using namespace kb;
int main(int argc, char **argv) {
    const int T = argc > 1 ? atoi(argv[1]) : 400;
    const int nx = argc > 2 ? atoi(argv[2]) : 400, ny = argc > 3 ? atoi(argv[3]) : 250;
    const int sigmag = argc > 4 ? atoi(argv[4]) : 0;
    std::vector<double> times(T);
    // 8 per night, 0.005 d apart
    for (int t = 0; t < T; ++t) times[t] = (t / 8) * 1.0 + (t % 8) * 0.005;
    std::vector<float> vx, vy;
    for (int j = 0; j < ny; ++j) for (int i = 0; i < nx; ++i) { vx.push_back(-4.0f + 8.0f * i / nx); vy.push_back(-2.5f + 5.0f * j / ny); }
    std::vector<int2> runs;
    for (int t = 0; t + 8 <= T; t += 8) { runs.push_back(make_int2(t, 8)); runs.push_back(make_int2(t, 4)); runs.push_back(make_int2(t + 4, 4)); for (int k = 0; k < 8; k += 2) runs.push_back(make_int2(t + k, 2)); }
    StackView sv{};
    sv.T = T; sv.H = 2048; sv.W = 2048; sv.Wp = 2048; sv.nb = 4; sv.screen = (uint32_t *)0x1000; sv.q_step_psi = 1e-3f; sv.q_step_phi = 1e-3f;
    sv.all_valid = 1; sv.n_runs = (int)runs.size();
    kb_params p{};
    p.results_per_pixel = 8; p.do_sigmag_filter = sigmag; p.min_observations = T / 2; p.min_lh = 10.0f;
    p.x_start_min = 0; p.x_start_max = 2048; p.y_start_min = 0; p.y_start_max = 2048;
    std::string why;
    auto t0 = std::chrono::steady_clock::now();
    try {
        TiledPlan *pl = tiled_plan_create(sv, p, vx.data(), vy.data(), vx.size(), 0, &why, times.data(), runs.data());
        printf("plan %p why=%s\n", (void *)pl, why.c_str());
    } catch (const std::exception &e) {
        printf("stopped at the first CUDA call: %s\n", e.what());
    }
    printf("host time until then: %.3f s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    return 0;
}
