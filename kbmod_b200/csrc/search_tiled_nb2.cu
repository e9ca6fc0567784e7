// search_tiled_nb2.cu -- instantiates the tiled search and clip kernels for uint16 stacks.
#include "search_tiled_kernel.cuh"

namespace kb {

void launch_tiled_nb2(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st) {
    launch_nb<2>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
}

void launch_clip_nb2(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                     cudaStream_t st) {
    launch_clip<2>(sv, p, d_vx, d_vy, clip, st);
}

void launch_exact_nb2(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                      cudaStream_t st) {
    launch_exact<2>(sv, p, d_vx, d_vy, clip, st);
}

}  // namespace kb
