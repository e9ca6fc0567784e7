// search_tiled_nb1.cu -- instantiates the tiled search and clip kernels for uint8 stacks.
#include "search_tiled_kernel.cuh"

namespace kb {

void launch_tiled_nb1(TiledPlan *plan, const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy,
                      Slot *d_slots, dim3 grid, cudaStream_t st) {
    launch_nb<1>(plan, sv, p, d_vx, d_vy, d_slots, grid, st);
}

void launch_clip_nb1(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                     cudaStream_t st) {
    launch_clip<1>(sv, p, d_vx, d_vy, clip, st);
}

void launch_exact_nb1(const StackView &sv, const kb_params &p, const float *d_vx, const float *d_vy, ClipScratch &clip,
                      cudaStream_t st) {
    launch_exact<1>(sv, p, d_vx, d_vy, clip, st);
}

}  // namespace kb
