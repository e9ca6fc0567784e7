// search_tiled.cu -- placeholder until the TMA-staged kernel lands: always defers to the generic kernel.
#include "common.cuh"
namespace kb {
struct TiledPlan {};
TiledPlan *tiled_plan_create(const StackView &, const kb_params &, const float *, const float *, uint64_t,
                             cudaStream_t, std::string *why_not) {
    if (why_not) *why_not = "tiled kernel not built";
    return nullptr;
}
int launch_search_tiled(TiledPlan *, const StackView &, const kb_params &, const float *, const float *, Slot *,
                        cudaStream_t) {
    throw Error("tiled kernel not built");
}
void tiled_plan_destroy(TiledPlan *p) { delete p; }
}  // namespace kb
