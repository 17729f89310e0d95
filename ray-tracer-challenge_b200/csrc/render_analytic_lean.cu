// Instantiations of the render kernel for scenes WITHOUT triangle meshes and nothing but solid patterns, no CSG, no
// guard boxes (render_kernel.cuh LEAN): cover.json.
#include "render_launch.cuh"

namespace rtc {

int launch_render_analytic_lean(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    return launch_render_impl<false, true>(S, P, lpp, grid, smem, stream);
}
int occupancy_render_analytic_lean(int* blocks_per_sm, size_t smem) { return occupancy_render_impl<false, true>(blocks_per_sm, smem); }

}  // namespace rtc
