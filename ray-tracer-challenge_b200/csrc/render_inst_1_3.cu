// render_kernel<MESH = 1, LPP, FEAT = 3> (render_kernel.cuh; FEAT bits: 1 = CSG / guards, 2 = general patterns).
#include "render_launch.cuh"

namespace rtc {
RTC_DEFINE_INST(1, 3)
}  // namespace rtc
