// render_kernel<MESH = 0, LPP, FEAT = 0> (render_kernel.cuh; FEAT bits: 1 = CSG / guards, 2 = general patterns).
#include "render_launch.cuh"

namespace rtc {
RTC_DEFINE_INST(0, 0)
}  // namespace rtc
