// World tree -> pointer-free blob (include/rtc_blob.h).  This is the walk the Zig side of the
// boundary performs over World(f64).objects (INTEGRATION.md shows the Zig twin); here it runs over
// the C++ mirror of that tree so the boundary can be exercised without a Zig toolchain.
#pragma once
#include <cstdint>
#include <vector>

#include "../../../include/rtc_b200.h"
#include "scene.hpp"

namespace rtc {

std::vector<uint8_t> flatten_world(const World& world);
rtc_camera flatten_camera(const Camera& cam);

}  // namespace rtc
