// Host BVH builder for one mesh: binned-SAH, two children per node, <= 4 triangles per leaf.
// Replaces (for traversal only) the reference's divide(8) group tree (shape.zig:372-399,
// group.zig:85-135): boxes here are supersets of their triangles, so culling by them cannot change
// which triangles a ray hits; tie-break order is carried by DTri.leaf instead of by tree shape.
#pragma once
#include <vector>

#include "device_scene.h"

namespace rtc {

struct BvhBuild {
    std::vector<BvhNode> nodes;  // nodes[0] is the root
    std::vector<DTri> tris;      // leaf order
    int max_depth = 0;
    float lo[3], hi[3];          // root box
};

// tris: object-space triangle records of the mesh (any order).
BvhBuild build_bvh(const std::vector<DTri>& tris);
// The same contract, built on the current CUDA device (bvh_gpu.cu): an LBVH.  Throws std::runtime_error on a CUDA error.
BvhBuild build_bvh_gpu(const std::vector<DTri>& tris);
// RTC_BVH=host|gpu|auto (default auto: meshes of >= 65,536 triangles are built on the device when there is one)
bool bvh_build_on_gpu(size_t n_tris);

}  // namespace rtc
