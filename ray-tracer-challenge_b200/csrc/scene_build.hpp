// Host-side staging of the device scene (see scene_build.cpp).
#pragma once
#include <vector_functions.h>

#include <cstddef>
#include <vector>

#include "device_scene.h"

namespace rtc {

struct HostScene {
    std::vector<rtc_light> lights;
    std::vector<rtc_material> materials;
    std::vector<rtc_pattern> patterns;
    std::vector<rtc_uv_pattern> uvs;
    std::vector<rtc_image> images;
    std::vector<uint8_t> texels;
    std::vector<rtc_transform> transforms;
    std::vector<rtc_leaf> leaves;
    std::vector<rtc_tri_normals> normals;
    std::vector<rtc_triangle> blob_tris;
    std::vector<rtc_guard> guards;
    std::vector<rtc_node> nodes;
    std::vector<DPrim> prims;
    std::vector<DMesh> meshes;
    std::vector<BvhNode> bvh;
    std::vector<DTri> tris;
    std::vector<int32_t> mesh_leaf;  // DMesh.leaf_map tables, one after the other
    int n_geometries = 0;            // distinct BVHs (<= meshes.size(): instances share them)
    std::vector<int4> csg_prog;
    std::vector<DCsgRoot> csg_roots;
    int bvh_max_depth = 0;
    float world_abs[3] = {0, 0, 0};
    int csg_max_segments = 0;
};

// Throws std::runtime_error on a malformed blob.
void build_host_scene(const void* blob, size_t nbytes, HostScene& out);

}  // namespace rtc
