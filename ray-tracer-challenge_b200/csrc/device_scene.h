// Device-resident scene: what rtc_scene_create builds out of a blob (include/rtc_blob.h) and what
// the render kernels read.  Layout rules: every record is a multiple of 16 bytes and is fetched
// with 16-byte loads (double2 / float4 / int4); arrays are 256-byte aligned cudaMalloc regions of
// one slab, so one H2D copy uploads a scene.
#pragma once
#include <vector_types.h>

#include <cstdint>

#include "../../include/rtc_b200.h"

namespace rtc {

// Triangle record in BVH leaf order, object space, f64 (triangle.zig:22-27).  80 bytes = 5 x 16.
struct DTri {
    double p1[3], e1[3], e2[3];
    int32_t leaf;  // rank of the triangle in its geometry's canonical order: DScene.mesh_leaf[DMesh.leaf_map + leaf] is
                   // the rtc_leaf index (material / normals / tie-break order) in the instance being traversed
    int32_t pad;
};
static_assert(sizeof(DTri) == 80, "DTri must be 80 bytes");

// Two-child BVH node with both child boxes inline, f32 rounded OUTWARD from the f64 triangle
// bounds.  64 bytes = 4 x float4.  n == 0: child c is an inner node; n > 0: c is the first of n
// DTri records.  An unused child has an inverted box (lo=+inf, hi=-inf) and n = 0, c = -1.
struct BvhNode {
    float lo0[3], hi0[3], lo1[3], hi1[3];
    int32_t c0, c1, n0, n1;  // c: inner node index, or ~((first triangle << 3) | count) for a leaf child; n: its triangle count
};
static_assert(sizeof(BvhNode) == 64, "BvhNode must be 64 bytes");

// One mesh INSTANCE = all non-CSG triangles sharing a world->object matrix and a casts_shadow flag.  Instances with the
// same object-space triangles share one geometry (BVH nodes + DTri records, scene_build.cpp): `root` is then the same.
struct DMesh {
    int32_t transform;   // rtc_transform index
    uint32_t flags;      // RTC_CASTS_SHADOW
    int32_t root;        // BvhNode index of the root
    int32_t n_tris;
    float lo[3], hi[3];  // object-space root box (outward rounded)
    float bound_abs[3];  // max |coordinate| of the root box per axis: scales the f32 slab error bound
    int32_t leaf_map;    // first entry of this instance in DScene.mesh_leaf (n_tris entries)
    float pad[2];
    float wlo[3], whi[3];  // conservative WORLD-space box of the mesh, f32 rounded outward (pre-test before
    float pad2[2];         // the ray transform); +-inf when the transform is degenerate
};
static_assert(sizeof(DMesh) % 16 == 0, "DMesh alignment");

// Top-level analytic leaf (or stand-alone triangle): conservative f32 WORLD box + the rtc_leaf it
// guards.  32 bytes = 2 x 16: the box is tested before the 96-byte rtc_leaf is touched.
struct DPrim {
    float lo[3], hi[3];  // ignored when RTC_UNBOUNDED is set in flags
    int32_t leaf;
    uint32_t flags;  // copy of rtc_leaf.flags
};
static_assert(sizeof(DPrim) == 32, "DPrim must be 32 bytes");

// CSG evaluation program (compiled from rtc_node trees): int4 words {op, a, b, c}.
enum CsgOpcode {
    CSG_OP_LEAF = 0,    // a = leaf: append its intersections as a new segment
    CSG_OP_BBOX = 1,    // a = rtc_node index, b = pc to jump to (after pushing an empty segment) on a miss
    CSG_OP_MERGE = 2,   // a = n: fuse the top n segments (group.zig:52-58)
    CSG_OP_FILTER = 3,  // a = rtc_csg_op, b,c = leaf range of the LEFT operand: fuse top 2, sort, filter
    CSG_OP_END = 4
};
struct DCsgRoot {
    int32_t pc;                    // first program word
    int32_t leaf_begin, leaf_end;  // leaves owned by this object
    int32_t pad;
};

struct DScene {
    const rtc_light* lights;
    const rtc_material* materials;
    const rtc_pattern* patterns;
    const rtc_uv_pattern* uvs;
    const rtc_image* images;
    const uint8_t* texels;
    const rtc_transform* transforms;
    const rtc_leaf* leaves;
    const rtc_tri_normals* normals;  // indexed by rtc_leaf.triangle
    const rtc_triangle* blob_tris;   // original order (CSG operands, guarded triangles)
    const rtc_guard* guards;
    const rtc_node* nodes;
    const DPrim* prims;  // top-level analytic leaves (and triangles kept out of meshes)
    const DMesh* meshes;
    const BvhNode* bvh;
    const DTri* tris;
    const int32_t* mesh_leaf;  // per instance: canonical triangle rank -> rtc_leaf index
    const int4* csg_prog;
    const DCsgRoot* csg_roots;
    int32_t n_lights, n_prims, n_meshes, n_csg;
    int32_t n_leaves, bvh_max_depth;
    float world_abs[3];  // max finite |coordinate| of any world pre-test box per axis (f32 slab error bound)
    int32_t ri_all_one;  // every material has refractive_index == 1.0: n1 = n2 = 1 whatever the containers are
    int32_t pad_ri[3];
    int32_t stack_cap;   // entries per thread of the shared-memory BVH traversal stack (bvh_max_depth + 2)
};

}  // namespace rtc
