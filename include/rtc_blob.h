/* rtc_blob.h — the flattened, pointer-free scene ("blob") that crosses the drop-in boundary.
 *
 * The host side of the boundary (Zig: a walk over World(f64).objects, src/raytracer/world.zig:24-25;
 * here: ray-tracer-challenge_b200/csrc/host/flatten.cpp over the C++ mirror of that tree) emits one
 * contiguous buffer: header, then sections of fixed-size little-endian records.  Every section
 * offset is a multiple of 16 bytes.  Records only hold indices into other sections, never pointers.
 *
 * What is kept from the reference tree (src/raytracer/shapes/shape.zig:99-131):
 *   - LEAVES in depth-first order of World.objects (groups dissolved).  The leaf index IS the
 *     tie-break order the reference gets from "concatenate children, stable-sort by t" at every
 *     level (group.zig:49-58, world.zig:74-81): equal t  =>  lower leaf index first.
 *   - each leaf's own world->object matrix (groups/CSG push theirs down, shape.zig:286-310),
 *     material, casts_shadow, cylinder/cone min/max/closed, triangle data;
 *   - CSG subtrees verbatim (csg.zig:26-31): operation, operands, and the CSG node's `_bbox`,
 *     which the reference never re-transforms (shape.zig:298-302) and which is therefore applied
 *     with the reference's own slab rule instead of being treated as a conservative hint;
 *   - GUARDS: group bounding boxes that are NOT supersets of their members (bounding_box.zig:24-32:
 *     @min/@max drop the NaN corners an unbounded child produces).  A leaf with a guard chain is
 *     only intersected when the ray passes every box of the chain, as in group.zig:46-50.
 *   Group boxes that are supersets are dropped: culling by them cannot change the hit set.
 */
#ifndef RTC_BLOB_H
#define RTC_BLOB_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTC_BLOB_MAGIC 0x42435452u /* "RTCB" */
#define RTC_BLOB_VERSION 2u

/* Shape.Variant tags that can be leaves (shape.zig:99-111) */
enum rtc_leaf_kind {
    RTC_SPHERE = 0,
    RTC_PLANE = 1,
    RTC_CUBE = 2,
    RTC_CYLINDER = 3,
    RTC_CONE = 4,
    RTC_TRIANGLE = 5,
    RTC_SMOOTH_TRIANGLE = 6,
    RTC_TEST_SHAPE = 7 /* never intersects (shape.zig:404-439) */
};
enum rtc_leaf_flags { RTC_CASTS_SHADOW = 1u, RTC_CLOSED = 2u, RTC_UNBOUNDED = 4u };

/* Pattern.Variant (patterns/pattern.zig:32-43) */
enum rtc_pattern_kind {
    RTC_PAT_SOLID = 0,
    RTC_PAT_BLEND = 1,
    RTC_PAT_PERTURB = 2,
    RTC_PAT_STRIPES = 3,
    RTC_PAT_GRADIENT = 4,
    RTC_PAT_RADIAL_GRADIENT = 5,
    RTC_PAT_RINGS = 6,
    RTC_PAT_CHECKERS = 7,
    RTC_PAT_TEXTURE_MAP = 8,
    RTC_PAT_TEST = 9
};
enum rtc_mapping { RTC_MAP_SPHERICAL = 0, RTC_MAP_PLANAR = 1, RTC_MAP_CYLINDRICAL = 2, RTC_MAP_CUBIC = 3 };
enum rtc_uv_kind { RTC_UV_ALIGN_CHECK = 0, RTC_UV_CHECKERS = 1, RTC_UV_IMAGE = 2, RTC_UV_TEST = 3 };
enum rtc_image_format { RTC_IMG_RGBA8 = 0 /* zigimg path */, RTC_IMG_RGB_F64 = 1 /* Canvas.fromPpm */ };
enum rtc_csg_op { RTC_CSG_UNION = 0, RTC_CSG_INTERSECTION = 1, RTC_CSG_DIFFERENCE = 2 };
enum rtc_node_kind { RTC_NODE_LEAF = 0, RTC_NODE_GROUP = 1, RTC_NODE_CSG = 2 };

typedef struct rtc_light { /* light.zig:10-23 */
    double position[3];
    double intensity[3];
} rtc_light;

typedef struct rtc_material { /* material.zig:14-25 */
    double ambient, diffuse, specular, shininess, reflective, transparency, refractive_index;
    int32_t pattern; /* root pattern record */
    int32_t pad;
} rtc_material;

typedef struct rtc_pattern {
    int32_t kind;     /* rtc_pattern_kind */
    int32_t a, b;     /* child pattern records (-1 = none) */
    int32_t mapping;  /* rtc_mapping, texture maps only */
    int32_t uv[6];    /* uv pattern records: [0] or front,back,left,right,up,down (texture_map.zig:216) */
    int32_t octaves;  /* perturb.zig:14-18 */
    int32_t pad;
    double inverse[12]; /* rows 0..2 of the pattern's inverse transform (row 3 is 0,0,0,1) */
    double color[3];    /* solid */
    double perturb_scale, perturb_persistence;
    double pad2;
} rtc_pattern;

typedef struct rtc_uv_pattern { /* texture_map.zig:21-105 */
    int32_t kind;               /* rtc_uv_kind */
    int32_t image;              /* image record */
    int32_t bilinear;
    int32_t sub[5];             /* align-check: central,ul,ur,bl,br ; checkers: a,b */
    double width, height;
} rtc_uv_pattern;

typedef struct rtc_image {
    int32_t width, height;
    int32_t format; /* rtc_image_format */
    int32_t pad;
    uint64_t offset; /* bytes from the start of the texel section, 16-byte aligned */
    uint64_t pad2;
} rtc_image;

typedef struct rtc_transform {
    double inverse[12]; /* rows 0..2 of Shape._inverse_transform (row 3 is 0,0,0,1 for every
                           transform the scene schema can express, scene.zig:214-241) */
} rtc_transform;

typedef struct rtc_leaf {
    int32_t kind;      /* rtc_leaf_kind */
    int32_t transform; /* rtc_transform record */
    int32_t material;  /* rtc_material record */
    uint32_t flags;    /* rtc_leaf_flags */
    double minimum, maximum; /* cylinder.zig:26-28, cone.zig:26-28 */
    int32_t triangle;  /* rtc_triangle record or -1 */
    int32_t guard;     /* first rtc_guard of the chain or -1 */
    double wmin[3], wmax[3]; /* world-space bounds = parentSpaceBounds (shape.zig:364-370);
                                ignored when RTC_UNBOUNDED is set */
    double pad;
} rtc_leaf;

typedef struct rtc_triangle { /* triangle.zig:22-27, :216-223 ; object space */
    double p1[3], e1[3], e2[3];
    int32_t leaf; /* back reference = tie-break order */
    int32_t pad;
} rtc_triangle;

typedef struct rtc_tri_normals { /* flat: n[0..2] = normal ; smooth: n1,n2,n3 */
    double n[9];
    double pad;
} rtc_tri_normals;

typedef struct rtc_guard {
    double bmin[3], bmax[3];
    int32_t next; /* enclosing guard or -1 */
    int32_t pad;
} rtc_guard;

/* Reference tree below (and including) each top-level CSG shape. */
typedef struct rtc_node {
    int32_t kind;  /* rtc_node_kind */
    int32_t op;    /* rtc_csg_op (CSG) */
    int32_t a, b;  /* LEAF: a = leaf ; CSG: a = left node, b = right node ;
                      GROUP: a = first entry in the child-index section, b = child count */
    int32_t leaf_begin, leaf_end; /* leaves of this subtree = [leaf_begin, leaf_end) */
    int32_t use_bbox; /* CSG: always 1 ; GROUP: 1 when its box is not a superset of its members */
    int32_t pad;
    double bmin[3], bmax[3]; /* `_bbox` as the reference holds it at render time */
} rtc_node;

typedef struct rtc_blob_header {
    uint32_t magic, version;
    uint64_t total_bytes;
    uint32_t n_lights, n_materials, n_patterns, n_uv_patterns, n_images, n_transforms;
    uint32_t n_leaves, n_triangles, n_guards, n_nodes, n_children, n_csg_roots;
    uint64_t texel_bytes;
    uint64_t off_lights, off_materials, off_patterns, off_uv_patterns, off_images, off_transforms;
    uint64_t off_leaves, off_triangles, off_tri_normals, off_guards, off_nodes, off_children;
    uint64_t off_csg_roots; /* int32 node index per top-level CSG object */
    uint64_t off_texels;
} rtc_blob_header;

#ifdef __cplusplus
}
#endif
#endif
