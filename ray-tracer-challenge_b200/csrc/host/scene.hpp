// Host scene model: the World/Shape/Material/Pattern tree exactly as the reference's parsing layer
// hands it to Camera.render (src/raytracer/camera.zig:80). It is DATA plus the build-time
// operations (setTransform push-down, bounds, divide(8)); no ray is ever traced here.
// The device flattener (flatten.cpp) consumes it; the oracle (oracle/) traces it on the CPU.
//
//   Shape / setTransform / divide      src/raytracer/shapes/shape.zig:99-131, 286-310, 372-399
//   BoundingBox                        src/raytracer/shapes/bounding_box.zig:17-110
//   Group addChild/partition/subgroup  src/raytracer/shapes/group.zig:75-135
//   Csg construction                   src/raytracer/shapes/shape.zig:254-281
//   Material / Light / Pattern / Uv    src/raytracer/material.zig:14-25, light.zig:10-23,
//                                      patterns/pattern.zig:21-110, patterns/texture_map.zig
//   Camera.new / setTransform          src/raytracer/camera.zig:33-61
#pragma once
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "rtc_math.hpp"

namespace rtc {

constexpr double kInf = std::numeric_limits<double>::infinity();

// ---- images (Canvas used as a texture, src/raytracer/canvas.zig:18-46) ---------------------------
struct Image {
    int width = 0, height = 0;
    std::vector<uint8_t> rgb8;   // PNG-sourced (zigimg path): 3 bytes per texel
    std::vector<double> rgbf;    // PPM-sourced (Canvas.fromPpm) : 3 doubles per texel
    std::string name;
    // zigimg hands out f32 colour channels (u8/255 as f32) which Canvas.fromImage widens to T
    // (canvas.zig:39-42).  See DESIGN.md "third-party arithmetic": parity unpinned for this step.
    Color texel(int x, int y) const {
        size_t i = ((size_t)y * width + x) * 3;
        if (!rgbf.empty()) return {rgbf[i], rgbf[i + 1], rgbf[i + 2]};
        return {(double)((float)rgb8[i] / 255.0f), (double)((float)rgb8[i + 1] / 255.0f),
                (double)((float)rgb8[i + 2] / 255.0f)};
    }
};

// ---- patterns ------------------------------------------------------------------------------------
struct Pattern;
using PatternP = std::shared_ptr<const Pattern>;

struct UvPattern {
    enum Kind { AlignCheck = 0, Checkers = 1, ImageMap = 2, Test = 3 };
    Kind kind = Test;
    double width = 1, height = 1;      // uv checkers
    PatternP sub[5];                   // align-check: central, ul, ur, bl, br ; checkers: a, b
    std::shared_ptr<const Image> image;
    bool bilinear = false;
};

struct Pattern {
    enum Kind { Solid = 0, Blend, Perturb, Stripes, Gradient, RadialGradient, Rings, Checkers, TextureMap, Test };
    enum Mapping { Spherical = 0, Planar, Cylindrical, Cubic };
    Kind kind = Solid;
    Mat4 transform = Mat4::identity();
    Mat4 inverse = Mat4::identity();
    Color color{1, 1, 1};              // solid
    PatternP a, b;                     // higher-order children
    double perturb_scale = 0.3;        // perturb.zig:14-18
    int perturb_octaves = 3;
    double perturb_persistence = 0.8;
    Mapping mapping = Spherical;       // texture map
    UvPattern uv[6];                   // [0] for non-cubic; cubic order front,back,left,right,up,down

    static Pattern solid(Color c) {
        Pattern p;
        p.kind = Solid;
        p.color = c;
        return p;
    }
    static Pattern binary(Kind k, PatternP a, PatternP b) {
        Pattern p;
        p.kind = k;
        p.a = std::move(a);
        p.b = std::move(b);
        return p;
    }
    void set_transform(const Mat4& m) {
        transform = m;
        inverse = m.inverse();
    }
};
inline PatternP share(Pattern p) { return std::make_shared<const Pattern>(std::move(p)); }

struct Material {
    // (one immutable default pattern for every Material: a scene constructs hundreds of thousands of Materials -- one per
    // triangle, overwritten when the object's material is stamped on -- and allocating a pattern for each was a tenth of
    // the load time)
    static const PatternP& default_pattern() {
        static const PatternP p = share(Pattern::solid({1, 1, 1}));
        return p;
    }
    PatternP pattern = default_pattern();
    double ambient = 0.1, diffuse = 0.9, specular = 0.9, shininess = 200.0;
    double reflective = 0.0, transparency = 0.0, refractive_index = 1.0;
};

struct Light {
    Tuple position;
    Color intensity;
};

// ---- bounding boxes ------------------------------------------------------------------------------
struct BBox {
    Tuple min = Tuple::point(kInf, kInf, kInf);
    Tuple max = Tuple::point(-kInf, -kInf, -kInf);
    // Zig @min/@max return the non-NaN operand -> fmin/fmax (bounding_box.zig:24-32)
    void add(const Tuple& p) {
        min.x = std::fmin(min.x, p.x);
        min.y = std::fmin(min.y, p.y);
        min.z = std::fmin(min.z, p.z);
        max.x = std::fmax(max.x, p.x);
        max.y = std::fmax(max.y, p.y);
        max.z = std::fmax(max.z, p.z);
    }
    bool contains_point(const Tuple& p) const {
        return min.x <= p.x && p.x <= max.x && min.y <= p.y && p.y <= max.y && min.z <= p.z && p.z <= max.z;
    }
    bool contains_box(const BBox& o) const { return contains_point(o.min) && contains_point(o.max); }
    void merge(const BBox& o) {
        add(o.min);
        add(o.max);
    }
    BBox transform(const Mat4& m) const;              // 8 corners (bounding_box.zig:49-70)
    void split(BBox& left, BBox& right) const;        // bounding_box.zig:72-110
};

// ---- shapes --------------------------------------------------------------------------------------
enum class ShapeKind { Test, Sphere, Cube, Cylinder, Cone, Plane, Triangle, SmoothTriangle, Group, Csg };
enum class CsgOp { Union, Intersection, Difference };

struct Shape {
    size_t id;
    ShapeKind kind;
    Mat4 transform = Mat4::identity();
    Mat4 inverse = Mat4::identity();
    Mat4 inverse_transpose = Mat4::identity();
    Material material;
    bool casts_shadow = true;
    // cylinder / cone
    double minimum = -kInf, maximum = kInf;
    bool closed = false;
    // triangles
    Tuple p1, p2, p3, e1, e2, normal, n1, n2, n3;
    // group
    std::vector<Shape> children;
    // csg
    std::shared_ptr<Shape> left, right;
    CsgOp op = CsgOp::Union;
    // group + csg : `_bbox`
    BBox bbox;
    // cache of parent_space_bounds() for triangles
    mutable BBox psb;
    mutable bool psb_valid = false;

    static size_t next_id();
    static Shape make(ShapeKind k) {
        Shape s;
        s.id = next_id();
        s.kind = k;
        return s;
    }
    static Shape sphere() { return make(ShapeKind::Sphere); }
    static Shape glass_sphere() {
        Shape s = sphere();
        s.material.transparency = 1.0;
        s.material.refractive_index = 1.5;
        return s;
    }
    static Shape test_shape() { return make(ShapeKind::Test); }
    static Shape cube() { return make(ShapeKind::Cube); }
    static Shape plane() { return make(ShapeKind::Plane); }
    static Shape cylinder() { return make(ShapeKind::Cylinder); }
    static Shape cone() { return make(ShapeKind::Cone); }
    static Shape triangle(const Tuple& p1, const Tuple& p2, const Tuple& p3);
    static Shape smooth_triangle(const Tuple& p1, const Tuple& p2, const Tuple& p3, const Tuple& n1,
                                 const Tuple& n2, const Tuple& n3);
    static Shape group();
    static Shape csg(CsgOp op, Shape left, Shape right);

    bool is_leaf() const { return kind != ShapeKind::Group && kind != ShapeKind::Csg; }
    void set_transform(const Mat4& m);       // shape.zig:286-310 (groups/CSG push down)
    BBox bounds() const;                     // object-space bounds per variant
    // shape.zig:364-370.  For triangles (the only kind a scene holds by the hundred thousand, and whose
    // bounds are asked for again at every level of divide()) the result is kept; set_transform resets it.
    BBox parent_space_bounds() const {
        if (kind != ShapeKind::Triangle && kind != ShapeKind::SmoothTriangle) return bounds().transform(transform);
        if (!psb_valid) {
            psb = bounds().transform(transform);
            psb_valid = true;
        }
        return psb;
    }
    void add_child(Shape child);             // group.zig:75-78
    void divide(size_t threshold);           // shape.zig:372-399
    void partition_children(std::vector<Shape>& left, std::vector<Shape>& right);  // group.zig:85-115
    void make_subgroup(std::vector<Shape>& members);                               // group.zig:117-135
    Tuple world_to_object(const Tuple& p) const { return inverse.tuple_mul(p); }
    Tuple normal_to_world(const Tuple& n) const {
        Tuple r = inverse_transpose.tuple_mul(n);
        r.w = 0.0;
        return r.normalized();
    }

   private:
    Shape() = default;
};

struct World {
    std::vector<Shape> objects;
    std::vector<Light> lights;
    static World default_world();            // world.zig:40-62
};

struct Camera {
    size_t hsize = 0, vsize = 0;
    double fov = 0, half_width = 0, half_height = 0, pixel_size = 0;
    Tuple from, to, up;                      // _saved_from_to_up (lib.zig camera orbit)
    Mat4 transform = Mat4::identity();
    Mat4 inverse = Mat4::identity();
    static Camera make(size_t hsize, size_t vsize, double fov);  // camera.zig:33-52
    void set_transform(const Mat4& m) {
        transform = m;
        inverse = m.inverse();
    }
};

struct SceneInfo {
    Camera camera;
    World world;
};

}  // namespace rtc
