// World tree -> blob.  See flatten.hpp and include/rtc_blob.h for the format and the rationale.
#include "flatten.hpp"

#include <cstring>
#include <map>
#include <unordered_map>
#include <stdexcept>

namespace rtc {
namespace {

struct Builder {
    std::vector<rtc_light> lights;
    std::vector<rtc_material> materials;
    std::vector<rtc_pattern> patterns;
    std::vector<rtc_uv_pattern> uvs;
    std::vector<rtc_image> images;
    std::vector<rtc_transform> transforms;
    std::vector<rtc_leaf> leaves;
    std::vector<rtc_triangle> triangles;
    std::vector<rtc_tri_normals> normals;
    std::vector<rtc_guard> guards;
    std::vector<rtc_node> nodes;
    std::vector<int32_t> children;
    std::vector<int32_t> csg_roots;
    std::vector<uint8_t> texels;

    std::map<const Pattern*, int32_t> pattern_ids;
    std::map<const Image*, int32_t> image_ids;
    std::map<std::vector<uint64_t>, int32_t> transform_ids;
    std::map<std::vector<uint64_t>, int32_t> material_ids;

    static uint64_t bits(double d) {
        uint64_t u;
        std::memcpy(&u, &d, 8);
        return u;
    }

    static void rows3(const Mat4& m, double out[12]) {
        // Row 3 of every matrix the schema can build is (0,0,0,1) up to the sign of zero; the
        // device relies on it (3x4 storage), so check instead of assuming.
        if (m.d[3][0] != 0.0 || m.d[3][1] != 0.0 || m.d[3][2] != 0.0 || m.d[3][3] != 1.0)
            throw std::runtime_error("flatten: non-affine transform (row 3 != 0,0,0,1)");
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 4; c++) out[r * 4 + c] = m.d[r][c];
    }

    int32_t add_image(const std::shared_ptr<const Image>& im) {
        auto it = image_ids.find(im.get());
        if (it != image_ids.end()) return it->second;
        rtc_image rec{};
        rec.width = im->width;
        rec.height = im->height;
        while (texels.size() % 16) texels.push_back(0);
        rec.offset = texels.size();
        size_t px = (size_t)im->width * im->height;
        if (!im->rgbf.empty()) {
            rec.format = RTC_IMG_RGB_F64;
            size_t at = texels.size();
            texels.resize(at + px * 3 * sizeof(double));
            std::memcpy(&texels[at], im->rgbf.data(), px * 3 * sizeof(double));
        } else {
            rec.format = RTC_IMG_RGBA8;
            size_t at = texels.size();
            texels.resize(at + px * 4);
            const uint8_t* src = im->rgb8.data();
            uint8_t* dst = texels.data() + at;  // (16-byte aligned offset of a vector's storage: whole words)
            for (size_t i = 0; i < px; i++) {
                const uint32_t v = (uint32_t)src[i * 3] | ((uint32_t)src[i * 3 + 1] << 8) | ((uint32_t)src[i * 3 + 2] << 16) | 0xff000000u;
                std::memcpy(dst + i * 4, &v, 4);
            }
        }
        int32_t id = (int32_t)images.size();
        images.push_back(rec);
        image_ids[im.get()] = id;
        return id;
    }

    int32_t add_pattern(const PatternP& p);

    int32_t add_uv(const UvPattern& uv) {
        rtc_uv_pattern rec{};
        rec.kind = uv.kind == UvPattern::AlignCheck ? RTC_UV_ALIGN_CHECK
                   : uv.kind == UvPattern::Checkers ? RTC_UV_CHECKERS
                   : uv.kind == UvPattern::ImageMap ? RTC_UV_IMAGE
                                                    : RTC_UV_TEST;
        rec.image = -1;
        rec.bilinear = uv.bilinear ? 1 : 0;
        rec.width = uv.width;
        rec.height = uv.height;
        for (int i = 0; i < 5; i++) rec.sub[i] = uv.sub[i] ? add_pattern(uv.sub[i]) : -1;
        if (uv.kind == UvPattern::ImageMap) {
            if (!uv.image || uv.image->width <= 0) throw std::runtime_error("flatten: uv image missing");
            rec.image = add_image(uv.image);
        }
        uvs.push_back(rec);
        return (int32_t)uvs.size() - 1;
    }

    // (the leaves of a mesh come one after the other with the same matrix and the same material: the last answer is
    // kept, so that 140,940 triangles do not mean 281,880 heap-allocated keys and map look-ups)
    rtc_transform last_transform{};
    int32_t last_transform_id = -1;
    Material last_material;
    int32_t last_material_id = -1;

    int32_t add_transform(const Mat4& inverse) {
        rtc_transform rec{};
        rows3(inverse, rec.inverse);
        if (last_transform_id >= 0 && std::memcmp(rec.inverse, last_transform.inverse, sizeof rec.inverse) == 0) return last_transform_id;
        const int32_t found = find_transform(rec);
        last_transform = rec;
        last_transform_id = found;
        return found;
    }
    int32_t find_transform(const rtc_transform& rec) {
        std::vector<uint64_t> key(12);
        for (int i = 0; i < 12; i++) key[i] = bits(rec.inverse[i]);
        auto it = transform_ids.find(key);
        if (it != transform_ids.end()) return it->second;
        int32_t id = (int32_t)transforms.size();
        transforms.push_back(rec);
        transform_ids[key] = id;
        return id;
    }

    int32_t add_material(const Material& m) {
        if (last_material_id >= 0 && m.pattern.get() == last_material.pattern.get() && bits(m.ambient) == bits(last_material.ambient) &&
            bits(m.diffuse) == bits(last_material.diffuse) && bits(m.specular) == bits(last_material.specular) &&
            bits(m.shininess) == bits(last_material.shininess) && bits(m.reflective) == bits(last_material.reflective) &&
            bits(m.transparency) == bits(last_material.transparency) && bits(m.refractive_index) == bits(last_material.refractive_index))
            return last_material_id;
        const int32_t found = find_material(m);
        last_material = m;
        last_material_id = found;
        return found;
    }
    int32_t find_material(const Material& m) {
        int32_t pat = add_pattern(m.pattern);
        std::vector<uint64_t> key = {bits(m.ambient),      bits(m.diffuse),      bits(m.specular),
                                     bits(m.shininess),    bits(m.reflective),   bits(m.transparency),
                                     bits(m.refractive_index), (uint64_t)pat};
        auto it = material_ids.find(key);
        if (it != material_ids.end()) return it->second;
        rtc_material rec{};
        rec.ambient = m.ambient;
        rec.diffuse = m.diffuse;
        rec.specular = m.specular;
        rec.shininess = m.shininess;
        rec.reflective = m.reflective;
        rec.transparency = m.transparency;
        rec.refractive_index = m.refractive_index;
        rec.pattern = pat;
        int32_t id = (int32_t)materials.size();
        materials.push_back(rec);
        material_ids[key] = id;
        return id;
    }
};

int32_t Builder::add_pattern(const PatternP& p) {
    auto it = pattern_ids.find(p.get());
    if (it != pattern_ids.end()) return it->second;
    rtc_pattern rec{};
    switch (p->kind) {
        case Pattern::Solid: rec.kind = RTC_PAT_SOLID; break;
        case Pattern::Blend: rec.kind = RTC_PAT_BLEND; break;
        case Pattern::Perturb: rec.kind = RTC_PAT_PERTURB; break;
        case Pattern::Stripes: rec.kind = RTC_PAT_STRIPES; break;
        case Pattern::Gradient: rec.kind = RTC_PAT_GRADIENT; break;
        case Pattern::RadialGradient: rec.kind = RTC_PAT_RADIAL_GRADIENT; break;
        case Pattern::Rings: rec.kind = RTC_PAT_RINGS; break;
        case Pattern::Checkers: rec.kind = RTC_PAT_CHECKERS; break;
        case Pattern::TextureMap: rec.kind = RTC_PAT_TEXTURE_MAP; break;
        case Pattern::Test: rec.kind = RTC_PAT_TEST; break;
    }
    rec.a = p->a ? add_pattern(p->a) : -1;
    rec.b = p->b ? add_pattern(p->b) : -1;
    rec.mapping = (int32_t)p->mapping;
    for (int i = 0; i < 6; i++) rec.uv[i] = -1;
    if (p->kind == Pattern::TextureMap) {
        int n = p->mapping == Pattern::Cubic ? 6 : 1;
        for (int i = 0; i < n; i++) rec.uv[i] = add_uv(p->uv[i]);
    }
    rec.octaves = p->perturb_octaves;
    rows3(p->inverse, rec.inverse);
    rec.color[0] = p->color.r;
    rec.color[1] = p->color.g;
    rec.color[2] = p->color.b;
    rec.perturb_scale = p->perturb_scale;
    rec.perturb_persistence = p->perturb_persistence;
    int32_t id = (int32_t)patterns.size();
    patterns.push_back(rec);
    pattern_ids[p.get()] = id;
    return id;
}

bool finite3(const Tuple& t) { return std::isfinite(t.x) && std::isfinite(t.y) && std::isfinite(t.z); }

// Object-space bounds are finite?  (planes, untruncated cylinders / cones are not)
bool leaf_bounded(const Shape& s) {
    BBox b = s.bounds();
    return finite3(b.min) && finite3(b.max);
}

bool always_pass(const BBox& b) {
    // Per axis, the untouched extent (+inf,-inf) and the all-covering one (-inf,+inf) both come out
    // of the reference slab test (bounding_box.zig:112-136) as (-inf,+inf): such a box never culls.
    auto axis = [](double lo, double hi) { return (lo == kInf && hi == -kInf) || (lo == -kInf && hi == kInf); };
    return axis(b.min.x, b.max.x) && axis(b.min.y, b.max.y) && axis(b.min.z, b.max.z);
}

// Is `box` a superset of the true extent of every leaf below `s`?  Per leaf the test is
//   le(box.min, w.min) && le(w.max, box.max) on every axis,  le(a, b) = a <= b + 1e-9 (|a| + |b| + 1e-300)
// (box-of-box vs direct transform differ by rounding only; a few ulps of slack), w = the leaf's parent-space bounds.
// le is monotone in either argument (b + 1e-9 |b| grows with b, a - 1e-9 |a| with a, and rounding keeps order), so
// "for every leaf" equals the same test on the component-wise min of the w.min and max of the w.max of the subtree:
// the extents are gathered ONCE per node, bottom-up (a dragon's 23,490 triangles sit under ~8,000 groups 13 deep; asking
// every group for all its leaves again was 0.5 s of the 0.75 s flattening of dragons.json).
struct Extent {
    bool ok = true;   // every leaf is bounded with finite bounds
    size_t leaves = 0;
    double mn[3] = {kInf, kInf, kInf}, mx[3] = {-kInf, -kInf, -kInf};
    void merge(const Extent& e) {
        ok = ok && e.ok;
        leaves += e.leaves;
        for (int k = 0; k < 3; k++) {
            mn[k] = std::fmin(mn[k], e.mn[k]);
            mx[k] = std::fmax(mx[k], e.mx[k]);
        }
    }
};
thread_local std::unordered_map<const Shape*, Extent> g_extent;  // groups and CSG nodes of the world being flattened

Extent extent_of(const Shape& s) {
    if (s.kind == ShapeKind::Group || s.kind == ShapeKind::Csg) {
        auto it = g_extent.find(&s);
        if (it != g_extent.end()) return it->second;
        Extent e;
        if (s.kind == ShapeKind::Group) {
            for (const Shape& c : s.children) e.merge(extent_of(c));
        } else {
            e.merge(extent_of(*s.left));
            e.merge(extent_of(*s.right));
        }
        g_extent.emplace(&s, e);
        return e;
    }
    Extent e;
    if (s.kind == ShapeKind::Test) return e;
    e.leaves = 1;
    if (!leaf_bounded(s)) {
        e.ok = false;
        return e;
    }
    BBox w = s.parent_space_bounds();
    if (!finite3(w.min) || !finite3(w.max)) {
        e.ok = false;
        return e;
    }
    e.mn[0] = w.min.x, e.mn[1] = w.min.y, e.mn[2] = w.min.z;
    e.mx[0] = w.max.x, e.mx[1] = w.max.y, e.mx[2] = w.max.z;
    return e;
}

bool covers(const BBox& box, const Shape& s) {
    const Extent e = extent_of(s);
    if (!e.ok) return false;
    if (e.leaves == 0) return true;
    auto le = [](double a, double b) { return a <= b + 1e-9 * (std::fabs(a) + std::fabs(b) + 1e-300); };
    return le(box.min.x, e.mn[0]) && le(box.min.y, e.mn[1]) && le(box.min.z, e.mn[2]) && le(e.mx[0], box.max.x) &&
           le(e.mx[1], box.max.y) && le(e.mx[2], box.max.z);
}

struct Walker {
    Builder& b;
    int32_t guard = -1;  // innermost active guard

    int32_t emit_leaf(const Shape& s) {
        rtc_leaf rec{};
        switch (s.kind) {
            case ShapeKind::Sphere: rec.kind = RTC_SPHERE; break;
            case ShapeKind::Plane: rec.kind = RTC_PLANE; break;
            case ShapeKind::Cube: rec.kind = RTC_CUBE; break;
            case ShapeKind::Cylinder: rec.kind = RTC_CYLINDER; break;
            case ShapeKind::Cone: rec.kind = RTC_CONE; break;
            case ShapeKind::Triangle: rec.kind = RTC_TRIANGLE; break;
            case ShapeKind::SmoothTriangle: rec.kind = RTC_SMOOTH_TRIANGLE; break;
            case ShapeKind::Test: rec.kind = RTC_TEST_SHAPE; break;
            default: throw std::runtime_error("flatten: not a leaf");
        }
        rec.transform = b.add_transform(s.inverse);
        rec.material = b.add_material(s.material);
        rec.flags = (s.casts_shadow ? RTC_CASTS_SHADOW : 0u) | (s.closed ? RTC_CLOSED : 0u);
        rec.minimum = s.minimum;
        rec.maximum = s.maximum;
        rec.triangle = -1;
        rec.guard = guard;
        int32_t id = (int32_t)b.leaves.size();
        BBox w = s.parent_space_bounds();
        if (leaf_bounded(s) && finite3(w.min) && finite3(w.max)) {
            rec.wmin[0] = w.min.x, rec.wmin[1] = w.min.y, rec.wmin[2] = w.min.z;
            rec.wmax[0] = w.max.x, rec.wmax[1] = w.max.y, rec.wmax[2] = w.max.z;
        } else {
            rec.flags |= RTC_UNBOUNDED;
            for (int i = 0; i < 3; i++) rec.wmin[i] = -kInf, rec.wmax[i] = kInf;
        }
        if (s.kind == ShapeKind::Triangle || s.kind == ShapeKind::SmoothTriangle) {
            rtc_triangle t{};
            const Tuple* src[3] = {&s.p1, &s.e1, &s.e2};
            double* dst[3] = {t.p1, t.e1, t.e2};
            for (int k = 0; k < 3; k++) dst[k][0] = src[k]->x, dst[k][1] = src[k]->y, dst[k][2] = src[k]->z;
            t.leaf = id;
            rtc_tri_normals n{};
            if (s.kind == ShapeKind::Triangle) {
                n.n[0] = s.normal.x, n.n[1] = s.normal.y, n.n[2] = s.normal.z;
            } else {
                const Tuple* ns[3] = {&s.n1, &s.n2, &s.n3};
                for (int k = 0; k < 3; k++) n.n[k * 3] = ns[k]->x, n.n[k * 3 + 1] = ns[k]->y, n.n[k * 3 + 2] = ns[k]->z;
            }
            rec.triangle = (int32_t)b.triangles.size();
            b.triangles.push_back(t);
            b.normals.push_back(n);
        }
        b.leaves.push_back(rec);
        return id;
    }

    static void put_box(const BBox& box, double mn[3], double mx[3]) {
        mn[0] = box.min.x, mn[1] = box.min.y, mn[2] = box.min.z;
        mx[0] = box.max.x, mx[1] = box.max.y, mx[2] = box.max.z;
    }

    // Outside CSG: groups dissolve, non-superset boxes become guards.
    void walk(const Shape& s) {
        if (s.kind == ShapeKind::Group) {
            int32_t saved = guard;
            if (!always_pass(s.bbox) && !covers(s.bbox, s)) {
                rtc_guard g{};
                put_box(s.bbox, g.bmin, g.bmax);
                g.next = guard;
                b.guards.push_back(g);
                guard = (int32_t)b.guards.size() - 1;
            }
            for (const Shape& c : s.children) walk(c);
            guard = saved;
        } else if (s.kind == ShapeKind::Csg) {
            b.csg_roots.push_back(node(s));
        } else {
            emit_leaf(s);
        }
    }

    // At and below a CSG: keep the tree.
    int32_t node(const Shape& s) {
        rtc_node n{};
        n.leaf_begin = (int32_t)b.leaves.size();
        if (s.kind == ShapeKind::Group) {
            n.kind = RTC_NODE_GROUP;
            n.use_bbox = (!always_pass(s.bbox) && !covers(s.bbox, s)) ? 1 : 0;
            put_box(s.bbox, n.bmin, n.bmax);
            std::vector<int32_t> kids;
            for (const Shape& c : s.children) kids.push_back(node(c));
            n.a = (int32_t)b.children.size();
            n.b = (int32_t)kids.size();
            b.children.insert(b.children.end(), kids.begin(), kids.end());
        } else if (s.kind == ShapeKind::Csg) {
            n.kind = RTC_NODE_CSG;
            n.op = s.op == CsgOp::Union ? RTC_CSG_UNION : s.op == CsgOp::Intersection ? RTC_CSG_INTERSECTION : RTC_CSG_DIFFERENCE;
            n.use_bbox = always_pass(s.bbox) ? 0 : 1;
            put_box(s.bbox, n.bmin, n.bmax);
            n.a = node(*s.left);
            n.b = node(*s.right);
        } else {
            n.kind = RTC_NODE_LEAF;
            n.a = emit_leaf(s);
        }
        n.leaf_end = (int32_t)b.leaves.size();
        b.nodes.push_back(n);
        return (int32_t)b.nodes.size() - 1;
    }
};

template <class T>
uint64_t put(std::vector<uint8_t>& out, const std::vector<T>& v) {
    while (out.size() % 16) out.push_back(0);
    uint64_t off = out.size();
    if (!v.empty()) {
        out.resize(off + v.size() * sizeof(T));
        std::memcpy(&out[off], v.data(), v.size() * sizeof(T));
    }
    return off;
}

}  // namespace

std::vector<uint8_t> flatten_world(const World& world) {
    Builder b;
    for (const Light& l : world.lights) {
        rtc_light rec{};
        rec.position[0] = l.position.x, rec.position[1] = l.position.y, rec.position[2] = l.position.z;
        rec.intensity[0] = l.intensity.r, rec.intensity[1] = l.intensity.g, rec.intensity[2] = l.intensity.b;
        b.lights.push_back(rec);
    }
    Walker w{b};
    g_extent.clear();
    {   // how many leaves there will be (the extents are wanted anyway and are kept): no regrowing of 100,000-record vectors
        size_t n_leaves = 0;
        for (const Shape& s : world.objects) n_leaves += extent_of(s).leaves;
        b.leaves.reserve(n_leaves);
        b.triangles.reserve(n_leaves);
        b.normals.reserve(n_leaves);
    }
    for (const Shape& s : world.objects) w.walk(s);
    g_extent.clear();

    rtc_blob_header h{};
    h.magic = RTC_BLOB_MAGIC;
    h.version = RTC_BLOB_VERSION;
    h.n_lights = (uint32_t)b.lights.size();
    h.n_materials = (uint32_t)b.materials.size();
    h.n_patterns = (uint32_t)b.patterns.size();
    h.n_uv_patterns = (uint32_t)b.uvs.size();
    h.n_images = (uint32_t)b.images.size();
    h.n_transforms = (uint32_t)b.transforms.size();
    h.n_leaves = (uint32_t)b.leaves.size();
    h.n_triangles = (uint32_t)b.triangles.size();
    h.n_guards = (uint32_t)b.guards.size();
    h.n_nodes = (uint32_t)b.nodes.size();
    h.n_children = (uint32_t)b.children.size();
    h.n_csg_roots = (uint32_t)b.csg_roots.size();
    h.texel_bytes = b.texels.size();

    std::vector<uint8_t> out;
    out.reserve(sizeof(rtc_blob_header) + 16 * 16 + b.lights.size() * sizeof(rtc_light) + b.materials.size() * sizeof(rtc_material) +
                b.patterns.size() * sizeof(rtc_pattern) + b.uvs.size() * sizeof(rtc_uv_pattern) + b.images.size() * sizeof(rtc_image) +
                b.transforms.size() * sizeof(rtc_transform) + b.leaves.size() * sizeof(rtc_leaf) + b.triangles.size() * sizeof(rtc_triangle) +
                b.normals.size() * sizeof(rtc_tri_normals) + b.guards.size() * sizeof(rtc_guard) + b.nodes.size() * sizeof(rtc_node) +
                b.children.size() * sizeof(int32_t) + b.csg_roots.size() * sizeof(int32_t) + b.texels.size());
    out.resize(sizeof(rtc_blob_header));
    h.off_lights = put(out, b.lights);
    h.off_materials = put(out, b.materials);
    h.off_patterns = put(out, b.patterns);
    h.off_uv_patterns = put(out, b.uvs);
    h.off_images = put(out, b.images);
    h.off_transforms = put(out, b.transforms);
    h.off_leaves = put(out, b.leaves);
    h.off_triangles = put(out, b.triangles);
    h.off_tri_normals = put(out, b.normals);
    h.off_guards = put(out, b.guards);
    h.off_nodes = put(out, b.nodes);
    h.off_children = put(out, b.children);
    h.off_csg_roots = put(out, b.csg_roots);
    h.off_texels = put(out, b.texels);
    while (out.size() % 16) out.push_back(0);
    h.total_bytes = out.size();
    std::memcpy(out.data(), &h, sizeof h);
    return out;
}

rtc_camera flatten_camera(const Camera& cam) {
    rtc_camera c{};
    c.hsize = cam.hsize;
    c.vsize = cam.vsize;
    c.half_width = cam.half_width;
    c.half_height = cam.half_height;
    c.pixel_size = cam.pixel_size;
    for (int r = 0; r < 4; r++)
        for (int col = 0; col < 4; col++) c.inverse[r * 4 + col] = cam.inverse.d[r][col];
    return c;
}

}  // namespace rtc
