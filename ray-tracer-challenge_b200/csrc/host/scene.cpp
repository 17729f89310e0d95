// Build-time operations on the host scene model (see scene.hpp for the reference map).
#include <cstdint>
#include <memory>
#include "scene.hpp"

#include <cstring>
#include <atomic>

namespace rtc {

size_t Shape::next_id() {  // shape.zig:122-131 : process-wide counter
    static std::atomic<size_t> counter{0};
    return counter.fetch_add(1);
}

// ---- BBox ----------------------------------------------------------------------------------------
BBox BBox::transform(const Mat4& m) const {
    // Corner order as in bounding_box.zig:50-57.  Infinite extents meet zero matrix entries as
    // inf*0 = NaN, and add() (fmin/fmax) then ignores that corner -- the reference quirk that can
    // leave the box "empty" (min=+inf,max=-inf), which its slab test always passes.
    const Tuple c[8] = {min,
                        Tuple::point(min.x, min.y, max.z),
                        Tuple::point(min.x, max.y, min.z),
                        Tuple::point(min.x, max.y, max.z),
                        Tuple::point(max.x, min.y, min.z),
                        Tuple::point(max.x, min.y, max.z),
                        Tuple::point(max.x, max.y, min.z),
                        max};
    BBox out;
    for (const Tuple& p : c) out.add(m.tuple_mul(p));
    return out;
}

void BBox::split(BBox& left, BBox& right) const {
    double dx = max.x - min.x, dy = max.y - min.y, dz = max.z - min.z;
    double greatest = std::fmax(dx, std::fmax(dy, dz));
    double x0 = min.x, y0 = min.y, z0 = min.z, x1 = max.x, y1 = max.y, z1 = max.z;
    if (greatest == dx) {
        x0 = x0 + dx / 2.0;
        x1 = x0;
    } else if (greatest == dy) {
        y0 = y0 + dy / 2.0;
        y1 = y0;
    } else {
        z0 = z0 + dz / 2.0;
        z1 = z0;
    }
    left.min = min;
    left.max = Tuple::point(x1, y1, z1);
    right.min = Tuple::point(x0, y0, z0);
    right.max = max;
}

// ---- Shape construction --------------------------------------------------------------------------
Shape Shape::triangle(const Tuple& p1, const Tuple& p2, const Tuple& p3) {  // shape.zig:186-206
    Shape s = make(ShapeKind::Triangle);
    s.p1 = p1;
    s.p2 = p2;
    s.p3 = p3;
    s.e1 = p2.sub(p1);
    s.e2 = p3.sub(p1);
    s.normal = s.e2.cross(s.e1).normalized();
    return s;
}

Shape Shape::smooth_triangle(const Tuple& p1, const Tuple& p2, const Tuple& p3, const Tuple& n1,
                             const Tuple& n2, const Tuple& n3) {  // shape.zig:208-229
    Shape s = make(ShapeKind::SmoothTriangle);
    s.p1 = p1;
    s.p2 = p2;
    s.p3 = p3;
    s.e1 = p2.sub(p1);
    s.e2 = p3.sub(p1);
    s.n1 = n1;
    s.n2 = n2;
    s.n3 = n3;
    return s;
}

Shape Shape::group() {  // shape.zig:237-252 : the bbox is itself a Shape there and takes an id
    (void)next_id();
    return make(ShapeKind::Group);
}

Shape Shape::csg(CsgOp op, Shape left, Shape right) {  // shape.zig:254-281
    BBox box = left.parent_space_bounds();
    box.merge(right.parent_space_bounds());
    Shape s = make(ShapeKind::Csg);
    s.bbox = box;
    s.op = op;
    s.left = std::shared_ptr<Shape>(new Shape(std::move(left)));
    s.right = std::shared_ptr<Shape>(new Shape(std::move(right)));
    return s;
}

void Shape::set_transform(const Mat4& m) {
    switch (kind) {
        case ShapeKind::Group:
            for (Shape& c : children) c.set_transform(m.mul(c.transform));
            bbox = bbox.transform(m);
            break;
        case ShapeKind::Csg:
            // NB: `_bbox` is NOT updated here in the reference (shape.zig:298-302).
            left->set_transform(m.mul(left->transform));
            right->set_transform(m.mul(right->transform));
            break;
        default: {
            transform = m;
            psb_valid = false;
            // The triangles of a mesh all receive the same matrix: same input, same inverse (matrix.zig:202-217
            // is deterministic), so the last result is reused instead of 140,940 x nesting-depth inversions.
            static thread_local Mat4 last_m, last_inv, last_inv_t;
            static thread_local bool have_last = false;
            if (!have_last || std::memcmp(&last_m, &m, sizeof(Mat4)) != 0) {
                last_inv = m.inverse();
                last_inv_t = last_inv.transpose();
                last_m = m;
                have_last = true;
            }
            inverse = last_inv;
            inverse_transpose = last_inv_t;
        }
    }
}

BBox Shape::bounds() const {
    BBox b;
    switch (kind) {
        case ShapeKind::Test:
        case ShapeKind::Sphere:
        case ShapeKind::Cube:
            b.min = Tuple::point(-1, -1, -1);
            b.max = Tuple::point(1, 1, 1);
            break;
        case ShapeKind::Plane:  // plane.zig:45-53
            b.min = Tuple::point(-kInf, 0, -kInf);
            b.max = Tuple::point(kInf, 0, kInf);
            break;
        case ShapeKind::Cylinder:  // cylinder.zig:114-120
            b.min = Tuple::point(-1, minimum, -1);
            b.max = Tuple::point(1, maximum, 1);
            break;
        case ShapeKind::Cone: {  // cone.zig:129-137
            double limit = std::fmax(std::fabs(minimum), std::fabs(maximum));
            b.min = Tuple::point(-limit, minimum, -limit);
            b.max = Tuple::point(limit, maximum, limit);
            break;
        }
        case ShapeKind::Triangle:
        case ShapeKind::SmoothTriangle:
            b.add(p1);
            b.add(p2);
            b.add(p3);
            break;
        case ShapeKind::Group:
        case ShapeKind::Csg:
            b = bbox;
            break;
    }
    return b;
}

void Shape::add_child(Shape child) {
    bbox.merge(child.parent_space_bounds());
    children.push_back(std::move(child));
}

void Shape::partition_children(std::vector<Shape>& left, std::vector<Shape>& right) {
    BBox lb, rb;
    bbox.split(lb, rb);
    // classify first (same tests, same order as group.zig:92-112), then move every child exactly once into a vector of
    // exactly its final size; a group whose children all straddle the split is left as it is
    std::vector<unsigned char> side(children.size());
    size_t n_left = 0, n_right = 0;
    for (size_t i = 0; i < children.size(); i++) {
        BBox cb = children[i].parent_space_bounds();
        side[i] = lb.contains_box(cb) ? 0 : rb.contains_box(cb) ? 1 : 2;
        n_left += side[i] == 0;
        n_right += side[i] == 1;
    }
    if (n_left == 0 && n_right == 0) return;
    std::vector<Shape> keep;
    left.reserve(n_left);
    right.reserve(n_right);
    keep.reserve(children.size() - n_left - n_right);
    for (size_t i = 0; i < children.size(); i++)
        (side[i] == 0 ? left : side[i] == 1 ? right : keep).push_back(std::move(children[i]));
    children = std::move(keep);
}

void Shape::make_subgroup(std::vector<Shape>& members) {
    // add_child (group.zig:75-78) for every member, in order: the box is merged in the same order (same bits);
    // the members themselves change hands as one vector instead of one Shape move at a time
    Shape sub = Shape::group();
    for (const Shape& m : members) sub.bbox.merge(m.parent_space_bounds());
    sub.children = std::move(members);
    members.clear();
    bbox.merge(sub.parent_space_bounds());
    children.push_back(std::move(sub));
}

// divide() of a whole subtree of NEW subgroups, planned on child indices first and built afterwards, so that a leaf
// changes hands once instead of once per level (a Shape is ~0.9 kB; a dragon's 23,490 triangles end up 13 levels deep:
// moving them level by level was half of the scene set-up time).  Same tests in the same order as the level-by-level
// form -- partition_children / make_subgroup above, which stay for the reference's own tests of them: a subgroup's box
// is the merge of its members' parent-space bounds in child order (its transform is the identity, so those bounds are
// the ones the parent saw), it is split only when it holds >= threshold members, and its children are the members
// that straddle the split, in order, followed by the left and the right subgroup.
namespace {
struct DivPlan {
    std::vector<uint32_t> keep;
    std::unique_ptr<DivPlan> left, right;
    BBox left_box, right_box;
};

std::unique_ptr<DivPlan> plan_divide(const std::vector<BBox>& pb, std::vector<uint32_t> members, const BBox& box, size_t threshold) {
    auto n = std::make_unique<DivPlan>();
    if (members.size() < threshold) {
        n->keep = std::move(members);
        return n;
    }
    BBox lb, rb;
    box.split(lb, rb);
    std::vector<uint32_t> l, r;
    l.reserve(members.size());
    r.reserve(members.size());
    for (uint32_t i : members) {
        if (lb.contains_box(pb[i])) l.push_back(i);
        else if (rb.contains_box(pb[i])) r.push_back(i);
        else n->keep.push_back(i);
    }
    if (!l.empty()) {
        for (uint32_t i : l) n->left_box.merge(pb[i]);
        n->left = plan_divide(pb, std::move(l), n->left_box, threshold);
    }
    if (!r.empty()) {
        for (uint32_t i : r) n->right_box.merge(pb[i]);
        n->right = plan_divide(pb, std::move(r), n->right_box, threshold);
    }
    return n;
}

void build_divided(Shape& g, const DivPlan& n, const BBox& lbox, const BBox& rbox, std::vector<Shape>& pool, size_t threshold);

void add_planned_subgroup(Shape& g, const DivPlan& sub_plan, const BBox& box, std::vector<Shape>& pool, size_t threshold) {
    Shape sub = Shape::group();
    sub.bbox = box;
    build_divided(sub, sub_plan, sub_plan.left_box, sub_plan.right_box, pool, threshold);
    g.bbox.merge(sub.parent_space_bounds());  // add_child (group.zig:75-78)
    g.children.push_back(std::move(sub));
}

void build_divided(Shape& g, const DivPlan& n, const BBox& lbox, const BBox& rbox, std::vector<Shape>& pool, size_t threshold) {
    g.children.reserve(n.keep.size() + (n.left ? 1 : 0) + (n.right ? 1 : 0));
    for (uint32_t i : n.keep) g.children.push_back(std::move(pool[i]));
    if (n.left) add_planned_subgroup(g, *n.left, lbox, pool, threshold);
    if (n.right) add_planned_subgroup(g, *n.right, rbox, pool, threshold);
    // the members that stayed may be groups or CSG objects of their own (the new subgroups are planned already)
    for (size_t k = 0; k < n.keep.size(); k++) g.children[k].divide(threshold);
}
}  // namespace

void Shape::divide(size_t threshold) {
    if (kind == ShapeKind::Group) {
        if (children.size() >= threshold) {
            std::vector<BBox> pb(children.size());
            std::vector<uint32_t> all(children.size());
            for (size_t i = 0; i < children.size(); i++) {
                pb[i] = children[i].parent_space_bounds();
                all[i] = (uint32_t)i;
            }
            std::unique_ptr<DivPlan> plan = plan_divide(pb, std::move(all), bbox, threshold);
            if (plan->left || plan->right) {
                std::vector<Shape> pool = std::move(children);
                children.clear();
                build_divided(*this, *plan, plan->left_box, plan->right_box, pool, threshold);
                return;
            }
        }
        for (Shape& c : children) c.divide(threshold);
    } else if (kind == ShapeKind::Csg) {
        left->divide(threshold);
        right->divide(threshold);
    }
}

// ---- World / Camera ------------------------------------------------------------------------------
World World::default_world() {
    World w;
    Shape s1 = Shape::sphere();
    Material m;
    m.pattern = share(Pattern::solid({0.8, 1.0, 0.6}));
    m.diffuse = 0.7;
    m.specular = 0.2;
    s1.material = m;
    Shape s2 = Shape::sphere();
    s2.set_transform(Mat4::identity().scale(0.5, 0.5, 0.5));
    w.objects.push_back(std::move(s1));
    w.objects.push_back(std::move(s2));
    w.lights.push_back({Tuple::point(-10, 10, -10), {1, 1, 1}});
    return w;
}

Camera Camera::make(size_t hsize, size_t vsize, double fov) {
    Camera c;
    double half_view = std::tan(fov / 2.0);
    double aspect = (double)hsize / (double)vsize;
    c.half_width = half_view * aspect;
    c.half_height = half_view;
    if (aspect >= 1.0) {
        c.half_width = half_view;
        c.half_height = half_view / aspect;
    }
    c.hsize = hsize;
    c.vsize = vsize;
    c.fov = fov;
    c.pixel_size = (c.half_width * 2.0) / (double)hsize;
    return c;
}

}  // namespace rtc
