// JSON scene description -> Camera + World, host mirror of src/parsing/scene.zig:164-661.
// Order of operations inside parse_object is the reference's (scene.zig:440-591):
//   inherit -> construct variant (children parsed WITHOUT the parent transform) -> setTransform
//   (push-down) -> material / casts_shadow -> divide(8)   ... at EVERY nesting level.
#include <algorithm>
#include <atomic>
#include <exception>
#include <map>
#include <thread>

#include "json.hpp"
#include "loader.hpp"

namespace rtc {
namespace {

struct Inherited {  // InheritedState (scene.zig:430-436)
    bool has_material = false;
    Material material;
    Mat4 transform = Mat4::identity();
    int casts_shadow = -1;
};

struct Ctx {
    const FileLoader& load;
    std::map<std::string, const Json*> definitions;
    std::vector<std::shared_ptr<const Image>> images;
    std::vector<std::shared_ptr<Image>> missing;  // stand-ins filled after the whole scene is read
    std::map<std::string, Shape> obj_groups;      // OBJ files parsed so far, by "file|normalize"
    // Images met so far, not yet decoded: nothing in the parse looks inside an image, so the files are inflated side by
    // side once the scene is read (skybox.json: five 2048x2048 PNGs) instead of one after the other where they appear.
    struct Pending {
        std::shared_ptr<Image> img;
        std::string bytes, file;
    };
    std::vector<Pending> pending;
};

// Decodes cx.pending on all host cores; the error reported is the one of the FIRST failing file in scene order, which is
// the one a decode-as-you-go parse would have stopped at.
void decode_pending(Ctx& cx) {
    const size_t n = cx.pending.size();
    if (n == 0) return;
    std::vector<std::exception_ptr> errors(n);
    std::atomic<size_t> next{0};
    auto work = [&] {
        for (size_t i = next++; i < n; i = next++) {
            try {
                std::shared_ptr<Image> d = decode_image(cx.pending[i].bytes, cx.pending[i].file);
                *cx.pending[i].img = std::move(*d);
            } catch (...) {
                errors[i] = std::current_exception();
            }
        }
    };
    const size_t n_threads = std::min<size_t>(n, std::max(1u, std::thread::hardware_concurrency()));
    std::vector<std::thread> pool;
    for (size_t t = 1; t < n_threads; t++) pool.emplace_back(work);
    work();
    for (std::thread& t : pool) t.join();
    cx.pending.clear();
    for (const std::exception_ptr& e : errors)
        if (e) std::rethrow_exception(e);
}

void stamp_triangles(Shape& s, const Material& mat, bool casts_shadow) {
    if (s.kind == ShapeKind::Group) {
        for (Shape& c : s.children) stamp_triangles(c, mat, casts_shadow);
    } else {
        s.material = mat;
        s.casts_shadow = casts_shadow;
    }
}

void vec3(const Json& j, double out[3]) {
    const auto& a = j.array();
    if (a.size() != 3) throw JsonError("expected [x,y,z]");
    for (int i = 0; i < 3; i++) out[i] = a[i].number();
}

// parseTransform (scene.zig:214-241): ops applied in list order, each one left-multiplied.
Mat4 parse_transform(const Json& list) {
    Mat4 m = Mat4::identity();
    for (const Json& op : list.array()) {
        const auto& kv = op.single();
        const std::string& k = kv.first;
        const Json& v = kv.second;
        double a[3];
        if (k == "translate") {
            vec3(v, a);
            m = m.translate(a[0], a[1], a[2]);
        } else if (k == "scale") {
            vec3(v, a);
            m = m.scale(a[0], a[1], a[2]);
        } else if (k == "rotate-x") {
            m = m.rotate_x(v.number());
        } else if (k == "rotate-y") {
            m = m.rotate_y(v.number());
        } else if (k == "rotate-z") {
            m = m.rotate_z(v.number());
        } else if (k == "shear") {
            auto g = [&](const char* n) {
                const Json* f = v.find(n);
                return f ? f->number() : 0.0;
            };
            m = m.shear(g("xy"), g("xz"), g("yx"), g("yz"), g("zx"), g("zy"));
        } else {
            throw JsonError("unknown transform '" + k + "'");
        }
    }
    return m;
}

Pattern parse_pattern(Ctx& cx, const Json& j);

UvPattern parse_uv_pattern(Ctx& cx, const Json& j) {  // scene.zig:243-298
    const auto& kv = j.single();
    const Json& v = kv.second;
    UvPattern uv;
    if (kv.first == "align-check") {
        uv.kind = UvPattern::AlignCheck;
        const char* names[5] = {"central", "upper-left", "upper-right", "bottom-left", "bottom-right"};
        for (int i = 0; i < 5; i++) uv.sub[i] = share(parse_pattern(cx, v.at(names[i])));
    } else if (kv.first == "checkers") {
        uv.kind = UvPattern::Checkers;
        uv.width = v.at("width").number();
        uv.height = v.at("height").number();
        const auto& ps = v.at("patterns").array();
        if (ps.size() != 2) throw JsonError("uv checkers needs 2 patterns");
        uv.sub[0] = share(parse_pattern(cx, ps[0]));
        uv.sub[1] = share(parse_pattern(cx, ps[1]));
    } else if (kv.first == "image") {
        uv.kind = UvPattern::ImageMap;
        const std::string& file = v.at("file").string();
        const Json* interp = v.find("interpolation");
        if (interp) {
            if (interp->string() == "bilinear") uv.bilinear = true;
            else if (interp->string() != "none") throw JsonError("unknown interpolation");
        }
        std::string bytes;
        bool found = true;
        try {
            bytes = cx.load(file);
        } catch (const std::exception&) {
            found = false;
        }
        if (found) {
            auto img = std::make_shared<Image>();
            img->name = file;
            cx.images.push_back(img);
            cx.pending.push_back({img, std::move(bytes), file});
            uv.image = img;
        } else {
            // SURVEY F6 (data/lancellotti-posy.png absent from the mount): placeholder, filled later.
            auto img = std::make_shared<Image>();
            img->name = file;
            cx.missing.push_back(img);
            uv.image = img;
        }
    } else {
        throw JsonError("unknown uv pattern '" + kv.first + "'");
    }
    return uv;
}

Pattern parse_pattern(Ctx& cx, const Json& j) {  // scene.zig:300-405
    const auto& kv = j.at("type").single();
    const std::string& k = kv.first;
    const Json& v = kv.second;
    Pattern p;
    auto two = [&](Pattern::Kind kind) {
        const auto& a = v.array();
        if (a.size() != 2) throw JsonError("pattern needs 2 sub-patterns");
        PatternP p1 = share(parse_pattern(cx, a[0]));
        PatternP p2 = share(parse_pattern(cx, a[1]));
        return Pattern::binary(kind, p1, p2);
    };
    if (k == "solid") {
        double c[3];
        vec3(v, c);
        p = Pattern::solid({c[0], c[1], c[2]});
    } else if (k == "stripes") p = two(Pattern::Stripes);
    else if (k == "rings") p = two(Pattern::Rings);
    else if (k == "gradient") p = two(Pattern::Gradient);
    else if (k == "radial-gradient") p = two(Pattern::RadialGradient);
    else if (k == "checkers") p = two(Pattern::Checkers);
    else if (k == "blend") p = two(Pattern::Blend);
    else if (k == "perturb") {
        p.kind = Pattern::Perturb;
        p.a = share(parse_pattern(cx, v));
    } else if (k == "texture-map") {
        p.kind = Pattern::TextureMap;
        const auto& mk = v.single();
        if (mk.first == "spherical" || mk.first == "planar" || mk.first == "cylindrical") {
            p.mapping = mk.first == "spherical" ? Pattern::Spherical
                        : mk.first == "planar"  ? Pattern::Planar
                                                : Pattern::Cylindrical;
            p.uv[0] = parse_uv_pattern(cx, mk.second.at("uv-pattern"));
        } else if (mk.first == "cubic") {
            p.mapping = Pattern::Cubic;
            const char* faces[6] = {"front", "back", "left", "right", "up", "down"};
            for (int i = 0; i < 6; i++) p.uv[i] = parse_uv_pattern(cx, mk.second.at(faces[i]));
        } else {
            throw JsonError("UnknownMapping");
        }
    } else {
        throw JsonError("unknown pattern '" + k + "'");
    }
    if (const Json* t = j.find("transform"))
        if (t->type != Json::Null) p.set_transform(parse_transform(*t));
    return p;
}

Material parse_material(Ctx& cx, const Json& j, bool has_base, const Material& base) {  // scene.zig:407-428
    Material m = has_base ? base : Material();
    auto opt = [&](const char* name, double& field) {
        const Json* f = j.find(name);
        if (f && f->type != Json::Null) field = f->number();
    };
    if (const Json* p = j.find("pattern"))
        if (p->type != Json::Null) m.pattern = share(parse_pattern(cx, *p));
    opt("ambient", m.ambient);
    opt("diffuse", m.diffuse);
    opt("specular", m.specular);
    opt("shininess", m.shininess);
    opt("reflective", m.reflective);
    opt("transparency", m.transparency);
    opt("refractive-index", m.refractive_index);
    return m;
}

Inherited inherit(Ctx& cx, const Json& obj, const Inherited& in) {  // ObjectConfig.inherit (scene.zig:164-190)
    Inherited out;
    const Json* m = obj.find("material");
    if (m && m->type != Json::Null) {
        out.material = parse_material(cx, *m, in.has_material, in.material);
        out.has_material = true;
    } else {
        out.material = in.material;
        out.has_material = in.has_material;
    }
    const Json* t = obj.find("transform");
    out.transform = (t && t->type != Json::Null) ? parse_transform(*t).mul(in.transform) : in.transform;
    const Json* s = obj.find("casts-shadow");
    out.casts_shadow = (s && s->type != Json::Null) ? (s->boolean() ? 1 : 0) : in.casts_shadow;
    return out;
}

double opt_num(const Json& v, const char* name, double dflt) {
    const Json* f = v.find(name);
    return f ? f->number() : dflt;
}
bool opt_bool(const Json& v, const char* name, bool dflt) {
    const Json* f = v.find(name);
    return f ? f->boolean() : dflt;
}

Shape parse_object(Ctx& cx, const Json& obj, const Inherited& inherited) {  // scene.zig:440-591
    Inherited info = inherit(cx, obj, inherited);
    const auto& kv = obj.at("type").single();
    const std::string& k = kv.first;
    const Json& v = kv.second;

    Inherited child_state;  // what groups / csg pass down: material + shadow flag, NOT the transform
    child_state.has_material = info.has_material;
    child_state.material = info.material;
    child_state.casts_shadow = info.casts_shadow;

    auto build = [&]() -> Shape {
        if (k == "from-definition") {
            auto it = cx.definitions.find(v.string());
            if (it == cx.definitions.end()) throw JsonError("UnknownDefinition");
            Inherited down = child_state;
            down.transform = inherited.transform;
            Shape parent = parse_object(cx, *it->second, down);
            Inherited parent_state;
            parent_state.has_material = true;
            parent_state.material = parent.material;
            parent_state.transform = parent.transform;  // identity for groups / CSG (never set there)
            parent_state.casts_shadow = parent.casts_shadow ? 1 : 0;
            info = inherit(cx, obj, parent_state);
            return parent;
        }
        if (k == "from-obj") {
            // The reference reads and parses the file again for every object made from it (scene.zig:455-488: six
            // times data/dragon.obj in dragons.json).  Here the file is parsed once per (file, normalize) with no
            // inherited state; every object takes a copy of that group and stamps its own material / casts-shadow on
            // the triangles -- the only two things obj.zig:101-160 takes from the inherited state.
            const std::string file = v.at("file").string();
            const bool normalize = opt_bool(v, "normalize", true);
            const std::string key = file + (normalize ? "|n" : "|r");
            auto it = cx.obj_groups.find(key);
            if (it == cx.obj_groups.end()) {
                ObjParser parser;
                parser.load_obj(cx.load(file), ObjInherited{}, normalize);
                it = cx.obj_groups.emplace(key, parser.to_group()).first;
            }
            Shape g = it->second;
            const Material mat = info.has_material ? info.material : Material();
            const bool casts = info.casts_shadow < 0 ? true : (info.casts_shadow != 0);
            stamp_triangles(g, mat, casts);
            return g;
        }
        if (k == "sphere") return Shape::sphere();
        if (k == "plane") return Shape::plane();
        if (k == "cube") return Shape::cube();
        if (k == "cylinder" || k == "cone") {
            Shape c = k == "cylinder" ? Shape::cylinder() : Shape::cone();
            c.minimum = opt_num(v, "min", -kInf);
            c.maximum = opt_num(v, "max", kInf);
            c.closed = opt_bool(v, "closed", false);
            return c;
        }
        if (k == "triangle") {
            double a[3], b[3], c[3];
            vec3(v.at("p1"), a);
            vec3(v.at("p2"), b);
            vec3(v.at("p3"), c);
            return Shape::triangle(Tuple::point(a[0], a[1], a[2]), Tuple::point(b[0], b[1], b[2]),
                                   Tuple::point(c[0], c[1], c[2]));
        }
        if (k == "group") {
            Shape g = Shape::group();
            for (const Json& child : v.array()) g.add_child(parse_object(cx, child, child_state));
            return g;
        }
        if (k == "csg") {
            Shape l = parse_object(cx, v.at("left"), child_state);
            Shape r = parse_object(cx, v.at("right"), child_state);
            const std::string& o = v.at("operation").string();
            CsgOp op;
            if (o == "union") op = CsgOp::Union;
            else if (o == "intersection") op = CsgOp::Intersection;
            else if (o == "difference") op = CsgOp::Difference;
            else throw JsonError("unknown csg operation");
            return Shape::csg(op, std::move(l), std::move(r));
        }
        throw JsonError("UnknownShape '" + k + "'");
    };

    Shape shape = build();
    shape.set_transform(info.transform);
    if (info.has_material) shape.material = info.material;
    if (info.casts_shadow >= 0) shape.casts_shadow = info.casts_shadow != 0;
    shape.divide(8);
    return shape;
}

}  // namespace

SceneInfo parse_scene(const std::string& json_text, const FileLoader& load, long width_override,
                      long height_override) {
    Json root = JsonParser(json_text).parse();
    Ctx cx{load, {}, {}, {}, {}, {}};
    if (const Json* defs = root.find("shape-definitions"))
        for (const Json& d : defs->array()) cx.definitions[d.at("name").string()] = &d.at("value");

    const Json& cam = root.at("camera");
    size_t w = (size_t)cam.at("width").number(), h = (size_t)cam.at("height").number();
    if (width_override > 0) w = (size_t)width_override;
    if (height_override > 0) h = (size_t)height_override;
    SceneInfo out;
    out.camera = Camera::make(w, h, cam.at("field-of-view").number());
    double f[3], t[3], u[3];
    vec3(cam.at("from"), f);
    vec3(cam.at("to"), t);
    vec3(cam.at("up"), u);
    out.camera.from = Tuple::point(f[0], f[1], f[2]);
    out.camera.to = Tuple::point(t[0], t[1], t[2]);
    out.camera.up = Tuple::vec3(u[0], u[1], u[2]);
    out.camera.set_transform(Mat4::view_transform(out.camera.from, out.camera.to, out.camera.up));

    try {
        for (const Json& o : root.at("objects").array()) out.world.objects.push_back(parse_object(cx, o, Inherited{}));
        for (const Json& l : root.at("lights").array()) {
            const auto& kv = l.single();
            if (kv.first != "point-light") throw JsonError("unknown light");
            double p[3], i[3];
            vec3(kv.second.at("position"), p);
            vec3(kv.second.at("intensity"), i);
            out.world.lights.push_back({Tuple::point(p[0], p[1], p[2]), {i[0], i[1], i[2]}});
        }
    } catch (...) {
        decode_pending(cx);  // an image met BEFORE the failing element that does not decode is the earlier error
        throw;
    }
    decode_pending(cx);
    for (auto& m : cx.missing) {
        auto s = standin_image(cx.images, m->name);
        m->width = s->width;
        m->height = s->height;
        m->rgb8 = s->rgb8;
        m->name = s->name;
    }
    return out;
}

}  // namespace rtc
