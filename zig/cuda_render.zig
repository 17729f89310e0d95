//! Zig side of the drop-in boundary: flattens a `World(f64)` into the blob of include/rtc_blob.h and
//! calls the C ABI of include/rtc_b200.h in place of the body of `Camera.render`
//! (src/raytracer/camera.zig:80-125).
//!
//! STATUS: source only.  The build container has no Zig toolchain (SURVEY F4), so this file has
//! never been compiled; the authoritative, tested implementation of the same walk is
//! ray-tracer-challenge_b200/csrc/host/flatten.cpp.  Written against the reference's Zig
//! 0.12-dev-era API (build.zig:26-60).  Copy into src/raytracer/ of the reference to use.
const std = @import("std");
const Allocator = std.mem.Allocator;

const Tuple = @import("tuple.zig").Tuple;
const Matrix = @import("matrix.zig").Matrix;
const Color = @import("color.zig").Color;
const Canvas = @import("canvas.zig").Canvas;
const Camera = @import("camera.zig").Camera;
const World = @import("world.zig").World;
const Shape = @import("shapes/shape.zig").Shape;
const Material = @import("material.zig").Material;
const Pattern = @import("patterns/pattern.zig").Pattern;

pub const c = @cImport({
    @cInclude("rtc_b200.h");
});

const T = f64;

/// The reference's only parallelism knob is `n_threads` (lib.zig:157-164; camera.zig:88-97 hands one row to each
/// worker).  Its twin here: how many GPUs of this process render a frame.  Call before the first `render`; scenes
/// created afterwards are built once and replicated device to device, and every frame is split over the devices
/// (include/rtc_b200.h rtc_set_devices).  Nothing else in this file changes.
pub fn setDevices(n_gpus: usize) !void {
    if (c.rtc_set_devices(@intCast(n_gpus)) != 0) return error.CudaRenderFailed;
}

fn Section(comptime R: type) type {
    return std.ArrayList(R);
}

const Builder = struct {
    allocator: Allocator,
    lights: Section(c.rtc_light),
    materials: Section(c.rtc_material),
    patterns: Section(c.rtc_pattern),
    uvs: Section(c.rtc_uv_pattern),
    images: Section(c.rtc_image),
    transforms: Section(c.rtc_transform),
    leaves: Section(c.rtc_leaf),
    triangles: Section(c.rtc_triangle),
    normals: Section(c.rtc_tri_normals),
    guards: Section(c.rtc_guard),
    nodes: Section(c.rtc_node),
    children: Section(i32),
    csg_roots: Section(i32),
    texels: std.ArrayList(u8),
    guard: i32 = -1,

    fn init(allocator: Allocator) Builder {
        return .{
            .allocator = allocator,
            .lights = Section(c.rtc_light).init(allocator),
            .materials = Section(c.rtc_material).init(allocator),
            .patterns = Section(c.rtc_pattern).init(allocator),
            .uvs = Section(c.rtc_uv_pattern).init(allocator),
            .images = Section(c.rtc_image).init(allocator),
            .transforms = Section(c.rtc_transform).init(allocator),
            .leaves = Section(c.rtc_leaf).init(allocator),
            .triangles = Section(c.rtc_triangle).init(allocator),
            .normals = Section(c.rtc_tri_normals).init(allocator),
            .guards = Section(c.rtc_guard).init(allocator),
            .nodes = Section(c.rtc_node).init(allocator),
            .children = Section(i32).init(allocator),
            .csg_roots = Section(i32).init(allocator),
            .texels = std.ArrayList(u8).init(allocator),
        };
    }

    fn rows3(m: Matrix(T, 4), out: *[12]f64) void {
        for (0..3) |r| {
            for (0..4) |col| out[r * 4 + col] = m.data[r][col];
        }
    }

    fn addTransform(self: *Builder, inverse: Matrix(T, 4)) !i32 {
        var rec: c.rtc_transform = undefined;
        rows3(inverse, &rec.inverse);
        // (the C++ flattener deduplicates identical matrices; optional)
        try self.transforms.append(rec);
        return @intCast(self.transforms.items.len - 1);
    }

    fn addPattern(self: *Builder, p: *const Pattern(T)) anyerror!i32 {
        var rec = std.mem.zeroes(c.rtc_pattern);
        rec.a = -1;
        rec.b = -1;
        for (&rec.uv) |*u| u.* = -1;
        rows3(p._inverse_transform, &rec.inverse);
        switch (p.variant) {
            .solid => |s| {
                rec.kind = c.RTC_PAT_SOLID;
                rec.color = .{ s.a.r, s.a.g, s.a.b };
            },
            .stripes => |s| {
                rec.kind = c.RTC_PAT_STRIPES;
                rec.a = try self.addPattern(s.a);
                rec.b = try self.addPattern(s.b);
            },
            .rings => |s| {
                rec.kind = c.RTC_PAT_RINGS;
                rec.a = try self.addPattern(s.a);
                rec.b = try self.addPattern(s.b);
            },
            .checkers => |s| {
                rec.kind = c.RTC_PAT_CHECKERS;
                rec.a = try self.addPattern(s.a);
                rec.b = try self.addPattern(s.b);
            },
            .gradient => |s| {
                rec.kind = c.RTC_PAT_GRADIENT;
                rec.a = try self.addPattern(s.a);
                rec.b = try self.addPattern(s.b);
            },
            .radial_gradient => |s| {
                rec.kind = c.RTC_PAT_RADIAL_GRADIENT;
                rec.a = try self.addPattern(s.a);
                rec.b = try self.addPattern(s.b);
            },
            .blend => |s| {
                rec.kind = c.RTC_PAT_BLEND;
                rec.a = try self.addPattern(s.a);
                rec.b = try self.addPattern(s.b);
            },
            .perturb => |s| {
                rec.kind = c.RTC_PAT_PERTURB;
                rec.a = try self.addPattern(s.a);
                rec.octaves = @intCast(s.info.octaves);
                rec.perturb_scale = s.info.scale_value;
                rec.perturb_persistence = s.info.persistence;
            },
            .texture_map => |m| try self.addTextureMap(&rec, m),
            .test_pattern => rec.kind = c.RTC_PAT_TEST,
        }
        try self.patterns.append(rec);
        return @intCast(self.patterns.items.len - 1);
    }

    // texture maps: one uv record (spherical / planar / cylindrical) or six (cubic, order
    // front, back, left, right, up, down as in texture_map.zig:216); images are appended to the
    // texel section as RGBA8, 16-byte aligned.  Same structure as flatten.cpp add_uv/add_image.
    fn addTextureMap(self: *Builder, rec: *c.rtc_pattern, m: anytype) !void {
        _ = self;
        _ = m;
        rec.kind = c.RTC_PAT_TEXTURE_MAP;
        return error.TextureMapFlatteningNotWrittenInZigYet; // see flatten.cpp:80-97
    }

    fn addMaterial(self: *Builder, m: Material(T)) !i32 {
        var rec = std.mem.zeroes(c.rtc_material);
        rec.ambient = m.ambient;
        rec.diffuse = m.diffuse;
        rec.specular = m.specular;
        rec.shininess = m.shininess;
        rec.reflective = m.reflective;
        rec.transparency = m.transparency;
        rec.refractive_index = m.refractive_index;
        rec.pattern = try self.addPattern(&m.pattern);
        try self.materials.append(rec);
        return @intCast(self.materials.items.len - 1);
    }

    fn finite(t: Tuple(T)) bool {
        return std.math.isFinite(t.x) and std.math.isFinite(t.y) and std.math.isFinite(t.z);
    }

    fn emitLeaf(self: *Builder, s: *const Shape(T)) !i32 {
        var rec = std.mem.zeroes(c.rtc_leaf);
        rec.transform = try self.addTransform(s._inverse_transform);
        rec.material = try self.addMaterial(s.material);
        rec.flags = if (s.casts_shadow) c.RTC_CASTS_SHADOW else 0;
        rec.minimum = -std.math.inf(T);
        rec.maximum = std.math.inf(T);
        rec.triangle = -1;
        rec.guard = self.guard;
        const id: i32 = @intCast(self.leaves.items.len);
        switch (s.variant) {
            .sphere => rec.kind = c.RTC_SPHERE,
            .plane => rec.kind = c.RTC_PLANE,
            .cube => rec.kind = c.RTC_CUBE,
            .cylinder => |cy| {
                rec.kind = c.RTC_CYLINDER;
                rec.minimum = cy.min;
                rec.maximum = cy.max;
                if (cy.closed) rec.flags |= c.RTC_CLOSED;
            },
            .cone => |co| {
                rec.kind = c.RTC_CONE;
                rec.minimum = co.min;
                rec.maximum = co.max;
                if (co.closed) rec.flags |= c.RTC_CLOSED;
            },
            .triangle => |t| {
                rec.kind = c.RTC_TRIANGLE;
                rec.triangle = try self.emitTriangle(t.p1, t.e1, t.e2, id, .{ t.normal, t.normal, t.normal });
            },
            .smooth_triangle => |t| {
                rec.kind = c.RTC_SMOOTH_TRIANGLE;
                rec.triangle = try self.emitTriangle(t.p1, t.e1, t.e2, id, .{ t.n1, t.n2, t.n3 });
            },
            else => rec.kind = c.RTC_TEST_SHAPE,
        }
        const w = s.parentSpaceBounds().variant.bounding_box;
        const ob = s.bounds().variant.bounding_box;
        if (finite(ob.min) and finite(ob.max) and finite(w.min) and finite(w.max)) {
            rec.wmin = .{ w.min.x, w.min.y, w.min.z };
            rec.wmax = .{ w.max.x, w.max.y, w.max.z };
        } else {
            rec.flags |= c.RTC_UNBOUNDED;
        }
        try self.leaves.append(rec);
        return id;
    }

    fn emitTriangle(self: *Builder, p1: Tuple(T), e1: Tuple(T), e2: Tuple(T), leaf: i32, n: [3]Tuple(T)) !i32 {
        var t = std.mem.zeroes(c.rtc_triangle);
        t.p1 = .{ p1.x, p1.y, p1.z };
        t.e1 = .{ e1.x, e1.y, e1.z };
        t.e2 = .{ e2.x, e2.y, e2.z };
        t.leaf = leaf;
        var nr = std.mem.zeroes(c.rtc_tri_normals);
        for (n, 0..) |v, k| {
            nr.n[k * 3] = v.x;
            nr.n[k * 3 + 1] = v.y;
            nr.n[k * 3 + 2] = v.z;
        }
        try self.triangles.append(t);
        try self.normals.append(nr);
        return @intCast(self.triangles.items.len - 1);
    }

    /// Outside CSG: groups dissolve (their superset boxes cannot change a hit set).
    fn walk(self: *Builder, s: *const Shape(T)) anyerror!void {
        switch (s.variant) {
            .group => |g| {
                // a non-superset, culling-capable `_bbox` would be pushed as an rtc_guard here
                // (flatten.cpp Walker::walk); the scene schema cannot produce one.
                for (g.children.items) |*child| try self.walk(child);
            },
            .csg => try self.csg_roots.append(try self.node(s)),
            else => _ = try self.emitLeaf(s),
        }
    }

    /// At and below a CSG: keep the tree (post-order node records).
    fn node(self: *Builder, s: *const Shape(T)) anyerror!i32 {
        var n = std.mem.zeroes(c.rtc_node);
        n.leaf_begin = @intCast(self.leaves.items.len);
        switch (s.variant) {
            .group => |g| {
                n.kind = c.RTC_NODE_GROUP;
                var kids = std.ArrayList(i32).init(self.allocator);
                defer kids.deinit();
                for (g.children.items) |*child| try kids.append(try self.node(child));
                n.a = @intCast(self.children.items.len);
                n.b = @intCast(kids.items.len);
                try self.children.appendSlice(kids.items);
            },
            .csg => |cs| {
                n.kind = c.RTC_NODE_CSG;
                n.op = switch (cs.operation) {
                    .Union => c.RTC_CSG_UNION,
                    .Intersection => c.RTC_CSG_INTERSECTION,
                    .Difference => c.RTC_CSG_DIFFERENCE,
                };
                const b = cs._bbox.variant.bounding_box; // verbatim: never re-transformed (shape.zig:298-302)
                n.bmin = .{ b.min.x, b.min.y, b.min.z };
                n.bmax = .{ b.max.x, b.max.y, b.max.z };
                n.use_bbox = 1;
                n.a = try self.node(cs.left);
                n.b = try self.node(cs.right);
            },
            else => {
                n.kind = c.RTC_NODE_LEAF;
                n.a = try self.emitLeaf(s);
            },
        }
        n.leaf_end = @intCast(self.leaves.items.len);
        try self.nodes.append(n);
        return @intCast(self.nodes.items.len - 1);
    }

    fn put(out: *std.ArrayList(u8), comptime R: type, items: []const R) !u64 {
        while (out.items.len % 16 != 0) try out.append(0);
        const off: u64 = out.items.len;
        try out.appendSlice(std.mem.sliceAsBytes(items));
        return off;
    }

    fn finish(self: *Builder) ![]u8 {
        var out = std.ArrayList(u8).init(self.allocator);
        try out.appendNTimes(0, @sizeOf(c.rtc_blob_header));
        var h = std.mem.zeroes(c.rtc_blob_header);
        h.magic = c.RTC_BLOB_MAGIC;
        h.version = c.RTC_BLOB_VERSION;
        h.n_lights = @intCast(self.lights.items.len);
        h.n_materials = @intCast(self.materials.items.len);
        h.n_patterns = @intCast(self.patterns.items.len);
        h.n_uv_patterns = @intCast(self.uvs.items.len);
        h.n_images = @intCast(self.images.items.len);
        h.n_transforms = @intCast(self.transforms.items.len);
        h.n_leaves = @intCast(self.leaves.items.len);
        h.n_triangles = @intCast(self.triangles.items.len);
        h.n_guards = @intCast(self.guards.items.len);
        h.n_nodes = @intCast(self.nodes.items.len);
        h.n_children = @intCast(self.children.items.len);
        h.n_csg_roots = @intCast(self.csg_roots.items.len);
        h.texel_bytes = self.texels.items.len;
        h.off_lights = try put(&out, c.rtc_light, self.lights.items);
        h.off_materials = try put(&out, c.rtc_material, self.materials.items);
        h.off_patterns = try put(&out, c.rtc_pattern, self.patterns.items);
        h.off_uv_patterns = try put(&out, c.rtc_uv_pattern, self.uvs.items);
        h.off_images = try put(&out, c.rtc_image, self.images.items);
        h.off_transforms = try put(&out, c.rtc_transform, self.transforms.items);
        h.off_leaves = try put(&out, c.rtc_leaf, self.leaves.items);
        h.off_triangles = try put(&out, c.rtc_triangle, self.triangles.items);
        h.off_tri_normals = try put(&out, c.rtc_tri_normals, self.normals.items);
        h.off_guards = try put(&out, c.rtc_guard, self.guards.items);
        h.off_nodes = try put(&out, c.rtc_node, self.nodes.items);
        h.off_children = try put(&out, i32, self.children.items);
        h.off_csg_roots = try put(&out, i32, self.csg_roots.items);
        h.off_texels = try put(&out, u8, self.texels.items);
        while (out.items.len % 16 != 0) try out.append(0);
        h.total_bytes = out.items.len;
        @memcpy(out.items[0..@sizeOf(c.rtc_blob_header)], std.mem.asBytes(&h));
        return out.toOwnedSlice();
    }
};

pub fn flattenWorld(allocator: Allocator, world: World(T)) ![]u8 {
    var b = Builder.init(allocator);
    for (world.lights.items) |l| {
        try b.lights.append(.{
            .position = .{ l.position.x, l.position.y, l.position.z },
            .intensity = .{ l.intensity.r, l.intensity.g, l.intensity.b },
        });
    }
    for (world.objects.items) |*object| try b.walk(object);
    return b.finish();
}

/// Drop-in body for Camera(f64).render (camera.zig:80).
pub fn render(camera: Camera(T), allocator: Allocator, world: World(T)) !Canvas(T) {
    var image = try Canvas(T).new(allocator, camera.hsize, camera.vsize);
    errdefer image.destroy();

    const blob = try flattenWorld(allocator, world);
    defer allocator.free(blob);
    var scene: ?*c.rtc_scene = null;
    if (c.rtc_scene_create(blob.ptr, blob.len, &scene) != 0) return error.CudaRenderFailed;
    defer c.rtc_scene_destroy(scene);

    var cam = std.mem.zeroes(c.rtc_camera);
    cam.hsize = camera.hsize;
    cam.vsize = camera.vsize;
    cam.half_width = camera.half_width;
    cam.half_height = camera.half_height;
    cam.pixel_size = camera.pixel_size;
    for (0..4) |r| {
        for (0..4) |col| cam.inverse[r * 4 + col] = camera._inverse_transform.data[r][col];
    }

    // page-locked staging canvas: the kernel stores finished pixels straight into it over PCIe
    const n = camera.hsize * camera.vsize * 3;
    const raw = c.rtc_host_alloc(n * @sizeOf(f64)) orelse return error.OutOfMemory;
    defer c.rtc_host_free(raw);
    const rgb: [*]f64 = @ptrCast(@alignCast(raw));
    var stats: c.rtc_stats = undefined;
    if (c.rtc_render(scene, &cam, 5, null, rgb, &stats) != 0) return error.CudaRenderFailed;
    // Color(f64) is a packed struct (color.zig:9-15): copy, never alias.
    for (image.pixels, 0..) |*p, i| p.* = Color(T).new(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
    return image;
}
