// Host mirror of the reference parsing layer (src/parsing/scene.zig, src/parsing/obj.zig) plus the
// image decoders that zigimg / Canvas.fromPpm provide there.
#pragma once
#include <functional>
#include <string>

#include "scene.hpp"

namespace rtc {

// load_file_data callback of parseScene (scene.zig:612-618): file name -> bytes.
using FileLoader = std::function<std::string(const std::string&)>;
FileLoader directory_loader(const std::string& dir);
std::string read_file(const std::string& path);

// parseScene (scene.zig:612-661).  width/height > 0 replace camera.width/height in the JSON, the
// only way the reference can change resolution.
SceneInfo parse_scene(const std::string& json_text, const FileLoader& load, long width_override = 0,
                      long height_override = 0);

// ObjParser (obj.zig:10-283)
struct ObjInherited {
    bool has_material = false;
    Material material;
    int casts_shadow = -1;  // -1 = null
};
struct ObjParser {
    Shape default_group = Shape::group();
    std::vector<size_t> active_path;  // child index of the active named group; empty = default
    std::vector<std::pair<std::string, size_t>> named_groups;
    Tuple offset = Tuple::vec3(0, 0, 0);
    double scale = 1.0;
    std::vector<Tuple> vertices, normals;
    size_t lines_ignored = 0;
    void load_obj(const std::string& text, const ObjInherited& st, bool normalize);
    Shape to_group() const { return default_group; }
    const Shape* named_group(const std::string& name) const;

   private:
    Shape& active();
    bool handle_line(const std::string& line, const ObjInherited& st);
};

// PNG (8-bit, colour types 0/2/3/4/6, non-interlaced; zlib inflate) and P3 PPM (canvas.zig:48-121).
std::shared_ptr<Image> decode_png(const std::string& bytes, const std::string& name);
std::shared_ptr<Image> decode_ppm(const std::string& text);
std::shared_ptr<Image> decode_image(const std::string& bytes, const std::string& name);
// data/lancellotti-posy.png is absent from the reference mount.  Stand-in = uniform
// texel holding the per-channel mean of the other images decoded so far for the same scene.
std::shared_ptr<Image> standin_image(const std::vector<std::shared_ptr<const Image>>& others,
                                     const std::string& name);

}  // namespace rtc
