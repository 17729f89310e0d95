// The reference's existing C ABI (src/lib.zig:233-309) on top of the CUDA renderer: include/rtc_renderer.h.
// Host logic only; every pixel comes from rtc_render_rgba8 (rtc_api.cu).
#include "../../include/rtc_renderer.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>

#include "host/flatten.hpp"
#include "host/loader.hpp"
#include "../../include/rtc_b200.h"

namespace {

using namespace rtc;

// Renderer(T) (lib.zig:65-191) with the CPU thread pool replaced by a device-resident scene.
struct Renderer {
    rtc_scene* scene = nullptr;
    rtc_camera cam{};
    Tuple from, to, up;  // camera._saved_from_to_up
    uint8_t* pixels = nullptr;
    size_t width = 0, height = 0;
    std::thread render_thread;
    bool render_thread_live = false;
    std::atomic<bool> render_done{false};
    rtc_stats stats{};
    bool have_stats = false;
    int render_rc = 0;

    ~Renderer() {
        if (render_thread_live) render_thread.join();
        if (scene) rtc_scene_destroy(scene);
        if (pixels) rtc_host_free(pixels);
    }
};

std::unique_ptr<Renderer> g_renderer;  // `var renderer: ?Renderer(f64) = null` (lib.zig:194)
std::string g_asset_dir;
bool g_asset_dir_set = false;

// init_renderer_result / init_renderer_thread / init_renderer_thread_done (lib.zig:213-215)
bool g_init_ok = false;
std::string g_init_err;
std::thread g_init_thread;
bool g_init_thread_live = false;
std::atomic<bool> g_init_done{false};

std::string asset_dir() {
    if (g_asset_dir_set) return g_asset_dir;
    const char* e = std::getenv("RTC_ASSET_DIR");
    return e ? e : ".";
}

// Renderer.new (lib.zig:76-106)
void init_worker(std::string scene_text) {
    std::unique_ptr<Renderer> r(new Renderer);
    try {
        SceneInfo si = parse_scene(scene_text, directory_loader(asset_dir()));
        std::vector<uint8_t> blob = flatten_world(si.world);
        r->cam = flatten_camera(si.camera);
        r->from = si.camera.from;
        r->to = si.camera.to;
        r->up = si.camera.up;
        r->width = (size_t)r->cam.hsize;
        r->height = (size_t)r->cam.vsize;
        if (rtc_scene_create(blob.data(), blob.size(), &r->scene) != RTC_OK) throw std::runtime_error(rtc_last_error());
        size_t n = 4 * r->width * r->height;
        r->pixels = static_cast<uint8_t*>(rtc_host_alloc(n));
        if (!r->pixels) throw std::runtime_error(rtc_last_error());
        std::memset(r->pixels, 0, n);  // lib.zig:88-97
        g_renderer = std::move(r);
        g_init_ok = true;
    } catch (const std::exception& e) {
        g_init_ok = false;
        g_init_err = e.what();
    }
    g_init_done.store(true, std::memory_order_release);
}

[[noreturn]] void uninitialised() {
    std::fputs("Renderer is uninitialized\n", stderr);  // the reference's panic message
    std::abort();
}

// camera.setTransform(viewTransform(from, to, up)) + _saved_from_to_up (lib.zig:176-177, 188-189)
void set_view(Renderer& r, const Tuple& from) {
    const double f[3] = {from.x, from.y, from.z}, t[3] = {r.to.x, r.to.y, r.to.z}, u[3] = {r.up.x, r.up.y, r.up.z};
    if (rtc_camera_look_at(&r.cam, f, t, u) != RTC_OK) {
        std::fprintf(stderr, "%s\n", rtc_last_error());  // the reference panics with the error name
        std::abort();
    }
    r.from = from;
}

}  // namespace

extern "C" {

void rtcRendererSetAssetDir(const char* dir) {
    g_asset_dir = dir ? dir : ".";
    g_asset_dir_set = true;
}

void startInitRenderer(const char* scene, size_t n_threads) {
    (void)n_threads;
    if (g_init_thread_live) {
        g_init_thread.join();
        g_init_thread_live = false;
    }
    g_renderer.reset();
    g_init_done.store(false, std::memory_order_release);
    g_init_thread = std::thread(init_worker, std::string(scene ? scene : ""));
    g_init_thread_live = true;
}

bool tryFinishInitRenderer(void) {
    bool done = g_init_done.load(std::memory_order_acquire);
    if (done && g_init_thread_live) {
        g_init_thread.join();
        g_init_thread_live = false;
    }
    return done;
}

bool initRendererIsOk(void) { return g_init_ok; }
const uint8_t* initRendererGetPixels(void) { return g_renderer ? g_renderer->pixels : nullptr; }
size_t initRendererGetWidth(void) { return g_renderer ? g_renderer->width : 0; }
size_t initRendererGetHeight(void) { return g_renderer ? g_renderer->height : 0; }
const char* initRendererGetErr(void) { return g_init_err.c_str(); }

void deinitRenderer(void) { g_renderer.reset(); }

void startRender(void) {
    if (!g_renderer) uninitialised();
    Renderer* r = g_renderer.get();
    if (r->render_thread_live) {
        r->render_thread.join();
        r->render_thread_live = false;
    }
    r->render_done.store(false, std::memory_order_release);
    // one worker instead of a pool with a task per row (lib.zig:150-156): the rows are the GPU's business
    r->render_thread = std::thread([r] {
        r->render_rc = rtc_render_rgba8(r->scene, &r->cam, 5, nullptr, r->pixels, &r->stats);
        if (r->render_rc != RTC_OK) {
            std::fprintf(stderr, "%s\n", rtc_last_error());  // worker errors panic in the reference (lib.zig:143)
            std::abort();
        }
        r->have_stats = true;
        r->render_done.store(true, std::memory_order_release);
    });
    r->render_thread_live = true;
}

bool tryFinishRender(void) {
    if (!g_renderer) uninitialised();
    Renderer* r = g_renderer.get();
    bool done = r->render_done.load(std::memory_order_acquire);
    if (done && r->render_thread_live) {
        r->render_thread.join();
        r->render_thread_live = false;
    }
    return done;
}

void rotateCamera(double angle) {
    if (!g_renderer) uninitialised();
    Renderer& r = *g_renderer;
    // lib.zig:166-178, operation for operation
    Tuple from = r.from;
    const Tuple delta = Tuple::point(0.0, 0.0, 0.0).sub(r.to);
    from = from.add(delta);
    from = Mat4::identity().rotate(r.up, angle).tuple_mul(from);
    from = from.sub(delta);
    set_view(r, from);
}

void moveCamera(double distance) {
    if (!g_renderer) uninitialised();
    Renderer& r = *g_renderer;
    // lib.zig:180-190
    const Tuple delta = r.to.sub(r.from).mul(distance);
    set_view(r, r.from.add(delta));
}

int rtcRendererLastStats(void* stats) {
    if (!g_renderer || !g_renderer->have_stats || !stats) return -1;
    std::memcpy(stats, &g_renderer->stats, sizeof(rtc_stats));
    return 0;
}

}  // extern "C"
