/* rtc_b200.h — C ABI of the B200 renderer that replaces the reference's CPU render loop.
 *
 * Replaces:  Camera(T).render(self, allocator, world) !Canvas(T)      src/raytracer/camera.zig:80-125
 *            Renderer.startRender / renderWorker (RGBA8 twin)          src/lib.zig:135-164
 * and everything those call per pixel (World.colorAt, world.zig:111).  Plain pointers and sizes
 * only; no allocator, no Zig or torch type crosses this boundary.  Every function returns 0 on
 * success or a negative rtc_status; rtc_last_error() gives the message for the calling thread.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with
 * RTC_ERR_CUDA.
 */
#ifndef RTC_B200_H
#define RTC_B200_H
#include <stddef.h>
#include <stdint.h>

#include "rtc_blob.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum rtc_status {
    RTC_OK = 0,
    RTC_ERR_INVALID = -1,  /* bad argument / malformed blob */
    RTC_ERR_CUDA = -2,     /* no device, launch or copy failure */
    RTC_ERR_PARSE = -3,    /* scene JSON / OBJ / image error (host loader) */
    RTC_ERR_CAPACITY = -4  /* a per-ray CSG interval list or traversal stack overflowed */
} rtc_status;

typedef struct rtc_scene rtc_scene; /* device-resident scene, opaque */

/* Camera(T) fields read by rayForPixel (camera.zig:22-30, :64-76). */
typedef struct rtc_camera {
    uint64_t hsize, vsize;
    double half_width, half_height, pixel_size;
    double inverse[16]; /* Camera._inverse_transform, row-major */
} rtc_camera;

/* Which rows of the frame this call renders: interleaved scanline tiles.  Row y belongs to the
 * call iff (y / tile_rows) % n_parts == part.  {1,0,1} = whole frame. */
typedef struct rtc_partition {
    uint32_t tile_rows, part, n_parts;
} rtc_partition;

typedef struct rtc_stats {
    uint64_t primary_rays, shadow_rays, secondary_rays; /* world.zig:91-92,160-165,178-187 */
    uint64_t pixels;
    float kernel_ms; /* CUDA-event time of the kernels of this call (render kernel + tile-order kernels, or the wavefront launches); replicated scenes: the slowest device */
    uint32_t kernel; /* which kernel rendered the frame: 0 megakernel, 1 wavefront, 2 megakernel with several lanes per pixel (small partitions) */
    uint64_t launches; /* kernels launched by this call */
    /* shadow rays (counted in shadow_rays, as the reference casts them) that were NOT traced because
     * the light is behind the surface: lighting() returns ambient either way (material.zig:55-63) */
    uint64_t shadow_rays_culled;
} rtc_stats;

const char* rtc_last_error(void);
int rtc_device_count(void);
/* Selects the (first) CUDA device of scenes created afterwards (default 0; process-wide). */
int rtc_set_device(int device);
/* How many GPUs render a frame: scenes created afterwards are replicated on devices d .. d+n-1 (d = rtc_set_device).
 * The blob is parsed and its BVHs built ONCE; the scene's slab is copied device to device.  A whole-frame render call on
 * such a scene -- rtc_render, rtc_render_rgba8, rtc_render_device*, the lib.zig twin -- hands the interleaved 8-row
 * tiles k, k+n, ... to device k and returns when all are in; frames are bit-identical to the one-device frame.  This is
 * the knob that n_threads is in the reference (lib.zig:157-164 spawns n_threads workers; camera.zig:88-97 one task per
 * row): the caller chooses how many workers, the library splits the rows.  Default 1. */
int rtc_set_devices(int n);

/* What rtc_scene_create builds out of a blob, without touching a device (host-side validation + mesh instancing +
 * BVH build only; runs anywhere): a World(f64) handed to Camera.render (world.zig:20-25) becomes n_prims analytic
 * top-level leaves, n_csg CSG programs and n_meshes triangle-mesh INSTANCES over n_geometries distinct BVHs (objects
 * made from the same OBJ data share one: scene.zig:455-488 re-parses the file per object instead). */
typedef struct rtc_scene_info {
    uint32_t n_prims, n_meshes, n_geometries, n_csg;
    uint64_t n_bvh_nodes, n_mesh_triangles; /* of the distinct geometries */
    uint64_t device_bytes;                  /* size of the one slab the scene occupies in HBM */
    uint32_t bvh_max_depth, n_leaves;
    double build_seconds;
} rtc_scene_info;
int rtc_blob_inspect(const void* blob, size_t nbytes, rtc_scene_info* info_out);

/* Scene lifetime.  The blob is copied; the caller may free it as soon as the call returns. */
int rtc_scene_create(const void* blob, size_t nbytes, rtc_scene** out);
void rtc_scene_destroy(rtc_scene* scene);

/* camera.zig:80 twin.  rgb_out: caller-owned HOST memory, hsize*vsize*3 doubles, row-major, y down,
 * linear RGB exactly as Canvas.pixels holds it.  Rows outside `part` (NULL = whole frame) are left
 * untouched.  max_depth is the `remaining_recursions` seed (5 at camera.zig:118). */
int rtc_render(rtc_scene* scene, const rtc_camera* cam, int max_depth, const rtc_partition* part,
               double* rgb_out, rtc_stats* stats);
/* Page-locked, device-mapped host memory for canvases.  rtc_render / rtc_render_rgba8 accept any host
 * pointer; when it comes from rtc_host_alloc (or is otherwise pinned and mapped) the frame reaches it WHILE it is
 * still rendering -- whole frames: the kernel announces every finished row of tiles and the calling thread hands
 * them to the copy engine; row partitions: warps of the kernel move them -- instead of being staged in device
 * memory and copied afterwards (Output).  The Zig shim allocates its staging canvas here. */
void* rtc_host_alloc(size_t nbytes);
void rtc_host_free(void* p);
/* lib.zig:135-155 twin: clamp()-quantised RGBA8 (color.zig:61-71), alpha 255, 4 bytes per pixel. */
int rtc_render_rgba8(rtc_scene* scene, const rtc_camera* cam, int max_depth, const rtc_partition* part,
                     uint8_t* rgba_out, rtc_stats* stats);
/* Same render, result left in DEVICE memory owned by the caller (e.g. a torch tensor used as an
 * NCCL send buffer): d_rgb holds the rows of `part` only, packed in increasing y,
 * rtc_partition_rows(...) * hsize * 3 doubles.  `stream` is a cudaStream_t (NULL = default). */
int rtc_render_device(rtc_scene* scene, const rtc_camera* cam, int max_depth, const rtc_partition* part,
                      double* d_rgb, void* stream, rtc_stats* stats);
/* Multi-GPU twin: d_frame is the WHOLE frame (hsize*vsize*3 doubles) and only the rows of `part` are
 * written, each at its frame position.  d_frame may be memory of ANOTHER GPU mapped into this
 * device's address space (CUDA peer access / symmetric memory over NVLink): every rank then renders
 * its interleaved tiles straight into rank 0's canvas and no gather step is needed (camera.zig:116-121:
 * pixels are independent, so the only exchange of the path is the framebuffer itself). */
int rtc_render_device_frame(rtc_scene* scene, const rtc_camera* cam, int max_depth, const rtc_partition* part,
                            double* d_frame, void* stream, rtc_stats* stats);
/* rtc_render_device_frame without the wait: the frame is only put on `stream` (render kernel, tile-order kernels, a
 * 64-byte copy of the ray counters), nothing is synchronised and no statistics are returned.  rtc_stats_fetch waits for
 * the frames enqueued so far and hands out their statistics, oldest first (stats_out may be NULL to just drain; at most
 * 16 frames are kept -- older ones are dropped); it returns the first error any of those frames raised.  The first
 * frames of a (scene, frame size) are measured to pick the kernel -- five, eleven for small partitions, fifteen for
 * frames shared tile by tile (six heavy-tile thresholds, two frames each) -- and therefore complete before the call
 * returns.  Frames of one scene are ordered among themselves even when put on different streams. */
int rtc_render_device_async(rtc_scene* scene, const rtc_camera* cam, int max_depth, const rtc_partition* part,
                            double* d_frame, void* stream);
int rtc_stats_fetch(rtc_scene* scene, rtc_stats* stats_out, uint32_t max_frames, uint32_t* n_out);
/* rtc_render_device for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  The whole frame,
 * hsize*vsize*3 doubles, into caller-owned device memory d_rgb, by the megakernel, bit-identical to rtc_render_device's.
 * The call makes no cudaMalloc / cudaFree / allocating library call: the kernel's scratch (ray counters, per-tile costs,
 * the tile order made from the previous call's costs) lives in a fixed array in the library's device image -- 3.1 MB that
 * every GPU which loads the library holds until the process exits.  LIMIT: frames of at most rtc_static_max_tiles() =
 * 262,144 tiles of 8x4 pixels (3840x2160 fits); larger frames are refused with RTC_ERR_INVALID.
 * The array is shared by EVERY call on a device and keeps the last call's contents (that is how a re-render of a view
 * gets its tile order); the call clears what must start at zero.  Calls that overlap on the GPU (different streams without
 * an event between them, or different host threads) race on it: one call at a time per device.  It is independent of the
 * scratch of the other entry points, which may be used on the same scene in between.
 * stats != NULL: the call waits for the frame and reports rays, kernel time and capacity errors like rtc_render_device.
 * stats == NULL: the work (two memsets, the render kernel and, when the call before recorded tile costs, the two order kernels) is only
 * put on `stream`;
 * nothing is synchronised and capacity errors are not reported.  Replicated scenes (rtc_set_devices) render on their
 * first device. */
uint64_t rtc_static_max_tiles(void);
int rtc_render_device_static(rtc_scene* scene, const rtc_camera* cam, int max_depth, double* d_rgb, void* stream,
                             rtc_stats* stats);
/* One frame shared by several renderers, one PROCESS per GPU (rank r of n_ranks; the layout torchrun / MPI hosts use).
 * Instead of fixed row partitions the renderers take the tiles of ONE order -- the frame's 8x4-pixel tiles sorted into
 * buckets by what they cost in the previous frame -- in turn: entry r, r + n, r + 2n, ...: every renderer gets its
 * share of the expensive tiles.  To make that order, every renderer needs every tile's cost: each adds the cost of the
 * tiles it renders to the buffers of ALL renderers (cost[b][r] = buffer b of renderer r as the calling device
 * addresses it: peer-mapped / symmetric memory over NVLink; two buffers per renderer, `words` 32-bit words each,
 * >= rtc_tile_share_words(hsize, vsize), allocated by the caller).
 * Protocol: (1) every rank calls rtc_scene_share_tiles, then a barrier; (2) every rank renders frame k with
 * rtc_render_device_frame / rtc_render_device_async(part = NULL, d_frame = the shared whole frame); (3) a device-side
 * barrier across the ranks on the streams used (it is also what tells rank 0 that the frame is complete) -- then frame
 * k + 1.  Without (3) the renderers' copies of the costs can differ and tiles can be rendered twice or not at all.
 * rtc_scene_share_tiles(scene, NULL) ends the sharing. */
typedef struct rtc_tile_share {
    uint32_t rank, n_ranks; /* n_ranks <= 8 */
    uint64_t words;
    void* cost[2][8];
} rtc_tile_share;
uint64_t rtc_tile_share_words(uint64_t hsize, uint64_t vsize);
int rtc_scene_share_tiles(rtc_scene* scene, const rtc_tile_share* share);
uint64_t rtc_partition_rows(uint64_t vsize, const rtc_partition* part);

/* Host loader: the C++ mirror of src/parsing (scene.zig:612, obj.zig:195) + flattener, for hosts
 * that do not have the Zig parsing layer at hand (tests, bench, Python).  width/height > 0 replace
 * camera.width/height (the reference's only resolution knob, scene.zig:629-631).
 * rtc_blob_free releases the buffer. */
int rtc_load_scene_json(const char* json_text, const char* asset_dir, long width, long height,
                        void** blob_out, size_t* nbytes_out, rtc_camera* cam_out);
void rtc_blob_free(void* blob);
/* The step after the render.  rtc_canvas_ppm: Canvas.ppm (canvas.zig:181-254) of an f64 frame, character for
 * character (clamp() quantisation, lines broken before column 70); *text_out is NUL-terminated, release it with
 * rtc_blob_free.  rtc_write_png: the rgb24 PNG main.zig:95-99 writes from Canvas.toImage (canvas.zig:150-176),
 * from the RGBA8 frame of rtc_render_rgba8 (alpha dropped). */
int rtc_canvas_ppm(const double* rgb, uint64_t width, uint64_t height, char** text_out, size_t* nbytes_out);
int rtc_write_png(const char* path, const uint8_t* rgba, uint64_t width, uint64_t height);
/* lib.zig:166-190 camera orbit: rebuilds cam->inverse from from/to/up (matrix.zig:54-67). */
int rtc_camera_look_at(rtc_camera* cam, const double from[3], const double to[3], const double up[3]);

#ifdef __cplusplus
}
#endif
#endif
