/* rtc_renderer.h — the reference's EXISTING C ABI, served by the B200 renderer.
 *
 * src/lib.zig:233-309 exports twelve functions around one global Renderer(f64) (lib.zig:194): parse a
 * scene description, render it to an RGBA8 buffer the library owns, orbit / dolly the camera and render
 * again; the caller polls instead of blocking (www/index.js:69-80, 104-144, 253-298).  This header
 * declares the same twelve names with the same signatures and meaning, so a native host written against
 * the reference's library links against librtc_b200.so unchanged.  Differences, all behind the boundary:
 *   - the frame is rendered by the CUDA path (rtc_render_rgba8, include/rtc_b200.h); the scene is parsed
 *     once and stays resident on the device, so rotateCamera / moveCamera + startRender is a re-render
 *     ("preheated", README.md:122-125), not a reload;
 *   - n_threads is accepted and ignored (it sized the reference's CPU thread pool, lib.zig:150);
 *   - files named by the scene (OBJ, PNG, PPM) are read from the directory set with
 *     rtcRendererSetAssetDir (default: $RTC_ASSET_DIR, else "."), the stand-in for the host-provided
 *     loadFileData import (lib.zig:121-123);
 *   - where the reference panics ("Renderer is uninitialized", lib.zig:279,291,301,309) these print the same
 *     message to stderr and abort().
 * One renderer per process, like the reference.  Not thread-safe beyond the start/tryFinish protocol.
 */
#ifndef RTC_RENDERER_H
#define RTC_RENDERER_H
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* lib.zig:233-237.  `scene` is the JSON scene description (NUL-terminated); it must stay valid until
 * tryFinishInitRenderer() has returned true.  Parsing, flattening and the upload run on a worker thread. */
void startInitRenderer(const char* scene, size_t n_threads);
/* lib.zig:239-244: true once the worker is done (and joined). */
bool tryFinishInitRenderer(void);
/* lib.zig:246-251 */
bool initRendererIsOk(void);
/* lib.zig:253-263: valid when initRendererIsOk().  width*height*4 bytes, RGBA, row-major, y down, all 0 until
 * the first render (lib.zig:88-97); owned by the library until deinitRenderer(). */
const uint8_t* initRendererGetPixels(void);
size_t initRendererGetWidth(void);
size_t initRendererGetHeight(void);
/* lib.zig:265-267: valid when !initRendererIsOk(); NUL-terminated message. */
const char* initRendererGetErr(void);
/* lib.zig:269-273 */
void deinitRenderer(void);
/* lib.zig:275-281: starts rendering the frame (depth 5, lib.zig:143) into the pixel buffer. */
void startRender(void);
/* lib.zig:283-293: true once the frame is complete. */
bool tryFinishRender(void);
/* lib.zig:166-178, 295-301: orbit `from` about the axis `up` through `to` by `angle` radians. */
void rotateCamera(double angle);
/* lib.zig:180-190, 303-309: move `from` towards `to` by `distance` x |to - from|. */
void moveCamera(double distance);

/* Not in the reference: where the scene's asset files are read from (see above). */
void rtcRendererSetAssetDir(const char* dir);
/* Not in the reference: kernel time and ray counts of the last finished frame (0 = ok, -1 = no frame yet).
 * `stats` is an rtc_stats (include/rtc_b200.h). */
int rtcRendererLastStats(void* stats);

#ifdef __cplusplus
}
#endif
#endif
