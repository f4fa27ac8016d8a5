/* tests/c_driver/driver.c — the drop-in boundary driven from plain C, the way the reference's Swift package would bind
 * it through a module map: RayTracerEngine.init(from:) -> render(cameraIndex:) -> RGBA8 -> image file
 * (Sources/RayTracer/RayTracer.swift:30-49,115-205), using nothing but include/rtcuda.h and include/rtcuda_host.h.
 * Also goes below the engine once: the flattened RTContext (rt_scene_desc) straight into rt_scene_create / rt_render.
 *
 *   driver <scene.rtscene> <out.ppm|out.png> [devices]
 * prints "frame W H fnv=<hash of the RGBA bytes> rays=<n>" for the engine path and "module fnv=<hash>" for the direct
 * one; exit code 0 only when both frames are identical.
 * devices >= 2: the scene once more through rt_scene_create_multi on GPUs 0..devices-1, one rt_render call into a
 * page-locked buffer of rt_host_alloc: "multi <n> fnv=<hash> rays=<n>"; that frame must equal the others too.
 * Always: one frame with a progress callback that counts its calls ("progress calls=<k> last=<rows>"), and one that
 * cancels at its first in-flight report ("cancel rc=<code> ms=<time from the request to the return>"). */
#define _POSIX_C_SOURCE 199309L
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <time.h>

#include "rtcuda_host.h"

static double now_ms(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

struct progress_state { int calls, last, total, cancel_at_call; double cancel_ms; };
static int on_progress(int32_t rows, int32_t total, void* user) {
    struct progress_state* p = (struct progress_state*)user;
    p->calls++; p->last = rows; p->total = total;
    if (p->cancel_at_call > 0 && p->calls >= p->cancel_at_call && rows > 0 && rows < total) { p->cancel_ms = now_ms(); return 1; }
    return 0;
}

static uint64_t fnv1a(const uint8_t* p, size_t n) {
    uint64_t h = 1469598103934665603ULL;
    for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ULL; }
    return h;
}

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s scene out-image\n", argv[0]); return 2; }
    if (rt_abi_version() != RT_ABI_VERSION) { fprintf(stderr, "header / library ABI mismatch\n"); return 2; }

    rth_engine_t* eng = rth_engine_create(argv[1], 0);
    if (!eng || !rth_engine_ready(eng)) { fprintf(stderr, "engine: %s\n", rth_last_error()); return 3; }
    rt_camera cam;
    if (rth_engine_camera(eng, 0, &cam) != RT_OK) { fprintf(stderr, "camera: %s\n", rth_last_error()); return 3; }
    const size_t bytes = (size_t)cam.width * cam.height * 4;
    uint8_t* rgba = (uint8_t*)calloc(bytes, 1);
    uint8_t* rgba2 = (uint8_t*)calloc(bytes, 1);
    rt_stats st;
    int32_t rc = rth_engine_render(eng, 0, NULL, rgba, NULL, NULL, &st);
    if (rc != RT_OK) { fprintf(stderr, "render: %d %s\n", rc, rth_last_error()); return 4; }
    printf("frame %d %d fnv=%016" PRIx64 " rays=%" PRId64 "\n", cam.width, cam.height, fnv1a(rgba, bytes), st.rays);
    if (rth_image_write(argv[2], rgba, cam.width, cam.height) != RT_OK) { fprintf(stderr, "cannot write %s\n", argv[2]); return 5; }
    /* the reference's error for a camera that does not exist (RayTracer.swift:160-168) */
    if (rth_engine_render(eng, 99, NULL, rgba2, NULL, NULL, NULL) != RT_ERR_BAD_CAMERA) { fprintf(stderr, "camera 99 accepted\n"); return 6; }

    /* below the engine: RTContext(scene:) flattened -> the render module */
    rth_context_t* ctx = rth_context_create(argv[1]);
    rt_scene_desc desc;
    if (!ctx || rth_context_describe(ctx, &desc) != RT_OK) { fprintf(stderr, "context: %s\n", rth_last_error()); return 7; }
    rt_scene_t* scene = rt_scene_create(&desc, 0);
    if (!scene) { fprintf(stderr, "scene: %s\n", rt_last_error()); return 7; }
    rc = rt_render(scene, &cam, NULL, rgba2, NULL, NULL, NULL);
    if (rc != RT_OK) { fprintf(stderr, "rt_render: %d %s\n", rc, rt_last_error()); return 8; }
    printf("module fnv=%016" PRIx64 "\n", fnv1a(rgba2, bytes));
    int same = memcmp(rgba, rgba2, bytes) == 0;

    /* progress while the frame is in flight, then cancellation (RayTracer.swift:172-181) */
    {
        struct progress_state ps = {0, 0, 0, 0, 0.0};
        rt_render_params prm;
        memset(&prm, 0, sizeof prm);
        prm.progress = on_progress; prm.progressUser = &ps;
        rc = rt_render(scene, &cam, &prm, rgba2, NULL, NULL, NULL);
        if (rc != RT_OK || ps.last != cam.height || ps.total != cam.height) { fprintf(stderr, "progress: rc %d last %d\n", rc, ps.last); return 10; }
        same = same && memcmp(rgba, rgba2, bytes) == 0;
        printf("progress calls=%d last=%d\n", ps.calls, ps.last);
        struct progress_state pc = {0, 0, 0, 2, 0.0};       /* call 1 is the report of row 0 before the launch */
        prm.progressUser = &pc;
        rc = rt_render(scene, &cam, &prm, rgba2, NULL, NULL, NULL);
        const double t1 = now_ms();
        /* a frame too short to be caught in flight completes (rc 0); otherwise it must be cut short */
        if (pc.cancel_ms > 0 && rc != RT_ERR_CANCELLED) { fprintf(stderr, "cancel: rc %d\n", rc); return 11; }
        printf("cancel rc=%d ms=%.3f\n", rc, pc.cancel_ms > 0 ? t1 - pc.cancel_ms : -1.0);
        rc = rt_render(scene, &cam, NULL, rgba2, NULL, NULL, NULL);      /* the handle is usable after a cancelled frame */
        if (rc != RT_OK || memcmp(rgba, rgba2, bytes) != 0) { fprintf(stderr, "frame after cancel differs (rc %d)\n", rc); return 12; }
    }

    const int ndev = argc > 3 ? atoi(argv[3]) : 1;
    if (ndev >= 2) {
        int32_t devs[64];
        for (int k = 0; k < ndev && k < 64; k++) devs[k] = k;
        rt_scene_t* multi = rt_scene_create_multi(&desc, devs, ndev);
        if (!multi || rt_scene_device_count(multi) != ndev) { fprintf(stderr, "multi: %s\n", rt_last_error()); return 13; }
        uint8_t* pinned = (uint8_t*)rt_host_alloc(bytes);
        if (!pinned) { fprintf(stderr, "rt_host_alloc: %s\n", rt_last_error()); return 13; }
        memset(pinned, 0, bytes);
        rt_stats ms;
        for (int k = 0; k < 2; k++) {                                     /* second frame: learnt tile order */
            rc = rt_render(multi, &cam, NULL, pinned, NULL, NULL, &ms);
            if (rc != RT_OK) { fprintf(stderr, "multi rt_render: %d %s\n", rc, rt_last_error()); return 14; }
            same = same && memcmp(rgba, pinned, bytes) == 0;
        }
        printf("multi %d fnv=%016" PRIx64 " rays=%" PRId64 "\n", ndev, fnv1a(pinned, bytes), ms.rays);
        same = same && ms.rays == st.rays;
        rt_host_free(pinned);
        rt_scene_destroy(multi);
    }
    rt_scene_destroy(scene);
    rth_context_destroy(ctx);
    rth_engine_destroy(eng);
    free(rgba); free(rgba2);
    return same ? 0 : 9;
}
