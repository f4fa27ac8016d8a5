/* tests/c_driver/driver.c — the drop-in boundary driven from plain C, the way the reference's Swift package would bind
 * it through a module map: RayTracerEngine.init(from:) -> render(cameraIndex:) -> RGBA8 -> image file
 * (Sources/RayTracer/RayTracer.swift:30-49,115-205), using nothing but include/rtcuda.h and include/rtcuda_host.h.
 * Also goes below the engine once: the flattened RTContext (rt_scene_desc) straight into rt_scene_create / rt_render.
 *
 *   driver <scene.rtscene> <out.ppm|out.png>
 * prints "frame W H fnv=<hash of the RGBA bytes> rays=<n>" for the engine path and "module fnv=<hash>" for the direct
 * one; exit code 0 only when both frames are identical. */
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rtcuda_host.h"

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
    const int same = memcmp(rgba, rgba2, bytes) == 0;
    rt_scene_destroy(scene);
    rth_context_destroy(ctx);
    rth_engine_destroy(eng);
    free(rgba); free(rgba2);
    return same ? 0 : 9;
}
