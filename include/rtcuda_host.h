/*
 * rtcuda_host.h — C surface of the C++ host-side mirror (myraytracer_b200/host):
 * RayTracerEngine (Sources/RayTracer/RayTracer.swift:25-205) over the render module of rtcuda.h.
 * Exists because the Swift toolchain is absent: Python tests and bench.py drive the mirror
 * through these calls exactly as Swift code would drive RayTracerEngine.
 */
#ifndef RTCUDA_HOST_H
#define RTCUDA_HOST_H

#include "rtcuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rth_engine rth_engine_t;    /* RayTracerEngine */
typedef struct rth_context rth_context_t;  /* Scene + RTContext(scene:) without a renderer (no GPU needed) */

const char* rth_last_error(void);

/* RayTracerEngine.init(from:) — RayTracer.swift:30-49.  Never NULL unless out of memory: like the
 * reference, a scene that fails to decode leaves an empty engine (rth_engine_ready() == 0). */
rth_engine_t* rth_engine_create(const char* sceneUrl, int32_t device);
/* The same engine with its scene resident on several GPUs (rt_scene_create_multi): render() shards every frame. */
rth_engine_t* rth_engine_create_multi(const char* sceneUrl, const int32_t* devices, int32_t nDevices);
void rth_engine_destroy(rth_engine_t* engine);
int32_t rth_engine_ready(const rth_engine_t* engine);
int32_t rth_engine_camera_count(const rth_engine_t* engine);
int32_t rth_engine_camera(const rth_engine_t* engine, int32_t cameraIndex, rt_camera* out);
rt_scene_t* rth_engine_scene(const rth_engine_t* engine);
int32_t rth_engine_describe(const rth_engine_t* engine, rt_scene_desc* out);
int32_t rth_engine_instance_object_ids(const rth_engine_t* engine, int32_t* out, int64_t capacity);

/* RayTracerEngine.render(format:cameraIndex:progress:) — RayTracer.swift:115-131; error codes
 * -20 / -21 / -10 as the reference's NSErrors (:142-154). */
int32_t rth_engine_render(rth_engine_t* engine, int32_t cameraIndex, const rt_render_params* params, uint8_t* rgbaOut,
                          double* linearOut, rt_hit_id* idsOut, rt_stats* statsOut);

/* RayTracerEngine.inspect(scene:format:) — RayTracer.swift:52-67: counts4 = meshes, triangles, spheres, planes
 * (sceneMeshAndTriangleCounts, :208-227); the CameraSpec fields come from rth_engine_camera / rth_engine_camera_name. */
int32_t rth_engine_inspect(const rth_engine_t* engine, int64_t* counts4, int32_t* camerasOut);
int32_t rth_engine_camera_name(const rth_engine_t* engine, int32_t cameraIndex, char* imageNameOut, int32_t capacity,
                               int32_t* idOut);

/* Animation.  The reference's batch driver builds a new engine per frame file
 * (BatchRender/BatchRender.swift:26-40) and its RTContext.refitTLAS (RTContext.swift:883-928) is never called; here a
 * frame is a new set of instance transforms on a resident scene: n = instance count, 16 doubles (column-major
 * localToWorld) per instance in RTContext order (analytic primitives, mesh instances, base meshes).  worldToLocal,
 * normal matrix and world bounds are re-derived as refitTLAS does, then rt_scene_update_instances refits the TLAS.
 * boundsOnDevice != 0: the world bounds of mesh instances (a union over every triangle) come from the device kernel
 * instead of the host loop — same values (tested), O(instances) host work. */
int32_t rth_engine_set_instance_transforms(rth_engine_t* engine, const double* localToWorld, int64_t n, int32_t boundsOnDevice);

/* Headless replacement of makeCGImage + PNG output (Helpers/Image.swift; ImageIO is Apple-only): writes the RGBA8
 * frame as .png (8-bit RGBA, stored deflate) or .ppm (binary P6), chosen by the extension.  Row 0 = top. */
int32_t rth_image_write(const char* path, const uint8_t* rgba, int32_t width, int32_t height);

rth_context_t* rth_context_create(const char* sceneUrl);
void rth_context_destroy(rth_context_t* ctx);
int32_t rth_context_describe(const rth_context_t* ctx, rt_scene_desc* out);

#ifdef __cplusplus
}
#endif
#endif /* RTCUDA_HOST_H */
