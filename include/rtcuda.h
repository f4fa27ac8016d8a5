/*
 * rtcuda.h — C ABI of the B200-native render module ("CRTCuda").
 *
 * Drop-in boundary for MyRayTracer's render loop: it replaces the body of
 *   Renderer.render(scene:cameraIndex:progressRow:) -> [Vec3]
 *       (Sources/RayTracer/Extensions/Object+Extension.swift:52-379)
 * and the RGBA8 conversion that follows it
 *       (Sources/RayTracer/RayTracer.swift:186-195),
 * fed by the arrays of RTContext (Sources/RayTracer/Models/RTContext.swift:74-92).
 *
 * Conventions copied from the reference's only native module, CPly
 * (Sources/CPly/include/PLYReaderWrapper.h:18-69, Sources/CPly/wrapper.cpp:17-56):
 * pure C, <stdint.h> only, opaque handle, create() returns NULL on failure and
 * never throws, destroy(NULL) is safe, every call tolerates a NULL handle,
 * caller-owned output buffers, borrowed input pointers (valid for the duration
 * of the call only), enums cross as int.  Status: 0 = ok, negative = error;
 * rt_last_error() returns a thread-local message.
 *
 * All vectors are FP64 as in the reference (Vec3 = SIMD3<Double>,
 * RTContext.swift:13-16); matrices are column-major like simd_double4x4.
 */
#ifndef RTCUDA_H
#define RTCUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2

/* status codes; -10/-20/-21 mirror the reference's NSError codes (RayTracer.swift:121-123,142-154) */
enum {
    RT_OK = 0,
    RT_ERR_INVALID = -1,       /* bad argument / inconsistent description */
    RT_ERR_CUDA = -2,          /* CUDA runtime failure (message in rt_last_error) */
    RT_ERR_UNSUPPORTED = -3,   /* feature outside the implemented path (fails loudly, never falls back) */
    RT_ERR_CANCELLED = -4,     /* progress callback asked to stop */
    RT_ERR_BAD_CAMERA = -10,   /* "Invalid camera index" */
    RT_ERR_NO_SCENE = -20,     /* "No scene loaded. Can't render." */
    RT_ERR_NO_RENDERER = -21   /* "Renderer not initialized." */
};

/* Material.type strings of the reference ("mirror", "dielectric", "conductor", anything else);
 * Object+Extension.swift:189,207,252 */
enum { RT_MAT_PLAIN = 0, RT_MAT_MIRROR = 1, RT_MAT_DIELECTRIC = 2, RT_MAT_CONDUCTOR = 3 };

/* GeometryType raw values, Accelearion/BVH.swift:33-38 */
enum { RT_GEO_MESH = 0, RT_GEO_TRIANGLE = 1, RT_GEO_SPHERE = 2, RT_GEO_PLANE = 3 };

/* HitKind (ParsingKit): .none/.triangle/.sphere/.plane */
enum { RT_HIT_NONE = 0, RT_HIT_TRIANGLE = 1, RT_HIT_SPHERE = 2, RT_HIT_PLANE = 3 };

/* ParsingKit Material fields read by trace (Object+Extension.swift:105-274) */
typedef struct rt_material {
    int32_t type;
    int32_t reserved;
    double ambient[3], diffuse[3], specular[3], mirror[3], absorption[3];
    double phong, ior, absorptionIndex, roughness;
} rt_material;

typedef struct rt_point_light { double position[3], intensity[3]; } rt_point_light;       /* :118-140 */
typedef struct rt_area_light { double position[3], normal[3], size, radiance[3]; } rt_area_light; /* :145-186 */
typedef struct rt_sphere { double center[3], radius; } rt_sphere;                         /* RTContext.swift:513-527 */
typedef struct rt_plane { double center[3], normal[3]; } rt_plane;                         /* RTContext.swift:530-538 */

/* PrimitiveInfo (BVH.swift:42-64) minus bounds/centroid, which are recomputed */
typedef struct rt_prim {
    int32_t type;            /* RT_GEO_TRIANGLE / SPHERE / PLANE */
    int32_t smooth;          /* ShadingMode: 0 flat, 1 smooth */
    int64_t primitiveIndex;  /* into triangles / spheres / planes */
    int64_t materialID;
} rt_prim;

/* BLAS.prims (RTContext.swift:34-42): a contiguous range of rt_scene_desc.prims.
 * The reference's own BVH nodes are NOT passed: the module builds its GPU layout itself. */
typedef struct rt_blas { int64_t firstPrim, primCount; } rt_blas;

/* Instance (RTContext.swift:43-61) */
typedef struct rt_instance {
    int32_t blas;                 /* index into rt_scene_desc.blas */
    int32_t hasMaterialOverride;  /* materialOverride != nil */
    int64_t materialOverride;
    double localToWorld[16], worldToLocal[16];
    double normalMatrix[9];       /* simd_double3x3, column-major */
    double worldBounds[6];        /* minP, maxP (Instance.worldBounds) */
    double instanceMotion[3];
} rt_instance;

/* Flattened RTContext + the Scene fields trace reads (Object+Extension.swift:88,97,102,116) */
typedef struct rt_scene_desc {
    uint32_t abiVersion;          /* RT_ABI_VERSION */
    uint32_t reserved;
    int64_t nTriangles;
    const double* triVertices;    /* nTriangles*9: v0,v1,v2 (Triangle.v0/v1/v2) */
    const double* triNormals;     /* nTriangles*9: n0,n1,n2 */
    const double* triMotionBlur;  /* nTriangles*3 or NULL (all zero) */
    int64_t nSpheres;  const rt_sphere* spheres;
    int64_t nPlanes;   const rt_plane* planes;
    int64_t nPrims;    const rt_prim* prims;
    int64_t nBlas;     const rt_blas* blas;
    int64_t nInstances; const rt_instance* instances;
    int64_t nMaterials; const rt_material* materials;
    int64_t nPointLights; const rt_point_light* pointLights;
    int64_t nAreaLights;  const rt_area_light* areaLights;
    double ambient[3];            /* scene.lights.ambient */
    double background[3];         /* scene.backgroundColor */
    double intersectionTestEpsilon, shadowRayEpsilon;
    int32_t maxRecursionDepth;
    int32_t buildMode;            /* RT_BUILD_* : who builds the mesh BVHs */
} rt_scene_desc;

/* ParsingKit Camera fields read by render/makeCameraBasis (Object+Extension.swift:58-63,385-425) */
typedef struct rt_camera {
    int32_t width, height;        /* imageResolution */
    int32_t numSamples;
    int32_t isLookAt;             /* type?.lowercased() == "lookat" */
    int32_t hasFovy;
    int32_t reserved;
    double position[3], gaze[3], gazePoint[3], up[3];
    double nearPlane[4];          /* l, r, b, t */
    double nearDistance, fovy;
    double apertureSize, focusDistance;
} rt_camera;

/* progress: rows finished so far (multiples of the 8-row chunk of Object+Extension.swift:75); return non-zero to
 * cancel (RayTracer.swift:178-180).  Called from the thread that called rt_render, WHILE the frame is in flight, like
 * ProgressRelay reports finished chunks (Object+Extension.swift:22-41): rt_render reads the kernels' work counter
 * through a side stream every ~100 us and reports the rows the finished share of the work amounts to (the kernels draw
 * tiles in cost order, not top to bottom, so "rows" is the reference's unit, not a scanline position).  A non-zero
 * return stops the frame: the work counter is pushed past its end, every warp exits at its next draw (< 1 ms), and
 * rt_render returns RT_ERR_CANCELLED with the output buffers undefined. */
typedef int (*rt_progress_fn)(int32_t rowsDone, int32_t rowsTotal, void* user);

/* rt_scene_desc.buildMode.  The reference builds on the CPU (Accelearion/BVH.swift:78-250); the image does not depend
 * on the tree.  SAH (default): this module's binned-SAH tree (16 bins, leaf-cost rule, <= 4 triangles per leaf), best
 * traversal speed — built by the device builder for meshes of RT_BUILD_AUTO_DEVICE_TRIANGLES triangles and more (3x
 * faster on a million triangles and the SAME node array byte for byte), by the multi-threaded host builder below
 * that.  DEVICE_LBVH: Morton-order radix tree built by CUDA kernels, faster still to build, 66 % slower to traverse; a
 * tree too deep for the traversal stack falls back to the SAH builder for that mesh.  DEVICE_SAH: every mesh through
 * the device SAH builder (the host builder's own splits and leaves, made level by level by CUDA kernels, per-bin
 * boxes through order-preserving 64-bit atomics).  HOST_SAH: every mesh through the host builder.  The environment
 * variable RTCUDA_BUILD=auto|device|device-sah|host overrides the field. */
#define RT_BUILD_SAH         0
#define RT_BUILD_DEVICE_LBVH 1
#define RT_BUILD_DEVICE_SAH  2
#define RT_BUILD_HOST_SAH    3
#define RT_BUILD_AUTO_DEVICE_TRIANGLES 65536

#define RT_SHARD_ROWS  0
#define RT_SHARD_TILES 1
#define RT_SCHEDULE_ADAPTIVE 0
#define RT_SCHEDULE_SPATIAL  1

typedef struct rt_render_params {
    /* One shard of the frame is rendered, the other pixels are left untouched; shardCount <= 1 renders the whole
     * frame.  (Multi-process sharding: one rank per GPU.)  Which pixels: see shardMode below. */
    int32_t shardIndex, shardCount;
    int32_t collectCounters;      /* 1: also fill rt_stats box/primitive test counts (slower kernel variant) */
    int32_t kernelVariant;        /* 0 = automatic.  For measurements only: 1 = fast kernel (lane = (pixel, sample), warp
                                   * packets at depth 0; the default when no material is glossy; RT_ERR_UNSUPPORTED on
                                   * a scene with roughness != 0), 2 = general kernel (lane = pixel, sequential RNG). */
    rt_progress_fn progress;      /* may be NULL */
    void* progressUser;
    /* Order in which the warps draw the 8x4-pixel tiles of the frame.  RT_SCHEDULE_ADAPTIVE (0, default): every frame
     * records the time spent per tile, and the next frame with the same image size, sample count, shard and kernel
     * draws the longest tiles first, so that it ends on its cheapest work instead of waiting for a few slow packets
     * (rt_stats.tilesScheduled tells whether a cost order was used).  The first frame of a geometry — the reference's
     * usage: one render per camera, RayTracer.swift:169 — runs in the spatial order (RTCUDA_PROBE=1: it takes its order
     * from one probe ray per tile traced just before the frame; measured slower than the spatial order).
     * RT_SCHEDULE_SPATIAL (1): always top to bottom.  The image is the same either way. */
    int32_t scheduleMode;
    /* RT_SHARD_ROWS (0): the 8-row chunks c (Object+Extension.swift:75) with c % shardCount == shardIndex.
     * RT_SHARD_TILES (1): the 8x4-pixel tiles (tx, ty) with (tx + ty) % shardCount == shardIndex — every shard samples
     * the whole image, so the shards finish together whatever the image shows (for frames written in place into one
     * shared buffer, rt_shared_buffer_*). */
    int32_t shardMode;
} rt_render_params;

/* identity of one primary ray's closest hit (the reference's Hit has no primitive id) */
typedef struct rt_hit_id {
    int32_t inst;                 /* instance index, -1 = miss */
    int32_t kind;                 /* RT_HIT_* */
    int64_t prim;                 /* primitiveIndex within its kind's array, -1 = miss */
    double t;
} rt_hit_id;

/* RenderStats (Models/RenderStats.swift:8-23) plus the ray counts the reference never fills (RayTracer.swift:168) */
typedef struct rt_stats {
    int64_t meshes, triangles, spheres, planes;
    int64_t rays;                 /* raysPrimary + raysShadow + raysSecondary */
    int64_t milliseconds;         /* wall clock of the call, like RenderStats.milliseconds */
    int64_t raysPrimary, raysShadow, raysSecondary;
    double deviceMs;              /* CUDA-event time of the render kernels */
    int64_t kernelLaunches;       /* kernels launched by this call */
    /* only with collectCounters: */
    int64_t boxTests, triTests, sphereTests, planeTests, instEntries, maxStack;
    int64_t smoothHits;           /* closest hits resolved on smooth-shaded triangles (3 normals fetched) */
    int64_t tilesScheduled;       /* tiles drawn in cost order — measured on the previous frame, or estimated by the probe rays of a first frame (0: spatial order) */
} rt_stats;

typedef struct rt_scene_info {
    int64_t triangles, spheres, planes, instances, blasCount;
    int64_t blasNodes, tlasNodes, maxLeafSize, maxDepth;
    int64_t deviceBytes;
    double buildMs, uploadMs;
    int32_t device;
    int32_t deviceBuiltBlas;      /* mesh BVHs built by a device builder (RT_BUILD_DEVICE_*) */
    int32_t deviceCount;          /* GPUs the scene is resident on (rt_scene_create_multi), 1 otherwise */
    int32_t wideNodes;            /* nodes of the wide BVHs the beam walk of coherent packets uses (0: none built) */
    double replicateMs;           /* device-to-device copies of the scene onto the other GPUs */
} rt_scene_info;

typedef struct rt_scene rt_scene_t;

int32_t rt_abi_version(void);
int32_t rt_device_count(void);
const char* rt_last_error(void);

/* Build the GPU acceleration structures from the flattened context and upload them to `device`.
 * Replaces what Renderer.init(ctx:) keeps (Object+Extension.swift:13-19).  NULL on failure. */
rt_scene_t* rt_scene_create(const rt_scene_desc* desc, int32_t device);
void rt_scene_destroy(rt_scene_t* scene);

/* The same scene resident on several GPUs of this process ("rt_render is blocking, internally
 * asynchronous on CUDA streams across all selected GPUs"; 8e).  The acceleration structures are built ONCE and uploaded
 * to devices[0]; the other GPUs receive device-to-device copies (NVLink peer copies), nothing is rebuilt.  rt_render on
 * such a handle deals the frame's 8-row chunks (Object+Extension.swift:75-82) round-robin to the GPUs, runs every
 * GPU's kernel on its own stream from its own host thread, and every GPU copies ITS rows straight into the caller's
 * host buffers over its own PCIe link; no collective, no gather through one GPU.  rt_scene_update_instances updates
 * every replica.  rt_render_device / rt_render_finish address devices[0] only.  NULL on failure. */
rt_scene_t* rt_scene_create_multi(const rt_scene_desc* desc, const int32_t* devices, int32_t nDevices);
int32_t rt_scene_device_count(const rt_scene_t* scene);

/* Page-locked host memory for rt_render's output buffers (optional): every GPU of a multi-device scene then writes its
 * rows by asynchronous DMA; pageable buffers work too, through the driver's staging.  Swift side:
 * UnsafeMutableRawPointer, freed with rt_host_free (like the caller-owned buffers of ply_reader_extract_properties,
 * Sources/CPly/include/PLYReaderWrapper.h:49-53, but allocated by the module). */
void* rt_host_alloc(uint64_t bytes);
void rt_host_free(void* p);

int32_t rt_scene_get_info(const rt_scene_t* scene, rt_scene_info* out);
/* Animation: new placements for the instances of a resident scene.  Replaces
 * RTContext.refitTLAS (Models/RTContext.swift:883-928) + BVHBuilder.refit (Accelearion/BVH.swift:254-292): `instances`
 * has the scene's instance count, every entry keeps its `blas`; localToWorld / worldToLocal / normalMatrix /
 * materialOverride are taken as given.  worldBounds too — except where the box is inverted (minP > maxP, e.g.
 * +inf / -inf): that mesh instance's bounds are derived ON THE DEVICE from the BLAS's resident triangles (the union of
 * their transformed boxes, bit-identical to the loop of refitTLAS step 1, which costs the reference O(triangles) host
 * work per instance and frame) and written back into instances[i].worldBounds.  The instance records go to the
 * device and the TLAS boxes are refitted there, bottom-up, with the topology kept (kernels on `cudaStream`, NULL =
 * default stream); the mesh BVHs, triangles and materials are untouched.  Blocking.  Scenes with instanceMotion are
 * refused (RT_ERR_UNSUPPORTED): their TLAS is the reference's own topology, rebuilt per scene. */
int32_t rt_scene_update_instances(rt_scene_t* scene, rt_instance* instances, int64_t nInstances, void* cudaStream);

/* The tile order the next frame of the last rendered geometry will use (rt_render_params.scheduleMode): returns the
 * number of tiles (0: nothing learnt yet, < 0: error) and copies min(n, capacity) entries — slot k draws tile
 * orderOut[k], tiles numbered row-major within the shard (8x4 pixels each).  No reference analogue; for tests. */
int64_t rt_scene_get_schedule(rt_scene_t* scene, int32_t* orderOut, int64_t capacity);

/* The mesh BVHs as they lie on the device: 64-byte node records (two conservative FP32 child boxes + two child
 * references, csrc/gpu_layout.h: GNode), all BLASes concatenated.  Returns the size in bytes (< 0: error) and copies
 * min(size, capacityBytes).  No reference analogue; for tests (the builders of rt_scene_desc.buildMode compared). */
int64_t rt_scene_read_blas_nodes(rt_scene_t* scene, void* nodesOut, int64_t capacityBytes);
/* The wide BVHs of the beam walk (built when RTCUDA_WIDE=1; csrc/gpu_layout.h: WNode, 16 children of 32 bytes: FP32
 * box, child reference, sort axis in child 0), all BLASes concatenated; rt_scene_info.wideNodes of them.  Same calling
 * convention.  No reference analogue; for tests (the collapse of the binary tree checked leaf by leaf). */
int64_t rt_scene_read_wide_nodes(rt_scene_t* scene, void* nodesOut, int64_t capacityBytes);

/* Render one camera.  Blocking; one render at a time per scene handle (not re-entrant), different scenes may render
 * from different host threads at once (each scene has its own non-blocking stream).  Host buffers, all optional
 * except one of rgbaOut/linearOut:
 *   rgbaOut   width*height*4 bytes, row 0 = top (RayTracer.swift:186-195)
 *   linearOut width*height*3 doubles, the [Vec3] Renderer.render returns (0..255 scale)
 *   idsOut    width*height*n*n records, n = floor(sqrt(max(1,numSamples)))
 * Device->host copies are part of the call. */
int32_t rt_render(rt_scene_t* scene, const rt_camera* camera, const rt_render_params* params, uint8_t* rgbaOut,
                  double* linearOut, rt_hit_id* idsOut, rt_stats* statsOut);

/* Same, but outputs stay in DEVICE memory owned by the caller (pointers valid on the scene's device)
 * and the launch is enqueued on `cudaStream` (a cudaStream_t, NULL = default stream) without
 * synchronising; statsOut->deviceMs and ray counts are filled by rt_render_finish(). */
int32_t rt_render_device(rt_scene_t* scene, const rt_camera* camera, const rt_render_params* params,
                         void* dRgbaOut, void* dLinearOut, void* dIdsOut, void* cudaStream);
int32_t rt_render_finish(rt_scene_t* scene, rt_stats* statsOut);

/* PLY -> device ingest.  The reference reads a mesh through miniply behind the CPly C ABI
 * (ply_reader_load_element / ply_reader_extract_properties / ply_reader_extract_list, Sources/CPly/include/
 * PLYReaderWrapper.h:38-66, driven by Helpers/PLYReader.swift:54-210) and turns it into Triangle objects in the mesh
 * branch of RTContext.init (Models/RTContext.swift:251-363), all on one thread.  Here the caller hands over the RAW
 * element blocks of a binary_little_endian file plus their layout, and kernels produce the per-triangle FP64 arrays
 * rt_scene_desc takes — bit-identical to the reference's loops (positions read as Float and widened, file normals
 * normalised, quads split 0,1,3 / 2,3,1, smooth vertex normals = sum of unit face normals in face order,
 * flat normals normalised twice). */
typedef struct rt_ply_layout {
    int64_t vertexCount, faceCount;
    int32_t vertexStride;         /* bytes of one vertex record */
    int32_t posOffset[3];         /* byte offsets of the float32 properties x, y, z in a vertex record */
    int32_t normalOffset[3];      /* float32 nx, ny, nz; -1: the file has no normals */
    int32_t faceLeadBytes;        /* scalar properties before the index list in a face record */
    int32_t faceCountBytes;       /* size of the list's count (uchar = 1) */
    int32_t faceIndexBytes;       /* size of one index (int / uint = 4) */
    int32_t faceTrailBytes;       /* scalar properties after the index list */
    int32_t verticesPerFace;      /* 3 or 4: every face record has this count (verified on the device) */
    int32_t reserved;
} rt_ply_layout;

typedef struct rt_ply_stats { int64_t triangles, kernelLaunches; double deviceMs; } rt_ply_stats;

/* triVerticesOut / triNormalsOut: faceCount * (verticesPerFace == 4 ? 2 : 1) * 9 doubles each, caller-owned host
 * buffers.  smooth: the mesh's ShadingMode.  RT_ERR_INVALID: a face index out of range or a face with another vertex
 * count (message in rt_last_error).  Blocking. */
int32_t rt_ply_ingest(const rt_ply_layout* layout, const void* vertexBlock, const void* faceBlock, int32_t smooth,
                      int32_t device, double* triVerticesOut, double* triNormalsOut, rt_ply_stats* statsOut);

/* rt_ply_ingest without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernels, same results bit for bit, but:
 *   - dVertexBlock / dFaceBlock / dTriVerticesOut / dTriNormalsOut are DEVICE pointers owned by the caller (valid on
 *     `device`; sizes as for rt_ply_ingest), the kernels run on `cudaStream` (a cudaStream_t, NULL = default stream);
 *   - every intermediate lives in one fixed device array inside the library (RT_PLY_STATIC_SCRATCH_BYTES, held by every
 *     GPU that loads the library until the process exits), so a mesh may have at most RT_PLY_STATIC_MAX_VERTICES
 *     vertices and RT_PLY_STATIC_MAX_TRIANGLES triangles (quads count twice): larger meshes are refused with
 *     RT_ERR_UNSUPPORTED and a message — use rt_ply_ingest for them;
 *   - the call allocates and frees nothing (no cudaMalloc / cudaFree / event or stream creation); the first call on a
 *     device looks the array's address up once;
 *   - all calls share the array, which keeps the last call's contents (the call clears what must start at zero): two
 *     calls that overlap on one GPU — from two threads or two streams — RACE on it.  One call at a time per GPU.
 * Blocking (the face indices are validated before anything dereferences them).  statsOut->deviceMs is 0. */
#define RT_PLY_STATIC_MAX_VERTICES 65536
#define RT_PLY_STATIC_MAX_TRIANGLES 131072
#define RT_PLY_STATIC_SCRATCH_BYTES (10u << 20)
int32_t rt_ply_ingest_static(const rt_ply_layout* layout, const void* dVertexBlock, const void* dFaceBlock, int32_t smooth,
                             int32_t device, double* dTriVerticesOut, double* dTriNormalsOut, void* cudaStream,
                             rt_ply_stats* statsOut);

/* Multi-GPU framebuffer (one process per GPU): the reference merges finished 8-row chunks into
 * one [Vec3] array by plain copies (Object+Extension.swift:363-373).  Here rank 0 owns the RGBA8 frame in device
 * memory and exports it; every other rank maps it over NVLink (CUDA IPC, peer access) and passes the mapped
 * pointer as dRgbaOut to rt_render_device, so its kernel stores finished pixels straight into rank 0's frame —
 * the gather is fused into the render kernel's epilogue, no staging copy and no collective on the data path.
 *   create: cudaMalloc on `device`, zero-filled; handleOut receives the 64-byte IPC handle to send to the peers
 *   open:   map a peer's buffer into this process on `device`
 *   close:  unmap (opened) or free (created) */
#define RT_IPC_HANDLE_BYTES 64
int32_t rt_shared_buffer_create(int32_t device, uint64_t bytes, void** devPtrOut, uint8_t* handleOut);
int32_t rt_shared_buffer_open(int32_t device, const uint8_t* handle, void** devPtrOut);
int32_t rt_shared_buffer_close(int32_t device, void* devPtr, int32_t wasOpened);

#ifdef __cplusplus
}
#endif
#endif /* RTCUDA_H */
