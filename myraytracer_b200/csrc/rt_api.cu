// rt_api.cu — implementation of the C ABI declared in include/rtcuda.h.
//
// rt_scene_create : flattened RTContext (Models/RTContext.swift:74-92) -> GPU layout
//                   (replaces BVHBuilder for meshes, Accelearion/BVH.swift:78-250, and buildTLAS,
//                   RTContext.swift:459-474, with the builders of bvh_build.cpp)
// rt_render       : Renderer.render (Extensions/Object+Extension.swift:52-379) + RGBA8 conversion
//                   (RayTracer.swift:186-195); camera set-up = makeCameraBasis (:382-427) on the host.
// There is no CPU fallback: without a CUDA device every entry point fails with RT_ERR_CUDA.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rtcuda.h"
#include "bvh_build.h"
#include "rt_kernels.h"

using namespace rtg;

namespace rtg {   // ply_ingest.cu
int plyIngestDevice(const rt_ply_layout& L, const void* vertexBlock, const void* faceBlock, int smooth, double* triVertsOut,
                    double* triNormalsOut, const char** msg, int64_t* launches, double* deviceMs);
int plyIngestStatic(const rt_ply_layout& L, const void* dVertexBlock, const void* dFaceBlock, int smooth, int device,
                    double* dTriVertsOut, double* dTriNormalsOut, cudaStream_t stream, const char** msg, int64_t* launches);
}

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }

#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                                \
            return RT_ERR_CUDA;                                                                        \
        }                                                                                              \
    } while (0)

// One device array of a scene: where its pointer lives inside rt_scene (byte offset) and its size — what
// rt_scene_create_multi needs to replicate the scene on another GPU by device-to-device copies.
struct SceneArray { size_t slotOffset, bytes; };

template <class T>
int upload(const std::vector<T>& v, T** d, int64_t& bytes, std::vector<SceneArray>* reg = nullptr, const void* owner = nullptr) {
    *d = nullptr;
    size_t n = v.size() * sizeof(T);
    CK(cudaMalloc((void**)d, n ? n : 16));
    if (n) CK(cudaMemcpy(*d, v.data(), n, cudaMemcpyHostToDevice));
    bytes += (int64_t)n;
    if (reg) reg->push_back(SceneArray{(size_t)((const char*)d - (const char*)owner), n ? n : 16});
    return RT_OK;
}

template <class T>
int uploadRaw(const T* src, size_t count, T** d, int64_t& bytes, std::vector<SceneArray>* reg, const void* owner) {
    *d = nullptr;
    const size_t n = count * sizeof(T);
    CK(cudaMalloc((void**)d, n ? n : 16));
    if (n) CK(cudaMemcpy(*d, src, n, cudaMemcpyHostToDevice));
    bytes += (int64_t)n;
    if (reg) reg->push_back(SceneArray{(size_t)((const char*)d - (const char*)owner), n ? n : 16});
    return RT_OK;
}

// an empty device array of `count` elements, registered like upload()'s
template <class T>
int deviceArray(T** d, size_t count, int64_t& bytes, std::vector<SceneArray>* reg, const void* owner) {
    *d = nullptr;
    const size_t n = count * sizeof(T);
    CK(cudaMalloc((void**)d, n ? n : 16));
    bytes += (int64_t)n;
    if (reg) reg->push_back(SceneArray{(size_t)((const char*)d - (const char*)owner), n ? n : 16});
    return RT_OK;
}

double nowMs() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// RTCUDA_TRACE=1: phase times of rt_scene_create on stderr (diagnostics)
bool traceOn() { static const bool on = std::getenv("RTCUDA_TRACE") != nullptr; return on; }
void trace(const char* what, double t0) { if (traceOn()) std::fprintf(stderr, "[rtcuda trace] %-32s %8.2f ms\n", what, nowMs() - t0); }

bool isIdentity34(const double* m16) {
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++)
            if (m16[c * 4 + r] != ((c == r) ? 1.0 : 0.0)) return false;
    return true;
}

}  // namespace

static const size_t kFrameStateBytes = 16 * sizeof(unsigned long long) + 64;

// What an instance record needs from its BLAS (kept by the scene: rt_scene_update_instances re-derives the records)
struct BlasInfo {
    int kind = INST_MESH; int64_t prim = 0; int64_t matId = 0; int smooth = 0;
    int32_t nodeBase = 0, triBase = 0, rootRef = kEmptyRef; int f64 = 0;
    int32_t wideRoot = kEmptyRef;      // root in DeviceScene::wnodes, kEmptyRef = this mesh has no wide BVH
    int64_t triCount = 0;              // triangles of a mesh BLAS (slots triBase .. triBase + triCount of tris / trisD)
    double lo[3], hi[3], centre[3];
    double motion[3] = {0, 0, 0};      // Triangle.motionBlur of the mesh (RTContext.swift:280,345)
    double payload[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // sphere c.xyz, r | plane c.xyz, n.xyz
};

// Worker threads of a multi-device scene: one per replica; run(job) executes job(replica index) on all of them.
struct ReplicaPool {
    std::vector<std::thread> threads;
    std::mutex m;
    std::condition_variable cvStart, cvDone;
    std::function<int(int)> job;
    std::vector<int> rc;
    std::vector<std::string> err;
    int generation = 0, remaining = 0;
    bool quit = false;
    void start(int n);
    void launch(std::function<int(int)> j);     // returns at once; wait() joins the generation
    void wait();
    ~ReplicaPool();
};

struct rt_scene {
    int device = 0;
    DeviceScene ds{};
    std::vector<SceneArray> arrays;    // every device array of the scene (what a replica copies)
    // multi-device scenes (rt_scene_create_multi): replicas[0] == this, the others own a copy of the scene on their GPU
    std::vector<rt_scene*> replicas;
    ReplicaPool* pool = nullptr;
    bool isReplica = false;
    // rt_render's own streams: the frame (non-blocking, so scenes rendered from different host threads do not
    // serialise on the legacy default stream) and the progress / cancel side channel
    cudaStream_t stream = nullptr, poll = nullptr;
    cudaEvent_t evReset = nullptr;     // the frame's counters have been zeroed: the work counter may be polled
    int* hPoll = nullptr;              // pinned: [0] work counter read back, [1] the value that cancels a frame
    int64_t pendingUnits = 0;          // work units of the frame in flight: passUnits[0] (counting pass of an area-light frame, else 0) + passUnits[1]
    int64_t passUnits[2] = {0, 0}, passWarps[2] = {0, 0};   // per pass: units / warps drawing them (a warp's last, failing draw counts too)
    rt_scene_info info{};
    int64_t meshes = 0;
    // device allocations
    WNode* dWNodes = nullptr;          // wide BVHs of the beam walk (collapsed from dNodes)
    GNode* dNodes = nullptr; GTri* dTris = nullptr; GTriD* dTrisD = nullptr; double* dNormals = nullptr;
    GNode* dTlas = nullptr; GInstance* dInst = nullptr; GMaterial* dMat = nullptr; GPointLight* dLights = nullptr;
    GAreaLight* dAreaLights = nullptr; double* dJitter = nullptr;
    RefNode* dRefTlas = nullptr; int32_t* dRefTlasPrim = nullptr;
    bool motion = false;               // Mesh.motionBlur or MeshInstance.motionBlur somewhere: ray time matters
    // animation (rt_scene_update_instances): BLAS facts, TLAS topology and its bottom-up order
    std::vector<BlasInfo> blasInfo;
    std::vector<int32_t> instBlas;     // rt_instance.blas of every instance
    bool instanceMotion = false;
    TNode* dTlasD = nullptr;           // FP64 TLAS nodes (topology + boxes), refitted on the device
    int* dRefitOrder = nullptr;        // node indices, deepest level first
    std::vector<int> refitLevelStart;  // level l = dRefitOrder[refitLevelStart[l] .. refitLevelStart[l + 1])
    int* dRefitLevelStart = nullptr;
    double* dInstBounds = nullptr;     // Instance.worldBounds, 6 doubles per instance
    GInstance* hInst = nullptr; double* hInstBounds = nullptr;   // pinned staging for the per-frame upload
    BoundsChunk* dBoundChunks = nullptr; int nBoundChunks = 0;   // (instance, triangle range) work items of launchInstanceBounds
    int32_t* dDerive = nullptr; int32_t* hDerive = nullptr;      // per instance: 1 = derive worldBounds on the device
    unsigned long long* dBoundsEnc = nullptr; unsigned long long* hBoundsEnc = nullptr;
    int* dDefer = nullptr; size_t deferBytes = 0;                // two-launch frames: [0] number of deferred units, then the units
    unsigned* dAreaIndex = nullptr; size_t areaIndexBytes = 0;   // per-pixel jitter index of an area-light frame
    // render state (re-used across calls)
    int* dTileCounter = nullptr; unsigned long long* dCounters = nullptr;
    // tile schedule learnt from the previous frame of the same geometry (rt_kernels.cu: scheduleKernel)
    unsigned* dTileCost = nullptr; size_t tileCostBytes = 0;
    int* dTileOrder = nullptr; size_t tileOrderBytes = 0;
    unsigned* dSchedWork = nullptr; size_t schedWorkBytes = 0;
    int schedKey[8] = {}; bool schedValid = false; int64_t pendingScheduled = 0;
    bool pendingProbe = false;         // the frame in flight took its tile order from probe rays (first frame of a geometry)
    double* dPathStack = nullptr; size_t pathStackBytes = 0;
    uint8_t* dRgba = nullptr; size_t rgbaBytes = 0;
    double* dLinear = nullptr; size_t linearBytes = 0;
    HitId* dIds = nullptr; size_t idsBytes = 0;
    unsigned long long* hCounters = nullptr;  // pinned
    cudaEvent_t evStart = nullptr, evStop = nullptr, evSide = nullptr;
    cudaStream_t side = nullptr;       // epilogue of a frame (counters to the host, tile order of the next frame)
    bool sideBusy = false;
    int blocks[KV_COUNT][2] = {};      // [variant][counters]
    bool glossy = false;               // some material has roughness != 0: RNG draws inside trace
    int64_t bigBlasTriangles = kBigBlasTriangles;   // meshes above this size are walked per lane by incoherent packets (RTCUDA_BIG_BLAS overrides: tests)
    bool recurse = false;              // maxRecursionDepth > 0 and a mirror / dielectric / conductor material: child rays exist
    bool smallGeometry = false;        // no mesh above kBigBlasTriangles: analytic primitives and small meshes (rt_render_kernel.cuh: BLOCKS)
    bool nonFinite = false;            // a light / material / ambient constant is inf or NaN: the non-finite guard of `trace` can fire
    int pendingVariant = 0;
    bool pending = false; bool pendingCounters = false;
    cudaStream_t pendingStream = nullptr;
    int64_t launches = 0;
    int64_t updateLaunches = 0;        // kernels of the last rt_scene_update_instances
};


// Instance (RTContext.swift:43-61) -> device record
static void fillInstance(const rt_instance& in, const BlasInfo& bi, GInstance& g, int64_t bigBlasTriangles = kBigBlasTriangles) {
    std::memset(&g, 0, sizeof g);
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 3; r++) { g.w2l[c * 3 + r] = in.worldToLocal[c * 4 + r]; g.l2w[c * 3 + r] = in.localToWorld[c * 4 + r]; }
    std::memcpy(g.nmat, in.normalMatrix, sizeof g.nmat);
    std::memcpy(g.rootLo, bi.lo, 24); std::memcpy(g.rootHi, bi.hi, 24); std::memcpy(g.centre, bi.centre, 24);
    g.kind = bi.kind;
    g.identity = (isIdentity34(in.worldToLocal) && isIdentity34(in.localToWorld)) ? 1 : 0;
    // simd_determinant of the upper 3x3 (RTContext.swift:691-698), same expansion as the oracle
    const double* m = in.localToWorld;
    const double c0[3] = {m[0], m[1], m[2]}, c1[3] = {m[4], m[5], m[6]}, c2[3] = {m[8], m[9], m[10]};
    const double cx = c1[1] * c2[2] - c1[2] * c2[1], cy = c1[2] * c2[0] - c1[0] * c2[2], cz = c1[0] * c2[1] - c1[1] * c2[0];
    g.flipNormal = ((c0[0] * cx + c0[1] * cy) + c0[2] * cz) < 0.0 ? 1 : 0;
    g.smooth = bi.smooth;
    g.nodeBase = bi.nodeBase; g.triBase = bi.triBase; g.triF64 = bi.f64; g.rootRef = bi.rootRef;
    g.bigBlas = (bi.kind == INST_MESH && bi.triCount > bigBlasTriangles) ? 1 : 0;
    g.wideRoot = bi.kind == INST_MESH ? bi.wideRoot : kEmptyRef;
    {
        double M = 0;
        for (int a = 0; a < 3; a++) M = std::fmax(M, std::fmax(std::fabs(bi.lo[a] - bi.centre[a]), std::fabs(bi.hi[a] - bi.centre[a])));
        g.beamSlack = std::ldexp(M, -15);
    }
    g.matId = in.hasMaterialOverride ? in.materialOverride : bi.matId;  // RTContext.swift:704
    g.primIndex = bi.prim;
    const bool moves = in.instanceMotion[0] != 0 || in.instanceMotion[1] != 0 || in.instanceMotion[2] != 0;
    std::memcpy(g.motion, in.instanceMotion, 24); std::memcpy(g.meshMotion, bi.motion, 24);
    g.hasMotion = (moves || bi.motion[0] != 0 || bi.motion[1] != 0 || bi.motion[2] != 0) ? 1 : 0;
    std::memcpy(g.prim, bi.payload, sizeof g.prim);
}

// Scene bounds, their centre and the padding of the FP32 TLAS boxes from the root node
static double tlasFrame(const TNode& r, double* sLo, double* sHi, double* sC) {
    double M = 0;
    for (int a = 0; a < 3; a++) {
        sLo[a] = std::fmin(r.lo0[a], r.lo1[a]); sHi[a] = std::fmax(r.hi0[a], r.hi1[a]); sC[a] = 0.5 * (sLo[a] + sHi[a]);
        M = std::fmax(M, std::fmax(std::fabs(sLo[a] - sC[a]), std::fabs(sHi[a] - sC[a])));
    }
    return std::fmax(M * std::ldexp(1.0, -19), 1e-30);
}

// Per-device state that is not scene data: frame counters, events, streams, pinned staging (a replica gets its own).
static int allocFrameState(rt_scene* sc, size_t nInstances) {
    const size_t n = nInstances ? nInstances : 1;
    CK(cudaMallocHost((void**)&sc->hInst, n * sizeof(GInstance)));
    CK(cudaMallocHost((void**)&sc->hInstBounds, n * 6 * sizeof(double)));
    CK(cudaMallocHost((void**)&sc->hDerive, n * sizeof(int32_t)));
    CK(cudaMallocHost((void**)&sc->hBoundsEnc, n * 6 * sizeof(unsigned long long)));
    CK(cudaMalloc((void**)&sc->dDerive, n * sizeof(int32_t)));
    CK(cudaMalloc((void**)&sc->dBoundsEnc, n * 6 * sizeof(unsigned long long)));
    // ray / test counters (16 x u64) and the work counter behind them: one allocation, one memset per frame
    CK(cudaMalloc((void**)&sc->dCounters, kFrameStateBytes));
    sc->dTileCounter = reinterpret_cast<int*>(sc->dCounters + 16);
    CK(cudaMallocHost((void**)&sc->hCounters, 16 * sizeof(unsigned long long)));
    CK(cudaMallocHost((void**)&sc->hPoll, 2 * sizeof(int)));
    sc->hPoll[0] = 0; sc->hPoll[1] = 1 << 30;
    CK(cudaEventCreate(&sc->evStart));
    CK(cudaEventCreate(&sc->evStop));
    CK(cudaEventCreateWithFlags(&sc->evSide, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&sc->evReset, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&sc->side, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&sc->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&sc->poll, cudaStreamNonBlocking));
    return RT_OK;
}

// The pointers of DeviceScene are the scene's own device arrays (again after a replica copied them).
static void bindDeviceScene(rt_scene* sc) {
    DeviceScene& S = sc->ds;
    S.nodes = sc->dNodes; S.tris = sc->dTris; S.trisD = sc->dTrisD; S.triNormals = sc->dNormals; S.triVertsD = nullptr;
    S.tlas = sc->dTlas; S.instances = sc->dInst; S.materials = sc->dMat; S.pointLights = sc->dLights;
    S.areaLights = sc->dAreaLights; S.jitter = sc->dJitter;
    S.refTlas = sc->dRefTlas; S.refTlasPrim = sc->dRefTlasPrim;
    S.wnodes = sc->dWNodes;
}

extern "C" {

int32_t rt_abi_version(void) { return RT_ABI_VERSION; }

int32_t rt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* rt_last_error(void) { return g_err.c_str(); }

static int createImpl(const rt_scene_desc* d, int device, rt_scene* sc) {
    if (!d) return fail(RT_ERR_INVALID, "null scene description");
    if (d->abiVersion != RT_ABI_VERSION) return fail(RT_ERR_INVALID, "rt_scene_desc.abiVersion mismatch");
    if (d->nMaterials < 1 || !d->materials) return fail(RT_ERR_INVALID, "at least one material is required");
    if (d->nAreaLights < 0 || (d->nAreaLights > 0 && !d->areaLights)) return fail(RT_ERR_INVALID, "area light array missing");
    // every count non-negative, every array present when its count is not zero — before anything is dereferenced
    if (d->nTriangles < 0 || d->nSpheres < 0 || d->nPlanes < 0 || d->nPrims < 0 || d->nBlas < 0 || d->nInstances < 0 || d->nPointLights < 0)
        return fail(RT_ERR_INVALID, "negative element count in rt_scene_desc");
    if (d->nTriangles > 0 && (!d->triVertices || !d->triNormals)) return fail(RT_ERR_INVALID, "triangle arrays missing");
    if (d->nSpheres > 0 && !d->spheres) return fail(RT_ERR_INVALID, "sphere array missing");
    if (d->nPlanes > 0 && !d->planes) return fail(RT_ERR_INVALID, "plane array missing");
    if (d->nPrims > 0 && !d->prims) return fail(RT_ERR_INVALID, "prim array missing");
    if (d->nBlas > 0 && !d->blas) return fail(RT_ERR_INVALID, "blas array missing");
    if (d->nInstances > 0 && !d->instances) return fail(RT_ERR_INVALID, "instance array missing");
    if (d->nPointLights > 0 && !d->pointLights) return fail(RT_ERR_INVALID, "point light array missing");
    if (d->maxRecursionDepth < 0 || d->maxRecursionDepth > 64) return fail(RT_ERR_INVALID, "maxRecursionDepth outside 0..64");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(RT_ERR_INVALID, "invalid CUDA device index");
    CK(cudaSetDevice(device));
    sc->device = device;
    const double t0 = nowMs();

    // ---- who builds the mesh BVHs (rt_scene_desc.buildMode, overridden by RTCUDA_BUILD)
    if (d->buildMode < RT_BUILD_SAH || d->buildMode > RT_BUILD_HOST_SAH) return fail(RT_ERR_INVALID, "rt_scene_desc.buildMode outside 0..3");
    int buildMode = d->buildMode;
    if (const char* e = std::getenv("RTCUDA_BUILD"))
        buildMode = std::strcmp(e, "device") == 0 ? RT_BUILD_DEVICE_LBVH : std::strcmp(e, "device-sah") == 0 ? RT_BUILD_DEVICE_SAH :
                    std::strcmp(e, "host") == 0 ? RT_BUILD_HOST_SAH : RT_BUILD_SAH;
    bool deviceBuild = buildMode == RT_BUILD_DEVICE_LBVH || buildMode == RT_BUILD_DEVICE_SAH;
    if (buildMode == RT_BUILD_SAH)          // the SAH tree, big meshes on the device (same node array as the host builder's)
        for (int64_t b = 0; b < d->nBlas && d->blas; b++) deviceBuild = deviceBuild || d->blas[b].primCount >= RT_BUILD_AUTO_DEVICE_TRIANGLES;
    if (const char* e = std::getenv("RTCUDA_BIG_BLAS")) sc->bigBlasTriangles = std::atoll(e);
    struct DevVerts {       // the scene's triangle vertices on the device, for the device builder only
        double* p = nullptr;
        ~DevVerts() { cudaFree(p); }
    } dVerts;
    if (deviceBuild && d->nTriangles > 0) {
        CK(cudaMalloc((void**)&dVerts.p, (size_t)d->nTriangles * 9 * sizeof(double)));
        CK(cudaMemcpy(dVerts.p, d->triVertices, (size_t)d->nTriangles * 9 * sizeof(double), cudaMemcpyHostToDevice));
        trace("create: vertices -> device", t0);
    }
    int deviceBuilt = 0;

    // ---- BLAS build (one per rt_blas, shared by instances)
    std::vector<BlasInfo>& blasInfo = sc->blasInfo;
    blasInfo.assign((size_t)d->nBlas, BlasInfo());
    // The scene-wide node / triangle arrays are assembled ON THE DEVICE from per-BLAS pieces: a piece the device builder
    // left in device memory (RT_BUILD_SAH / RT_BUILD_DEVICE_SAH) is rebased by a kernel and copied device to device —
    // it never visits the host; a host-built piece is rebased here and copied up into its slice.
    struct Piece {
        int64_t nodeBase = 0, triBase = 0, nNodes = 0, nTris = 0;
        bool f64 = false;
        std::vector<GNode> nodes; std::vector<GTri> tris; std::vector<GTriD> trisD;     // host-built
        void* dNodes = nullptr; void* dTris = nullptr;                                   // device-built, kept on the device
    };
    struct Pieces : std::vector<Piece> {
        ~Pieces() { for (Piece& p : *this) { cudaFree(p.dNodes); cudaFree(p.dTris); } }
    } pieces;
    int64_t nNodes = 0, nTris = 0, nTrisD = 0;
    std::vector<WNode> wnodes;
    int64_t maxDepth = 0, maxLeaf = 0;
    const char* residentEnv = std::getenv("RTCUDA_BUILD_RESIDENT");
    const bool resident = !(residentEnv && std::atoi(residentEnv) == 0);
    // wide BVHs for the beam walk of coherent packets (rt_beam.cuh): RTCUDA_WIDE=1 builds them (first B200 run: C4 12.4 ms
    // against 10.8 ms for the binary packet walk, profiles/r2q — off until it wins)
    const char* wideEnv = std::getenv("RTCUDA_WIDE");
    const bool wantWide = wideEnv && std::atoi(wideEnv) != 0;
    const int64_t wideMinTris = 64;
    for (int64_t b = 0; b < d->nBlas; b++) {
        const rt_blas& bl = d->blas[b];
        BlasInfo& bi = blasInfo[b];
        if (bl.primCount < 1 || bl.firstPrim < 0 || bl.firstPrim + bl.primCount > d->nPrims)
            return fail(RT_ERR_INVALID, "rt_blas range outside prims");
        const rt_prim* pr = d->prims + bl.firstPrim;
        bi.matId = pr[0].materialID;
        bi.smooth = pr[0].smooth;
        if (bl.primCount == 1 && pr[0].type == RT_GEO_SPHERE) {
            if (pr[0].primitiveIndex < 0 || pr[0].primitiveIndex >= d->nSpheres) return fail(RT_ERR_INVALID, "sphere index out of range");
            const rt_sphere& s = d->spheres[pr[0].primitiveIndex];
            bi.kind = INST_SPHERE; bi.prim = pr[0].primitiveIndex;
            bi.payload[0] = s.center[0]; bi.payload[1] = s.center[1]; bi.payload[2] = s.center[2]; bi.payload[3] = s.radius;
            for (int a = 0; a < 3; a++) { bi.lo[a] = s.center[a] - s.radius; bi.hi[a] = s.center[a] + s.radius; bi.centre[a] = s.center[a]; }
            continue;
        }
        if (bl.primCount == 1 && pr[0].type == RT_GEO_PLANE) {
            if (pr[0].primitiveIndex < 0 || pr[0].primitiveIndex >= d->nPlanes) return fail(RT_ERR_INVALID, "plane index out of range");
            bi.kind = INST_PLANE; bi.prim = pr[0].primitiveIndex;
            { const rt_plane& pl = d->planes[bi.prim]; for (int a = 0; a < 3; a++) { bi.payload[a] = pl.center[a]; bi.payload[3 + a] = pl.normal[a]; } }
            for (int a = 0; a < 3; a++) { bi.lo[a] = -1e5; bi.hi[a] = 1e5; bi.centre[a] = 0; }  // RTContext.swift:172-174
            continue;
        }
        std::vector<int64_t> idx((size_t)bl.primCount);
        for (int64_t k = 0; k < bl.primCount; k++) {
            if (pr[k].type != RT_GEO_TRIANGLE) return fail(RT_ERR_UNSUPPORTED, "a multi-primitive BLAS must hold triangles only");
            if (pr[k].primitiveIndex < 0 || pr[k].primitiveIndex >= d->nTriangles) return fail(RT_ERR_INVALID, "triangle index out of range");
            if (pr[k].materialID != bi.matId || pr[k].smooth != bi.smooth)
                return fail(RT_ERR_UNSUPPORTED, "per-triangle material / shading mode inside one BLAS");
            idx[k] = pr[k].primitiveIndex;
        }
        BlasBuildInput in{};
        in.triVertices = d->triVertices; in.triIndices = idx.data(); in.count = bl.primCount; in.threads = 0;
        in.keepOnDevice = resident && !(wantWide && bl.primCount >= wideMinTris);     // the wide collapse reads the nodes on the host
        bool blur = false;
        if (d->triMotionBlur) {            // one motion vector per mesh; the boxes grow by it (RTContext.swift:283-289)
            const double* m0 = d->triMotionBlur + idx[0] * 3;
            for (int64_t k = 0; k < bl.primCount; k++)
                if (std::memcmp(d->triMotionBlur + idx[k] * 3, m0, 24) != 0)
                    return fail(RT_ERR_UNSUPPORTED, "per-triangle motion vectors inside one BLAS");
            std::memcpy(in.motionBlur, m0, 24); std::memcpy(bi.motion, m0, 24);
            blur = m0[0] != 0 || m0[1] != 0 || m0[2] != 0;
            if (blur) sc->motion = true;
        }
        BlasBuildOutput out;
        bool built = false;
        const double tBlas = nowMs();
        if ((buildMode == RT_BUILD_DEVICE_SAH && bl.primCount >= 2) || (buildMode == RT_BUILD_SAH && bl.primCount >= RT_BUILD_AUTO_DEVICE_TRIANGLES)) {
            CK(buildBlasDeviceSah(dVerts.p, in, out, renderStackCapacity() - 2, nullptr));
            built = out.maxDepth <= renderStackCapacity() - 2;      // deeper than the stacks: buildBlas reports it below
            if (built) deviceBuilt++;
        } else if (buildMode == RT_BUILD_DEVICE_LBVH && bl.primCount >= 2 && !blur) {   // the radix-tree builder does not grow boxes by a motion vector
            CK(buildBlasDevice(dVerts.p, in, out));
            built = out.maxDepth <= renderStackCapacity() - 2;      // too deep for the stacks: the SAH builder takes over
            if (built) deviceBuilt++;
        }
        if (!built) {
            cudaFree(out.dNodes); cudaFree(out.dTris);          // a device attempt that was given up (tree too deep for the stacks)
            buildBlas(in, out);
        }
        if (bl.primCount >= 4096) trace(built ? "create: BLAS (device builder)" : "create: BLAS (host builder)", tBlas);
        if (out.maxDepth > renderStackCapacity() - 2) return fail(RT_ERR_UNSUPPORTED, "BLAS deeper than the traversal stack");
        bi.kind = bl.primCount == 1 ? INST_TRI : INST_MESH;
        bi.prim = pr[0].primitiveIndex;
        bi.f64 = out.f64 ? 1 : 0;
        bi.nodeBase = (int32_t)nNodes;
        bi.triBase = (int32_t)(out.f64 ? nTrisD : nTris);
        const bool onDevice = out.dTris != nullptr;
        const int64_t pieceNodes = onDevice ? out.nNodesDevice : (int64_t)out.nodes.size();
        bi.triCount = bl.primCount;
        bi.rootRef = out.rootRef;
        std::memcpy(bi.lo, out.rootLo, 24); std::memcpy(bi.hi, out.rootHi, 24); std::memcpy(bi.centre, out.centre, 24);
        // device references are absolute (gpu_layout.h): rebase the builder's BLAS-relative ones
        auto rebase = [&](int32_t ref) -> int32_t {
            if (ref == kEmptyRef) return ref;
            if (ref >= 0) return ref + bi.nodeBase;
            const uint32_t r = ~(uint32_t)ref;
            return ~(int32_t)((((r >> 3) + (uint32_t)bi.triBase) << 4) | (out.f64 ? 8u : 0u) | (r & 7u));
        };
        if ((int64_t)bi.triBase + bl.primCount >= (int64_t)1 << 27 || nNodes + pieceNodes >= kEmptyRef)
            return fail(RT_ERR_UNSUPPORTED, "scene exceeds the 2^27 triangles / 2^31 nodes of the device references");
        if (wantWide && bl.primCount >= wideMinTris && out.rootRef >= 0 && out.rootRef != kEmptyRef) {
            WideBuildOutput wide;
            collapseWide(out.nodes, out.rootRef, wide);
            // every visit pushes at most kWide entries and pops one: the beam stack must hold a root-to-leaf path of them
            if (wide.rootRef == 0 && (wide.depth - 1) * (kWide - 1) + kWide + 1 <= beamStackCapacity() &&
                wnodes.size() + wide.nodes.size() < (size_t)kEmptyRef / (size_t)kWide) {
                const int32_t wbase = (int32_t)wnodes.size();
                for (WNode& w : wide.nodes)
                    for (int k = 0; k < kWide; k++) {
                        int32_t& r = w.c[k].ref;
                        if (r == kEmptyRef) continue;
                        r = r >= 0 ? r + wbase : rebase(r);
                    }
                bi.wideRoot = wbase;
                wnodes.insert(wnodes.end(), wide.nodes.begin(), wide.nodes.end());
            }
        }
        for (GNode& g : out.nodes) { g.child0 = rebase(g.child0); g.child1 = rebase(g.child1); }
        bi.rootRef = rebase(out.rootRef);
        pieces.emplace_back();
        Piece& pc = pieces.back();
        pc.nodeBase = bi.nodeBase; pc.triBase = bi.triBase; pc.nNodes = pieceNodes; pc.nTris = bl.primCount; pc.f64 = out.f64;
        if (onDevice) { pc.dNodes = out.dNodes; pc.dTris = out.dTris; out.dNodes = out.dTris = nullptr; }
        else { pc.nodes = std::move(out.nodes); pc.tris = std::move(out.tris); pc.trisD = std::move(out.trisD); }
        nNodes += pieceNodes;
        if (out.f64) nTrisD += bl.primCount; else nTris += bl.primCount;
        if (out.maxDepth > maxDepth) maxDepth = out.maxDepth;
        if (out.maxLeaf > maxLeaf) maxLeaf = out.maxLeaf;
        if (bi.kind == INST_MESH) sc->meshes++;
    }

    // ---- instances
    std::vector<GInstance> inst((size_t)d->nInstances);
    std::vector<double> wb((size_t)d->nInstances * 6);
    bool instanceMotion = false;
    for (int64_t i = 0; i < d->nInstances; i++) {
        const rt_instance& in = d->instances[i];
        if (in.blas < 0 || in.blas >= d->nBlas) return fail(RT_ERR_INVALID, "rt_instance.blas out of range");
        const bool moves = in.instanceMotion[0] != 0 || in.instanceMotion[1] != 0 || in.instanceMotion[2] != 0;
        if (moves) { sc->motion = true; instanceMotion = true; }
        fillInstance(in, blasInfo[in.blas], inst[i], sc->bigBlasTriangles);
        sc->instBlas.push_back(in.blas);
        std::memcpy(&wb[i * 6], in.worldBounds, 48);
    }
    sc->smallGeometry = true;
    for (const BlasInfo& bi : blasInfo) if (bi.kind == INST_MESH && bi.triCount > kBigBlasTriangles) sc->smallGeometry = false;
    if (const char* e = std::getenv("RTCUDA_SMALL_GEOMETRY")) sc->smallGeometry = std::atoi(e) != 0;      // measurements / tests
    TlasBuildOutput tl;
    buildTlas(wb.data(), d->nInstances, tl);
    if (tl.maxDepth > 38) return fail(RT_ERR_UNSUPPORTED, "TLAS deeper than the TLAS stack");
    // instanceMotion: the reference's instance boxes do not follow the motion, so which displaced instances a ray
    // reaches depends on the reference's own TLAS topology — rebuilt exactly and walked in FP64 (rt_trace.cuh)
    RefTlasOutput refTl;
    if (instanceMotion) buildReferenceTlas(wb.data(), d->nInstances, refTl);

    // ---- materials / lights
    std::vector<GMaterial> mats((size_t)d->nMaterials);
    for (int64_t i = 0; i < d->nMaterials; i++) {
        const rt_material& m = d->materials[i];
        GMaterial& g = mats[i];
        std::memcpy(g.ambient, m.ambient, 24); std::memcpy(g.diffuse, m.diffuse, 24); std::memcpy(g.specular, m.specular, 24);
        std::memcpy(g.mirror, m.mirror, 24); std::memcpy(g.absorption, m.absorption, 24);
        g.phong = m.phong; g.ior = m.ior; g.absorptionIndex = m.absorptionIndex; g.roughness = m.roughness;
        g.type = (m.type >= RT_MAT_MIRROR && m.type <= RT_MAT_CONDUCTOR) ? m.type : RT_MAT_PLAIN;
        g.hasAbsorption = (m.absorption[0] != 0 || m.absorption[1] != 0 || m.absorption[2] != 0) ? 1 : 0;
        if (g.type != RT_MAT_PLAIN && m.roughness != 0) sc->glossy = true;
        if (g.type != RT_MAT_PLAIN && d->maxRecursionDepth > 0) sc->recurse = true;
        const double* coef[5] = {m.ambient, m.diffuse, m.specular, m.mirror, m.absorption};
        for (int k = 0; k < 5; k++)
            for (int a = 0; a < 3; a++) if (!std::isfinite(coef[k][a])) sc->nonFinite = true;
        if (!std::isfinite(m.phong) || !std::isfinite(m.ior) || !std::isfinite(m.absorptionIndex)) sc->nonFinite = true;
    }
    for (int a = 0; a < 3; a++) if (!std::isfinite(d->ambient[a]) || !std::isfinite(d->background[a])) sc->nonFinite = true;
    for (int64_t i = 0; i < d->nPointLights; i++)
        for (int a = 0; a < 3; a++) if (!std::isfinite(d->pointLights[i].intensity[a]) || !std::isfinite(d->pointLights[i].position[a])) sc->nonFinite = true;
    for (int64_t i = 0; i < d->nAreaLights; i++)
        for (int a = 0; a < 3; a++) if (!std::isfinite(d->areaLights[i].radiance[a])) sc->nonFinite = true;
    std::vector<GPointLight> lights((size_t)d->nPointLights);
    for (int64_t i = 0; i < d->nPointLights; i++) {
        std::memcpy(lights[i].position, d->pointLights[i].position, 24);
        std::memcpy(lights[i].intensity, d->pointLights[i].intensity, 24);
    }
    // area lights (:145-186) and the 10x10 stratified jitter table every frame rebuilds from a fixed seed (:92-93,646-659)
    std::vector<GAreaLight> areaLights((size_t)d->nAreaLights);
    for (int64_t i = 0; i < d->nAreaLights; i++) {
        std::memcpy(areaLights[i].position, d->areaLights[i].position, 24);
        std::memcpy(areaLights[i].normal, d->areaLights[i].normal, 24);
        areaLights[i].size = d->areaLights[i].size;
        std::memcpy(areaLights[i].radiance, d->areaLights[i].radiance, 24);
    }
    std::vector<double> jitter;
    if (d->nAreaLights > 0) {
        jitter.resize(200);
        // PCG32(seed:) of the reference (:556-589): state = 0, inc = (seed << 1) | 1, next, state += golden, next
        uint64_t state = 0, inc = (0x123456789ABCDEFULL << 1) | 1;
        auto next = [&]() -> uint32_t {
            const uint64_t old = state;
            state = old * 6364136223846793005ULL + inc;
            const uint32_t xs = (uint32_t)(((old >> 18) ^ old) >> 27), rot = (uint32_t)(old >> 59);
            return (xs >> rot) | (xs << ((~rot + 1u) & 31u));
        };
        next(); state += 0x9E3779B97F4A7C15ULL; next();
        int k = 0;
        for (int gy = 0; gy < 10; gy++)
            for (int gx = 0; gx < 10; gx++, k++) {
                jitter[k] = (double)gx + (double)next() * 2.3283064365386963e-10;
                jitter[100 + k] = (double)gy + (double)next() * 2.3283064365386963e-10;
            }
    }
    const double t1 = nowMs();
    trace("create: build phase total", t0);

    // ---- upload
    int64_t bytes = 0;
    int rc;
    if ((rc = deviceArray(&sc->dNodes, (size_t)nNodes, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(wnodes, &sc->dWNodes, bytes, &sc->arrays, sc))) return rc;
    if ((rc = deviceArray(&sc->dTris, (size_t)nTris, bytes, &sc->arrays, sc))) return rc;
    if ((rc = deviceArray(&sc->dTrisD, (size_t)nTrisD, bytes, &sc->arrays, sc))) return rc;
    for (Piece& pc : pieces) {
        const size_t triBytes = (size_t)pc.nTris * (pc.f64 ? sizeof(GTriD) : sizeof(GTri));
        void* triDst = pc.f64 ? (void*)(sc->dTrisD + pc.triBase) : (void*)(sc->dTris + pc.triBase);
        if (pc.dTris) {
            CK(launchRebaseNodes(sc->dNodes + pc.nodeBase, (const GNode*)pc.dNodes, (int)pc.nNodes, (int)pc.nodeBase, (int)pc.triBase, pc.f64 ? 1 : 0, 0));
            if (triBytes) CK(cudaMemcpyAsync(triDst, pc.dTris, triBytes, cudaMemcpyDeviceToDevice, 0));
        } else {
            if (pc.nNodes) CK(cudaMemcpy(sc->dNodes + pc.nodeBase, pc.nodes.data(), (size_t)pc.nNodes * sizeof(GNode), cudaMemcpyHostToDevice));
            if (triBytes) CK(cudaMemcpy(triDst, pc.f64 ? (const void*)pc.trisD.data() : (const void*)pc.tris.data(), triBytes, cudaMemcpyHostToDevice));
        }
    }
    CK(cudaDeviceSynchronize());          // the pieces are freed when `pieces` goes out of scope
    if ((rc = uploadRaw(d->triNormals, (size_t)d->nTriangles * 9, &sc->dNormals, bytes, &sc->arrays, sc))) return rc;     // straight from the caller's array
    // TLAS records: the BLAS node format — FP32 boxes relative to the centre of the scene bounds, padded by
    // M * 2^-19 and rounded outward (conservative: the instance's own exact test decides)
    std::vector<GNode> tlasF(tl.nodes.size());
    double sLo[3] = {0, 0, 0}, sHi[3] = {0, 0, 0}, sC[3] = {0, 0, 0};
    if (!tl.nodes.empty()) {
        const double P = tlasFrame(tl.nodes[0], sLo, sHi, sC);
        auto down = [](double x) { float f = (float)x; return (double)f > x ? std::nextafterf(f, -INFINITY) : f; };
        auto up = [](double x) { float f = (float)x; return (double)f < x ? std::nextafterf(f, INFINITY) : f; };
        for (size_t i = 0; i < tl.nodes.size(); i++) {
            const TNode& t = tl.nodes[i];
            GNode& g = tlasF[i];
            float lo0[3], hi0[3], lo1[3], hi1[3];
            for (int a = 0; a < 3; a++) {
                lo0[a] = down(t.lo0[a] - sC[a] - P); hi0[a] = up(t.hi0[a] - sC[a] + P);
                lo1[a] = down(t.lo1[a] - sC[a] - P); hi1[a] = up(t.hi1[a] - sC[a] + P);
            }
            g.q0[0] = lo0[0]; g.q0[1] = lo0[1]; g.q0[2] = lo0[2]; g.q0[3] = hi0[0];
            g.q1[0] = hi0[1]; g.q1[1] = hi0[2]; g.q1[2] = lo1[0]; g.q1[3] = lo1[1];
            g.q2[0] = lo1[2]; g.q2[1] = hi1[0]; g.q2[2] = hi1[1]; g.q2[3] = hi1[2];
            g.child0 = t.child0; g.child1 = t.child1; g.pad0 = g.pad1 = 0;
        }
    }
    // animation support: the FP64 TLAS, its nodes grouped by depth (deepest first) for the bottom-up refit kernel,
    // the instances' world bounds, pinned staging for the per-frame instance upload
    {
        std::vector<int> depth(tl.nodes.size(), 0), order(tl.nodes.size());
        int maxD = 0;
        for (size_t i = 0; i < tl.nodes.size(); i++) {       // pre-order: a child's index is above its parent's
            const int32_t ch[2] = {tl.nodes[i].child0, tl.nodes[i].child1};
            for (int c = 0; c < 2; c++)
                if (ch[c] >= 0 && ch[c] != kEmptyRef) { depth[ch[c]] = depth[i] + 1; if (depth[ch[c]] > maxD) maxD = depth[ch[c]]; }
        }
        sc->refitLevelStart.assign(1, 0);
        size_t k = 0;
        for (int lv = maxD; lv >= 0 && !tl.nodes.empty(); lv--) {
            for (size_t i = 0; i < tl.nodes.size(); i++) if (depth[i] == lv) order[k++] = (int)i;
            sc->refitLevelStart.push_back((int)k);
        }
        if ((rc = upload(tl.nodes, &sc->dTlasD, bytes, &sc->arrays, sc))) return rc;
        if ((rc = upload(order, &sc->dRefitOrder, bytes, &sc->arrays, sc))) return rc;
        if ((rc = upload(sc->refitLevelStart, &sc->dRefitLevelStart, bytes, &sc->arrays, sc))) return rc;
        if ((rc = upload(wb, &sc->dInstBounds, bytes, &sc->arrays, sc))) return rc;
        std::vector<BoundsChunk> chunks;                      // 4096 triangles per block of the bounds kernel
        for (size_t i = 0; i < (size_t)d->nInstances; i++) {
            const BlasInfo& bi = blasInfo[sc->instBlas[i]];
            if (bi.kind != INST_MESH) continue;
            for (int64_t f = 0; f < bi.triCount; f += 4096)
                chunks.push_back(BoundsChunk{(int32_t)i, (int32_t)(bi.triBase + f), (int32_t)std::min<int64_t>(4096, bi.triCount - f), bi.f64});
        }
        sc->nBoundChunks = (int)chunks.size();
        if ((rc = upload(chunks, &sc->dBoundChunks, bytes, &sc->arrays, sc))) return rc;
        sc->instanceMotion = instanceMotion;
    }
    if ((rc = upload(tlasF, &sc->dTlas, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(inst, &sc->dInst, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(mats, &sc->dMat, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(lights, &sc->dLights, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(areaLights, &sc->dAreaLights, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(jitter, &sc->dJitter, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(refTl.nodes, &sc->dRefTlas, bytes, &sc->arrays, sc))) return rc;
    if ((rc = upload(refTl.primIdx, &sc->dRefTlasPrim, bytes, &sc->arrays, sc))) return rc;
    if ((rc = allocFrameState(sc, (size_t)d->nInstances))) return rc;
    CK(cudaDeviceSynchronize());
    const double t2 = nowMs();
    trace("create: upload", t1);

    DeviceScene& S = sc->ds;
    bindDeviceScene(sc);
    S.tlasRootRef = tl.rootRef;
    std::memcpy(S.tlasLo, sLo, 24); std::memcpy(S.tlasHi, sHi, 24); std::memcpy(S.tlasCentre, sC, 24);
    S.nInstances = (int32_t)d->nInstances; S.nMaterials = (int32_t)d->nMaterials; S.nPointLights = (int32_t)d->nPointLights;
    S.nAreaLights = (int32_t)d->nAreaLights;
    S.refTlasNodes = (int32_t)refTl.nodes.size();
    S.wideMode = wantWide ? std::atoi(wideEnv) : 0;
    S.maxDepth = d->maxRecursionDepth;
    std::memcpy(S.ambient, d->ambient, 24); std::memcpy(S.background, d->background, 24);
    // a scene without instances has no TLAS, and `trace` then returns .zero, not the background
    // (`guard let tlas = ctx.tlas else { return .zero }`, Object+Extension.swift:98; RTContext.swift:416)
    if (d->nInstances == 0) S.background[0] = S.background[1] = S.background[2] = 0.0;
    S.eps = d->intersectionTestEpsilon; S.shadowEps = d->shadowRayEpsilon;

    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    for (int v = 0; v < KV_COUNT; v++)
        for (int c = 0; c < 2; c++) {
            int bps = 0;
            CK(renderOccupancy(v, &bps, c != 0));
            sc->blocks[v][c] = prop.multiProcessorCount * (bps > 0 ? bps : 1);
        }

    rt_scene_info& I = sc->info;
    I.triangles = d->nTriangles; I.spheres = d->nSpheres; I.planes = d->nPlanes; I.instances = d->nInstances; I.blasCount = d->nBlas;
    I.blasNodes = nNodes; I.tlasNodes = (int64_t)tl.nodes.size(); I.maxLeafSize = maxLeaf; I.maxDepth = maxDepth;
    I.deviceBytes = bytes; I.buildMs = t1 - t0; I.uploadMs = t2 - t1; I.device = device; I.deviceBuiltBlas = deviceBuilt; I.deviceCount = 1;
    I.wideNodes = (int32_t)wnodes.size();
    return RT_OK;
}

void rt_scene_destroy(rt_scene_t* sc) {
    if (!sc) return;
    if (!sc->isReplica) {
        delete sc->pool;                                  // joins the worker threads
        sc->pool = nullptr;
        for (size_t r = 1; r < sc->replicas.size(); r++) rt_scene_destroy(sc->replicas[r]);
        sc->replicas.clear();
    }
    cudaSetDevice(sc->device);
    if (sc->stream) cudaStreamSynchronize(sc->stream);
    if (sc->side) cudaStreamSynchronize(sc->side);
    cudaFree(sc->dWNodes);
    cudaFree(sc->dNodes); cudaFree(sc->dTris); cudaFree(sc->dTrisD); cudaFree(sc->dNormals); cudaFree(sc->dTlas);
    cudaFree(sc->dInst); cudaFree(sc->dMat); cudaFree(sc->dLights); cudaFree(sc->dCounters);
    cudaFree(sc->dAreaLights); cudaFree(sc->dJitter); cudaFree(sc->dAreaIndex); cudaFree(sc->dDefer);
    cudaFree(sc->dRefTlas); cudaFree(sc->dRefTlasPrim);
    cudaFree(sc->dTlasD); cudaFree(sc->dRefitOrder); cudaFree(sc->dRefitLevelStart); cudaFree(sc->dInstBounds);
    if (sc->hInst) cudaFreeHost(sc->hInst);
    if (sc->hInstBounds) cudaFreeHost(sc->hInstBounds);
    if (sc->hDerive) cudaFreeHost(sc->hDerive);
    if (sc->hBoundsEnc) cudaFreeHost(sc->hBoundsEnc);
    cudaFree(sc->dDerive); cudaFree(sc->dBoundsEnc); cudaFree(sc->dBoundChunks);
    cudaFree(sc->dTileCost); cudaFree(sc->dTileOrder); cudaFree(sc->dSchedWork);
    cudaFree(sc->dPathStack); cudaFree(sc->dRgba); cudaFree(sc->dLinear); cudaFree(sc->dIds);
    if (sc->hCounters) cudaFreeHost(sc->hCounters);
    if (sc->hPoll) cudaFreeHost(sc->hPoll);
    if (sc->evStart) cudaEventDestroy(sc->evStart);
    if (sc->evStop) cudaEventDestroy(sc->evStop);
    if (sc->evSide) cudaEventDestroy(sc->evSide);
    if (sc->evReset) cudaEventDestroy(sc->evReset);
    if (sc->side) cudaStreamDestroy(sc->side);
    if (sc->stream) cudaStreamDestroy(sc->stream);
    if (sc->poll) cudaStreamDestroy(sc->poll);
    delete sc;
}

rt_scene_t* rt_scene_create(const rt_scene_desc* desc, int32_t device) {
    rt_scene* sc = nullptr;
    try {
        sc = new rt_scene();
        int rc = createImpl(desc, device, sc);
        if (rc != RT_OK) { rt_scene_destroy(sc); return nullptr; }
        sc->replicas.assign(1, sc);
        return sc;
    } catch (const std::exception& e) {
        g_err = e.what();
    } catch (...) {
        g_err = "unknown exception";
    }
    if (sc) rt_scene_destroy(sc);
    return nullptr;
}

// A copy of `src` on another GPU of this process: the scene's device arrays travel device to device (NVLink peer copy),
// nothing is rebuilt or re-uploaded from the host; frame state (counters, streams, events, staging) is the replica's own.
static int cloneScene(const rt_scene* src, int device, rt_scene* dst) {
    dst->device = device;
    dst->isReplica = true;
    dst->ds = src->ds; dst->info = src->info; dst->info.device = device; dst->meshes = src->meshes;
    dst->motion = src->motion; dst->glossy = src->glossy; dst->nonFinite = src->nonFinite; dst->recurse = src->recurse; dst->smallGeometry = src->smallGeometry; dst->bigBlasTriangles = src->bigBlasTriangles; dst->instanceMotion = src->instanceMotion;
    dst->blasInfo = src->blasInfo; dst->instBlas = src->instBlas; dst->refitLevelStart = src->refitLevelStart;
    dst->nBoundChunks = src->nBoundChunks;
    dst->arrays = src->arrays;
    CK(cudaSetDevice(device));
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, device, src->device));
    if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(src->device, 0); if (e != cudaSuccess) cudaGetLastError(); }
    for (const SceneArray& a : src->arrays) {
        void* const from = *reinterpret_cast<void* const*>(reinterpret_cast<const char*>(src) + a.slotOffset);
        void** to = reinterpret_cast<void**>(reinterpret_cast<char*>(dst) + a.slotOffset);
        CK(cudaMalloc(to, a.bytes));
        CK(cudaMemcpyPeerAsync(*to, device, from, src->device, a.bytes, nullptr));
    }
    int rc = allocFrameState(dst, dst->instBlas.size());
    if (rc) return rc;
    bindDeviceScene(dst);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    for (int v = 0; v < KV_COUNT; v++)
        for (int c = 0; c < 2; c++) {
            int bps = 0;
            CK(renderOccupancy(v, &bps, c != 0));
            dst->blocks[v][c] = prop.multiProcessorCount * (bps > 0 ? bps : 1);
        }
    CK(cudaDeviceSynchronize());
    return RT_OK;
}

rt_scene_t* rt_scene_create_multi(const rt_scene_desc* desc, const int32_t* devices, int32_t nDevices) {
    if (!devices || nDevices < 1 || nDevices > 64) { fail(RT_ERR_INVALID, "rt_scene_create_multi: 1..64 devices are required"); return nullptr; }
    for (int a = 0; a < nDevices; a++)
        for (int b = a + 1; b < nDevices; b++)
            if (devices[a] == devices[b]) { fail(RT_ERR_INVALID, "rt_scene_create_multi: a device is listed twice"); return nullptr; }
    rt_scene* sc = rt_scene_create(desc, devices[0]);
    if (!sc || nDevices == 1) return sc;
    try {
        const double t0 = nowMs();
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess) ndev = 0;
        for (int k = 1; k < nDevices; k++) {
            if (devices[k] < 0 || devices[k] >= ndev) { fail(RT_ERR_INVALID, "invalid CUDA device index"); rt_scene_destroy(sc); return nullptr; }
            rt_scene* rep = new rt_scene();
            sc->replicas.push_back(rep);                  // owned from here on: rt_scene_destroy(sc) frees it
            if (cloneScene(sc, devices[k], rep) != RT_OK) { rt_scene_destroy(sc); return nullptr; }
        }
        sc->info.replicateMs = nowMs() - t0;
        sc->info.deviceCount = nDevices;
        sc->pool = new ReplicaPool();
        sc->pool->start(nDevices);
        cudaSetDevice(sc->device);
        return sc;
    } catch (const std::exception& e) {
        g_err = e.what();
    } catch (...) {
        g_err = "unknown exception";
    }
    rt_scene_destroy(sc);
    return nullptr;
}

int32_t rt_scene_device_count(const rt_scene_t* sc) { return sc ? (int32_t)sc->replicas.size() : 0; }

void* rt_host_alloc(uint64_t bytes) {
    void* p = nullptr;
    if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); g_err = "rt_host_alloc: cudaHostAlloc failed"; return nullptr; }
    return p;
}
void rt_host_free(void* p) { if (p) cudaFreeHost(p); }

static int updateInstancesOne(rt_scene_t* sc, rt_instance* instances, int64_t nInstances, void* cudaStream);

int32_t rt_scene_update_instances(rt_scene_t* sc, rt_instance* instances, int64_t nInstances, void* cudaStream) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    int rc = updateInstancesOne(sc, instances, nInstances, cudaStream);
    // replicas: the same records and the same refit on their own GPU (their own stream: `cudaStream` belongs to device 0)
    for (size_t r = 1; r < sc->replicas.size() && rc == RT_OK; r++) rc = updateInstancesOne(sc->replicas[r], instances, nInstances, nullptr);
    if (sc->replicas.size() > 1) cudaSetDevice(sc->device);
    return rc;
}

}  // extern "C"

void ReplicaPool::start(int n) {
    rc.assign((size_t)n, 0); err.assign((size_t)n, std::string());
    for (int i = 0; i < n; i++)
        threads.emplace_back([this, i]() {
            int seen = 0;
            for (;;) {
                std::function<int(int)> j;
                {
                    std::unique_lock<std::mutex> lk(m);
                    cvStart.wait(lk, [&] { return quit || generation != seen; });
                    if (quit) return;
                    seen = generation;
                    j = job;
                }
                int r;
                try { r = j(i); } catch (...) { r = RT_ERR_CUDA; g_err = "exception in a replica worker"; }
                {
                    std::lock_guard<std::mutex> lk(m);
                    rc[(size_t)i] = r;
                    if (r != RT_OK) err[(size_t)i] = g_err;      // g_err is thread-local: hand the message to the caller's thread
                    if (--remaining == 0) cvDone.notify_all();
                }
            }
        });
}
void ReplicaPool::launch(std::function<int(int)> j) {
    std::lock_guard<std::mutex> lk(m);
    job = std::move(j);
    remaining = (int)threads.size();
    generation++;
    cvStart.notify_all();
}
void ReplicaPool::wait() {
    std::unique_lock<std::mutex> lk(m);
    cvDone.wait(lk, [&] { return remaining == 0; });
}
ReplicaPool::~ReplicaPool() {
    { std::lock_guard<std::mutex> lk(m); quit = true; }
    cvStart.notify_all();
    for (std::thread& t : threads) if (t.joinable()) t.join();
}

extern "C" {

static int updateInstancesOne(rt_scene_t* sc, rt_instance* instances, int64_t nInstances, void* cudaStream) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    if (sc->pending) return fail(RT_ERR_INVALID, "a render is in flight on this scene (call rt_render_finish)");
    if (!instances || nInstances != (int64_t)sc->instBlas.size())
        return fail(RT_ERR_INVALID, "rt_scene_update_instances: the instance count of a scene is fixed");
    if (sc->instanceMotion)
        return fail(RT_ERR_UNSUPPORTED, "scenes with instanceMotion walk the reference's own TLAS topology: rebuild the scene instead");
    CK(cudaSetDevice(sc->device));
    cudaStream_t st = (cudaStream_t)cudaStream;
    const size_t n = (size_t)nInstances;
    if (sc->sideBusy) CK(cudaEventSynchronize(sc->evSide));      // the previous frame's epilogue is done
    bool anyDerived = false;
    for (size_t i = 0; i < n; i++) {
        const rt_instance& in = instances[i];
        if (in.blas != sc->instBlas[i]) return fail(RT_ERR_INVALID, "rt_scene_update_instances: an instance keeps its BLAS");
        if (in.instanceMotion[0] != 0 || in.instanceMotion[1] != 0 || in.instanceMotion[2] != 0)
            return fail(RT_ERR_UNSUPPORTED, "rt_scene_update_instances: instanceMotion cannot be introduced by an update");
        const BlasInfo& bi = sc->blasInfo[in.blas];
        fillInstance(in, bi, sc->hInst[i], sc->bigBlasTriangles);
        // an inverted box (minP > maxP) asks for Instance.worldBounds to be derived from the BLAS's triangles
        const bool derive = !(in.worldBounds[0] <= in.worldBounds[3]) || !(in.worldBounds[1] <= in.worldBounds[4]) ||
                            !(in.worldBounds[2] <= in.worldBounds[5]);
        if (derive && bi.kind != INST_MESH)
            return fail(RT_ERR_INVALID, "rt_scene_update_instances: only mesh instances can have their bounds derived");
        sc->hDerive[i] = derive ? 1 : 0;
        anyDerived = anyDerived || derive;
        for (int a = 0; a < 3; a++) {
            sc->hBoundsEnc[i * 6 + a] = encodeOrdered(derive ? (double)INFINITY : in.worldBounds[a]);
            sc->hBoundsEnc[i * 6 + 3 + a] = encodeOrdered(derive ? -(double)INFINITY : in.worldBounds[3 + a]);
        }
    }
    CK(cudaMemcpyAsync(sc->dInst, sc->hInst, n * sizeof(GInstance), cudaMemcpyHostToDevice, st));
    if (anyDerived) {
        CK(cudaMemcpyAsync(sc->dDerive, sc->hDerive, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(sc->dBoundsEnc, sc->hBoundsEnc, n * 6 * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
        CK(launchInstanceBounds(sc->ds, sc->dBoundChunks, sc->nBoundChunks, sc->dDerive, sc->dBoundsEnc, st));
        CK(cudaMemcpyAsync(sc->hBoundsEnc, sc->dBoundsEnc, n * 6 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    for (size_t i = 0; i < n; i++)
        for (int a = 0; a < 6; a++) {
            sc->hInstBounds[i * 6 + a] = decodeOrdered(sc->hBoundsEnc[i * 6 + a]);
            if (sc->hDerive[i]) instances[i].worldBounds[a] = sc->hInstBounds[i * 6 + a];     // handed back to the caller
        }
    if (sc->info.tlasNodes > 0) {
        // scene bounds (the union of the instance bounds), their centre and the box padding: host, O(instances)
        TNode root{};
        for (int a = 0; a < 3; a++) { root.lo0[a] = root.lo1[a] = INFINITY; root.hi0[a] = root.hi1[a] = -INFINITY; }
        for (size_t i = 0; i < n; i++)
            for (int a = 0; a < 3; a++) {
                root.lo0[a] = std::fmin(root.lo0[a], sc->hInstBounds[i * 6 + a]);
                root.hi0[a] = std::fmax(root.hi0[a], sc->hInstBounds[i * 6 + 3 + a]);
            }
        std::memcpy(root.lo1, root.lo0, 24); std::memcpy(root.hi1, root.hi0, 24);
        double sLo[3], sHi[3], sC[3];
        const double P = tlasFrame(root, sLo, sHi, sC);
        CK(cudaMemcpyAsync(sc->dInstBounds, sc->hInstBounds, n * 6 * sizeof(double), cudaMemcpyHostToDevice, st));
        CK(launchTlasRefit(sc->dTlasD, sc->dTlas, sc->dInstBounds, sc->dRefitOrder, sc->dRefitLevelStart,
                           (int)sc->refitLevelStart.size() - 1, sC, P, st));
        std::memcpy(sc->ds.tlasLo, sLo, 24); std::memcpy(sc->ds.tlasHi, sHi, 24); std::memcpy(sc->ds.tlasCentre, sC, 24);
    }
    CK(cudaStreamSynchronize(st));       // the staging buffers are free again when this returns
    sc->updateLaunches = (anyDerived ? 1 : 0) + (sc->info.tlasNodes > 0 ? 1 : 0);
    return RT_OK;
}

int32_t rt_ply_ingest(const rt_ply_layout* L, const void* vertexBlock, const void* faceBlock, int32_t smooth, int32_t device,
                      double* triVerticesOut, double* triNormalsOut, rt_ply_stats* statsOut) {
    if (!L || !triVerticesOut || !triNormalsOut) return fail(RT_ERR_INVALID, "rt_ply_ingest: null argument");
    if (L->vertexCount < 0 || L->faceCount < 0 || L->vertexStride < 12 || (L->verticesPerFace != 3 && L->verticesPerFace != 4) ||
        L->faceCountBytes < 1 || L->faceCountBytes > 4 || (L->faceIndexBytes != 1 && L->faceIndexBytes != 2 && L->faceIndexBytes != 4) ||
        L->faceLeadBytes < 0 || L->faceTrailBytes < 0)
        return fail(RT_ERR_INVALID, "rt_ply_ingest: inconsistent layout");
    for (int a = 0; a < 3; a++)
        if (L->posOffset[a] < 0 || L->posOffset[a] + 4 > L->vertexStride || L->normalOffset[a] + 4 > L->vertexStride)
            return fail(RT_ERR_INVALID, "rt_ply_ingest: property offset outside the vertex record");
    if ((L->vertexCount > 0 && !vertexBlock) || (L->faceCount > 0 && !faceBlock)) return fail(RT_ERR_INVALID, "rt_ply_ingest: null block");
    if (L->vertexCount >= ((int64_t)1 << 31) || L->faceCount >= ((int64_t)1 << 30)) return fail(RT_ERR_UNSUPPORTED, "rt_ply_ingest: mesh too large for 32-bit indices");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(RT_ERR_INVALID, "invalid CUDA device index");
    CK(cudaSetDevice(device));
    const char* msg = "";
    int64_t launches = 0;
    double ms = 0;
    if (plyIngestDevice(*L, vertexBlock, faceBlock, smooth, triVerticesOut, triNormalsOut, &msg, &launches, &ms) != 0) {
        cudaGetLastError();
        return fail(RT_ERR_INVALID, std::string("rt_ply_ingest: ") + msg);
    }
    if (statsOut) { statsOut->triangles = L->faceCount * (L->verticesPerFace == 4 ? 2 : 1); statsOut->kernelLaunches = launches; statsOut->deviceMs = ms; }
    return RT_OK;
}

// rt_ply_ingest with every intermediate in a fixed device array inside the library: allocates and frees nothing
// (ply_ingest.cu: plyIngestStatic).  No counterpart in the reference.
int32_t rt_ply_ingest_static(const rt_ply_layout* L, const void* dVertexBlock, const void* dFaceBlock, int32_t smooth, int32_t device,
                             double* dTriVerticesOut, double* dTriNormalsOut, void* cudaStream, rt_ply_stats* statsOut) {
    if (!L || !dTriVerticesOut || !dTriNormalsOut) return fail(RT_ERR_INVALID, "rt_ply_ingest_static: null argument");
    if (L->vertexCount < 0 || L->faceCount < 0 || L->vertexStride < 12 || (L->verticesPerFace != 3 && L->verticesPerFace != 4) ||
        L->faceCountBytes < 1 || L->faceCountBytes > 4 || (L->faceIndexBytes != 1 && L->faceIndexBytes != 2 && L->faceIndexBytes != 4) ||
        L->faceLeadBytes < 0 || L->faceTrailBytes < 0)
        return fail(RT_ERR_INVALID, "rt_ply_ingest_static: inconsistent layout");
    for (int a = 0; a < 3; a++)
        if (L->posOffset[a] < 0 || L->posOffset[a] + 4 > L->vertexStride || L->normalOffset[a] + 4 > L->vertexStride)
            return fail(RT_ERR_INVALID, "rt_ply_ingest_static: property offset outside the vertex record");
    if ((L->vertexCount > 0 && !dVertexBlock) || (L->faceCount > 0 && !dFaceBlock)) return fail(RT_ERR_INVALID, "rt_ply_ingest_static: null block");
    const int64_t nTris = L->faceCount * (L->verticesPerFace == 4 ? 2 : 1);
    if (L->vertexCount > RT_PLY_STATIC_MAX_VERTICES || nTris > RT_PLY_STATIC_MAX_TRIANGLES)
        return fail(RT_ERR_UNSUPPORTED, "rt_ply_ingest_static: the mesh has " + std::to_string(L->vertexCount) + " vertices and " +
                    std::to_string(nTris) + " triangles, the static scratch holds " + std::to_string(RT_PLY_STATIC_MAX_VERTICES) + " and " +
                    std::to_string(RT_PLY_STATIC_MAX_TRIANGLES) + ": use rt_ply_ingest");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(RT_ERR_INVALID, "invalid CUDA device index");
    CK(cudaSetDevice(device));
    const char* msg = "";
    int64_t launches = 0;
    const int rc = plyIngestStatic(*L, dVertexBlock, dFaceBlock, smooth, device, dTriVerticesOut, dTriNormalsOut, (cudaStream_t)cudaStream,
                                   &msg, &launches);
    if (rc != 0) {
        cudaGetLastError();
        return fail(rc == -2 ? RT_ERR_UNSUPPORTED : RT_ERR_INVALID, std::string("rt_ply_ingest_static: ") + msg);
    }
    if (statsOut) { statsOut->triangles = nTris; statsOut->kernelLaunches = launches; statsOut->deviceMs = 0; }
    return RT_OK;
}

int32_t rt_scene_get_info(const rt_scene_t* sc, rt_scene_info* out) {
    if (!sc || !out) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    *out = sc->info;
    return RT_OK;
}

int64_t rt_scene_get_schedule(rt_scene_t* sc, int32_t* orderOut, int64_t capacity) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    if (sc->pending) return fail(RT_ERR_INVALID, "a render is in flight on this scene (call rt_render_finish)");
    if (!sc->schedValid) return 0;
    const int64_t n = sc->schedKey[7];
    if (orderOut && capacity > 0) {
        if (cudaSetDevice(sc->device) != cudaSuccess ||
            cudaMemcpy(orderOut, sc->dTileOrder, (size_t)(n < capacity ? n : capacity) * sizeof(int32_t), cudaMemcpyDeviceToHost) != cudaSuccess)
            return fail(RT_ERR_CUDA, "rt_scene_get_schedule: copy failed");
    }
    return n;
}

int64_t rt_scene_read_wide_nodes(rt_scene_t* sc, void* nodesOut, int64_t capacityBytes) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    const int64_t bytes = (int64_t)sc->info.wideNodes * (int64_t)sizeof(WNode);
    if (nodesOut && capacityBytes > 0 && bytes > 0) {
        if (cudaSetDevice(sc->device) != cudaSuccess ||
            cudaMemcpy(nodesOut, sc->dWNodes, (size_t)(bytes < capacityBytes ? bytes : capacityBytes), cudaMemcpyDeviceToHost) != cudaSuccess)
            return fail(RT_ERR_CUDA, "rt_scene_read_wide_nodes: copy failed");
    }
    return bytes;
}

int64_t rt_scene_read_blas_nodes(rt_scene_t* sc, void* nodesOut, int64_t capacityBytes) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    const int64_t bytes = sc->info.blasNodes * (int64_t)sizeof(GNode);
    if (nodesOut && capacityBytes > 0 && bytes > 0) {
        if (cudaSetDevice(sc->device) != cudaSuccess ||
            cudaMemcpy(nodesOut, sc->dNodes, (size_t)(bytes < capacityBytes ? bytes : capacityBytes), cudaMemcpyDeviceToHost) != cudaSuccess)
            return fail(RT_ERR_CUDA, "rt_scene_read_blas_nodes: copy failed");
    }
    return bytes;
}

}  // extern "C"

#include "rt_api_render.inl"
