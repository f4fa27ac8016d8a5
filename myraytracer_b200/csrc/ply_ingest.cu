// ply_ingest.cu — PLY -> device ingest.   COMPILE WITH -fmad=false.
//
// The step right before the hot path.  The reference goes PLY -> [Float] -> [Vec3] (FP64) -> Triangle objects on one
// thread: miniply through the CPly C ABI (Helpers/PLYReader.swift:54-210), then the mesh branch of RTContext.init
// (Models/RTContext.swift:251-363): face normals, smooth vertex normals as the sum of the UNIT face normals of the
// faces around a vertex (:305-326), one Triangle per face with v0,v1,v2,n0,n1,n2.  Here the raw binary-little-endian
// vertex and face blocks of the file go to the GPU as they are and kernels produce the same per-triangle FP64 arrays
// (rt_scene_desc.triVertices / triNormals), bit for bit:
//   * positions are read as Float and widened (PLYReader.swift:85-101); file normals are normalised (:104-122)
//   * quads split 0,1,3 / 2,3,1 like miniply (miniply.cpp:1976-1985)
//   * a vertex normal is summed in (face, corner) order — the order of the reference's loop — so the FP64 sum is the
//     same: the faces around each vertex are gathered into a CSR list (count, exclusive scan, fill), each short list is
//     sorted, then added up sequentially by one thread
//   * flat shading normalises the unit face normal once more (:343)
// No warp-level primitives: atomics, a three-kernel scan and plain per-element kernels — HBM-bound byte and FP64 work.
#include <math_constants.h>

#include <mutex>

#include "../../include/rtcuda.h"
#include "rt_kernels.h"
#include "rt_math.cuh"

namespace rtg {

#define RT_PLY_THREADS 256

struct PlyDev {
    rt_ply_layout L;
    const unsigned char* vblock;
    const unsigned char* fblock;
    int32_t faceRecord;          // bytes of one face record
    int32_t trisPerFace;         // 1 or 2
    double* pos;                 // vertexCount * 3
    double* vnormFile;           // vertexCount * 3 (file normals, normalised) or null
    int32_t* idx;                // nTris * 3
    double* faceN;               // nTris * 3 (unit)
    unsigned* count;             // vertexCount + 1: faces-around-vertex counts, then their exclusive prefix
    unsigned* cursor;            // vertexCount
    unsigned* entries;           // nTris * 3: corner ids (t * 3 + c) grouped by vertex
    double* vnorm;               // vertexCount * 3: smooth vertex normals
    int* error;                  // 1: a face record with another vertex count, 2: index out of range
};

__device__ __forceinline__ float loadF32(const unsigned char* p) {
    unsigned u = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
    return __uint_as_float(u);
}
__device__ __forceinline__ long long loadInt(const unsigned char* p, int bytes) {
    unsigned u = 0;
    for (int k = 0; k < bytes; k++) u |= (unsigned)p[k] << (8 * k);
    if (bytes == 4) return (long long)(int)u;             // Int32 indices (PLYReader.swift:186-198)
    return (long long)u;
}

__global__ void __launch_bounds__(RT_PLY_THREADS) plyVertices(PlyDev P) {
    const long long v = (long long)blockIdx.x * RT_PLY_THREADS + threadIdx.x;
    if (v >= P.L.vertexCount) return;
    const unsigned char* r = P.vblock + v * P.L.vertexStride;
    for (int a = 0; a < 3; a++) P.pos[v * 3 + a] = (double)loadF32(r + P.L.posOffset[a]);
    P.count[v] = 0;
    if (v == 0) P.count[P.L.vertexCount] = 0;
    P.cursor[v] = 0;
    if (P.vnormFile) {
        const double a = (double)loadF32(r + P.L.normalOffset[0]), b = (double)loadF32(r + P.L.normalOffset[1]),
                     c = (double)loadF32(r + P.L.normalOffset[2]);
        const double inv = 1.0 / sqrt((a * a + b * b) + c * c);
        P.vnormFile[v * 3] = a * inv; P.vnormFile[v * 3 + 1] = b * inv; P.vnormFile[v * 3 + 2] = c * inv;
    }
}

__global__ void __launch_bounds__(RT_PLY_THREADS) plyFaces(PlyDev P) {
    const long long f = (long long)blockIdx.x * RT_PLY_THREADS + threadIdx.x;
    if (f >= P.L.faceCount) return;
    const unsigned char* r = P.fblock + f * P.faceRecord + P.L.faceLeadBytes;
    const int k = P.L.verticesPerFace;
    if (loadInt(r, P.L.faceCountBytes) != k) { *P.error = 1; return; }
    long long q[4] = {0, 0, 0, 0};
    bool ok = true;
    for (int c = 0; c < k; c++) {
        q[c] = loadInt(r + P.L.faceCountBytes + c * P.L.faceIndexBytes, P.L.faceIndexBytes);
        ok = ok && q[c] >= 0 && q[c] < P.L.vertexCount;
    }
    if (!ok) { *P.error = 2; return; }
    int32_t* o = P.idx + f * P.trisPerFace * 3;
    if (k == 3) { o[0] = (int32_t)q[0]; o[1] = (int32_t)q[1]; o[2] = (int32_t)q[2]; }
    else { o[0] = (int32_t)q[0]; o[1] = (int32_t)q[1]; o[2] = (int32_t)q[3]; o[3] = (int32_t)q[2]; o[4] = (int32_t)q[3]; o[5] = (int32_t)q[1]; }
}

// unit face normal (RTContext.swift:307-314) and, for smooth meshes, the faces-around-vertex counts
__global__ void __launch_bounds__(RT_PLY_THREADS) plyFaceNormals(PlyDev P, long long nTris, int smooth) {
    const long long t = (long long)blockIdx.x * RT_PLY_THREADS + threadIdx.x;
    if (t >= nTris) return;
    const int32_t* i3 = P.idx + t * 3;
    const V3 v0 = ld3(P.pos + (long long)i3[0] * 3), v1 = ld3(P.pos + (long long)i3[1] * 3), v2 = ld3(P.pos + (long long)i3[2] * 3);
    const V3 n = normalize(cross(v1 - v0, v2 - v0));
    P.faceN[t * 3] = n.x; P.faceN[t * 3 + 1] = n.y; P.faceN[t * 3 + 2] = n.z;
    if (smooth)
        for (int c = 0; c < 3; c++) atomicAdd(P.count + i3[c], 1u);
}

// ---- exclusive scan of count[0 .. n] in place: per-block scan, scan of the block sums, add-back ---------------------
#define RT_SCAN_ITEMS 8
#define RT_SCAN_BLOCK (RT_PLY_THREADS * RT_SCAN_ITEMS)
__global__ void __launch_bounds__(RT_PLY_THREADS) scanBlocks(unsigned* a, long long n, unsigned* blockSums) {
    __shared__ unsigned part[RT_PLY_THREADS];
    const long long base = (long long)blockIdx.x * RT_SCAN_BLOCK + (long long)threadIdx.x * RT_SCAN_ITEMS;
    unsigned v[RT_SCAN_ITEMS], sum = 0;
    for (int k = 0; k < RT_SCAN_ITEMS; k++) { v[k] = base + k < n ? a[base + k] : 0u; sum += v[k]; }
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int d = 1; d < RT_PLY_THREADS; d <<= 1) {               // Hillis-Steele inclusive scan of the thread sums
        const unsigned add = threadIdx.x >= (unsigned)d ? part[threadIdx.x - d] : 0u;
        __syncthreads();
        part[threadIdx.x] += add;
        __syncthreads();
    }
    unsigned run = part[threadIdx.x] - sum;
    for (int k = 0; k < RT_SCAN_ITEMS; k++) { if (base + k < n) a[base + k] = run; run += v[k]; }
    if (threadIdx.x == RT_PLY_THREADS - 1) blockSums[blockIdx.x] = part[threadIdx.x];
}
__global__ void scanSums(unsigned* blockSums, int nBlocks) {      // a few thousand entries at most: one thread
    unsigned run = 0;
    for (int b = 0; b < nBlocks; b++) { const unsigned v = blockSums[b]; blockSums[b] = run; run += v; }
}
__global__ void __launch_bounds__(RT_PLY_THREADS) scanAdd(unsigned* a, long long n, const unsigned* blockSums) {
    const long long base = (long long)blockIdx.x * RT_SCAN_BLOCK + (long long)threadIdx.x * RT_SCAN_ITEMS;
    const unsigned off = blockSums[blockIdx.x];
    for (int k = 0; k < RT_SCAN_ITEMS; k++) if (base + k < n) a[base + k] += off;
}

// In-place exclusive scan of a[0 .. n); blockSums holds ceil(n / 2048) entries.  Also used by bvh_build_sah_gpu.cu.
cudaError_t launchExclusiveScan(unsigned* a, long long n, unsigned* blockSums, cudaStream_t stream) {
    const int blocks = (int)((n + RT_SCAN_BLOCK - 1) / RT_SCAN_BLOCK);
    if (blocks <= 0) return cudaSuccess;
    RT_LAUNCH(scanBlocks, blocks, RT_PLY_THREADS, stream, a, n, blockSums);
    RT_LAUNCH(scanSums, 1, 1, stream, blockSums, blocks);
    RT_LAUNCH(scanAdd, blocks, RT_PLY_THREADS, stream, a, n, blockSums);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(RT_PLY_THREADS) plyFillAdjacency(PlyDev P, long long nTris) {
    const long long t = (long long)blockIdx.x * RT_PLY_THREADS + threadIdx.x;
    if (t >= nTris) return;
    for (int c = 0; c < 3; c++) {
        const int32_t v = P.idx[t * 3 + c];
        P.entries[P.count[v] + atomicAdd(P.cursor + v, 1u)] = (unsigned)(t * 3 + c);
    }
}

// Smooth vertex normal (:316-326): the unit face normals around the vertex added in (face, corner) order — the order of
// `for t in faces { for c in 0..<3 { vtxNormals[idx] += faceNormals[t] } }` — then normalised.  A vertex used by no
// face gets normalize(0) = NaN, as in the reference.
__global__ void __launch_bounds__(RT_PLY_THREADS) plyVertexNormals(PlyDev P) {
    const long long v = (long long)blockIdx.x * RT_PLY_THREADS + threadIdx.x;
    if (v >= P.L.vertexCount) return;
    unsigned* e = P.entries + P.count[v];
    const int n = (int)(P.count[v + 1] - P.count[v]);
    for (int i = 1; i < n; i++) {                                 // insertion sort: the valence is small
        const unsigned key = e[i];
        int j = i - 1;
        while (j >= 0 && e[j] > key) { e[j + 1] = e[j]; j--; }
        e[j + 1] = key;
    }
    V3 a = mk(0, 0, 0);
    for (int i = 0; i < n; i++) a = a + ld3(P.faceN + (long long)(e[i] / 3u) * 3);
    a = normalize(a);
    P.vnorm[v * 3] = a.x; P.vnorm[v * 3 + 1] = a.y; P.vnorm[v * 3 + 2] = a.z;
}

// One Triangle per face: v0, v1, v2 and n0, n1, n2 (:270-281, :327-346)
__global__ void __launch_bounds__(RT_PLY_THREADS) plyEmit(PlyDev P, long long nTris, int smooth, double* __restrict__ triVerts,
                                                          double* __restrict__ triNormals) {
    const long long t = (long long)blockIdx.x * RT_PLY_THREADS + threadIdx.x;
    if (t >= nTris) return;
    const int32_t* i3 = P.idx + t * 3;
    for (int c = 0; c < 3; c++)
        for (int a = 0; a < 3; a++) triVerts[t * 9 + c * 3 + a] = P.pos[(long long)i3[c] * 3 + a];
    if (P.vnormFile) {
        for (int c = 0; c < 3; c++)
            for (int a = 0; a < 3; a++) triNormals[t * 9 + c * 3 + a] = P.vnormFile[(long long)i3[c] * 3 + a];
    } else if (smooth) {
        for (int c = 0; c < 3; c++)
            for (int a = 0; a < 3; a++) triNormals[t * 9 + c * 3 + a] = P.vnorm[(long long)i3[c] * 3 + a];
    } else {
        const V3 n = normalize(ld3(P.faceN + t * 3));              // :343 (normalised twice)
        for (int c = 0; c < 3; c++) { triNormals[t * 9 + c * 3] = n.x; triNormals[t * 9 + c * 3 + 1] = n.y; triNormals[t * 9 + c * 3 + 2] = n.z; }
    }
}

struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct Piece {                       // a slice of the ingest's one allocation
    void* p = nullptr;
    size_t off = 0;
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Returns 0, or -1 with `msg` set (CUDA error or malformed data); *launches counts the kernels run.
int plyIngestDevice(const rt_ply_layout& L, const void* vertexBlock, const void* faceBlock, int smooth, double* triVertsOut,
                    double* triNormalsOut, const char** msg, int64_t* launches, double* deviceMs) {
    *msg = "";
#define PCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { *msg = cudaGetErrorString(e_); return -1; } } while (0)
    const int k = L.verticesPerFace;
    const long long nV = L.vertexCount, nF = L.faceCount, nT = nF * (k == 4 ? 2 : 1);
    const bool fileNormals = L.normalOffset[0] >= 0 && L.normalOffset[1] >= 0 && L.normalOffset[2] >= 0;
    const bool needSmooth = smooth && !fileNormals;
    PlyDev P{};
    P.L = L;
    P.faceRecord = L.faceLeadBytes + L.faceCountBytes + k * L.faceIndexBytes + L.faceTrailBytes;
    P.trisPerFace = k == 4 ? 2 : 1;
    // every array in ONE allocation, carved into 256-byte aligned pieces: a driver memory call costs more than these
    // kernels do, and now and then a hundred times that (profiles/r2an_build_trace.txt)
    const int scanBlocksN = (int)((nV + 1 + RT_SCAN_BLOCK - 1) / RT_SCAN_BLOCK);
    Piece vb, fb, pos, vnf, idx, fn, cnt, cur, ent, vn, err, sums, tv, tn;
    size_t total = 0;
    auto place = [&](Piece& pc, size_t bytes) { pc.off = total; total += (bytes + 255) & ~(size_t)255; };
    place(vb, (size_t)nV * L.vertexStride); place(fb, (size_t)nF * P.faceRecord);
    place(pos, (size_t)nV * 24); place(idx, (size_t)nT * 12); place(fn, (size_t)nT * 24);
    place(cnt, (size_t)(nV + 1) * 4); place(cur, (size_t)nV * 4); place(err, 4);
    place(tv, (size_t)nT * 72); place(tn, (size_t)nT * 72);
    if (fileNormals) place(vnf, (size_t)nV * 24);
    if (needSmooth) { place(ent, (size_t)nT * 12); place(vn, (size_t)nV * 24); place(sums, (size_t)scanBlocksN * 4); }
    DevBuf arena;
    PCK(arena.alloc(total));
    for (Piece* pc : {&vb, &fb, &pos, &idx, &fn, &cnt, &cur, &err, &tv, &tn}) pc->p = static_cast<char*>(arena.p) + pc->off;
    if (fileNormals) vnf.p = static_cast<char*>(arena.p) + vnf.off;
    if (needSmooth) for (Piece* pc : {&ent, &vn, &sums}) pc->p = static_cast<char*>(arena.p) + pc->off;
    struct Ev {                      // destroyed on every return path
        cudaEvent_t e = nullptr;
        ~Ev() { if (e) cudaEventDestroy(e); }
    } ev0, ev1;
    PCK(cudaEventCreate(&ev0.e)); PCK(cudaEventCreate(&ev1.e));
    const cudaEvent_t e0 = ev0.e, e1 = ev1.e;
    PCK(cudaMemcpyAsync(vb.p, vertexBlock, (size_t)nV * L.vertexStride, cudaMemcpyHostToDevice, 0));
    PCK(cudaMemcpyAsync(fb.p, faceBlock, (size_t)nF * P.faceRecord, cudaMemcpyHostToDevice, 0));
    PCK(cudaMemsetAsync(err.p, 0, 4, 0));
    P.vblock = vb.as<unsigned char>(); P.fblock = fb.as<unsigned char>();
    P.pos = pos.as<double>(); P.vnormFile = fileNormals ? vnf.as<double>() : nullptr; P.idx = idx.as<int32_t>();
    P.faceN = fn.as<double>(); P.count = cnt.as<unsigned>(); P.cursor = cur.as<unsigned>(); P.entries = ent.as<unsigned>();
    P.vnorm = vn.as<double>(); P.error = err.as<int>();
    const int gV = (int)((nV + RT_PLY_THREADS - 1) / RT_PLY_THREADS), gF = (int)((nF + RT_PLY_THREADS - 1) / RT_PLY_THREADS),
              gT = (int)((nT + RT_PLY_THREADS - 1) / RT_PLY_THREADS);
    int64_t n = 0;
    PCK(cudaEventRecord(e0, 0));
    if (gV > 0) { RT_LAUNCH(plyVertices, gV, RT_PLY_THREADS, 0, P); n++; }
    if (gF > 0) { RT_LAUNCH(plyFaces, gF, RT_PLY_THREADS, 0, P); n++; }
    int herr = 0;
    PCK(cudaMemcpy(&herr, err.p, 4, cudaMemcpyDeviceToHost));      // indices are validated before anything dereferences them
    if (herr == 1) { *msg = "face records with differing vertex counts (device ingest takes uniform triangle or quad lists)"; return -1; }
    if (herr == 2) { *msg = "mesh face index out of range"; return -1; }
    if (gT > 0 && !fileNormals) { RT_LAUNCH(plyFaceNormals, gT, RT_PLY_THREADS, 0, P, nT, needSmooth ? 1 : 0); n++; }
    if (needSmooth && gT > 0) {
        PCK(launchExclusiveScan(P.count, nV + 1, sums.as<unsigned>(), 0));
        RT_LAUNCH(plyFillAdjacency, gT, RT_PLY_THREADS, 0, P, nT);
        RT_LAUNCH(plyVertexNormals, gV, RT_PLY_THREADS, 0, P);
        n += 5;
    }
    if (gT > 0) { RT_LAUNCH(plyEmit, gT, RT_PLY_THREADS, 0, P, nT, needSmooth ? 1 : 0, tv.as<double>(), tn.as<double>()); n++; }
    PCK(cudaEventRecord(e1, 0));
    PCK(cudaGetLastError());
    PCK(cudaMemcpy(triVertsOut, tv.p, (size_t)nT * 72, cudaMemcpyDeviceToHost));
    PCK(cudaMemcpy(triNormalsOut, tn.p, (size_t)nT * 72, cudaMemcpyDeviceToHost));
    float ms = 0;
    PCK(cudaEventElapsedTime(&ms, e0, e1));
    if (launches) *launches = n;
    if (deviceMs) *deviceMs = ms;
    return 0;
#undef PCK
}

// ---- rt_ply_ingest_static: the same kernels with NO allocation ------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Every intermediate of the pipeline
// above (positions, normals, indices, face normals, adjacency lists, scan sums, error flag) lives in ONE file-scope
// device array of fixed size, which every GPU that loads the library holds until exit; inputs and outputs are the
// caller's DEVICE buffers.  Limits: kPlyStaticMaxVertices vertices, kPlyStaticMaxTriangles triangles.
// All calls share the array and it keeps the last call's contents: whatever must start at zero (the faces-around-vertex
// counts and cursors, the error flag) is cleared by the call itself; two calls that overlap on one GPU race on it.
static const long long kPlyStaticMaxVertices = RT_PLY_STATIC_MAX_VERTICES;
static const long long kPlyStaticMaxTriangles = RT_PLY_STATIC_MAX_TRIANGLES;
static const size_t kPlyAlign = 256;
static constexpr size_t plyUp(size_t n) { return (n + kPlyAlign - 1) / kPlyAlign * kPlyAlign; }
// layout of the array: byte offsets of the intermediates at the maximum size
static constexpr size_t kOffPos = 0;                                                           // nV * 3 doubles
static constexpr size_t kOffVn = kOffPos + plyUp((size_t)RT_PLY_STATIC_MAX_VERTICES * 24);     // nV * 3 doubles: file OR smooth normals
static constexpr size_t kOffCnt = kOffVn + plyUp((size_t)RT_PLY_STATIC_MAX_VERTICES * 24);     // nV + 1 counts
static constexpr size_t kOffCur = kOffCnt + plyUp(((size_t)RT_PLY_STATIC_MAX_VERTICES + 1) * 4);
static constexpr size_t kOffSums = kOffCur + plyUp((size_t)RT_PLY_STATIC_MAX_VERTICES * 4);    // scan block sums
static constexpr size_t kOffErr = kOffSums + plyUp((((size_t)RT_PLY_STATIC_MAX_VERTICES + 1 + RT_SCAN_BLOCK - 1) / RT_SCAN_BLOCK) * 4);
static constexpr size_t kOffIdx = kOffErr + kPlyAlign;                                         // nT * 3 int32
static constexpr size_t kOffFaceN = kOffIdx + plyUp((size_t)RT_PLY_STATIC_MAX_TRIANGLES * 12); // nT * 3 doubles
static constexpr size_t kOffEnt = kOffFaceN + plyUp((size_t)RT_PLY_STATIC_MAX_TRIANGLES * 24); // nT * 3 corner ids
static constexpr size_t kPlyStaticBytes = kOffEnt + plyUp((size_t)RT_PLY_STATIC_MAX_TRIANGLES * 12);
static_assert(kPlyStaticBytes <= RT_PLY_STATIC_SCRATCH_BYTES, "the intermediates of the largest mesh fit the static array");

static __device__ __align__(256) unsigned char g_plyScratch[RT_PLY_STATIC_SCRATCH_BYTES];

// In host code `g_plyScratch` is a host-side stand-in, not a device address: the address the kernels need comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static const int kPlyMaxDevices = 64;
static unsigned char* g_plyScratchPtr[kPlyMaxDevices];
static bool g_plyScratchKnown[kPlyMaxDevices];

static cudaError_t plyScratchFor(int device, unsigned char** out) {
    if (!g_plyScratchKnown[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_plyScratch);
        if (e != cudaSuccess) return e;
        g_plyScratchPtr[device] = static_cast<unsigned char*>(p);
        g_plyScratchKnown[device] = true;
    }
    *out = g_plyScratchPtr[device];
    return cudaSuccess;
}

// The pipeline of plyIngestDevice on `stream` with every intermediate in g_plyScratch.  dVertexBlock / dFaceBlock /
// dTriVertsOut / dTriNormalsOut are DEVICE pointers of the caller (valid on `device`, which must be current).
// Calls no cudaMalloc / cudaFree / event or stream creation.  Blocking: the face indices are validated (one 4-byte
// read-back) before any kernel dereferences them, and the call returns when the outputs are complete.
// Returns 0, -1 with `msg` set (CUDA error or malformed data), -2 when the mesh exceeds the static limits.
int plyIngestStatic(const rt_ply_layout& L, const void* dVertexBlock, const void* dFaceBlock, int smooth, int device,
                    double* dTriVertsOut, double* dTriNormalsOut, cudaStream_t stream, const char** msg, int64_t* launches) {
    *msg = "";
#define PCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { *msg = cudaGetErrorString(e_); return -1; } } while (0)
    const int k = L.verticesPerFace;
    const long long nV = L.vertexCount, nF = L.faceCount, nT = nF * (k == 4 ? 2 : 1);
    if (device < 0 || device >= kPlyMaxDevices) { *msg = "device index outside 0..63"; return -1; }
    if (nV > kPlyStaticMaxVertices || nT > kPlyStaticMaxTriangles) {
        *msg = "mesh exceeds the static scratch of rt_ply_ingest_static (RT_PLY_STATIC_MAX_VERTICES vertices, "
               "RT_PLY_STATIC_MAX_TRIANGLES triangles): use rt_ply_ingest";
        return -2;
    }
    const bool fileNormals = L.normalOffset[0] >= 0 && L.normalOffset[1] >= 0 && L.normalOffset[2] >= 0;
    const bool needSmooth = smooth && !fileNormals;
    unsigned char* base = nullptr;
    {
        static std::mutex m;                        // the per-device address table is filled once per device
        std::lock_guard<std::mutex> lk(m);
        PCK(plyScratchFor(device, &base));
    }
    PlyDev P{};
    P.L = L;
    P.faceRecord = L.faceLeadBytes + L.faceCountBytes + k * L.faceIndexBytes + L.faceTrailBytes;
    P.trisPerFace = k == 4 ? 2 : 1;
    P.vblock = static_cast<const unsigned char*>(dVertexBlock);
    P.fblock = static_cast<const unsigned char*>(dFaceBlock);
    P.pos = reinterpret_cast<double*>(base + kOffPos);
    P.vnormFile = fileNormals ? reinterpret_cast<double*>(base + kOffVn) : nullptr;
    P.vnorm = reinterpret_cast<double*>(base + kOffVn);           // file normals and smooth normals exclude each other
    P.count = reinterpret_cast<unsigned*>(base + kOffCnt);
    P.cursor = reinterpret_cast<unsigned*>(base + kOffCur);
    unsigned* sums = reinterpret_cast<unsigned*>(base + kOffSums);
    P.error = reinterpret_cast<int*>(base + kOffErr);
    P.idx = reinterpret_cast<int32_t*>(base + kOffIdx);
    P.faceN = reinterpret_cast<double*>(base + kOffFaceN);
    P.entries = reinterpret_cast<unsigned*>(base + kOffEnt);
    const int gV = (int)((nV + RT_PLY_THREADS - 1) / RT_PLY_THREADS), gF = (int)((nF + RT_PLY_THREADS - 1) / RT_PLY_THREADS),
              gT = (int)((nT + RT_PLY_THREADS - 1) / RT_PLY_THREADS);
    int64_t n = 0;
    // what must start at zero: the error flag here; counts and cursors of vertices 0 .. nV by plyVertices
    PCK(cudaMemsetAsync(P.error, 0, 4, stream));
    if (gV > 0) { RT_LAUNCH(plyVertices, gV, RT_PLY_THREADS, stream, P); n++; }
    if (gF > 0) { RT_LAUNCH(plyFaces, gF, RT_PLY_THREADS, stream, P); n++; }
    int herr = 0;
    PCK(cudaMemcpyAsync(&herr, P.error, 4, cudaMemcpyDeviceToHost, stream));
    PCK(cudaStreamSynchronize(stream));                            // indices are validated before anything dereferences them
    if (herr == 1) { *msg = "face records with differing vertex counts (device ingest takes uniform triangle or quad lists)"; return -1; }
    if (herr == 2) { *msg = "mesh face index out of range"; return -1; }
    if (gT > 0 && !fileNormals) { RT_LAUNCH(plyFaceNormals, gT, RT_PLY_THREADS, stream, P, nT, needSmooth ? 1 : 0); n++; }
    if (needSmooth && gT > 0) {
        PCK(launchExclusiveScan(P.count, nV + 1, sums, stream));
        RT_LAUNCH(plyFillAdjacency, gT, RT_PLY_THREADS, stream, P, nT);
        RT_LAUNCH(plyVertexNormals, gV, RT_PLY_THREADS, stream, P);
        n += 5;
    }
    if (gT > 0) { RT_LAUNCH(plyEmit, gT, RT_PLY_THREADS, stream, P, nT, needSmooth ? 1 : 0, dTriVertsOut, dTriNormalsOut); n++; }
    PCK(cudaGetLastError());
    PCK(cudaStreamSynchronize(stream));
    if (launches) *launches = n;
    return 0;
#undef PCK
}

}  // namespace rtg
