// rt_api_render.inl — render entry points (included by rt_api.cu).

namespace {

struct H3 { double x, y, z; };
inline H3 h3(const double* p) { return {p[0], p[1], p[2]}; }
inline H3 operator-(H3 a, H3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline H3 operator+(H3 a, H3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline H3 operator*(H3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline H3 neg(H3 a) { return {-a.x, -a.y, -a.z}; }
inline double hdot(H3 a, H3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline H3 hcross(H3 a, H3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline H3 hnorm(H3 a) { return a * (1.0 / std::sqrt(hdot(a, a))); }
inline void put(double* d, H3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

// makeCameraBasis (Object+Extension.swift:382-427) + du, dv, m, q00 (:67-70)
int buildCamera(const rt_camera* c, const rt_render_params* prm, CameraParams& cp) {
    std::memset(&cp, 0, sizeof cp);
    const int W = c->width > 1 ? c->width : 1, H = c->height > 1 ? c->height : 1;  // max(1, ...) :62-63
    const double aspect = (double)W / (double)H;
    const H3 e = h3(c->position);
    const double nd = c->nearDistance;
    double l, r, b, t;
    H3 w, u, v;
    if (c->isLookAt) {
        const H3 gazeNorm = hnorm(h3(c->gazePoint) - h3(c->position));
        w = neg(gazeNorm);
        const H3 upNorm = hnorm(h3(c->up));
        u = hnorm(hcross(upNorm, w));
        v = hnorm(hcross(w, u));
        if (c->hasFovy) {
            const double fovYRad = (c->fovy * M_PI) / (2.0 * 180.0);
            t = nd * std::tan(fovYRad);
        } else {
            t = nd * 0.5;
        }
        b = -t; r = t * aspect; l = -r;
    } else {
        l = c->nearPlane[0]; r = c->nearPlane[1]; b = c->nearPlane[2]; t = c->nearPlane[3];
        const H3 gazeNorm = hnorm(h3(c->gaze));
        w = neg(gazeNorm);
        const H3 upNorm = hnorm(h3(c->up));
        u = hnorm(hcross(upNorm, w));
        v = hnorm(hcross(w, u));
    }
    const H3 m = e - w * nd;
    const H3 q00 = (m + u * l) + v * t;
    put(cp.eye, e); put(cp.u, u); put(cp.v, v); put(cp.w, w); put(cp.q00, q00); put(cp.m, m);
    cp.du = (r - l) / (double)W;
    cp.dv = (t - b) / (double)H;
    cp.aperture = c->apertureSize; cp.focus = c->focusDistance;
    cp.dof = (c->apertureSize > 0 && c->focusDistance > 0) ? 1 : 0;
    cp.width = W; cp.height = H;
    cp.samples = c->numSamples > 1 ? c->numSamples : 1;
    cp.n = (int)std::sqrt((double)cp.samples);
    cp.shardIndex = 0; cp.shardCount = 1; cp.shardTiles = 0;
    if (prm && prm->shardCount > 1) {
        if (prm->shardIndex < 0 || prm->shardIndex >= prm->shardCount) return fail(RT_ERR_INVALID, "shardIndex outside [0, shardCount)");
        if (prm->shardMode < 0 || prm->shardMode > RT_SHARD_TILES) return fail(RT_ERR_INVALID, "shardMode outside 0..1");
        cp.shardIndex = prm->shardIndex; cp.shardCount = prm->shardCount; cp.shardTiles = prm->shardMode == RT_SHARD_TILES ? 1 : 0;
    }
    return RT_OK;
}

template <class T>
int ensure(T** p, size_t* have, size_t need) {
    if (*have >= need && *p) return RT_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    CK(cudaMalloc((void**)p, need));
    *have = need;
    return RT_OK;
}

}  // namespace

extern "C" {

#ifdef RT_WARPTIMES
static unsigned long long* g_wt = nullptr;
static size_t g_wtBytes = 0;
static unsigned long long g_wtWarps = 0;
#endif

int32_t rt_render_device(rt_scene_t* sc, const rt_camera* camera, const rt_render_params* params, void* dRgbaOut,
                         void* dLinearOut, void* dIdsOut, void* cudaStream) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    if (!camera) return fail(RT_ERR_BAD_CAMERA, "Invalid camera index");
    if (sc->pending) return fail(RT_ERR_INVALID, "a render is already in flight on this scene (call rt_render_finish)");
    CK(cudaSetDevice(sc->device));
    CameraParams cp;
    int rc = buildCamera(camera, params, cp);
    if (rc) return rc;
    const bool counters = params && params->collectCounters;
    const bool area = sc->ds.nAreaLights > 0;
    int variant = (sc->glossy || area || sc->motion || sc->nonFinite) ? KV_GENERAL : KV_FAST;
    if (params && params->kernelVariant != 0) {
        if (params->kernelVariant < 1 || params->kernelVariant > KV_COUNT) return fail(RT_ERR_INVALID, "kernelVariant outside 0..2");
        variant = params->kernelVariant - 1;
        if (sc->glossy && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "the fast kernel needs roughness == 0 on every material (RNG skip-ahead)");
        if (area && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "the fast kernel does not carry the area lights' per-chunk jitter index");
        if (sc->motion && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "the fast kernel does not carry the ray time of motion blur");
        if (sc->nonFinite && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "a light, material or ambient term is not finite: the fast kernel has no non-finite guard (Object+Extension.swift:277-280)");
    }
    // the area lights' jitter index runs through a whole 8-row chunk (Object+Extension.swift:288): chunks stay whole
    if (area && cp.shardCount > 1 && cp.shardTiles)
        return fail(RT_ERR_UNSUPPORTED, "scenes with area lights shard by 8-row chunks only (RT_SHARD_ROWS)");
    const int blocks = sc->blocks[variant][counters ? 1 : 0];
    cudaStream_t st = (cudaStream_t)cudaStream;
    // scenes with recursion whose child rays are rare (big meshes, a few reflective objects) render in two launches;
    // RTCUDA_SPLIT=0|1 overrides the choice (measurements, tests)
    bool split = variant == KV_FAST && sc->recurse && !area && !sc->smallGeometry;
    if (const char* e = std::getenv("RTCUDA_SPLIT")) split = variant == KV_FAST && sc->recurse && !area && std::atoi(e) != 0;

    const size_t threads = (size_t)blocks * renderBlockSize(variant);
    // general kernel: one depth-first stack of maxDepth + 1 entries per thread; fast kernel: per warp a shared queue of
    // 32 x (maxDepth + 3 + 2 per extra pass of a batch) entries plus maxDepth + 2 private spill-over entries per lane
    // (rt_shade.cuh: WarpQueue)
    // an area-light frame of a scene the fast kernel can follow counts its jitter-table draws there (pass 1), then shades
    // in the general kernel (pass 2): the path stack holds the larger of the two
    const bool fastCount = area && variant == KV_GENERAL && !sc->glossy && !sc->motion && !std::getenv("RTCUDA_AREA_COUNT_GENERAL");
    const int blocksCount = sc->blocks[KV_FAST][0];
    const size_t threadsCount = fastCount ? (size_t)blocksCount * renderBlockSize(KV_FAST) : 0;
    const size_t need = (size_t)(2 * sc->ds.maxDepth + 5 + 2 * (RT_BATCH_PASSES - 1)) * kPathFields * std::max(threads, threadsCount) * sizeof(double);
    if ((rc = ensure(&sc->dPathStack, &sc->pathStackBytes, need))) return rc;

    RenderTargets rt{};
    rt.rgba = (uint8_t*)dRgbaOut; rt.linear = (double*)dLinearOut; rt.ids = (HitId*)dIdsOut;
    rt.pathStack = sc->dPathStack; rt.counters = sc->dCounters;
    rt.tileCounter = sc->dTileCounter;
    rt.motion = sc->motion ? 1 : 0;
    rt.recurse = sc->recurse ? (sc->smallGeometry ? 2 : 1) : 0;
    if (std::getenv("RTCUDA_FORCE_RECURSE")) rt.recurse = 1;       // measurements only: the queue instantiation on a scene without child rays
    // Tile schedule: the frame records its cost per tile, the next frame of the same geometry draws the longest
    // tiles first.
    const int nTiles = ShardTiles(cp.width, cp.height, cp.shardIndex, cp.shardCount, cp.shardTiles).count;
    const bool adaptive = !(params && params->scheduleMode == RT_SCHEDULE_SPATIAL) && nTiles > 0;
    if (params && (params->scheduleMode < 0 || params->scheduleMode > RT_SCHEDULE_SPATIAL)) return fail(RT_ERR_INVALID, "scheduleMode outside 0..1");
    sc->pendingScheduled = 0; sc->pendingProbe = false;
    if (adaptive) {
        const int key[8] = {cp.width, cp.height, cp.n, cp.dof, cp.shardIndex, cp.shardCount * 2 + cp.shardTiles, variant, nTiles};
        const bool grown = sc->tileCostBytes < (size_t)nTiles * 4 || sc->tileOrderBytes < (size_t)nTiles * 4;
        if ((rc = ensure(&sc->dTileCost, &sc->tileCostBytes, (size_t)nTiles * 4))) return rc;
        if ((rc = ensure(&sc->dTileOrder, &sc->tileOrderBytes, (size_t)nTiles * 4))) return rc;
        if ((rc = ensure(&sc->dSchedWork, &sc->schedWorkBytes, scheduleWorkBytes(nTiles)))) return rc;
        if (grown || std::memcmp(key, sc->schedKey, sizeof key) != 0) sc->schedValid = false;
        std::memcpy(sc->schedKey, key, sizeof key);
        bool probed = false;
        // RTCUDA_PROBE=1: the first frame of a geometry takes its order from one probe ray per tile; =2 (measurements): EVERY
        // frame does, as if it were the first.  Off by default: at 8 GPUs a probe-ordered C4 frame takes 1.73 ms against
        // 1.68 ms in the spatial order and 1.51 ms in the measured order (profiles/r2z) — three more launches and an
        // estimate from one ray per tile cost more than the order they produce gains.
        const char* pe = std::getenv("RTCUDA_PROBE");
        const int probeMode = pe ? std::atoi(pe) : 0;
        if (!sc->schedValid || probeMode == 2) {
            if (sc->sideBusy) CK(cudaStreamWaitEvent(st, sc->evSide, 0));
            // no frame of this geometry has measured its tiles yet: one probe ray per tile can estimate them
            if (probeMode != 0 && !sc->motion && nTiles >= 256) {
                CK(launchTileProbe(sc->ds, cp, sc->dTileCost, nTiles, st));
                CK(launchSchedule(sc->dTileCost, sc->dTileOrder, sc->dSchedWork, nTiles, st));      // leaves tileCost zeroed
                probed = true;
            } else {
                CK(cudaMemsetAsync(sc->dTileCost, 0, (size_t)nTiles * 4, st));
            }
        }
        rt.tileCost = sc->dTileCost;
        rt.tileOrder = (sc->schedValid || probed) ? sc->dTileOrder : nullptr;
        sc->pendingScheduled = (sc->schedValid || probed) ? nTiles : 0;
        sc->pendingProbe = probed;
    }
#ifdef RT_WARPTIMES
    {   // profiling build: per-warp timestamps, dumped by rt_render_finish to $RT_WARPTIMES_OUT
        static unsigned long long* dWT = nullptr; static size_t wtBytes = 0;
        const size_t needWT = (threads / 32 * 4 + (size_t)((cp.width + 7) / 8) * ((cp.height + 7) / 8 * 2) * 16 * 4) * sizeof(unsigned long long);
        if ((rc = ensure(&dWT, &wtBytes, needWT))) return rc;
        CK(cudaMemsetAsync(dWT, 0, needWT, st));
        rt.warpTimes = dWT; g_wt = dWT; g_wtBytes = needWT; g_wtWarps = threads / 32;
    }
#endif
    // The frame on the caller's stream is [memset, render kernel]; what follows the kernel — counters to the host,
    // tile order of the next frame — runs on a side stream, off the caller's critical path (a frame split over 8
    // GPUs lasts 1.5 ms: five more stream operations were 2 % of it).  The next frame waits for that epilogue.
    if (sc->sideBusy) CK(cudaStreamWaitEvent(st, sc->evSide, 0));
    CK(cudaEventRecord(sc->evStart, st));
    CK(cudaMemsetAsync(sc->dCounters, 0, kFrameStateBytes, st));
    CK(cudaEventRecord(sc->evReset, st));
    {   // what the work counter counts up to (progress / cancel of rt_render): see the kernels' unit rules
        const int ns = cp.n * cp.n;
        const int unitsPerTile = variant == KV_FAST ? FastUnits(ns, sc->recurse).perTile : 1;
        sc->passUnits[1] = (int64_t)nTiles * unitsPerTile;
        sc->passWarps[1] = (int64_t)(threads / 32);
        // area lights: two passes over the tiles (the work counter is reset in between); pass 0 is the counting pass
        sc->passUnits[0] = !area ? 0 : fastCount ? (int64_t)nTiles * FastUnits(ns, sc->recurse).perTile : sc->passUnits[1];
        sc->passWarps[0] = !area ? 0 : fastCount ? (int64_t)(threadsCount / 32) : sc->passWarps[1];
        if (split) {        // the first launch draws every unit, the second the deferred ones (their number is known on the device only)
            sc->passUnits[0] = sc->passUnits[1]; sc->passWarps[0] = sc->passWarps[1];
            sc->passUnits[1] = sc->passUnits[0] / 16 + 1;
            if ((rc = ensure(&sc->dDefer, &sc->deferBytes, ((size_t)sc->passUnits[0] + 1) * sizeof(int)))) return rc;
            CK(cudaMemsetAsync(sc->dDefer, 0, sizeof(int), st));
        }
        sc->pendingUnits = sc->passUnits[0] + sc->passUnits[1];
    }
    if (area) {
        // pass 1 follows every path and counts its jitter-table draws per pixel; the prefix kernel turns the counts
        // into each pixel's first index within its chunk; pass 2 (below) is the frame
        if ((rc = ensure(&sc->dAreaIndex, &sc->areaIndexBytes, (size_t)cp.width * cp.height * sizeof(unsigned)))) return rc;
        rt.areaIndex = sc->dAreaIndex;
        rt.areaPass = 1;
        if (fastCount) CK(launchRender(KV_FAST, sc->ds, cp, rt, blocksCount, false, st));
        else CK(launchRender(variant, sc->ds, cp, rt, blocks, counters, st));
        CK(launchAreaPrefix(sc->dAreaIndex, cp.width, cp.height, cp.shardIndex, cp.shardCount, st));
        CK(cudaMemsetAsync(sc->dTileCounter, 0, sizeof(int), st));
        rt.areaPass = 2;
    }
    if (split) {
        // two launches (rt_render_kernel.cuh: DEFER): depth 0 of every unit in the lean instantiation, which defers the units
        // whose primary hits spawn child rays; then the queue instantiation over the deferred units
        rt.deferCount = sc->dDefer; rt.deferList = sc->dDefer + 1;
        rt.deferPhase = 1; rt.recurse = 0;
        CK(launchRender(variant, sc->ds, cp, rt, blocks, counters, st));
        CK(cudaMemsetAsync(sc->dTileCounter, 0, sizeof(int), st));
        rt.deferPhase = 2; rt.recurse = sc->smallGeometry ? 2 : 1;
    }
    CK(launchRender(variant, sc->ds, cp, rt, blocks, counters, st));
    CK(cudaEventRecord(sc->evStop, st));
    CK(cudaStreamWaitEvent(sc->side, sc->evStop, 0));
    CK(cudaMemcpyAsync(sc->hCounters, sc->dCounters, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, sc->side));
    if (adaptive) {
        CK(launchSchedule(sc->dTileCost, sc->dTileOrder, sc->dSchedWork, nTiles, sc->side));   // costs of this frame -> order of the next
        sc->schedValid = true;
    }
    CK(cudaEventRecord(sc->evSide, sc->side));
    sc->sideBusy = true;
    sc->pending = true; sc->pendingCounters = counters; sc->pendingStream = st;
    sc->launches = (adaptive ? 3 : 1) + (area ? 2 : 0) + (adaptive && sc->pendingProbe ? 3 : 0) + (split ? 1 : 0);
    return RT_OK;
}

int32_t rt_render_finish(rt_scene_t* sc, rt_stats* st) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    if (!sc->pending) return fail(RT_ERR_INVALID, "no render in flight");
    CK(cudaSetDevice(sc->device));
    sc->pending = false;
    CK(cudaEventSynchronize(sc->evStop));
    CK(cudaEventSynchronize(sc->evSide));       // counters are on the host, the next frame's tile order is ready
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, sc->evStart, sc->evStop));
#ifdef RT_WARPTIMES
    if (const char* path = getenv("RT_WARPTIMES_OUT")) {
        std::vector<unsigned long long> h(g_wtBytes / 8);
        CK(cudaMemcpy(h.data(), g_wt, g_wtBytes, cudaMemcpyDeviceToHost));
        if (FILE* f = fopen(path, "wb")) { fwrite(&g_wtWarps, 8, 1, f); fwrite(h.data(), 8, h.size(), f); fclose(f); }
    }
#endif
    if (st) {
        std::memset(st, 0, sizeof *st);
        st->meshes = sc->meshes; st->triangles = sc->info.triangles; st->spheres = sc->info.spheres; st->planes = sc->info.planes;
        st->raysPrimary = (int64_t)sc->hCounters[0]; st->raysShadow = (int64_t)sc->hCounters[1];
        st->raysSecondary = (int64_t)sc->hCounters[2];
        st->rays = st->raysPrimary + st->raysShadow + st->raysSecondary;
        st->deviceMs = ms;
        st->kernelLaunches = sc->launches;
        st->tilesScheduled = sc->pendingScheduled;
        if (sc->pendingCounters) {
            st->boxTests = (int64_t)sc->hCounters[3]; st->triTests = (int64_t)sc->hCounters[4];
            st->sphereTests = (int64_t)sc->hCounters[5]; st->planeTests = (int64_t)sc->hCounters[6];
            st->instEntries = (int64_t)sc->hCounters[7]; st->maxStack = (int64_t)sc->hCounters[8];
            st->smoothHits = (int64_t)sc->hCounters[9];
        }
    }
    return RT_OK;
}

// ---- rt_render: one frame into host buffers, on the scene's own stream(s) ------------------------------------------

namespace {

struct FrameShape { size_t W, H, nn, chunks; };

// Device -> host copy of the rows replica r of N renders: its 8-row chunks c = r, r + N, ... are one strided 2-D copy
// (pitch = N chunks) plus, when the image's last chunk is ragged and belongs to r, one copy of that chunk's rows.
int copyOwnedRows(void* host, const void* dev, size_t rowBytes, const FrameShape& f, int r, int N, cudaStream_t st) {
    if (N <= 1) { CK(cudaMemcpyAsync(host, dev, rowBytes * f.H, cudaMemcpyDeviceToHost, st)); return RT_OK; }
    const size_t chunkBytes = rowBytes * 8, pitch = chunkBytes * (size_t)N;
    const size_t fullChunks = f.H / 8;                                   // chunks with all 8 rows
    const size_t mineFull = (size_t)r < fullChunks ? (fullChunks - (size_t)r + (size_t)N - 1) / (size_t)N : 0;
    char* h = (char*)host + chunkBytes * (size_t)r;
    const char* d = (const char*)dev + chunkBytes * (size_t)r;
    if (mineFull > 0) CK(cudaMemcpy2DAsync(h, pitch, d, pitch, chunkBytes, mineFull, cudaMemcpyDeviceToHost, st));
    if (f.chunks > fullChunks && (int)(fullChunks % (size_t)N) == r)       // the ragged last chunk is this replica's
        CK(cudaMemcpyAsync((char*)host + chunkBytes * fullChunks, (const char*)dev + chunkBytes * fullChunks,
                           rowBytes * (f.H - fullChunks * 8), cudaMemcpyDeviceToHost, st));
    return RT_OK;
}

int copyFrame(rt_scene* rep, int r, int N, const FrameShape& f, uint8_t* rgbaOut, double* linearOut, rt_hit_id* idsOut);

// Everything one GPU does for a frame of rt_render: buffers, launch, copies of its rows.  Enqueues only — except that a
// device -> host copy into PAGEABLE memory blocks the calling thread until the kernel before it has finished, which is
// why a caller that wants to report progress from this thread defers the copies (copyFrame) until the kernel is done.
int enqueueFrame(rt_scene* rep, int r, int N, const rt_camera* camera, const rt_render_params* params, const FrameShape& f,
                 uint8_t* rgbaOut, double* linearOut, rt_hit_id* idsOut, bool deferCopies = false) {
    CK(cudaSetDevice(rep->device));
    int rc;
    if (rgbaOut && (rc = ensure(&rep->dRgba, &rep->rgbaBytes, f.W * f.H * 4))) return rc;
    if (linearOut && (rc = ensure(&rep->dLinear, &rep->linearBytes, f.W * f.H * 3 * sizeof(double)))) return rc;
    if (idsOut && (rc = ensure(&rep->dIds, &rep->idsBytes, f.W * f.H * f.nn * sizeof(HitId)))) return rc;
    rt_render_params prm{};
    if (params) prm = *params;
    if (N > 1) { prm.shardIndex = r; prm.shardCount = N; prm.shardMode = RT_SHARD_ROWS; }
    // rows outside the caller's own shard (multi-process sharding) are defined as zero in the host buffers
    if (N == 1 && prm.shardCount > 1) {
        if (rgbaOut) CK(cudaMemsetAsync(rep->dRgba, 0, f.W * f.H * 4, rep->stream));
        if (linearOut) CK(cudaMemsetAsync(rep->dLinear, 0, f.W * f.H * 3 * sizeof(double), rep->stream));
        if (idsOut) CK(cudaMemsetAsync(rep->dIds, 0, f.W * f.H * f.nn * sizeof(HitId), rep->stream));
    }
    rc = rt_render_device(rep, camera, &prm, rgbaOut ? rep->dRgba : nullptr, linearOut ? rep->dLinear : nullptr,
                          idsOut ? rep->dIds : nullptr, rep->stream);
    if (rc) return rc;
    if (deferCopies) return RT_OK;
    if ((rc = copyFrame(rep, r, N, f, rgbaOut, linearOut, idsOut))) { rt_render_finish(rep, nullptr); return rc; }
    return RT_OK;
}

int copyFrame(rt_scene* rep, int r, int N, const FrameShape& f, uint8_t* rgbaOut, double* linearOut, rt_hit_id* idsOut) {
    CK(cudaSetDevice(rep->device));
    cudaError_t ce = cudaSuccess;
    auto copy = [&](void* h, const void* d, size_t rowBytes) {
        if (!h || ce != cudaSuccess) return;
        if (copyOwnedRows(h, d, rowBytes, f, r, N, rep->stream) != RT_OK) ce = cudaErrorUnknown;
    };
    copy(rgbaOut, rep->dRgba, f.W * 4);
    copy(linearOut, rep->dLinear, f.W * 3 * sizeof(double));
    copy(idsOut, rep->dIds, f.W * f.nn * sizeof(HitId));
    return ce != cudaSuccess ? RT_ERR_CUDA : RT_OK;
}

// Share of the frame's work units already FINISHED on one GPU, from its work counter (drawn units minus the ones the
// resident warps may still hold); -1 while the frame has not started.  cancel: push the counter past its end.
double pollProgress(rt_scene* rep, bool cancel, int& lastDrawn, int64_t& base) {
    if (cudaSetDevice(rep->device) != cudaSuccess) return -1;
    if (cudaEventQuery(rep->evReset) != cudaSuccess) { cudaGetLastError(); return -1; }
    if (cancel) cudaMemcpyAsync(rep->dTileCounter, rep->hPoll + 1, sizeof(int), cudaMemcpyHostToDevice, rep->poll);
    cudaMemcpyAsync(rep->hPoll, rep->dTileCounter, sizeof(int), cudaMemcpyDeviceToHost, rep->poll);
    if (cudaStreamSynchronize(rep->poll) != cudaSuccess) { cudaGetLastError(); return -1; }
    const int drawn = rep->hPoll[0];
    if (drawn < lastDrawn && base == 0) base = rep->passUnits[0];   // area lights: the second pass started (counter reset)
    lastDrawn = drawn;
    const int pass = (rep->passUnits[0] > 0 && base == 0) ? 0 : 1;
    const int64_t passUnits = rep->passUnits[pass];
    int64_t done = (int64_t)drawn - rep->passWarps[pass];
    if (done < 0) done = 0;
    if (done > passUnits) done = passUnits;
    return rep->pendingUnits > 0 ? (double)(base + done) / (double)rep->pendingUnits : 1.0;
}

}  // namespace

int32_t rt_render(rt_scene_t* sc, const rt_camera* camera, const rt_render_params* params, uint8_t* rgbaOut,
                  double* linearOut, rt_hit_id* idsOut, rt_stats* statsOut) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    if (!camera) return fail(RT_ERR_BAD_CAMERA, "Invalid camera index");
    if (!rgbaOut && !linearOut) return fail(RT_ERR_INVALID, "one of rgbaOut / linearOut is required");
    const double w0 = nowMs();
    const int N = (int)sc->replicas.size();
    if (N > 1 && params && params->shardCount > 1)
        return fail(RT_ERR_INVALID, "a multi-device scene shards the frame itself: shardCount must be 0 or 1");
    FrameShape f;
    f.W = camera->width > 1 ? camera->width : 1; f.H = camera->height > 1 ? camera->height : 1;
    const int samples = camera->numSamples > 1 ? camera->numSamples : 1;
    const size_t n = (size_t)std::sqrt((double)samples);
    f.nn = n * n; f.chunks = (f.H + 7) / 8;
    const int32_t H = (int32_t)f.H;
    rt_progress_fn progress = params ? params->progress : nullptr;
    void* user = params ? params->progressUser : nullptr;
    if (progress && progress(0, H, user)) return fail(RT_ERR_CANCELLED, "cancelled");

    // ---- enqueue on every GPU (N > 1: each from its own host thread, so the launches do not queue up behind one another)
    std::vector<rt_stats> stats((size_t)N);
    auto frameJob = [&](int r) -> int {
        rt_scene* rep = sc->replicas[(size_t)r];
        int rc = enqueueFrame(rep, r, N, camera, params, f, rgbaOut, linearOut, idsOut);
        if (rc) return rc;
        rc = rt_render_finish(rep, &stats[(size_t)r]);       // waits for the kernels; the copies follow on the same stream
        cudaError_t ce = cudaStreamSynchronize(rep->stream);
        if (rc) return rc;
        CK(ce);
        return RT_OK;
    };
    int rc = RT_OK;
    bool cancelled = false;
    if (N > 1) sc->pool->launch(frameJob);
    else if ((rc = enqueueFrame(sc, 0, 1, camera, params, f, rgbaOut, linearOut, idsOut, progress != nullptr))) return rc;

    // ---- while the frame is in flight: progress and cancellation (RayTracer.swift:172-181)
    if (progress) {
        std::vector<int> lastDrawn((size_t)N, 0);
        std::vector<int64_t> base((size_t)N, 0);
        int32_t reported = 0;
        for (;;) {
            bool running = false;
            if (N > 1) { std::lock_guard<std::mutex> lk(sc->pool->m); running = sc->pool->remaining > 0; }
            else { cudaSetDevice(sc->device); running = cudaEventQuery(sc->evStop) == cudaErrorNotReady; cudaGetLastError(); }
            if (!running) break;
            double sum = 0;
            bool started = true;
            for (int r = 0; r < N; r++) {
                rt_scene* rep = sc->replicas[(size_t)r];
                if (N > 1 && !rep->pending) { started = false; continue; }       // its worker has not launched yet / is done
                const double p = pollProgress(rep, cancelled, lastDrawn[(size_t)r], base[(size_t)r]);
                if (p < 0) started = false; else sum += p;
            }
            if (started && !cancelled) {
                const int32_t rows = std::min<int32_t>(H, (int32_t)((sum / N) * (double)f.chunks) * 8);
                if (rows > reported) {
                    reported = rows;
                    if (progress(rows, H, user)) cancelled = true;                // takes effect at the next poll
                }
            }
            std::this_thread::sleep_for(std::chrono::microseconds(cancelled ? 20 : 100));
        }
    }
    if (N > 1) {
        sc->pool->wait();
        for (int r = 0; r < N; r++)
            if (sc->pool->rc[(size_t)r] != RT_OK && rc == RT_OK) { rc = sc->pool->rc[(size_t)r]; g_err = sc->pool->err[(size_t)r]; }
        cudaSetDevice(sc->device);
    } else {
        if (progress && !cancelled && (rc = copyFrame(sc, 0, 1, f, rgbaOut, linearOut, idsOut))) { rt_render_finish(sc, nullptr); return rc; }
        rc = rt_render_finish(sc, &stats[0]);
        cudaError_t ce = cudaStreamSynchronize(sc->stream);
        if (rc == RT_OK && ce != cudaSuccess) { g_err = std::string("rt_render: ") + cudaGetErrorString(ce); rc = RT_ERR_CUDA; }
    }
    if (cancelled) {
        for (rt_scene* rep : sc->replicas) rep->schedValid = false;       // the costs of a cut-short frame order nothing
        return fail(RT_ERR_CANCELLED, "cancelled");
    }
    if (rc) return rc;
    if (statsOut) {
        rt_stats& o = *statsOut;
        o = stats[0];
        for (int r = 1; r < N; r++) {
            const rt_stats& s2 = stats[(size_t)r];
            o.rays += s2.rays; o.raysPrimary += s2.raysPrimary; o.raysShadow += s2.raysShadow; o.raysSecondary += s2.raysSecondary;
            o.kernelLaunches += s2.kernelLaunches; o.tilesScheduled += s2.tilesScheduled;
            o.boxTests += s2.boxTests; o.triTests += s2.triTests; o.sphereTests += s2.sphereTests; o.planeTests += s2.planeTests;
            o.instEntries += s2.instEntries; o.smoothHits += s2.smoothHits;
            if (s2.maxStack > o.maxStack) o.maxStack = s2.maxStack;
            if (s2.deviceMs > o.deviceMs) o.deviceMs = s2.deviceMs;      // the GPUs run side by side: the frame takes the slowest
        }
        o.milliseconds = (int64_t)(nowMs() - w0);
    }
    if (progress) progress(H, H, user);
    return RT_OK;
}

int32_t rt_shared_buffer_create(int32_t device, uint64_t bytes, void** devPtrOut, uint8_t* handleOut) {
    if (!devPtrOut || !handleOut || bytes == 0) return fail(RT_ERR_INVALID, "rt_shared_buffer_create: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "IPC handle size");
    CK(cudaSetDevice(device));
    void* p = nullptr;
    CK(cudaMalloc(&p, bytes));
    CK(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); g_err = std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e); return RT_ERR_CUDA; }
    std::memcpy(handleOut, &h, sizeof h);
    *devPtrOut = p;
    return RT_OK;
}

int32_t rt_shared_buffer_open(int32_t device, const uint8_t* handle, void** devPtrOut) {
    if (!devPtrOut || !handle) return fail(RT_ERR_INVALID, "rt_shared_buffer_open: bad argument");
    CK(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    void* p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *devPtrOut = p;
    return RT_OK;
}

int32_t rt_shared_buffer_close(int32_t device, void* devPtr, int32_t wasOpened) {
    if (!devPtr) return RT_OK;
    CK(cudaSetDevice(device));
    if (wasOpened) CK(cudaIpcCloseMemHandle(devPtr));
    else CK(cudaFree(devPtr));
    return RT_OK;
}

}  // extern "C"
