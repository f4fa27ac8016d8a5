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
    int variant = (sc->glossy || area || sc->motion) ? KV_GENERAL : KV_FAST;
    if (params && params->kernelVariant != 0) {
        if (params->kernelVariant < 1 || params->kernelVariant > KV_COUNT) return fail(RT_ERR_INVALID, "kernelVariant outside 0..2");
        variant = params->kernelVariant - 1;
        if (sc->glossy && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "the fast kernel needs roughness == 0 on every material (RNG skip-ahead)");
        if (area && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "the fast kernel does not carry the area lights' per-chunk jitter index");
        if (sc->motion && variant == KV_FAST)
            return fail(RT_ERR_UNSUPPORTED, "the fast kernel does not carry the ray time of motion blur");
    }
    // the area lights' jitter index runs through a whole 8-row chunk (Object+Extension.swift:288): chunks stay whole
    if (area && cp.shardCount > 1 && cp.shardTiles)
        return fail(RT_ERR_UNSUPPORTED, "scenes with area lights shard by 8-row chunks only (RT_SHARD_ROWS)");
    const int blocks = sc->blocks[variant][counters ? 1 : 0];
    cudaStream_t st = (cudaStream_t)cudaStream;

    const size_t threads = (size_t)blocks * renderBlockSize(variant);
    const size_t need = (size_t)(sc->ds.maxDepth + 1) * kPathFields * threads * sizeof(double);
    if ((rc = ensure(&sc->dPathStack, &sc->pathStackBytes, need))) return rc;

    RenderTargets rt{};
    rt.rgba = (uint8_t*)dRgbaOut; rt.linear = (double*)dLinearOut; rt.ids = (HitId*)dIdsOut;
    rt.pathStack = sc->dPathStack; rt.counters = sc->dCounters;
    rt.tileCounter = sc->dTileCounter;
    rt.motion = sc->motion ? 1 : 0;
    // Tile schedule: the frame records its cost per tile, the next frame of the same geometry draws the longest
    // tiles first.
    const int nTiles = ShardTiles(cp.width, cp.height, cp.shardIndex, cp.shardCount, cp.shardTiles).count;
    const bool adaptive = !(params && params->scheduleMode == RT_SCHEDULE_SPATIAL) && nTiles > 0;
    if (params && (params->scheduleMode < 0 || params->scheduleMode > RT_SCHEDULE_SPATIAL)) return fail(RT_ERR_INVALID, "scheduleMode outside 0..1");
    sc->pendingScheduled = 0;
    if (adaptive) {
        const int key[8] = {cp.width, cp.height, cp.n, cp.dof, cp.shardIndex, cp.shardCount * 2 + cp.shardTiles, variant, nTiles};
        const bool grown = sc->tileCostBytes < (size_t)nTiles * 4 || sc->tileOrderBytes < (size_t)nTiles * 4;
        if ((rc = ensure(&sc->dTileCost, &sc->tileCostBytes, (size_t)nTiles * 4))) return rc;
        if ((rc = ensure(&sc->dTileOrder, &sc->tileOrderBytes, (size_t)nTiles * 4))) return rc;
        if ((rc = ensure(&sc->dSchedWork, &sc->schedWorkBytes, scheduleWorkBytes(nTiles)))) return rc;
        if (grown || std::memcmp(key, sc->schedKey, sizeof key) != 0) sc->schedValid = false;
        std::memcpy(sc->schedKey, key, sizeof key);
        if (!sc->schedValid) {
            if (sc->sideBusy) CK(cudaStreamWaitEvent(st, sc->evSide, 0));
            CK(cudaMemsetAsync(sc->dTileCost, 0, (size_t)nTiles * 4, st));
        }
        rt.tileCost = sc->dTileCost;
        rt.tileOrder = sc->schedValid ? sc->dTileOrder : nullptr;
        sc->pendingScheduled = sc->schedValid ? nTiles : 0;
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
    if (area) {
        // pass 1 follows every path and counts its jitter-table draws per pixel; the prefix kernel turns the counts
        // into each pixel's first index within its chunk; pass 2 (below) is the frame
        if ((rc = ensure(&sc->dAreaIndex, &sc->areaIndexBytes, (size_t)cp.width * cp.height * sizeof(unsigned)))) return rc;
        rt.areaIndex = sc->dAreaIndex;
        rt.areaPass = 1;
        CK(launchRender(variant, sc->ds, cp, rt, blocks, counters, st));
        CK(launchAreaPrefix(sc->dAreaIndex, cp.width, cp.height, cp.shardIndex, cp.shardCount, st));
        CK(cudaMemsetAsync(sc->dTileCounter, 0, sizeof(int), st));
        rt.areaPass = 2;
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
    sc->launches = (adaptive ? 3 : 1) + (area ? 2 : 0);
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

int32_t rt_render(rt_scene_t* sc, const rt_camera* camera, const rt_render_params* params, uint8_t* rgbaOut,
                  double* linearOut, rt_hit_id* idsOut, rt_stats* statsOut) {
    if (!sc) return fail(RT_ERR_NO_SCENE, "No scene loaded. Can't render.");
    if (!camera) return fail(RT_ERR_BAD_CAMERA, "Invalid camera index");
    if (!rgbaOut && !linearOut) return fail(RT_ERR_INVALID, "one of rgbaOut / linearOut is required");
    const double w0 = nowMs();
    CK(cudaSetDevice(sc->device));
    const size_t W = camera->width > 1 ? camera->width : 1, H = camera->height > 1 ? camera->height : 1;
    const int samples = camera->numSamples > 1 ? camera->numSamples : 1;
    const size_t n = (size_t)std::sqrt((double)samples);
    int rc;
    if (rgbaOut && (rc = ensure(&sc->dRgba, &sc->rgbaBytes, W * H * 4))) return rc;
    if (linearOut && (rc = ensure(&sc->dLinear, &sc->linearBytes, W * H * 3 * sizeof(double)))) return rc;
    if (idsOut && (rc = ensure(&sc->dIds, &sc->idsBytes, W * H * n * n * sizeof(HitId)))) return rc;
    // rows outside this rank's shard are defined as zero in the host buffers
    const bool sharded = params && params->shardCount > 1;
    if (sharded) {
        if (rgbaOut) CK(cudaMemsetAsync(sc->dRgba, 0, W * H * 4, 0));
        if (linearOut) CK(cudaMemsetAsync(sc->dLinear, 0, W * H * 3 * sizeof(double), 0));
        if (idsOut) CK(cudaMemsetAsync(sc->dIds, 0, W * H * n * n * sizeof(HitId), 0));
    }
    if (params && params->progress && params->progress(0, (int32_t)H, params->progressUser)) return fail(RT_ERR_CANCELLED, "cancelled");
    rc = rt_render_device(sc, camera, params, rgbaOut ? sc->dRgba : nullptr, linearOut ? sc->dLinear : nullptr,
                          idsOut ? sc->dIds : nullptr, nullptr);
    if (rc) return rc;
    cudaError_t ce = cudaSuccess;
    if (rgbaOut) ce = cudaMemcpyAsync(rgbaOut, sc->dRgba, W * H * 4, cudaMemcpyDeviceToHost, 0);
    if (linearOut && ce == cudaSuccess) ce = cudaMemcpyAsync(linearOut, sc->dLinear, W * H * 3 * sizeof(double), cudaMemcpyDeviceToHost, 0);
    if (idsOut && ce == cudaSuccess) ce = cudaMemcpyAsync(idsOut, sc->dIds, W * H * n * n * sizeof(HitId), cudaMemcpyDeviceToHost, 0);
    rc = rt_render_finish(sc, statsOut);        // also on a failed copy: the frame must not stay "in flight" on the handle
    CK(ce);
    if (rc) return rc;
    CK(cudaStreamSynchronize(0));
    if (statsOut) statsOut->milliseconds = (int64_t)(nowMs() - w0);
    if (params && params->progress) params->progress((int32_t)H, (int32_t)H, params->progressUser);
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
