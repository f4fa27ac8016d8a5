// rt_kernels.cu — the render megakernel for sm_100a.   COMPILE WITH -fmad=false.
//
// One persistent kernel replaces the reference's whole per-frame loop:
//   Renderer.render pixel/sample loops     Extensions/Object+Extension.swift:285-361
//   primary-ray generation                 :292-344      (PCG32 :556-589)
//   trace closure (shading + recursion)    :96-283
//   Fresnel / reflect / Beer helpers       :467-505, orthonormalBasis :530-552
//   sample average                         :354
//   RGBA8 conversion                       RayTracer.swift:186-195
// Work distribution: warps pull 8x4-pixel tiles of the rank's 8-row chunks
// (the reference's chunk unit, :75-82) from a global atomic counter (persistent
// threads); one lane = one pixel, looping over that pixel's samples so the
// per-pixel PCG32 stream and FP64 summation order match the reference.
// The trace recursion is an explicit per-thread stack of (ray, throughput)
// entries processed depth-first, reflection before refraction, which keeps the
// reference's RNG draw order.
#include <cuda_runtime.h>
#include <math_constants.h>

#include "rt_kernels.h"
#include "rt_trace.cuh"

namespace rtg {

#define PATH_FIELDS 11

struct PathEntry {
    V3 o, d, w;
    long long beerMat;   // Hit.mat of the dielectric surface this refraction ray entered
    int depth;
    int beer;            // 1: apply Beer attenuation at the next hit (Object+Extension.swift:245-248)
};

__device__ __forceinline__ void pathStore(double* base, size_t stride, size_t gtid, int level, const PathEntry& e) {
    double* p = base + ((size_t)level * PATH_FIELDS) * stride + gtid;
    p[0 * stride] = e.o.x; p[1 * stride] = e.o.y; p[2 * stride] = e.o.z;
    p[3 * stride] = e.d.x; p[4 * stride] = e.d.y; p[5 * stride] = e.d.z;
    p[6 * stride] = e.w.x; p[7 * stride] = e.w.y; p[8 * stride] = e.w.z;
    p[9 * stride] = __longlong_as_double(e.beerMat);
    p[10 * stride] = __longlong_as_double((long long)e.depth | ((long long)e.beer << 32));
}
__device__ __forceinline__ void pathLoad(const double* base, size_t stride, size_t gtid, int level, PathEntry& e) {
    const double* p = base + ((size_t)level * PATH_FIELDS) * stride + gtid;
    e.o = mk(p[0 * stride], p[1 * stride], p[2 * stride]);
    e.d = mk(p[3 * stride], p[4 * stride], p[5 * stride]);
    e.w = mk(p[6 * stride], p[7 * stride], p[8 * stride]);
    e.beerMat = __double_as_longlong(p[9 * stride]);
    long long m = __double_as_longlong(p[10 * stride]);
    e.depth = (int)(m & 0xffffffffLL);
    e.beer = (int)(m >> 32);
}

__device__ __forceinline__ V3 reflectDir(V3 d, V3 n) { return d - 2.0 * dot(d, n) * n; }  // :467

// :530-552
__device__ __forceinline__ void orthonormalBasis(V3 n, V3& tangent, V3& bitangent) {
    double sign = n.z >= 0 ? 1.0 : -1.0;
    double a = -1.0 / (sign + n.z);
    double b = n.x * n.y * a;
    tangent = normalize(mk(1.0 + sign * n.x * n.x * a, sign * b, -sign * n.x));
    bitangent = normalize(mk(b, sign + n.y * n.y * a, -n.y));
}

// :476-490
__device__ __forceinline__ void fresnelDielectric(double n1, double n2, double cosI, double& R, bool& hasCosT, double& cosT,
                                                  double& sin2T) {
    double cosTheta = dmax(0.0, dmin(1.0, fabs(cosI)));
    double eta = n1 / n2;
    sin2T = eta * eta * dmax(0.0, 1.0 - cosTheta * cosTheta);
    if (sin2T > 1.0) { R = 1.0; hasCosT = false; cosT = 0; return; }
    double cosPhi = sqrt(dmax(0.0, 1.0 - sin2T));
    double Rs = (n2 * cosTheta - n1 * cosPhi) / (n2 * cosTheta + n1 * cosPhi);
    double Rp = (n1 * cosTheta - n2 * cosPhi) / (n1 * cosTheta + n2 * cosPhi);
    R = 0.5 * (Rs * Rs + Rp * Rp);
    hasCosT = true; cosT = cosPhi;
}

// :492-505
__device__ __forceinline__ double fresnelConductor(double eta, double k, double cosI_) {
    double cosI = dmax(0.0, dmin(1.0, fabs(cosI_)));
    double cos2 = cosI * cosI;
    double eta2k2 = eta * eta + k * k;
    double twoEtaCos = 2.0 * eta * cosI;
    double Rs = (eta2k2 - twoEtaCos + cos2) / (eta2k2 + twoEtaCos + cos2);
    double Rp = ((eta2k2 * cos2) - twoEtaCos + 1.0) / ((eta2k2 * cos2) + twoEtaCos + 1.0);
    return 0.5 * (Rs + Rp);
}

__device__ __forceinline__ unsigned char quantize(double c) {  // RayTracer.swift:189-193
    double x = dmin(dmax(c, 0.0), 255.0);
    return (unsigned char)(int)x;
}

template <bool COUNT>
__global__ void __launch_bounds__(RT_BLOCK) renderKernel(DeviceScene S, CameraParams cam, RenderTargets out) {
    __shared__ int32_t sstack[RT_STACK * RT_BLOCK];
    const int lane = threadIdx.x & 31;
    const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;

    Counters C;
    C.box = C.tri = C.sphere = C.plane = C.inst = C.smooth = 0; C.maxStack = 0;
    unsigned long long nPrimary = 0, nShadow = 0, nSecondary = 0;

    const int W = cam.width, H = cam.height;
    const int tilesX = (W + 7) >> 3;
    const int nChunks = (H + 7) >> 3;
    const int shardCount = cam.shardCount < 1 ? 1 : cam.shardCount;
    const int owned = (nChunks - cam.shardIndex + shardCount - 1) / shardCount;
    const int totalTiles = owned * tilesX * 2;
    const V3 eye = ld3(cam.eye), cu = ld3(cam.u), cv = ld3(cam.v), cw = ld3(cam.w), q00 = ld3(cam.q00), cm = ld3(cam.m);
    const int n = cam.n, samples = cam.samples;
    const V3 ambient = ld3(S.ambient), background = ld3(S.background);
    const double shadowEps = S.shadowEps;

    for (;;) {
        int tile = 0;
        if (lane == 0) tile = atomicAdd(out.tileCounter, 1);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= totalTiles) break;
        const int k = tile / (tilesX * 2), rem = tile - k * (tilesX * 2);
        const int ty = rem / tilesX, tx = rem - ty * tilesX;
        const int i = tx * 8 + (lane & 7);
        const int j = (k * shardCount + cam.shardIndex) * 8 + ty * 4 + (lane >> 3);
        if (i >= W || j >= H) continue;

        PCG32 rng;
        rng.seed((((uint64_t)j << 32) ^ (uint64_t)i) + 0x9E3779B97F4A7C15ULL);  // :294 (Swift precedence)
        V3 pixel = mk(0, 0, 0);
        int sampleIndex = 0;
        for (int sy = 0; sy < n && sampleIndex < samples; sy++) {
            for (int sx = 0; sx < n && sampleIndex < samples; sx++) {
                // ---- primary ray, :307-344
                const double xi1 = rng.nextFloat();
                const double xi2 = rng.nextFloat();
                const double iOffset = ((double)sx + xi1) / (double)n;
                const double jOffset = ((double)sy + xi2) / (double)n;
                const double currentI = (double)i + iOffset;
                const double currentJ = (double)j + jOffset;
                const V3 vOff = cv * (currentJ * cam.dv);
                const V3 rowTopLeft = q00 - vOff;
                const V3 uOff = cu * (currentI * cam.du);
                const V3 s = rowTopLeft + uOff;
                const V3 dir0 = normalize(s - eye);
                V3 dir = dir0, camEye = eye;
                if (cam.dof) {  // :325-338
                    const V3 forward = -cw;
                    const double denom = dot(dir0, forward);
                    const double tFocus = fabs(denom) < 1e-6 ? cam.focus : (cam.focus / denom);
                    const V3 pFocus = eye + dir0 * tFocus;
                    const double uRand = rng.nextFloat() - 0.5;
                    const double vRand = rng.nextFloat() - 0.5;
                    const V3 lensOffset = (uRand * cu + vRand * cv) * cam.aperture;
                    const V3 a = eye + lensOffset;
                    dir = normalize(pFocus - a);
                    camEye = a;
                }
                (void)rng.nextFloat();  // ray time: always drawn (:340); no motion in supported scenes
                const double denomW = dot(dir, cw);
                const double tImg = dot(cm - camEye, cw) / (denomW == 0.0 ? 4.9406564584124654e-324 : denomW);  // :342
                const double tMinPrimary = dmax(tImg, 0.0);
                nPrimary++;

                // ---- trace: explicit depth-first stack (reflection before refraction)
                V3 sampleColor = mk(0, 0, 0);
                PathEntry cur;
                cur.o = camEye; cur.d = dir; cur.w = mk(1, 1, 1); cur.beerMat = 0; cur.depth = 0; cur.beer = 0;
                int level = 0;
                bool havePath = true;
                bool primary = true;
                while (havePath) {
                    HitRec hit;
                    hit.t = CUDART_INF; hit.u = hit.v = 0; hit.inst = -1; hit.slot = 0;
                    traceScene<false, COUNT>(S, cur.o, cur.d, primary ? tMinPrimary : 0.0, hit, sstack, C);
                    Surface sf;
                    if (hit.inst >= 0) resolveHit(S, hit, cur.o, cur.d, sf);
                    if (COUNT && hit.inst >= 0 && sf.kind == 1 && S.instances[hit.inst].smooth) C.smooth++;
                    if (primary && out.ids) {
                        HitId& h = out.ids[((size_t)j * W + i) * (size_t)(n * n) + (size_t)(sy * n + sx)];
                        h.inst = hit.inst;
                        h.kind = hit.inst >= 0 ? sf.kind : 0;
                        h.prim = hit.inst >= 0 ? sf.prim : -1;
                        h.t = hit.t;
                    }
                    primary = false;
                    int spawned = 0;
                    if (hit.inst < 0) {
                        sampleColor = sampleColor + cur.w * background;            // :101-103
                    } else {
                        long long mi = sf.mat - 1;                                  // :105
                        if (mi > (long long)S.nMaterials - 1) mi = (long long)S.nMaterials - 1;
                        if (mi < 0) mi = 0;
                        const GMaterial& mat = S.materials[mi];
                        V3 w = cur.w;
                        if (cur.beer && sf.mat == cur.beerMat) {                    // :245-248
                            const double dist = dmax(hit.t, 0.0);
                            if (isfinite(dist) && dist > 0)
                                w = w * mk(exp(-mat.absorption[0] * dist), exp(-mat.absorption[1] * dist), exp(-mat.absorption[2] * dist));
                        }
                        const V3 p = sf.p, Ngeo = sf.n;
                        const bool frontFacing = dot(cur.d, Ngeo) < 0;               // :109
                        const V3 N = frontFacing ? Ngeo : -Ngeo;
                        const bool hasRefraction = mat.ior > 0;
                        const bool computeDirect = !hasRefraction || frontFacing;   // :115
                        V3 Lo = computeDirect ? ambient * ld3(mat.ambient) : mk(0, 0, 0);
                        if (computeDirect) {
                            for (int l = 0; l < S.nPointLights; l++) {             // :118-143
                                const GPointLight& pl = S.pointLights[l];
                                V3 wi = ld3(pl.position) - p;
                                const double dist = length(wi);
                                wi = normalize(wi);
                                HitRec sh;
                                sh.t = dist; sh.u = sh.v = 0; sh.inst = -1; sh.slot = 0;   // tMax = dist (:124)
                                nShadow++;
                                const bool blocked = traceScene<true, COUNT>(S, p + wi * shadowEps, wi, 0.0, sh, sstack, C);
                                if (!blocked) {
                                    const double NdotL = dmax(0.0, dot(N, wi));
                                    if (NdotL > 0) {
                                        const double shininess = dmax(1.0, mat.phong);
                                        const V3 Ld = ld3(mat.diffuse) * NdotL;
                                        const V3 view = normalize(-cur.d);
                                        const V3 h = normalize(wi + view);
                                        const double NdotH = dmax(0.0, dot(N, h));
                                        const V3 Ls = ld3(mat.specular) * pow(NdotH, shininess);
                                        const V3 atten = ld3(pl.intensity) / dmax(dist * dist, 1e-12);
                                        Lo = Lo + (Ld + Ls) * atten;
                                    }
                                }
                            }
                        }
                        if (finite3(Lo)) {                                          // NaN guard, :277-280
                            sampleColor = sampleColor + w * Lo;
                            if (cur.depth < S.maxDepth && mat.type != 0) {
                                PathEntry c1, c2;
                                bool has2 = false;
                                c1.depth = c2.depth = cur.depth + 1;
                                c1.beer = c2.beer = 0; c1.beerMat = c2.beerMat = 0;
                                if (mat.type == 1) {                                // mirror :189-206
                                    V3 rd = normalize(reflectDir(cur.d, N));
                                    if (mat.roughness != 0) {
                                        V3 t, b;
                                        orthonormalBasis(rd, t, b);
                                        const double rand1 = rng.nextFloat() - 0.5;
                                        const double rand2 = rng.nextFloat() - 0.5;
                                        rd = rd + (mat.roughness * rand1) * b + (mat.roughness * rand2) * t;
                                        rd = normalize(rd);
                                    }
                                    c1.o = p + N * shadowEps; c1.d = rd; c1.w = w * ld3(mat.mirror);
                                } else if (mat.type == 2) {                         // dielectric :207-251
                                    V3 mN = N;
                                    if (mat.roughness != 0) {
                                        V3 t, b;
                                        orthonormalBasis(N, t, b);
                                        const double r1 = (rng.nextFloat() * 2.0) - 1.0;
                                        const double r2 = (rng.nextFloat() * 2.0) - 1.0;
                                        mN = normalize(N + (t * r1 * mat.roughness) + (b * r2 * mat.roughness));
                                    }
                                    const bool entering = frontFacing;
                                    const double n1 = entering ? 1.0 : mat.ior;
                                    const double n2 = entering ? mat.ior : 1.0;
                                    const double cosI = -dot(cur.d, mN);
                                    double R, cosT, sin2T;
                                    bool hasCosT;
                                    fresnelDielectric(n1, n2, cosI, R, hasCosT, cosT, sin2T);
                                    const V3 rd = normalize(reflectDir(cur.d, mN));
                                    c1.o = p + rd * shadowEps; c1.d = rd;
                                    if (!hasCosT || sin2T > 1) {
                                        c1.w = w;                                   // total internal reflection :235-236
                                    } else {
                                        c1.w = w * R;
                                        const double eta = n1 / n2;
                                        const V3 td = normalize((cur.d * eta) + (mN * (eta * cosI - cosT)));
                                        c2.o = p + td * shadowEps; c2.d = td; c2.w = w * (1.0 - R);
                                        c2.beer = (entering && mat.hasAbsorption) ? 1 : 0;
                                        c2.beerMat = sf.mat;
                                        has2 = true;
                                    }
                                } else {                                            // conductor :252-275
                                    const double cosI = dmax(0.0, -dot(cur.d, N));
                                    const double Rf = fresnelConductor(mat.ior, mat.absorptionIndex, cosI);
                                    V3 rd = normalize(reflectDir(cur.d, N));
                                    if (mat.roughness != 0) {
                                        V3 t, b;
                                        orthonormalBasis(rd, t, b);
                                        const double rand1 = rng.nextFloat() - 0.5;
                                        const double rand2 = rng.nextFloat() - 0.5;
                                        rd = rd + (mat.roughness * rand1) * b + (mat.roughness * rand2) * t;
                                        rd = normalize(rd);
                                    }
                                    c1.o = p + N * shadowEps; c1.d = rd; c1.w = w * (mk(Rf, Rf, Rf) * ld3(mat.mirror));
                                }
                                if (has2) { pathStore(out.pathStack, stride, gtid, level, c2); level++; nSecondary++; }
                                cur = c1;
                                nSecondary++;
                                spawned = 1;
                            }
                        }
                    }
                    if (!spawned) {
                        if (level > 0) { level--; pathLoad(out.pathStack, stride, gtid, level, cur); }
                        else havePath = false;
                    }
                }
                pixel = pixel + sampleColor;
                sampleIndex++;
            }
        }
        const V3 c = pixel / (double)samples;                                       // :354
        const size_t pix = (size_t)j * W + i;
        if (out.linear) { out.linear[pix * 3] = c.x; out.linear[pix * 3 + 1] = c.y; out.linear[pix * 3 + 2] = c.z; }
        if (out.rgba) {
            uchar4 q;
            q.x = quantize(c.x); q.y = quantize(c.y); q.z = quantize(c.z); q.w = 255;
            reinterpret_cast<uchar4*>(out.rgba)[pix] = q;
        }
    }

    // ray counters: warp reduce, one atomic per warp
    for (int off = 16; off > 0; off >>= 1) {
        nPrimary += __shfl_down_sync(0xffffffffu, nPrimary, off);
        nShadow += __shfl_down_sync(0xffffffffu, nShadow, off);
        nSecondary += __shfl_down_sync(0xffffffffu, nSecondary, off);
        if (COUNT) {
            C.box += __shfl_down_sync(0xffffffffu, C.box, off);
            C.tri += __shfl_down_sync(0xffffffffu, C.tri, off);
            C.sphere += __shfl_down_sync(0xffffffffu, C.sphere, off);
            C.plane += __shfl_down_sync(0xffffffffu, C.plane, off);
            C.inst += __shfl_down_sync(0xffffffffu, C.inst, off);
            C.smooth += __shfl_down_sync(0xffffffffu, C.smooth, off);
            int m = __shfl_down_sync(0xffffffffu, C.maxStack, off);
            if (m > C.maxStack) C.maxStack = m;
        }
    }
    if (lane == 0) {
        atomicAdd(&out.counters[0], nPrimary);
        atomicAdd(&out.counters[1], nShadow);
        atomicAdd(&out.counters[2], nSecondary);
        if (COUNT) {
            atomicAdd(&out.counters[3], C.box);
            atomicAdd(&out.counters[4], C.tri);
            atomicAdd(&out.counters[5], C.sphere);
            atomicAdd(&out.counters[6], C.plane);
            atomicAdd(&out.counters[7], C.inst);
            atomicMax(&out.counters[8], (unsigned long long)C.maxStack);
            atomicAdd(&out.counters[9], C.smooth);
        }
    }
}

cudaError_t launchRender(const DeviceScene& S, const CameraParams& cam, const RenderTargets& out, int blocks, bool counters,
                         cudaStream_t stream) {
    if (counters) renderKernel<true><<<blocks, RT_BLOCK, 0, stream>>>(S, cam, out);
    else renderKernel<false><<<blocks, RT_BLOCK, 0, stream>>>(S, cam, out);
    return cudaGetLastError();
}

int renderBlockSize() { return RT_BLOCK; }
int renderStackCapacity() { return RT_STACK; }

cudaError_t renderOccupancy(int* blocksPerSm, bool counters) {
    if (counters) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocksPerSm, renderKernel<true>, RT_BLOCK, 0);
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocksPerSm, renderKernel<false>, RT_BLOCK, 0);
}

}  // namespace rtg
