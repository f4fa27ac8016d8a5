// oracle/render_impl.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
// trace closure (Object+Extension.swift:96-283) and the pixel loops (:285-361).
#pragma once
#include "render.hpp"

namespace orc {

inline Vec3 Renderer::trace(Ray& inRay, int depth, PCG32& rng, int64_t& jitterIndex, Counters& C) const {
    if (depth > maxDepth) return background;                              // :97
    if (!ctx.hasTlas) return Vec3();                                      // :98
    intersectTLAS(ctx, inRay, eps, C);                                    // :99
    if (inRay.hit.kind == HIT_NONE) return background;                    // :101-103

    int64_t nMat = (int64_t)ctx.materials.size();
    int64_t matIndex = std::max<int64_t>(0, std::min<int64_t>(nMat - 1, inRay.hit.mat - 1));  // :105
    const Material& mat = ctx.materials[matIndex];
    const Vec3 p = inRay.hit.p;
    const Vec3 Ngeo = inRay.hit.n;
    const bool frontFacing = dot(inRay.dir, Ngeo) < 0;                    // :109
    const Vec3 N = frontFacing ? Ngeo : -Ngeo;

    const bool hasRefraction = mat.ior > 0;                               // :112-113
    const bool computeDirectLight = !hasRefraction || frontFacing;        // :115
    Vec3 Lo = computeDirectLight ? ambient * mat.ambient : Vec3();        // :116
    if (computeDirectLight) {
        for (const PointLight& pl : ctx.pointLights) {                    // :118-143
            Vec3 wi = pl.position - p;
            double dist = length(wi);
            wi = normalize(wi);
            Ray sRay(p + wi * shadowEps, wi, inRay.time);
            sRay.tMax = dist;
            bool blocked = occluded(sRay, C);
            if (!blocked) {
                double NdotL = dmax(0.0, dot(N, wi));
                if (NdotL > 0) {
                    double shininess = dmax(1.0, mat.phong);
                    Vec3 Ld = mat.diffuse * NdotL;
                    Vec3 view = normalize(-inRay.dir);
                    Vec3 h = normalize(wi + view);
                    double NdotH = dmax(0.0, dot(N, h));
                    Vec3 Ls = mat.specular * std::pow(NdotH, shininess);
                    Vec3 atten = pl.intensity / dmax(dist * dist, 1e-12);
                    Lo += (Ld + Ls) * atten;
                }
            }
        }
        for (const AreaLight& al : ctx.areaLights) {                      // :145-186
            Vec3 nL = normalize(al.normal);
            Vec3 t, b;
            orthonormalBasis(nL, t, b);
            double size = al.size;
            double area = size * size;
            double r1 = jitterX[jitterIndex % 100] / 10.0 - 0.5;
            double r2 = jitterY[jitterIndex % 100] / 10.0 - 0.5;
            jitterIndex += 1;
            Vec3 samplePos = al.position + t * (r1 * size) + b * (r2 * size);
            Vec3 wi = samplePos - p;
            double dist2 = length_squared(wi);
            double dist = std::sqrt(dist2);
            wi = wi / dist;
            double NdotL = dot(N, wi);
            if (NdotL <= 0) continue;
            double Ln = std::fabs(dot(nL, -wi));
            if (Ln <= 0) continue;
            Ray sRay(p + wi * shadowEps, wi, inRay.time);
            sRay.tMax = dist - shadowEps;
            if (occluded(sRay, C)) continue;
            Vec3 view = normalize(-inRay.dir);
            Vec3 h = normalize(wi + view);
            Vec3 Ld = mat.diffuse * NdotL;
            Vec3 Ls = mat.specular * std::pow(dmax(0.0, dot(N, h)), mat.phong);
            Vec3 brdf = Ld + Ls;
            Vec3 contrib = brdf * al.radiance * (Ln / dist2) * area;
            Lo += contrib;
        }
    }

    if (mat.type == pk::MAT_MIRROR && depth < maxDepth) {                 // :189-206
        Vec3 rd = normalize(reflect(inRay.dir, N));
        if (mat.roughness != 0) {
            Vec3 t, b;
            orthonormalBasis(rd, t, b);
            double rand1 = rng.nextFloat() - 0.5;
            double rand2 = rng.nextFloat() - 0.5;
            rd = rd + (mat.roughness * rand1) * b + (mat.roughness * rand2) * t;
            rd = normalize(rd);
        }
        Ray rRay = inRay;
        rRay.tMin = 0;
        rRay.origin = p + N * shadowEps;
        rRay.dir = rd;
        rRay.invDir = 1.0 / rd;
        rRay.hit = Hit();
        C.raysSecondary++;
        Vec3 Li = trace(rRay, depth + 1, rng, jitterIndex, C);
        Lo += mat.mirror * Li;
    } else if (mat.type == pk::MAT_DIELECTRIC && depth < maxDepth) {      // :207-251
        Vec3 microfacetN = N;
        if (mat.roughness != 0) {
            Vec3 t, b;
            orthonormalBasis(N, t, b);
            double r1 = (rng.nextFloat() * 2.0) - 1.0;
            double r2 = (rng.nextFloat() * 2.0) - 1.0;
            microfacetN = normalize(N + (t * r1 * mat.roughness) + (b * r2 * mat.roughness));
        }
        const bool entering = frontFacing;
        double n1 = entering ? 1.0 : mat.ior;
        double n2 = entering ? mat.ior : 1.0;
        double cosI = -dot(inRay.dir, microfacetN);
        double R, cosT, sin2T;
        bool hasCosT;
        fresnelDielectric(n1, n2, cosI, R, hasCosT, cosT, sin2T);
        Vec3 rd = normalize(reflect(inRay.dir, microfacetN));
        Ray rRay = inRay;
        rRay.tMin = 0;
        rRay.origin = p + rd * shadowEps;
        rRay.dir = rd;
        rRay.invDir = 1.0 / rd;
        rRay.hit = Hit();
        C.raysSecondary++;
        Vec3 LiR = trace(rRay, depth + 1, rng, jitterIndex, C);
        if (!hasCosT || sin2T > 1) {                                      // :235-236
            Lo += LiR;
        } else {
            double eta = n1 / n2;
            Vec3 td = normalize((inRay.dir * eta) + (microfacetN * (eta * cosI - cosT)));
            Ray tRay(p + td * shadowEps, td, inRay.time);
            C.raysSecondary++;
            Vec3 LiT = trace(tRay, depth + 1, rng, jitterIndex, C);
            if (entering && mat.absorption != Vec3() && tRay.hit.kind != HIT_NONE && tRay.hit.mat == inRay.hit.mat) {
                double d = dmax(tRay.hit.t, 0.0);                         // :246
                if (std::isfinite(d) && d > 0) {                          // beerAttenuate :470-474
                    Vec3 att(std::exp(-mat.absorption.x * d), std::exp(-mat.absorption.y * d),
                             std::exp(-mat.absorption.z * d));
                    LiT *= att;
                } else {
                    LiT *= LiT;                                           // guard path returns Li -> LiT *= LiT
                }
            }
            Lo += LiR * R + LiT * (1.0 - R);
        }
    } else if (mat.type == pk::MAT_CONDUCTOR && depth < maxDepth) {       // :252-275
        double cosI = dmax(0.0, -dot(inRay.dir, N));
        Vec3 Rf = fresnelConductorRGB(mat.ior, mat.absorptionIndex, cosI);
        Vec3 rd = normalize(reflect(inRay.dir, N));
        if (mat.roughness != 0) {
            Vec3 t, b;
            orthonormalBasis(rd, t, b);
            double rand1 = rng.nextFloat() - 0.5;
            double rand2 = rng.nextFloat() - 0.5;
            rd = rd + (mat.roughness * rand1) * b + (mat.roughness * rand2) * t;
            rd = normalize(rd);
        }
        Ray rRay = inRay;
        rRay.tMin = 0;
        rRay.origin = p + N * shadowEps;
        rRay.dir = rd;
        rRay.invDir = 1.0 / rd;
        rRay.hit = Hit();
        C.raysSecondary++;
        Vec3 Li = trace(rRay, depth + 1, rng, jitterIndex, C);
        Lo += (Rf * mat.mirror) * Li;
    }

    if (!(std::isfinite(Lo.x) && std::isfinite(Lo.y) && std::isfinite(Lo.z))) return Vec3(0, 0, 0);  // :277-280
    return Lo;
}

// one 8-row chunk: Object+Extension.swift:287-360
inline void Renderer::renderChunk(const pk::Camera& cam, const CameraBasis& cb, int width, int startRow, int endRow,
                                  Vec3* out, PrimaryHit* ids, Counters& C) const {
    const int height = std::max(1, cam.height);
    const double du = (cb.r - cb.l) / (double)width;                     // :67
    const double dv = (cb.t - cb.b) / (double)height;                    // :68
    const Vec3 eye = cb.eye, u = cb.u, v = cb.v, w = cb.w;
    const Vec3 m = eye - w * cb.nd;                                       // :69
    const Vec3 q00 = m + u * cb.l + v * cb.t;                             // :70
    const int samples = std::max(1, cam.numSamples);                      // :298
    const int n = (int)std::sqrt((double)samples);                        // :300
    int64_t jitterIndex = 0;                                              // :288
    for (int j = startRow; j < endRow; j++) {
        for (int i = 0; i < width; i++) {
            PCG32 rng((((uint64_t)j << 32) ^ (uint64_t)i) + 0x9E3779B97F4A7C15ULL);  // :294 (Swift precedence)
            Vec3 pixelColor;
            int sampleIndex = 0;
            for (int sy = 0; sy < n; sy++) {
                for (int sx = 0; sx < n; sx++) {
                    double xi1 = rng.nextFloat();
                    double xi2 = rng.nextFloat();
                    double iOffset = ((double)sx + xi1) / (double)n;
                    double jOffset = ((double)sy + xi2) / (double)n;
                    double currentI = (double)i + iOffset;
                    double currentJ = (double)j + jOffset;
                    Vec3 vOff = v * (currentJ * dv);
                    Vec3 rowTopLeft = q00 - vOff;
                    Vec3 uOff = u * (currentI * du);
                    Vec3 s = rowTopLeft + uOff;
                    Vec3 dir0 = normalize(s - eye);
                    Vec3 dir = dir0;
                    Vec3 camEye = eye;
                    if (cam.apertureSize > 0 && cam.focusDistance > 0) {  // :325-338
                        Vec3 forward = -w;
                        double denom = dot(dir0, forward);
                        double tFocus = std::fabs(denom) < 1e-6 ? cam.focusDistance : (cam.focusDistance / denom);
                        Vec3 pFocus = eye + dir0 * tFocus;
                        double uRand = rng.nextFloat() - 0.5;
                        double vRand = rng.nextFloat() - 0.5;
                        Vec3 lensOffset = (uRand * u + vRand * v) * cam.apertureSize;
                        Vec3 a = eye + lensOffset;
                        dir = normalize(pFocus - a);
                        camEye = a;
                    }
                    Ray ray(camEye, dir, rng.nextFloat());                // :340
                    double denom = dot(dir, w);
                    double tImg = dot((eye - w * cb.nd) - camEye, w) /
                                  (denom == 0 ? std::numeric_limits<double>::denorm_min() : denom);  // :342
                    ray.tMin = dmax(tImg, 0.0);
                    C.raysPrimary++;
                    Vec3 sampleColor = trace(ray, 0, rng, jitterIndex, C);
                    if (ids) {
                        PrimaryHit& h = ids[((size_t)j * width + i) * (size_t)(n * n) + (size_t)(sy * n + sx)];
                        h.inst = ray.hit.kind == HIT_NONE ? -1 : (int32_t)ray.hit.inst;
                        h.kind = ray.hit.kind;
                        h.prim = ray.hit.kind == HIT_NONE ? -1 : ray.hit.prim;
                        h.t = ray.hit.t;
                    }
                    pixelColor += sampleColor;
                    sampleIndex += 1;
                    if (sampleIndex >= samples) break;
                }
                if (sampleIndex >= samples) break;
            }
            out[(size_t)j * width + i] = pixelColor / (double)samples;    // :354
        }
    }
}

inline void Renderer::render(int cameraIndex, Vec3* out, PrimaryHit* ids, Counters& total,
                             const RenderOptions& opt) const {
    const pk::Camera& cam = scene.cameras[cameraIndex];
    const int width = std::max(1, cam.width), height = std::max(1, cam.height);  // :62-63
    const CameraBasis cb = makeCameraBasis(cam, (double)width / (double)height);  // :65
    const int CH = 8;                                                     // :75
    const int nChunks = (height + CH - 1) / CH;
    int T = opt.threads > 0 ? opt.threads : (int)std::thread::hardware_concurrency();
    if (T < 1) T = 1;
    std::atomic<int> next{0};
    std::vector<Counters> perThread(T);
    auto worker = [&](int tid) {
        for (;;) {
            int c = next.fetch_add(1);
            if (c >= nChunks) break;
            if (opt.chunkStride > 1 && (c % opt.chunkStride) != opt.chunkOffset) continue;
            int startRow = c * CH, endRow = std::min(startRow + CH, height);
            renderChunk(cam, cb, width, startRow, endRow, out, ids, perThread[tid]);
        }
    };
    if (T == 1) {
        worker(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back(worker, t);
        for (auto& t : th) t.join();
    }
    for (const Counters& c : perThread) total.add(c);
}

// RayTracer.swift:186-195: clamp to [0,255], truncate, alpha 255.
inline void toRGBA8(const Vec3* linear, size_t count, uint8_t* rgba) {
    for (size_t i = 0; i < count; i++) {
        auto q = [](double c) -> uint8_t {
            double x = dmin(dmax(c, 0.0), 255.0);  // simd_clamp = min(max(x, lo), hi)
            return (uint8_t)x;
        };
        rgba[4 * i + 0] = q(linear[i].x);
        rgba[4 * i + 1] = q(linear[i].y);
        rgba[4 * i + 2] = q(linear[i].z);
        rgba[4 * i + 3] = 255;
    }
}

}  // namespace orc
