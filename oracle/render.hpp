// oracle/render.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// Restates the reference's render loop:
//   Renderer.render + trace closure   Extensions/Object+Extension.swift:52-379
//   makeCameraBasis                   :382-427
//   occluded                          :429-433
//   reflect / beerAttenuate / fresnelDielectric / fresnelConductorRGB   :467-505
//   orthonormalBasis                  :530-552
//   PCG32                             :556-589
//   buildStratifiedJitter             :646-659
//   RGBA8 conversion                  RayTracer.swift:186-195
// The Swift TaskGroup over 8-row chunks (:75-82,285-361) is a std::thread pool
// pulling the same chunks from an atomic counter.
#pragma once
#include <atomic>
#include <thread>

#include "traverse.hpp"

namespace orc {

// Object+Extension.swift:556-589
struct PCG32 {
    uint64_t state, inc;
    explicit PCG32(uint64_t seed) {
        state = 0;
        inc = (seed << 1) | 1;
        next();
        state += 0x9E3779B97F4A7C15ULL;
        next();
    }
    uint32_t next() {
        uint64_t old = state;
        state = old * 6364136223846793005ULL + inc;
        uint32_t xorshifted = (uint32_t)(((old >> 18) ^ old) >> 27);
        uint32_t rot = (uint32_t)(old >> 59);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31u));
    }
    double nextFloat() { return (double)next() * 2.3283064365386963e-10; }
};

// Object+Extension.swift:646-659
inline void buildStratifiedJitter(PCG32& rng, double xs[100], double ys[100]) {
    int k = 0;
    for (int gy = 0; gy < 10; gy++)
        for (int gx = 0; gx < 10; gx++) {
            xs[k] = (double)gx + rng.nextFloat();
            ys[k] = (double)gy + rng.nextFloat();
            k++;
        }
}

inline Vec3 reflect(Vec3 d, Vec3 n) { return d - 2.0 * dot(d, n) * n; }  // :467

// :530-552
inline void orthonormalBasis(Vec3 n, Vec3& tangent, Vec3& bitangent) {
    double sign = n.z >= 0 ? 1.0 : -1.0;
    double a = -1.0 / (sign + n.z);
    double b = n.x * n.y * a;
    tangent = normalize(Vec3(1.0 + sign * n.x * n.x * a, sign * b, -sign * n.x));
    bitangent = normalize(Vec3(b, sign + n.y * n.y * a, -n.y));
}

// :476-490.  hasCosT == false stands for cosT == nil.
inline void fresnelDielectric(double n1, double n2, double cosI, double& R, bool& hasCosT, double& cosT, double& sin2T) {
    double cosTheta = dmax(0.0, dmin(1.0, std::fabs(cosI)));
    double eta = n1 / n2;
    sin2T = eta * eta * dmax(0.0, 1.0 - cosTheta * cosTheta);
    if (sin2T > 1.0) { R = 1.0; hasCosT = false; cosT = 0; return; }
    double cosPhi = std::sqrt(dmax(0.0, 1.0 - sin2T));
    double Rs = (n2 * cosTheta - n1 * cosPhi) / (n2 * cosTheta + n1 * cosPhi);
    double Rp = (n1 * cosTheta - n2 * cosPhi) / (n1 * cosTheta + n2 * cosPhi);
    R = 0.5 * (Rs * Rs + Rp * Rp);
    hasCosT = true; cosT = cosPhi;
}

// :492-505 (all three channels equal: eta and k are scalars)
inline Vec3 fresnelConductorRGB(double eta, double k, double cosI_) {
    double cosI = dmax(0.0, dmin(1.0, std::fabs(cosI_)));
    double cos2 = cosI * cosI;
    double eta2 = eta * eta, k2 = k * k;
    double eta2k2 = eta2 + k2;
    double twoEtaCos = 2.0 * eta * cosI;
    double Rs = (eta2k2 - twoEtaCos + cos2) / (eta2k2 + twoEtaCos + cos2);
    double Rp = ((eta2k2 * cos2) - twoEtaCos + 1.0) / ((eta2k2 * cos2) + twoEtaCos + 1.0);
    double r = 0.5 * (Rs + Rp);
    return Vec3(r, r, r);
}

struct CameraBasis {
    Vec3 eye, u, v, w;
    double l, r, b, t, nd;
};

// :382-427
inline CameraBasis makeCameraBasis(const pk::Camera& cam, double aspect) {
    CameraBasis cb;
    cb.eye = toV(cam.position);
    cb.nd = cam.nearDistance;
    Vec3 gazeNorm;
    if (cam.isLookAt) {
        gazeNorm = normalize(toV(cam.gazePoint) - toV(cam.position));
        cb.w = -gazeNorm;
        Vec3 upNorm = normalize(toV(cam.up));
        cb.u = normalize(cross(upNorm, cb.w));
        cb.v = normalize(cross(cb.w, cb.u));
        if (cam.hasFovy) {
            double fovYRad = (cam.fovy * M_PI) / (2.0 * 180.0);
            cb.t = cb.nd * std::tan(fovYRad);
        } else {
            cb.t = cb.nd * 0.5;
        }
        cb.b = -cb.t;
        cb.r = cb.t * aspect;
        cb.l = -cb.r;
    } else {
        cb.l = cam.nearPlane[0]; cb.r = cam.nearPlane[1]; cb.b = cam.nearPlane[2]; cb.t = cam.nearPlane[3];
        gazeNorm = normalize(toV(cam.gaze));
        cb.w = -gazeNorm;
        Vec3 upNorm = normalize(toV(cam.up));
        cb.u = normalize(cross(upNorm, cb.w));
        cb.v = normalize(cross(cb.w, cb.u));
    }
    return cb;
}

struct PrimaryHit {  // oracle extra: identity of each primary ray's closest hit
    int32_t inst;    // instance index, -1 = miss
    int32_t kind;    // HitKind
    int64_t prim;    // primitiveIndex within its kind's global array
    double t;
};

struct RenderOptions {
    int threads = 0;         // 0 = hardware_concurrency
    int chunkStride = 1;     // render only chunks with index % stride == offset (bounded CPU baseline)
    int chunkOffset = 0;
};

struct Renderer {
    const RTContext& ctx;
    const pk::Scene& scene;
    double jitterX[100], jitterY[100];
    Vec3 background, ambient;
    int maxDepth;
    double eps, shadowEps;

    Renderer(const RTContext& c, const pk::Scene& s) : ctx(c), scene(s) {
        PCG32 rng(0x123456789ABCDEFULL);                                   // :92
        buildStratifiedJitter(rng, jitterX, jitterY);
        background = toV(s.backgroundColor);
        ambient = toV(s.ambient);
        maxDepth = s.maxRecursionDepth;                                   // :88
        eps = c.intersectionTestEpsilon;                                  // :87
        shadowEps = c.shadowRayEpsilon;                                   // :89
    }

    bool occluded(const Ray& r, Counters& C) const {                      // :429-433
        if (!ctx.hasTlas) return false;
        C.raysShadow++;
        return occludedTLAS(ctx, r, eps, C);
    }

    Vec3 trace(Ray& inRay, int depth, PCG32& rng, int64_t& jitterIndex, Counters& C) const;
    void renderChunk(const pk::Camera& cam, const CameraBasis& cb, int width, int startRow, int endRow, Vec3* out,
                     PrimaryHit* ids, Counters& C) const;
    // out: width*height Vec3 (0..255 scale); ids: width*height*n*n or null.
    void render(int cameraIndex, Vec3* out, PrimaryHit* ids, Counters& total, const RenderOptions& opt) const;
};

}  // namespace orc
