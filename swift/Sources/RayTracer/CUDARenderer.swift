//
//  CUDARenderer.swift — drop-in body for `Renderer` (Extensions/Object+Extension.swift:13-19, 52-379)
//
//  The binding a maintainer adds to the reference: same class name, same `init(ctx:)`, same
//  `render(scene:cameraIndex:progressRow:) async throws -> [Vec3]`; the per-pixel loop, `trace`, the
//  TLAS/BLAS traversal and the primitive tests run in the CRTCuda module (hand-written CUDA for sm_100a)
//  instead of the Swift TaskGroup.  `renderRGBA8(...)` additionally hands back the RGBA8 frame the kernel
//  already produced, so `renderRGBA8Async` (RayTracer.swift:186-195) can skip its conversion loop.
//
//  NOT COMPILED in this repository: no Swift toolchain exists in the build container or on the GPU box, and
//  the reference itself needs Apple's `simd` and the un-vendored ParsingKit.  The identical C ABI is
//  exercised by the C++ mirror in myraytracer_b200/host/ and by tests/ (see INTEGRATION.md).
//

import Foundation
import simd
import CRTCuda
@preconcurrency import ParsingKit

final class Renderer: @unchecked Sendable {
    public let ctx: RTContext
    private var handle: OpaquePointer?            // rt_scene_t*
    private var blasOfInstance = [Int32]()        // rt_instance.blas of every instance (fixed for the scene's life)

    /// Renderer(ctx:) — flattens RTContext (Models/RTContext.swift:74-92) into rt_scene_desc and uploads it.
    /// `devices`: the GPUs the scene is made resident on.  nil = every GPU of the box (rt_device_count): the BVHs are
    /// built once, the other GPUs get device-to-device copies, and rt_render deals the frame's 8-row chunks to all of
    /// them (rt_scene_create_multi); one entry = that GPU alone (rt_scene_create).
    public init(ctx: RTContext, scene: Scene, devices: [Int32]? = nil) {
        let gpus: [Int32] = devices ?? (0..<max(1, rt_device_count())).map { Int32($0) }
        self.ctx = ctx
        var triV = [Double](); triV.reserveCapacity(ctx.triangles.count * 9)
        var triN = [Double](); triN.reserveCapacity(ctx.triangles.count * 9)
        var triM = [Double](); triM.reserveCapacity(ctx.triangles.count * 3)   // Triangle.motionBlur (RTContext.swift:280,345)
        for t in ctx.triangles {
            for v in [t.v0, t.v1, t.v2] { triV += [v.x, v.y, v.z] }
            for n in [t.n0, t.n1, t.n2] { triN += [n.x, n.y, n.z] }
            triM += [t.motionBlur.x, t.motionBlur.y, t.motionBlur.z]
        }
        var areaLights = ctx.areaLights.map {                                    // Object+Extension.swift:145-186
            rt_area_light(position: ($0.position.x, $0.position.y, $0.position.z), normal: ($0.normal.x, $0.normal.y, $0.normal.z),
                          size: $0.size, radiance: ($0.radiance.x, $0.radiance.y, $0.radiance.z))
        }
        var spheres = ctx.spheres.map { rt_sphere(center: ($0.center.x, $0.center.y, $0.center.z), radius: $0.radius) }
        var planes = ctx.planes.map { rt_plane(center: ($0.center.x, $0.center.y, $0.center.z),
                                               normal: ($0.normal.x, $0.normal.y, $0.normal.z)) }
        // BLAS.prims (RTContext.swift:34-42) become contiguous ranges of one prim array
        var prims = [rt_prim](), blas = [rt_blas](), blasIndex = [ObjectIdentifier: Int32]()
        var instances = [rt_instance]()
        for inst in ctx.instances {
            let key = ObjectIdentifier(inst.blas)
            if blasIndex[key] == nil {
                blasIndex[key] = Int32(blas.count)
                blas.append(rt_blas(firstPrim: Int64(prims.count), primCount: Int64(inst.blas.prims.count)))
                for p in inst.blas.prims {
                    prims.append(rt_prim(type: Int32(p.type.rawValue), smooth: p.shadingMode == .smooth ? 1 : 0,
                                         primitiveIndex: Int64(p.primitiveIndex), materialID: Int64(p.materialID)))
                }
            }
            var ri = rt_instance()
            ri.blas = blasIndex[key]!
            blasOfInstance.append(ri.blas)
            ri.hasMaterialOverride = inst.materialOverride == nil ? 0 : 1
            ri.materialOverride = Int64(inst.materialOverride ?? 0)
            withUnsafeMutableBytes(of: &ri.localToWorld) { $0.copyBytes(from: Self.bytes(inst.localToWorld)) }
            withUnsafeMutableBytes(of: &ri.worldToLocal) { $0.copyBytes(from: Self.bytes(inst.worldToLocal)) }
            withUnsafeMutableBytes(of: &ri.normalMatrix) { $0.copyBytes(from: Self.bytes3(inst.normalMatrix)) }
            ri.worldBounds = (inst.worldBounds.minP.x, inst.worldBounds.minP.y, inst.worldBounds.minP.z,
                              inst.worldBounds.maxP.x, inst.worldBounds.maxP.y, inst.worldBounds.maxP.z)
            ri.instanceMotion = (inst.instanceMotion.x, inst.instanceMotion.y, inst.instanceMotion.z)
            instances.append(ri)
        }
        var materials = scene.materials.map { m -> rt_material in
            var r = rt_material()
            switch m.type { case "mirror": r.type = 1; case "dielectric": r.type = 2; case "conductor": r.type = 3; default: r.type = 0 }
            r.ambient = (m.ambient.x, m.ambient.y, m.ambient.z); r.diffuse = (m.diffuse.x, m.diffuse.y, m.diffuse.z)
            r.specular = (m.specular.x, m.specular.y, m.specular.z); r.mirror = (m.mirror.x, m.mirror.y, m.mirror.z)
            r.absorption = (m.absorption.x, m.absorption.y, m.absorption.z)
            r.phong = m.phong; r.ior = m.ior; r.absorptionIndex = m.absorptionIndex; r.roughness = m.roughness
            return r
        }
        var lights = scene.lights.points.compactMap { $0 as? PointLight }.map {
            rt_point_light(position: ($0.position.x, $0.position.y, $0.position.z),
                           intensity: ($0.intensity.x, $0.intensity.y, $0.intensity.z))
        }
        var desc = rt_scene_desc()
        desc.abiVersion = UInt32(RT_ABI_VERSION)
        desc.nTriangles = Int64(ctx.triangles.count)
        desc.nSpheres = Int64(spheres.count); desc.nPlanes = Int64(planes.count)
        desc.nPrims = Int64(prims.count); desc.nBlas = Int64(blas.count); desc.nInstances = Int64(instances.count)
        desc.nMaterials = Int64(materials.count); desc.nPointLights = Int64(lights.count)
        desc.nAreaLights = Int64(areaLights.count)
        desc.ambient = (scene.lights.ambient.x, scene.lights.ambient.y, scene.lights.ambient.z)
        desc.background = (scene.backgroundColor.x, scene.backgroundColor.y, scene.backgroundColor.z)
        desc.intersectionTestEpsilon = scene.intersectionTestEpsilon
        desc.shadowRayEpsilon = scene.shadowRayEpsilon
        desc.maxRecursionDepth = Int32(scene.maxRecursionDepth)
        // borrowed pointers: valid for the duration of rt_scene_create only (the module copies to the device)
        handle = triV.withUnsafeBufferPointer { v in triN.withUnsafeBufferPointer { n in
            spheres.withUnsafeBufferPointer { s in planes.withUnsafeBufferPointer { p in
            prims.withUnsafeBufferPointer { pr in blas.withUnsafeBufferPointer { b in
            instances.withUnsafeBufferPointer { i in materials.withUnsafeBufferPointer { m in
            lights.withUnsafeBufferPointer { l -> OpaquePointer? in
            triM.withUnsafeBufferPointer { tm in areaLights.withUnsafeBufferPointer { al -> OpaquePointer? in
                desc.triMotionBlur = tm.baseAddress; desc.areaLights = al.baseAddress
                desc.triVertices = v.baseAddress; desc.triNormals = n.baseAddress
                desc.spheres = s.baseAddress; desc.planes = p.baseAddress; desc.prims = pr.baseAddress
                desc.blas = b.baseAddress; desc.instances = i.baseAddress; desc.materials = m.baseAddress
                desc.pointLights = l.baseAddress
                // NULL on failure, never throws (CPly convention)
                return gpus.count > 1 ? gpus.withUnsafeBufferPointer { rt_scene_create_multi(&desc, $0.baseAddress, Int32(gpus.count)) }
                                      : rt_scene_create(&desc, gpus[0])
            } } } } } } } } } } }
        if handle == nil { print("❌ CRTCuda: \(String(cString: rt_last_error()))") }
    }

    deinit { rt_scene_destroy(handle) }              // destroy(NULL) is safe

    /// Animation: the counterpart of RTContext.refitTLAS (Models/RTContext.swift:883-928), which the reference defines
    /// but never calls — its batch driver builds a new engine per frame (BatchRender.swift:26-40).  New
    /// localToWorld matrices for the instances of the resident scene; worldBounds of mesh instances are left inverted
    /// (+inf / -inf) so that the module derives them on the device; the TLAS is refitted there.
    public func refit(localToWorld: [Mat4]) throws {
        guard let handle else { throw NSError(domain: "Render", code: -21, userInfo: [NSLocalizedDescriptionKey: "Renderer not initialized."]) }
        precondition(localToWorld.count == ctx.instances.count)
        var updated = [rt_instance]()
        for (i, inst) in ctx.instances.enumerated() {
            var ri = rt_instance()
            ri.blas = blasOfInstance[i]
            ri.hasMaterialOverride = inst.materialOverride == nil ? 0 : 1
            ri.materialOverride = Int64(inst.materialOverride ?? 0)
            let M = localToWorld[i]
            withUnsafeMutableBytes(of: &ri.localToWorld) { $0.copyBytes(from: Self.bytes(M)) }
            withUnsafeMutableBytes(of: &ri.worldToLocal) { $0.copyBytes(from: Self.bytes(M.inverse)) }
            withUnsafeMutableBytes(of: &ri.normalMatrix) { $0.copyBytes(from: Self.bytes3(normalTransformMatrix(from: M))) }
            if inst.blas.prims.count > 1 {
                ri.worldBounds = (.infinity, .infinity, .infinity, -.infinity, -.infinity, -.infinity)      // derive on the device
            } else {
                let b = inst.blas.prims[0].bounds.transformed(by: M)
                ri.worldBounds = (b.minP.x, b.minP.y, b.minP.z, b.maxP.x, b.maxP.y, b.maxP.z)
            }
            updated.append(ri)
        }
        let status = updated.withUnsafeMutableBufferPointer { rt_scene_update_instances(handle, $0.baseAddress, Int64($0.count), nil) }
        guard status == RT_OK else {
            throw NSError(domain: "Render", code: Int(status), userInfo: [NSLocalizedDescriptionKey: String(cString: rt_last_error())])
        }
    }

    /// render(scene:cameraIndex:progressRow:) — same signature and result as the reference (:52-56): W*H Vec3, 0..255 scale.
    @Sendable public func render(scene: Scene, cameraIndex: Int = 0,
                                 progressRow: (@Sendable (Int) -> Void)? = nil) async throws -> [Vec3] {
        let (linear, _, _) = try renderBoth(scene: scene, cameraIndex: cameraIndex, wantLinear: true, wantRGBA: false, progressRow: progressRow)
        return linear
    }

    /// RGBA8 straight from the kernel epilogue (clamp + truncate, alpha 255: RayTracer.swift:186-195) + ray counts.
    public func renderRGBA8(scene: Scene, cameraIndex: Int = 0,
                            progressRow: (@Sendable (Int) -> Void)? = nil) throws -> ([UInt8], rt_stats) {
        let (_, rgba, st) = try renderBoth(scene: scene, cameraIndex: cameraIndex, wantLinear: false, wantRGBA: true, progressRow: progressRow)
        return (rgba, st)
    }

    private func renderBoth(scene: Scene, cameraIndex: Int, wantLinear: Bool, wantRGBA: Bool,
                            progressRow: (@Sendable (Int) -> Void)?) throws -> ([Vec3], [UInt8], rt_stats) {
        guard let handle else { throw NSError(domain: "Render", code: -21, userInfo: [NSLocalizedDescriptionKey: "Renderer not initialized."]) }
        let cam = scene.cameras[cameraIndex]
        let width = max(1, cam.imageResolution.0), height = max(1, cam.imageResolution.1)
        var rc = rt_camera()
        rc.width = Int32(width); rc.height = Int32(height); rc.numSamples = Int32(cam.numSamples)
        rc.isLookAt = (cam.type?.lowercased() == "lookat") ? 1 : 0
        rc.hasFovy = cam.fovy == nil ? 0 : 1; rc.fovy = cam.fovy ?? 0
        rc.position = (cam.position.x, cam.position.y, cam.position.z); rc.gaze = (cam.gaze.x, cam.gaze.y, cam.gaze.z)
        rc.gazePoint = (cam.gazePoint.x, cam.gazePoint.y, cam.gazePoint.z); rc.up = (cam.up.x, cam.up.y, cam.up.z)
        rc.nearPlane = (cam.nearPlane[0], cam.nearPlane[1], cam.nearPlane[2], cam.nearPlane[3])
        rc.nearDistance = cam.nearDistance; rc.apertureSize = cam.apertureSize; rc.focusDistance = cam.focusDistance
        var linear = [Double](repeating: 0, count: wantLinear ? width * height * 3 : 0)
        var rgba = [UInt8](repeating: 0, count: wantRGBA ? width * height * 4 : 0)
        var stats = rt_stats()
        var params = rt_render_params()               // whole frame on the scene's GPU(s), automatic kernel choice
        // progress / cancellation while the frame is in flight (RayTracer.swift:172-181): the module calls back from this
        // thread with the rows the finished share of the work amounts to; a cancelled Task stops the frame (RT_ERR_CANCELLED)
        final class Relay { let row: (@Sendable (Int) -> Void)?; init(_ r: (@Sendable (Int) -> Void)?) { row = r } }
        let relay = Relay(progressRow)
        params.progressUser = Unmanaged.passUnretained(relay).toOpaque()
        params.progress = { rowsDone, _, user in
            let r = Unmanaged<Relay>.fromOpaque(user!).takeUnretainedValue()
            r.row?(Int(rowsDone) - 1)
            return Task.isCancelled ? 1 : 0
        }
        let status = withExtendedLifetime(relay) { linear.withUnsafeMutableBufferPointer { lp in rgba.withUnsafeMutableBufferPointer { rp in
            rt_render(handle, &rc, &params, wantRGBA ? rp.baseAddress : nil, wantLinear ? lp.baseAddress : nil, nil, &stats)
        } } }
        if status == RT_ERR_CANCELLED { throw CancellationError() }      // like `try Task.checkCancellation()` in the reference's chunk loop
        guard status == RT_OK else {
            throw NSError(domain: "Render", code: Int(status), userInfo: [NSLocalizedDescriptionKey: String(cString: rt_last_error())])
        }
        var out = [Vec3](); out.reserveCapacity(width * height)
        if wantLinear { for p in 0..<(width * height) { out.append(Vec3(linear[3 * p], linear[3 * p + 1], linear[3 * p + 2])) } }
        return (out, rgba, stats)
    }

    private static func bytes(_ m: Mat4) -> [UInt8] { withUnsafeBytes(of: m) { Array($0) } }          // simd_double4x4 is column-major
    private static func bytes3(_ m: simd_double3x3) -> [UInt8] {                                       // 3 columns padded to 32 B each
        var out = [Double](); for c in [m.columns.0, m.columns.1, m.columns.2] { out += [c.x, c.y, c.z] }
        return out.withUnsafeBytes { Array($0) }
    }
}
