// From SceneDesc to what crosses the C ABI: BVH build + flattening into
// pbrtgpu_scene (RenderOptions::MakeScene, api.cpp:1763-1772), and the
// camera/film/sampler/integrator parameters into pbrtgpu_render_params
// (RenderOptions::MakeIntegrator/MakeCamera, api.cpp:1774-1847).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <functional>
#include <cstring>

#include "api.h"
#include "setup.h"
#include "xform.h"

namespace pg {

pbrtgpu_scene FlatScene::view() const {
    pbrtgpu_scene s;
    memset(&s, 0, sizeof(s));
    s.n_nodes = (int32_t)bvhOffset.size();
    s.bvh_bounds = bvhBounds.data();
    s.bvh_offset = bvhOffset.data();
    s.bvh_nprims = bvhNPrims.data();
    s.bvh_axis = bvhAxis.data();
    s.n_prims = (int32_t)primType.size();
    s.prim_type = primType.data();
    s.prim_flags = primFlags.data();
    s.prim_shape = primShape.data();
    s.prim_material = primMaterial.data();
    s.prim_light = primLight.data();
    s.prim_orig_id = primOrigId.data();
    s.tri_p = triP.data();
    s.tri_n = triN.empty() ? nullptr : triN.data();
    s.tri_uv = triUV.empty() ? nullptr : triUV.data();
    s.tri_s = triS.empty() ? nullptr : triS.data();
    s.n_planes = (int32_t)(planeParams.size() / PBRTGPU_PLANE_STRIDE);
    s.plane_params = planeParams.data();
    s.n_spheres = (int32_t)(sphereParams.size() / PBRTGPU_SPHERE_STRIDE);
    s.sphere_params = sphereParams.data();
    s.n_materials = (int32_t)matTag.size();
    s.mat_tag = matTag.data();
    s.mat_params = matParams.data();
    s.n_lights = (int32_t)lightTag.size();
    s.light_tag = lightTag.data();
    s.light_prim = lightPrim.data();
    s.light_strategy = lightStrategy.data();
    s.light_portal_begin = lightPortalBegin.data();
    s.light_params = lightParams.data();
    s.light_nsamples = lightNSamples.data();
    s.n_portals = (int32_t)(portalParams.size() / PBRTGPU_PORTAL_STRIDE);
    s.portal_params = portalParams.data();
    memcpy(s.world_center, worldCenter, sizeof(worldCenter));
    s.world_radius = worldRadius;
    return s;
}

namespace {
pbrthost_hlbvh_builder gHlbvhBuilder = nullptr;
void *gHlbvhUser = nullptr;
double gBuildWallMs = 0;
float gBuilderMs = 0;
}  // namespace
void setHlbvhBuilder(pbrthost_hlbvh_builder fn, void *user) { gHlbvhBuilder = fn, gHlbvhUser = user; }
void lastBvhBuildMs(double *wallMs, float *builderMs) {
    if (wallMs) *wallMs = gBuildWallMs;
    if (builderMs) *builderMs = gBuilderMs;
}

bool flattenScene(const SceneDesc &desc, FlatScene *out, std::string *err) {
    FlatScene &fs = *out;
    // MakeAccelerator("bvh"), accelerators/bvh.cpp:740-760
    if (desc.acceleratorName != "bvh") {
        if (err) *err = "Accelerator \"" + desc.acceleratorName + "\" is not supported by the GPU path integrator (bvh only)";
        return false;
    }
    std::string splitName = desc.acceleratorParams.findOneString("splitmethod", "sah");
    SplitMethod sm;
    if (splitName == "sah") sm = SplitMethod::SAH;
    else if (splitName == "hlbvh") sm = SplitMethod::HLBVH;
    else if (splitName == "middle") sm = SplitMethod::Middle;
    else if (splitName == "equal") sm = SplitMethod::EqualCounts;
    else {
        hostWarning("BVH split method \"%s\" unknown.  Using \"sah\".", splitName.c_str());
        sm = SplitMethod::SAH;
    }
    int maxPrimsInNode = desc.acceleratorParams.findOneInt("maxnodeprims", 4);

    size_t n = desc.prims.size();
    std::vector<Bounds3> pb(n);
    for (size_t i = 0; i < n; ++i) pb[i] = primWorldBound(desc, desc.prims[i]);
    BuiltBVH bvh;
    auto t0 = std::chrono::steady_clock::now();
    gBuilderMs = 0;
    if (sm == SplitMethod::HLBVH && gHlbvhBuilder && n > 0) {
        // the treelets are built by the installed builder (libpbrtgpu's pbrtgpu_build_hlbvh); same tree, node for node
        static_assert(sizeof(Bounds3) == 6 * sizeof(float), "Bounds3 is six floats");
        size_t cap = 2 * n;
        bvh.bounds.resize(6 * cap), bvh.offset.resize(cap), bvh.nPrims.resize(cap), bvh.axis.resize(cap), bvh.orderedPrims.resize(n);
        int32_t nNodes = 0;
        int rc = gHlbvhBuilder(gHlbvhUser, &pb[0].pMin.x, (int32_t)n, maxPrimsInNode, (int32_t)cap, bvh.bounds.data(), bvh.offset.data(),
                               bvh.nPrims.data(), bvh.axis.data(), bvh.orderedPrims.data(), &nNodes, &gBuilderMs);
        if (rc != 0 || nNodes <= 0) {
            if (err) *err = "the installed HLBVH builder failed (status " + std::to_string(rc) + ")";
            return false;
        }
        bvh.bounds.resize(6 * (size_t)nNodes), bvh.offset.resize(nNodes), bvh.nPrims.resize(nNodes), bvh.axis.resize(nNodes);
    } else
        bvh = buildBVH(pb, maxPrimsInNode, sm);
    gBuildWallMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    fs.bvhBounds = std::move(bvh.bounds);
    fs.bvhOffset = std::move(bvh.offset);
    fs.bvhNPrims = std::move(bvh.nPrims);
    fs.bvhAxis = std::move(bvh.axis);

    bool anyN = false, anyUV = false, anyS = false;
    for (auto &m : desc.meshes) anyN |= !m.n.empty(), anyUV |= !m.uv.empty(), anyS |= !m.s.empty();
    fs.primType.resize(n); fs.primFlags.resize(n); fs.primShape.resize(n); fs.primMaterial.resize(n);
    fs.primLight.resize(n); fs.primOrigId.resize(n);
    fs.triP.assign(9 * n, 0.f);
    if (anyN) fs.triN.assign(9 * n, 0.f);
    if (anyUV) fs.triUV.assign(6 * n, 0.f);
    if (anyS) fs.triS.assign(9 * n, 0.f);
    std::vector<int> slotOfPrim(n, -1);
    for (size_t slot = 0; slot < n; ++slot) {
        int pi = bvh.orderedPrims[slot];
        slotOfPrim[pi] = (int)slot;
        const PrimDesc &p = desc.prims[pi];
        fs.primType[slot] = p.type;
        fs.primShape[slot] = p.shape;
        fs.primMaterial[slot] = p.material;
        fs.primLight[slot] = p.light;
        fs.primOrigId[slot] = pi;
        uint8_t flags = 0;
        if (p.type == PBRTGPU_PRIM_TRIANGLE) {
            const MeshData &m = desc.meshes[p.shape];
            if (m.reverseOrientation) flags |= PBRTGPU_PF_REVERSE_ORIENTATION;
            if (m.swapsHandedness) flags |= PBRTGPU_PF_SWAPS_HANDEDNESS;
            for (int k = 0; k < 3; ++k) {
                int vi = m.indices[3 * p.tri + k];
                fs.triP[9 * slot + 3 * k] = m.p[vi].x, fs.triP[9 * slot + 3 * k + 1] = m.p[vi].y, fs.triP[9 * slot + 3 * k + 2] = m.p[vi].z;
                if (!m.n.empty()) fs.triN[9 * slot + 3 * k] = m.n[vi].x, fs.triN[9 * slot + 3 * k + 1] = m.n[vi].y, fs.triN[9 * slot + 3 * k + 2] = m.n[vi].z;
                if (!m.s.empty()) fs.triS[9 * slot + 3 * k] = m.s[vi].x, fs.triS[9 * slot + 3 * k + 1] = m.s[vi].y, fs.triS[9 * slot + 3 * k + 2] = m.s[vi].z;
                if (!m.uv.empty()) fs.triUV[6 * slot + 2 * k] = m.uv[2 * vi], fs.triUV[6 * slot + 2 * k + 1] = m.uv[2 * vi + 1];
            }
            if (!m.n.empty()) flags |= PBRTGPU_PF_HAS_NORMALS;
            if (!m.uv.empty()) flags |= PBRTGPU_PF_HAS_UV;
            if (!m.s.empty()) flags |= PBRTGPU_PF_HAS_TANGENTS;
        } else if (p.type == PBRTGPU_PRIM_SPHERE) {
            if (desc.spheres[p.shape].reverseOrientation) flags |= PBRTGPU_PF_REVERSE_ORIENTATION;
            if (desc.spheres[p.shape].swapsHandedness) flags |= PBRTGPU_PF_SWAPS_HANDEDNESS;
        } else {
            if (desc.planes[p.shape].reverseOrientation) flags |= PBRTGPU_PF_REVERSE_ORIENTATION;
            if (desc.planes[p.shape].swapsHandedness) flags |= PBRTGPU_PF_SWAPS_HANDEDNESS;
        }
        fs.primFlags[slot] = flags;
    }

    for (auto &pl : desc.planes) {
        float rec[PBRTGPU_PLANE_STRIDE] = {0};
        rec[0] = pl.lo.x, rec[1] = pl.lo.y, rec[2] = pl.lo.z, rec[3] = pl.hi.x, rec[4] = pl.hi.y, rec[5] = pl.hi.z;
        rec[6] = (float)pl.ax, rec[7] = pl.facingFw ? 1.f : 0.f, rec[8] = pl.area;
        memcpy(rec + 16, pl.o2w.m.m, 64);
        memcpy(rec + 32, pl.o2w.mInv.m, 64);
        fs.planeParams.insert(fs.planeParams.end(), rec, rec + PBRTGPU_PLANE_STRIDE);
    }
    for (auto &sp : desc.spheres) {
        float rec[PBRTGPU_SPHERE_STRIDE] = {0};
        rec[0] = sp.radius, rec[1] = sp.zMin, rec[2] = sp.zMax, rec[3] = sp.thetaMin, rec[4] = sp.thetaMax, rec[5] = sp.phiMax;
        memcpy(rec + 16, sp.o2w.m.m, 64);
        memcpy(rec + 32, sp.o2w.mInv.m, 64);
        fs.sphereParams.insert(fs.sphereParams.end(), rec, rec + PBRTGPU_SPHERE_STRIDE);
    }
    for (auto &m : desc.materials) {
        fs.matTag.push_back(m.tag);
        fs.matParams.insert(fs.matParams.end(), m.params, m.params + PBRTGPU_MAT_STRIDE);
    }
    fs.lightPortalBegin.push_back(0);
    for (auto &l : desc.lights) {
        fs.lightTag.push_back(l.tag);
        fs.lightPrim.push_back(l.prim >= 0 ? slotOfPrim[l.prim] : -1);
        fs.lightStrategy.push_back(l.strategy);
        fs.lightNSamples.push_back(l.nSamples);
        fs.lightParams.insert(fs.lightParams.end(), l.params, l.params + PBRTGPU_LIGHT_STRIDE);
        for (auto &p : l.portals) {
            float rec[PBRTGPU_PORTAL_STRIDE] = {0};
            rec[0] = p.lo.x, rec[1] = p.lo.y, rec[2] = p.lo.z, rec[3] = p.hi.x, rec[4] = p.hi.y, rec[5] = p.hi.z;
            rec[6] = (float)p.ax, rec[7] = p.facingFw ? 1.f : 0.f, rec[8] = p.area;
            fs.portalParams.insert(fs.portalParams.end(), rec, rec + PBRTGPU_PORTAL_STRIDE);
        }
        fs.lightPortalBegin.push_back((int32_t)(fs.portalParams.size() / PBRTGPU_PORTAL_STRIDE));
    }
    // Scene::WorldBound() = aggregate bound; BoundingSphere (geometry.h:960-963)
    if (!fs.bvhOffset.empty()) {
        fs.worldBound.pMin = v3(fs.bvhBounds[0], fs.bvhBounds[1], fs.bvhBounds[2]);
        fs.worldBound.pMax = v3(fs.bvhBounds[3], fs.bvhBounds[4], fs.bvhBounds[5]);
        V3 c = (fs.worldBound.pMin + fs.worldBound.pMax) / 2;
        fs.worldCenter[0] = c.x, fs.worldCenter[1] = c.y, fs.worldCenter[2] = c.z;
        const Bounds3 &b = fs.worldBound;
        bool inside = c.x >= b.pMin.x && c.x <= b.pMax.x && c.y >= b.pMin.y && c.y <= b.pMax.y && c.z >= b.pMin.z && c.z <= b.pMax.z;
        fs.worldRadius = inside ? distance(c, b.pMax) : 0;
    }
    return true;
}

// ---- filters (src/filters/*.cpp), evaluated on the host only to fill the
// 16x16 table of Film's constructor (film.cpp:68-77).
static bool makeFilterTable(const SceneDesc &d, float radius[2], float table[256]) {
    const ParamSet &ps = d.filterParams;
    const std::string &name = d.filterName;
    std::function<float(float, float)> eval;
    float xw, yw;
    if (name == "box") {
        xw = ps.findOneFloat("xwidth", 0.5f), yw = ps.findOneFloat("ywidth", 0.5f);
        eval = [](float, float) { return 1.f; };
    } else if (name == "gaussian") {
        xw = ps.findOneFloat("xwidth", 2.f), yw = ps.findOneFloat("ywidth", 2.f);
        float alpha = ps.findOneFloat("alpha", 2.f);
        float expX = std::exp(-alpha * xw * xw), expY = std::exp(-alpha * yw * yw);
        eval = [=](float x, float y) {
            auto G = [alpha](float dd, float expv) { return std::max((float)0, float(std::exp(-alpha * dd * dd) - expv)); };
            return G(x, expX) * G(y, expY);
        };
    } else if (name == "mitchell") {
        xw = ps.findOneFloat("xwidth", 2.f), yw = ps.findOneFloat("ywidth", 2.f);
        float B = ps.findOneFloat("B", 1.f / 3.f), C = ps.findOneFloat("C", 1.f / 3.f);
        float invX = 1.f / xw, invY = 1.f / yw;
        eval = [=](float px, float py) {
            auto M = [B, C](float x) {
                x = std::abs(2 * x);
                if (x > 1)
                    return ((-B - 6 * C) * x * x * x + (6 * B + 30 * C) * x * x + (-12 * B - 48 * C) * x + (8 * B + 24 * C)) * (1.f / 6.f);
                return ((12 - 9 * B - 6 * C) * x * x * x + (-18 + 12 * B + 6 * C) * x * x + (6 - 2 * B)) * (1.f / 6.f);
            };
            return M(px * invX) * M(py * invY);
        };
    } else if (name == "sinc") {
        xw = ps.findOneFloat("xwidth", 4.f), yw = ps.findOneFloat("ywidth", 4.f);
        float tau = ps.findOneFloat("tau", 3.f);
        eval = [=](float px, float py) {
            auto Sinc = [](float x) {
                x = std::abs(x);
                if (x < 1e-5) return 1.f;
                return std::sin(kPi * x) / (kPi * x);
            };
            auto W = [&](float x, float r) {
                x = std::abs(x);
                if (x > r) return 0.f;
                float lanczos = Sinc(x / tau);
                return Sinc(x) * lanczos;
            };
            return W(px, xw) * W(py, yw);
        };
    } else if (name == "triangle") {
        xw = ps.findOneFloat("xwidth", 2.f), yw = ps.findOneFloat("ywidth", 2.f);
        eval = [=](float px, float py) { return std::max((float)0, xw - std::abs(px)) * std::max((float)0, yw - std::abs(py)); };
    } else {
        hostError("Filter \"%s\" unknown.", name.c_str());
        return false;
    }
    ps.reportUnused();
    radius[0] = xw, radius[1] = yw;
    int offset = 0;
    for (int y = 0; y < 16; ++y)
        for (int x = 0; x < 16; ++x, ++offset) {
            float px = (x + 0.5f) * xw / 16, py = (y + 0.5f) * yw / 16;
            table[offset] = eval(px, py);
        }
    return true;
}

bool makeRenderSetup(const SceneDesc &d, RenderSetup *out) {
    pbrtgpu_render_params &rp = out->params;
    memset(&rp, 0, sizeof(rp));
    if (d.integratorName != "path" && d.integratorName != "directlighting") {
        hostError("Integrator \"%s\" is not a plugin this library replaces (Integrator \"path\" and \"directlighting\")", d.integratorName.c_str());
        return false;
    }
    if (d.cameraName != "perspective") {
        hostError("Camera \"%s\" is not supported by the GPU path integrator (perspective only)", d.cameraName.c_str());
        return false;
    }
    if (d.samplerName != "halton") {
        hostError("Sampler \"%s\" is not supported by the GPU path integrator (halton only)", d.samplerName.c_str());
        return false;
    }
    if (d.filmName != "image") {
        hostError("Film \"%s\" unknown.", d.filmName.c_str());
        return false;
    }
    if (!makeFilterTable(d, rp.filter_radius, rp.filter_table)) return false;

    // CreateFilm, film.cpp:213-252
    const ParamSet &fp = d.filmParams;
    out->filename = fp.findOneString("filename", "pbrt.exr");
    int xres = fp.findOneInt("xresolution", 1280), yres = fp.findOneInt("yresolution", 720);
    float crop[4] = {0, 1, 0, 1};
    if (auto cr = fp.findFloats("cropwindow")) {
        if (cr->size() == 4) {
            crop[0] = clampf(std::min((*cr)[0], (*cr)[1]), 0.f, 1.f);
            crop[1] = clampf(std::max((*cr)[0], (*cr)[1]), 0.f, 1.f);
            crop[2] = clampf(std::min((*cr)[2], (*cr)[3]), 0.f, 1.f);
            crop[3] = clampf(std::max((*cr)[2], (*cr)[3]), 0.f, 1.f);
        } else
            hostError("%d values supplied for \"cropwindow\". Expected 4.", (int)cr->size());
    }
    out->scale = fp.findOneFloat("scale", 1.);
    fp.findOneFloat("diagonal", 35.);
    rp.max_sample_luminance = fp.findOneFloat("maxsampleluminance", INFINITY);
    fp.reportUnused();
    rp.full_resolution[0] = xres, rp.full_resolution[1] = yres;
    rp.cropped_bounds[0] = (int)std::ceil(xres * crop[0]);
    rp.cropped_bounds[1] = (int)std::ceil(yres * crop[2]);
    rp.cropped_bounds[2] = (int)std::ceil(xres * crop[1]);
    rp.cropped_bounds[3] = (int)std::ceil(yres * crop[3]);
    // Film::GetSampleBounds, film.cpp:80-86
    rp.sample_bounds[0] = (int)std::floor(float(rp.cropped_bounds[0]) + 0.5f - rp.filter_radius[0]);
    rp.sample_bounds[1] = (int)std::floor(float(rp.cropped_bounds[1]) + 0.5f - rp.filter_radius[1]);
    rp.sample_bounds[2] = (int)std::ceil(float(rp.cropped_bounds[2]) - 0.5f + rp.filter_radius[0]);
    rp.sample_bounds[3] = (int)std::ceil(float(rp.cropped_bounds[3]) - 0.5f + rp.filter_radius[1]);

    // CreatePerspectiveCamera + ProjectiveCamera ctor
    const ParamSet &cp = d.cameraParams;
    float shutteropen = cp.findOneFloat("shutteropen", 0.f), shutterclose = cp.findOneFloat("shutterclose", 1.f);
    if (shutterclose < shutteropen) {
        hostWarning("Shutter close time [%f] < shutter open [%f].  Swapping them.", shutterclose, shutteropen);
        std::swap(shutterclose, shutteropen);
    }
    rp.shutter_open = shutteropen, rp.shutter_close = shutterclose;
    rp.lens_radius = cp.findOneFloat("lensradius", 0.f);
    rp.focal_distance = cp.findOneFloat("focaldistance", 1e6);
    float frame = cp.findOneFloat("frameaspectratio", float(xres) / float(yres));
    float screen[4];  // xmin xmax ymin ymax
    if (frame > 1.f) screen[0] = -frame, screen[1] = frame, screen[2] = -1.f, screen[3] = 1.f;
    else screen[0] = -1.f, screen[1] = 1.f, screen[2] = -1.f / frame, screen[3] = 1.f / frame;
    if (auto sw = cp.findFloats("screenwindow")) {
        if (sw->size() == 4) for (int i = 0; i < 4; ++i) screen[i] = (*sw)[i];
        else hostError("\"screenwindow\" should have four values");
    }
    float fov = cp.findOneFloat("fov", 90.);
    float halffov = cp.findOneFloat("halffov", -1.f);
    if (halffov > 0.f) fov = 2.f * halffov;
    cp.reportUnused();
    Xform cameraToScreen = xfPerspective(fov, 1e-2f, 1000.f);
    Xform screenToRaster = xfMul(xfMul(xfScale((float)xres, (float)yres, 1),
                                       xfScale(1 / (screen[1] - screen[0]), 1 / (screen[2] - screen[3]), 1)),
                                 xfTranslate(v3(-screen[0], -screen[3], 0)));
    Xform rasterToScreen = xfInverse(screenToRaster);
    Xform rasterToCamera = xfMul(xfInverse(cameraToScreen), rasterToScreen);
    memcpy(rp.raster_to_camera, rasterToCamera.m.m, 64);
    memcpy(rp.camera_to_world, d.cameraToWorld.m.m, 64);

    // CreateHaltonSampler, halton.cpp:133-139
    rp.spp = d.samplerParams.findOneInt("pixelsamples", 16);
    rp.sample_at_pixel_center = d.samplerParams.findOneBool("samplepixelcenter", false) ? 1 : 0;
    d.samplerParams.reportUnused();

    // CreatePathIntegrator, path.cpp:191-214
    const ParamSet &ip = d.integratorParams;
    rp.max_depth = ip.findOneInt("maxdepth", 5);
    memcpy(rp.pixel_bounds, rp.sample_bounds, sizeof(rp.pixel_bounds));
    if (auto pb = ip.findInts("pixelbounds")) {
        if (pb->size() != 4)
            hostError("Expected four values for \"pixelbounds\" parameter. Got %d.", (int)pb->size());
        else {
            rp.pixel_bounds[0] = std::max(rp.pixel_bounds[0], (*pb)[0]);
            rp.pixel_bounds[2] = std::min(rp.pixel_bounds[2], (*pb)[1]);
            rp.pixel_bounds[1] = std::max(rp.pixel_bounds[1], (*pb)[2]);
            rp.pixel_bounds[3] = std::min(rp.pixel_bounds[3], (*pb)[3]);
            if ((rp.pixel_bounds[2] - rp.pixel_bounds[0]) * (rp.pixel_bounds[3] - rp.pixel_bounds[1]) == 0)
                hostError("Degenerate \"pixelbounds\" specified.");
        }
    }
    if (d.integratorName == "directlighting") {
        // CreateDirectLightingIntegrator, directlighting.cpp:100-137
        rp.integrator = PBRTGPU_INTEGRATOR_DIRECT;
        // With allowMultipleLobes = false (integrator.cpp:615) glass is SpecularReflection + SpecularTransmission and
        // Li recurses into BOTH (directlighting.cpp:91-95).  The device walks that tree depth first with a per-path
        // branch stack (integrator_stages.h); it evaluates the transmission branch when it is pushed, which is only
        // the reference's result if the sample drawn later cannot change it -- i.e. one specular-transmission lobe.
        for (auto &m : d.materials) {
            bool twoTransLobes = m.tag == PBRTGPU_MAT_UBER && (m.params[9] != 0 || m.params[10] != 0 || m.params[11] != 0) &&
                                 (m.params[12] < 1 || m.params[13] < 1 || m.params[14] < 1);
            if (m.tag == PBRTGPU_MAT_DISPERSIVE_GLASS || twoTransLobes) {
                hostError("Integrator \"directlighting\": %s is not supported by the GPU integrator",
                          twoTransLobes ? "Material \"uber\" with both Kt and opacity < 1 (two specular transmission lobes)"
                                        : "Material \"dispersive_glass\"");
                return false;
            }
        }
        std::string st = ip.findOneString("strategy", "all");
        if (st == "one") rp.direct_strategy = PBRTGPU_DIRECT_ONE;
        else if (st == "all") rp.direct_strategy = PBRTGPU_DIRECT_ALL;
        else {
            hostWarning("Strategy \"%s\" for direct lighting unknown. Using \"all\".", st.c_str());
            rp.direct_strategy = PBRTGPU_DIRECT_ALL;
        }
        rp.rr_threshold = 1.f;
    } else {
        rp.integrator = PBRTGPU_INTEGRATOR_PATH;
        rp.rr_threshold = ip.findOneFloat("rrthreshold", 1.);
        std::string ls = ip.findOneString("lightsamplestrategy", "uniform");
        if (ls == "uniform") rp.light_strategy = 0;
        else if (ls == "power") rp.light_strategy = 1;
        else {
            hostError("Light sample distribution type \"%s\" is not supported by the GPU path integrator. Using \"uniform\".", ls.c_str());
            rp.light_strategy = 0;
        }
    }
    ip.reportUnused();
    rp.shard_index = 0, rp.shard_count = 1;
    rp.spp_begin = 0, rp.spp_end = 0;
    return true;
}

// Film::WriteImage's per-pixel arithmetic (film.cpp:169-207), splats excluded.
void filmToRGB(const float *xyzw, int nPixels, float scale, float *rgb) {
    for (int i = 0; i < nPixels; ++i) {
        float c[3];
        xyzToRGB(&xyzw[4 * i], c);
        float w = xyzw[4 * i + 3];
        if (w != 0) {
            float invWt = 1.f / w;
            c[0] = std::max(0.f, c[0] * invWt);
            c[1] = std::max(0.f, c[1] * invWt);
            c[2] = std::max(0.f, c[2] * invWt);
        }
        // (splat term: + splatScale * 0)
        c[0] += 0.f, c[1] += 0.f, c[2] += 0.f;
        rgb[3 * i] = c[0] * scale, rgb[3 * i + 1] = c[1] * scale, rgb[3 * i + 2] = c[2] * scale;
    }
}

}  // namespace pg
