// Second half of integrator.h: shade, next-event resolve, film.
#ifndef PG_DEVICE_INTEGRATOR_STAGES_H
#define PG_DEVICE_INTEGRATOR_STAGES_H

namespace pg {

// Debug builds only: -DPG_HOST_TRACE (tests/hostsim with g++) or -DPG_DEVICE_TRACE (an A/B build of the CUDA library,
// tools/build_variants.py) print every vertex; render ONE sample (pixel_bounds of one pixel, one sample number) with them.
#if !defined(__CUDACC__) && defined(PG_HOST_TRACE)
#define PG_TRACE(...) fprintf(stderr, __VA_ARGS__)
#elif defined(__CUDA_ARCH__) && defined(PG_DEVICE_TRACE)
#define PG_TRACE(...) printf(__VA_ARGS__)
#else
#define PG_TRACE(...)
#endif

// ---- stage 2b: a path ray that escaped.  Environment light on the camera ray
// or after a specular bounce (path.cpp:91-101), then the path ends (path.cpp:104).
// Integrator "directlighting" over materials with specular transmission: Li recurses into reflection, then
// transmission (directlighting.cpp:91-95).  The tree is walked depth first, one vertex per wavefront
// iteration: the path follows the reflection branch and keeps the transmission branch of every vertex on a
// per-slot stack; when the branch it follows ends (a miss, no specular lobe, the depth limit) the most recent
// pending branch continues.  Each stack entry also stands for SpecularTransmit's sampler.Get2D()
// (integrator.cpp:684), which the reference draws after the whole reflection subtree: popping an entry
// advances the sampler dimension by two whether or not the entry carries a ray.
PG_HD void popBranch(const WaveParams &w, int depth, int cur, int slot) {
    int flags = w.flags[slot];
    int count = (flags >> 26) & 31;
    if (count == 0) return;
    int dim = w.dim[slot];
    F4 *stack = w.branch + 3 * (size_t)slot * w.branchCapacity;
    F4 a = f4(0, 0, 0, 0), b = a, c = a;
    bool found = false;
    while (count > 0 && !found) {
        --count;
        if (dim + 1 >= kArrayStartDim && dim < w.hal.arrayEndDim) dim = w.hal.arrayEndDim;  // GlobalSampler::Get2D
        dim += 2;
        a = stack[3 * count], b = stack[3 * count + 1], c = stack[3 * count + 2];
        found = a.w != 0.f;
    }
    w.dim[slot] = dim;
    flags = (flags & ~(31 << 26)) | (count << 26);
    if (found) {
        int bounces = (int)b.w;
        flags = (flags & ~0xff) | bounces | 256;
        F4 bq = w.beta[slot];
        w.beta[slot] = f4(c.x, c.y, c.z, bq.w);
        Ray r;
        r.o = v3(a.x, a.y, a.z), r.d = v3(b.x, b.y, b.z), r.tMax = kInfinity;
        int np = atomicIncInt(&w.cnt->nActive[depth + 1]);
        storeRay(w.rays[cur ^ 1], np, r);
        w.qSlot[cur ^ 1][np] = slot;
    }
    w.flags[slot] = flags;
}

PG_HD void stageMiss(const WaveParams &w, int depth, int cur, int pos) {
    int slot = w.qSlot[cur][pos];
    int flags = w.flags[slot];
    int bounces = flags & 255;
    bool specularBounce = (flags >> 8) & 1;
    if (bounces == 0 || specularBounce) {
        bool any = false;
        for (int i = 0; i < w.sc.nLights; ++i) any |= PG_LD(w.sc.lightTag + i) == PBRTGPU_LIGHT_INFINITE;
        if (any) {
            Ray ray = loadRay(w.rays[cur], pos);
            F4 Lq = w.L[slot], bq = w.beta[slot];
            Spec L = spec(Lq.x, Lq.y, Lq.z), beta = spec(bq.x, bq.y, bq.z);
            for (int i = 0; i < w.sc.nLights; ++i)
                if (PG_LD(w.sc.lightTag + i) == PBRTGPU_LIGHT_INFINITE)
                    L += beta * infiniteLe(w.sc.lightParams + PBRTGPU_LIGHT_STRIDE * (size_t)i, ray.d);
            w.L[slot] = f4(L.r, L.g, L.b, Lq.w);
        }
    }
    if (w.directTree) popBranch(w, depth, cur, slot);
}

// ---- stage 3: one path vertex (the body of path.cpp:81-187 after scene.Intersect found a hit).
// pos = position in the depth-d closest-hit queue; cur = which of the two
// ping-pong queues holds depth d.
//
// The body is written as a straight sequence of PHASES, each guarded by
// per-thread predicates, with sync() between them.  On the GPU sync() is
// __syncthreads(): the shading code is ~300 KB of SASS, ten times the SM's
// instruction cache, and warps that drift apart in it stall on instruction
// fetch (ncu: 74 % of stall samples "no instruction", i-cache hit rate 45 %).
// Keeping the warps of a block inside the same phase makes one warp's fetch
// serve them all.  Every thread of the block calls this function, `active`
// or not.  On the CPU (tests/hostsim) sync() does nothing.
struct NoSync {
    PG_HD void operator()(int) const {}
};

// The two records that live across all phases and are handed by reference to the out-of-line leaves.  On the
// GPU k_shade_sync keeps one per thread in SHARED memory: as locals they sit on the per-thread stack, and
// 1024 threads x ~1 KB of stack is several times the L1 (ncu r01e: three quarters of the kernel's L2 traffic
// and most of its DRAM traffic were local-memory sectors).  37 words: odd, so the threads of a warp hit 32
// different banks when they read the same member.
struct ShadeScratch {
    SurfInt it;
    BSDF bsdf;
    int pad;
};
static_assert(sizeof(ShadeScratch) == 37 * 4, "ShadeScratch must stay an odd number of words");

// LAMBERT: every vertex handed in lies on a material whose BSDF is one LambertianReflection (MatLobes::shadeClass 0,
// the front run of the hit queue) or on a primitive without a material; the BSDF calls are the closed forms of
// bsdf_eval.h (same arithmetic, bit-identical results), and nothing of the lobe machinery is compiled in.
// PORTAL = false: the scene has no portal light; the portal estimators are not compiled in.
// SPEC: every vertex handed in lies on a material all of whose lobes are specular (MatLobes::shadeClass 2: smooth glass, mirror).
// UniformSampleOneLight is not called for it (path.cpp:121: NumComponents(BSDF_ALL & ~BSDF_SPECULAR) == 0), so phases 2-6 -- and
// their barriers -- are not compiled in; what remains is the general code, expression for expression.
template <class Sync, bool EXT = true, bool LAMBERT = false, bool PORTAL = true, bool SPEC = false>
PG_HD void stageShadePhased(const WaveParams &w, int depth, int cur, int pos, bool active, Sync sync, ShadeScratch &scr) {
    const int flagsNS = BSDF_ALL & ~BSDF_SPECULAR;
    // ---- phase 0: path state, the hit record -> SurfaceInteraction, emitted light (path.cpp:91-96)
    int slot = 0, prim = 0, flags = 0, bounces = 0, mat = -1;
    bool specularBounce = false, touchL = false;
    Ray ray, next;
    Spec L = spec(0.f), beta = spec(0.f);
    float wvl0 = 0, etaScale = 0;
    SurfInt &it = scr.it;
    if (active) {
        slot = w.qSlot[cur][pos];
        ray = loadRay(w.rays[cur], pos);
        F4 hit = w.hits[pos];
        prim = f2i(hit.x);
        F4 bq = w.beta[slot];
        beta = spec(bq.x, bq.y, bq.z);
        etaScale = bq.w;
        flags = w.flags[slot];
        bounces = flags & 255;
        specularBounce = (flags >> 8) & 1;
        // The path's radiance record {L, hero wavelength} is read and written back only by the vertices that need it: those
        // that add emitted light (path.cpp:91-96) and, in scenes with a dispersive material, all of them (the wavelength).
        // Everywhere else the 16-byte gather and the 16-byte scatter per vertex would move an unchanged value.
        touchL = !w.skipWavelength || ((bounces == 0 || specularBounce) && w.sc.primLight && PG_LD(w.sc.primLight + prim) >= 0);
        if (touchL) {
            F4 Lq = w.L[slot];
            L = spec(Lq.x, Lq.y, Lq.z);
            wvl0 = Lq.w;
        }
        it = hitInteraction(w.sc, prim, hit.y, hit.z, hit.w, ray);
        PG_TRACE("shade d=%d slot=%d prim=%d o=(%.9g %.9g %.9g) d=(%.9g %.9g %.9g) p=(%.9g %.9g %.9g) n=(%g %g %g) bounces=%d\n", depth, slot, prim, ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, it.p.x, it.p.y, it.p.z, it.n.x, it.n.y, it.n.z, bounces);
        if (bounces == 0 || specularBounce) L += beta * hitLe(w.sc, prim, it.n, -ray.d);
    }
    bool alive = active && bounces < w.maxDepth;
    if (alive) {
        if (it.wvlsLost) wvl0 = 0.f;  // Transform::operator()(SurfaceInteraction) leaves wvls default-constructed
        mat = w.sc.primMaterial ? PG_LD(w.sc.primMaterial + prim) : -1;
        // no material: skip the boundary, same bounce count (path.cpp:108-113)
        if (mat < 0) next = spawnRay(it.p, it.pError, it.n, ray.d);
    }
    const bool scatter = alive && mat >= 0;
    sync(0);

    // ---- phase 1: the material's BSDF (path.cpp:106-107)
    PathSampler smp;
    smp.index = 0, smp.dim = 0;
    BSDF &bsdf = scr.bsdf;
    bsdf.nLobes = 0;
    Spec lambertR = spec(0.f);
    if (scatter) {
        smp.index = w.hIndex[slot];
        smp.dim = w.dim[slot];
        makeBSDF(it, w.sc.matLobes, mat, w.sc.matParams + PBRTGPU_MAT_STRIDE * (size_t)mat, wvl0, &bsdf);
        if (LAMBERT) {
            const float *r = &w.sc.matLobes[mat].lobes[0].R.r;
            lambertR = spec(PG_LD(r), PG_LD(r + 1), PG_LD(r + 2));
        }
    }
    sync(1);

    // ---- phase 2: UniformSampleOneLight's samples and light pick (integrator.cpp:100-122)
    bool nee = !SPEC && scatter && (LAMBERT || bsdfNumComponents(bsdf, flagsNS) > 0) && w.sc.nLights > 0;
    int li = 0, lightTag = 0;
    float pickPdf = 0, uL0 = 0, uL1 = 0, uS0 = 0, uS1 = 0;
    if (nee) {
        float uPick = get1D(w.hal, smp);
        li = findIntervalCdf(w.sc.lightCdf, w.sc.nLights + 1, uPick);
        pickPdf = (w.sc.lightFuncInt > 0) ? PG_LD(w.sc.lightFunc + li) / (w.sc.lightFuncInt * w.sc.nLights) : 0;
        if (pickPdf == 0) nee = false;
    }
    if (nee) {
        get2D(w.hal, smp, &uL0, &uL1);
        get2D(w.hal, smp, &uS0, &uS1);
        lightTag = PG_LD(w.sc.lightTag + li);
    }
    if (!SPEC) sync(2);

    // ---- phase 3: sample the light (EstimateDirect, integrator.cpp:133-141); portal lights take their own estimator
    const bool portal = PORTAL && nee && lightTag == PBRTGPU_LIGHT_PORTAL_AREA;
    const bool stock = nee && !portal;
    const bool isDelta = lightTag == PBRTGPU_LIGHT_POINT || lightTag == PBRTGPU_LIGHT_DISTANT || lightTag == PBRTGPU_LIGHT_SPOT;
    V3 wi = v3(0, 0, 0);  // Vector3f() is zero-initialised
    float lightPdf = 0, scatteringPdf = 0;
    ShapeSample pL;
    Spec Li = spec(0.f);
    if (portal) {
        NeeRequest r;
        r.hasShadow = r.hasClosest = false;
        r.light = -1;
        r.mode = NEE_MIS_LIGHT;
        estimateDirectPortal<EXT>(w.sc, li, it, bsdf, uS0, uS1, uL0, uL1, spec(1.f), &r);
        if (r.hasClosest) {
            Spec c = beta * (r.closestContrib / pickPdf);
            int np = atomicIncInt(&w.cnt->nNee[depth]);
            storeRay(w.neeRays, np, r.closestRay);
            w.neeContrib[np] = f4(c.r, c.g, c.b, bitsToFloat((unsigned)slot));
            w.neeInfo[2 * (size_t)np] = r.light;
            w.neeInfo[2 * (size_t)np + 1] = r.mode;
        }
    }
    if (stock) Li = lightSampleLi(w.sc, li, it, uL0, uL1, &wi, &lightPdf, &pL);
    if (stock) PG_TRACE("  light d=%d slot=%d dim=%d li=%d uL=(%.9g %.9g) uS=(%.9g %.9g) wi=(%.9g %.9g %.9g) pdf=%.9g Li=(%.9g %.9g %.9g) pL=(%.9g %.9g %.9g)\n", depth, slot, smp.dim, li, uL0, uL1, uS0, uS1, wi.x, wi.y, wi.z, lightPdf, Li.r, Li.g, Li.b, pL.p.x, pL.p.y, pL.p.z);
    if (!SPEC) sync(3);

    // ---- phase 4: BSDF value and pdf for the light sample; the shadow ray (integrator.cpp:142-184)
    if (stock && lightPdf > 0 && !isBlack(Li)) {
        Spec f = (LAMBERT ? lambertF(bsdf, lambertR, it.wo, wi) : bsdfF(bsdf, it.wo, wi, flagsNS)) * absDot(wi, it.ns);
        scatteringPdf = LAMBERT ? lambertPdf(bsdf, it.wo, wi) : bsdfPdf(bsdf, it.wo, wi, flagsNS);
        if (!isBlack(f)) {
            Spec Ld;
            if (isDelta) Ld = f * Li / lightPdf;
            else Ld = f * Li * powerHeuristic(lightPdf, scatteringPdf) / lightPdf;
            Spec c = beta * (Ld / pickPdf);
            Ray sr = spawnRayTo(it.p, it.pError, it.n, pL.p, pL.pError, pL.n);
            int sp = atomicIncInt(&w.cnt->nShadow[depth]);
            storeRay(w.shadowRays, sp, sr);
            w.shadowContrib[sp] = f4(c.r, c.g, c.b, bitsToFloat((unsigned)slot));
            PG_TRACE("  shadow d=%d slot=%d f=(%.9g %.9g %.9g) spdf=%.9g c=(%.9g %.9g %.9g) o=(%.9g %.9g %.9g) d=(%.9g %.9g %.9g)\n", depth, slot, f.r, f.g, f.b, scatteringPdf, c.r, c.g, c.b, sr.o.x, sr.o.y, sr.o.z, sr.d.x, sr.d.y, sr.d.z);
        }
    }
    if (!SPEC) sync(4);

    // ---- phase 5: sample the BSDF for MIS (integrator.cpp:187-207)
    bool misRay = stock && !isDelta;
    Spec fMis = spec(0.f);
    bool sampledSpecular = false;
    if (misRay) {
        int sampledType;
        fMis = LAMBERT ? lambertSampleF(bsdf, lambertR, it.wo, &wi, uS0, uS1, &scatteringPdf, &sampledType)
                       : bsdfSampleF(bsdf, it.wo, &wi, uS0, uS1, &scatteringPdf, flagsNS, &sampledType);
        fMis *= absDot(wi, it.ns);
        sampledSpecular = (sampledType & BSDF_SPECULAR) != 0;
        misRay = !isBlack(fMis) && scatteringPdf > 0;
    }
    if (!SPEC) sync(5);

    // ---- phase 6: the light's pdf for that direction; the MIS ray (integrator.cpp:208-254)
    if (misRay) {
        float weight = 1;
        if (!sampledSpecular) {
            lightPdf = lightPdfLi(w.sc, li, it, wi);
            if (lightPdf == 0) misRay = false;
            else weight = powerHeuristic(scatteringPdf, lightPdf);
        }
        Ray nr;
        if (misRay) {
            nr = spawnRay(it.p, it.pError, it.n, wi);
            // the ray can only count if its closest hit is the light's own shape: skip it when it misses that shape
            if (lightTag == PBRTGPU_LIGHT_DIFFUSE_AREA && !primitiveAccepts(w.sc.bvh, PG_LD(w.sc.lightPrim + li), nr)) {
                misRay = false;
                atomicIncInt(&w.cnt->nNeeCulled[depth]);
            }
        }
        if (misRay) {
            Spec c = beta * ((fMis * weight / scatteringPdf) / pickPdf);
            int np = atomicIncInt(&w.cnt->nNee[depth]);
            storeRay(w.neeRays, np, nr);
            w.neeContrib[np] = f4(c.r, c.g, c.b, bitsToFloat((unsigned)slot));
            w.neeInfo[2 * (size_t)np] = li;
            w.neeInfo[2 * (size_t)np + 1] = NEE_MIS_LIGHT;
        }
    }
    if (!SPEC) sync(6);

    // ---- phase 7: sample the BSDF for the next direction (path.cpp:131-151)
    V3 wo = v3(0, 0, 0);
    float pdf = 0;
    int sampled = 0;
    Spec f = spec(0.f);
    if (scatter) {
        wo = -ray.d;
        wi = v3(0, 0, 0);
        float u0, u1;
        get2D(w.hal, smp, &u0, &u1);
        f = LAMBERT ? lambertSampleF(bsdf, lambertR, wo, &wi, u0, u1, &pdf, &sampled) : bsdfSampleF(bsdf, wo, &wi, u0, u1, &pdf, BSDF_ALL, &sampled);
    }
    sync(7);

    // ---- phase 8: throughput, Russian roulette (path.cpp:177-184), the next ray
    if (scatter) {
        if (isBlack(f) || pdf == 0.f) alive = false;
        else {
            beta *= f * absDot(wi, it.ns) / pdf;
            specularBounce = (sampled & BSDF_SPECULAR) != 0;
            if ((sampled & BSDF_SPECULAR) && (sampled & BSDF_TRANSMISSION)) {
                float eta = bsdf.eta;
                etaScale *= (dot(wo, it.n) > 0) ? (eta * eta) : 1 / (eta * eta);
            }
            next = spawnRay(it.p, it.pError, it.n, wi);
            Spec rrBeta = beta * etaScale;
            if (maxComponentValue(rrBeta) < w.rrThreshold && bounces > 3) {
                float q = maxf(.05f, 1 - maxComponentValue(rrBeta));
                if (get1D(w.hal, smp) < q) alive = false;
                else beta /= 1 - q;
            }
            ++bounces;
        }
        w.dim[slot] = smp.dim;
    }
    if (active && touchL) w.L[slot] = f4(L.r, L.g, L.b, wvl0);
    if (alive) {
        w.beta[slot] = f4(beta.r, beta.g, beta.b, etaScale);
        w.flags[slot] = (flags & ~0x1ff) | bounces | (specularBounce ? 256 : 0);
        int np = atomicIncInt(&w.cnt->nActive[depth + 1]);
        storeRay(w.rays[cur ^ 1], np, next);
        w.qSlot[cur ^ 1][np] = slot;
    }
}

PG_HD void stageShade(const WaveParams &w, int depth, int cur, int pos) {
    ShadeScratch scr;
    stageShadePhased(w, depth, cur, pos, true, NoSync(), scr);
}
// the Lambert instantiation, for callers that know the vertex's material class (tests/hostsim with PBRTGPU_HOSTSIM_LAMBERT)
PG_HD void stageShadeLambert(const WaveParams &w, int depth, int cur, int pos) {
    ShadeScratch scr;
    stageShadePhased<NoSync, true, true, true>(w, depth, cur, pos, true, NoSync(), scr);
}
// the all-specular instantiation (MatLobes::shadeClass 2)
PG_HD void stageShadeSpecular(const WaveParams &w, int depth, int cur, int pos) {
    ShadeScratch scr;
    stageShadePhased<NoSync, true, false, true, true>(w, depth, cur, pos, true, NoSync(), scr);
}

// ---- stage 3 for Integrator "directlighting": DirectLightingIntegrator::Li (directlighting.cpp:63-98)
// at one vertex.  Direct light from all lights (UniformSampleAllLights, integrator.cpp:69-98, with the
// sampler's 2D arrays) or from one (UniformSampleOneLight without a light distribution,
// integrator.cpp:100-122), then SpecularReflect (integrator.cpp:639-675).  SpecularTransmit is not
// followed: materials with specular transmission are refused at load (host/setup.cpp); for all other
// materials its Sample_f finds no lobe and returns black, and the Get2D it draws comes after everything
// this path still uses.  The recursion f * Li(...) * |cos| / pdf is carried as throughput.
PG_HD void pushDirectRays(const WaveParams &w, int depth, int slot, const NeeRequest &r) {
    if (r.hasShadow) {
        int sp = atomicIncInt(&w.cnt->nShadow[depth]);
        if (sp < w.neeCapacity) {
            storeRay(w.shadowRays, sp, r.shadowRay);
            w.shadowContrib[sp] = f4(r.shadowContrib.r, r.shadowContrib.g, r.shadowContrib.b, bitsToFloat((unsigned)slot));
        } else
            atomicAddU64(&w.cnt->dropped, 1ull);  // never with queues of nSlots x light samples per vertex; reported, not silent
    }
    if (r.hasClosest) {
        int np = atomicIncInt(&w.cnt->nNee[depth]);
        if (np < w.neeCapacity) {
            storeRay(w.neeRays, np, r.closestRay);
            w.neeContrib[np] = f4(r.closestContrib.r, r.closestContrib.g, r.closestContrib.b, bitsToFloat((unsigned)slot));
            w.neeInfo[2 * (size_t)np] = r.light;
            w.neeInfo[2 * (size_t)np + 1] = r.mode;
        } else
            atomicAddU64(&w.cnt->dropped, 1ull);
    }
}

PG_HD void stageShadeDirect(const WaveParams &w, int depth, int cur, int pos) {
    int slot = w.qSlot[cur][pos];
    Ray ray = loadRay(w.rays[cur], pos);
    F4 hit = w.hits[pos];
    int prim = f2i(hit.x);
    F4 Lq = w.L[slot], bq = w.beta[slot];
    Spec L = spec(Lq.x, Lq.y, Lq.z), beta = spec(bq.x, bq.y, bq.z);
    float wvl0 = Lq.w;
    int flags = w.flags[slot];
    int bounces = flags & 255, vertices = (flags >> 10) & 0xffff;
    SurfInt it = hitInteraction(w.sc, prim, hit.y, hit.z, hit.w, ray);
    if (it.wvlsLost) wvl0 = 0.f;
    int mat = w.sc.primMaterial ? PG_LD(w.sc.primMaterial + prim) : -1;
    bool alive = false;
    Ray next;
    if (mat < 0) {
        // no BSDF: Li(isect.SpawnRay(ray.d), ..., depth) -- before emission is added (directlighting.cpp:77-79)
        next = spawnRay(it.p, it.pError, it.n, ray.d);
        alive = true;
    } else {
        PathSampler smp;
        smp.index = w.hIndex[slot];
        smp.dim = w.dim[slot];
        BSDF bsdf;
        makeBSDF(it, w.sc.matLobes + w.sc.nMaterials, mat, w.sc.matParams + PBRTGPU_MAT_STRIDE * (size_t)mat, wvl0, &bsdf);  // allowMultipleLobes = false
        L += beta * hitLe(w.sc, prim, it.n, -ray.d);
        const int nLights = w.sc.nLights;
        if (nLights > 0) {
            NeeRequest r;
            bool culled;
            if (w.directStrategy == PBRTGPU_DIRECT_ONE) {
                int li = (int)(get1D(w.hal, smp) * nLights);
                if (li > nLights - 1) li = nLights - 1;
                float lightPdf = 1.f / nLights;
                float uL0, uL1, uS0, uS1;
                get2D(w.hal, smp, &uL0, &uL1);
                get2D(w.hal, smp, &uS0, &uS1);
                estimateDirectRays(w.sc, li, it, bsdf, uL0, uL1, uS0, uS1, spec(1.f), &r, &culled);
                // EstimateDirect(...) / lightPdf, then the throughput
                if (r.hasShadow) r.shadowContrib = beta * (r.shadowContrib / lightPdf);
                if (r.hasClosest) r.closestContrib = beta * (r.closestContrib / lightPdf);
                pushDirectRays(w, depth, slot, r);
            } else {
                int px, py, sampleNum;
                slotToSample(w, slot, &px, &py, &sampleNum);
                for (int j = 0; j < nLights; ++j) {
                    int n = w.sc.lightNSamples ? PG_LD(w.sc.lightNSamples + j) : 1;
                    int a = 2 * (vertices * nLights + j);  // Get2DArray is called twice per light per vertex
                    if (a + 1 < w.nSampleArrays) {
                        for (int k = 0; k < n; ++k) {
                            float uL0, uL1, uS0, uS1;
                            arraySample2D(w.hal, smp.index, sampleNum, a, n, k, &uL0, &uL1);
                            arraySample2D(w.hal, smp.index, sampleNum, a + 1, n, k, &uS0, &uS1);
                            estimateDirectRays(w.sc, j, it, bsdf, uL0, uL1, uS0, uS1, spec(1.f), &r, &culled);
                            // L += Ld / nSamples with Ld the sum over k; here term by term
                            if (r.hasShadow) r.shadowContrib = beta * (r.shadowContrib / (float)n);
                            if (r.hasClosest) r.closestContrib = beta * (r.closestContrib / (float)n);
                            pushDirectRays(w, depth, slot, r);
                        }
                    } else {  // the arrays are used up (vertices behind material-less boundaries): single samples
                        float uL0, uL1, uS0, uS1;
                        get2D(w.hal, smp, &uL0, &uL1);
                        get2D(w.hal, smp, &uS0, &uS1);
                        estimateDirectRays(w.sc, j, it, bsdf, uL0, uL1, uS0, uS1, spec(1.f), &r, &culled);
                        if (r.hasShadow) r.shadowContrib = beta * r.shadowContrib;
                        if (r.hasClosest) r.closestContrib = beta * r.closestContrib;
                        pushDirectRays(w, depth, slot, r);
                    }
                }
            }
        }
        int branches = (flags >> 26) & 31;
        if (bounces + 1 < w.maxDepth) {
            V3 wo = -ray.d, wi = v3(0, 0, 0);
            float pdf, u0, u1;
            int sampled;
            get2D(w.hal, smp, &u0, &u1);
            Spec f = bsdfSampleF(bsdf, wo, &wi, u0, u1, &pdf, BSDF_REFLECTION | BSDF_SPECULAR, &sampled);
            const bool reflects = pdf > 0.f && !isBlack(f) && absDot(wi, it.ns) != 0.f;
            if (w.directTree) {
                // SpecularTransmit (integrator.cpp:677-755).  One specular-transmission lobe per material (checked at
                // setup), so its Sample_f does not depend on the sample it is given.
                V3 wiT = v3(0, 0, 0);
                float pdfT, uT0 = 0.5f, uT1 = 0.5f;
                int sampledT;
                if (!reflects) get2D(w.hal, smp, &uT0, &uT1);  // no reflection subtree in between: drawn right away
                Spec fT = bsdfSampleF(bsdf, wo, &wiT, uT0, uT1, &pdfT, BSDF_TRANSMISSION | BSDF_SPECULAR, &sampledT);
                const bool transmits = pdfT > 0.f && !isBlack(fT) && absDot(wiT, it.ns) != 0.f;
                if (reflects) {
                    if (branches < w.branchCapacity) {
                        F4 *e = w.branch + 3 * ((size_t)slot * w.branchCapacity + branches);
                        Ray rt;
                        rt.o = rt.d = v3(0, 0, 0);
                        Spec bt = spec(0.f);
                        if (transmits) {
                            rt = spawnRay(it.p, it.pError, it.n, wiT);
                            bt = beta * (fT * absDot(wiT, it.ns) / pdfT);
                        }
                        e[0] = f4(rt.o.x, rt.o.y, rt.o.z, transmits ? 1.f : 0.f);
                        e[1] = f4(rt.d.x, rt.d.y, rt.d.z, (float)(bounces + 1));
                        e[2] = f4(bt.r, bt.g, bt.b, 0.f);
                        ++branches;
                    }
                } else if (transmits) {
                    beta *= fT * absDot(wiT, it.ns) / pdfT;
                    next = spawnRay(it.p, it.pError, it.n, wiT);
                    ++bounces;
                    alive = true;
                }
            }
            if (reflects) {
                beta *= f * absDot(wi, it.ns) / pdf;
                next = spawnRay(it.p, it.pError, it.n, wi);
                ++bounces;
                alive = true;
            }
        }
        flags = (flags & ~(31 << 26)) | (branches << 26);
        w.dim[slot] = smp.dim;
    }
    w.L[slot] = f4(L.r, L.g, L.b, wvl0);
    if (alive) {
        w.beta[slot] = f4(beta.r, beta.g, beta.b, bq.w);
        // bit 8 (specularBounce) stays set: every vertex of this integrator adds emission, and so does a miss
        w.flags[slot] = (flags & ~0x3fffdff) | bounces | 256 | (((vertices + (mat >= 0 ? 1 : 0)) & 0xffff) << 10);
        int np = atomicIncInt(&w.cnt->nActive[depth + 1]);
        storeRay(w.rays[cur ^ 1], np, next);
        w.qSlot[cur ^ 1][np] = slot;
    } else if (w.directTree) {
        w.flags[slot] = (flags & ~0x3fffc00) | (((vertices + (mat >= 0 ? 1 : 0)) & 0xffff) << 10);
        popBranch(w, depth, cur, slot);
    }
}

// L[slot] += (r, g, b) from a resolve kernel.  One queue entry per slot and kernel (path integrator,
// directlighting "one"): a plain read-modify-write, nobody else touches the slot.  Several entries per
// slot (directlighting "all": one per light sample): float atomics.
PG_HD void addRadiance(const WaveParams &w, int slot, float r, float g, float b) {
    if (w.neeCapacity > w.nSlots) {
        float *Lp = (float *)&w.L[slot];
        atomicAddFloat(Lp, r);
        atomicAddFloat(Lp + 1, g);
        atomicAddFloat(Lp + 2, b);
    } else {
        F4 Lq = w.L[slot];
        Lq.x += r, Lq.y += g, Lq.z += b;
        w.L[slot] = Lq;
    }
}

// ---- stage 4a: shadow rays.  Unoccluded -> the light sample counts.
PG_HD void stageShadowResolve(const WaveParams &w, int pos, bool occluded) {
    PG_TRACE("  shadow resolve slot=%d occluded=%d c=(%.9g %.9g %.9g)\n", (int)floatToBits(w.shadowContrib[pos].w), (int)occluded, w.shadowContrib[pos].x, w.shadowContrib[pos].y, w.shadowContrib[pos].z);
    if (occluded) return;
    F4 c = w.shadowContrib[pos];
    int slot = (int)floatToBits(c.w);
    addRadiance(w, slot, c.x, c.y, c.z);
}

// ---- stage 4b: closest-hit next-event rays (BSDF-sampled MIS ray, portal rays).
PG_HD void stageNeeResolve(const WaveParams &w, int pos) {
    F4 c = w.neeContrib[pos];
    int slot = (int)floatToBits(c.w);
    int light = w.neeInfo[2 * (size_t)pos], mode = w.neeInfo[2 * (size_t)pos + 1];
    Ray ray = loadRay(w.neeRays, pos);
    F4 hit = w.neeHits[pos];
    int prim = f2i(hit.x);
    Spec Li = spec(0.f);
    if (prim >= 0) {
        int hitLight = w.sc.primLight ? PG_LD(w.sc.primLight + prim) : -1;
        if (hitLight >= 0 && (mode != NEE_MIS_LIGHT || hitLight == light)) {
            SurfInt li = hitInteraction(w.sc, prim, hit.y, hit.z, hit.w, ray);
            Li = areaLightL(w.sc, hitLight, li.n, -ray.d);
        }
    } else if (mode == NEE_MIS_LIGHT) {
        if (PG_LD(w.sc.lightTag + light) == PBRTGPU_LIGHT_INFINITE)
            Li = infiniteLe(w.sc.lightParams + PBRTGPU_LIGHT_STRIDE * (size_t)light, ray.d);
    } else if (mode == NEE_ANY_OR_LIGHT)
        Li = loadSpec(w.sc.lightParams + PBRTGPU_LIGHT_STRIDE * (size_t)light);
    PG_TRACE("  nee resolve slot=%d prim=%d Li=(%g %g %g) b=(%g %g %g)\n", slot, prim, Li.r, Li.g, Li.b, hit.y, hit.z, hit.w);
    if (isBlack(Li)) return;
    addRadiance(w, slot, c.x * Li.r, c.y * Li.g, c.z * Li.b);
}

// ---- stage 5: scrub (integrator.cpp:592-613) + FilmTile::AddSample with atomics
PG_HD void stageFilm(const WaveParams &w, int slot) {
    if (!(w.flags[slot] & (1 << 9))) return;
    F4 Lq = w.L[slot];
    Spec L = spec(Lq.x, Lq.y, Lq.z);
    bool bad = hasNaNs(L);
    if (!bad && luminance(L) < -1e-5) bad = true;
    if (!bad && isInfF(luminance(L))) bad = true;
    if (bad) {
        L = spec(0.f);
        atomicAddU64(&w.cnt->scrubbed, 1ull);
    }
    const FilmParams &fp = w.film;
    if (luminance(L) > fp.maxSampleLuminance) L *= fp.maxSampleLuminance / luminance(L);
    float dx = w.pFilm[2 * (size_t)slot] - 0.5f, dy = w.pFilm[2 * (size_t)slot + 1] - 0.5f;
    int p0x = (int)ceilf(dx - fp.radius[0]), p0y = (int)ceilf(dy - fp.radius[1]);
    int p1x = (int)floorf(dx + fp.radius[0]) + 1, p1y = (int)floorf(dy + fp.radius[1]) + 1;
    p0x = p0x > fp.cropped[0] ? p0x : fp.cropped[0];
    p0y = p0y > fp.cropped[1] ? p0y : fp.cropped[1];
    p1x = p1x < fp.cropped[2] ? p1x : fp.cropped[2];
    p1y = p1y < fp.cropped[3] ? p1y : fp.cropped[3];
    int width = fp.cropped[2] - fp.cropped[0];
    for (int y = p0y; y < p1y; ++y) {
        float fy = fabsf((y - dy) * fp.invRadius[1] * 16);
        int iy = (int)floorf(fy);
        iy = iy < 15 ? iy : 15;
        for (int x = p0x; x < p1x; ++x) {
            float fx = fabsf((x - dx) * fp.invRadius[0] * 16);
            int ix = (int)floorf(fx);
            ix = ix < 15 ? ix : 15;
            float fw = PG_LD(fp.filterTable + iy * 16 + ix);
            float *px = fp.rgbw + 4 * ((size_t)(y - fp.cropped[1]) * width + (x - fp.cropped[0]));
            // pixel.contribSum += L * sampleWeight * filterWeight; sampleWeight (the camera ray weight) is 1
            Spec c = L * 1.f * fw;
            atomicAddFloat4(px, c.r, c.g, c.b, fw);
        }
    }
}

}  // namespace pg
#endif
