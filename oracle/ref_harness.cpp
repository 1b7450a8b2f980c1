// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// ref_harness: drives the REFERENCE's own scene construction and BVH code
// (compiled in place from /root/reference by oracle/Makefile) to produce
// ground truth for the parity tests:
//
//   ref_harness dump      scene.pbrt out_prefix
//       -> out_prefix.nodes.bin  raw LinearBVHNode[] (32 B each, bvh.cpp:95-104)
//          out_prefix.prims.bin  per BVH-ordered primitive: int32 origId,
//                                int32 type (0 tri, 1 aaplane, 2 sphere, 3 other),
//                                float[9] triangle vertices (world space)
//          out_prefix.lights.txt one line per scene.lights entry
//   ref_harness intersect scene.pbrt rays.bin out.bin closest|any [nthreads]
//       rays.bin: n x 8 float32 {ox,oy,oz,tMax,dx,dy,dz,pad}
//       closest -> out.bin: n x {int32 origId (-1 miss), float32 t}
//       any     -> out.bin: n x uint8
//       prints one JSON line with the Mrays/s of BVHAccel::Intersect(P) run
//       inside the reference's ParallelFor.
//   ref_harness halton    scene.pbrt queries.bin out.bin
//       the scene's own Sampler "halton" (built by MakeSampler over the film's sample bounds, halton.cpp:65-139).
//       queries.bin: n x {int32 x, int32 y, int32 dim, int32 count, int64 sampleNum}.
//       count == 0 -> one float: HaltonSampler::SampleDimension(GetIndexForSample(sampleNum), dim) at pixel (x, y);
//       count  > 0 -> `count` floats: StartPixel, SetSampleNumber(sampleNum), then Get1D() `count` times (`dim` unused).
//
// origId is the index of the primitive in RenderOptions::primitives BEFORE
// the BVH reorders them (api.cpp:1380-1436), i.e. creation order.
//
// The reference keeps RenderOptions, LinearBVHNode and BVHAccel's arrays
// private to their .cpp files, so this file textually includes those two
// translation units (they are compiled from where they lie; nothing is
// copied) and substitutes its own pbrtWorldEnd for the reference's.
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#define private public
#define protected public
#define pbrtWorldEnd pbrtWorldEnd_reference
#include "core/api.cpp"
#include "accelerators/bvh.cpp"
#undef pbrtWorldEnd
#undef private
#undef protected

#include "shapes/plane.h"
#include "shapes/sphere.h"
#include "shapes/triangle.h"

namespace {
std::string gCommand, gOutA, gRays, gMode;
int gThreads = 1;
}  // namespace

namespace pbrt {

static int shapeType(const Shape *s) {
    if (dynamic_cast<const Triangle *>(s)) return 0;
    if (dynamic_cast<const AAPlaneShape *>(s)) return 1;
    if (dynamic_cast<const Sphere *>(s)) return 2;
    return 3;
}

static void harnessDump(const BVHAccel *bvh,
                        const std::map<const Primitive *, int> &origId,
                        const Scene &scene, int totalNodes) {
    {
        FILE *f = fopen((gOutA + ".nodes.bin").c_str(), "wb");
        fwrite(bvh->nodes, sizeof(LinearBVHNode), totalNodes, f);
        fclose(f);
    }
    {
        FILE *f = fopen((gOutA + ".prims.bin").c_str(), "wb");
        for (const auto &p : bvh->primitives) {
            int32_t hdr[2] = {origId.at(p.get()), 3};
            float v[9] = {0};
            auto gp = dynamic_cast<const GeometricPrimitive *>(p.get());
            if (gp) {
                hdr[1] = shapeType(gp->shape.get());
                if (auto tri = dynamic_cast<const Triangle *>(gp->shape.get())) {
                    for (int k = 0; k < 3; ++k) {
                        const Point3f &q = tri->mesh->p[tri->v[k]];
                        v[3 * k] = q.x, v[3 * k + 1] = q.y, v[3 * k + 2] = q.z;
                    }
                }
            }
            fwrite(hdr, sizeof(hdr), 1, f);
            fwrite(v, sizeof(v), 1, f);
        }
        fclose(f);
    }
    {
        std::ofstream f(gOutA + ".lights.txt");
        for (const auto &l : scene.lights) {
            const Light *lp = l.get();
            const char *kind = "other";
            if (dynamic_cast<const PortalArealight *>(lp)) kind = "portal";
            else if (dynamic_cast<const DiffuseAreaLight *>(lp)) kind = "diffuse";
            else if (dynamic_cast<const InfiniteAreaLight *>(lp)) kind = "infinite";
            else if (dynamic_cast<const PointLight *>(lp)) kind = "point";
            Spectrum pw = lp->Power();
            Float rgb[3];
            pw.ToRGB(rgb);
            f << kind << " power " << rgb[0] << " " << rgb[1] << " " << rgb[2]
              << "\n";
        }
    }
}

static int countNodes(const BVHAccel *bvh) {
    // nodes are DFS-ordered; walk to find the total count.
    int maxIdx = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int i = stack.back();
        stack.pop_back();
        maxIdx = std::max(maxIdx, i);
        const LinearBVHNode &n = bvh->nodes[i];
        if (n.nPrimitives == 0) {
            stack.push_back(i + 1);
            stack.push_back(n.secondChildOffset);
        }
    }
    return maxIdx + 1;
}

static void harnessIntersect(const Scene &scene,
                             const std::map<const Primitive *, int> &origId) {
    std::vector<float> rays;
    {
        FILE *f = fopen(gRays.c_str(), "rb");
        if (!f) { fprintf(stderr, "cannot open %s\n", gRays.c_str()); exit(1); }
        fseek(f, 0, SEEK_END);
        long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        rays.resize(sz / sizeof(float));
        if (fread(rays.data(), 1, sz, f) != (size_t)sz) exit(1);
        fclose(f);
    }
    int64_t n = rays.size() / 8;
    bool any = gMode == "any";
    std::vector<int32_t> ids(any ? 0 : n);
    std::vector<float> ts(any ? 0 : n);
    std::vector<uint8_t> occ(any ? n : 0);
    auto t0 = std::chrono::steady_clock::now();
    ParallelFor([&](int64_t i) {
        const float *r = &rays[8 * i];
        Ray ray(Point3f(r[0], r[1], r[2]), Vector3f(r[4], r[5], r[6]),
                Vector4f(0.f), r[3]);
        if (any) {
            occ[i] = scene.IntersectP(ray) ? 1 : 0;
        } else {
            SurfaceInteraction isect;
            if (scene.Intersect(ray, &isect)) {
                ids[i] = origId.at(isect.primitive);
                ts[i] = ray.tMax;
            } else {
                ids[i] = -1;
                ts[i] = ray.tMax;
            }
        }
    }, n, 4096);
    auto t1 = std::chrono::steady_clock::now();
    double sec = std::chrono::duration<double>(t1 - t0).count();
    FILE *f = fopen(gOutA.c_str(), "wb");
    if (any) fwrite(occ.data(), 1, n, f);
    else
        for (int64_t i = 0; i < n; ++i) {
            fwrite(&ids[i], 4, 1, f);
            fwrite(&ts[i], 4, 1, f);
        }
    fclose(f);
    printf("{\"rays\": %lld, \"seconds\": %.6f, \"mrays_per_s\": %.4f, "
           "\"threads\": %d, \"mode\": \"%s\"}\n",
           (long long)n, sec, n / sec / 1e6, gThreads, gMode.c_str());
}

static void harnessHalton() {
    std::unique_ptr<Camera> camera(renderOptions->MakeCamera());
    std::shared_ptr<Sampler> sampler = MakeSampler(renderOptions->SamplerName, renderOptions->SamplerParams, camera->film);
    HaltonSampler *hs = dynamic_cast<HaltonSampler *>(sampler.get());
    if (!hs) { fprintf(stderr, "the scene's sampler is not \"halton\"\n"); exit(1); }
    struct Query { int32_t x, y, dim, count; int64_t num; };
    std::vector<Query> q;
    {
        FILE *f = fopen(gRays.c_str(), "rb");
        if (!f) { fprintf(stderr, "cannot open %s\n", gRays.c_str()); exit(1); }
        Query one;
        while (fread(&one, sizeof(one), 1, f) == 1) q.push_back(one);
        fclose(f);
    }
    FILE *f = fopen(gOutA.c_str(), "wb");
    for (const Query &k : q) {
        hs->StartPixel(Point2i(k.x, k.y));
        hs->SetSampleNumber(k.num);
        if (k.count == 0) {
            float v = hs->SampleDimension(hs->GetIndexForSample(k.num), k.dim);
            fwrite(&v, 4, 1, f);
        } else
            for (int i = 0; i < k.count; ++i) {
                float v = hs->Get1D();
                fwrite(&v, 4, 1, f);
            }
    }
    fclose(f);
}

void pbrtWorldEnd() {
    if (gCommand == "halton") {
        harnessHalton();
        fflush(stdout);
        _exit(0);
    }
    std::map<const Primitive *, int> origId;
    for (size_t i = 0; i < renderOptions->primitives.size(); ++i)
        origId[renderOptions->primitives[i].get()] = (int)i;
    std::unique_ptr<Integrator> integrator(renderOptions->MakeIntegrator());
    std::unique_ptr<Scene> scene(renderOptions->MakeScene());
    const BVHAccel *bvh =
        dynamic_cast<const BVHAccel *>(scene->aggregate.get());
    if (!bvh) { fprintf(stderr, "aggregate is not a BVHAccel\n"); exit(1); }
    if (gCommand == "dump")
        harnessDump(bvh, origId, *scene, countNodes(bvh));
    else if (gCommand == "intersect")
        harnessIntersect(*scene, origId);
    fflush(stdout);
    _exit(0);
}

}  // namespace pbrt

int main(int argc, char **argv) {
    if (argc < 4) {
        fprintf(stderr,
                "usage: ref_harness dump scene.pbrt out_prefix\n"
                "       ref_harness intersect scene.pbrt rays.bin out.bin "
                "closest|any [nthreads]\n"
                "       ref_harness halton scene.pbrt queries.bin out.bin\n");
        return 2;
    }
    gCommand = argv[1];
    std::string scene = argv[2];
    pbrt::Options options;
    options.quiet = true;
    if (gCommand == "dump") {
        gOutA = argv[3];
    } else if (gCommand == "halton" && argc >= 5) {
        gRays = argv[3];
        gOutA = argv[4];
    } else if (gCommand == "intersect" && argc >= 6) {
        gRays = argv[3];
        gOutA = argv[4];
        gMode = argv[5];
        if (argc >= 7) gThreads = atoi(argv[6]);
    } else {
        return 2;
    }
    options.nThreads = gThreads;
    FLAGS_minloglevel = 2;
    pbrt::pbrtInit(options);
    pbrt::pbrtParseFile(scene);
    pbrt::pbrtCleanup();
    return 0;
}
