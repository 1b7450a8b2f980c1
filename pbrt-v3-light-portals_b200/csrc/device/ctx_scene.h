// The scene as the kernels see it (device pointers; host pointers in tests/hostsim).
#ifndef PG_DEVICE_CTX_SCENE_H
#define PG_DEVICE_CTX_SCENE_H

#include "bsdf.h"
#include "bvh.h"

namespace pg {

// Everything a kernel needs to know about the scene; passed by value.
struct DevScene {
    DevBVH bvh;
    int nPrims, nLights, nMaterials;
    const int *primOrigId;
    const int *primMaterial;
    const int *primLight;
    const unsigned char *primFlags;
    const float *triN;   // 9 per slot or null
    const float *triUV;  // 6 per slot or null
    const float *triS;   // 9 per slot or null
    const int *matTag;
    const float *matParams;
    const MatLobes *matLobes;  // nMaterials entries, built at upload
    const int *lightTag;
    const int *lightPrim;
    const int *lightStrategy;
    const int *lightPortalBegin;
    const int *lightNSamples;  // Light::nSamples, or null = 1 each (directlighting "all" only)
    const float *lightParams;
    const float *portalParams;
    const float *lightCdf;  // nLights+1 (power strategy) or null
    float lightFuncInt;
    const float *lightFunc;
    float worldCenter[3], worldRadius;
};

}  // namespace pg
#endif
