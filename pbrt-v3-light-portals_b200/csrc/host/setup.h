// Render-time setup derived from a SceneDesc (see setup.cpp).
#ifndef PG_HOST_SETUP_H
#define PG_HOST_SETUP_H

#include <string>

#include "../../../include/pbrthost.h"
#include "scene.h"

namespace pg {

struct RenderSetup {
    pbrtgpu_render_params params;
    std::string filename;  // Film "filename"
    float scale = 1.f;     // Film "scale"
};

bool makeRenderSetup(const SceneDesc &d, RenderSetup *out);
void filmToRGB(const float *xyzw, int nPixels, float scale, float *rgb);

// imageio.cpp: WriteImage (src/core/imageio.cpp:81-122); .pfm float32, .exr half RGB
bool writeImage(const std::string &name, const float *rgb, const int croppedBounds[4], const int fullRes[2]);
bool readPFM(const std::string &name, std::vector<float> *rgb, int *w, int *h);

void setHlbvhBuilder(pbrthost_hlbvh_builder fn, void *user);  // setup.cpp: who builds "splitmethod" "hlbvh" trees
void lastBvhBuildMs(double *wallMs, float *builderMs);

void setHostQuiet(bool q);
int hostErrorCount();

}  // namespace pg
#endif
