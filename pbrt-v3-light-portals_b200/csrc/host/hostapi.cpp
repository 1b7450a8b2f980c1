// extern "C" surface of libpbrthost.so (include/pbrthost.h).
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/pbrthost.h"
#include "../common/halton_tables.h"
#include "api.h"
#include "setup.h"

struct pbrthost_scene {
    pg::SceneDesc desc;
    pg::FlatScene flat;
    pg::RenderSetup setup;
    pbrtgpu_scene view;
};

static int finish(std::unique_ptr<pbrthost_scene> s, bool parsed, pbrthost_scene **out) {
    if (!parsed || !s->desc.worldEnded) {
        if (parsed) pg::hostError("scene description has no WorldEnd");
        return PBRTGPU_ERR_INVALID;
    }
    std::string err;
    if (!pg::flattenScene(s->desc, &s->flat, &err)) {
        pg::hostError("%s", err.c_str());
        return PBRTGPU_ERR_INVALID;
    }
    if (!pg::makeRenderSetup(s->desc, &s->setup)) return PBRTGPU_ERR_INVALID;
    s->view = s->flat.view();
    *out = s.release();
    return PBRTGPU_OK;
}

extern "C" {

int pbrthost_load_file(const char *path, int quiet, pbrthost_scene **out) {
    if (!path || !out) return PBRTGPU_ERR_INVALID;
    pg::setHostQuiet(quiet != 0);
    auto s = std::make_unique<pbrthost_scene>();
    pg::SceneApi api(&s->desc);
    bool ok = pg::parseFile(path, api, s->desc);
    return finish(std::move(s), ok, out);
}

int pbrthost_load_string(const char *text, const char *base_dir, int quiet, pbrthost_scene **out) {
    if (!text || !out) return PBRTGPU_ERR_INVALID;
    pg::setHostQuiet(quiet != 0);
    auto s = std::make_unique<pbrthost_scene>();
    s->desc.sceneDir = base_dir ? base_dir : "";
    pg::SceneApi api(&s->desc);
    bool ok = pg::parseString(text, api, s->desc);
    return finish(std::move(s), ok, out);
}

void pbrthost_free(pbrthost_scene *s) { delete s; }

const pbrtgpu_scene *pbrthost_scene_view(const pbrthost_scene *s) { return s ? &s->view : nullptr; }
const pbrtgpu_render_params *pbrthost_render_params(const pbrthost_scene *s) { return s ? &s->setup.params : nullptr; }
const char *pbrthost_output_filename(const pbrthost_scene *s) { return s ? s->setup.filename.c_str() : nullptr; }
const char *pbrthost_integrator_name(const pbrthost_scene *s) { return s ? s->desc.integratorName.c_str() : nullptr; }

void pbrthost_set_spp(pbrthost_scene *s, int spp) {
    if (s && spp > 0) s->setup.params.spp = spp;
}

int pbrthost_film_to_rgb(const pbrthost_scene *s, const float *film_xyzw, float *rgb) {
    if (!s || !film_xyzw || !rgb) return PBRTGPU_ERR_INVALID;
    const int *cb = s->setup.params.cropped_bounds;
    pg::filmToRGB(film_xyzw, (cb[2] - cb[0]) * (cb[3] - cb[1]), s->setup.scale, rgb);
    return PBRTGPU_OK;
}

int pbrthost_write_image(const pbrthost_scene *s, const char *filename, const float *rgb) {
    if (!s || !filename || !rgb) return PBRTGPU_ERR_INVALID;
    return pg::writeImage(filename, rgb, s->setup.params.cropped_bounds, s->setup.params.full_resolution)
               ? PBRTGPU_OK
               : PBRTGPU_ERR_INVALID;
}

void pbrthost_set_hlbvh_builder(pbrthost_hlbvh_builder fn, void *user) { pg::setHlbvhBuilder(fn, user); }
void pbrthost_last_bvh_build_ms(double *wall_ms, float *builder_ms) { pg::lastBvhBuildMs(wall_ms, builder_ms); }

int pbrthost_halton_permutations(uint16_t *out, int64_t n) {
    std::vector<uint16_t> perms = pg::radicalInversePermutations(pg::firstPrimes());
    if (!out || n != (int64_t)perms.size()) return PBRTGPU_ERR_INVALID;
    memcpy(out, perms.data(), perms.size() * sizeof(uint16_t));
    return PBRTGPU_OK;
}

}  // extern "C"
