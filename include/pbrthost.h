/* pbrthost.h -- C ABI of libpbrthost.so, the host front end that stays
 * reference-shaped: pbrt's scene-file grammar, plugin names and parameters,
 * BVHAccel construction and Film output.  It owns no GPU state; it produces
 * the flat arrays and render parameters that pbrtgpu.h consumes, and turns
 * the film that comes back into an image file.
 *
 * Replaces, on the reference side: pbrtParseFile (src/core/parser.cpp:1094),
 * RenderOptions::MakeScene / MakeIntegrator (src/core/api.cpp:1763-1832),
 * Film::WriteImage (src/core/film.cpp:169-211).
 */
#ifndef PBRTHOST_H
#define PBRTHOST_H

#include "pbrtgpu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pbrthost_scene pbrthost_scene; /* opaque */

/* Parse a .pbrt file (or text; base_dir resolves Include / plymesh paths),
 * build the BVH the way BVHAccel does and flatten.  Returns 0 or a negative
 * pbrtgpu_status; details go to stderr unless quiet. */
int pbrthost_load_file(const char *path, int quiet, pbrthost_scene **out);
int pbrthost_load_string(const char *text, const char *base_dir, int quiet, pbrthost_scene **out);
void pbrthost_free(pbrthost_scene *s);

/* Borrowed views, valid until pbrthost_free. */
const pbrtgpu_scene *pbrthost_scene_view(const pbrthost_scene *s);
const pbrtgpu_render_params *pbrthost_render_params(const pbrthost_scene *s);
const char *pbrthost_output_filename(const pbrthost_scene *s); /* Film "filename" */
const char *pbrthost_integrator_name(const pbrthost_scene *s);

/* Overrides the CLI offers: --outfile, and Sampler "pixelsamples". */
void pbrthost_set_spp(pbrthost_scene *s, int spp);

/* Film::WriteImage: film_xyzw (cropped W*H*4: sum X,Y,Z, sum w) -> rgb
 * (W*H*3), then write by extension (.exr half RGB, .pfm float32). */
int pbrthost_film_to_rgb(const pbrthost_scene *s, const float *film_xyzw, float *rgb);
int pbrthost_write_image(const pbrthost_scene *s, const char *filename, const float *rgb);

/* Accelerator "bvh" "string splitmethod" "hlbvh" can be built by someone else: a function with the signature of
 * pbrtgpu_build_hlbvh (pbrtgpu.h; `user` = its ctx) replaces the host's HLBVHBuild for every scene loaded afterwards.  The library
 * itself stays free of GPU state -- an embedding application connects the two:
 *     pbrthost_set_hlbvh_builder((pbrthost_hlbvh_builder)pbrtgpu_build_hlbvh, ctx);
 * fn == NULL restores the host build.  A builder that fails makes the load fail (no silent fallback).  The other split
 * methods are always built on the host. */
typedef int (*pbrthost_hlbvh_builder)(void *user, const float *prim_bounds, int32_t n, int32_t max_prims_in_node, int32_t node_capacity,
                                      float *bvh_bounds, int32_t *bvh_offset, uint16_t *bvh_nprims, uint8_t *bvh_axis, int32_t *ordered_prims,
                                      int32_t *n_nodes_out, float *elapsed_ms);
void pbrthost_set_hlbvh_builder(pbrthost_hlbvh_builder fn, void *user);
/* Wall-clock milliseconds the last load spent building its BVH, and the builder's own device time (0 for a host build). */
void pbrthost_last_bvh_build_ms(double *wall_ms, float *builder_ms);

/* Host-side pieces exposed for parity tests. */
int pbrthost_halton_permutations(uint16_t *out, int64_t n); /* ComputeRadicalInversePermutations, lowdiscrepancy.cpp:2490-2504; n = 3682913 */

#ifdef __cplusplus
}
#endif
#endif /* PBRTHOST_H */
