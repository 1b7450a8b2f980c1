// pbrt_gpu: command-line front end, the counterpart of the reference's
// main/pbrt.cpp:47-174 for scenes whose Integrator is "path".  Everything goes
// through the two C ABIs (pbrthost.h, pbrtgpu.h), as an embedding application
// would; there is no CPU rendering path here.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pbrtgpu.h"
#include "pbrthost.h"

static void usage(const char *msg = nullptr) {
    if (msg) fprintf(stderr, "pbrt_gpu: %s\n\n", msg);
    fprintf(stderr,
            "usage: pbrt_gpu [<options>] <filename.pbrt...>\n"
            "Rendering options:\n"
            "  --help               Print this help text.\n"
            "  --gpu <num>          CUDA device to render on. (Default: 0)\n"
            "  --outfile <filename> Write the final image to the given filename.\n"
            "  --spp <num>          Override the sampler's pixelsamples.\n"
            "  --quick              Render with 1/4 of the pixel samples (at least 1).\n"
            "  --quiet              Suppress all text output other than error messages.\n"
            "  --stats              Print ray counts and the device time after rendering.\n");
    exit(msg ? 1 : 0);
}

int main(int argc, char *argv[]) {
    int device = 0, spp = 0;
    bool quiet = false, quick = false, printStats = false;
    std::string outfile;
    std::vector<std::string> filenames;
    for (int i = 1; i < argc; ++i) {
        auto is = [&](const char *name) { return !strcmp(argv[i], name) || !strcmp(argv[i], name + 1); };  // --opt or -opt
        if (is("--gpu")) {
            if (i + 1 == argc) usage("missing value after --gpu argument");
            device = atoi(argv[++i]);
        } else if (is("--outfile")) {
            if (i + 1 == argc) usage("missing value after --outfile argument");
            outfile = argv[++i];
        } else if (is("--spp")) {
            if (i + 1 == argc) usage("missing value after --spp argument");
            spp = atoi(argv[++i]);
        } else if (is("--quick")) quick = true;
        else if (is("--quiet")) quiet = true;
        else if (is("--stats")) printStats = true;
        else if (is("--help") || !strcmp(argv[i], "-h")) usage();
        else if (argv[i][0] == '-') usage((std::string("unknown option ") + argv[i]).c_str());
        else filenames.push_back(argv[i]);
    }
    if (filenames.empty()) usage("no scene file given");

    pbrtgpu_ctx *ctx = nullptr;
    if (pbrtgpu_create(device, &ctx) != PBRTGPU_OK) {
        fprintf(stderr, "pbrt_gpu: %s\n", pbrtgpu_last_error(nullptr));
        return 1;
    }
    int rc = 0;
    for (const std::string &f : filenames) {
        pbrthost_scene *hs = nullptr;
        if (pbrthost_load_file(f.c_str(), quiet ? 1 : 0, &hs) != 0 || !hs) {
            fprintf(stderr, "pbrt_gpu: could not load \"%s\"\n", f.c_str());
            rc = 1;
            continue;
        }
        const pbrtgpu_render_params *rp = pbrthost_render_params(hs);
        if (spp > 0) pbrthost_set_spp(hs, spp);
        else if (quick) pbrthost_set_spp(hs, rp->spp / 4 > 0 ? rp->spp / 4 : 1);
        rp = pbrthost_render_params(hs);
        int w = rp->cropped_bounds[2] - rp->cropped_bounds[0], h = rp->cropped_bounds[3] - rp->cropped_bounds[1];
        std::vector<float> film(4 * (size_t)w * h), rgb(3 * (size_t)w * h);
        pbrtgpu_stats st;
        memset(&st, 0, sizeof(st));
        auto t0 = std::chrono::steady_clock::now();
        if (pbrtgpu_upload_scene(ctx, pbrthost_scene_view(hs)) != PBRTGPU_OK || pbrtgpu_render(ctx, rp, film.data(), &st) != PBRTGPU_OK) {
            fprintf(stderr, "pbrt_gpu: %s\n", pbrtgpu_last_error(ctx));
            pbrthost_free(hs);
            rc = 1;
            continue;
        }
        double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        pbrthost_film_to_rgb(hs, film.data(), rgb.data());
        const char *name = outfile.empty() ? pbrthost_output_filename(hs) : outfile.c_str();
        if (pbrthost_write_image(hs, name, rgb.data()) != 0) {
            fprintf(stderr, "pbrt_gpu: could not write \"%s\"\n", name);
            rc = 1;
        }
        if (printStats && !quiet) {
            double samples = (double)st.camera_rays;
            printf("%s: %d x %d, %d spp -> %s\n", f.c_str(), w, h, rp->spp, name);
            printf("  camera rays %lld, closest-hit rays %lld, shadow rays %lld, scrubbed samples %lld\n", (long long)st.camera_rays,
                   (long long)st.closest_rays, (long long)st.shadow_rays, (long long)st.scrubbed);
            printf("  device %.1f ms (%.1f Msamples/s), upload + render + readback %.1f ms\n", st.device_ms,
                   st.device_ms > 0 ? samples / st.device_ms / 1e3 : 0.0, wall * 1e3);
        }
        pbrthost_free(hs);
    }
    pbrtgpu_destroy(ctx);
    return rc;
}
