// host/gpu_path_integrator.cpp -- Integrator "path" on the B200.
//
// Product host code.  Implements the reference's Integrator interface
// (src/core/integrator.h:53-58: `virtual void Render(const Scene &)`) on top of
// the C ABI in include/sptx.h: flatten the parsed scene, upload, render every
// sample on the device, read the spectral film back and hand it to the
// reference's own Film::WriteImage (src/core/film.cpp:180-310), so the .dat /
// .exr writers and everything before WorldEnd stay the reference's host C++.
//
// Binding.  The reference resolves integrator names in a static if-chain
// (RenderOptions::MakeIntegrator, src/core/api.cpp:1750-1771: name == "path"
// -> CreatePathIntegrator).  A maintainer adds one branch there
// (INTEGRATION.md).  This in-place build does not edit the reference: it
// defines pbrt::CreatePathIntegrator itself and links ahead of the archive, so
// the CPU PathIntegrator object is never pulled in.  There is no CPU fallback:
// if the device path fails, Render reports the error and aborts.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <glog/logging.h>

#define private public
#define protected public
#include "camera.h"
#include "film.h"
#include "integrator.h"
#include "integrators/metadata.h"
#include "integrators/path.h"
#include "integrators/spectralpath.h"
#include "paramset.h"
#include "sampler.h"
#include "scene.h"
#include "spectrum.h"
#undef private
#undef protected

#include "error.h"
#include "flatten.h"
#include "sptx.h"

namespace pbrt {

class GpuPathIntegrator : public Integrator {
  public:
    GpuPathIntegrator(const sptx_host::PathParams &params,
                      std::shared_ptr<const Camera> camera,
                      std::shared_ptr<Sampler> sampler)
        : params(params), camera(camera), sampler(sampler) {}
    void Render(const Scene &scene) override;

  private:
    sptx_host::PathParams params;
    std::shared_ptr<const Camera> camera;
    std::shared_ptr<Sampler> sampler;
};

// Leaves without running static destructors: pbrt's worker threads are still
// parked on their condition variable, and std::exit() would wait for them.
[[noreturn]] static void ExitNow() {
    fflush(nullptr);
    std::_Exit(1);
}

static void Die(sptx_ctx *ctx, const char *what) {
    // No CPU fallback (north star): a failed device path is fatal, reported
    // the way the reference reports unrecoverable errors.
    Error("GpuPathIntegrator: %s failed: %s", what, sptx_last_error(ctx));
    ExitNow();
}

void GpuPathIntegrator::Render(const Scene &scene) {
    sptx_host::FlatScene flat;
    std::string err;
    if (!sptx_host::FlattenScene(scene, *camera, *sampler, params, &flat,
                                 &err)) {
        Error("GpuPathIntegrator: scene is outside the GPU hot path: %s",
              err.c_str());
        ExitNow();
    }
    if (const char *dump = std::getenv("SPTX_DUMP_SCENE")) {
        if (sptx_scene_save(dump, &flat.desc) != 0) Die(nullptr, "scene save");
        if (std::getenv("SPTX_DUMP_ONLY")) return;
    }

    // One context per GPU.  SPTX_DEVICES=0,1,... (or SPTX_DEVICE=n for one) names them; every
    // context gets the whole scene and a contiguous block of the sample numbers of every pixel
    // (HaltonSampler::GetIndexForSample is a pure function of the sample number), each driven
    // by its own host thread -- the analogue of ParallelFor2D over tiles
    // (src/core/integrator.cpp:235-241) -- and the per-GPU films are summed into the first one
    // through peer access (Film::MergeFilmTile, src/core/film.cpp:124-142).
    std::vector<int> devices;
    if (const char *ds = std::getenv("SPTX_DEVICES")) {
        for (const char *p = ds; *p;) {
            char *end = nullptr;
            long v = std::strtol(p, &end, 10);
            if (end == p) break;
            devices.push_back((int)v);
            p = (*end == ',') ? end + 1 : end;
        }
    }
    if (devices.empty()) {
        int device = 0;
        if (const char *d = std::getenv("SPTX_DEVICE")) device = std::atoi(d);
        devices.push_back(device);
    }
    const int spp = (int)sampler->samplesPerPixel;
    if ((int)devices.size() > 8 || (int)devices.size() > std::max(spp, 1)) {
        Error("GpuPathIntegrator: SPTX_DEVICES names %d contexts (at most 8, and at most one per sample)",
              (int)devices.size());
        ExitNow();
    }
    const int nCtx = (int)devices.size();
    if (nCtx > 1 && (flat.desc.sampler.type == SPTX_SAMPLER_RANDOM ||
                     flat.desc.sampler.type == SPTX_SAMPLER_ZEROTWO)) {
        Error("GpuPathIntegrator: a per-tile RNG stream (random, 02sequence) cannot be cut by sample number; "
              "use one device");
        ExitNow();
    }
    std::vector<sptx_ctx *> ctxs((size_t)nCtx, nullptr);
    for (int k = 0; k < nCtx; ++k) {
        if (sptx_create(devices[k], &ctxs[k]) != 0) Die(nullptr, "sptx_create");
        if (sptx_upload_scene(ctxs[k], &flat.desc) != 0) Die(ctxs[k], "sptx_upload_scene");
        if (sptx_film_clear(ctxs[k]) != 0) Die(ctxs[k], "sptx_film_clear");
    }
    sptx_ctx *ctx = ctxs[0];
    auto t0 = std::chrono::steady_clock::now();
    {
        std::vector<std::thread> workers;
        std::vector<int> rcs((size_t)nCtx, 0);
        const int base = spp / nCtx, extra = spp % nCtx;
        int s0 = 0;
        for (int k = 0; k < nCtx; ++k) {
            const int s1 = s0 + base + (k < extra ? 1 : 0);
            workers.emplace_back([&, k, s0, s1] { rcs[k] = sptx_render(ctxs[k], nullptr, s0, s1); });
            s0 = s1;
        }
        for (auto &w : workers) w.join();
        for (int k = 0; k < nCtx; ++k)
            if (rcs[k] != 0) Die(ctxs[k], "sptx_render");
    }
    if (nCtx > 1 && sptx_film_merge_contexts(ctxs.data(), nCtx, 0) != 0) Die(ctx, "sptx_film_merge_contexts");
    auto t1 = std::chrono::steady_clock::now();

    // Film pixels back into the reference's Film, then its own writer.
    Film *film = camera->film;
    Bounds2i cb = film->croppedPixelBounds;
    int w = cb.pMax.x - cb.pMin.x, h = cb.pMax.y - cb.pMin.y;
    size_t nPix = (size_t)std::max(0, w) * std::max(0, h);
    std::vector<float> contrib(nPix * SPTX_NBINS), weight(nPix);
    if (sptx_film_download(ctx, contrib.data(), weight.data()) != 0)
        Die(ctx, "sptx_film_download");
    for (size_t i = 0; i < nPix; ++i) {
        Film::Pixel &px = film->pixels[i];
        Spectrum L;
        for (int b = 0; b < SPTX_NBINS; ++b) L[b] = contrib[i * SPTX_NBINS + b];
        // MergeFilmTile (src/core/film.cpp:124-142): XYZ for the RGB writer,
        // the raw spectrum for the spectral one.
        Float xyz[3];
        L.ToXYZ(xyz);
        for (int c = 0; c < 3; ++c) px.xyz[c] += xyz[c];
        px.filterWeightSum += weight[i];
        if (film->spectralFlag) px.L += L;
    }
    sptx_counters c;
    std::memset(&c, 0, sizeof(c));
    for (int k = 0; k < nCtx; ++k) {   // rays summed over the contexts, device time = the slowest
        sptx_counters ck;
        std::memset(&ck, 0, sizeof(ck));
        sptx_get_counters(ctxs[k], &ck);
        c.camera_rays += ck.camera_rays;
        c.closest_hit_rays += ck.closest_hit_rays;
        c.shadow_rays += ck.shadow_rays;
        c.render_ms = std::max(c.render_ms, ck.render_ms);
    }
    double wall = std::chrono::duration<double>(t1 - t0).count();
    double nSamples = (double)c.camera_rays;
    fprintf(stderr,
            "GpuPathIntegrator: %d GPU context(s), %.0f camera samples, %.3f s device "
            "(%.2f Msamples/s), %.3f s wall; rays: %llu closest-hit, %llu "
            "shadow\n",
            nCtx, nSamples, c.render_ms * 1e-3,
            c.render_ms > 0 ? nSamples / (c.render_ms * 1e3) : 0.0, wall,
            (unsigned long long)c.closest_hit_rays,
            (unsigned long long)c.shadow_rays);
    for (sptx_ctx *k : ctxs) sptx_destroy(k);
    film->WriteImage();
}

// Same parameter handling as the reference factory
// (src/integrators/path.cpp:190-213).  The return type is dictated by the
// declaration api.cpp was compiled against; api.cpp only ever converts the
// result to Integrator* (a no-op for single inheritance).
static sptx_host::PathParams ReadPathParams(const ParamSet &params,
                                            const Camera &cameraRef) {
    const Camera *camera = &cameraRef;
    sptx_host::PathParams pp;
    pp.maxDepth = params.FindOneInt("maxdepth", 5);
    int np;
    const int *pb = params.FindInt("pixelbounds", &np);
    Bounds2i pixelBounds = camera->film->GetSampleBounds();
    if (pb) {
        if (np != 4)
            Error("Expected four values for \"pixelbounds\" parameter. Got %d.",
                  np);
        else {
            pixelBounds = Intersect(pixelBounds,
                                    Bounds2i{{pb[0], pb[2]}, {pb[1], pb[3]}});
            if (pixelBounds.Area() == 0)
                Error("Degenerate \"pixelbounds\" specified.");
        }
    }
    pp.pixelBounds[0] = pixelBounds.pMin.x;
    pp.pixelBounds[1] = pixelBounds.pMin.y;
    pp.pixelBounds[2] = pixelBounds.pMax.x;
    pp.pixelBounds[3] = pixelBounds.pMax.y;
    pp.rrThreshold = params.FindOneFloat("rrthreshold", 1.);
    pp.lightSampleStrategy =
        params.FindOneString("lightsamplestrategy", "spatial");
    return pp;
}

PathIntegrator *CreatePathIntegrator(const ParamSet &params,
                                     std::shared_ptr<Sampler> sampler,
                                     std::shared_ptr<const Camera> camera) {
    sptx_host::PathParams pp = ReadPathParams(params, *camera);
    Integrator *gpu = new GpuPathIntegrator(pp, camera, sampler);
    return reinterpret_cast<PathIntegrator *>(gpu);
}

// Integrator "spectralpath" (RenderOptions::MakeIntegrator,
// src/core/api.cpp: name == "spectralpath" -> CreateSpectralPathIntegrator;
// factory src/integrators/spectralpath.cpp:345-378): the same parameters plus
// numCABands (default 4), the number of wavelength bands traced per camera
// sample.  Bound like CreatePathIntegrator: defined here, linked ahead of the
// archive, so the CPU SpectralPathIntegrator object is never pulled in.
SpectralPathIntegrator *CreateSpectralPathIntegrator(
    const ParamSet &params, std::shared_ptr<Sampler> sampler,
    std::shared_ptr<const Camera> camera) {
    sptx_host::PathParams pp = ReadPathParams(params, *camera);
    pp.numCABands = params.FindOneInt("numCABands", 4);
    if (pp.numCABands != 1)
        Warning(
            "Using spectral rendering. For every pixel sample we will trace "
            "%dx more rays. Rendering will be %d times slower.",
            pp.numCABands, pp.numCABands);
    if (pp.numCABands < 1) {
        // the reference's band loop then runs zero times: every sample is black
        Error("numCABands must be >= 1. Got %d.", pp.numCABands);
        return nullptr;
    }
    Integrator *gpu = new GpuPathIntegrator(pp, camera, sampler);
    return reinterpret_cast<SpectralPathIntegrator *>(gpu);
}

// Integrator "metadata" (src/integrators/metadata.cpp:91-130): the standard
// sample loop with an Li that returns the first hit's depth, materialId,
// instanceId or world position in place of radiance (ISET3d's depth maps and
// label images).  Same binding as the two factories above.
MetadataIntegrator *CreateMetadataIntegrator(
    const ParamSet &params, std::shared_ptr<Sampler> sampler,
    std::shared_ptr<const Camera> camera) {
    sptx_host::PathParams pp = ReadPathParams(params, *camera);
    std::string st = params.FindOneString("strategy", "depth");
    if (st == "depth")
        pp.metadataStrategy = SPTX_METADATA_DEPTH;
    else if (st == "material")
        pp.metadataStrategy = SPTX_METADATA_MATERIAL;
    else if (st == "mesh")
        pp.metadataStrategy = SPTX_METADATA_MESH;
    else if (st == "coordinates")
        pp.metadataStrategy = SPTX_METADATA_COORDINATES;
    else {
        Warning(
            "Strategy \"%s\" for metadata unknown. "
            "Using \"depth\".",
            st.c_str());
        pp.metadataStrategy = SPTX_METADATA_DEPTH;
    }
    Integrator *gpu = new GpuPathIntegrator(pp, camera, sampler);
    return reinterpret_cast<MetadataIntegrator *>(gpu);
}

}  // namespace pbrt
