// host/shim/shim_stubs.cpp -- build shim for compiling the reference front end in place.
//
// Link-time stand-ins for the pieces of the reference that host/Makefile
// does NOT compile because they need libraries absent from this image:
//   * src/cameras/omni.cpp, src/cameras/realisticEye.cpp  (GSL)
//   * src/textures/ptex.cpp                               (Ptex)
//   * src/core/imageio.cpp                                (OpenEXR)
// None of them is on the PathIntegrator hot path.  The
// factories report an error the way the reference's own factories do
// (src/core/api.cpp:1752-1762) and return nullptr.  WriteImage keeps a PFM
// writer so that RGB (spectralFlag=false) renders of the analytic scenes
// (src/tests/analytic_scenes.cpp:281) can still be read back by tests.
#include <glog/logging.h>

#include <cmath>
#include <cstdio>
#include <memory>
#include <string>
#include <vector>

#include "cameras/omni.h"
#include "cameras/realisticEye.h"
#include "error.h"
#include "imageio.h"
#include "spectrum.h"
#include "textures/ptex.h"

int FLAGS_stderrthreshold = 1;
int FLAGS_minloglevel = 1;
int FLAGS_v = 0;
bool FLAGS_logtostderr = false;
std::string FLAGS_log_dir;

namespace sptx_glog_shim {
int min_log_level = 2;  // ERROR and above reach stderr
}

namespace pbrt {

OmniCamera *CreateOmniCamera(const ParamSet &, const AnimatedTransform &,
                             Film *, const Medium *) {
    Error("Camera \"omni\" needs GSL: not built into this in-place build.");
    return nullptr;
}

RealisticEye *CreateRealisticEye(const ParamSet &, const AnimatedTransform &,
                                 Film *, const Medium *) {
    Error("Camera \"realisticEye\" needs GSL: not built into this in-place build.");
    return nullptr;
}

PtexTexture<Float> *CreatePtexFloatTexture(const Transform &,
                                           const TextureParams &) {
    Error("Ptex textures are not built into this in-place build.");
    return nullptr;
}

PtexTexture<Spectrum> *CreatePtexSpectrumTexture(const Transform &,
                                                 const TextureParams &) {
    Error("Ptex textures are not built into this in-place build.");
    return nullptr;
}

// PFM only (what the infinite-light fixtures use): "PF" / "Pf", width, height, scale (negative =
// little endian), then float32 scanlines stored bottom to top, flipped to top-down -- the
// semantics of the reference's ReadImagePFM (src/core/imageio.cpp:350-440).  Other formats need
// src/core/imageio.cpp (lodepng / OpenEXR), which this in-place build leaves out.
std::unique_ptr<RGBSpectrum[]> ReadImage(const std::string &name, Point2i *res) {
    const size_t dot = name.rfind('.');
    if (dot == std::string::npos || name.substr(dot) != ".pfm") {
        Error("ReadImage(\"%s\"): only .pfm image reading is built into this in-place build.", name.c_str());
        return nullptr;
    }
    FILE *fp = fopen(name.c_str(), "rb");
    if (!fp) {
        Error("Unable to open \"%s\".", name.c_str());
        return nullptr;
    }
    char tag[8] = {0};
    int w = 0, h = 0;
    float scale = 0.f;
    if (fscanf(fp, "%7s %d %d %f", tag, &w, &h, &scale) != 4 || w <= 0 || h <= 0 ||
        (std::string(tag) != "PF" && std::string(tag) != "Pf")) {
        fclose(fp);
        Error("Error reading PFM file \"%s\"", name.c_str());
        return nullptr;
    }
    fgetc(fp);  // the single whitespace byte that ends the header
    const int nc = std::string(tag) == "PF" ? 3 : 1;
    std::vector<float> data((size_t)nc * w * h);
    const bool ok = fread(data.data(), sizeof(float), data.size(), fp) == data.size();
    fclose(fp);
    if (!ok) {
        Error("Error reading PFM file \"%s\"", name.c_str());
        return nullptr;
    }
    if (scale > 0)  // big-endian file on a little-endian host
        for (float &f : data) {
            unsigned char *b = reinterpret_cast<unsigned char *>(&f);
            std::swap(b[0], b[3]);
            std::swap(b[1], b[2]);
        }
    const float s = std::abs(scale);
    std::unique_ptr<RGBSpectrum[]> rgb(new RGBSpectrum[(size_t)w * h]);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const float *src = &data[(size_t)nc * ((size_t)(h - 1 - y) * w + x)];  // flip in y
            Float c[3] = {s * src[0], s * src[nc == 3 ? 1 : 0], s * src[nc == 3 ? 2 : 0]};
            rgb[(size_t)y * w + x] = RGBSpectrum::FromRGB(c);
        }
    res->x = w;
    res->y = h;
    return rgb;
}

RGBSpectrum *ReadImageEXR(const std::string &name, int *, int *, Bounds2i *,
                          Bounds2i *) {
    Error("ReadImageEXR(\"%s\"): OpenEXR is not built into this in-place build.",
          name.c_str());
    return nullptr;
}

// Whatever the extension, the pixels are written as a little-endian PFM
// (bottom-to-top scanlines, 3 x float32) to "<name>.pfm".
void WriteImage(const std::string &name, const Float *rgb,
                const Bounds2i &outputBounds, const Point2i &) {
    Vector2i res = outputBounds.Diagonal();
    std::string out = name + ".pfm";
    FILE *f = fopen(out.c_str(), "wb");
    if (!f) {
        Error("Unable to open \"%s\" for writing.", out.c_str());
        return;
    }
    fprintf(f, "PF\n%d %d\n-1.0\n", res.x, res.y);
    std::vector<float> row(3 * res.x);
    for (int y = res.y - 1; y >= 0; --y) {
        for (int i = 0; i < 3 * res.x; ++i)
            row[i] = (float)rgb[3 * res.x * y + i];
        fwrite(row.data(), sizeof(float), row.size(), f);
    }
    fclose(f);
}

}  // namespace pbrt
