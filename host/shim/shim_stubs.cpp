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

std::unique_ptr<RGBSpectrum[]> ReadImage(const std::string &name, Point2i *) {
    Error("ReadImage(\"%s\"): image reading is not built into this in-place build.",
          name.c_str());
    return nullptr;
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
