// host/flatten.h -- scene flattening for the B200 path-tracing core.
//
// Product host code.  Turns the reference's pointer graph (Scene, Camera,
// Sampler, PathIntegrator parameters) into the POD tables of include/sptx.h.
// Compiled against the reference's own headers where they lie; nothing of the
// reference is copied.
#ifndef SPTX_HOST_FLATTEN_H
#define SPTX_HOST_FLATTEN_H

#include <cstdint>
#include <string>
#include <vector>

#include "sptx.h"

namespace pbrt {
class Scene;
class Camera;
class Sampler;
}  // namespace pbrt

namespace sptx_host {

// Integrator parameters as CreatePathIntegrator reads them
// (reference src/integrators/path.cpp:190-213).
struct PathParams {
    int maxDepth = 5;
    float rrThreshold = 1.f;
    int pixelBounds[4] = {0, 0, 0, 0};
    std::string lightSampleStrategy = "spatial";
    // 0: Integrator "path".  >= 1: Integrator "spectralpath", numCABands
    // (src/integrators/spectralpath.cpp:345-378)
    int numCABands = 0;
    // 0: radiance.  SPTX_METADATA_*: Integrator "metadata"
    // (src/integrators/metadata.cpp:91-130)
    int metadataStrategy = 0;
};

// Owns the flattened arrays; desc points into them.
struct FlatScene {
    sptx_scene_desc desc;
    std::vector<sptx_bvh_node> nodes;
    std::vector<sptx_bvh> bvhs;
    std::vector<sptx_prim> prims;
    std::vector<sptx_mesh> meshes;
    std::vector<sptx_triangle> triangles;
    std::vector<float> positions, normals, tangents, uvs;
    std::vector<sptx_sphere> spheres;
    std::vector<sptx_instance> instances;
    std::vector<sptx_material> materials;
    std::vector<sptx_lobe> lobes;
    std::vector<float> spectra;
    std::vector<sptx_light> lights;
    std::vector<float> lensElements, exitPupilBounds;
    std::vector<int32_t> primes, primeSums;
    std::vector<uint16_t> perms;
    std::vector<float> envFloats;  // infinite lights' blocks (sptx_light)
    std::vector<float> texFloats;  // image textures' blocks (sptx_scene_desc::tex_floats)
    void Finish();  // points desc at the vectors
};

// Returns false and sets *err if the scene uses anything outside the hot
// path (non-constant textures, other materials/lights/shapes, media, animated
// transforms, a non-BVH aggregate ...): reported, never approximated.
bool FlattenScene(const pbrt::Scene &scene, const pbrt::Camera &camera,
                  pbrt::Sampler &sampler, const PathParams &params,
                  FlatScene *out, std::string *err);

}  // namespace sptx_host

#endif  // SPTX_HOST_FLATTEN_H
