// wavefront.cuh -- path state in HBM and the wavefront kernels' shared types.
//
// PathIntegrator::Li (integrators/path.cpp:64-188) restructured as a wavefront:
//   generate -> [ intersect -> shade -> resolve ]* -> film
// One "pass" carries nPix * S paths (S consecutive sample numbers of every
// pixel of the render region); path id = s * nPix + pixel, so the film kernel
// can add a pixel's samples in sample order, exactly like FilmTile::AddSample.
// All per-path arrays are structure-of-arrays; spectra are path-major ([P][32] floats, one
// 128-byte line per path and quantity) so a thread moves its row as eight float4.
#pragma once
#include "light.cuh"
#include "sampler.cuh"

namespace sptx {

struct PathArrays {
    int64_t capacity;  // P
    // current path ray: {o.xyz, d.x} {d.y, d.z, etaScale, rayWeight}
    float4 *rayA, *rayB;
    // closest hit of the path ray: {prim, inst, b0, b1} {b2, t, -, -}
    float4 *hitA, *hitB;
    float *beta[2];  // [P][32], ping-pong per bounce
    float *L;        // [P][32]
    uint64_t *sindex;  // Halton index of the path's sample
    uint32_t *info;    // bounces | specularBounce << 8 | dim << 16
    float2 *pfilm;
    // next-event estimation, pending until the resolve kernel
    float4 *shA, *shB;    // shadow ray {o.xyz, d.x} {d.y, d.z, tMax, lightPdfSel}
    float4 *misA, *misB;  // MIS ray {o.xyz, d.x} {d.y, d.z, bits(light), bits(flags)}
    float *fa, *fb;       // [P][32] contributions of the two MIS strategies
    // queues of path ids
    int *q[2];
    int *qNee;
    int *qMis;  // subset of qNee that also has a BSDF-sampled MIS ray (counters[4])
    // the active queue again, split by shading class (written by K2, read by K3):
    // warps of K3 then run one kind of material code
    int *qc;        // [kShadeClasses][P]
    int *counters;  // [0],[1]: sizes of q[0], q[1]; [2]: qNee size; [3]: error flags;
                    // [8 .. 8+kShadeClasses): sizes of the class queues
    uint8_t *spill;  // per path: film sample needs the atomic (wide filter) pass
};

constexpr int kShadeClasses = 8;  // class 0 = no hit
constexpr int kClassCounter0 = 8;
enum { NEE_HAS_SHADOW = 1, NEE_HAS_MIS = 2, NEE_MIS_REACHED_LIGHT = 4 };
enum { ERR_DIM_OVERFLOW = 1, ERR_INTERACTION = 2, ERR_TERMS = 4 };

// The kernel variants: lobe types a variant's code can evaluate, and how many non-specular lobes
// (terms of BSDF::f) one of its materials may have.  sptx_upload_scene maps every shading class to
// the first variant that covers its materials (shade_variant_for).
constexpr int kNumShadeVariants = 5;
constexpr int kVariantLobes[kNumShadeVariants] = {
    1 << SPTX_LOBE_LAMBERT,                                    // matte
    kSpecularLobes,                                            // glass, mirror: no light sampling
    (1 << SPTX_LOBE_LAMBERT) | (1 << SPTX_LOBE_MICROFACET_REFL),  // plastic, metal
    kAllLobes,
    // matte / plastic with an imagemap Kd (the only variant that evaluates textures)
    (1 << SPTX_LOBE_LAMBERT) | (1 << SPTX_LOBE_MICROFACET_REFL) | kTexturedLobes};
constexpr int kVariantTerms[kNumShadeVariants] = {1, 1, 2, kMaxTerms, 2};

// Which variant K3 runs for each shading class (-1: no primitive of the scene is in the class).
struct ShadePlan {
    int variant[kShadeClasses];
};
// First variant whose lobe set covers `lobeMask` (lobe types, | kTexturedLobes when a material of the
// class has a textured lobe) and whose term list holds `nonSpecular` terms; -1: none does (textured
// materials with lobes outside the textured variant).
inline int shade_variant_for(int lobeMask, int nonSpecular) {
    for (int v = 0; v < kNumShadeVariants; ++v)
        if ((lobeMask & ~kVariantLobes[v]) == 0 && nonSpecular <= kVariantTerms[v]) return v;
    return (lobeMask & kTexturedLobes) ? -1 : 3;
}

struct RenderParams {
    int nPix;          // pixels in the render region
    int samplesInPass; // S
    int sampleBegin;   // first sample number of this pass
    int maxDepth;
    float rrThreshold;
    const int2 *pixels;  // [nPix] pixel coordinates
    float wavelength;    // ray.wavelength (550 for PathIntegrator)
    int ignoreRayWeight; // MetadataIntegrator::IgnoreRayWeight (integrators/metadata.h:64): the film
                         // weight of a sample is 1 (core/integrator.cpp:320-321); the camera's
                         // weight still gates Li
    int band, nBands;    // SpectralPathIntegrator: this launch traces band `band` of `nBands`
    const int *sampleList;  // debug: explicit sample number per path (or NULL)
    int betaOnes;  // this bounce's incoming throughput is 1 in every bin (the camera vertex): its rows are
                   // neither written by K1 nor read by K3 / K4b
};

struct DeviceCounters {
    unsigned long long cameraRays, closestRays, shadowRays, vertices, nodes[2], tests[2];
};

SPTX_DEV uint32_t pack_info(int bounces, int specular, int dim) {
    return (uint32_t)bounces | ((uint32_t)specular << 8) | ((uint32_t)dim << 16);
}

// SampledSpectrum::y() tail, core/spectrum.h:415-421.
SPTX_DEV float y_finish(const DScene &S, float yy) {
    yy = (yy < 0) ? 0.f : yy;
    return yy * S.y_scale_num / S.y_scale_den;
}

// Warp-aggregated append: one atomicAdd per warp (ballot + popc prefix).
SPTX_DEV void queue_push(int *queue, int *counter, int value, bool pred) {
    unsigned active = __activemask();
    unsigned mask = __ballot_sync(active, pred);
    if (mask == 0) return;
    int leader = __ffs(mask) - 1;
    int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(active, base, leader);
    if (pred) queue[base + __popc(mask & ((1u << lane) - 1u))] = value;
}

}  // namespace sptx
