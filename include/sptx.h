/* sptx.h -- C ABI of the B200 spectral path-tracing core.
 *
 * This is the drop-in boundary for pbrt-v3-spectral's PathIntegrator hot path.
 * The reference has no FFI: `pbrtWorldEnd` calls `Integrator::Render(scene)`
 * in-process (reference src/core/api.cpp:1695-1707, src/core/integrator.h:53-58).
 * The host C++ layer (host/gpu_path_integrator.cpp) implements that interface,
 * flattens the reference's scene graph into the POD tables below and drives
 * these entry points.  Every struct is plain data: no C++ types, no torch
 * types, pointers + counts only.  All functions return 0 on success and a
 * non-zero SPTX_E* code on failure; nothing throws across this boundary.
 *
 * Arithmetic conventions (needed for parity with the reference's x86-64 build,
 * CMakeLists.txt:73-76: -O3, no FMA contraction): all device math is IEEE
 * float32 with fused multiply-add DISABLED, double where the reference uses
 * double (Cross, Quadratic, triangle edge fallback).
 */
#ifndef SPTX_H
#define SPTX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPTX_NBINS 31          /* nSpectralSamples, src/core/spectrum.h:50 */
#define SPTX_SPECTRUM_STRIDE 32 /* floats per row of the spectra table (128 B) */
#define SPTX_FILTER_TABLE_WIDTH 16 /* Film::filterTableWidth, src/core/film.h:98 */
#define SPTX_ABI_VERSION 1

/* ---- error codes ------------------------------------------------------- */
enum {
    SPTX_OK = 0,
    SPTX_EINVAL = 1,       /* bad argument / malformed scene tables          */
    SPTX_ECUDA = 2,        /* CUDA runtime error (see sptx_last_error)       */
    SPTX_EUNSUPPORTED = 3, /* scene uses a feature outside the hot path      */
    SPTX_ENOMEM = 4,
    SPTX_EIO = 5,
    SPTX_ESTATE = 6        /* call order (e.g. render before upload)         */
};

/* ---- geometry tables ---------------------------------------------------- */

/* Bit-identical to LinearBVHNode, src/accelerators/bvh.cpp:95-104 (32 B). */
typedef struct sptx_bvh_node {
    float bmin[3];
    float bmax[3];
    int32_t offset;   /* leaf: primitivesOffset; interior: secondChildOffset */
    uint16_t n_prims; /* 0 -> interior                                        */
    uint8_t axis;
    uint8_t pad;
} sptx_bvh_node;

/* One BVHAccel.  bvhs[0] is the scene aggregate; the others are the nested
 * aggregates object instances point at (src/core/api.cpp:1590-1613).
 * Node and primitive offsets stored inside nodes are relative to the bases. */
typedef struct sptx_bvh {
    int32_t node_base;
    int32_t n_nodes;
    int32_t prim_base;
    int32_t n_prims;
} sptx_bvh;

enum { SPTX_PRIM_TRIANGLE = 0, SPTX_PRIM_SPHERE = 1, SPTX_PRIM_INSTANCE = 2 };

/* Entry of BVHAccel::primitives (post-build order).  GeometricPrimitive binds
 * shape+material+areaLight (src/core/primitive.cpp:119-152);
 * TransformedPrimitive is an instance (src/core/primitive.cpp:78-99). */
typedef struct sptx_prim {
    int32_t kind;     /* SPTX_PRIM_*                                         */
    int32_t shape;    /* index into triangles / spheres / instances          */
    int32_t material; /* index into materials, -1 = none                     */
    int32_t light;    /* index into lights (area light), -1 = none           */
} sptx_prim;

/* TriangleMesh, src/shapes/triangle.h:51-69 (positions are world space). */
typedef struct sptx_mesh {
    int32_t p_base;  /* first vertex of this mesh in positions[]             */
    int32_t n_base;  /* first normal in normals[], -1 if the mesh has none   */
    int32_t s_base;  /* first tangent in tangents[], -1 if none              */
    int32_t uv_base; /* first uv in uvs[], -1 if none                        */
    int32_t flip;    /* reverseOrientation ^ transformSwapsHandedness        */
    int32_t n_vertices;
} sptx_mesh;

typedef struct sptx_triangle {
    int32_t v[3]; /* mesh-local vertex indices                               */
    int32_t mesh;
} sptx_triangle;

/* Sphere, src/shapes/sphere.h.  Matrices are row-major m[r][c]. */
typedef struct sptx_sphere {
    float o2w[16];     /* ObjectToWorld->m                                   */
    float w2o[16];     /* ObjectToWorld->mInv (== WorldToObject->m)          */
    float radius, z_min, z_max, theta_min, theta_max, phi_max;
    int32_t reverse_orientation;
    int32_t swaps_handedness;
} sptx_sphere;

typedef struct sptx_instance {
    float p2w[16]; /* PrimitiveToWorld.m (static transforms only)            */
    float w2p[16]; /* its inverse (mInv)                                      */
    int32_t bvh;   /* index into bvhs[]                                       */
    int32_t instance_id;
    int32_t is_identity;
    int32_t pad;
} sptx_instance;

/* ---- shading tables ------------------------------------------------------ */

enum {
    SPTX_LOBE_LAMBERT = 0,         /* LambertianReflection                   */
    SPTX_LOBE_OREN_NAYAR = 1,      /* OrenNayar (A, B precomputed)           */
    SPTX_LOBE_SPECULAR_REFL = 2,   /* SpecularReflection                     */
    SPTX_LOBE_SPECULAR_TRANS = 3,  /* SpecularTransmission                   */
    SPTX_LOBE_FRESNEL_SPECULAR = 4,/* FresnelSpecular                        */
    SPTX_LOBE_MICROFACET_REFL = 5, /* MicrofacetReflection + TrowbridgeReitz */
    SPTX_LOBE_LAMBERT_TRANS = 6,   /* LambertianTransmission (T)             */
    SPTX_LOBE_MICROFACET_TRANS = 7,/* MicrofacetTransmission + TrowbridgeReitz
                                      (T, eta_a, eta_b, alphas)              */
    SPTX_LOBE_FRESNEL_BLEND = 8    /* FresnelBlend + TrowbridgeReitz
                                      (R = Rd, T = Rs, alphas)               */
};
enum { SPTX_FRESNEL_NOOP = 0, SPTX_FRESNEL_DIELECTRIC = 1, SPTX_FRESNEL_CONDUCTOR = 2 };

/* One BxDF of the BSDF a material builds for constant textures
 * (src/materials/{matte,plastic,glass,metal,mirror}.cpp).  Spectra are row
 * indices into spectra[] (already clamped the way the material clamps them). */
typedef struct sptx_lobe {
    int32_t type;    /* SPTX_LOBE_*                                          */
    int32_t fresnel; /* SPTX_FRESNEL_* for SPECULAR_REFL / MICROFACET_REFL   */
    int32_t R;       /* reflectance spectrum row, -1 if unused, <= -2: image texture
                        SPTX_TEX_ROW(offset) (LAMBERT / OREN_NAYAR lobes only)   */
    int32_t T;       /* transmittance spectrum row, -1 if unused             */
    int32_t fr_eta_i, fr_eta_t, fr_k; /* conductor Fresnel spectra rows      */
    float eta_a, eta_b;   /* dielectric: Fresnel etaI/etaT or lobe etaA/etaB */
    float alpha_x, alpha_y; /* Trowbridge-Reitz alphas (after remap)         */
    float A, B;             /* Oren-Nayar                                    */
    int32_t sample_visible; /* MicrofacetDistribution::sampleVisibleArea     */
} sptx_lobe;

typedef struct sptx_material {
    int32_t first_lobe;
    int32_t n_lobes; /* <= 8, BSDF::MaxBxDFs, src/core/reflection.h:199      */
    float eta;       /* BSDF::eta                                            */
    int32_t material_id;
} sptx_material;

enum { SPTX_LIGHT_DIFFUSE_AREA = 0, SPTX_LIGHT_POINT = 1, SPTX_LIGHT_INFINITE = 2 };
enum { SPTX_METADATA_NONE = 0, SPTX_METADATA_DEPTH = 1, SPTX_METADATA_MATERIAL = 2, SPTX_METADATA_MESH = 3,
       SPTX_METADATA_COORDINATES = 4 };

/* DiffuseAreaLight (src/lights/diffuse.cpp), PointLight (src/lights/point.cpp) or InfiniteAreaLight
 * (src/lights/infinite.cpp).  For an infinite light `shape` is the offset (in floats) of its block in
 * env_floats[], `spectrum` the first of the seven consecutive rows rgbIllum2Spect{White, Cyan,
 * Magenta, Yellow, Red, Green, Blue} (src/core/spectrum.cpp:363-392), `area` = worldRadius and
 * `pos` = worldCenter (InfiniteAreaLight::Preprocess).  Block layout (integers as float values):
 *   [0] W [1] H of the radiance map's level 0 (MIPMap::pyramid[0]), [2] nu [3] nv of the sampling
 *   distribution, [4] Power().y(), [5..8) unused, [8..24) LightToWorld.m, [24..40) WorldToLight.m,
 *   W*H RGB texels (row v, column u), then Distribution2D: the marginal's func[nv], cdf[nv+1],
 *   funcInt, then per row v: func[nu], cdf[nu+1], funcInt (src/core/sampling.h:55-150). */
typedef struct sptx_light {
    int32_t type;
    int32_t shape_kind; /* SPTX_PRIM_TRIANGLE / SPTX_PRIM_SPHERE (area)      */
    int32_t shape;      /* triangle / sphere index                           */
    int32_t two_sided;
    int32_t spectrum;   /* Lemit (area) or I (point) row in spectra[]        */
    float area;         /* shape->Area() as the reference computed it        */
    float pos[3];       /* point light position                              */
} sptx_light;

/* ---- sensor -------------------------------------------------------------- */

enum { SPTX_CAMERA_PERSPECTIVE = 0, SPTX_CAMERA_REALISTIC = 1 };

typedef struct sptx_camera {
    int32_t type;
    float raster_to_camera[16]; /* ProjectiveCamera::RasterToCamera.m         */
    float camera_to_world[16];  /* CameraToWorld start transform .m           */
    float lens_radius, focal_distance;
    float shutter_open, shutter_close;
    /* RealisticCamera (src/cameras/realistic.cpp) */
    int32_t n_elements;      /* rows of lens_elements[] (4 floats each)       */
    int32_t n_pupil_bounds;  /* rows of exit_pupil_bounds[] (4 floats each)   */
    int32_t simple_weighting;
    int32_t ca_flag;         /* chromaticAberrationEnabled                    */
    float film_diagonal;     /* Film::diagonal (metres)                       */
    float pad[3];
} sptx_camera;

enum {
    SPTX_SAMPLER_HALTON = 0,
    SPTX_SAMPLER_RANDOM = 1,
    SPTX_SAMPLER_SOBOL = 2,
    SPTX_SAMPLER_ZEROTWO = 3 /* ZeroTwoSequenceSampler (src/samplers/zerotwosequence.cpp): spp is the
                                rounded-up power of two, n_dims = nSampledDimensions; no tables */
};

/* HaltonSampler state, src/samplers/halton.cpp:65-92. */
typedef struct sptx_sampler {
    int32_t type;
    int32_t spp;
    int32_t sample_at_center;
    int32_t base_scales[2];
    int32_t base_exponents[2];
    int32_t sample_stride;
    int32_t mult_inverse[2];
    int32_t n_dims; /* dimensions covered by primes[]/prime_sums[]/perms[]    */
    int32_t log2_resolution; /* SobolSampler::log2Resolution, src/samplers/sobol.h:69 */
} sptx_sampler;

/* Film, src/core/film.cpp:50-92. Bounds are {x0, y0, x1, y1}. */
typedef struct sptx_film {
    int32_t full_res[2];
    int32_t cropped[4];       /* croppedPixelBounds                           */
    int32_t sample_bounds[4]; /* GetSampleBounds()                            */
    float filter_radius[2];
    float filter_table[SPTX_FILTER_TABLE_WIDTH * SPTX_FILTER_TABLE_WIDTH];
    float max_sample_luminance;
    float scale;
    float diagonal;
    int32_t spectral_flag;
} sptx_film;

enum { SPTX_LIGHTS_UNIFORM = 0, SPTX_LIGHTS_POWER = 1, SPTX_LIGHTS_SPATIAL = 2 };

/* PathIntegrator parameters, src/integrators/path.cpp:190-213. */
typedef struct sptx_integrator {
    int32_t max_depth;
    float rr_threshold;
    int32_t pixel_bounds[4];
    int32_t light_strategy; /* as named in the scene file; the 1-light       */
                            /* fallback to uniform is applied by the core     */
    int32_t num_ca_bands;   /* 0 or 1: Integrator "path".  >= 1 from Integrator  */
                            /* "spectralpath" (numCABands): per camera sample  */
                            /* one path per wavelength band, each band's bins   */
                            /* taken from its own path, src/integrators/        */
                            /* spectralpath.cpp:198-342                         */
} sptx_integrator;

/* Everything Render() needs, as flat host arrays.  The caller keeps ownership
 * and may free the arrays once sptx_upload_scene returns. */
typedef struct sptx_scene_desc {
    uint32_t abi_version;
    uint32_t metadata_strategy; /* 0: radiance (Integrator "path" / "spectralpath").      */
                                /* SPTX_METADATA_*: Integrator "metadata", whose Li is the */
                                /* first hit's depth / materialId / instanceId / position  */
                                /* (src/integrators/metadata.cpp:53-89)                    */
    sptx_camera camera;
    sptx_sampler sampler;
    sptx_film film;
    sptx_integrator integrator;
    float cie_y[SPTX_SPECTRUM_STRIDE]; /* SampledSpectrum::Y                  */
    float cie_y_integral;               /* CIE_Y_integral, spectrum.h:87       */
    int32_t pad0;

    int64_t n_nodes;      const sptx_bvh_node *nodes;
    int64_t n_bvhs;       const sptx_bvh *bvhs;
    int64_t n_prims;      const sptx_prim *prims;
    int64_t n_meshes;     const sptx_mesh *meshes;
    int64_t n_triangles;  const sptx_triangle *triangles;
    int64_t n_positions;  const float *positions; /* xyz                      */
    int64_t n_normals;    const float *normals;   /* xyz                      */
    int64_t n_tangents;   const float *tangents;  /* xyz                      */
    int64_t n_uvs;        const float *uvs;       /* uv                       */
    int64_t n_spheres;    const sptx_sphere *spheres;
    int64_t n_instances;  const sptx_instance *instances;
    int64_t n_materials;  const sptx_material *materials;
    int64_t n_lobes;      const sptx_lobe *lobes;
    int64_t n_spectra;    const float *spectra;   /* rows of 32 floats        */
    int64_t n_lights;     const sptx_light *lights;
    int64_t n_lens_floats;  const float *lens_elements;     /* 4 per element  */
    int64_t n_pupil_floats; const float *exit_pupil_bounds; /* 4 per bound    */
    /* Sampler tables.  Halton: Primes / PrimeSums / radicalInversePermutations.
     * Sobol (SPTX_SAMPLER_SOBOL) reuses the first two slots: `primes` holds
     * SobolMatrices32[0 .. n_dims * 52) (bit patterns), `prime_sums` holds
     * VdCSobolMatrices[m - 1][0..52) then VdCSobolMatricesInv[m - 1][0..52)
     * (m = log2_resolution) as little-endian (lo, hi) int32 pairs; `perms` is empty
     * (src/core/sobolmatrices.h:47-52, src/core/lowdiscrepancy.h:229-274). */
    int64_t n_primes;     const int32_t *primes;     /* Primes[0..n_dims)     */
    int64_t n_prime_sums; const int32_t *prime_sums; /* PrimeSums[0..n_dims)  */
    int64_t n_perms;      const uint16_t *perms; /* radicalInversePermutations */
    /* Infinite lights' blocks (see sptx_light).  Last member: scene files written before it existed
     * are still read (sptx_scene_load fills it with 0 / NULL). */
    int64_t n_env_floats; const float *env_floats;
    /* Image textures (ImageTexture<RGBSpectrum, Spectrum>, src/textures/imagemap.h:80-137, with a
     * UVMapping2D).  A lobe whose R is the value of such a texture has R = SPTX_TEX_ROW(offset),
     * offset = where the texture's block starts in tex_floats[] (integers stored as bit patterns):
     *   [0] levels [1] wrap (0 Repeat, 1 Black, 2 Clamp) [2] doTrilinear [3] maxAnisotropy (float)
     *   [4..8) su sv du dv (float) [8] first of the seven rgbIllum2Spect* rows in spectra[]
     *       (ImageTexture::convertOut's FromRGB(rgb) defaults to SpectrumType::Illuminant)
     *   [9] offset of MIPMap::weightLut (128 floats) in tex_floats[] [10..16) unused,
     *   then per level {uSize, vSize, offset of its RGB texels from the block start},
     *   then the texels, level after level, row t, column s (MIPMap::pyramid, src/core/mipmap.h).
     * Also a trailing member (older scene files are still read). */
    int64_t n_tex_floats; const float *tex_floats;
} sptx_scene_desc;
#define SPTX_TEX_ROW(offset) (-2 - (int32_t)(offset))

/* ---- ray batches (parity test #1 and the traversal benchmark) ----------- */

typedef struct sptx_ray {
    float o[3];
    float t_max;
    float d[3];
    float time;
} sptx_ray; /* 32 B */

typedef struct sptx_hit {
    int32_t prim;     /* global index into prims[] of the hit primitive, -1 = miss */
    int32_t instance; /* index into instances[] of the enclosing instance, -1 = none */
    float t;          /* ray.tMax after the traversal                         */
    float b0, b1, b2; /* triangle: barycentrics at the hit; sphere: its object-space hit point */
    int32_t n_nodes;  /* BVH nodes visited (only when counters are enabled)   */
    int32_t n_tests;  /* primitive tests made (only when counters are enabled) */
} sptx_hit; /* 32 B */

typedef struct sptx_counters {
    uint64_t camera_rays;
    uint64_t closest_hit_rays; /* "Regular ray intersection tests"            */
    uint64_t shadow_rays;      /* "Shadow ray intersection tests"             */
    uint64_t path_vertices;
    uint64_t kernel_launches;
    double render_ms;          /* device time of the last sptx_render         */
    double traverse_ms;        /* of which traversal kernels                  */
    double shade_ms;
    double film_ms;
    /* per kernel type, CUDA-event time and launches since the last reset:
     * 0 generate, 1 intersect, 2 shade, 3 resolve, 4 film, 5 film spill, 6 ray batch */
    double kernel_ms[8];
    uint64_t kernel_count[8];
    /* visit counts of the wavefront traversal kernels (only when enabled with
     * sptx_set_option("count_visits", 1)): BVH nodes fetched / primitive tests */
    uint64_t nodes_visited[2]; /* [0] intersect kernel, [1] resolve kernel    */
    uint64_t prims_tested[2];
} sptx_counters;

typedef struct sptx_ctx sptx_ctx;

/* ---- entry points -------------------------------------------------------- */

/* Creates a context on CUDA device `device`.  Replaces: construction of the
 * integrator in RenderOptions::MakeIntegrator (src/core/api.cpp:1750-1771). */
int sptx_create(int device, sptx_ctx **out);
void sptx_destroy(sptx_ctx *ctx);
/* Message of the last failure on ctx (or of the last failed sptx_create /
 * scene-file call when ctx is NULL).  Replaces Error() (src/core/error.cpp:89-103). */
const char *sptx_last_error(const sptx_ctx *ctx);

/* Copies the scene tables to HBM and builds device-side derived tables
 * (per-leaf triangle vertex blocks, Halton pixel offsets, light-selection
 * tables).  Replaces Preprocess + the pointer graph Render walks
 * (src/core/integrator.cpp:229, src/integrators/path.cpp:59-62). */
int sptx_upload_scene(sptx_ctx *ctx, const sptx_scene_desc *scene);

/* BVHAccel::Intersect (any_hit = 0, src/accelerators/bvh.cpp:662-700) or
 * BVHAccel::IntersectP (any_hit = 1, :702-738) for n rays in HOST memory. */
int sptx_intersect_batch(sptx_ctx *ctx, const sptx_ray *rays, int64_t n,
                         sptx_hit *hits, int any_hit, int with_counters);
/* Same with rays and hits already resident in device memory; *ms (optional)
 * receives the CUDA-event time of the traversal kernel alone. */
int sptx_intersect_batch_device(sptx_ctx *ctx, const void *d_rays, int64_t n,
                                void *d_hits, int any_hit, int with_counters,
                                float *ms);

/* sptx_intersect_batch (closest hit / any hit for n rays in HOST memory) for callers that
 * cannot allocate device memory: the device copies of rays and hits live in a fixed file-scope
 * device array (SPTX_NOALLOC_MAX_RAYS * 64 bytes = 4 MiB, held by every GPU that loads the
 * library until the process exits) instead of cudaMalloc'd scratch, and the call makes no
 * cudaMalloc / cudaFree / allocating library call.  n > SPTX_NOALLOC_MAX_RAYS is rejected with
 * SPTX_EINVAL.  All calls share the array: calls that overlap on one GPU (two contexts, two
 * threads) race and must be serialised by the caller.  Visit counters are not recorded
 * (n_nodes = n_tests = 0).
 * No counterpart in the reference. */
#define SPTX_NOALLOC_MAX_RAYS 65536
int sptx_intersect_batch_noalloc(sptx_ctx *ctx, const sptx_ray *rays, int64_t n,
                                 sptx_hit *hits, int any_hit);

/* sptx_set_option(ctx, "device_bvh", 1): the next sptx_upload_scene ignores the scene's node array
 * and builds every aggregate's BVH on the device (Morton codes, radix sort, parallel LBVH; csrc/lbvh.cuh)
 * instead of taking the host build of BVHAccel::BVHAccel (src/accelerators/bvh.cpp:183-225, HLBVHBuild
 * :404-638).  Primitive ids reported by sptx_intersect_batch stay the caller's.  Device time of the
 * last build in milliseconds (0 when the host's tree was used): */
double sptx_bvh_build_ms(const sptx_ctx *ctx);

/* Measurement hook (no counterpart in the reference): read bandwidth of an L2-resident buffer of
 * `bytes`, swept `reps` times, in GB/s -- the roofline denominator of traversal when the BVH
 * working set lives in L2 (bench.py, tools/bench_configs.py). */
int sptx_measure_l2(sptx_ctx *ctx, int64_t bytes, int reps, double *gbs);

/* SamplerIntegrator::Render's pixel/sample loop (src/core/integrator.cpp:
 * 228-334) for sample numbers [sample_begin, sample_end) of every pixel of
 * `pixel_bounds` (NULL = the film's sample bounds), accumulated into the
 * device film.  Does not clear the film and does not write files. */
int sptx_render(sptx_ctx *ctx, const int32_t *pixel_bounds, int sample_begin,
                int sample_end);
/* Rank-interleaved variant for multi-GPU pixel partitioning: renders only the
 * 16x16 tiles t with t % n_parts == part (tile order of integrator.cpp:235-249). */
int sptx_render_tiles(sptx_ctx *ctx, int part, int n_parts, int sample_begin,
                      int sample_end);

/* Film::Clear (src/core/film.cpp:114-122). */
int sptx_film_clear(sptx_ctx *ctx);
/* Device film: contrib is float32 [H][W][32] (31 bins + filterWeightSum in
 * slot 31) over croppedPixelBounds.  Exposed so the host can sum films across
 * GPUs with NCCL (the reference's MergeFilmTile, src/core/film.cpp:124-142). */
int sptx_film_device_ptr(sptx_ctx *ctx, void **d_film, int64_t *n_floats);
/* Multi-GPU film merge over NVLink peer memory (the reference's MergeFilmTile,
 * src/core/film.cpp:124-142, across processes).  Every rank exports its film
 * (a 64-byte CUDA IPC handle, to be passed to the merging rank by any host
 * channel); the merging rank maps the peers' films and ONE kernel sums
 * film[0] + film[1] + ... in rank order into its own film -- reading the peers
 * straight over NVLink -- and, if `planes` is given, also writes the .dat
 * payload of sptx_film_pack_dat from the same pass (float64 [31][H][W] on the
 * host).  `handles` holds n_ranks handles of SPTX_IPC_HANDLE_BYTES each; slot
 * `self` is ignored.  The caller synchronises the ranks around the call (all
 * renders done before, peers keep their contexts alive until it returns). */
#define SPTX_IPC_HANDLE_BYTES 64
int sptx_film_export(sptx_ctx *ctx, void *handle);
int sptx_film_merge_peers(sptx_ctx *ctx, const void *handles, int n_ranks, int self,
                          double *planes);
/* The same merge shaped as a reduce-scatter: EVERY rank calls it (between two barriers of the
 * caller) and sums its own 1/n_ranks slice of the film over all ranks, in rank order (same result
 * bit for bit), writing the sums into rank `dst`'s film over NVLink.  Each link carries 1/n_ranks of
 * a film per peer instead of rank dst pulling n_ranks - 1 whole films through its own ingress.
 * Mapped peer films stay open in the context between calls. */
int sptx_film_reduce_scatter(sptx_ctx *ctx, const void *handles, int n_ranks, int self, int dst);
/* One process driving one context per GPU (GpuPathIntegrator::Render with SPTX_DEVICES; SURVEY
 * 8(b) "threading"): the films of ctxs[0..n) are summed in that order into ctxs[dst]'s film, each
 * device summing its own slice through peer access (no IPC handles, no NCCL).  Replaces
 * Film::MergeFilmTile (src/core/film.cpp:124-142) across the render threads' films. */
int sptx_film_merge_contexts(sptx_ctx *const *ctxs, int n, int dst);
/* SURVEY 8(b)'s NCCL form of the exchange step, for hosts that already hold a communicator:
 * ncclReduce(sum, float32, W*H*32 floats) of the device film onto `root`, on the context's
 * stream; `nccl_comm` is an ncclComm_t.  libnccl.so.2 is looked up at run time (no link-time
 * dependency).  A single thread driving several devices wraps its calls in ncclGroupStart / End. */
int sptx_film_reduce(sptx_ctx *ctx, void *nccl_comm, int root);
/* Film pixels to host: contrib [H][W][31] float32, weight [H][W] float32. */
int sptx_film_download(sptx_ctx *ctx, float *contrib, float *weight);
/* Film::WriteImage's spectral branch (src/core/film.cpp:226-306) up to the
 * file: L * scale as float64, bin-major [31][H][W], into host memory. */
int sptx_film_pack_dat(sptx_ctx *ctx, double *planes);
/* The .dat writer itself: header "W H 31\nv3 \n" + planes (film.cpp:267-306). */
int sptx_write_dat(const char *path, int width, int height, const double *planes);

int sptx_get_counters(sptx_ctx *ctx, sptx_counters *out);
int sptx_reset_counters(sptx_ctx *ctx);
/* Named integer options: "count_visits" (0/1), "device_bvh" (0/1), "shade_exact" (0/1), and
 * "ray_sort" for sptx_intersect_batch / _device: 0 = rays are walked in the caller's order, 1 = in the
 * order of a key made of the origin's cell in the scene bounds and the direction's octant (one radix
 * sort on the device; results stay at the ray's own index and are identical), 2 (default) = like 1
 * for batches of >= 65536 rays (smaller ones do not repay the sort).  sptx_intersect_batch_noalloc
 * never sorts (the sort's scratch is allocated). */
int sptx_set_option(sptx_ctx *ctx, const char *name, int64_t value);
/* "sparse_voxels" 1 (takes effect at the next upload): the spatial light distribution's per-voxel tables
 * (SpatialLightDistribution, src/core/lightdistrib.cpp:91-300) are built only for the voxels that can hold
 * a surface point; the default does so when the full table would pass 256 MB (many lights).
 * sptx_light_table_rows: number of voxel rows the last upload built (0: not the spatial strategy). */
int64_t sptx_light_table_rows(sptx_ctx *ctx);
/* 1 when the last sptx_intersect_batch / _device call walked its rays in sorted order. */
int sptx_last_batch_sorted(sptx_ctx *ctx);
/* Max paths in flight per wavefront pass (0 = default). */
int sptx_set_wavefront_size(sptx_ctx *ctx, int64_t max_paths);

/* Test hooks: the device sampler and camera evaluated stand-alone.
 * values[i] = SampleDimension(index of (px[i],py[i],sample[i]), dim[i])
 * (src/samplers/halton.cpp:96-127). */
int sptx_sampler_eval(sptx_ctx *ctx, const int32_t *px, const int32_t *py,
                      const int32_t *sample, const int32_t *dim, int64_t n,
                      float *values);
/* rays[i], weights[i] = Camera::GenerateRayDifferential's main ray for
 * CameraSample{pFilm, time, pLens} = cs[5*i .. 5*i+4] at `wavelength`. */
int sptx_camera_rays(sptx_ctx *ctx, const float *cs, int64_t n, float wavelength,
                     sptx_ray *rays, float *weights);

/* L[i][31] = PathIntegrator::Li (src/integrators/path.cpp:64-188) of sample
 * number sample[i] of pixel (px[i], py[i]), following the per-sample sequence
 * of SamplerIntegrator::Render (src/core/integrator.cpp:264-292); nothing is
 * added to the film.  pfilm_weight[i] = {pFilm.x, pFilm.y, rayWeight} (optional). */
int sptx_debug_radiance(sptx_ctx *ctx, const int32_t *px, const int32_t *py,
                        const int32_t *sample, int64_t n, float *L, float *pfilm_weight);

/* Flattened-scene files (what host/flatten.cpp writes). */
int sptx_scene_save(const char *path, const sptx_scene_desc *scene);
int sptx_scene_load(const char *path, sptx_scene_desc **out);
void sptx_scene_free(sptx_scene_desc *scene);
const char *sptx_io_last_error(void);

int sptx_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SPTX_H */
