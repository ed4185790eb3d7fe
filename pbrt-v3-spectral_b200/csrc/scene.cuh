// scene.cuh -- the scene as it lives in HBM.
//
// Layout (DESIGN.md "Data layout in HBM"):
//   nodes    : LinearBVHNode array, 32 B each, all aggregates concatenated,
//              read as two 128-bit loads per node.
//   leafv    : one 48 B block per primitive IN BVH PRIMITIVE ORDER (so the
//              triangles of a leaf are contiguous): three float4 =
//              {p0.xyz, p1.x} {p1.yz, p2.xy} {p2.z, bits(shape), 0, bits(kind)}.
//              Traversal touches only nodes + leafv.
//   the shading tables (meshes, normals, uvs, materials, lobes, spectra,
//   lights) are read by the shading kernels through the read-only path.
#pragma once
#include "common.cuh"

namespace sptx {

struct DCamera;

struct DScene {
    const float4 *nodes;  // 2 x float4 per node
    const sptx_bvh *bvhs;
    const float4 *leafv;  // 3 x float4 per primitive
    const sptx_prim *prims;
    const uint8_t *prim_class;  // shading class of the primitive's material (1..kShadeClasses-1)
    const int *prim_ids;  // device-built BVHs reorder the primitives: index the caller knows (or NULL)
    const sptx_mesh *meshes;
    const sptx_triangle *triangles;
    const float *positions, *normals, *tangents, *uvs;
    const sptx_sphere *spheres;
    const sptx_instance *instances;
    const sptx_material *materials;
    const sptx_lobe *lobes;
    const float *spectra;
    const sptx_light *lights;
    int n_lights;
    const float *env;       // infinite lights' blocks (sptx_light), or NULL
    int n_infinite;         // number of infinite lights; their indices in lights[]:
    int infinite[4];
    int n_bvhs;
    // every primitive that carries an area light sits in the top-level aggregate (the reference drops area
    // lights inside object instances, api.cpp:1377-1386): its shape can be tested with a world-space ray
    int lights_top_level;
    // image textures' blocks (sptx_scene_desc::tex_floats), or NULL; the camera (device copy), which
    // the textured shading variant needs for the camera ray's differentials
    const float *tex;
    const DCamera *camera;
    // light selection
    int light_strategy;      // effective (uniform when n_lights == 1)
    const float *light_func; // uniform/power: func[n], then cdf[n+1]
    float light_func_int;
    // spatial: per-voxel rows of (n_lights func) + (n_lights+1 cdf) + funcInt
    const float *voxel_tables;
    const int *voxel_rows;   // sparse tables: row of a voxel (-1: no surface in it), or NULL (row = voxel)
    int n_voxels[3];
    float wb_min[3], wb_max[3];  // scene.WorldBound()
    // colour matching
    float cie_y[SPTX_SPECTRUM_STRIDE];
    float y_scale_num, y_scale_den;  // 310 and CIE_Y_integral * 31 as floats
};

struct DCamera {
    sptx_camera c;
    const float *lens;   // 4 floats per element interface
    const float *pupil;  // 4 floats per exit-pupil bound
    float film_extent[4];  // Film::GetPhysicalExtent(): x0,y0,x1,y1
    int full_res[2];
};

struct DSampler {
    sptx_sampler s;
    const int32_t *primes;
    const int32_t *prime_sums;
    const uint16_t *perms;
    // per dimension: {base, floor((2^32-1)/base), offset into perms, bits(1.f / base)} and the
    // constant tail invBase * perm[0] / (1 - invBase) of ScrambledRadicalInverse
    const uint4 *dim_tab;
    const float *dim_tail;
    const uint64_t *pixel_offset;  // per pixel of sample_bounds (Halton)
    int sb[4];                     // sample bounds x0,y0,x1,y1
};

struct DFilm {
    sptx_film f;
    float *pixels;  // [H][W][32]: 31 bins + filterWeightSum
    int width, height;
};

}  // namespace sptx
