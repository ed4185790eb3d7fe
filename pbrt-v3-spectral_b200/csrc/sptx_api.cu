// sptx_api.cu -- the C ABI of include/sptx.h: context, scene upload, render
// driver, film access.  The only translation unit with device code; the
// kernels live in the .cuh files it includes.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <chrono>
#include <vector>

#include "setup.cuh"
#include "serial.cuh"
#include "lbvh.cuh"

using namespace sptx;

namespace sptx {  // shade.cu, compiled once per arithmetic mode, with and without Light "infinite"
#define SPTX_DECLARE_SHADE(ns)                                                                                  \
    namespace ns {                                                                                              \
    int launch_shade(cudaStream_t stream, int smCount, int64_t nPaths, const DScene &hS, const DSampler &hH,   \
                     const PathArrays &A, const RenderParams &R, int qi, int betaIn, const ShadePlan &plan,    \
                     DeviceCounters *cnt);                                                                      \
    }
SPTX_DECLARE_SHADE(shade_exact)
SPTX_DECLARE_SHADE(shade_fast)
SPTX_DECLARE_SHADE(shade_exact_env)
SPTX_DECLARE_SHADE(shade_fast_env)
#undef SPTX_DECLARE_SHADE
}  // namespace sptx

static std::string g_last_error;

struct sptx_ctx {
    int device = 0;
    std::string err;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
    bool haveScene = false;
    std::vector<void *> sceneAllocs;
    std::vector<void *> pathAllocs;
    // Device buffers of the previous scene, kept across the free -> allocate cycle of a re-upload
    // (sptx_upload_scene): a scene of the same shape takes them back instead of going through
    // cudaFree / cudaMalloc, which are device-wide synchronisation points (measured: sporadic
    // 60-110 ms stalls per upload).  Whatever is not taken back is freed when the upload ends.
    struct Spare {
        void *ptr;
        size_t bytes;
    };
    std::vector<Spare> spares;
    std::vector<Spare> sizes;  // byte size of every live allocation made through dmalloc()
    DScene S;
    DCamera C;
    DSampler H;
    // device copies of S and H: the large kernels take pointers, so that handing the
    // descriptors to non-inlined device functions does not copy them to each thread's stack
    DScene *dS = nullptr;
    DSampler *dH = nullptr;
    DCamera *dC = nullptr;
    // pinned staging for film downloads, and the device pixel list of the last render region
    float *dFilmSplit = nullptr;  // [P][31] contrib + [P] weight, the download layout
    size_t dFilmSplitFloats = 0;
    int metadataStrategy = 0;     // SPTX_METADATA_*: Integrator "metadata" instead of radiance
    float *dLacc = nullptr;       // [P][32] merged spectrum of a sample's bands (num_ca_bands > 1)
    int64_t dLaccPaths = 0;
    int2 *dPixels = nullptr;
    int64_t dPixelsCap = 0;
    int64_t nPixels = 0;          // entries of dPixels in use
    int pixelsKey[12] = {-1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // what dPixels holds: {kind, bounds / partition ...}
    // grow-only device scratch of the entry points that stage host data (ray batches, .dat planes):
    // no cudaMalloc / cudaFree per call (both are device-wide synchronisation points)
    // peer films mapped through CUDA IPC, kept open between merges (opening a handle costs
    // milliseconds; the handle of a film stays the same while its allocation lives)
    struct PeerMap {
        unsigned char handle[SPTX_IPC_HANDLE_BYTES];
        void *ptr;
    };
    std::vector<PeerMap> peerMaps;
    void *scratch[4] = {nullptr, nullptr, nullptr, nullptr};  // [3]: ray-sort keys, order and cub's scratch
    size_t scratchBytes[4] = {0, 0, 0, 0};
    DFilm F;
    PathArrays A;
    sptx_integrator integ;
    int64_t maxPaths = 0;       // requested wavefront size
    int64_t allocatedPaths = 0; // capacity of A
    DeviceCounters *dCnt = nullptr;
    sptx_counters counters;
    bool profileKernels = false;
    bool countVisits = false;
    // K3's arithmetic: the reference's roundings, or FMA contraction + approximate division /
    // square root (csrc/shade.cu).  SPTX_EXACT=1 in the environment makes exact the default.
    bool shadeExact = false;
    ShadePlan shadePlan;
    // "device_bvh": sptx_upload_scene ignores the scene's node array and builds the BVHs on the
    // device (csrc/lbvh.cuh); build_ms = device time of the last build
    bool deviceBvh = false;
    double bvhBuildMs = 0.0;
    // "ray_sort": 0 never, 1 always, 2 (default) batches of kRaySortMin rays and more
    int raySort = 2;
    int lastBatchSorted = 0;
    // "sparse_voxels": 1 forces the sparse form of the spatial light tables (default: above 256 MB)
    bool sparseVoxels = false;
    int64_t voxelRows = 0;
    std::vector<cudaEvent_t> evPool;
    int smCount = 148;
};

static int fail(sptx_ctx *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx)
        ctx->err = buf;
    else
        g_last_error = buf;
    return code;
}

#define CK(call)                                                                      \
    do {                                                                              \
        cudaError_t e_ = (call);                                                      \
        if (e_ != cudaSuccess)                                                        \
            return fail(ctx, SPTX_ECUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                          \
    } while (0)

static int grid_for(const sptx_ctx *ctx, int64_t n, int block = kBlock) {
    int64_t blocks = (n + block - 1) / block;
    int64_t cap = (int64_t)ctx->smCount * 16;  // a multiple of the SM count
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// cudaMalloc, or a spare buffer of exactly this size.
static cudaError_t dmalloc(sptx_ctx *ctx, void **p, size_t bytes) {
    for (size_t i = 0; i < ctx->spares.size(); ++i)
        if (ctx->spares[i].bytes == bytes) {
            *p = ctx->spares[i].ptr;
            ctx->spares.erase(ctx->spares.begin() + (long)i);
            ctx->sizes.push_back({*p, bytes});
            return cudaSuccess;
        }
    cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaSuccess) ctx->sizes.push_back({*p, bytes});
    return e;
}
static size_t forget_size(sptx_ctx *ctx, void *p) {
    for (size_t i = 0; i < ctx->sizes.size(); ++i)
        if (ctx->sizes[i].ptr == p) {
            const size_t b = ctx->sizes[i].bytes;
            ctx->sizes[i] = ctx->sizes.back();
            ctx->sizes.pop_back();
            return b;
        }
    return 0;
}

template <typename T>
static int upload(sptx_ctx *ctx, std::vector<void *> &pool, const T *src, int64_t n, const T **out) {
    *out = nullptr;
    if (n <= 0) {
        // keep kernels safe: a 1-element dummy so pointers are never null
        void *p = nullptr;
        CK(dmalloc(ctx, &p, sizeof(T) > 16 ? sizeof(T) : 16));
        CK(cudaMemsetAsync(p, 0, sizeof(T) > 16 ? sizeof(T) : 16, ctx->stream));
        pool.push_back(p);
        *out = (const T *)p;
        return SPTX_OK;
    }
    void *p = nullptr;
    CK(dmalloc(ctx, &p, (size_t)n * sizeof(T)));
    pool.push_back(p);
    CK(cudaMemcpyAsync(p, src, (size_t)n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *out = (const T *)p;
    return SPTX_OK;
}

template <typename T>
static int dalloc(sptx_ctx *ctx, std::vector<void *> &pool, int64_t n, T **out) {
    void *p = nullptr;
    if (n < 1) n = 1;
    CK(dmalloc(ctx, &p, (size_t)n * sizeof(T)));
    pool.push_back(p);
    *out = (T *)p;
    return SPTX_OK;
}

static void free_pool(sptx_ctx *ctx, std::vector<void *> &pool) {
    for (void *p : pool) {
        forget_size(ctx, p);
        cudaFree(p);
    }
    pool.clear();
}
// The pool's buffers become spares for the allocations that follow; drop_spares() ends the cycle.
static void recycle_pool(sptx_ctx *ctx, std::vector<void *> &pool) {
    for (void *p : pool) {
        const size_t b = forget_size(ctx, p);
        if (b)
            ctx->spares.push_back({p, b});
        else
            cudaFree(p);
    }
    pool.clear();
}
static void drop_spares(sptx_ctx *ctx) {
    for (auto &s : ctx->spares) cudaFree(s.ptr);
    ctx->spares.clear();
}

static int ctx_scratch(sptx_ctx *ctx, int slot, size_t bytes, void **out) {
    if (bytes > ctx->scratchBytes[slot]) {
        CK(cudaStreamSynchronize(ctx->stream));
        if (ctx->scratch[slot]) cudaFree(ctx->scratch[slot]);
        ctx->scratch[slot] = nullptr;
        ctx->scratchBytes[slot] = 0;
        CK(cudaMalloc(&ctx->scratch[slot], bytes));
        ctx->scratchBytes[slot] = bytes;
    }
    *out = ctx->scratch[slot];
    return SPTX_OK;
}

extern "C" int sptx_abi_version(void) { return SPTX_ABI_VERSION; }

extern "C" const char *sptx_last_error(const sptx_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_last_error.c_str();
}

extern "C" int sptx_create(int device, sptx_ctx **out) {
    sptx_ctx *ctx = nullptr;
    if (!out) return fail(nullptr, SPTX_EINVAL, "sptx_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, SPTX_ECUDA,
                    "sptx_create: no CUDA device (%s); there is no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count)
        return fail(nullptr, SPTX_EINVAL, "sptx_create: device %d out of range [0,%d)", device, count);
    ctx = new sptx_ctx();
    ctx->device = device;
    memset(&ctx->S, 0, sizeof(ctx->S));
    memset(&ctx->C, 0, sizeof(ctx->C));
    memset(&ctx->H, 0, sizeof(ctx->H));
    memset(&ctx->F, 0, sizeof(ctx->F));
    memset(&ctx->A, 0, sizeof(ctx->A));
    memset(&ctx->counters, 0, sizeof(ctx->counters));
    auto bail = [&](cudaError_t ce, const char *what) {
        int rc = fail(nullptr, SPTX_ECUDA, "sptx_create: %s: %s", what, cudaGetErrorString(ce));
        delete ctx;
        return rc;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail(e, "props");
    ctx->smCount = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess)
        return bail(e, "cudaStreamCreate");
    cudaEventCreate(&ctx->ev0);
    cudaEventCreate(&ctx->ev1);
    cudaEventCreate(&ctx->ev2);
    cudaEventCreate(&ctx->ev3);
    if ((e = cudaMalloc((void **)&ctx->dCnt, sizeof(DeviceCounters))) != cudaSuccess)
        return bail(e, "cudaMalloc");
    cudaMemsetAsync(ctx->dCnt, 0, sizeof(DeviceCounters), ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    // Traversal keeps a 64-entry stack per thread in local memory.
    cudaDeviceSetLimit(cudaLimitStackSize, 2048);
    ctx->profileKernels = getenv("SPTX_PROFILE_KERNELS") != nullptr;
    {
        const char *ex = getenv("SPTX_EXACT");
        ctx->shadeExact = ex && ex[0] && ex[0] != '0';
    }
    for (int c = 0; c < kShadeClasses; ++c) ctx->shadePlan.variant[c] = -1;
    *out = ctx;
    return SPTX_OK;
}

extern "C" void sptx_destroy(sptx_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    free_pool(ctx, ctx->sceneAllocs);
    free_pool(ctx, ctx->pathAllocs);
    drop_spares(ctx);
    if (ctx->dCnt) cudaFree(ctx->dCnt);
    if (ctx->dS) cudaFree(ctx->dS);
    if (ctx->dFilmSplit) cudaFree(ctx->dFilmSplit);
    if (ctx->dLacc) cudaFree(ctx->dLacc);
    if (ctx->dPixels) cudaFree(ctx->dPixels);
    for (auto &m : ctx->peerMaps) cudaIpcCloseMemHandle(m.ptr);
    for (void *s : ctx->scratch)
        if (s) cudaFree(s);
    if (ctx->dH) cudaFree(ctx->dH);
    if (ctx->dC) cudaFree(ctx->dC);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaEventDestroy(ctx->ev2);
    cudaEventDestroy(ctx->ev3);
    for (cudaEvent_t e : ctx->evPool) cudaEventDestroy(e);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int sptx_set_option(sptx_ctx *ctx, const char *name, int64_t value) {
    if (!ctx || !name) return fail(ctx, SPTX_EINVAL, "sptx_set_option: NULL argument");
    if (!strcmp(name, "count_visits")) {
        ctx->countVisits = value != 0;
        return SPTX_OK;
    }
    if (!strcmp(name, "device_bvh")) {  // takes effect at the next sptx_upload_scene
        ctx->deviceBvh = value != 0;
        return SPTX_OK;
    }
    if (!strcmp(name, "sparse_voxels")) {  // takes effect at the next sptx_upload_scene
        ctx->sparseVoxels = value != 0;
        return SPTX_OK;
    }
    if (!strcmp(name, "ray_sort")) {  // sptx_intersect_batch*: order the rays by origin cell / direction first
        if (value < 0 || value > 2) return fail(ctx, SPTX_EINVAL, "sptx_set_option: ray_sort is 0, 1 or 2");
        ctx->raySort = (int)value;
        return SPTX_OK;
    }
    if (!strcmp(name, "shade_exact")) {  // 1: K3 with the reference's roundings; 0: the fast build of K3
        ctx->shadeExact = value != 0;
        return SPTX_OK;
    }
    return fail(ctx, SPTX_EINVAL, "sptx_set_option: unknown option %s", name);
}

extern "C" double sptx_bvh_build_ms(const sptx_ctx *ctx) { return ctx ? ctx->bvhBuildMs : 0.0; }

extern "C" int sptx_set_wavefront_size(sptx_ctx *ctx, int64_t max_paths) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "ctx is NULL");
    if (max_paths < 0) return fail(ctx, SPTX_EINVAL, "wavefront size must be >= 0");
    ctx->maxPaths = max_paths;
    return SPTX_OK;
}

// ---- scene upload ---------------------------------------------------------------

static int validate_scene(sptx_ctx *ctx, const sptx_scene_desc *sc) {
#define REQ(cond, msg)                                                        \
    do {                                                                      \
        if (!(cond)) return fail(ctx, SPTX_EINVAL, "sptx_upload_scene: %s", msg); \
    } while (0)
    REQ(sc->abi_version == SPTX_ABI_VERSION, "ABI version mismatch");
    REQ(sc->n_bvhs >= 1 && sc->bvhs, "no aggregate");
    REQ(sc->n_prims == 0 || sc->prims, "prims missing");
    REQ(sc->n_nodes == 0 || sc->nodes, "nodes missing");
    for (int64_t i = 0; i < sc->n_bvhs; ++i) {
        const sptx_bvh &b = sc->bvhs[i];
        REQ(b.node_base >= 0 && b.n_nodes >= 0 && (int64_t)b.node_base + b.n_nodes <= sc->n_nodes,
            "bvh node range out of bounds");
        REQ(b.prim_base >= 0 && b.n_prims >= 0 && (int64_t)b.prim_base + b.n_prims <= sc->n_prims,
            "bvh primitive range out of bounds");
        // children come after their parent (checked below), so one forward pass gives every
        // node's depth; the device's traversal stack holds 64 entries like the reference's
        // (accelerators/bvh.cpp:671)
        std::vector<uint8_t> depth((size_t)std::max(b.n_nodes, 1), 0);
        for (int k = 0; k < b.n_nodes; ++k) {
            const sptx_bvh_node &n = sc->nodes[b.node_base + k];
            if (n.n_prims > 0)
                REQ(n.offset >= 0 && n.offset + (int)n.n_prims <= b.n_prims, "leaf range out of bounds");
            else {
                REQ(n.offset > k && n.offset < b.n_nodes && k + 1 < b.n_nodes && n.axis < 3,
                    "interior node malformed");
                REQ(depth[k] < 63, "BVH deeper than the 64-entry traversal stack");
                depth[k + 1] = std::max<uint8_t>(depth[k + 1], (uint8_t)(depth[k] + 1));
                depth[n.offset] = std::max<uint8_t>(depth[n.offset], (uint8_t)(depth[k] + 1));
            }
        }
        if (i > 0)  // an instance's aggregate holds shapes only (no nested instances on the device)
            for (int k = 0; k < b.n_prims; ++k)
                if (sc->prims[b.prim_base + k].kind == SPTX_PRIM_INSTANCE)
                    return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: nested object instances");
    }
    REQ(sc->n_prims < ((int64_t)1 << 31) && sc->n_nodes < ((int64_t)1 << 31), "scene too large for 32-bit ids");
    REQ(sc->integrator.max_depth <= 255,
        "maxdepth above 255 (the path state keeps the bounce count in 8 bits)");
    for (int64_t i = 0; i < sc->n_prims; ++i) {
        const sptx_prim &p = sc->prims[i];
        REQ(p.material >= -1 && p.material < sc->n_materials, "material index out of range");
        REQ(p.light >= -1 && p.light < sc->n_lights, "light index out of range");
        if (p.kind == SPTX_PRIM_TRIANGLE)
            REQ(p.shape >= 0 && p.shape < sc->n_triangles, "triangle index out of range");
        else if (p.kind == SPTX_PRIM_SPHERE)
            REQ(p.shape >= 0 && p.shape < sc->n_spheres, "sphere index out of range");
        else if (p.kind == SPTX_PRIM_INSTANCE)
            REQ(p.shape >= 0 && p.shape < sc->n_instances, "instance index out of range");
        else
            return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: primitive kind %d", p.kind);
    }
    for (int64_t i = 0; i < sc->n_instances; ++i)
        REQ(sc->instances[i].bvh > 0 && sc->instances[i].bvh < sc->n_bvhs, "instance bvh out of range");
    for (int64_t i = 0; i < sc->n_triangles; ++i) {
        const sptx_triangle &t = sc->triangles[i];
        REQ(t.mesh >= 0 && t.mesh < sc->n_meshes, "mesh index out of range");
        const sptx_mesh &m = sc->meshes[t.mesh];
        for (int k = 0; k < 3; ++k) REQ(t.v[k] >= 0 && t.v[k] < m.n_vertices, "vertex index out of range");
    }
    for (int64_t i = 0; i < sc->n_meshes; ++i) {
        const sptx_mesh &m = sc->meshes[i];
        REQ(m.p_base >= 0 && (int64_t)m.p_base + m.n_vertices <= sc->n_positions, "mesh positions");
        REQ(m.n_base < 0 || (int64_t)m.n_base + m.n_vertices <= sc->n_normals, "mesh normals");
        REQ(m.s_base < 0 || (int64_t)m.s_base + m.n_vertices <= sc->n_tangents, "mesh tangents");
        REQ(m.uv_base < 0 || (int64_t)m.uv_base + m.n_vertices <= sc->n_uvs, "mesh uvs");
    }
    for (int64_t i = 0; i < sc->n_materials; ++i) {
        const sptx_material &m = sc->materials[i];
        REQ(m.n_lobes >= 0 && m.n_lobes <= 8 && m.first_lobe >= 0 &&
                (int64_t)m.first_lobe + m.n_lobes <= sc->n_lobes,
            "material lobe range");
        // BSDF::f sums the non-specular lobes only (specular lobes evaluate to zero)
        int nonSpecular = 0;
        for (int l = 0; l < m.n_lobes; ++l) {
            const int ty = sc->lobes[m.first_lobe + l].type;
            if (ty != SPTX_LOBE_SPECULAR_REFL && ty != SPTX_LOBE_SPECULAR_TRANS &&
                ty != SPTX_LOBE_FRESNEL_SPECULAR)
                ++nonSpecular;
        }
        if (nonSpecular > kMaxTerms)
            return fail(ctx, SPTX_EUNSUPPORTED,
                        "sptx_upload_scene: material with %d non-specular lobes (max %d)", nonSpecular,
                        kMaxTerms);
        // image textures: the value of a material's FIRST lobe, a Lambert lobe (the Kd of matte / plastic)
        for (int l = 0; l < m.n_lobes; ++l) {
            const sptx_lobe &lb = sc->lobes[m.first_lobe + l];
            if (lb.R > -2) continue;
            if (l != 0 || lb.type != SPTX_LOBE_LAMBERT)
                return fail(ctx, SPTX_EUNSUPPORTED,
                            "sptx_upload_scene: an image texture on a lobe that is not the material's first, Lambert lobe");
            if (sc->sampler.type == SPTX_SAMPLER_RANDOM || sc->sampler.type == SPTX_SAMPLER_ZEROTWO)
                return fail(ctx, SPTX_EUNSUPPORTED,
                            "sptx_upload_scene: image textures with the tile-serial samplers (random, 02sequence)");
            const int64_t off = -2 - (int64_t)lb.R;
            REQ(sc->tex_floats && off >= 0 && off + 16 <= sc->n_tex_floats, "texture block offset");
            auto geti = [&](int64_t k) {
                int32_t v;
                memcpy(&v, sc->tex_floats + off + k, 4);
                return v;
            };
            const int levels = geti(0);
            REQ(levels >= 1 && levels <= 32 && off + 16 + 3 * (int64_t)levels <= sc->n_tex_floats, "texture levels");
            REQ(geti(1) >= 0 && geti(1) <= 2, "texture wrap mode");
            REQ(geti(8) >= 0 && (int64_t)geti(8) + 7 <= sc->n_spectra, "texture without its seven RGB basis rows");
            REQ(geti(9) >= 0 && (int64_t)geti(9) + 128 <= sc->n_tex_floats, "texture EWA weight table");
            for (int lv = 0; lv < levels; ++lv) {
                const int64_t us = geti(16 + 3 * lv), vs = geti(16 + 3 * lv + 1), to = geti(16 + 3 * lv + 2);
                REQ(us >= 1 && vs >= 1 && to >= 16 && off + to + 3 * us * vs <= sc->n_tex_floats, "texture level");
            }
        }
    }
    for (int64_t i = 0; i < sc->n_lobes; ++i) {
        const sptx_lobe &l = sc->lobes[i];
        REQ(l.type >= SPTX_LOBE_LAMBERT && l.type <= SPTX_LOBE_FRESNEL_BLEND, "lobe type");
        const bool textured = l.R <= -2;  // R = SPTX_TEX_ROW(offset): checked with the materials below
        REQ(l.T >= -1 && l.fr_eta_i >= -1 && l.fr_eta_t >= -1 && l.fr_k >= -1 &&
                l.R < sc->n_spectra && l.T < sc->n_spectra && l.fr_eta_i < sc->n_spectra &&
                l.fr_eta_t < sc->n_spectra && l.fr_k < sc->n_spectra,
            "lobe spectrum row out of range");
        // the rows a lobe of this type reads must exist
        const bool needsT = l.type == SPTX_LOBE_SPECULAR_TRANS || l.type == SPTX_LOBE_FRESNEL_SPECULAR ||
                            l.type == SPTX_LOBE_LAMBERT_TRANS || l.type == SPTX_LOBE_MICROFACET_TRANS ||
                            l.type == SPTX_LOBE_FRESNEL_BLEND;
        const bool needsR = l.type != SPTX_LOBE_SPECULAR_TRANS && l.type != SPTX_LOBE_LAMBERT_TRANS &&
                            l.type != SPTX_LOBE_MICROFACET_TRANS;
        REQ(!needsR || l.R >= 0 || textured, "lobe without its reflectance row");
        REQ(!needsT || l.T >= 0, "lobe without its transmittance row");
        if (l.fresnel == SPTX_FRESNEL_CONDUCTOR &&
            (l.type == SPTX_LOBE_SPECULAR_REFL || l.type == SPTX_LOBE_MICROFACET_REFL))
            REQ(l.fr_eta_i >= 0 && l.fr_eta_t >= 0 && l.fr_k >= 0, "conductor lobe without its Fresnel rows");
    }
    for (int64_t i = 0; i < sc->n_lights; ++i) {
        const sptx_light &l = sc->lights[i];
        REQ(l.spectrum >= 0 && l.spectrum < sc->n_spectra, "light spectrum row out of range");
        if (l.type == SPTX_LIGHT_DIFFUSE_AREA) {
            if (l.shape_kind == SPTX_PRIM_TRIANGLE)
                REQ(l.shape >= 0 && l.shape < sc->n_triangles, "light triangle out of range");
            else if (l.shape_kind == SPTX_PRIM_SPHERE)
                REQ(l.shape >= 0 && l.shape < sc->n_spheres, "light sphere out of range");
            else
                return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: area light shape kind");
        } else if (l.type == SPTX_LIGHT_INFINITE) {
            // the block header, the two matrices, the texels and the distribution must lie inside env_floats
            REQ(sc->env_floats && l.shape >= 0 && (int64_t)l.shape + 40 <= sc->n_env_floats, "infinite light block");
            const float *e = sc->env_floats + l.shape;
            const int64_t W = (int64_t)e[0], H = (int64_t)e[1], nu = (int64_t)e[2], nv = (int64_t)e[3];
            REQ(W >= 1 && H >= 1 && nu >= 1 && nv >= 1 && W <= 65536 && H <= 65536 && nu <= 131072 && nv <= 131072,
                "infinite light map size");
            REQ((int64_t)l.shape + 40 + 3 * W * H + (2 * nv + 2) + nv * (2 * nu + 2) <= sc->n_env_floats,
                "infinite light block truncated");
            REQ(l.spectrum + 7 <= sc->n_spectra, "infinite light without its seven RGB basis rows");
            REQ(l.area > 0, "infinite light without a world radius");
        } else if (l.type != SPTX_LIGHT_POINT)
            return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: light type %d", l.type);
    }
    const sptx_film &f = sc->film;
    REQ(f.cropped[2] >= f.cropped[0] && f.cropped[3] >= f.cropped[1], "film bounds");
    REQ(f.filter_radius[0] > 0 && f.filter_radius[1] > 0, "filter radius");
    if (sc->sampler.type == SPTX_SAMPLER_HALTON) {
        REQ(sc->sampler.n_dims >= 5 && sc->n_primes >= sc->sampler.n_dims &&
                sc->n_prime_sums >= sc->sampler.n_dims && sc->perms,
            "Halton tables missing");
        const int last = sc->sampler.n_dims - 1;
        REQ((int64_t)sc->prime_sums[last] + sc->primes[last] <= sc->n_perms, "Halton permutations short");
    } else if (sc->sampler.type == SPTX_SAMPLER_SOBOL) {
        REQ(sc->sampler.n_dims >= 5 && sc->sampler.n_dims <= 1024 &&
                sc->n_primes >= (int64_t)sc->sampler.n_dims * 52 && sc->n_prime_sums >= 4 * 52 &&
                sc->sampler.log2_resolution >= 0 && sc->sampler.log2_resolution < 26,
            "Sobol tables missing");
    } else if (sc->sampler.type == SPTX_SAMPLER_RANDOM) {
        REQ(sc->sampler.spp > 0, "RandomSampler needs samplesPerPixel");
    } else if (sc->sampler.type == SPTX_SAMPLER_ZEROTWO) {
        REQ(sc->sampler.spp > 0 && (sc->sampler.spp & (sc->sampler.spp - 1)) == 0 && sc->sampler.n_dims >= 0 &&
                sc->sampler.n_dims <= 64,
            "ZeroTwoSequenceSampler needs a power-of-two samplesPerPixel and 0..64 sampled dimensions");
    } else
        return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: sampler type %d", sc->sampler.type);
    if (sc->camera.type == SPTX_CAMERA_REALISTIC)
        REQ(sc->camera.n_elements > 0 && sc->n_lens_floats >= 4 * sc->camera.n_elements &&
                sc->camera.n_pupil_bounds > 0 && sc->n_pupil_floats >= 4 * sc->camera.n_pupil_bounds,
            "realistic camera tables missing");
    else if (sc->camera.type != SPTX_CAMERA_PERSPECTIVE)
        return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: camera type %d", sc->camera.type);
#undef REQ
    return SPTX_OK;
}

// ---- BVH build on the device (csrc/lbvh.cuh) ------------------------------------------------------
// Builds aggregate `b` (primitives [primBase, primBase + n) of the scene tables) into `nodesOut`
// (2n - 1 records), permuting the primitive tables into Morton order.  Scratch comes from `pool`
// (released with the scene).  Returns the root bounds in rootBounds[6].
static int lbvh_build_one(sptx_ctx *ctx, std::vector<void *> &pool, DScene &S, int primBase, int n,
                          const float *dInstBounds, sptx_bvh_node *nodesOut, sptx_prim *primsTmp, float4 *leafvTmp,
                          uint8_t *classTmp, int *idsTmp, float rootBounds[6]) {
    int rc;
    float *bmin, *bmax, *nbmin, *nbmax;
    uint32_t *codes, *codesSorted;
    int *order, *orderSorted, *left, *right, *parent, *flag, *size, *axis, *dfs, *frontier[2], *scal;
    const int nn = 2 * n - 1;
    if ((rc = dalloc(ctx, pool, 3 * (int64_t)n, &bmin)) || (rc = dalloc(ctx, pool, 3 * (int64_t)n, &bmax)) ||
        (rc = dalloc(ctx, pool, 3 * (int64_t)nn, &nbmin)) || (rc = dalloc(ctx, pool, 3 * (int64_t)nn, &nbmax)) ||
        (rc = dalloc(ctx, pool, (int64_t)n, &codes)) || (rc = dalloc(ctx, pool, (int64_t)n, &codesSorted)) ||
        (rc = dalloc(ctx, pool, (int64_t)n, &order)) || (rc = dalloc(ctx, pool, (int64_t)n, &orderSorted)) ||
        (rc = dalloc(ctx, pool, (int64_t)n, &left)) || (rc = dalloc(ctx, pool, (int64_t)n, &right)) ||
        (rc = dalloc(ctx, pool, (int64_t)nn, &parent)) || (rc = dalloc(ctx, pool, (int64_t)n, &flag)) ||
        (rc = dalloc(ctx, pool, (int64_t)nn, &size)) || (rc = dalloc(ctx, pool, (int64_t)n, &axis)) ||
        (rc = dalloc(ctx, pool, (int64_t)nn, &dfs)) || (rc = dalloc(ctx, pool, (int64_t)n, &frontier[0])) ||
        (rc = dalloc(ctx, pool, (int64_t)n, &frontier[1])) || (rc = dalloc(ctx, pool, 8, &scal)))
        return rc;
    cudaStream_t st = ctx->stream;
    const int B = 256, G = (n + B - 1) / B;
    // centroid bounds as ordered ints: min = +inf, max = -inf
    const int init[8] = {0x7f800000, 0x7f800000, 0x7f800000, (int)(0xff800000u ^ 0x7fffffffu),
                         (int)(0xff800000u ^ 0x7fffffffu), (int)(0xff800000u ^ 0x7fffffffu), 0, 0};
    CK(cudaMemcpyAsync(scal, init, sizeof(init), cudaMemcpyHostToDevice, st));
    k_lbvh_prim_bounds<<<G, B, 0, st>>>(S, primBase, n, dInstBounds, bmin, bmax, scal, scal + 3);
    k_lbvh_morton<<<G, B, 0, st>>>(n, bmin, bmax, scal, scal + 3, codes, order);
    size_t tmpBytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, codes, codesSorted, order, orderSorted, n, 0, 30, st);
    void *tmp = nullptr;
    if ((rc = dalloc(ctx, pool, (int64_t)tmpBytes + 16, (char **)&tmp))) return rc;
    cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, codes, codesSorted, order, orderSorted, n, 0, 30, st);
    CK(cudaMemsetAsync(flag, 0, (size_t)n * sizeof(int), st));
    k_lbvh_tree<<<G, B, 0, st>>>(n, codesSorted, left, right, parent, axis);
    k_lbvh_fit<<<G, B, 0, st>>>(n, orderSorted, bmin, bmax, left, right, parent, flag, nbmin, nbmax, size);
    // depth-first numbering, one launch per level
    const int zero = 0;
    CK(cudaMemcpyAsync(dfs, &zero, sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(frontier[0], &zero, sizeof(int), cudaMemcpyHostToDevice, st));
    int nFront = 1, cur = 0;
    for (int level = 0; nFront > 0; ++level) {
        if (level > 128) return fail(ctx, SPTX_ECUDA, "device BVH build: tree deeper than 128 levels");
        CK(cudaMemsetAsync(scal + 6, 0, sizeof(int), st));
        k_lbvh_number<<<(nFront + B - 1) / B, B, 0, st>>>(n, nFront, frontier[cur], frontier[cur ^ 1], scal + 6, left,
                                                         right, size, dfs);
        CK(cudaMemcpyAsync(&nFront, scal + 6, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        cur ^= 1;
    }
    k_lbvh_emit<<<(nn + B - 1) / B, B, 0, st>>>(n, left, right, axis, dfs, nbmin, nbmax, nodesOut);
    k_lbvh_permute<<<G, B, 0, st>>>(n, primBase, orderSorted, S.prims, S.leafv, S.prim_class, S.prim_ids, primsTmp,
                                    leafvTmp, classTmp, idsTmp);
    // permuted tables back in place (this aggregate's range only)
    CK(cudaMemcpyAsync(const_cast<sptx_prim *>(S.prims) + primBase, primsTmp + primBase, (size_t)n * sizeof(sptx_prim),
                       cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(const_cast<float4 *>(S.leafv) + 3 * (size_t)primBase, leafvTmp + 3 * (size_t)primBase,
                       (size_t)n * 3 * sizeof(float4), cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(const_cast<uint8_t *>(S.prim_class) + primBase, classTmp + primBase, (size_t)n,
                       cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(const_cast<int *>(S.prim_ids) + primBase, idsTmp + primBase, (size_t)n * sizeof(int),
                       cudaMemcpyDeviceToDevice, st));
    float rb[6];
    CK(cudaMemcpyAsync(rb, nbmin, 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(rb + 3, nbmax, 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    memcpy(rootBounds, rb, sizeof(rb));
    return SPTX_OK;
}

__global__ void k_iota(int *p, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = (int)i;
}
// All aggregates of the scene, object BVHs first (the top level needs their root bounds).
// Replaces S.nodes / S.bvhs; sets S.prim_ids; S.wb_* = the top level's root bounds.
static int lbvh_build_scene(sptx_ctx *ctx, std::vector<void *> &pool, DScene &S, const sptx_scene_desc *sc) {
    int rc;
    const int64_t nPrims = sc->n_prims;
    int *ids;
    if ((rc = dalloc(ctx, pool, std::max<int64_t>(nPrims, 1), &ids))) return rc;
    k_iota<<<grid_for(ctx, std::max<int64_t>(nPrims, 1)), kBlock, 0, ctx->stream>>>(ids, nPrims);
    S.prim_ids = ids;
    sptx_prim *primsTmp;
    float4 *leafvTmp;
    uint8_t *classTmp;
    int *idsTmp;
    if ((rc = dalloc(ctx, pool, std::max<int64_t>(nPrims, 1), &primsTmp)) ||
        (rc = dalloc(ctx, pool, 3 * std::max<int64_t>(nPrims, 1), &leafvTmp)) ||
        (rc = dalloc(ctx, pool, std::max<int64_t>(nPrims, 1), &classTmp)) ||
        (rc = dalloc(ctx, pool, std::max<int64_t>(nPrims, 1), &idsTmp)))
        return rc;
    std::vector<sptx_bvh> bvhs(sc->bvhs, sc->bvhs + sc->n_bvhs);
    int64_t totalNodes = 0;
    for (auto &b : bvhs) {
        b.node_base = (int32_t)totalNodes;
        b.n_nodes = b.n_prims >= 2 ? 2 * b.n_prims - 1 : 0;  // 0 or 1 primitive: node-less aggregate
        totalNodes += b.n_nodes;
    }
    sptx_bvh_node *nodes;
    if ((rc = dalloc(ctx, pool, std::max<int64_t>(totalNodes, 1), &nodes))) return rc;
    std::vector<float> instBounds(6 * (size_t)sc->n_bvhs, 0.f);
    float *dInstBounds;
    if ((rc = dalloc(ctx, pool, 6 * (int64_t)sc->n_bvhs, &dInstBounds))) return rc;
    cudaEventRecord(ctx->ev2, ctx->stream);
    for (int pass = 0; pass < 2; ++pass)   // object aggregates (index > 0), then the scene's
        for (int64_t k = (pass == 0 ? 1 : 0); k < (pass == 0 ? sc->n_bvhs : 1); ++k) {
            const sptx_bvh &b = bvhs[(size_t)k];
            float rb[6] = {0, 0, 0, 0, 0, 0};
            if (pass == 1)
                CK(cudaMemcpyAsync(dInstBounds, instBounds.data(), instBounds.size() * sizeof(float),
                                   cudaMemcpyHostToDevice, ctx->stream));
            if (b.n_prims >= 2) {
                if ((rc = lbvh_build_one(ctx, pool, S, b.prim_base, b.n_prims, dInstBounds, nodes + b.node_base,
                                         primsTmp, leafvTmp, classTmp, idsTmp, rb)))
                    return rc;
            } else if (b.n_prims == 1) {
                // one primitive: its bounds through the same kernel (scratch of 2 x 3 floats + 6 ints)
                float *bb;
                int *cc;
                if ((rc = dalloc(ctx, pool, 6, &bb)) || (rc = dalloc(ctx, pool, 6, &cc))) return rc;
                k_lbvh_prim_bounds<<<1, 32, 0, ctx->stream>>>(S, b.prim_base, 1, dInstBounds, bb, bb + 3, cc, cc + 3);
                CK(cudaMemcpyAsync(rb, bb, 6 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
                CK(cudaStreamSynchronize(ctx->stream));
            }
            for (int a = 0; a < 6; ++a) instBounds[6 * (size_t)k + a] = rb[a];
        }
    cudaEventRecord(ctx->ev3, ctx->stream);
    CK(cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3);
    ctx->bvhBuildMs = ms;
    S.nodes = reinterpret_cast<const float4 *>(nodes);
    const sptx_bvh *dB;
    if ((rc = upload(ctx, pool, bvhs.data(), (int64_t)bvhs.size(), &dB))) return rc;
    S.bvhs = dB;
    for (int a = 0; a < 3; ++a) {
        S.wb_min[a] = instBounds[a];
        S.wb_max[a] = instBounds[3 + a];
    }
    return SPTX_OK;
}

// SampledSpectrum::y() on the host (core/spectrum.h:415-421).
static float host_y(const sptx_scene_desc *sc, const float *row, float scale) {
    float yy = 0.f;
    for (int i = 0; i < SPTX_NBINS; ++i) yy += sc->cie_y[i] * (row[i] * scale);
    yy = (yy < 0) ? 0 : yy;
    float num = (float)(705 - 395);
    float den = (float)(sc->cie_y_integral * (float)SPTX_NBINS);
    return yy * num / den;
}

extern "C" int sptx_upload_scene(sptx_ctx *ctx, const sptx_scene_desc *sc) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_upload_scene: ctx is NULL");
    if (!sc) return fail(ctx, SPTX_EINVAL, "sptx_upload_scene: scene is NULL");
    // SPTX_TRACE_UPLOAD=1: wall time of each phase (stream drained at every mark), to stderr
    const char *traceEnv = getenv("SPTX_TRACE_UPLOAD");  // "2": host time only, no extra syncs
    const bool trace = traceEnv != nullptr;
    const bool traceSync = trace && traceEnv[0] != '2';
    auto tPrev = std::chrono::steady_clock::now();
    int rc = validate_scene(ctx, sc);
    if (rc) return rc;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    auto mark = [&](const char *what) {
        if (!trace) return;
        if (traceSync) cudaStreamSynchronize(ctx->stream);
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "sptx_upload_scene: %-18s %8.3f ms\n", what,
                std::chrono::duration<double, std::milli>(now - tPrev).count());
        tPrev = now;
    };
    mark("validate+drain");
    recycle_pool(ctx, ctx->sceneAllocs);
    ctx->haveScene = false;
    // (the pixel list of the previous scene stays: its key holds every bound it was made from)
    mark("free");
    std::vector<void *> &pool = ctx->sceneAllocs;
    DScene &S = ctx->S;
    memset(&S, 0, sizeof(S));

    const sptx_bvh_node *dNodes;
    if ((rc = upload(ctx, pool, sc->nodes, sc->n_nodes, &dNodes))) return rc;
    S.nodes = reinterpret_cast<const float4 *>(dNodes);
    if ((rc = upload(ctx, pool, sc->bvhs, sc->n_bvhs, &S.bvhs))) return rc;
    if ((rc = upload(ctx, pool, sc->prims, sc->n_prims, &S.prims))) return rc;
    if ((rc = upload(ctx, pool, sc->meshes, sc->n_meshes, &S.meshes))) return rc;
    if ((rc = upload(ctx, pool, sc->triangles, sc->n_triangles, &S.triangles))) return rc;
    if ((rc = upload(ctx, pool, sc->positions, 3 * sc->n_positions, &S.positions))) return rc;
    if ((rc = upload(ctx, pool, sc->normals, 3 * sc->n_normals, &S.normals))) return rc;
    if ((rc = upload(ctx, pool, sc->tangents, 3 * sc->n_tangents, &S.tangents))) return rc;
    if ((rc = upload(ctx, pool, sc->uvs, 2 * sc->n_uvs, &S.uvs))) return rc;
    if ((rc = upload(ctx, pool, sc->spheres, sc->n_spheres, &S.spheres))) return rc;
    if ((rc = upload(ctx, pool, sc->instances, sc->n_instances, &S.instances))) return rc;
    if ((rc = upload(ctx, pool, sc->materials, sc->n_materials, &S.materials))) return rc;
    if ((rc = upload(ctx, pool, sc->lobes, sc->n_lobes, &S.lobes))) return rc;
    {
        // shading class of a material = its list of (lobe type, fresnel kind); primitives
        // carry the class so K2 can sort the paths it hands to K3
        std::vector<std::vector<int>> sigs;
        std::vector<uint8_t> matClass((size_t)sc->n_materials, 1);
        for (int64_t m = 0; m < sc->n_materials; ++m) {
            std::vector<int> sig;
            for (int l = 0; l < sc->materials[m].n_lobes; ++l) {
                const sptx_lobe &L = sc->lobes[sc->materials[m].first_lobe + l];
                sig.push_back(L.type * 16 + L.fresnel + (L.R <= -2 ? 8 : 0));
            }
            size_t k = 0;
            while (k < sigs.size() && sigs[k] != sig) ++k;
            if (k == sigs.size()) sigs.push_back(sig);
            matClass[(size_t)m] = (uint8_t)(1 + k % (kShadeClasses - 1));
        }
        std::vector<uint8_t> primClass((size_t)std::max<int64_t>(sc->n_prims, 1), 1);
        // per class: the lobe types and the largest number of non-specular lobes of the materials
        // its primitives use -> the K3 variant that shades the class
        int classLobes[kShadeClasses] = {0}, classTerms[kShadeClasses] = {0};
        bool classUsed[kShadeClasses] = {false};
        for (int64_t i = 0; i < sc->n_prims; ++i) {
            const int m = sc->prims[i].material;
            const int cls = m >= 0 ? matClass[(size_t)m] : 1;
            primClass[(size_t)i] = (uint8_t)cls;
            if (sc->prims[i].kind == SPTX_PRIM_INSTANCE) continue;  // never the primitive of a hit
            classUsed[cls] = true;
            if (m < 0) continue;
            int nonSpecular = 0;
            for (int l = 0; l < sc->materials[m].n_lobes; ++l) {
                const int ty = sc->lobes[sc->materials[m].first_lobe + l].type;
                classLobes[cls] |= 1 << ty;
                if (sc->lobes[sc->materials[m].first_lobe + l].R <= -2) classLobes[cls] |= kTexturedLobes;
                if (!((kSpecularLobes >> ty) & 1)) ++nonSpecular;
            }
            classTerms[cls] = std::max(classTerms[cls], nonSpecular);
        }
        for (int c = 0; c < kShadeClasses; ++c) {
            ctx->shadePlan.variant[c] = (c > 0 && classUsed[c]) ? shade_variant_for(classLobes[c], classTerms[c]) : -1;
            if (c > 0 && classUsed[c] && ctx->shadePlan.variant[c] < 0)
                return fail(ctx, SPTX_EUNSUPPORTED,
                            "sptx_upload_scene: a textured material shares its shading class with lobe types the "
                            "textured variant does not evaluate (more than %d lobe signatures in the scene)",
                            kShadeClasses - 1);
        }
        if ((rc = upload(ctx, pool, primClass.data(), (int64_t)primClass.size(), &S.prim_class))) return rc;
    }
    if ((rc = upload(ctx, pool, sc->spectra, SPTX_SPECTRUM_STRIDE * sc->n_spectra, &S.spectra))) return rc;
    if ((rc = upload(ctx, pool, sc->lights, sc->n_lights, &S.lights))) return rc;
    S.n_lights = (int)sc->n_lights;
    if ((rc = upload(ctx, pool, sc->env_floats, sc->n_env_floats, &S.env))) return rc;
    if ((rc = upload(ctx, pool, sc->tex_floats, sc->n_tex_floats, &S.tex))) return rc;
    S.n_infinite = 0;
    for (int64_t i = 0; i < sc->n_lights; ++i)
        if (sc->lights[i].type == SPTX_LIGHT_INFINITE) {
            if (S.n_infinite == 4)
                return fail(ctx, SPTX_EUNSUPPORTED, "sptx_upload_scene: more than 4 infinite lights");
            S.infinite[S.n_infinite++] = (int)i;
        }
    S.n_bvhs = (int)sc->n_bvhs;
    S.lights_top_level = sc->n_bvhs > 0;
    for (int64_t p = 0; p < sc->n_prims && S.lights_top_level; ++p)
        if (sc->prims[p].light >= 0 &&
            (p < sc->bvhs[0].prim_base || p >= (int64_t)sc->bvhs[0].prim_base + sc->bvhs[0].n_prims))
            S.lights_top_level = 0;
    for (int i = 0; i < SPTX_SPECTRUM_STRIDE; ++i) S.cie_y[i] = i < SPTX_NBINS ? sc->cie_y[i] : 0.f;
    S.y_scale_num = (float)(705 - 395);  // sampledLambdaEnd - sampledLambdaStart
    S.y_scale_den = (float)(sc->cie_y_integral * (float)SPTX_NBINS);

    // leaf blocks
    float4 *leafv;
    if ((rc = dalloc(ctx, pool, 3 * std::max<int64_t>(sc->n_prims, 1), &leafv))) return rc;
    if (sc->n_prims > 0)
        k_build_leafv<<<grid_for(ctx, sc->n_prims), kBlock, 0, ctx->stream>>>(
            S.prims, sc->n_prims, S.triangles, S.meshes, S.positions, leafv);
    S.leafv = leafv;

    // scene.WorldBound() == the root node's bounds (accelerators/bvh.cpp:227-229)
    if (sc->bvhs[0].n_nodes > 0) {
        const sptx_bvh_node &root = sc->nodes[sc->bvhs[0].node_base];
        for (int i = 0; i < 3; ++i) {
            S.wb_min[i] = root.bmin[i];
            S.wb_max[i] = root.bmax[i];
        }
    }

    S.prim_ids = nullptr;
    ctx->bvhBuildMs = 0.0;
    if (ctx->deviceBvh) {
        if ((rc = lbvh_build_scene(ctx, pool, S, sc))) return rc;
        mark("device BVH build");
    }

    // ---- light selection (core/lightdistrib.cpp:48-66) ----
    ctx->integ = sc->integrator;
    if (ctx->integ.max_depth < 0) ctx->integ.max_depth = 0;  // (path.cpp:88-112: the first vertex is always visited)
    if (sc->metadata_strategy > SPTX_METADATA_COORDINATES)
        return fail(ctx, SPTX_EINVAL, "sptx_upload_scene: metadata strategy %u", sc->metadata_strategy);
    ctx->metadataStrategy = (int)sc->metadata_strategy;
    const int nL = S.n_lights;
    int strategy = sc->integrator.light_strategy;
    if (nL == 1) strategy = SPTX_LIGHTS_UNIFORM;
    S.light_strategy = strategy;
    S.n_voxels[0] = S.n_voxels[1] = S.n_voxels[2] = 1;
    {
        // Uniform: func = 1; Power: func = light->Power().y()
        std::vector<float> tab((size_t)2 * nL + 1, 0.f);
        for (int j = 0; j < nL; ++j) {
            float v = 1.f;
            if (strategy == SPTX_LIGHTS_POWER && sc->lights[j].type == SPTX_LIGHT_INFINITE)
                v = sc->env_floats[sc->lights[j].shape + 4];  // InfiniteAreaLight::Power().y(), from the flattener
            else if (strategy == SPTX_LIGHTS_POWER) {
                const sptx_light &l = sc->lights[j];
                const float *row = sc->spectra + (size_t)l.spectrum * SPTX_SPECTRUM_STRIDE;
                float tmp[SPTX_NBINS];
                for (int b = 0; b < SPTX_NBINS; ++b) {
                    if (l.type == SPTX_LIGHT_DIFFUSE_AREA)  // (twoSided ? 2 : 1) * Lemit * area * Pi
                        tmp[b] = row[b] * (l.two_sided ? 2.f : 1.f) * l.area * kPi;
                    else  // 4 * Pi * I
                        tmp[b] = row[b] * (4.f * kPi);
                }
                v = host_y(sc, tmp, 1.f);
            }
            tab[j] = v;
        }
        float funcInt = 0.f;
        if (nL > 0) {
            float *cdf = tab.data() + nL;
            cdf[0] = 0.f;
            for (int j = 1; j < nL + 1; ++j) cdf[j] = cdf[j - 1] + tab[j - 1] / (float)nL;
            funcInt = cdf[nL];
            if (funcInt == 0)
                for (int j = 1; j < nL + 1; ++j) cdf[j] = (float)j / (float)nL;
            else
                for (int j = 1; j < nL + 1; ++j) cdf[j] /= funcInt;
        }
        S.light_func_int = funcInt;
        if ((rc = upload(ctx, pool, tab.data(), (int64_t)tab.size(), &S.light_func))) return rc;
    }
    S.voxel_tables = S.light_func;
    if (strategy == SPTX_LIGHTS_SPATIAL && nL > 0) {
        // SpatialLightDistribution::SpatialLightDistribution, lightdistrib.cpp:91-111
        float diag[3] = {S.wb_max[0] - S.wb_min[0], S.wb_max[1] - S.wb_min[1], S.wb_max[2] - S.wb_min[2]};
        int me = (diag[0] > diag[1] && diag[0] > diag[2]) ? 0 : ((diag[1] > diag[2]) ? 1 : 2);
        float bmax = diag[me];
        const int maxVoxels = 64;
        int64_t nVox = 1;
        for (int i = 0; i < 3; ++i) {
            S.n_voxels[i] = std::max(1, (int)std::round(diag[i] / bmax * (float)maxVoxels));
            nVox *= S.n_voxels[i];
        }
        const int64_t rowFloats = 2 * (int64_t)nL + 2;
        // The reference fills a voxel's row when a path first asks for it (a hash table under a lock,
        // lightdistrib.cpp:135-219).  Here rows are built up front: for every voxel while that is small
        // (256 MB), else only for the voxels that can hold a surface point -- the ones the leaves of
        // the scene's BVH overlap (k_voxel_mark), typically a few per cent of the grid.
        int64_t nRows = nVox;
        const int *voxList = nullptr;
        S.voxel_rows = nullptr;
        const int64_t topNodes = sc->n_bvhs > 0 ? (ctx->deviceBvh ? std::max(0, 2 * sc->bvhs[0].n_prims - 1) : sc->bvhs[0].n_nodes) : 0;
        if (((double)nVox * rowFloats * 4.0 > 256e6 || ctx->sparseVoxels) && topNodes > 0) {
            int *flags, *scan, *rows, *list;
            if ((rc = dalloc(ctx, pool, nVox, &flags)) || (rc = dalloc(ctx, pool, nVox, &scan)) ||
                (rc = dalloc(ctx, pool, nVox, &rows)) || (rc = dalloc(ctx, pool, nVox, &list)))
                return rc;
            CK(cudaMemsetAsync(flags, 0, (size_t)nVox * sizeof(int), ctx->stream));
            k_voxel_mark<<<grid_for(ctx, topNodes), kBlock, 0, ctx->stream>>>(S, ctx->deviceBvh ? 0 : sc->bvhs[0].node_base, (int)topNodes, flags);
            size_t tmpBytes = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, tmpBytes, flags, scan, (int)nVox, ctx->stream);
            void *tmp;
            if ((rc = dalloc(ctx, pool, (int64_t)tmpBytes + 16, (unsigned char **)&tmp))) return rc;
            cub::DeviceScan::ExclusiveSum(tmp, tmpBytes, flags, scan, (int)nVox, ctx->stream);
            k_voxel_rows<<<grid_for(ctx, nVox), kBlock, 0, ctx->stream>>>(flags, scan, nVox, rows, list);
            int lastScan = 0, lastFlag = 0;
            CK(cudaMemcpyAsync(&lastScan, scan + nVox - 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaMemcpyAsync(&lastFlag, flags + nVox - 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            nRows = (int64_t)lastScan + lastFlag;
            voxList = list;
            S.voxel_rows = rows;
        }
        ctx->voxelRows = nRows;
        const double bytes = (double)nRows * rowFloats * 4.0;
        if (bytes > 16e9)
            return fail(ctx, SPTX_EUNSUPPORTED,
                        "sptx_upload_scene: spatial light distribution for %d lights needs %.1f GB "
                        "(%lld of %lld voxels hold a surface); use lightsamplestrategy \"power\" or \"uniform\"",
                        nL, bytes * 1e-9, (long long)nRows, (long long)nVox);
        // RadicalInverse(d, i), d = 0..4, i = 0..127 (bases 2, 3, 5, 7, 11)
        std::vector<float> h5(5 * 128);
        const uint32_t bases[5] = {2, 3, 5, 7, 11};
        for (int i = 0; i < 128; ++i) {
            h5[5 * i] = (float)((double)reverse_bits64((uint64_t)i) * 0x1p-64);
            for (int d = 1; d < 5; ++d) h5[5 * i + d] = radical_inverse_base((uint64_t)i, bases[d]);
        }
        const float *dH5;
        if ((rc = upload(ctx, pool, h5.data(), (int64_t)h5.size(), &dH5))) return rc;
        float *tables;
        if ((rc = dalloc(ctx, pool, std::max<int64_t>(nRows, 1) * rowFloats, &tables))) return rc;
        S.voxel_tables = tables;
        mark("arrays");
        if (nRows > 0)
            k_voxel_tables<<<grid_for(ctx, nRows, 64), 64, 0, ctx->stream>>>(S, dH5, tables, voxList, nRows);
        mark("voxel tables");
    }

    // ---- camera ----
    DCamera &C = ctx->C;
    memset(&C, 0, sizeof(C));
    C.c = sc->camera;
    C.full_res[0] = sc->film.full_res[0];
    C.full_res[1] = sc->film.full_res[1];
    if ((rc = upload(ctx, pool, sc->lens_elements, sc->n_lens_floats, &C.lens))) return rc;
    if ((rc = upload(ctx, pool, sc->exit_pupil_bounds, sc->n_pupil_floats, &C.pupil))) return rc;
    {
        // Film::GetPhysicalExtent, core/film.cpp:94-99
        float aspect = (float)sc->film.full_res[1] / (float)sc->film.full_res[0];
        float d = sc->film.diagonal;
        float x = std::sqrt(d * d / (1 + aspect * aspect));
        float y = aspect * x;
        C.film_extent[0] = -x / 2;
        C.film_extent[1] = -y / 2;
        C.film_extent[2] = x / 2;
        C.film_extent[3] = y / 2;
    }

    mark("camera");

    // ---- sampler ----
    DSampler &H = ctx->H;
    memset(&H, 0, sizeof(H));
    H.s = sc->sampler;
    for (int i = 0; i < 4; ++i) H.sb[i] = sc->film.sample_bounds[i];
    if (sc->sampler.type == SPTX_SAMPLER_HALTON) {
    if ((rc = upload(ctx, pool, sc->primes, (int64_t)sc->sampler.n_dims, &H.primes))) return rc;
    if ((rc = upload(ctx, pool, sc->prime_sums, (int64_t)sc->sampler.n_dims, &H.prime_sums))) return rc;
    if ((rc = upload(ctx, pool, sc->perms, sc->n_perms, &H.perms))) return rc;
    {
        const int nd = std::max(1, (int)sc->sampler.n_dims);
        uint4 *tab;
        float *tail;
        if ((rc = dalloc(ctx, pool, (int64_t)nd, &tab))) return rc;
        if ((rc = dalloc(ctx, pool, (int64_t)nd, &tail))) return rc;
        if (sc->sampler.n_dims > 0)
            k_halton_dims<<<(nd + 127) / 128, 128, 0, ctx->stream>>>(H.primes, H.prime_sums, H.perms,
                                                                      (int)sc->sampler.n_dims, tab, tail);
        H.dim_tab = tab;
        H.dim_tail = tail;
    }
    {
        int w = std::max(0, H.sb[2] - H.sb[0]), h = std::max(0, H.sb[3] - H.sb[1]);
        uint64_t *off;
        if ((rc = dalloc(ctx, pool, (int64_t)w * h, &off))) return rc;
        if ((int64_t)w * h > 0)
            k_pixel_offsets<<<grid_for(ctx, (int64_t)w * h), kBlock, 0, ctx->stream>>>(
                H.s, H.sb[0], H.sb[1], w, h, off);
        H.pixel_offset = off;
    }
    }  // Halton
    if (sc->sampler.type == SPTX_SAMPLER_SOBOL) {
        if ((rc = upload(ctx, pool, sc->primes, (int64_t)sc->sampler.n_dims * 52, &H.primes))) return rc;
        if ((rc = upload(ctx, pool, sc->prime_sums, (int64_t)4 * 52, &H.prime_sums))) return rc;
    }

    mark("sampler");

    // ---- film ----
    DFilm &F = ctx->F;
    memset(&F, 0, sizeof(F));
    F.f = sc->film;
    F.width = std::max(0, sc->film.cropped[2] - sc->film.cropped[0]);
    F.height = std::max(0, sc->film.cropped[3] - sc->film.cropped[1]);
    if ((rc = dalloc(ctx, pool, (int64_t)F.width * F.height * SPTX_SPECTRUM_STRIDE, &F.pixels))) return rc;
    CK(cudaMemsetAsync(F.pixels, 0, (size_t)std::max<int64_t>(1, (int64_t)F.width * F.height) *
                                        SPTX_SPECTRUM_STRIDE * sizeof(float), ctx->stream));
    if (!ctx->dS) {
        CK(cudaMalloc(&ctx->dS, sizeof(DScene)));
        CK(cudaMalloc(&ctx->dH, sizeof(DSampler)));
        CK(cudaMalloc(&ctx->dC, sizeof(DCamera)));
    }
    ctx->S.camera = ctx->dC;
    CK(cudaMemcpyAsync(ctx->dC, &ctx->C, sizeof(DCamera), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->dS, &ctx->S, sizeof(DScene), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->dH, &ctx->H, sizeof(DSampler), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    mark("film+drain");
    drop_spares(ctx);  // buffers of the previous scene this one did not take back
    mark("drop spares");
    ctx->haveScene = true;
    return SPTX_OK;
}

// ---- render driver ------------------------------------------------------------------

__global__ void k_set_counters(int *counters, int a, int va, int b, int vb) {
    if (a >= 0) counters[a] = va;
    if (b >= 0) counters[b] = vb;
}
// start of a bounce: empty the queues this bounce fills
__global__ void k_reset_bounce(int *counters, int qNext) {
    if (threadIdx.x < kShadeClasses) counters[kClassCounter0 + threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        counters[qNext] = 0;
        counters[2] = 0;
        counters[4] = 0;
    }
}

static int ensure_paths(sptx_ctx *ctx, int64_t P) {
    if (ctx->allocatedPaths >= P) return SPTX_OK;
    CK(cudaStreamSynchronize(ctx->stream));
    free_pool(ctx, ctx->pathAllocs);
    ctx->allocatedPaths = 0;
    std::vector<void *> &pool = ctx->pathAllocs;
    PathArrays &A = ctx->A;
    memset(&A, 0, sizeof(A));
    A.capacity = P;
    int rc;
#define DA(field, count) \
    if ((rc = dalloc(ctx, pool, (int64_t)(count), &A.field))) return rc
    DA(rayA, P);
    DA(rayB, P);
    DA(hitA, P);
    DA(hitB, P);
    DA(beta[0], P * SPTX_SPECTRUM_STRIDE);
    DA(beta[1], P * SPTX_SPECTRUM_STRIDE);
    DA(L, P * SPTX_SPECTRUM_STRIDE);
    DA(sindex, P);
    DA(info, P);
    DA(pfilm, P);
    DA(shA, P);
    DA(shB, P);
    DA(misA, P);
    DA(misB, P);
    DA(fa, P * SPTX_SPECTRUM_STRIDE);
    DA(fb, P * SPTX_SPECTRUM_STRIDE);
    DA(q[0], P);
    DA(q[1], P);
    DA(qNee, P);
    DA(qMis, P);
    DA(qc, P * kShadeClasses);
    DA(counters, 16);
    DA(spill, P);
#undef DA
    CK(cudaMemsetAsync(A.counters, 0, 16 * sizeof(int), ctx->stream));
    ctx->allocatedPaths = P;
    return SPTX_OK;
}

// Samplers whose values come from one RNG stream per tile (RandomSampler, ZeroTwoSequenceSampler).
static bool tile_serial(const sptx_ctx *ctx) {
    return ctx->H.s.type == SPTX_SAMPLER_RANDOM || ctx->H.s.type == SPTX_SAMPLER_ZEROTWO;
}

// RandomSampler: the tiles of SamplerIntegrator::Render, one single-lane warp each (serial.cuh).
// The stream of a tile is stateful, so only the scene's whole sample range can be rendered.
static int render_serial(sptx_ctx *ctx, const int32_t *bounds, int part, int nParts, int s0, int s1) {
    if (s0 != 0 || s1 != ctx->H.s.spp)
        return fail(ctx, SPTX_EUNSUPPORTED,
                    "render: RandomSampler / ZeroTwoSequenceSampler streams run through all %d samples of a tile; sample range "
                    "[%d,%d) cannot be rendered on its own", ctx->H.s.spp, s0, s1);
    if (ctx->metadataStrategy)
        return fail(ctx, SPTX_EUNSUPPORTED,
                    "render: Integrator \"metadata\" with the RandomSampler is not on the device path");
    if (ctx->integ.num_ca_bands > 1)
        return fail(ctx, SPTX_EUNSUPPORTED,
                    "render: SpectralPathIntegrator bands (numCABands = %d) with the RandomSampler are "
                    "not on the device path", (int)ctx->integ.num_ca_bands);
    SerialParams P;
    for (int i = 0; i < 4; ++i) {
        P.sb[i] = ctx->F.f.sample_bounds[i];
        P.pb[i] = ctx->integ.pixel_bounds[i];
    }
    if (bounds) {
        P.pb[0] = std::max(P.pb[0], bounds[0]);
        P.pb[1] = std::max(P.pb[1], bounds[1]);
        P.pb[2] = std::min(P.pb[2], bounds[2]);
        P.pb[3] = std::min(P.pb[3], bounds[3]);
    }
    P.nTilesX = (P.sb[2] - P.sb[0] + 15) / 16;
    P.nTilesY = (P.sb[3] - P.sb[1] + 15) / 16;
    P.spp = ctx->H.s.spp;
    P.maxDepth = ctx->integ.max_depth;
    P.rrThreshold = ctx->integ.rr_threshold;
    P.wavelength = 550.f;
    P.nDims = 0;
    P.tables = nullptr;
    const int nTiles = P.nTilesX * P.nTilesY;
    if (nTiles <= 0) return SPTX_OK;
    const bool zeroTwo = ctx->H.s.type == SPTX_SAMPLER_ZEROTWO;
    if (zeroTwo) {
        // per tile: the pixel's sample tables, [nDims][spp] floats + [nDims][spp] points
        P.nDims = ctx->H.s.n_dims;
        void *t = nullptr;
        int rc = ctx_scratch(ctx, 1, std::max<size_t>(16, (size_t)nTiles * P.nDims * P.spp * 3 * sizeof(float)), &t);
        if (rc) return rc;
        P.tables = (float *)t;
    }
    cudaEventRecord(ctx->ev0, ctx->stream);
    if (zeroTwo)
        k_render_serial<ZeroTwoSampler><<<nTiles, 32, 0, ctx->stream>>>(ctx->dS, ctx->C, ctx->F, P, ctx->dCnt, part, nParts);
    else
        k_render_serial<Pcg32><<<nTiles, 32, 0, ctx->stream>>>(ctx->dS, ctx->C, ctx->F, P, ctx->dCnt, part, nParts);
    cudaEventRecord(ctx->ev1, ctx->stream);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->counters.render_ms += ms;
    ctx->counters.kernel_launches += 1;
    return SPTX_OK;
}

// Device pixel list of a render region.  `key` says what the list describes (rectangle, tile
// partition): the list of the previous call is kept when the key is the same, so a caller that
// renders the same region sample range after sample range pays for it once.
static int set_pixels(sptx_ctx *ctx, const int key[12], const std::vector<int2> &pixels) {
    const int64_t n = (int64_t)pixels.size();
    if (n > ctx->dPixelsCap) {
        CK(cudaStreamSynchronize(ctx->stream));
        if (ctx->dPixels) cudaFree(ctx->dPixels);
        ctx->dPixels = nullptr;
        ctx->dPixelsCap = 0;
        ctx->pixelsKey[0] = -1;
        CK(cudaMalloc((void **)&ctx->dPixels, (size_t)n * sizeof(int2)));
        ctx->dPixelsCap = n;
    }
    if (n > 0) {
        CK(cudaMemcpyAsync(ctx->dPixels, pixels.data(), (size_t)n * sizeof(int2), cudaMemcpyHostToDevice,
                           ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    ctx->nPixels = n;
    memcpy(ctx->pixelsKey, key, sizeof(ctx->pixelsKey));
    return SPTX_OK;
}
static bool have_pixels(const sptx_ctx *ctx, const int key[12]) {
    return key[0] >= 0 && memcmp(ctx->pixelsKey, key, sizeof(ctx->pixelsKey)) == 0;
}

// Renders the pixels of ctx->dPixels for sample numbers [s0, s1).
static int render_pixels(sptx_ctx *ctx, int s0, int s1, const int *dSampleList = nullptr) {
    const int64_t nPixAll = ctx->nPixels;
    if (nPixAll <= 0 || s1 <= s0) return SPTX_OK;
    int rc;
    const int64_t want = ctx->maxPaths > 0 ? ctx->maxPaths : ((int64_t)1 << 22);
    const int64_t chunkPix = dSampleList ? nPixAll : std::min<int64_t>(nPixAll, want);
    int S_pass = (int)std::max<int64_t>(1, std::min<int64_t>(s1 - s0, want / chunkPix));
    if ((rc = ensure_paths(ctx, chunkPix * S_pass))) return rc;
    int2 *dPixels = ctx->dPixels;
    PathArrays &A = ctx->A;
    const int maxDepth = ctx->integ.max_depth;
    uint64_t launches = 0;
    // Per-launch CUDA events on the launching stream, read back after the sync.
    size_t evUsed = 0;
    std::vector<int> evKind;
    auto tic = [&](int kind) {
        if (evUsed + 2 > ctx->evPool.size()) {
            cudaEvent_t a, b;
            cudaEventCreate(&a);
            cudaEventCreate(&b);
            ctx->evPool.push_back(a);
            ctx->evPool.push_back(b);
        }
        evKind.push_back(kind);
        cudaEventRecord(ctx->evPool[evUsed], ctx->stream);
    };
    auto toc = [&]() {
        cudaEventRecord(ctx->evPool[evUsed + 1], ctx->stream);
        evUsed += 2;
        ++launches;
    };
    const bool cv = ctx->countVisits;
    // SpectralPathIntegrator's bands (the per-sample Li hook always traces the 550 nm path)
    const int nBands = dSampleList ? 1 : std::max(1, (int)ctx->integ.num_ca_bands);
    const int deltaIndex = (int)std::round((float)SPTX_NBINS / (float)nBands);
    const float deltaWave = (float)((705 - 395) / SPTX_NBINS);  // integer division, as in the reference
    const float deltaWaveCA = deltaWave * deltaIndex;
    if (nBands > 1 && ctx->dLaccPaths < A.capacity) {
        if (ctx->dLacc) cudaFree(ctx->dLacc);
        ctx->dLacc = nullptr;
        ctx->dLaccPaths = 0;
        CK(cudaMalloc((void **)&ctx->dLacc, (size_t)A.capacity * SPTX_SPECTRUM_STRIDE * sizeof(float)));
        ctx->dLaccPaths = A.capacity;
    }
    cudaEventRecord(ctx->ev0, ctx->stream);
    for (int64_t p0 = 0; p0 < nPixAll; p0 += chunkPix) {
        const int nPix = (int)std::min<int64_t>(chunkPix, nPixAll - p0);
        for (int s = s0; s < s1; s += S_pass) {
            RenderParams R;
            R.nPix = nPix;
            R.samplesInPass = std::min(S_pass, s1 - s);
            R.sampleBegin = s;
            R.maxDepth = maxDepth;
            R.rrThreshold = ctx->integ.rr_threshold;
            R.pixels = dPixels + p0;
            R.sampleList = dSampleList;
            R.ignoreRayWeight = ctx->metadataStrategy && !dSampleList;
            R.betaOnes = 0;
            const int64_t nPaths = (int64_t)nPix * R.samplesInPass;
            const int g = grid_for(ctx, nPaths);
            if (ctx->metadataStrategy && !dSampleList) {
                // MetadataIntegrator: camera ray, closest hit, the hit's attribute as "radiance"
                R.band = 0;
                R.nBands = 1;
                R.wavelength = 550.f;
                k_set_counters<<<1, 1, 0, ctx->stream>>>(A.counters, 0, 0, 1, 0);
                k_reset_bounce<<<1, 32, 0, ctx->stream>>>(A.counters, 1);
                launches += 2;
                tic(0);
                k_generate<<<g, kBlock, 0, ctx->stream>>>(ctx->S, ctx->C, ctx->H, A, R, ctx->dCnt);
                toc();
                tic(1);
                k_intersect<false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, 0, ctx->dCnt);
                toc();
                tic(2);
                k_metadata<<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, 0, ctx->metadataStrategy);
                toc();
            }
            for (int band = 0; band < (ctx->metadataStrategy && !dSampleList ? 0 : nBands); ++band) {
                // ray.wavelength = middle of the band (integrators/spectralpath.cpp:240-268);
                // one band = PathIntegrator at 550 nm
                R.band = band;
                R.nBands = nBands;
                R.wavelength = nBands > 1 ? 395.f + deltaWaveCA * band + (deltaWaveCA / 2) : 550.f;
                k_set_counters<<<1, 1, 0, ctx->stream>>>(A.counters, 0, 0, 1, 0);
                ++launches;
                tic(0);
                k_generate<<<g, kBlock, 0, ctx->stream>>>(ctx->S, ctx->C, ctx->H, A, R, ctx->dCnt);
                toc();
                int qi = 0, betaIn = 0;
                for (int bounce = 0; bounce <= maxDepth; ++bounce) {
                    k_reset_bounce<<<1, 32, 0, ctx->stream>>>(A.counters, qi ^ 1);
                    ++launches;
                    tic(1);
                    if (cv)
                        k_intersect<true><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, qi, ctx->dCnt);
                    else
                        k_intersect<false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, qi, ctx->dCnt);
                    toc();
                    tic(2);
                    R.betaOnes = bounce == 0;
                    const bool env = ctx->S.n_infinite > 0;
                    const int nk = (ctx->shadeExact ? (env ? shade_exact_env::launch_shade : shade_exact::launch_shade)
                                                    : (env ? shade_fast_env::launch_shade : shade_fast::launch_shade))(
                        ctx->stream, ctx->smCount, nPaths, ctx->S, ctx->H, A, R, qi, betaIn, ctx->shadePlan,
                        ctx->dCnt);
                    toc();
                    launches += nk - 1;
                    if (bounce < maxDepth) {
                        tic(3);
                        if (cv) {
                            k_mis<true><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, ctx->dCnt);
                            if (ctx->shadeExact)
                                k_resolve<true, false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, betaIn, bounce == 0, ctx->dCnt);
                            else
                                k_resolve<true, true><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, betaIn, bounce == 0, ctx->dCnt);
                        } else {
                            k_mis<false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, ctx->dCnt);
                            if (ctx->shadeExact)
                                k_resolve<false, false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, betaIn, bounce == 0, ctx->dCnt);
                            else
                                k_resolve<false, true><<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, betaIn, bounce == 0, ctx->dCnt);
                        }
                        ++launches;
                        toc();
                    }
                    qi ^= 1;
                    betaIn ^= 1;
                }
                if (nBands > 1) {
                    const int lo = deltaIndex * band;
                    const int hi = std::min(deltaIndex * (band + 1), (int)SPTX_NBINS);
                    tic(4);
                    k_band_commit<<<g, kBlock, 0, ctx->stream>>>(ctx->S, A, ctx->dLacc, nPaths, lo, hi,
                                                                 band == 0, band == nBands - 1);
                    toc();
                }
            }
            if (!dSampleList) {
                tic(4);
                k_film<<<grid_for(ctx, nPix), kBlock, 0, ctx->stream>>>(ctx->S, ctx->F, A, R);
                toc();
                tic(5);
                k_film_spill<<<g, kBlock, 0, ctx->stream>>>(ctx->S, ctx->F, A, R);
                toc();
            }
        }
    }
    cudaEventRecord(ctx->ev1, ctx->stream);
    cudaError_t ce = cudaStreamSynchronize(ctx->stream);
    if (ce != cudaSuccess) return fail(ctx, SPTX_ECUDA, "render: %s", cudaGetErrorString(ce));
    CK(cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->counters.render_ms += ms;
    for (size_t e = 0; e < evKind.size(); ++e) {
        float t = 0.f;
        cudaEventElapsedTime(&t, ctx->evPool[2 * e], ctx->evPool[2 * e + 1]);
        const int k = evKind[e];
        ctx->counters.kernel_ms[k] += t;
        ctx->counters.kernel_count[k] += 1;
        if (k == 1 || k == 3) ctx->counters.traverse_ms += t;
        else if (k == 0 || k == 2) ctx->counters.shade_ms += t;
        else ctx->counters.film_ms += t;
    }
    ctx->counters.kernel_launches += launches;
    int flags[8];
    CK(cudaMemcpyAsync(flags, A.counters, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (flags[3]) {
        int f = flags[3];
        cudaMemsetAsync(A.counters + 3, 0, sizeof(int), ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        if (f & ERR_DIM_OVERFLOW)
            return fail(ctx, SPTX_EINVAL,
                        "render: a path ran past the %d Halton dimensions uploaded "
                        "(the reference aborts at 1000, samplers/halton.h:71-76)", ctx->H.s.n_dims);
        return fail(ctx, SPTX_EINVAL, "render: device error flags 0x%x", f);
    }
    return SPTX_OK;
}

extern "C" int sptx_render(sptx_ctx *ctx, const int32_t *pixel_bounds, int sample_begin,
                           int sample_end) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_render: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_render: no scene uploaded");
    if (sample_begin < 0 || sample_end < sample_begin)
        return fail(ctx, SPTX_EINVAL, "sptx_render: bad sample range [%d,%d)", sample_begin, sample_end);
    CK(cudaSetDevice(ctx->device));
    if (tile_serial(ctx)) return render_serial(ctx, pixel_bounds, 0, 1, sample_begin, sample_end);
    int b[4];
    for (int i = 0; i < 4; ++i) b[i] = ctx->F.f.sample_bounds[i];
    auto clip = [&](const int32_t *o) {
        b[0] = std::max(b[0], o[0]);
        b[1] = std::max(b[1], o[1]);
        b[2] = std::min(b[2], o[2]);
        b[3] = std::min(b[3], o[3]);
    };
    clip(ctx->integ.pixel_bounds);  // the InsideExclusive(pixel, pixelBounds) test
    if (pixel_bounds) clip(pixel_bounds);
    const int key[12] = {1, b[0], b[1], b[2], b[3], 0, 0, 0, 0, 0, 0, 0};
    if (!have_pixels(ctx, key)) {
        std::vector<int2> pixels;
        pixels.reserve((size_t)std::max(0, b[2] - b[0]) * (size_t)std::max(0, b[3] - b[1]));
        for (int y = b[1]; y < b[3]; ++y)
            for (int x = b[0]; x < b[2]; ++x) pixels.push_back(make_int2(x, y));
        int rc = set_pixels(ctx, key, pixels);
        if (rc) return rc;
    }
    return render_pixels(ctx, sample_begin, sample_end);
}

extern "C" int sptx_render_tiles(sptx_ctx *ctx, int part, int n_parts, int sample_begin,
                                 int sample_end) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_render_tiles: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_render_tiles: no scene uploaded");
    if (n_parts < 1 || part < 0 || part >= n_parts)
        return fail(ctx, SPTX_EINVAL, "sptx_render_tiles: part %d of %d", part, n_parts);
    CK(cudaSetDevice(ctx->device));
    if (tile_serial(ctx)) return render_serial(ctx, nullptr, part, n_parts, sample_begin, sample_end);
    // Tiles as SamplerIntegrator::Render cuts them (core/integrator.cpp:232-257).
    const int *sb = ctx->F.f.sample_bounds;
    const int *pb = ctx->integ.pixel_bounds;
    const int tileSize = 16;
    const int ntx = (sb[2] - sb[0] + tileSize - 1) / tileSize;
    const int nty = (sb[3] - sb[1] + tileSize - 1) / tileSize;
    const int key[12] = {2, part, n_parts, sb[0], sb[1], sb[2], sb[3], pb[0], pb[1], pb[2], pb[3], 0};
    if (!have_pixels(ctx, key)) {
        std::vector<int2> pixels;
        for (int ty = 0; ty < nty; ++ty)
            for (int tx = 0; tx < ntx; ++tx) {
                if ((ty * ntx + tx) % n_parts != part) continue;
                int x0 = sb[0] + tx * tileSize, x1 = std::min(x0 + tileSize, sb[2]);
                int y0 = sb[1] + ty * tileSize, y1 = std::min(y0 + tileSize, sb[3]);
                for (int y = y0; y < y1; ++y)
                    for (int x = x0; x < x1; ++x)
                        if (x >= pb[0] && x < pb[2] && y >= pb[1] && y < pb[3])
                            pixels.push_back(make_int2(x, y));
            }
        int rc = set_pixels(ctx, key, pixels);
        if (rc) return rc;
    }
    return render_pixels(ctx, sample_begin, sample_end);
}

// ---- film -----------------------------------------------------------------------------

extern "C" int sptx_film_clear(sptx_ctx *ctx) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_film_clear: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_clear: no scene uploaded");
    CK(cudaSetDevice(ctx->device));
    size_t n = (size_t)ctx->F.width * ctx->F.height * SPTX_SPECTRUM_STRIDE;
    if (n) CK(cudaMemsetAsync(ctx->F.pixels, 0, n * sizeof(float), ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SPTX_OK;
}

extern "C" int sptx_film_device_ptr(sptx_ctx *ctx, void **d_film, int64_t *n_floats) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_film_device_ptr: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_device_ptr: no scene uploaded");
    if (d_film) *d_film = ctx->F.pixels;
    if (n_floats) *n_floats = (int64_t)ctx->F.width * ctx->F.height * SPTX_SPECTRUM_STRIDE;
    return SPTX_OK;
}

extern "C" int sptx_film_download(sptx_ctx *ctx, float *contrib, float *weight) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_film_download: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_download: no scene uploaded");
    CK(cudaSetDevice(ctx->device));
    const size_t nPix = (size_t)ctx->F.width * ctx->F.height;
    const size_t nFloats = nPix * SPTX_SPECTRUM_STRIDE;
    if (nPix == 0) return SPTX_OK;
    if (nFloats > ctx->dFilmSplitFloats) {
        if (ctx->dFilmSplit) cudaFree(ctx->dFilmSplit);
        ctx->dFilmSplit = nullptr;
        ctx->dFilmSplitFloats = 0;
        CK(cudaMalloc((void **)&ctx->dFilmSplit, nFloats * sizeof(float)));
        ctx->dFilmSplitFloats = nFloats;
    }
    k_film_split<<<grid_for(ctx, (int64_t)nFloats), kBlock, 0, ctx->stream>>>(ctx->F.pixels, (int64_t)nPix,
                                                                              ctx->dFilmSplit);
    if (contrib)
        CK(cudaMemcpyAsync(contrib, ctx->dFilmSplit, nPix * SPTX_NBINS * sizeof(float),
                           cudaMemcpyDeviceToHost, ctx->stream));
    if (weight)
        CK(cudaMemcpyAsync(weight, ctx->dFilmSplit + nPix * SPTX_NBINS, nPix * sizeof(float),
                           cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SPTX_OK;
}

static int map_peer_film(sptx_ctx *ctx, const unsigned char *handle, void **out);

extern "C" int sptx_film_export(sptx_ctx *ctx, void *handle) {
    if (!ctx || !handle) return fail(ctx, SPTX_EINVAL, "sptx_film_export: NULL argument");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_export: no scene uploaded");
    static_assert(sizeof(cudaIpcMemHandle_t) == SPTX_IPC_HANDLE_BYTES, "IPC handle size");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, ctx->F.pixels));
    memcpy(handle, &h, sizeof(h));
    return SPTX_OK;
}

extern "C" int sptx_film_merge_peers(sptx_ctx *ctx, const void *handles, int n_ranks, int self,
                                     double *planes) {
    if (!ctx || !handles) return fail(ctx, SPTX_EINVAL, "sptx_film_merge_peers: NULL argument");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_merge_peers: no scene uploaded");
    if (n_ranks < 1 || n_ranks > 8 || self < 0 || self >= n_ranks)
        return fail(ctx, SPTX_EINVAL, "sptx_film_merge_peers: rank %d of %d (at most 8 ranks)", self, n_ranks);
    CK(cudaSetDevice(ctx->device));
    const int64_t nPix = (int64_t)ctx->F.width * ctx->F.height;
    if (nPix == 0) return SPTX_OK;
    PeerFilms P;
    P.n = n_ranks;
    void *mapped[8] = {nullptr};
    int rc = SPTX_OK;
    for (int r = 0; r < n_ranks && rc == SPTX_OK; ++r) {
        if (r == self) {
            P.film[r] = ctx->F.pixels;
            continue;
        }
        rc = map_peer_film(ctx, (const unsigned char *)handles + (size_t)r * SPTX_IPC_HANDLE_BYTES, &mapped[r]);
        P.film[r] = (const float *)mapped[r];
    }
    double *dPlanes = nullptr;
    if (rc == SPTX_OK && planes) rc = ctx_scratch(ctx, 0, (size_t)nPix * SPTX_NBINS * sizeof(double), (void **)&dPlanes);
    if (rc == SPTX_OK) {
        cudaEventRecord(ctx->ev2, ctx->stream);
        k_film_merge_pack<<<ctx->smCount * 4, 256, 0, ctx->stream>>>(P, ctx->F.pixels, nPix, ctx->F.f.scale, dPlanes);
        cudaEventRecord(ctx->ev3, ctx->stream);
        cudaError_t e = cudaSuccess;
        if (planes)
            e = cudaMemcpyAsync(planes, dPlanes, (size_t)nPix * SPTX_NBINS * sizeof(double),
                                cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) rc = fail(ctx, SPTX_ECUDA, "sptx_film_merge_peers: %s", cudaGetErrorString(e));
        float ms = 0.f;
        if (rc == SPTX_OK && cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3) == cudaSuccess) ctx->counters.film_ms += ms;
        ctx->counters.kernel_launches += 1;
    }
    return rc;
}

// Maps the film behind `handle` (a peer rank's sptx_film_export) or returns the mapping made earlier.
static int map_peer_film(sptx_ctx *ctx, const unsigned char *handle, void **out) {
    for (auto &m : ctx->peerMaps)
        if (!memcmp(m.handle, handle, SPTX_IPC_HANDLE_BYTES)) {
            *out = m.ptr;
            return SPTX_OK;
        }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess)
        return fail(ctx, SPTX_ECUDA, "mapping a peer film: %s", cudaGetErrorString(e));
    sptx_ctx::PeerMap m;
    memcpy(m.handle, handle, SPTX_IPC_HANDLE_BYTES);
    m.ptr = p;
    ctx->peerMaps.push_back(m);
    *out = p;
    return SPTX_OK;
}

static int reduce_slice(sptx_ctx *ctx, const PeerFilms &P, float *dst, int n_ranks, int self) {
    const int64_t nQuads = (int64_t)ctx->F.width * ctx->F.height * SPTX_SPECTRUM_STRIDE / 4;
    const int64_t q0 = nQuads * self / n_ranks, q1 = nQuads * (self + 1) / n_ranks;
    if (q1 > q0) {
        cudaEventRecord(ctx->ev2, ctx->stream);
        k_film_reduce_slice<<<grid_for(ctx, q1 - q0, 256), 256, 0, ctx->stream>>>(P, dst, q0, q1);
        cudaEventRecord(ctx->ev3, ctx->stream);
    }
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    float ms = 0.f;
    if (q1 > q0 && cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3) == cudaSuccess) ctx->counters.film_ms += ms;
    ctx->counters.kernel_launches += 1;
    return SPTX_OK;
}

extern "C" int sptx_film_reduce_scatter(sptx_ctx *ctx, const void *handles, int n_ranks, int self, int dst) {
    if (!ctx || !handles) return fail(ctx, SPTX_EINVAL, "sptx_film_reduce_scatter: NULL argument");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_reduce_scatter: no scene uploaded");
    if (n_ranks < 1 || n_ranks > 8 || self < 0 || self >= n_ranks || dst < 0 || dst >= n_ranks)
        return fail(ctx, SPTX_EINVAL, "sptx_film_reduce_scatter: rank %d of %d to %d (at most 8 ranks)", self,
                    n_ranks, dst);
    CK(cudaSetDevice(ctx->device));
    PeerFilms P;
    P.n = n_ranks;
    for (int r = 0; r < n_ranks; ++r) {
        if (r == self) {
            P.film[r] = ctx->F.pixels;
            continue;
        }
        void *p = nullptr;
        int rc = map_peer_film(ctx, (const unsigned char *)handles + (size_t)r * SPTX_IPC_HANDLE_BYTES, &p);
        if (rc) return rc;
        P.film[r] = (const float *)p;
    }
    return reduce_slice(ctx, P, const_cast<float *>(P.film[dst]), n_ranks, self);
}

// One process, one context per GPU (what GpuPathIntegrator::Render does with SPTX_DEVICES): the
// films meet through peer access instead of IPC handles.
extern "C" int sptx_film_merge_contexts(sptx_ctx *const *ctxs, int n, int dst) {
    if (!ctxs || n < 1 || n > 8 || dst < 0 || dst >= n)
        return fail(nullptr, SPTX_EINVAL, "sptx_film_merge_contexts: %d contexts, destination %d (at most 8)", n, dst);
    PeerFilms P;
    P.n = n;
    for (int r = 0; r < n; ++r) {
        sptx_ctx *c = ctxs[r];
        if (!c || !c->haveScene) return fail(c, SPTX_ESTATE, "sptx_film_merge_contexts: context %d has no scene", r);
        if (c->F.width != ctxs[0]->F.width || c->F.height != ctxs[0]->F.height)
            return fail(c, SPTX_EINVAL, "sptx_film_merge_contexts: films differ in size");
        P.film[r] = c->F.pixels;
    }
    for (int r = 0; r < n; ++r) {   // every device may read every other device's film
        sptx_ctx *ctx = ctxs[r];
        CK(cudaSetDevice(ctx->device));
        CK(cudaStreamSynchronize(ctx->stream));
        for (int q = 0; q < n; ++q) {
            if (ctxs[q]->device == ctx->device) continue;
            int can = 0;
            CK(cudaDeviceCanAccessPeer(&can, ctx->device, ctxs[q]->device));
            if (!can)
                return fail(ctx, SPTX_EUNSUPPORTED, "sptx_film_merge_contexts: device %d cannot access device %d",
                            ctx->device, ctxs[q]->device);
            cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[q]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return fail(ctx, SPTX_ECUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            cudaGetLastError();
        }
    }
    for (int r = 0; r < n; ++r) {   // (slices are disjoint: the launches of different devices overlap)
        int rc = reduce_slice(ctxs[r], P, ctxs[dst]->F.pixels, n, r);
        if (rc) return rc;
    }
    return SPTX_OK;
}

// ---- NCCL film reduce (sptx_film_reduce(ctx, ncclComm_t, root)) ------------------------
// For hosts that already hold an NCCL communicator: one ncclReduce(sum, float32) of the film onto
// `root` on the context's stream.  libnccl is looked up at run time (the library has no link-time
// dependency on it); the peer-memory merges above need no NCCL at all.
#include <dlfcn.h>
extern "C" int sptx_film_reduce(sptx_ctx *ctx, void *nccl_comm, int root) {
    if (!ctx || !nccl_comm) return fail(ctx, SPTX_EINVAL, "sptx_film_reduce: NULL argument");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_reduce: no scene uploaded");
    typedef int (*reduce_fn)(const void *, void *, size_t, int, int, int, void *, cudaStream_t);
    typedef const char *(*err_fn)(int);
    static reduce_fn ncclReduceP = nullptr;
    static err_fn ncclErrP = nullptr;
    if (!ncclReduceP) {
        void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return fail(ctx, SPTX_EUNSUPPORTED, "sptx_film_reduce: libnccl.so.2 not found (%s)", dlerror());
        ncclReduceP = (reduce_fn)dlsym(lib, "ncclReduce");
        ncclErrP = (err_fn)dlsym(lib, "ncclGetErrorString");
        if (!ncclReduceP) return fail(ctx, SPTX_EUNSUPPORTED, "sptx_film_reduce: ncclReduce not in libnccl");
    }
    CK(cudaSetDevice(ctx->device));
    const size_t count = (size_t)ctx->F.width * ctx->F.height * SPTX_SPECTRUM_STRIDE;
    if (count == 0) return SPTX_OK;
    cudaEventRecord(ctx->ev2, ctx->stream);
    // ncclFloat32 = 7, ncclSum = 0 (nccl.h); in place on the root
    const int rc = ncclReduceP(ctx->F.pixels, ctx->F.pixels, count, 7, 0, root, nccl_comm, ctx->stream);
    cudaEventRecord(ctx->ev3, ctx->stream);
    if (rc != 0)
        return fail(ctx, SPTX_ECUDA, "sptx_film_reduce: ncclReduce: %s", ncclErrP ? ncclErrP(rc) : "error");
    CK(cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3) == cudaSuccess) ctx->counters.film_ms += ms;
    return SPTX_OK;
}

extern "C" int sptx_film_pack_dat(sptx_ctx *ctx, double *planes) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_film_pack_dat: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_film_pack_dat: no scene uploaded");
    if (!planes) return fail(ctx, SPTX_EINVAL, "sptx_film_pack_dat: planes is NULL");
    CK(cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)ctx->F.width * ctx->F.height * SPTX_NBINS;
    if (n == 0) return SPTX_OK;
    double *d = nullptr;
    int rc = ctx_scratch(ctx, 0, (size_t)n * sizeof(double), (void **)&d);
    if (rc) return rc;
    k_pack_dat<<<grid_for(ctx, n), kBlock, 0, ctx->stream>>>(ctx->F, d);
    cudaError_t e = cudaMemcpyAsync(planes, d, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return fail(ctx, SPTX_ECUDA, "sptx_film_pack_dat: %s", cudaGetErrorString(e));
    return SPTX_OK;
}

extern "C" int sptx_write_dat(const char *path, int width, int height, const double *planes) {
    // Film::WriteImage, core/film.cpp:267-306: text header, then binary planes.
    FILE *f = fopen(path, "wb");
    if (!f) return fail(nullptr, SPTX_EIO, "sptx_write_dat: cannot open %s", path);
    fprintf(f, "%d %d %d\n", width, height, SPTX_NBINS);
    fprintf(f, "v3 \n");
    size_t n = (size_t)width * height * SPTX_NBINS;
    size_t w = fwrite(planes, sizeof(double), n, f);
    fclose(f);
    if (w != n) return fail(nullptr, SPTX_EIO, "sptx_write_dat: short write to %s", path);
    return SPTX_OK;
}

// ---- counters ----------------------------------------------------------------------------

extern "C" int sptx_get_counters(sptx_ctx *ctx, sptx_counters *out) {
    if (!ctx || !out) return fail(ctx, SPTX_EINVAL, "sptx_get_counters: NULL argument");
    CK(cudaSetDevice(ctx->device));
    DeviceCounters dc;
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpyAsync(&dc, ctx->dCnt, sizeof(dc), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->counters.camera_rays = dc.cameraRays;
    ctx->counters.closest_hit_rays = dc.closestRays;
    ctx->counters.shadow_rays = dc.shadowRays;
    ctx->counters.path_vertices = dc.vertices;
    for (int k = 0; k < 2; ++k) {
        ctx->counters.nodes_visited[k] = dc.nodes[k];
        ctx->counters.prims_tested[k] = dc.tests[k];
    }
    *out = ctx->counters;
    return SPTX_OK;
}

extern "C" int sptx_reset_counters(sptx_ctx *ctx) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_reset_counters: ctx is NULL");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemsetAsync(ctx->dCnt, 0, sizeof(DeviceCounters), ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    memset(&ctx->counters, 0, sizeof(ctx->counters));
    return SPTX_OK;
}

// ---- ray batches ----------------------------------------------------------------------------

// Ray binning (north_star (2), "incoherent batches"): a thread-per-ray walk of a deep two-level BVH
// lives on L1 / L2 latency, and a warp of unrelated rays touches 32 unrelated node chains.  Large
// batches are therefore walked in the order of a 32-bit key -- the origin's cell in a 32^3 grid over
// the scene bounds (Morton order), then the direction's cell on a 16 x 16 octahedral map, then the
// origin's position inside its cell -- found with one cub radix sort of (key, ray index); each thread
// reads ray order[i] and writes hit order[i], so the caller's arrays keep their order and every ray's
// result (and visit counts) is what the unsorted walk gives.  Measured on C4, 2^22 rays, sort
// included: incoherent closest 5.81 -> 2.19 ms, any-hit 4.74 -> 1.86 ms; camera rays in scanline
// order (one origin: the key is the direction cell) 3.09 -> 2.69 ms.  Other key layouts tried:
// [256^3 origin][direction] 2.30 ms, [64^3 origin][direction][fine origin] 2.31 ms.
SPTX_DEV uint32_t interleave3(uint32_t x, uint32_t y, uint32_t z, int bits) {
    uint32_t m = 0;
    for (int bit = bits - 1; bit >= 0; --bit)
        m = (m << 3) | (((x >> bit) & 1u) << 2) | (((y >> bit) & 1u) << 1) | ((z >> bit) & 1u);
    return m;
}
__global__ void k_ray_keys(DScene S, const sptx_ray *rays, int64_t n, uint32_t *keys, int *order) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
        const float4 a = __ldg(rp), b = __ldg(rp + 1);
        const float o[3] = {a.x, a.y, a.z};
        uint32_t cell[3];
        for (int k = 0; k < 3; ++k) {
            const float ext = S.wb_max[k] - S.wb_min[k];
            float t = ext > 0.f ? (o[k] - S.wb_min[k]) / ext : 0.f;
            t = fminf(fmaxf(t, 0.f), 1.f);  // (NaN -> 0)
            cell[k] = min((uint32_t)(t * 256.f), 255u);
        }
        const float l1 = fabsf(b.x) + fabsf(b.y) + fabsf(b.z);
        float px = l1 > 0.f ? b.x / l1 : 0.f, py = l1 > 0.f ? b.y / l1 : 0.f;
        if (b.z < 0.f) {  // fold the lower hemisphere outwards
            const float qx = (1.f - fabsf(py)) * (px >= 0.f ? 1.f : -1.f), qy = (1.f - fabsf(px)) * (py >= 0.f ? 1.f : -1.f);
            px = qx;
            py = qy;
        }
        const uint32_t du = min((uint32_t)(fminf(fmaxf(px * .5f + .5f, 0.f), 1.f) * 16.f), 15u);
        const uint32_t dv = min((uint32_t)(fminf(fmaxf(py * .5f + .5f, 0.f), 1.f) * 16.f), 15u);
        uint32_t dm = 0;
        for (int bit = 3; bit >= 0; --bit) dm = (dm << 2) | (((du >> bit) & 1u) << 1) | ((dv >> bit) & 1u);
        keys[i] = (interleave3(cell[0] >> 3, cell[1] >> 3, cell[2] >> 3, 5) << 17) | (dm << 9) |
                  interleave3(cell[0] & 7u, cell[1] & 7u, cell[2] & 7u, 3);
        order[i] = (int)i;
    }
}
constexpr int64_t kRaySortMin = 1 << 16;

static int intersect_batch_impl(sptx_ctx *ctx, const void *d_rays, int64_t n, void *d_hits, int any_hit,
                                int with_counters, float *ms, bool maySort) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_intersect_batch: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_intersect_batch: no scene uploaded");
    if (n < 0 || (n > 0 && (!d_rays || !d_hits))) return fail(ctx, SPTX_EINVAL, "sptx_intersect_batch: bad arguments");
    CK(cudaSetDevice(ctx->device));
    if (n == 0) {
        if (ms) *ms = 0.f;
        return SPTX_OK;
    }
    const sptx_ray *r = (const sptx_ray *)d_rays;
    sptx_hit *h = (sptx_hit *)d_hits;
    const int g = grid_for(ctx, n);
    const int *order = nullptr;
    int launches = 1;
    ctx->lastBatchSorted = 0;
    const bool trySort = maySort && n <= 0x7fffffff && (ctx->raySort == 1 || (ctx->raySort == 2 && n >= kRaySortMin));
    uint32_t *keys = nullptr, *keysSorted = nullptr;
    int *ord = nullptr, *ordSorted = nullptr;
    void *tmp = nullptr;
    size_t tmpBytes = 0;
    if (trySort) {  // (scratch first: growing it drains the stream)
        cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keys, keysSorted, ord, ordSorted, (int)n, 0, 32, ctx->stream);
        const size_t arr = ((size_t)n * 4 + 255) & ~(size_t)255;
        void *base = nullptr;
        int rc = ctx_scratch(ctx, 3, 4 * arr + tmpBytes, &base);
        if (rc) return rc;
        unsigned char *b = (unsigned char *)base;
        keys = (uint32_t *)b;
        keysSorted = (uint32_t *)(b + arr);
        ord = (int *)(b + 2 * arr);
        ordSorted = (int *)(b + 3 * arr);
        tmp = b + 4 * arr;
    }
    cudaEventRecord(ctx->ev0, ctx->stream);
    if (trySort) {
        k_ray_keys<<<g, kBlock, 0, ctx->stream>>>(ctx->S, r, n, keys, ord);
        cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, keys, keysSorted, ord, ordSorted, (int)n, 0, 32, ctx->stream);
        launches += 1 + 5;  // (cub: histogram + one pass per 8 key bits)
        order = ordSorted;
        ctx->lastBatchSorted = 1;
    }
    if (any_hit) {
        if (with_counters)
            k_ray_batch<true, true><<<g, kBlock, 0, ctx->stream>>>(ctx->S, r, order, n, h);
        else
            k_ray_batch<true, false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, r, order, n, h);
    } else {
        if (with_counters)
            k_ray_batch<false, true><<<g, kBlock, 0, ctx->stream>>>(ctx->S, r, order, n, h);
        else
            k_ray_batch<false, false><<<g, kBlock, 0, ctx->stream>>>(ctx->S, r, order, n, h);
    }
    cudaEventRecord(ctx->ev1, ctx->stream);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    float t = 0.f;
    cudaEventElapsedTime(&t, ctx->ev0, ctx->ev1);
    if (ms) *ms = t;
    ctx->counters.kernel_launches += launches;
    ctx->counters.traverse_ms += t;
    ctx->counters.kernel_ms[6] += t;
    ctx->counters.kernel_count[6] += 1;
    return SPTX_OK;
}

extern "C" int sptx_intersect_batch_device(sptx_ctx *ctx, const void *d_rays, int64_t n, void *d_hits,
                                           int any_hit, int with_counters, float *ms) {
    return intersect_batch_impl(ctx, d_rays, n, d_hits, any_hit, with_counters, ms, true);
}

extern "C" int sptx_last_batch_sorted(sptx_ctx *ctx) { return ctx ? ctx->lastBatchSorted : 0; }

extern "C" int64_t sptx_light_table_rows(sptx_ctx *ctx) { return ctx ? ctx->voxelRows : 0; }

extern "C" int sptx_intersect_batch(sptx_ctx *ctx, const sptx_ray *rays, int64_t n, sptx_hit *hits,
                                    int any_hit, int with_counters) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_intersect_batch: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_intersect_batch: no scene uploaded");
    if (n < 0 || (n > 0 && (!rays || !hits))) return fail(ctx, SPTX_EINVAL, "sptx_intersect_batch: bad arguments");
    if (n == 0) return SPTX_OK;
    CK(cudaSetDevice(ctx->device));
    void *dR = nullptr, *dH = nullptr;
    int rc = ctx_scratch(ctx, 1, (size_t)n * sizeof(sptx_ray), &dR);
    if (rc == SPTX_OK) rc = ctx_scratch(ctx, 2, (size_t)n * sizeof(sptx_hit), &dH);
    if (rc) return rc;
    CK(cudaMemcpyAsync(dR, rays, (size_t)n * sizeof(sptx_ray), cudaMemcpyHostToDevice, ctx->stream));
    rc = sptx_intersect_batch_device(ctx, dR, n, dH, any_hit, with_counters, nullptr);
    if (rc) return rc;
    CK(cudaMemcpyAsync(hits, dH, (size_t)n * sizeof(sptx_hit), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SPTX_OK;
}

// ---- ray batches without any allocation ------------------------------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture around them, serving loops,
// harnesses that forbid it): the device copies of the rays and of the hit records live in a
// file-scope device array of fixed size instead of cudaMalloc'd scratch.  Every GPU that loads the
// library holds the array until the process exits.  All calls share it and it keeps the previous
// call's contents; nothing in it has to start at zero (rays [0,n) are overwritten by the copy-in,
// hits [0,n) by the kernel, and only those are copied out).  Calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char g_noallocScratch[(size_t)SPTX_NOALLOC_MAX_RAYS *
                                                               (sizeof(sptx_ray) + sizeof(sptx_hit))];
// The array's name is a host-side stand-in in host code, not a device address: the address is
// asked for once per device, with that device current.
static void *g_noallocPtr[64] = {nullptr};

static int noalloc_scratch(sptx_ctx *ctx, void **out) {
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64) return fail(ctx, SPTX_EINVAL, "sptx_intersect_batch_noalloc: device %d", dev);
    if (!g_noallocPtr[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noallocScratch);
        if (e != cudaSuccess || !p)
            return fail(ctx, SPTX_ECUDA, "sptx_intersect_batch_noalloc: cudaGetSymbolAddress: %s",
                        cudaGetErrorString(e));
        g_noallocPtr[dev] = p;
    }
    *out = g_noallocPtr[dev];
    return SPTX_OK;
}

extern "C" int sptx_intersect_batch_noalloc(sptx_ctx *ctx, const sptx_ray *rays, int64_t n, sptx_hit *hits,
                                            int any_hit) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_intersect_batch_noalloc: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_intersect_batch_noalloc: no scene uploaded");
    if (n < 0 || (n > 0 && (!rays || !hits)))
        return fail(ctx, SPTX_EINVAL, "sptx_intersect_batch_noalloc: bad arguments");
    if (n > (int64_t)SPTX_NOALLOC_MAX_RAYS)
        return fail(ctx, SPTX_EINVAL,
                    "sptx_intersect_batch_noalloc: %lld rays exceed the fixed scratch of %d rays "
                    "(SPTX_NOALLOC_MAX_RAYS); split the batch or use sptx_intersect_batch",
                    (long long)n, (int)SPTX_NOALLOC_MAX_RAYS);
    if (n == 0) return SPTX_OK;
    CK(cudaSetDevice(ctx->device));
    void *base = nullptr;
    int rc = noalloc_scratch(ctx, &base);
    if (rc) return rc;
    sptx_ray *dR = (sptx_ray *)base;
    sptx_hit *dH = (sptx_hit *)((unsigned char *)base + (size_t)SPTX_NOALLOC_MAX_RAYS * sizeof(sptx_ray));
    CK(cudaMemcpyAsync(dR, rays, (size_t)n * sizeof(sptx_ray), cudaMemcpyHostToDevice, ctx->stream));
    rc = intersect_batch_impl(ctx, dR, n, dH, any_hit, 0, nullptr, false);  // (no ray sort: its scratch is cudaMalloc'd)
    if (rc) return rc;
    CK(cudaMemcpyAsync(hits, dH, (size_t)n * sizeof(sptx_hit), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SPTX_OK;
}

// ---- measurement hook: L2 read bandwidth ---------------------------------------------------------
// The denominators of the traversal rooflines: MEASURED_PEAKS.json holds the HBM copy bandwidth, but
// the BVH working sets of the configurations (100 KB .. 11 MB) live in the 126 MB L2.  Reads a buffer
// of `bytes` (resident in L2 after the first sweep) `reps` times with 128-bit loads and returns GB/s.
__global__ void k_l2_read(const float4 *buf, int64_t n4, int reps, float *sink) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < reps; ++r)
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
            const float4 v = __ldcg(buf + i);  // cache at L2 only: every sweep goes to L2, not L1
            acc.x += v.x;
            acc.y += v.y;
            acc.z += v.z;
            acc.w += v.w;
        }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) *sink = acc.x;
}

extern "C" int sptx_measure_l2(sptx_ctx *ctx, int64_t bytes, int reps, double *gbs) {
    if (!ctx || !gbs) return fail(ctx, SPTX_EINVAL, "sptx_measure_l2: NULL argument");
    if (bytes < 4096 || bytes > ((int64_t)1 << 32) || reps < 1) return fail(ctx, SPTX_EINVAL, "sptx_measure_l2: bad size");
    CK(cudaSetDevice(ctx->device));
    void *buf = nullptr;
    int rc = ctx_scratch(ctx, 1, (size_t)bytes + 16, &buf);
    if (rc) return rc;
    CK(cudaMemsetAsync(buf, 0, (size_t)bytes, ctx->stream));
    const int64_t n4 = bytes / 16;
    float *sink = (float *)((char *)buf + bytes);
    k_l2_read<<<ctx->smCount * 8, 256, 0, ctx->stream>>>((const float4 *)buf, n4, 2, sink);  // warm-up
    cudaEventRecord(ctx->ev0, ctx->stream);
    k_l2_read<<<ctx->smCount * 8, 256, 0, ctx->stream>>>((const float4 *)buf, n4, reps, sink);
    cudaEventRecord(ctx->ev1, ctx->stream);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    *gbs = (double)n4 * 16.0 * reps / (ms * 1e-3) * 1e-9;
    return SPTX_OK;
}

// ---- test hooks --------------------------------------------------------------------------------

extern "C" int sptx_sampler_eval(sptx_ctx *ctx, const int32_t *px, const int32_t *py, const int32_t *sample,
                                 const int32_t *dim, int64_t n, float *values) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_sampler_eval: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_sampler_eval: no scene uploaded");
    if (n <= 0) return SPTX_OK;
    if (tile_serial(ctx))
        return fail(ctx, SPTX_EUNSUPPORTED, "sptx_sampler_eval: only for the Halton and Sobol samplers (a RandomSampler value depends on the whole tile before it)");
    CK(cudaSetDevice(ctx->device));
    std::vector<void *> pool;
    const int32_t *dpx, *dpy, *ds, *dd;
    float *dv;
    int rc;
    if ((rc = upload(ctx, pool, px, n, &dpx)) || (rc = upload(ctx, pool, py, n, &dpy)) ||
        (rc = upload(ctx, pool, sample, n, &ds)) || (rc = upload(ctx, pool, dim, n, &dd)) ||
        (rc = dalloc(ctx, pool, n, &dv))) {
        free_pool(ctx, pool);
        return rc;
    }
    k_sampler_eval<<<grid_for(ctx, n), kBlock, 0, ctx->stream>>>(ctx->H, dpx, dpy, ds, dd, n, dv);
    cudaError_t e = cudaMemcpyAsync(values, dv, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    free_pool(ctx, pool);
    if (e != cudaSuccess) return fail(ctx, SPTX_ECUDA, "sptx_sampler_eval: %s", cudaGetErrorString(e));
    return SPTX_OK;
}

extern "C" int sptx_camera_rays(sptx_ctx *ctx, const float *cs, int64_t n, float wavelength,
                                sptx_ray *rays, float *weights) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_camera_rays: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_camera_rays: no scene uploaded");
    if (n <= 0) return SPTX_OK;
    CK(cudaSetDevice(ctx->device));
    std::vector<void *> pool;
    const float *dcs;
    sptx_ray *dr;
    float *dw;
    int rc;
    if ((rc = upload(ctx, pool, cs, 5 * n, &dcs)) || (rc = dalloc(ctx, pool, n, &dr)) ||
        (rc = dalloc(ctx, pool, n, &dw))) {
        free_pool(ctx, pool);
        return rc;
    }
    k_camera_rays<<<grid_for(ctx, n), kBlock, 0, ctx->stream>>>(ctx->C, dcs, n, wavelength, dr, dw);
    cudaError_t e = cudaMemcpyAsync(rays, dr, (size_t)n * sizeof(sptx_ray), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(weights, dw, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    free_pool(ctx, pool);
    if (e != cudaSuccess) return fail(ctx, SPTX_ECUDA, "sptx_camera_rays: %s", cudaGetErrorString(e));
    return SPTX_OK;
}

// Debug / parity hook: PathIntegrator::Li for explicit (pixel, sample) pairs,
// returned per sample instead of being added to the film.
extern "C" int sptx_debug_radiance(sptx_ctx *ctx, const int32_t *px, const int32_t *py,
                                   const int32_t *sample, int64_t n, float *L, float *pfilm_weight) {
    if (!ctx) return fail(nullptr, SPTX_EINVAL, "sptx_debug_radiance: ctx is NULL");
    if (!ctx->haveScene) return fail(ctx, SPTX_ESTATE, "sptx_debug_radiance: no scene uploaded");
    if (n <= 0) return SPTX_OK;
    if (tile_serial(ctx))
        return fail(ctx, SPTX_EUNSUPPORTED, "sptx_debug_radiance: only for the Halton and Sobol samplers (a RandomSampler value depends on the whole tile before it)");
    CK(cudaSetDevice(ctx->device));
    std::vector<int2> pixels((size_t)n);
    for (int64_t i = 0; i < n; ++i) pixels[i] = make_int2(px[i], py[i]);
    std::vector<void *> pool;
    const int32_t *dS;
    int rc = upload(ctx, pool, sample, n, &dS);
    const int noKey[12] = {-1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (rc == SPTX_OK) rc = set_pixels(ctx, noKey, pixels);
    if (rc == SPTX_OK) rc = render_pixels(ctx, 0, 1, dS);
    free_pool(ctx, pool);
    if (rc) return rc;
    const PathArrays &A = ctx->A;
    std::vector<float> tmp((size_t)n * SPTX_SPECTRUM_STRIDE);
    CK(cudaMemcpy(tmp.data(), A.L, tmp.size() * sizeof(float), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < n; ++i)
        for (int b = 0; b < SPTX_NBINS; ++b) L[i * SPTX_NBINS + b] = tmp[i * SPTX_SPECTRUM_STRIDE + b];
    if (pfilm_weight) {
        std::vector<float2> pf((size_t)n);
        std::vector<float4> rb((size_t)n);
        CK(cudaMemcpy(pf.data(), A.pfilm, (size_t)n * sizeof(float2), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(rb.data(), A.rayB, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost));
        for (int64_t i = 0; i < n; ++i) {
            pfilm_weight[3 * i] = pf[i].x;
            pfilm_weight[3 * i + 1] = pf[i].y;
            pfilm_weight[3 * i + 2] = rb[i].w;
        }
    }
    return SPTX_OK;
}
