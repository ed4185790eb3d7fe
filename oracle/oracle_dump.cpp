// oracle/oracle_dump.cpp -- TEST INFRASTRUCTURE ONLY.  Never linked into, loaded
// by or executed from the product path; only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may run it.
//
// The parity oracle: the UNMODIFIED reference (compiled in place from
// /root/reference by host/Makefile `ref`) driven over seeded inputs.  It parses
// a scene file exactly like `pbrt` does, then -- instead of rendering -- asks
// the reference's own objects the questions the GPU path must answer:
//   ORACLE_RAYS_IN / ORACLE_HITS_OUT     Scene::Intersect  (BVHAccel::Intersect,
//                                        src/accelerators/bvh.cpp:662-700)
//   ORACLE_RAYS_IN / ORACLE_ANYHIT_OUT   Scene::IntersectP (bvh.cpp:702-738)
//   ORACLE_RAYS_IN / ORACLE_SPAWN_OUT    SurfaceInteraction::SpawnRay / SpawnRayTo from the hits
//                                        (src/tests/shapes.cpp:154-205), ORACLE_SPAWN_K per kind
//   ORACLE_CAMSAMPLES_IN / ORACLE_CAMRAYS_OUT   Camera::GenerateRayDifferential
//   ORACLE_HALTON_IN / ORACLE_HALTON_OUT HaltonSampler::SampleDimension
//                                        (src/samplers/halton.cpp:96-127)
//   ORACLE_LI_IN / ORACLE_LI_OUT         PathIntegrator::Li for listed
//                                        (pixel, sample) pairs, following
//                                        SamplerIntegrator::Render's per-sample
//                                        sequence (src/core/integrator.cpp:264-323)
//   ORACLE_TIME_RAYS=<reps>              times the two traversal loops on all
//                                        host threads (CPU baseline for Mrays/s)
// With no ORACLE_* variable set it renders with the reference PathIntegrator.
//
// How it gets the Scene: the reference constructs scene and integrator inside
// pbrtWorldEnd (src/core/api.cpp:1695-1707) and offers no hook, so this file
// defines pbrt::CreatePathIntegrator (link flag --allow-multiple-definition,
// our object first) returning a wrapper around the reference's PathIntegrator.
#include <algorithm>
#include <atomic>
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
#include "accelerators/bvh.h"
#include "camera.h"
#include "film.h"
#include "integrator.h"
#include "integrators/path.h"
#include "interaction.h"
#include "paramset.h"
#include "primitive.h"
#include "sampler.h"
#include "samplers/halton.h"
#include "scene.h"
#include "spectrum.h"
#undef private
#undef protected
#include "error.h"
#include "memory.h"
#include "parallel.h"

namespace pbrt {

namespace {

struct RayIn {  // == sptx_ray
    float o[3], tMax, d[3], time;
};
struct HitOut {  // == sptx_hit
    int32_t prim, instance;
    float t, b0, b1, b2;
    int32_t nNodes, nTests;
};

template <typename T>
bool ReadAll(const char *path, std::vector<T> *out) {
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out->resize(n / sizeof(T));
    size_t r = fread(out->data(), sizeof(T), out->size(), f);
    fclose(f);
    return r == out->size();
}
template <typename T>
bool WriteAll(const char *path, const std::vector<T> &v) {
    FILE *f = fopen(path, "wb");
    if (!f) return false;
    size_t w = fwrite(v.data(), sizeof(T), v.size(), f);
    fclose(f);
    return w == v.size();
}

// Global primitive numbering: the scene aggregate's primitives in BVH order,
// then each instanced aggregate in order of first use (the same enumeration
// the host flattener documents in include/sptx.h).
struct PrimIndex {
    std::map<const Primitive *, int> topLevel;           // prim -> global index
    std::map<const Primitive *, int> nestedBase;         // aggregate -> first index
    std::map<uint32_t, int> instanceOfId;                // instanceId -> instance index
    std::map<uint32_t, const Primitive *> aggregateOfId; // instanceId -> aggregate
    std::map<const Primitive *, std::map<const Primitive *, int>> nested;
    void Build(const Scene &scene) {
        auto *bvh = dynamic_cast<const BVHAccel *>(scene.aggregate.get());
        if (!bvh) return;
        int next = (int)bvh->primitives.size();
        int nInst = 0;
        std::vector<const Primitive *> order;
        for (size_t i = 0; i < bvh->primitives.size(); ++i) {
            const Primitive *p = bvh->primitives[i].get();
            topLevel[p] = (int)i;
            if (auto *tp = dynamic_cast<const TransformedPrimitive *>(p)) {
                const Primitive *agg = tp->primitive.get();
                instanceOfId[tp->instanceId] = nInst++;
                aggregateOfId[tp->instanceId] = agg;
                if (!nestedBase.count(agg)) {
                    nestedBase[agg] = -1;
                    order.push_back(agg);
                }
            }
        }
        for (const Primitive *agg : order) {
            nestedBase[agg] = next;
            auto &m = nested[agg];
            if (auto *nb = dynamic_cast<const BVHAccel *>(agg)) {
                for (size_t i = 0; i < nb->primitives.size(); ++i)
                    m[nb->primitives[i].get()] = next + (int)i;
                next += (int)nb->primitives.size();
            } else {
                m[agg] = next;
                next += 1;
            }
        }
    }
    void Lookup(const SurfaceInteraction &isect, int32_t *prim, int32_t *inst) const {
        auto it = topLevel.find(isect.primitive);
        if (it != topLevel.end()) {
            *prim = it->second;
            *inst = -1;
            return;
        }
        *prim = -2;
        *inst = -2;
        auto ia = aggregateOfId.find(isect.instanceId);
        if (ia == aggregateOfId.end()) return;
        *inst = instanceOfId.at(isect.instanceId);
        const auto &m = nested.at(ia->second);
        auto ip = m.find(isect.primitive);
        if (ip != m.end()) *prim = ip->second;
    }
};

// ---- visit counts ------------------------------------------------------------------------
// The reference's BVHAccel::Intersect is not instrumented, and the reference is not edited.  This
// is that loop (src/accelerators/bvh.cpp:662-700) walked again over the reference's OWN node
// array, with the reference's own Bounds3::IntersectP and Primitive::Intersect doing every test,
// counting the nodes and primitives a ray visits (the per-ray counts the
// traversal roofline is built on).  Nested aggregates of object instances are entered the way
// TransformedPrimitive::Intersect does (src/core/primitive.cpp:78-99).  The walk's hit must
// equal Scene::Intersect's, which the caller checks.
struct NodeView {  // the bytes of the file-local LinearBVHNode, src/accelerators/bvh.cpp:95-104
    Bounds3f bounds;
    int offset;  // primitivesOffset (leaf) / secondChildOffset (interior)
    uint16_t nPrimitives;
    uint8_t axis;
    uint8_t pad[1];
};
static_assert(sizeof(NodeView) == 32, "LinearBVHNode is 32 bytes");

bool CountingWalk(const Primitive *agg, const Ray &ray, SurfaceInteraction *isect, int32_t *nNodes,
                  int32_t *nTests) {
    auto *bvh = dynamic_cast<const BVHAccel *>(agg);
    if (!bvh) {  // a bare primitive as the aggregate
        ++*nTests;
        return agg->Intersect(ray, isect);
    }
    if (!bvh->nodes) return false;
    const NodeView *nodes = reinterpret_cast<const NodeView *>(bvh->nodes);
    bool hit = false;
    Vector3f invDir(1 / ray.d.x, 1 / ray.d.y, 1 / ray.d.z);
    int dirIsNeg[3] = {invDir.x < 0, invDir.y < 0, invDir.z < 0};
    int toVisitOffset = 0, currentNodeIndex = 0;
    int nodesToVisit[64];
    while (true) {
        const NodeView *node = &nodes[currentNodeIndex];
        ++*nNodes;
        if (node->bounds.IntersectP(ray, invDir, dirIsNeg)) {
            if (node->nPrimitives > 0) {
                for (int i = 0; i < node->nPrimitives; ++i) {
                    const Primitive *p = bvh->primitives[node->offset + i].get();
                    ++*nTests;
                    if (auto *tp = dynamic_cast<const TransformedPrimitive *>(p)) {
                        Transform p2w;
                        tp->PrimitiveToWorld.Interpolate(ray.time, &p2w);
                        Ray r2 = Inverse(p2w)(ray);
                        if (CountingWalk(tp->primitive.get(), r2, isect, nNodes, nTests)) {
                            ray.tMax = r2.tMax;
                            if (!p2w.IsIdentity()) *isect = p2w(*isect);
                            isect->instanceId = tp->instanceId;
                            hit = true;
                        }
                    } else if (p->Intersect(ray, isect))
                        hit = true;
                }
                if (toVisitOffset == 0) break;
                currentNodeIndex = nodesToVisit[--toVisitOffset];
            } else {
                if (dirIsNeg[node->axis]) {
                    nodesToVisit[toVisitOffset++] = currentNodeIndex + 1;
                    currentNodeIndex = node->offset;
                } else {
                    nodesToVisit[toVisitOffset++] = node->offset;
                    currentNodeIndex = currentNodeIndex + 1;
                }
            }
        } else {
            if (toVisitOffset == 0) break;
            currentNodeIndex = nodesToVisit[--toVisitOffset];
        }
    }
    return hit;
}

class DumpIntegrator : public Integrator {
  public:
    DumpIntegrator(PathIntegrator *real) : real(real) {}
    void Render(const Scene &scene) override;

  private:
    std::unique_ptr<PathIntegrator> real;
};

void DumpIntegrator::Render(const Scene &scene) {
    const char *raysIn = getenv("ORACLE_RAYS_IN");
    const char *camIn = getenv("ORACLE_CAMSAMPLES_IN");
    const char *halIn = getenv("ORACLE_HALTON_IN");
    const char *liIn = getenv("ORACLE_LI_IN");
    if (!raysIn && !camIn && !halIn && !liIn) {
        real->Render(scene);
        return;
    }
    if (raysIn) {
        std::vector<RayIn> rays;
        if (!ReadAll(raysIn, &rays)) Error("oracle_dump: cannot read %s", raysIn);
        PrimIndex index;
        index.Build(scene);
        const int64_t n = (int64_t)rays.size();
        if (const char *out = getenv("ORACLE_HITS_OUT")) {
            std::vector<HitOut> hits(n);
            ParallelFor([&](int64_t i) {
                Ray r(Point3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]),
                      Vector3f(rays[i].d[0], rays[i].d[1], rays[i].d[2]), rays[i].tMax,
                      rays[i].time);
                SurfaceInteraction isect;
                HitOut h;
                memset(&h, 0, sizeof(h));
                h.prim = h.instance = -1;
                if (scene.Intersect(r, &isect)) index.Lookup(isect, &h.prim, &h.instance);
                h.t = r.tMax;
                if (getenv("ORACLE_COUNT_VISITS")) {
                    Ray rc(Point3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]),
                           Vector3f(rays[i].d[0], rays[i].d[1], rays[i].d[2]), rays[i].tMax,
                           rays[i].time);
                    SurfaceInteraction ic;
                    int32_t p2 = -1, i2 = -1;
                    if (CountingWalk(scene.aggregate.get(), rc, &ic, &h.nNodes, &h.nTests))
                        index.Lookup(ic, &p2, &i2);
                    if (p2 != h.prim || i2 != h.instance || rc.tMax != r.tMax) {
                        Error("oracle_dump: the counting walk disagrees with Scene::Intersect on ray %lld",
                              (long long)i);
                        fflush(nullptr);
                        std::_Exit(3);
                    }
                }
                hits[i] = h;
            }, n, 4096);
            WriteAll(out, hits);
        }
        if (const char *out = getenv("ORACLE_SPAWN_OUT")) {
            // src/tests/shapes.cpp:154-205 (Triangle.Reintersect) on a whole scene: for every input ray
            // that hits, ORACLE_SPAWN_K rays leave the hit through the reference's own
            // SurfaceInteraction::SpawnRay(w) (w uniform on the sphere, RNG(i) as in the test) and as
            // many through SpawnRayTo(a seeded point of the scene bound).  Output records: the
            // spawned ray (RayIn layout) followed by the parent's primitive / instance index.
            const int K = std::max(1, atoi(getenv("ORACLE_SPAWN_K") ? getenv("ORACLE_SPAWN_K") : "4"));
            struct SpawnOut {
                RayIn ray;
                int32_t parentPrim, parentInst;
            };
            std::vector<SpawnOut> spawned((size_t)n * 2 * K);
            std::vector<uint8_t> valid((size_t)n, 0);
            const Bounds3f wb = scene.WorldBound();
            ParallelFor([&](int64_t i) {
                Ray r(Point3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]),
                      Vector3f(rays[i].d[0], rays[i].d[1], rays[i].d[2]), rays[i].tMax, rays[i].time);
                SurfaceInteraction isect;
                if (!scene.Intersect(r, &isect)) return;
                int32_t pp = -1, pi = -1;
                index.Lookup(isect, &pp, &pi);
                valid[i] = 1;
                RNG rng((uint64_t)i);
                for (int j = 0; j < 2 * K; ++j) {
                    Point2f u;
                    u[0] = rng.UniformFloat();
                    u[1] = rng.UniformFloat();
                    Ray out;
                    if (j < K)
                        out = isect.SpawnRay(UniformSampleSphere(u));
                    else {
                        Float w = rng.UniformFloat();
                        Point3f p2 = wb.Lerp(Point3f(u[0], u[1], w));
                        out = isect.SpawnRayTo(p2);
                    }
                    SpawnOut &o = spawned[(size_t)i * 2 * K + j];
                    for (int c = 0; c < 3; ++c) {
                        o.ray.o[c] = out.o[c];
                        o.ray.d[c] = out.d[c];
                    }
                    o.ray.tMax = out.tMax;
                    o.ray.time = 0;
                    o.parentPrim = pp;
                    o.parentInst = pi;
                }
            }, n, 1024);
            std::vector<SpawnOut> kept;
            for (int64_t i = 0; i < n; ++i)
                if (valid[i])
                    kept.insert(kept.end(), spawned.begin() + (size_t)i * 2 * K,
                                spawned.begin() + (size_t)(i + 1) * 2 * K);
            WriteAll(out, kept);
        }
        if (const char *out = getenv("ORACLE_ANYHIT_OUT")) {
            std::vector<int32_t> any(n);
            ParallelFor([&](int64_t i) {
                Ray r(Point3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]),
                      Vector3f(rays[i].d[0], rays[i].d[1], rays[i].d[2]), rays[i].tMax,
                      rays[i].time);
                any[i] = scene.IntersectP(r) ? 1 : 0;
            }, n, 4096);
            WriteAll(out, any);
        }
        if (const char *reps = getenv("ORACLE_TIME_RAYS")) {
            int nreps = std::max(1, atoi(reps));
            for (int pass = 0; pass < 2; ++pass) {
                std::atomic<int64_t> nHit(0);
                auto t0 = std::chrono::steady_clock::now();
                for (int rep = 0; rep < nreps; ++rep)
                    ParallelFor([&](int64_t i) {
                        Ray r(Point3f(rays[i].o[0], rays[i].o[1], rays[i].o[2]),
                              Vector3f(rays[i].d[0], rays[i].d[1], rays[i].d[2]),
                              rays[i].tMax, rays[i].time);
                        bool hit;
                        if (pass == 0) {
                            SurfaceInteraction isect;
                            hit = scene.Intersect(r, &isect);
                        } else
                            hit = scene.IntersectP(r);
                        if (hit) nHit.fetch_add(1, std::memory_order_relaxed);
                    }, n, 4096);
                double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                printf("ORACLE_TIMING %s rays=%lld reps=%d seconds=%.6f mrays_per_s=%.4f threads=%d hits=%lld\n",
                       pass == 0 ? "closest" : "any", (long long)n, nreps, s,
                       (double)n * nreps / s * 1e-6, NumSystemCores(), (long long)nHit.load());
            }
        }
    }
    if (camIn) {
        std::vector<float> cs;
        if (!ReadAll(camIn, &cs)) Error("oracle_dump: cannot read %s", camIn);
        const size_t n = cs.size() / 5;
        std::vector<float> out(n * 8);
        const Camera &camera = *real->camera;
        for (size_t i = 0; i < n; ++i) {
            CameraSample s;
            s.pFilm = Point2f(cs[5 * i], cs[5 * i + 1]);
            s.time = cs[5 * i + 2];
            s.pLens = Point2f(cs[5 * i + 3], cs[5 * i + 4]);
            RayDifferential ray;
            // SpectralPathIntegrator sets the wavelength before generating the ray
            // (integrators/spectralpath.cpp:265-270); 550 is the Ray default
            if (const char *wl = getenv("ORACLE_WAVELENGTH")) ray.wavelength = (Float)atof(wl);
            Float w = camera.GenerateRayDifferential(s, &ray);
            float *o = &out[8 * i];
            o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.tMax;
            o[4] = ray.d.x; o[5] = ray.d.y; o[6] = ray.d.z; o[7] = w;
        }
        if (const char *p = getenv("ORACLE_CAMRAYS_OUT")) WriteAll(p, out);
    }
    if (halIn) {
        std::vector<int32_t> q;
        if (!ReadAll(halIn, &q)) Error("oracle_dump: cannot read %s", halIn);
        const size_t n = q.size() / 4;
        std::vector<float> out(n);
        // HaltonSampler and SobolSampler: GlobalSampler::GetIndexForSample / SampleDimension
        auto *h = dynamic_cast<GlobalSampler *>(real->sampler.get());
        if (!h) Error("oracle_dump: ORACLE_HALTON_IN needs a GlobalSampler (\"halton\", \"sobol\")");
        for (size_t i = 0; h && i < n; ++i) {
            h->StartPixel(Point2i(q[4 * i], q[4 * i + 1]));
            int64_t idx = h->GetIndexForSample(q[4 * i + 2]);
            out[i] = h->SampleDimension(idx, q[4 * i + 3]);
        }
        if (const char *p = getenv("ORACLE_HALTON_OUT")) WriteAll(p, out);
    }
    if (liIn) {
        std::vector<int32_t> q;
        if (!ReadAll(liIn, &q)) Error("oracle_dump: cannot read %s", liIn);
        const size_t n = q.size() / 3;
        std::vector<float> out(n * 34);
        real->Preprocess(scene, *real->sampler);
        std::unique_ptr<Sampler> smp = real->sampler->Clone(0);
        MemoryArena arena;
        for (size_t i = 0; i < n; ++i) {
            Point2i pixel(q[3 * i], q[3 * i + 1]);
            smp->StartPixel(pixel);
            smp->SetSampleNumber(q[3 * i + 2]);
            CameraSample cameraSample = smp->GetCameraSample(pixel);
            RayDifferential ray;
            Float rayWeight = real->camera->GenerateRayDifferential(cameraSample, &ray);
            ray.ScaleDifferentials(1 / std::sqrt((Float)smp->samplesPerPixel));
            Spectrum L(0.f);
            if (getenv("ORACLE_LI_TRACE") && (int)i < atoi(getenv("ORACLE_LI_TRACE"))) {
                // the camera vertex as the material sees it: parameterisation, differentials
                RayDifferential r2(ray);
                SurfaceInteraction si;
                if (scene.Intersect(r2, &si)) {
                    si.ComputeDifferentials(r2);
                    fprintf(stderr, "REFTRACE q %d %d %d p %.9g %.9g %.9g n %.9g %.9g %.9g uv %.9g %.9g\n", pixel.x,
                            pixel.y, q[3 * i + 2], si.p.x, si.p.y, si.p.z, si.n.x, si.n.y, si.n.z, si.uv[0], si.uv[1]);
                    fprintf(stderr, "REFTRACE   rxo %.9g %.9g %.9g rxd %.9g %.9g %.9g ryo %.9g %.9g %.9g ryd %.9g %.9g %.9g\n",
                            r2.rxOrigin.x, r2.rxOrigin.y, r2.rxOrigin.z, r2.rxDirection.x, r2.rxDirection.y,
                            r2.rxDirection.z, r2.ryOrigin.x, r2.ryOrigin.y, r2.ryOrigin.z, r2.ryDirection.x,
                            r2.ryDirection.y, r2.ryDirection.z);
                    fprintf(stderr, "REFTRACE   dpdu %.9g %.9g %.9g dpdv %.9g %.9g %.9g dudx %.9g dvdx %.9g dudy %.9g dvdy %.9g\n",
                            si.dpdu.x, si.dpdu.y, si.dpdu.z, si.dpdv.x, si.dpdv.y, si.dpdv.z, si.dudx, si.dvdx,
                            si.dudy, si.dvdy);
                    si.ComputeScatteringFunctions(r2, arena, true);
                    if (si.bsdf && si.bsdf->nBxDFs > 0) {
                        Spectrum f = si.bsdf->f(Vector3f(si.n), Vector3f(si.n)) * Pi;
                        fprintf(stderr, "REFTRACE   f*pi(n,n) %.9g %.9g %.9g %.9g nBxDFs %d\n", f[2], f[10], f[18], f[26],
                                si.bsdf->nBxDFs);
                    }
                }
            }
            if (rayWeight > 0) L = real->Li(ray, scene, *smp, arena, 0);
            float *o = &out[34 * i];
            for (int b = 0; b < nSpectralSamples; ++b) o[b] = L[b];
            o[31] = cameraSample.pFilm.x;
            o[32] = cameraSample.pFilm.y;
            o[33] = rayWeight;
            arena.Reset();
        }
        if (const char *p = getenv("ORACLE_LI_OUT")) WriteAll(p, out);
    }
}

}  // namespace

// Same parameters as the reference factory (src/integrators/path.cpp:190-213),
// handed to the reference's own PathIntegrator.
PathIntegrator *CreatePathIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler,
                                     std::shared_ptr<const Camera> camera) {
    int maxDepth = params.FindOneInt("maxdepth", 5);
    int np;
    const int *pb = params.FindInt("pixelbounds", &np);
    Bounds2i pixelBounds = camera->film->GetSampleBounds();
    if (pb && np == 4)
        pixelBounds = Intersect(pixelBounds, Bounds2i{{pb[0], pb[2]}, {pb[1], pb[3]}});
    Float rrThreshold = params.FindOneFloat("rrthreshold", 1.);
    std::string lightStrategy = params.FindOneString("lightsamplestrategy", "spatial");
    PathIntegrator *real =
        new PathIntegrator(maxDepth, camera, sampler, pixelBounds, rrThreshold, lightStrategy);
    Integrator *wrap = new DumpIntegrator(real);
    return reinterpret_cast<PathIntegrator *>(wrap);
}

}  // namespace pbrt

// main() is the reference's own CLI.
#include "main/pbrt.cpp"
