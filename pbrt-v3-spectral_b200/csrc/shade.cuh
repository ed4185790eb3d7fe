// shade.cuh -- K3: one path vertex of PathIntegrator::Li.
//
// integrators/path.cpp:88-184 for the vertex, core/integrator.cpp:85-215
// (UniformSampleOneLight, EstimateDirect) for next-event estimation.
//
// Two kinds of phases alternate, so that the 31-bin arithmetic never diverges:
//   scalar phase  one THREAD per path: interaction, BSDF frame, light choice,
//                 light / BSDF sampling, pdfs, MIS weights, spawned rays.  The
//                 spectral work it implies is written to shared memory as a
//                 "recipe" (formula ids + spectrum rows + scalars).
//   bin phase     the WARP walks its 32 recipes; lane b evaluates bin b.  Path
//                 spectra are path-major ([P][32] floats), so every access is
//                 one 128-byte line no matter how scattered the queue is, and
//                 all branches are warp-uniform.
// The two rays EstimateDirect traces (shadow ray, BSDF-sampled MIS ray) are
// written out with their per-bin contributions (fa, fb) and traced by K4.
#pragma once
#include "kernels.cuh"

namespace sptx {

// LightDistribution::Lookup + Distribution1D::SampleDiscrete.
SPTX_DEV int select_light(const DScene &S, V3 p, float u, float *pdf) {
    const int n = S.n_lights;
    if (S.light_strategy != SPTX_LIGHTS_SPATIAL)
        return sample_discrete(S.light_func, S.light_func + n, n, S.light_func_int, u, pdf);
    // SpatialLightDistribution::Lookup, core/lightdistrib.cpp:135-147
    V3 pmin = mk(S.wb_min[0], S.wb_min[1], S.wb_min[2]);
    V3 pmax = mk(S.wb_max[0], S.wb_max[1], S.wb_max[2]);
    V3 o = p - pmin;  // Bounds3::Offset
    if (pmax.x > pmin.x) o.x /= pmax.x - pmin.x;
    if (pmax.y > pmin.y) o.y /= pmax.y - pmin.y;
    if (pmax.z > pmin.z) o.z /= pmax.z - pmin.z;
    int pi0 = min(max((int)(o.x * (float)S.n_voxels[0]), 0), S.n_voxels[0] - 1);
    int pi1 = min(max((int)(o.y * (float)S.n_voxels[1]), 0), S.n_voxels[1] - 1);
    int pi2 = min(max((int)(o.z * (float)S.n_voxels[2]), 0), S.n_voxels[2] - 1);
    size_t voxel = ((size_t)pi0 * S.n_voxels[1] + pi1) * S.n_voxels[2] + pi2;
    const float *row = S.voxel_tables + voxel * (size_t)(2 * n + 2);
    return sample_discrete(row, row + n, n, __ldg(row + 2 * n + 1), u, pdf);
}

SPTX_DEV float power_heuristic(float fPdf, float gPdf) {
    float f = 1.f * fPdf, g = 1.f * gPdf;
    return (f * f) / (f * f + g * g);
}


// What one pass of the bin phase has to do for one path.
enum {
    W_EMIT = 1,      // L += beta * Le
    W_FA = 2,        // fa = f * |wi.ns| * Li [* weight] / lightPdf   (light-sampling strategy)
    W_FB = 4,        // fb = f * |wi.ns| * Le * weight / scatteringPdf (BSDF-sampling strategy)
    W_FC = 8,        // betaNext = beta * (f * |wi.ns| / pdf); max(betaNext * etaScale)
    W_FA_DELTA = 16  // delta light: no MIS weight
};

struct Recipe {
    int path;      // path id, -1 = idle lane
    int work;      // W_* flags of the pending pass
    int emitRow;   // Le spectrum row (W_EMIT)
    int row;       // W_FA: light-sample radiance row; W_FB: Lemit row of the MIS light
    float rowDiv;  // W_FA: 0, or divide the row by it (point light: I / d^2)
    float ad, weight, pdf;
    float etaScale;  // W_FC
    float rrDiv;     // epilogue: betaNext /= rrDiv
    // results of the pass
    int black;  // f (or Li) is black in every bin
    float maxBeta;
    FSum f;
};

// One pass of the bin phase: the warp walks its 32 recipes, lane = bin.
SPTX_NOINL void bin_pass(const DScene &S, const PathArrays &A, Recipe *wr, int lane,
                         const float *betaCur, float *betaNext) {
    const bool live = lane < SPTX_NBINS;
    unsigned todo = __ballot_sync(0xffffffffu, wr[lane].path >= 0 && wr[lane].work != 0);
    if (todo == 0) return;
    // the next path's throughput row is requested while this one is processed
    float betaAhead = betaCur[(size_t)wr[__ffs(todo) - 1].path * SPTX_SPECTRUM_STRIDE + lane];
    while (todo) {
        const int p = __ffs(todo) - 1;
        todo &= todo - 1;
        const Recipe &r = wr[p];
        const int work = r.work;
        const size_t at = (size_t)r.path * SPTX_SPECTRUM_STRIDE + lane;
        const float beta = betaAhead;
        if (todo) betaAhead = betaCur[(size_t)wr[__ffs(todo) - 1].path * SPTX_SPECTRUM_STRIDE + lane];
        if ((work & W_EMIT) && live) A.L[at] += beta * spec(S, r.emitRow, lane);
        int black = 0;
        float mb = 0.f;
        if (work & (W_FA | W_FB | W_FC)) {
            const float fs = live ? fsum_bin(S, r.f, lane) : 0.f;
            if (work & W_FA) {
                const float fv = fs * r.ad;
                float li = live ? spec(S, r.row, lane) : 0.f;
                if (r.rowDiv != 0.f) li = li / r.rowDiv;
                float c = fv * li;
                if (!(work & W_FA_DELTA)) c = c * r.weight;
                A.fa[at] = c / r.pdf;
                if (__ballot_sync(0xffffffffu, fv != 0.f) == 0) black = 1;
                if (__ballot_sync(0xffffffffu, li != 0.f) == 0) black = 1;
            } else if (work & W_FB) {
                const float fv = fs * r.ad;
                const float le = live ? spec(S, r.row, lane) : 0.f;
                A.fb[at] = fv * le * r.weight / r.pdf;
                if (__ballot_sync(0xffffffffu, fv != 0.f) == 0) black = 1;
            } else {
                const float nb = beta * (fs * r.ad / r.pdf);
                betaNext[at] = nb;
                if (__ballot_sync(0xffffffffu, fs != 0.f) == 0) black = 1;
                // MaxComponentValue(beta * etaScale): max does not depend on the order
                float v = live ? nb * r.etaScale : -kInf;
                for (int o = 16; o > 0; o >>= 1) {
                    float w = __shfl_xor_sync(0xffffffffu, v, o);
                    v = (v < w) ? w : v;
                }
                mb = v;
            }
        }
        if (lane == 0) {
            wr[p].black = black;
            wr[p].maxBeta = mb;
        }
    }
}

// The vertex is cut into stages with a block barrier between them.  The
// barriers are not needed for correctness (a recipe is only shared inside its
// warp): they keep the block's warps inside the same few KB of code, so that
// the SM's instruction cache serves all of them from one fetch (the kernel is
// ~200 KB of SASS; see DESIGN.md, "k_shade and the instruction cache").
constexpr int kShadeBlock = 1024;
constexpr int kShadeBlocksPerSM = 1;
#ifdef SPTX_STAGE_CLOCKS  // diagnosis build: per-stage warp clocks, printed by k_film
__device__ unsigned long long g_stageClocks[32];
#define SPTX_TICK(slot)                      \
    do {                                     \
        long long now_ = clock64();          \
        stageClk[slot] += now_ - tPrev;      \
        tPrev = now_;                        \
    } while (0)
#else
#define SPTX_TICK(slot)
#endif
#define SPTX_STAGE(slot)  \
    do {                  \
        SPTX_TICK(slot);  \
        __syncthreads();  \
        SPTX_TICK(slot + 1); \
    } while (0)

__global__ void __launch_bounds__(kShadeBlock, kShadeBlocksPerSM)
k_shade(DScene S, DSampler H, PathArrays A, RenderParams R, int qi, int betaIn,
        DeviceCounters *cnt) {
    extern __shared__ unsigned char smemRaw[];
    Recipe *recipes = reinterpret_cast<Recipe *>(smemRaw);
    // work list: the class queues one after the other, in block-sized chunks
    int classN[kShadeClasses], nChunks = 0, n = 0;
    for (int c = 0; c < kShadeClasses; ++c) {
        classN[c] = A.counters[kClassCounter0 + c];
        nChunks += (classN[c] + (int)blockDim.x - 1) / (int)blockDim.x;
        n += classN[c];
    }
    const float *betaCur = A.beta[betaIn];
    float *betaNext = A.beta[betaIn ^ 1];
    const int lane = threadIdx.x & 31;
    Recipe *wr = recipes + (threadIdx.x & ~31);
    Recipe &rc = recipes[threadIdx.x];
    const int noSpec = BSDF_ALL & ~BSDF_SPECULAR;
#ifdef SPTX_STAGE_CLOCKS
    long long stageClk[24] = {0};
    long long tPrev = clock64();
#endif
    for (int chunk = blockIdx.x; chunk < nChunks; chunk += gridDim.x) {
        int cls = 0, first = chunk;
        while (first >= (classN[cls] + (int)blockDim.x - 1) / (int)blockDim.x) {
            first -= (classN[cls] + (int)blockDim.x - 1) / (int)blockDim.x;
            ++cls;
        }
        const int *queue = A.qc + (size_t)cls * A.capacity;
        const int k = first * (int)blockDim.x + (int)threadIdx.x;
        bool alive = k < classN[cls];
        int i = 0;
        rc.path = -1;
        rc.work = 0;
        SampleStream st{0, 0};
        int bounces = 0, neeFlags = 0, lightNum = 0;
        float lightPdfSel = 0.f, etaScale = 1.f, rayWeight = 0.f;
        Isect it;
        Bsdf bs;
        Ray ray{mk(0, 0, 0), mk(0, 0, 1), kInf};
        // ---- stage 0: interaction, emission (path.cpp:91-112) ----
        if (alive) {
            i = queue[k];
            rc.path = i;
            // the bin passes read this path's throughput row much later: have it in L2 by then
            asm volatile("prefetch.global.L2 [%0];" ::"l"(betaCur + (size_t)i * SPTX_SPECTRUM_STRIDE));
            const float4 ra = A.rayA[i], rb = A.rayB[i];
            const float4 ha = A.hitA[i], hb = A.hitB[i];
            ray.o = mk(ra.x, ra.y, ra.z);
            ray.d = mk(ra.w, rb.x, rb.y);
            etaScale = rb.z;
            rayWeight = rb.w;
            const uint32_t info = A.info[i];
            bounces = (int)(info & 0xffu);
            const bool specularBounce = ((info >> 8) & 1u) != 0;
            st.index = A.sindex[i];
            st.dim = (int)(info >> 16);
            const int prim = __float_as_int(ha.x), inst = __float_as_int(ha.y);
            bool haveIsect = false;
            if (prim >= 0) {
                haveIsect = build_interaction(S, ray, prim, inst, ha.z, ha.w, hb.x, &it);
                if (!haveIsect) atomicOr(A.counters + 3, ERR_INTERACTION);
            }
            // isect.Le(-ray.d) -> DiffuseAreaLight::L(intr, w)
            if ((bounces == 0 || specularBounce) && haveIsect && it.light >= 0) {
                const sptx_light &Lt = S.lights[it.light];
                if (Lt.two_sided || dot(it.n, -ray.d) > 0) {
                    rc.work = W_EMIT;
                    rc.emitRow = Lt.spectrum;
                }
            }
            alive = haveIsect && bounces < R.maxDepth;
        }
        SPTX_STAGE(0);
        // ---- stage 1: BSDF frame, light choice (integrator.cpp:85-106) ----
        bool nee = false;
        V2 uLight{0, 0}, uScattering{0, 0};
        if (alive) {
            bsdf_init(S, it, &bs);
            if (bsdf_num_components(S, bs, noSpec) > 0 && S.n_lights > 0) {
                lightNum = select_light(S, it.p, get1d(H, st), &lightPdfSel);
                if (lightPdfSel != 0) {
                    uLight = get2d(H, st);
                    uScattering = get2d(H, st);
                    nee = true;
                }
            }
        }
        SPTX_STAGE(2);
        // ---- stage 2: light sampling strategy (integrator.cpp:119-163) ----
        bool delta = false, wantShadow = false, wantMis = false;
        V3 shO = mk(0, 0, 0), shD = mk(0, 0, 0), misO = mk(0, 0, 0), misD = mk(0, 0, 0);
        LiSample ls;
        ls.pdf = 0.f;
        ls.black = true;
        if (nee) {
            const sptx_light &Lt = S.lights[lightNum];
            delta = Lt.type == SPTX_LIGHT_POINT;
            RefPoint ref{it.p, it.pError, it.n};
            ls = light_sample_li(S, Lt, ref, uLight);
        }
        SPTX_STAGE(4);
        // (Li.IsBlack() is decided per bin by the pass: Recipe::black)
        if (nee && ls.pdf > 0 && !ls.black) {
            bsdf_f(S, bs, it.wo, ls.wi, noSpec, &rc.f);
            rc.ad = absdot(ls.wi, it.ns);
            const float scatteringPdf = bsdf_pdf(S, bs, it.wo, ls.wi, noSpec);
            rc.weight = delta ? 1.f : power_heuristic(ls.pdf, scatteringPdf);
            rc.pdf = ls.pdf;
            rc.row = ls.spectrum;
            rc.rowDiv = ls.divisor;
            rc.work |= W_FA | (delta ? W_FA_DELTA : 0);
            // VisibilityTester(ref, pShape) -> SpawnRayTo(Interaction)
            shO = offset_ray_origin(it.p, it.pError, it.n, ls.lp.p - it.p);
            V3 target = offset_ray_origin(ls.lp.p, ls.lp.pError, ls.lp.n, shO - ls.lp.p);
            shD = target - shO;
            wantShadow = true;
        }
        SPTX_STAGE(6);
        bin_pass(S, A, wr, lane, betaCur, betaNext);
        __syncwarp();
        if (wantShadow && !rc.black) neeFlags |= NEE_HAS_SHADOW;
        rc.work = 0;
        SPTX_STAGE(8);
        // ---- stage 3: BSDF sampling strategy (integrator.cpp:166-213) ----
        if (nee && !delta) {
            float sPdf;
            int sType;
            bsdf_sample_f(S, bs, it.wo, uScattering, noSpec, &misD, &sPdf, &sType, &rc.f);
            if (rc.f.n > 0 && sPdf > 0) {
                float weight = 1.f;
                bool skip = false;
                if (!(sType & BSDF_SPECULAR)) {
                    // (evaluated before the blackness of f is known: no side effects, and
                    //  a black f contributes nothing either way)
                    RefPoint ref{it.p, it.pError, it.n};
                    float lightPdf = light_pdf_li(S, S.lights[lightNum], ref, misD);
                    if (lightPdf == 0)
                        skip = true;
                    else
                        weight = power_heuristic(sPdf, lightPdf);
                }
                if (!skip) {
                    rc.ad = absdot(misD, it.ns);
                    rc.weight = weight;
                    rc.pdf = sPdf;
                    rc.row = S.lights[lightNum].spectrum;
                    rc.work = W_FB;
                    misO = offset_ray_origin(it.p, it.pError, it.n, misD);
                    wantMis = true;
                }
            }
        }
        SPTX_STAGE(10);
        bin_pass(S, A, wr, lane, betaCur, betaNext);
        __syncwarp();
        if (wantMis && !rc.black) neeFlags |= NEE_HAS_MIS;
        rc.work = 0;
        if (neeFlags) {
            A.shA[i] = make_float4(shO.x, shO.y, shO.z, shD.x);
            A.shB[i] = make_float4(shD.y, shD.z, 1.f - kShadowEpsilon, lightPdfSel);
            A.misA[i] = make_float4(misO.x, misO.y, misO.z, misD.x);
            A.misB[i] = make_float4(misD.y, misD.z, __int_as_float(lightNum), __int_as_float(neeFlags));
        }
        SPTX_STAGE(12);
        // ---- stage 4: BSDF sample for the next path segment (path.cpp:131-150) ----
        V3 nextO = mk(0, 0, 0), nextD = mk(0, 0, 0);
        int flagsC = 0;
        if (alive) {
            float pdf;
            bsdf_sample_f(S, bs, it.wo, get2d(H, st), BSDF_ALL, &nextD, &pdf, &flagsC, &rc.f);
            if (rc.f.n > 0 && pdf != 0.f) {
                rc.ad = absdot(nextD, it.ns);
                rc.pdf = pdf;
                if ((flagsC & BSDF_SPECULAR) && (flagsC & BSDF_TRANSMISSION)) {
                    float eta = bs.eta;
                    etaScale *= (dot(it.wo, it.n) > 0) ? (eta * eta) : 1.f / (eta * eta);
                }
                rc.etaScale = etaScale;
                rc.work = W_FC;
                nextO = offset_ray_origin(it.p, it.pError, it.n, nextD);
            } else
                alive = false;
        }
        SPTX_STAGE(14);
        bin_pass(S, A, wr, lane, betaCur, betaNext);
        __syncwarp();
        // ---- epilogue: Russian roulette (path.cpp:176-184), next ray ----
        rc.rrDiv = 0.f;
        if (alive && rc.black) alive = false;  // f.IsBlack() (path.cpp:137)
        if (alive) {
            const float maxBeta = rc.maxBeta;
            if (maxBeta < R.rrThreshold && bounces > 3) {
                float q = fmaxf(.05f, 1.f - maxBeta);
                if (get1d(H, st) < q)
                    alive = false;
                else
                    rc.rrDiv = 1.f - q;
            }
        }
        if (alive) {
            A.rayA[i] = make_float4(nextO.x, nextO.y, nextO.z, nextD.x);
            A.rayB[i] = make_float4(nextD.y, nextD.z, etaScale, rayWeight);
            if (st.dim >= H.s.n_dims) atomicOr(A.counters + 3, ERR_DIM_OVERFLOW);
            A.info[i] = pack_info(bounces + 1, (flagsC & BSDF_SPECULAR) ? 1 : 0, st.dim);
        }
        // beta /= 1 - q for the survivors of the roulette (only when bounces > 3)
        const unsigned needDiv = __ballot_sync(0xffffffffu, rc.path >= 0 && rc.rrDiv != 0.f);
        if (needDiv) {
            __syncwarp();
            for (unsigned m = needDiv; m; m &= m - 1) {
                const Recipe &r = wr[__ffs(m) - 1];
                const size_t at = (size_t)r.path * SPTX_SPECTRUM_STRIDE + lane;
                betaNext[at] = betaNext[at] / r.rrDiv;
            }
        }
        __syncwarp();
        queue_push(A.qNee, A.counters + 2, i, neeFlags != 0);
        queue_push(A.qMis, A.counters + 4, i, (neeFlags & NEE_HAS_MIS) != 0);
        queue_push(A.q[qi ^ 1], A.counters + (qi ^ 1), i, alive);
        SPTX_TICK(16);
    }
#ifdef SPTX_STAGE_CLOCKS
    if (lane == 0)
        for (int c = 0; c < 24; ++c) atomicAdd(&g_stageClocks[c], (unsigned long long)stageClk[c]);
#endif
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->vertices, (unsigned long long)n);
}

}  // namespace sptx
