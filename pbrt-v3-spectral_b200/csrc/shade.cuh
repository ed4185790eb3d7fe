// shade.cuh -- K3: one path vertex of PathIntegrator::Li.
//
// integrators/path.cpp:88-184 for the vertex, core/integrator.cpp:85-215
// (UniformSampleOneLight, EstimateDirect) for next-event estimation.
//
// Two phases per warp, so that the 31-bin arithmetic never diverges:
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

// What the bin phase has to do for one path.
enum {
    W_EMIT = 1,   // L += beta * Le
    W_FA = 2,     // fa = f * |wi.ns| * Li [* weight] / lightPdf   (light-sampling strategy)
    W_FB = 4,     // fb = f2 * |wi2.ns| * Le * weight / scatteringPdf (BSDF-sampling strategy)
    W_FC = 8,     // betaNext = beta * (f * |wi.ns| / pdf); max(betaNext * etaScale)
    W_FA_DELTA = 16,
    W_RRDIV = 32  // betaNext /= rrDiv (second bin phase)
};

struct Recipe {
    int path;       // path id, -1 = nothing to do
    int work;       // W_* flags
    int emitRow;    // Le spectrum row
    int liRow;      // light-sample radiance row
    float liDiv;    // 0, or divide the row by it (point light: I / d^2)
    float adA, weightA, pdfA;
    int lbRow;      // Lemit row of the MIS light
    float adB, weightB, pdfB;
    float adC, pdfC, etaScale;
    float rrDiv;
    // results of the bin phase
    int blackMask;   // bit 0: fa's f or Li black, bit 1: fb's f black, bit 2: fc's f black
    float maxBeta;
    FSum fa, fb, fc;
};

__global__ void __launch_bounds__(kBlock)
k_shade(DScene S, DSampler H, PathArrays A, RenderParams R, int qi, int betaIn,
        DeviceCounters *cnt) {
    extern __shared__ unsigned char smemRaw[];
    Recipe *recipes = reinterpret_cast<Recipe *>(smemRaw);
    const int n = A.counters[qi];
    const int *queue = A.q[qi];
    const float *betaCur = A.beta[betaIn];
    float *betaNext = A.beta[betaIn ^ 1];
    const int lane = threadIdx.x & 31;
    const int warpBase = threadIdx.x & ~31;
    for (int k0 = blockIdx.x * blockDim.x; k0 < n; k0 += gridDim.x * blockDim.x) {
        const int k = k0 + threadIdx.x;
        bool alive = k < n;
        int i = 0;
        Recipe &rc = recipes[threadIdx.x];
        rc.path = -1;
        rc.work = 0;
        // state that survives the bin phase
        SampleStream st{0, 0};
        int bounces = 0, neeFlags = 0, lightNum = 0, flagsC = 0;
        float lightPdfSel = 0.f, etaScale = 1.f, rayWeight = 0.f;
        bool wantShadow = false, wantMis = false;
        V3 shO = mk(0, 0, 0), shD = mk(0, 0, 0), misO = mk(0, 0, 0), misD = mk(0, 0, 0);
        V3 nextO = mk(0, 0, 0), nextD = mk(0, 0, 0);
        // ------------------------------ scalar phase ------------------------------
        if (alive) {
            i = queue[k];
            rc.path = i;
            const float4 ra = A.rayA[i], rb = A.rayB[i];
            const float4 ha = A.hitA[i], hb = A.hitB[i];
            Ray ray{mk(ra.x, ra.y, ra.z), mk(ra.w, rb.x, rb.y), kInf};
            etaScale = rb.z;
            rayWeight = rb.w;
            const uint32_t info = A.info[i];
            bounces = (int)(info & 0xffu);
            const bool specularBounce = ((info >> 8) & 1u) != 0;
            st.index = A.sindex[i];
            st.dim = (int)(info >> 16);
            const int prim = __float_as_int(ha.x), inst = __float_as_int(ha.y);
            Isect it;
            bool haveIsect = false;
            if (prim >= 0) {
                haveIsect = build_interaction(S, ray, prim, inst, ha.z, ha.w, hb.x, &it);
                if (!haveIsect) atomicOr(A.counters + 3, ERR_INTERACTION);
            }
            // Emission at the first vertex / after a specular bounce (path.cpp:91-101):
            // isect.Le(-ray.d) -> DiffuseAreaLight::L(intr, w)
            if ((bounces == 0 || specularBounce) && haveIsect && it.light >= 0) {
                const sptx_light &Lt = S.lights[it.light];
                if (Lt.two_sided || dot(it.n, -ray.d) > 0) {
                    rc.work |= W_EMIT;
                    rc.emitRow = Lt.spectrum;
                }
            }
            alive = haveIsect && bounces < R.maxDepth;
            if (alive) {
                Bsdf bs;
                bsdf_init(S, it, &bs);
                const V3 wo = it.wo;
                // ---- next-event estimation (UniformSampleOneLight, integrator.cpp:85-106) ----
                if (bsdf_num_components(S, bs, BSDF_ALL & ~BSDF_SPECULAR) > 0 && S.n_lights > 0) {
                    lightNum = select_light(S, it.p, get1d(H, st), &lightPdfSel);
                    if (lightPdfSel != 0) {
                        const V2 uLight = get2d(H, st);
                        const V2 uScattering = get2d(H, st);
                        const sptx_light &Lt = S.lights[lightNum];
                        const bool delta = Lt.type == SPTX_LIGHT_POINT;
                        const int noSpec = BSDF_ALL & ~BSDF_SPECULAR;
                        RefPoint ref{it.p, it.pError, it.n};
                        // light sampling strategy (integrator.cpp:119-163)
                        LiSample ls = light_sample_li(S, Lt, ref, uLight);
                        // (Li.IsBlack() is decided per bin below: blackMask bit 0)
                        if (ls.pdf > 0 && !ls.black) {
                            bsdf_f(S, bs, wo, ls.wi, noSpec, &rc.fa);
                            rc.adA = absdot(ls.wi, it.ns);
                            const float scatteringPdf = bsdf_pdf(S, bs, wo, ls.wi, noSpec);
                            rc.weightA = delta ? 1.f : power_heuristic(ls.pdf, scatteringPdf);
                            rc.pdfA = ls.pdf;
                            rc.liRow = ls.spectrum;
                            rc.liDiv = ls.divisor;
                            rc.work |= W_FA | (delta ? W_FA_DELTA : 0);
                            // VisibilityTester(ref, pShape) -> SpawnRayTo(Interaction)
                            shO = offset_ray_origin(it.p, it.pError, it.n, ls.lp.p - it.p);
                            V3 target = offset_ray_origin(ls.lp.p, ls.lp.pError, ls.lp.n, shO - ls.lp.p);
                            shD = target - shO;
                            wantShadow = true;
                        }
                        // BSDF sampling strategy (integrator.cpp:166-213)
                        if (!delta) {
                            float sPdf;
                            int sType;
                            bsdf_sample_f(S, bs, wo, uScattering, noSpec, &misD, &sPdf, &sType, &rc.fb);
                            if (rc.fb.n > 0 && sPdf > 0) {
                                float weight = 1.f;
                                bool skip = false;
                                if (!(sType & BSDF_SPECULAR)) {
                                    // (evaluated before the blackness of f is known; a black f
                                    //  contributes 0 either way)
                                    float lightPdf = light_pdf_li(S, Lt, ref, misD);
                                    if (lightPdf == 0)
                                        skip = true;
                                    else
                                        weight = power_heuristic(sPdf, lightPdf);
                                }
                                if (!skip) {
                                    rc.adB = absdot(misD, it.ns);
                                    rc.weightB = weight;
                                    rc.pdfB = sPdf;
                                    rc.lbRow = Lt.spectrum;
                                    rc.work |= W_FB;
                                    misO = offset_ray_origin(it.p, it.pError, it.n, misD);
                                    wantMis = true;
                                }
                            }
                        }
                    }
                }
                // ---- BSDF sample for the next path segment (path.cpp:131-150) ----
                float pdf;
                bsdf_sample_f(S, bs, wo, get2d(H, st), BSDF_ALL, &nextD, &pdf, &flagsC, &rc.fc);
                if (rc.fc.n > 0 && pdf != 0.f) {
                    rc.adC = absdot(nextD, it.ns);
                    rc.pdfC = pdf;
                    if ((flagsC & BSDF_SPECULAR) && (flagsC & BSDF_TRANSMISSION)) {
                        float eta = bs.eta;
                        etaScale *= (dot(wo, it.n) > 0) ? (eta * eta) : 1.f / (eta * eta);
                    }
                    rc.etaScale = etaScale;
                    rc.work |= W_FC;
                    nextO = offset_ray_origin(it.p, it.pError, it.n, nextD);
                } else
                    alive = false;
            }
        }
        __syncwarp();
        // ------------------------------ bin phase ------------------------------
        for (int p = 0; p < 32; ++p) {
            const Recipe &r = recipes[warpBase + p];
            const int work = r.work;
            if (r.path < 0 || work == 0) continue;
            const size_t base = (size_t)r.path * SPTX_SPECTRUM_STRIDE + lane;
            const bool live = lane < SPTX_NBINS;
            const float beta = betaCur[base];
            if (work & W_EMIT) {
                if (live) A.L[base] += beta * spec(S, r.emitRow, lane);
            }
            int black = 0;
            if (work & W_FA) {
                float fv = live ? fsum_bin(S, r.fa, lane) * r.adA : 0.f;
                float li = live ? spec(S, r.liRow, lane) : 0.f;
                if (r.liDiv != 0.f) li = li / r.liDiv;
                float c = fv * li;
                if (!(work & W_FA_DELTA)) c = c * r.weightA;
                A.fa[base] = c / r.pdfA;
                if (__ballot_sync(0xffffffffu, fv != 0.f) == 0) black |= 1;
                if (__ballot_sync(0xffffffffu, li != 0.f) == 0) black |= 1;
            }
            if (work & W_FB) {
                float fv = live ? fsum_bin(S, r.fb, lane) * r.adB : 0.f;
                float le = live ? spec(S, r.lbRow, lane) : 0.f;
                A.fb[base] = fv * le * r.weightB / r.pdfB;
                if (__ballot_sync(0xffffffffu, fv != 0.f) == 0) black |= 2;
            }
            float mb = 0.f;
            if (work & W_FC) {
                float fv = live ? fsum_bin(S, r.fc, lane) : 0.f;
                float nb = beta * (fv * r.adC / r.pdfC);
                betaNext[base] = nb;
                if (__ballot_sync(0xffffffffu, fv != 0.f) == 0) black |= 4;
                // MaxComponentValue(beta * etaScale): max is order-independent
                float v = live ? nb * r.etaScale : -kInf;
                for (int o = 16; o > 0; o >>= 1) {
                    float w = __shfl_xor_sync(0xffffffffu, v, o);
                    v = (v < w) ? w : v;
                }
                mb = v;
            }
            if (lane == 0) {
                recipes[warpBase + p].blackMask = black;
                recipes[warpBase + p].maxBeta = mb;
            }
        }
        __syncwarp();
        // ------------------------------ scalar epilogue ------------------------------
        bool wantNee = false;
        rc.rrDiv = 0.f;
        if (rc.path >= 0) {
            const int black = rc.blackMask;
            if (wantShadow && !(black & 1)) neeFlags |= NEE_HAS_SHADOW;
            if (wantMis && !(black & 2)) neeFlags |= NEE_HAS_MIS;
            if (neeFlags) {
                A.shA[i] = make_float4(shO.x, shO.y, shO.z, shD.x);
                A.shB[i] = make_float4(shD.y, shD.z, 1.f - kShadowEpsilon, lightPdfSel);
                A.misA[i] = make_float4(misO.x, misO.y, misO.z, misD.x);
                A.misB[i] = make_float4(misD.y, misD.z, __int_as_float(lightNum), __int_as_float(neeFlags));
                wantNee = true;
            }
            if (alive && (black & 4)) alive = false;  // f.IsBlack() (path.cpp:137)
            if (alive) {
                // Russian roulette (path.cpp:176-184)
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
        }
        // beta /= 1 - q for the survivors of the roulette (rare: bounces > 3)
        const unsigned needDiv = __ballot_sync(0xffffffffu, rc.path >= 0 && rc.rrDiv != 0.f);
        if (needDiv) {
            __syncwarp();
            for (unsigned m = needDiv; m; m &= m - 1) {
                const int p = __ffs(m) - 1;
                const Recipe &r = recipes[warpBase + p];
                const size_t base = (size_t)r.path * SPTX_SPECTRUM_STRIDE + lane;
                betaNext[base] = betaNext[base] / r.rrDiv;
            }
        }
        __syncwarp();
        queue_push(A.qNee, A.counters + 2, i, wantNee);
        queue_push(A.q[qi ^ 1], A.counters + (qi ^ 1), i, alive);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->vertices, (unsigned long long)n);
}

}  // namespace sptx
