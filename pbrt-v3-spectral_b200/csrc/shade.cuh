// shade.cuh -- K3: one path vertex of PathIntegrator::Li per thread.
//
// integrators/path.cpp:88-184 for the vertex, core/integrator.cpp:85-215
// (UniformSampleOneLight, EstimateDirect) for next-event estimation.  The two
// rays EstimateDirect traces (shadow ray, BSDF-sampled MIS ray) are not traced
// here: they are written out with their per-bin contributions (fa, fb) and
// traced by K4, which then adds beta * (fa + fb) / lightPdf to L.
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

__global__ void __launch_bounds__(kBlock)
k_shade(DScene S, DSampler H, PathArrays A, RenderParams R, int qi, int betaIn,
        DeviceCounters *cnt) {
    const int n = A.counters[qi];
    const int *queue = A.q[qi];
    const size_t P = (size_t)A.capacity;
    const float *betaCur = A.beta[betaIn];
    float *betaNext = A.beta[betaIn ^ 1];
    for (int k0 = blockIdx.x * blockDim.x; k0 < n; k0 += gridDim.x * blockDim.x) {
        const int k = k0 + threadIdx.x;
        bool alive = k < n;
        bool wantNee = false;
        int i = 0;
        if (alive) {
            i = queue[k];
            const float4 ra = A.rayA[i], rb = A.rayB[i];
            const float4 ha = A.hitA[i], hb = A.hitB[i];
            Ray ray{mk(ra.x, ra.y, ra.z), mk(ra.w, rb.x, rb.y), kInf};
            float etaScale = rb.z;
            const float rayWeight = rb.w;
            const uint32_t info = A.info[i];
            int bounces = (int)(info & 0xffu);
            const bool specularBounce = ((info >> 8) & 1u) != 0;
            SampleStream st{A.sindex[i], (int)(info >> 16)};
            const int prim = __float_as_int(ha.x), inst = __float_as_int(ha.y);
            const bool found = prim >= 0;
            Isect it;
            bool haveIsect = false;
            if (found) {
                haveIsect = build_interaction(S, ray, prim, inst, ha.z, ha.w, hb.x, &it);
                if (!haveIsect) atomicOr(A.counters + 3, ERR_INTERACTION);
            }
            // Emission at the first vertex / after a specular bounce (path.cpp:91-101)
            if ((bounces == 0 || specularBounce) && haveIsect && it.light >= 0) {
                const sptx_light &Lt = S.lights[it.light];
                // isect.Le(-ray.d) -> DiffuseAreaLight::L(intr, w)
                if (Lt.two_sided || dot(it.n, -ray.d) > 0)
                    for (int b = 0; b < SPTX_NBINS; ++b)
                        A.L[b * P + i] += betaCur[b * P + i] * spec(S, Lt.spectrum, b);
            }
            alive = haveIsect && bounces < R.maxDepth;
            if (alive) {
                Bsdf bs;
                bsdf_init(S, it, &bs);
                const V3 wo = it.wo;
                // ---- next-event estimation (UniformSampleOneLight) ----
                if (bsdf_num_components(S, bs, BSDF_ALL & ~BSDF_SPECULAR) > 0 && S.n_lights > 0) {
                    float lightPdfSel;
                    int lightNum = select_light(S, it.p, get1d(H, st), &lightPdfSel);
                    if (lightPdfSel != 0) {
                        const V2 uLight = get2d(H, st);
                        const V2 uScattering = get2d(H, st);
                        const sptx_light &Lt = S.lights[lightNum];
                        const bool delta = Lt.type == SPTX_LIGHT_POINT;
                        const int noSpec = BSDF_ALL & ~BSDF_SPECULAR;
                        RefPoint ref{it.p, it.pError, it.n};
                        int neeFlags = 0;
                        // light sampling strategy (integrator.cpp:119-163)
                        LiSample ls = light_sample_li(S, Lt, ref, uLight);
                        bool liBlack = ls.black;
                        if (!liBlack) {
                            liBlack = true;
                            for (int b = 0; b < SPTX_NBINS; ++b)
                                if (li_bin(S, ls, b) != 0.f) liBlack = false;
                        }
                        if (ls.pdf > 0 && !liBlack) {
                            FSum f;
                            bsdf_f(S, bs, wo, ls.wi, noSpec, &f);
                            const float ad = absdot(ls.wi, it.ns);
                            const float scatteringPdf = bsdf_pdf(S, bs, wo, ls.wi, noSpec);
                            const float weight = delta ? 1.f : power_heuristic(ls.pdf, scatteringPdf);
                            bool fBlack = true;
                            for (int b = 0; b < SPTX_NBINS; ++b) {
                                float fv = fsum_bin(S, f, b) * ad;
                                if (fv != 0.f) fBlack = false;
                                float c = fv * li_bin(S, ls, b);
                                if (!delta) c = c * weight;
                                A.fa[b * P + i] = c / ls.pdf;
                            }
                            if (!fBlack) {
                                // VisibilityTester(ref, pShape) -> SpawnRayTo(Interaction)
                                V3 origin = offset_ray_origin(it.p, it.pError, it.n, ls.lp.p - it.p);
                                V3 target =
                                    offset_ray_origin(ls.lp.p, ls.lp.pError, ls.lp.n, origin - ls.lp.p);
                                V3 d = target - origin;
                                A.shA[i] = make_float4(origin.x, origin.y, origin.z, d.x);
                                A.shB[i] = make_float4(d.y, d.z, 1.f - kShadowEpsilon, lightPdfSel);
                                neeFlags |= NEE_HAS_SHADOW;
                            }
                        }
                        // BSDF sampling strategy (integrator.cpp:166-213)
                        if (!delta) {
                            FSum f2;
                            V3 wi2;
                            float sPdf;
                            int sType;
                            bsdf_sample_f(S, bs, wo, uScattering, noSpec, &wi2, &sPdf, &sType, &f2);
                            if (f2.n > 0 && sPdf > 0) {
                                const float ad2 = absdot(wi2, it.ns);
                                bool f2Black = true;
                                for (int b = 0; b < SPTX_NBINS; ++b)
                                    if (fsum_bin(S, f2, b) * ad2 != 0.f) f2Black = false;
                                if (!f2Black) {
                                    float weight = 1.f;
                                    bool skip = false;
                                    if (!(sType & BSDF_SPECULAR)) {
                                        float lightPdf = light_pdf_li(S, Lt, ref, wi2);
                                        if (lightPdf == 0)
                                            skip = true;
                                        else
                                            weight = power_heuristic(sPdf, lightPdf);
                                    }
                                    if (!skip) {
                                        for (int b = 0; b < SPTX_NBINS; ++b) {
                                            float fv = fsum_bin(S, f2, b) * ad2;
                                            A.fb[b * P + i] =
                                                fv * spec(S, Lt.spectrum, b) * weight / sPdf;
                                        }
                                        V3 o2 = offset_ray_origin(it.p, it.pError, it.n, wi2);
                                        A.misA[i] = make_float4(o2.x, o2.y, o2.z, wi2.x);
                                        A.misB[i] = make_float4(wi2.y, wi2.z, __int_as_float(lightNum),
                                                                lightPdfSel);
                                        neeFlags |= NEE_HAS_MIS;
                                    }
                                }
                            }
                        }
                        if (neeFlags) {
                            A.shB[i].w = lightPdfSel;
                            A.misB[i].w = __int_as_float(neeFlags);
                            if (!(neeFlags & NEE_HAS_MIS)) A.misB[i].z = __int_as_float(lightNum);
                            wantNee = true;
                        }
                    }
                }
                // ---- BSDF sample for the next path segment (path.cpp:131-150) ----
                FSum f;
                V3 wi;
                float pdf;
                int flags;
                bsdf_sample_f(S, bs, wo, get2d(H, st), BSDF_ALL, &wi, &pdf, &flags, &f);
                bool fBlack = true;
                float maxBeta = 0.f;
                if (f.n > 0 && pdf != 0.f) {
                    const float ad = absdot(wi, it.ns);
                    for (int b = 0; b < SPTX_NBINS; ++b) {
                        float fv = fsum_bin(S, f, b);
                        if (fv != 0.f) fBlack = false;
                        float nb = betaCur[b * P + i] * (fv * ad / pdf);
                        betaNext[b * P + i] = nb;
                    }
                }
                if (fBlack || pdf == 0.f)
                    alive = false;
                else {
                    const bool spec = (flags & BSDF_SPECULAR) != 0;
                    if (spec && (flags & BSDF_TRANSMISSION)) {
                        float eta = bs.eta;
                        etaScale *= (dot(wo, it.n) > 0) ? (eta * eta) : 1.f / (eta * eta);
                    }
                    V3 o = offset_ray_origin(it.p, it.pError, it.n, wi);
                    // Russian roulette (path.cpp:176-184)
                    for (int b = 0; b < SPTX_NBINS; ++b) {
                        float v = betaNext[b * P + i] * etaScale;
                        maxBeta = (b == 0) ? v : ((maxBeta < v) ? v : maxBeta);
                    }
                    if (maxBeta < R.rrThreshold && bounces > 3) {
                        float q = fmaxf(.05f, 1.f - maxBeta);
                        if (get1d(H, st) < q)
                            alive = false;
                        else {
                            const float inv = 1.f - q;
                            for (int b = 0; b < SPTX_NBINS; ++b) betaNext[b * P + i] /= inv;
                        }
                    }
                    if (alive) {
                        A.rayA[i] = make_float4(o.x, o.y, o.z, wi.x);
                        A.rayB[i] = make_float4(wi.y, wi.z, etaScale, rayWeight);
                        if (st.dim >= H.s.n_dims) atomicOr(A.counters + 3, ERR_DIM_OVERFLOW);
                        A.info[i] = pack_info(bounces + 1, spec ? 1 : 0, st.dim);
                    }
                }
            }
        }
        queue_push(A.qNee, A.counters + 2, i, wantNee);
        queue_push(A.q[qi ^ 1], A.counters + (qi ^ 1), i, alive);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->vertices, (unsigned long long)n);
}

}  // namespace sptx
