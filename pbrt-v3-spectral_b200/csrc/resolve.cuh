// resolve.cuh -- K4 (shadow / MIS rays of EstimateDirect) and K5/K6 (film).
#pragma once
#include "kernels.cuh"

namespace sptx {

// ---- K4 ------------------------------------------------------------------------
// The two rays of EstimateDirect.  K4a traces the (few) BSDF-sampled MIS rays as
// closest-hit rays and records whether they reach the sampled light
// (core/integrator.cpp:190-212); K4b traces the shadow rays
// (VisibilityTester::Unoccluded, core/light.cpp:59-61) and adds
// L += beta * (Ld / lightPdf) (integrator.cpp:104-105, path.cpp:122-128).
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, SPTX_TRAV_BLOCKS)
k_mis(DScene S, PathArrays A, DeviceCounters *cnt) {
    TravCounters tc{0, 0};
    const int n = A.counters[4];
    unsigned long long nTraced = 0;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const int i = A.qMis[k];
        const float4 ma = A.misA[i], mb = A.misB[i];
        const int lightNum = __float_as_int(mb.z);
        Ray r{mk(ma.x, ma.y, ma.z), mk(ma.w, mb.x, mb.y), kInf};
        HitRec h;
        bool useB = false;
        // The ray counts only if its closest hit is the sampled light's own shape.  Sphere::Pdf gives every
        // direction the cone's density (shapes/sphere.cpp:294-306: no test of wi against the sphere), so for a
        // sphere light nearly all of these rays pass beside it: the walk's own sphere test, run on the
        // unshortened ray first, says so, and no closer hit can turn a miss of the shape into a hit of it.
        // (Triangle lights never get here with such a ray: Shape::Pdf intersects the triangle.)
#ifndef SPTX_TRACE_DEAD_SEGMENTS
        {
            const sptx_light &Ls = S.lights[lightNum];
            if (S.lights_top_level && Ls.type == SPTX_LIGHT_DIFFUSE_AREA && Ls.shape_kind == SPTX_PRIM_SPHERE) {
                SphereHit sh;
                if (!sphere_test(S.spheres[Ls.shape], r, &sh)) continue;
            }
        }
#endif
        ++nTraced;
        if (!scene_intersect<COUNT>(S, r, &h, &tc)) {
            // the sampled ray leaves the scene: an infinite light's Le(ray) (core/integrator.cpp:204-205)
            useB = S.lights[lightNum].type == SPTX_LIGHT_INFINITE;
        } else {
            if (S.prims[h.prim].light == lightNum) {
                const sptx_light &Lt = S.lights[lightNum];
                if (Lt.two_sided)
                    useB = true;
                else {
                    Isect it;
                    if (build_interaction(S, r, h.prim, h.inst, h.b0, h.b1, h.b2, &it))
                        useB = dot(it.n, -r.d) > 0;
                }
            }
        }
        if (useB) A.misB[i].w = __int_as_float(__float_as_int(mb.w) | NEE_MIS_REACHED_LIGHT);
    }
    if (COUNT) {
        atomicAdd(&cnt->nodes[1], (unsigned long long)tc.nodes);
        atomicAdd(&cnt->tests[1], (unsigned long long)tc.tests);
    }
    for (int o = 16; o > 0; o >>= 1) nTraced += __shfl_down_sync(0xffffffffu, nTraced, o);
    if ((threadIdx.x & 31) == 0 && nTraced) atomicAdd(&cnt->closestRays, nTraced);
}

// (any-hit needs fewer registers than closest-hit: 8 resident blocks at 64 registers measured
//  7 % faster than 6 at 80 on the bench scene, the same on C1)
// FAST (launched beside K3's fast build): Ld / lightPdf as a multiplication by the path's
// correctly rounded reciprocal instead of an IEEE division per bin (11 % of the kernel's instructions).
template <bool COUNT, bool FAST>
__global__ void __launch_bounds__(kBlock, SPTX_ANYHIT_BLOCKS)
k_resolve(DScene S, PathArrays A, int betaIn, int betaOnes, DeviceCounters *cnt) {
    TravCounters tc{0, 0};
    const int n = A.counters[2];
    const float *beta = A.beta[betaIn];
    const int lane = threadIdx.x & 31;
    unsigned long long nShadow = 0;
    for (int k0 = blockIdx.x * blockDim.x; k0 < n; k0 += gridDim.x * blockDim.x) {
        const int k = k0 + threadIdx.x;
        bool useA = false, useB = false;
        int i = 0;
        float lightPdfSel = 1.f;
        if (k < n) {
            i = A.qNee[k];
            const float4 sa = A.shA[i], sb = A.shB[i];
            const int flags = __float_as_int(A.misB[i].w);
            lightPdfSel = FAST ? __frcp_rn(sb.w) : sb.w;
            useB = (flags & NEE_MIS_REACHED_LIGHT) != 0;
            if (flags & NEE_HAS_SHADOW) {
                Ray r{mk(sa.x, sa.y, sa.z), mk(sa.w, sb.x, sb.y), sb.z};
                useA = !scene_intersect_p<COUNT>(S, r, &tc);
                ++nShadow;
            }
        }
        // bin phase: the warp walks its paths, lane = bin; a few paths per round so that
        // their rows are in flight together
        const unsigned mA = __ballot_sync(0xffffffffu, useA);
        const unsigned mB = __ballot_sync(0xffffffffu, useB);
        unsigned m = mA | mB;
        while (m) {
// paths per round (their rows are in flight together).  Under the any-hit walk's 64-register cap:
// K4 per bench step with 1 / 2 / 3 / 4 / 6 paths = 48.4 / 31.5 / 29.6 / 32.6 / 35.2 ms
#ifndef SPTX_NEE_ROUND
#define SPTX_NEE_ROUND 3
#endif
            constexpr int kRound = SPTX_NEE_ROUND;
            size_t at[kRound];
            float sel[kRound], va[kRound], vb[kRound], bt[kRound], Lv[kRound];
            bool on[kRound];
#pragma unroll
            for (int j = 0; j < kRound; ++j) {
                on[j] = m != 0;
                const int p = on[j] ? __ffs(m) - 1 : 0;
                m &= m - 1;
                const int ip = __shfl_sync(0xffffffffu, i, p);
                sel[j] = __shfl_sync(0xffffffffu, lightPdfSel, p);
                at[j] = (size_t)ip * SPTX_SPECTRUM_STRIDE + lane;
                va[j] = (on[j] && ((mA >> p) & 1u)) ? A.fa[at[j]] : 0.f;
                vb[j] = (on[j] && ((mB >> p) & 1u)) ? A.fb[at[j]] : 0.f;
                bt[j] = on[j] ? (betaOnes ? 1.f : beta[at[j]]) : 0.f;
                Lv[j] = on[j] ? A.L[at[j]] : 0.f;
                // Ld = Spectrum(0) + fa + fb in the order EstimateDirect adds them
                float Ld = 0.f;
                if ((mA >> p) & 1u) Ld += va[j];
                if ((mB >> p) & 1u) Ld += vb[j];
                va[j] = Ld;
            }
#pragma unroll
            for (int j = 0; j < kRound; ++j)
                // (lane 31 is the row's pad word: its 0 / sel would take the division's slow path
                //  alone, with the other 31 lanes waiting)
                if (on[j] && lane < SPTX_NBINS)
                    A.L[at[j]] = Lv[j] + bt[j] * (FAST ? va[j] * sel[j] : va[j] / sel[j]);
        }
    }
    if (COUNT) {
        atomicAdd(&cnt->nodes[1], (unsigned long long)tc.nodes);
        atomicAdd(&cnt->tests[1], (unsigned long long)tc.tests);
    }
    // one atomic per warp for the ray counter
    for (int o = 16; o > 0; o >>= 1) nShadow += __shfl_down_sync(0xffffffffu, nShadow, o);
    if (lane == 0 && nShadow) atomicAdd(&cnt->shadowRays, nShadow);
}

// ---- MetadataIntegrator::Li (integrators/metadata.cpp:53-89) ---------------------
// The first hit's depth |p - o|, materialId, instanceId (every bin), or world position
// (bins 0..2); paths that hit nothing keep L = 0.  Runs over the active queue after K2.
__global__ void k_metadata(DScene S, PathArrays A, int qi, int strategy) {
    const int n = A.counters[qi];
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const int i = A.q[qi][k];
        const float4 ra = A.rayA[i], rb = A.rayB[i], ha = A.hitA[i], hb = A.hitB[i];
        const int prim = __float_as_int(ha.x), inst = __float_as_int(ha.y);
        if (prim < 0) continue;
        Ray ray{mk(ra.x, ra.y, ra.z), mk(ra.w, rb.x, rb.y), kInf};
        Isect it;
        if (!build_interaction(S, ray, prim, inst, ha.z, ha.w, hb.x, &it)) {
            atomicOr(A.counters + 3, ERR_INTERACTION);
            continue;
        }
        float v = 0.f;
        if (strategy == SPTX_METADATA_DEPTH)
            v = length(it.p - ray.o);
        else if (strategy == SPTX_METADATA_MATERIAL)
            v = (float)(uint32_t)S.materials[it.material].material_id;
        else if (strategy == SPTX_METADATA_MESH)  // SurfaceInteraction::instanceId defaults to 0
            v = inst >= 0 ? (float)(uint32_t)S.instances[inst].instance_id : 0.f;
        float *Lp = A.L + (size_t)i * SPTX_SPECTRUM_STRIDE;
        if (strategy == SPTX_METADATA_COORDINATES) {
            Lp[0] = it.p.x;
            Lp[1] = it.p.y;
            Lp[2] = it.p.z;
        } else
            for (int b = 0; b < SPTX_NBINS; ++b) Lp[b] = v;
    }
}

// ---- band merge of SpectralPathIntegrator --------------------------------------
// integrators/spectralpath.cpp:284-318: the radiance of a band's path goes through the
// NaN / negative / infinite-luminance guard, then only the band's bins [lo, hi) are copied into
// the sample's spectrum.  One thread per path; after the last band the merged row replaces the
// path's L, which is what the film kernels add.
__global__ void k_band_commit(DScene S, PathArrays A, float *Lacc, int64_t nPaths, int lo, int hi,
                              int first, int last) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nPaths;
         i += (int64_t)gridDim.x * blockDim.x) {
        float4 *Lp = reinterpret_cast<float4 *>(A.L + i * SPTX_SPECTRUM_STRIDE);
        float4 *Ap = reinterpret_cast<float4 *>(Lacc + i * SPTX_SPECTRUM_STRIDE);
        float Lv[SPTX_SPECTRUM_STRIDE], acc[SPTX_SPECTRUM_STRIDE];
#pragma unroll
        for (int c = 0; c < SPTX_SPECTRUM_STRIDE / 4; ++c) {
            const float4 v = Lp[c];
            const float4 a = first ? make_float4(0.f, 0.f, 0.f, 0.f) : Ap[c];
            Lv[4 * c] = v.x, Lv[4 * c + 1] = v.y, Lv[4 * c + 2] = v.z, Lv[4 * c + 3] = v.w;
            acc[4 * c] = a.x, acc[4 * c + 1] = a.y, acc[4 * c + 2] = a.z, acc[4 * c + 3] = a.w;
        }
        float yy = 0.f;
        bool nan = false;
#pragma unroll
        for (int b = 0; b < SPTX_NBINS; ++b) {
            if (isnan(Lv[b])) nan = true;
            yy += S.cie_y[b] * Lv[b];
        }
        const float y = y_finish(S, yy);
        const bool black = nan || y < -1e-5f || isinf(y);
#pragma unroll
        for (int b = 0; b < SPTX_NBINS; ++b)
            if (b >= lo && b < hi) acc[b] = black ? 0.f : Lv[b];
#pragma unroll
        for (int c = 0; c < SPTX_SPECTRUM_STRIDE / 4; ++c) {
            const float4 a = make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
            Ap[c] = a;
            if (last) Lp[c] = a;
        }
    }
}

// ---- K5: film accumulation, one thread per pixel ------------------------------
// The NaN / infinite-luminance guard of SamplerIntegrator::Render
// (core/integrator.cpp:295-316) and FilmTile::AddSample (core/film.h:123-163).
// Samples whose filter footprint is exactly their own pixel (always the case
// for the default box filter of radius 0.5 unless the sample sits on a pixel
// edge) are summed here in sample order, in registers, on top of the pixel's
// current value -- the same float32 sequence as the reference.  Other samples
// are marked for K6.
SPTX_DEV bool sample_footprint(const DFilm &F, float2 pFilm, int *x0, int *y0, int *x1, int *y1) {
    float dx = pFilm.x - 0.5f, dy = pFilm.y - 0.5f;
    int p0x = (int)ceilf(dx - F.f.filter_radius[0]);
    int p0y = (int)ceilf(dy - F.f.filter_radius[1]);
    int p1x = (int)floorf(dx + F.f.filter_radius[0]) + 1;
    int p1y = (int)floorf(dy + F.f.filter_radius[1]) + 1;
    *x0 = max(p0x, F.f.cropped[0]);
    *y0 = max(p0y, F.f.cropped[1]);
    *x1 = min(p1x, F.f.cropped[2]);
    *y1 = min(p1y, F.f.cropped[3]);
    return *x0 < *x1 && *y0 < *y1;
}
SPTX_DEV float filter_weight(const DFilm &F, float2 pFilm, int x, int y) {
    const int W = SPTX_FILTER_TABLE_WIDTH;
    float dx = pFilm.x - 0.5f, dy = pFilm.y - 0.5f;
    float fx = fabsf(((float)x - dx) * (1.f / F.f.filter_radius[0]) * (float)W);
    float fy = fabsf(((float)y - dy) * (1.f / F.f.filter_radius[1]) * (float)W);
    int ix = min((int)floorf(fx), W - 1), iy = min((int)floorf(fy), W - 1);
    return F.f.filter_table[iy * W + ix];
}

// Applies the radiance guard and the luminance clamp; returns the scale to
// apply to L (0 = black it out) -- y is computed once.
SPTX_DEV float guard_scale(const DScene &S, const DFilm &F, const PathArrays &A, size_t i) {
    const float *Lp = A.L + i * SPTX_SPECTRUM_STRIDE;
    float yy = 0.f;
    bool nan = false;
    for (int b = 0; b < SPTX_NBINS; ++b) {
        float v = Lp[b];
        if (isnan(v)) nan = true;
        yy += S.cie_y[b] * v;
    }
    if (nan) return 0.f;
    float y = y_finish(S, yy);
    if (isinf(y)) return 0.f;
    if (y > F.f.max_sample_luminance) return F.f.max_sample_luminance / y;
    return 1.f;
}

__global__ void __launch_bounds__(kBlock)
k_film(DScene S, DFilm F, PathArrays A, RenderParams R) {
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < R.nPix;
         pix += gridDim.x * blockDim.x) {
        const int2 pxy = R.pixels[pix];
        const bool inFilm = pxy.x >= F.f.cropped[0] && pxy.x < F.f.cropped[2] &&
                            pxy.y >= F.f.cropped[1] && pxy.y < F.f.cropped[3];
        float acc[SPTX_NBINS + 1];
        float *px = nullptr;
        if (inFilm) {
            px = F.pixels + ((size_t)(pxy.y - F.f.cropped[1]) * F.width + (pxy.x - F.f.cropped[0])) *
                                SPTX_SPECTRUM_STRIDE;
#pragma unroll
            for (int c = 0; c < SPTX_SPECTRUM_STRIDE / 4; ++c) {
                const float4 v = reinterpret_cast<const float4 *>(px)[c];
                acc[4 * c] = v.x;
                acc[4 * c + 1] = v.y;
                acc[4 * c + 2] = v.z;
                acc[4 * c + 3] = v.w;
            }
        }
        bool touched = false;
        for (int s = 0; s < R.samplesInPass; ++s) {
            const size_t i = (size_t)s * R.nPix + pix;
            const float2 pf = A.pfilm[i];
            int x0, y0, x1, y1;
            if (!sample_footprint(F, pf, &x0, &y0, &x1, &y1)) continue;
            const bool own = inFilm && x0 == pxy.x && y0 == pxy.y && x1 == x0 + 1 && y1 == y0 + 1;
            if (!own) {
                A.spill[i] = 1;
                continue;
            }
            // the sample's radiance row, moved once (8 x 128 bit) and kept in registers
            float Lv[SPTX_SPECTRUM_STRIDE];
            {
                const float4 *Lp = reinterpret_cast<const float4 *>(A.L + i * SPTX_SPECTRUM_STRIDE);
#pragma unroll
                for (int c = 0; c < SPTX_SPECTRUM_STRIDE / 4; ++c) {
                    const float4 v = Lp[c];
                    Lv[4 * c] = v.x;
                    Lv[4 * c + 1] = v.y;
                    Lv[4 * c + 2] = v.z;
                    Lv[4 * c + 3] = v.w;
                }
            }
            // guard of SamplerIntegrator::Render (core/integrator.cpp:295-316), as guard_scale()
            float g = 1.f;
            {
                float yy = 0.f;
                bool nan = false;
#pragma unroll
                for (int b = 0; b < SPTX_NBINS; ++b) {
                    if (isnan(Lv[b])) nan = true;
                    yy += S.cie_y[b] * Lv[b];
                }
                const float y = y_finish(S, yy);
                if (nan || isinf(y))
                    g = 0.f;
                else if (y > F.f.max_sample_luminance)
                    g = F.f.max_sample_luminance / y;
            }
            const float w = A.rayB[i].w;
            const float fw = filter_weight(F, pf, x0, y0);
#pragma unroll
            for (int b = 0; b < SPTX_NBINS; ++b) {
                float v = Lv[b];
                v = (g == 1.f) ? v : ((g == 0.f) ? 0.f : v * g);
                acc[b] += v * w * fw;
            }
            acc[SPTX_NBINS] += fw;
            touched = true;
        }
        if (touched) {
#pragma unroll
            for (int c = 0; c < SPTX_SPECTRUM_STRIDE / 4; ++c)
                reinterpret_cast<float4 *>(px)[c] =
                    make_float4(acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
        }
    }
}

// ---- K6: samples whose footprint covers other pixels (wide filters, pixel
// edges): atomic splat, the device analogue of MergeFilmTile's overlap sums.
__global__ void k_film_spill(DScene S, DFilm F, PathArrays A, RenderParams R) {
    const int64_t nPaths = (int64_t)R.nPix * R.samplesInPass;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nPaths;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (!A.spill[i]) continue;
        const float2 pf = A.pfilm[i];
        int x0, y0, x1, y1;
        if (!sample_footprint(F, pf, &x0, &y0, &x1, &y1)) continue;
        const float g = guard_scale(S, F, A, (size_t)i);
        const float w = A.rayB[i].w;
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
                const float fw = filter_weight(F, pf, x, y);
                float *px = F.pixels + ((size_t)(y - F.f.cropped[1]) * F.width + (x - F.f.cropped[0])) *
                                           SPTX_SPECTRUM_STRIDE;
                for (int b = 0; b < SPTX_NBINS; ++b) {
                    float v = A.L[(size_t)i * SPTX_SPECTRUM_STRIDE + b];
                    v = (g == 1.f) ? v : ((g == 0.f) ? 0.f : v * g);
                    atomicAdd(px + b, v * w * fw);
                }
                atomicAdd(px + SPTX_NBINS, fw);
            }
    }
}

}  // namespace sptx
