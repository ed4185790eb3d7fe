// serial.cuh -- the RandomSampler and ZeroTwoSequenceSampler paths: one warp lane per 16x16 tile.
//
// RandomSampler is ONE PCG32 stream per tile (SamplerIntegrator::Render clones
// the sampler with seed = tile.y * nTiles.x + tile.x, core/integrator.cpp:
// 245-249; RandomSampler::Clone -> RNG::SetSequence, samplers/random.cpp:57-61),
// consumed by the tile's pixels in row-major order, sample after sample, with a
// path-dependent number of draws per sample.  A sample's numbers are therefore
// only known once every earlier sample of its tile has run: to reproduce the
// reference's stream a tile is rendered serially, path after path, by one lane.
// Tiles are independent, so the grid is one single-lane warp per tile (no
// divergence inside a warp; 4096 warps in flight for a 1024x1024 image).  The
// vertex code is the same device functions the wavefront uses.
#pragma once
#include "resolve.cuh"

namespace sptx {

// RNG, core/rng.h:64-144.
struct Pcg32 {
    uint64_t state, inc;
    SPTX_DEV uint32_t next() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    SPTX_DEV void set_sequence(uint64_t seq) {
        state = 0u;
        inc = (seq << 1u) | 1u;
        next();
        state += 0x853c49e6748fea9bULL;
        next();
    }
    SPTX_DEV float uniform_float() {
        float v = (float)next() * 0x1p-32f;
        return (kOneMinusEpsilon < v) ? kOneMinusEpsilon : v;  // std::min(OneMinusEpsilon, v)
    }
    SPTX_DEV V2 get2d() {
        V2 p;
        p.x = uniform_float();  // brace-init list: left to right
        p.y = uniform_float();
        return p;
    }
    SPTX_DEV float get1d() { return uniform_float(); }
    SPTX_DEV void start_pixel() {}  // RandomSampler::StartPixel draws nothing (no array requests)
    SPTX_DEV void start_sample(int) {}
    // RNG::UniformUInt32(b), core/rng.h:86-103
    SPTX_DEV uint32_t next_bounded(uint32_t b) {
        const uint32_t threshold = (~b + 1u) % b;
        while (true) {
            const uint32_t r = next();
            if (r >= threshold) return r % b;
        }
    }
};

// ZeroTwoSequenceSampler (samplers/zerotwosequence.cpp:43-70) + PixelSampler (core/sampler.cpp:
// 107-135): at the start of every pixel the first nDims 1D and nDims 2D dimensions of all of the
// pixel's samples are generated from the tile's RNG stream -- a scrambled van der Corput / 2D
// Sobol' (0,2)-sequence in Gray-code order (core/lowdiscrepancy.h:112-139, 153-228), shuffled with
// RNG::UniformUInt32(b) draws (core/sampling.h:151-157) -- and dimensions past nDims come straight
// from the RNG.  The tables ([nDims][spp] floats and [nDims][spp] points) live in a per-tile
// slice of a global scratch buffer.
struct ZeroTwoSampler {
    Pcg32 rng;
    int nDims, spp;
    float *t1;   // [nDims][spp]
    float2 *t2;  // [nDims][spp]
    int cur1, cur2, index;
    SPTX_DEV void set_sequence(uint64_t seq) { rng.set_sequence(seq); }
    SPTX_DEV static float to_float(uint32_t v) {
        const float f = (float)v * 0x1p-32f;
        return (kOneMinusEpsilon < f) ? kOneMinusEpsilon : f;
    }
    // Shuffle(samp, count, nDimensions = 1, rng)
    template <typename T>
    SPTX_DEV void shuffle(T *samp, int count) {
        for (int i = 0; i < count; ++i) {
            const int other = i + (int)rng.next_bounded((uint32_t)(count - i));
            const T tmp = samp[i];
            samp[i] = samp[other];
            samp[other] = tmp;
        }
    }
    SPTX_DEV void start_pixel() {
        for (int d = 0; d < nDims; ++d) {  // VanDerCorput(1, spp, ...)
            float *p = t1 + (size_t)d * spp;
            uint32_t v = rng.next();  // scramble
            for (uint32_t i = 0; i < (uint32_t)spp; ++i) {
                p[i] = to_float(v);
                v ^= 0x80000000u >> __ffs((int)(i + 1)) - 1;  // CVanDerCorput[CountTrailingZeros(i + 1)]
            }
            for (int i = 0; i < spp; ++i) (void)rng.next_bounded(1u);  // Shuffle(samples + i, 1, 1, rng)
            shuffle(p, spp);
        }
        for (int d = 0; d < nDims; ++d) {  // Sobol2D(1, spp, ...)
            float2 *p = t2 + (size_t)d * spp;
            uint32_t v0 = rng.next(), v1 = rng.next();
            for (uint32_t i = 0; i < (uint32_t)spp; ++i) {
                p[i] = make_float2(to_float(v0), to_float(v1));
                const int k = __ffs((int)(i + 1)) - 1;
                v0 ^= 0x80000000u >> k;  // CSobol[0][k]
                // CSobol[1][k]: column k of the second Sobol' matrix = row k of Pascal's triangle
                // mod 2, top bit first (core/lowdiscrepancy.h:207-219)
                uint32_t c1 = 0x80000000u;
                for (int j = 0; j < k; ++j) c1 ^= c1 >> 1;
                v1 ^= c1;
            }
            for (int i = 0; i < spp; ++i) (void)rng.next_bounded(1u);
            shuffle(p, spp);
        }
    }
    SPTX_DEV void start_sample(int s) {
        index = s;
        cur1 = cur2 = 0;
    }
    SPTX_DEV float get1d() {
        if (cur1 < nDims) return t1[(size_t)(cur1++) * spp + index];
        return rng.uniform_float();
    }
    SPTX_DEV V2 get2d() {
        if (cur2 < nDims) {
            const float2 v = t2[(size_t)(cur2++) * spp + index];
            return V2{v.x, v.y};
        }
        // PixelSampler::Get2D returns Point2f(rng.UniformFloat(), rng.UniformFloat()): a constructor
        // call, whose arguments the reference's compiler (g++, x86-64) evaluates right to left --
        // y takes the first draw.  (RandomSampler::Get2D is a brace list: left to right.)
        V2 p;
        p.y = rng.uniform_float();
        p.x = rng.uniform_float();
        return p;
    }
};

struct SerialParams {
    int sb[4];       // film sample bounds
    int pb[4];       // integrator pixel bounds
    int nTilesX, nTilesY;
    int spp;
    int maxDepth;
    float rrThreshold;
    float wavelength;
    int nDims;      // ZeroTwoSequenceSampler: nSampledDimensions
    float *tables;  // ZeroTwoSequenceSampler: [nTiles][nDims][spp] floats + [nTiles][nDims][spp] points
};

// PathIntegrator::Li, integrators/path.cpp:64-188, for one camera ray.
template <typename SMP>
SPTX_DEV void li_serial(const DScene &S, const SerialParams &P, SMP &rng, Ray ray, float *L,
                        unsigned long long *nClosest,
                        unsigned long long *nShadow, unsigned long long *nVert) {
    float beta[SPTX_SPECTRUM_STRIDE];
    for (int b = 0; b < SPTX_SPECTRUM_STRIDE; ++b) {
        beta[b] = 1.f;
        L[b] = 0.f;
    }
    bool specularBounce = false;
    float etaScale = 1.f;
    TravCounters tc{0, 0};
    const int noSpec = BSDF_ALL & ~BSDF_SPECULAR;
    for (int bounces = 0;; ++bounces) {
        HitRec h;
        ray.tMax = kInf;
        const bool found = scene_intersect<false>(S, ray, &h, &tc);
        ++*nClosest;
        ++*nVert;
        Isect it;
        bool haveIsect = false;
        if (found) haveIsect = build_interaction(S, ray, h.prim, h.inst, h.b0, h.b1, h.b2, &it);
        if ((bounces == 0 || specularBounce) && haveIsect && it.light >= 0) {
            const sptx_light &Lt = S.lights[it.light];
            if (Lt.two_sided || dot(it.n, -ray.d) > 0)
                for (int b = 0; b < SPTX_NBINS; ++b) L[b] += beta[b] * spec(S, Lt.spectrum, b);
        }
        if ((bounces == 0 || specularBounce) && !found)  // the infinite lights' Le(ray), path.cpp:96-99
            for (int j = 0; j < S.n_infinite; ++j) {
                const EnvRgb e = env_le(S, S.lights[S.infinite[j]], ray.d);
                for (int b = 0; b < SPTX_NBINS; ++b) L[b] += beta[b] * env_bin(S, e, b);
            }
        if (!haveIsect || bounces >= P.maxDepth) break;
        Bsdf bs;
        bsdf_init(S, it, &bs);
        const V3 wo = it.wo;
        // UniformSampleOneLight, core/integrator.cpp:85-106
        if (bsdf_num_components(S, bs, noSpec) > 0 && S.n_lights > 0) {
            float lightPdfSel;
            const int lightNum = select_light(S, it.p, rng.get1d(), &lightPdfSel);
            if (lightPdfSel != 0) {
                const V2 uLight = rng.get2d();
                const V2 uScattering = rng.get2d();
                const sptx_light &Lt = S.lights[lightNum];
                const bool delta = Lt.type == SPTX_LIGHT_POINT;
                RefPoint ref{it.p, it.pError, it.n};
                float Ld[SPTX_NBINS];
                for (int b = 0; b < SPTX_NBINS; ++b) Ld[b] = 0.f;
                // light sampling (integrator.cpp:119-163)
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
                    bool fBlack = true;
                    for (int b = 0; b < SPTX_NBINS; ++b)
                        if (fsum_bin(S, f, b) * ad != 0.f) fBlack = false;
                    if (!fBlack) {
                        V3 origin = offset_ray_origin(it.p, it.pError, it.n, ls.lp.p - it.p);
                        V3 target = offset_ray_origin(ls.lp.p, ls.lp.pError, ls.lp.n, origin - ls.lp.p);
                        Ray sh{origin, target - origin, 1.f - kShadowEpsilon};
                        ++*nShadow;
                        if (!scene_intersect_p<false>(S, sh, &tc)) {
                            const float weight = delta ? 1.f : power_heuristic(ls.pdf, scatteringPdf);
                            for (int b = 0; b < SPTX_NBINS; ++b) {
                                float c = (fsum_bin(S, f, b) * ad) * li_bin(S, ls, b);
                                if (!delta) c = c * weight;
                                Ld[b] += c / ls.pdf;
                            }
                        }
                    }
                }
                // BSDF sampling (integrator.cpp:166-213)
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
                                Ray mr{offset_ray_origin(it.p, it.pError, it.n, wi2), wi2, kInf};
                                HitRec mh;
                                ++*nClosest;
                                bool useB = false;
                                const bool mFound = scene_intersect<false>(S, mr, &mh, &tc);
                                if (!mFound)
                                    useB = Lt.type == SPTX_LIGHT_INFINITE;  // Li = light.Le(ray)
                                else if (S.prims[mh.prim].light == lightNum) {
                                    if (Lt.two_sided)
                                        useB = true;
                                    else {
                                        Isect li;
                                        if (build_interaction(S, mr, mh.prim, mh.inst, mh.b0, mh.b1, mh.b2, &li))
                                            useB = dot(li.n, -mr.d) > 0;
                                    }
                                }
                                if (useB && Lt.type == SPTX_LIGHT_INFINITE) {
                                    const EnvRgb e = env_le(S, Lt, wi2);
                                    for (int b = 0; b < SPTX_NBINS; ++b)
                                        Ld[b] += (fsum_bin(S, f2, b) * ad2) * env_bin(S, e, b) * weight / sPdf;
                                } else if (useB)
                                    for (int b = 0; b < SPTX_NBINS; ++b)
                                        Ld[b] += (fsum_bin(S, f2, b) * ad2) * spec(S, Lt.spectrum, b) * weight / sPdf;
                            }
                        }
                    }
                }
                for (int b = 0; b < SPTX_NBINS; ++b) L[b] += beta[b] * (Ld[b] / lightPdfSel);
            }
        }
        // next segment (path.cpp:131-150)
        FSum f;
        V3 wi;
        float pdf;
        int flags;
        bsdf_sample_f(S, bs, wo, rng.get2d(), BSDF_ALL, &wi, &pdf, &flags, &f);
        bool fBlack = true;
        if (f.n > 0)
            for (int b = 0; b < SPTX_NBINS; ++b)
                if (fsum_bin(S, f, b) != 0.f) fBlack = false;
        if (fBlack || pdf == 0.f) break;
        const float ad = absdot(wi, it.ns);
        for (int b = 0; b < SPTX_NBINS; ++b) beta[b] = beta[b] * (fsum_bin(S, f, b) * ad / pdf);
        specularBounce = (flags & BSDF_SPECULAR) != 0;
        if (specularBounce && (flags & BSDF_TRANSMISSION)) {
            float eta = bs.eta;
            etaScale *= (dot(wo, it.n) > 0) ? (eta * eta) : 1.f / (eta * eta);
        }
        ray.o = offset_ray_origin(it.p, it.pError, it.n, wi);
        ray.d = wi;
        // Russian roulette (path.cpp:176-184)
        float maxBeta = beta[0] * etaScale;
        for (int b = 1; b < SPTX_NBINS; ++b) {
            float v = beta[b] * etaScale;
            maxBeta = (maxBeta < v) ? v : maxBeta;
        }
        if (maxBeta < P.rrThreshold && bounces > 3) {
            float q = fmaxf(.05f, 1.f - maxBeta);
            if (rng.get1d() < q) break;
            const float inv = 1.f - q;
            for (int b = 0; b < SPTX_NBINS; ++b) beta[b] = beta[b] / inv;
        }
    }
}

// SamplerIntegrator::Render's tile loop (core/integrator.cpp:245-330) for one tile per warp.
template <typename SMP>
SPTX_DEV void sampler_init(SMP &, const SerialParams &, int) {}
SPTX_DEV void sampler_init(ZeroTwoSampler &s, const SerialParams &P, int tile) {
    s.nDims = P.nDims;
    s.spp = P.spp;
    s.t1 = P.tables + (size_t)tile * P.nDims * P.spp * 3;
    s.t2 = reinterpret_cast<float2 *>(s.t1 + (size_t)P.nDims * P.spp);
    s.cur1 = s.cur2 = s.index = 0;
}

template <typename SMP>
__global__ void __launch_bounds__(32)
k_render_serial(const DScene *__restrict__ Sp, DCamera C, DFilm F, SerialParams P, DeviceCounters *cnt,
                int part, int nParts) {
    if (threadIdx.x != 0) return;
    const DScene &S = *Sp;
    const int tile = blockIdx.x;
    if (tile % nParts != part) return;  // tiles are dealt to ranks as sptx_render_tiles does
    const int tx = tile % P.nTilesX, ty = tile / P.nTilesX;
    SMP rng;
    sampler_init(rng, P, tile);
    rng.set_sequence((uint64_t)(ty * P.nTilesX + tx));
    const int tileSize = 16;
    const int x0 = P.sb[0] + tx * tileSize, x1 = min(x0 + tileSize, P.sb[2]);
    const int y0 = P.sb[1] + ty * tileSize, y1 = min(y0 + tileSize, P.sb[3]);
    unsigned long long nCam = 0, nClosest = 0, nShadow = 0, nVert = 0;
    float L[SPTX_SPECTRUM_STRIDE];
    for (int y = y0; y < y1; ++y)
        for (int x = x0; x < x1; ++x) {
            // tileSampler->StartPixel(pixel) comes before the pixel-bounds test (core/integrator.cpp:
            // 262-271): a ZeroTwoSequenceSampler draws its tables for skipped pixels too
            rng.start_pixel();
            if (!(x >= P.pb[0] && x < P.pb[2] && y >= P.pb[1] && y < P.pb[3])) continue;
            for (int s = 0; s < P.spp; ++s) {
                rng.start_sample(s);
                // Sampler::GetCameraSample, core/sampler.cpp:46-52
                const V2 u = rng.get2d();
                const V2 pFilm{(float)x + u.x, (float)y + u.y};
                (void)rng.get1d();  // time
                const V2 pLens = rng.get2d();
                Ray ray{mk(0, 0, 0), mk(0, 0, 1), kInf};
                const float w = camera_ray(C, pFilm, pLens, P.wavelength, &ray);
                ++nCam;
                for (int b = 0; b < SPTX_SPECTRUM_STRIDE; ++b) L[b] = 0.f;
                if (w > 0) li_serial(S, P, rng, ray, L, &nClosest, &nShadow, &nVert);
                // radiance guard (core/integrator.cpp:295-316) and FilmTile::AddSample (film.h:123-163)
                float yy = 0.f;
                bool nan = false;
                for (int b = 0; b < SPTX_NBINS; ++b) {
                    if (isnan(L[b])) nan = true;
                    yy += S.cie_y[b] * L[b];
                }
                const float lum = y_finish(S, yy);
                float g = 1.f;
                if (nan || isinf(lum))
                    g = 0.f;
                else if (lum > F.f.max_sample_luminance)
                    g = F.f.max_sample_luminance / lum;
                int fx0, fy0, fx1, fy1;
                const float2 pf = make_float2(pFilm.x, pFilm.y);
                if (!sample_footprint(F, pf, &fx0, &fy0, &fx1, &fy1)) continue;
                for (int py = fy0; py < fy1; ++py)
                    for (int px = fx0; px < fx1; ++px) {
                        const float fw = filter_weight(F, pf, px, py);
                        float *dst = F.pixels + ((size_t)(py - F.f.cropped[1]) * F.width + (px - F.f.cropped[0])) *
                                                    SPTX_SPECTRUM_STRIDE;
                        for (int b = 0; b < SPTX_NBINS; ++b) {
                            float v = L[b];
                            v = (g == 1.f) ? v : ((g == 0.f) ? 0.f : v * g);
                            atomicAdd(dst + b, v * w * fw);
                        }
                        atomicAdd(dst + SPTX_NBINS, fw);
                    }
            }
        }
    atomicAdd(&cnt->cameraRays, nCam);
    atomicAdd(&cnt->closestRays, nClosest);
    atomicAdd(&cnt->shadowRays, nShadow);
    atomicAdd(&cnt->vertices, nVert);
}

}  // namespace sptx
