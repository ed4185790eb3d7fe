// shade.cuh -- K3: one path vertex of PathIntegrator::Li.
//
// integrators/path.cpp:88-184 for the vertex, core/integrator.cpp:85-215
// (UniformSampleOneLight, EstimateDirect) for next-event estimation.
//
// One thread per path.  K2 hands the paths over sorted by shading class (one class per
// material lobe signature), and K3 is launched once per class with the kernel VARIANT compiled
// for that class's lobe types: a matte vertex runs code that contains no specular, microfacet or
// Fresnel branch at all, a glass / mirror vertex code without next-event estimation.  A block
// walks its class queue in chunks; the scattered 76 bytes of path state of the NEXT chunk are
// copied into shared memory (cp.async, 40 KB) while the current one is shaded.  The scene and
// sampler descriptors are __grid_constant__ kernel parameters (table pointers come from the
// parameter bank).  Spectra are path-major ([P][32] floats): a thread moves its 128-byte rows as float4.
// The two rays EstimateDirect traces (shadow ray, BSDF-sampled MIS ray) are
// written out with their per-bin contributions (fa, fb) and traced by K4.
#pragma once
#include "tex.cuh"
#include "wavefront.cuh"

namespace sptx {

struct Row4 {
    float4 *p;
    SPTX_DEV float4 ld(int c) const { return p[c]; }
    // rows this kernel only reads (the incoming beta): the read-only path tells the compiler that
    // the stores to the other rows cannot alias them, so the eight loads issue together
    SPTX_DEV float4 ldro(int c) const { return __ldg(p + c); }
    SPTX_DEV void st(int c, float4 v) const { p[c] = v; }
};
SPTX_DEV Row4 row4(float *base, int i) {
    return Row4{reinterpret_cast<float4 *>(base + (size_t)i * SPTX_SPECTRUM_STRIDE)};
}
// bins of chunk c that exist (the last chunk has the pad word)
SPTX_DEV bool any_nonzero4(float4 v, int c) {
    return v.x != 0.f || v.y != 0.f || v.z != 0.f || (c < 7 && v.w != 0.f);
}

// 128-byte line into L2 without tying up registers (the row is consumed a few thousand
// instructions later, by which time it would otherwise be a DRAM round trip).
SPTX_DEV void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Asynchronous global -> shared copies of one thread's own slots (LDGSTS).  The copies of a thread
// become visible to THAT thread after its wait; no other thread reads the slots, so no barrier.
SPTX_DEV void cp_async16(void *smem, const void *g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g));
}
SPTX_DEV void cp_async8(void *smem, const void *g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g));
}
SPTX_DEV void cp_async4(void *smem, const void *g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g));
}
SPTX_DEV void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// The vertex is cut into stages.  In round 1 a block barrier between stages kept the SM's warps
// inside the same few KB of code (the kernel was ~200 KB of SASS, free-running warps had a 53 %
// instruction-cache hit rate -- DESIGN.md 5); the per-class variants of round 2 are a third of that
// and run without barriers (SPTX_STAGE_MASK below).  One 512-thread block per SM (128 registers).
#ifndef SPTX_SHADE_BLOCK
#define SPTX_SHADE_BLOCK 512
#endif
#ifndef SPTX_SHADE_BLOCKS_PER_SM
#define SPTX_SHADE_BLOCKS_PER_SM 1
#endif
constexpr int kShadeBlock = SPTX_SHADE_BLOCK;
constexpr int kShadeBlocksPerSM = SPTX_SHADE_BLOCKS_PER_SM;
#ifdef SPTX_STAGE_CLOCKS  // diagnosis build: per-stage warp clocks, printed by k_film
__device__ unsigned long long g_stageClocks[32];
#define SPTX_TICK(slot)                 \
    do {                                \
        long long now_ = clock64();     \
        stageClk[slot] += now_ - tPrev; \
        tPrev = now_;                   \
    } while (0)
#else
#define SPTX_TICK(slot)
#endif
// Which stage ends carry the barrier (bit s = after slot s).  Measured on one 4-spp pass of the
// bench scene (shade ms): none 11.97, 0x01 9.39, 0x04 9.24, 0x10 9.13, 0x40 8.85, 0x100 9.31,
// 0x11 9.56 (the previous setting), 0x44 9.21, 0x50 8.99, 0x140 9.40, 0x41 9.46 -- one barrier,
// after the f / pdf / fa-row stage, was best in round 1 (C1 -2.5 %, C3 -7.7 % with it as well).
// End of round 2 (per-class variants, state staged through shared memory, descriptors in the
// parameter bank, out-of-line microfacet helpers -- the code a warp walks per vertex is a third of
// what it was): shade ms per bench step with mask 0 / 0x04 / 0x10 / 0x40 = 78.3 / 81.0 / 80.2 / 79.7.
// No barrier.
#ifndef SPTX_STAGE_MASK
#define SPTX_STAGE_MASK 0
#endif
#define SPTX_STAGE(slot)                                       \
    do {                                                       \
        SPTX_TICK(slot);                                       \
        if ((SPTX_STAGE_MASK >> (slot)) & 1) __syncthreads();  \
        SPTX_TICK(slot + 1);                                   \
    } while (0)

// 1: the kernels handle Light "infinite" (radiance map lookups in next-event estimation and MIS);
// 0: compiled for scenes without one -- the matte variant is 14 % shorter and the bench scene's K3
// 6 % faster (98.7 -> 93 ms per step) without that code and its live state.
#ifndef SPTX_SHADE_ENV
#define SPTX_SHADE_ENV 1
#endif
constexpr bool kShadeEnv = SPTX_SHADE_ENV != 0;

#ifdef SPTX_SHADE_NS
namespace SPTX_SHADE_NS {  // (one copy of the kernels per arithmetic mode and ENV setting, see shade.cu)
#endif

// Terms of a textured lobe (R = SPTX_TEX_ROW(..)) get the texture's value at this hit.
template <int MAXT>
SPTX_DEV void tex_patch(FSumN<MAXT> *f, const EnvRgb &tv) {
#pragma unroll
    for (int j = 0; j < MAXT; ++j)
        if (j < f->n && f->t[j].R <= -2) {
            FTerm &t = f->t[j];
            t.kind = TERM_TEX_S;
            t.e0 = tv.row[0];
            t.e1 = tv.row[1];
            t.e2 = tv.row[2];
            t.s1 = tv.w[0];
            t.s2 = tv.w[1];
            t.s3 = tv.w[2];
        }
}
// The camera ray's offset rays, as SamplerIntegrator::Render hands them to Li: GenerateRayDifferential
// of the path's CameraSample, then ScaleDifferentials(1 / sqrt(spp)) (core/integrator.cpp:279-283).
// Recomputed here (camera-ray vertices of textured surfaces only) instead of carried in the path state.
SPTX_NOINL void camera_differentials(const DScene &S, const DSampler &H, uint64_t sindex, V2 pFilm,
                                     float wavelength, const Ray &ray, RayDiff *rd) {
    SampleStream cs{sindex, 3};  // CameraSample: pFilm (dims 0, 1), time (2), pLens (3, 4)
    const V2 pLens = get2d(H, cs);
    Ray again;
    camera_ray(*S.camera, pFilm, pLens, wavelength, &again, rd);
    scale_differentials(ray, 1.f / sqrtf((float)H.s.spp), rd);
}

template <int LOBES, int MAXT>
__global__ void __launch_bounds__(kShadeBlock, kShadeBlocksPerSM)
k_shade(const __grid_constant__ DScene S, const __grid_constant__ DSampler H, PathArrays A, RenderParams R,
        int qi, int betaIn, int cls, DeviceCounters *cnt) {
    // (S and H are read from the kernel's parameter bank: a table pointer or a sampler constant costs
    //  no load.  __grid_constant__ lets the out-of-line helpers take their address without a local copy.)
    const int n = A.counters[kClassCounter0 + cls];
    const int *queue = A.qc + (size_t)cls * A.capacity;
    const int noSpec = BSDF_ALL & ~BSDF_SPECULAR;
#ifdef SPTX_STAGE_CLOCKS
    long long stageClk[24] = {0};
    long long tPrev = clock64();
#endif
    // The path state of a chunk (ray, hit, info, sample index: 76 B per path, a scattered gather
    // behind the queue entry) is copied into shared memory while the block still shades the chunk
    // before it: after the stage barrier all warps start a chunk together, and they all used to
    // wait for these loads together.
    __shared__ float4 sState[5][kShadeBlock];
    auto stage_state = [&](int path) {
        const int t = (int)threadIdx.x;
        cp_async16(&sState[0][t], A.rayA + path);
        cp_async16(&sState[1][t], A.rayB + path);
        cp_async16(&sState[2][t], A.hitA + path);
        cp_async16(&sState[3][t], A.hitB + path);
        cp_async8(&sState[4][t], A.sindex + path);
        cp_async4(reinterpret_cast<char *>(&sState[4][t]) + 8, A.info + path);
    };
    int iCur = -1;
    {
        const int k0 = (int)blockIdx.x * kShadeBlock + (int)threadIdx.x;
        if (k0 < n) {
            iCur = queue[k0];
            stage_state(iCur);
        }
    }
    // (one block-uniform induction variable: the chunk's first queue entry)
    for (int c0 = (int)blockIdx.x * kShadeBlock; c0 < n; c0 += (int)gridDim.x * kShadeBlock) {
        const int k = c0 + (int)threadIdx.x;
        bool alive = k < n;
        const bool valid = alive;
        int i = 0, iNext = -1;
        int emitRow = -1;
        SampleStream st{0, 0};
        int bounces = 0, neeFlags = 0, lightNum = 0;
        float lightPdfSel = 0.f, etaScale = 1.f, rayWeight = 0.f;
        constexpr bool kTex = (LOBES & kTexturedLobes) != 0;
        // Variants whose terms carry per-bin Fresnel / microfacet code keep their row loops rolled: unrolled
        // eight times (two terms, every kind, four lanes) the plastic / metal variant was 25 176
        // instructions (393 KB) with an instruction-cache hit rate of 56 % and `no_instruction` its top stall.
        constexpr bool kHeavyTerms = (LOBES & (kMicrofacetLobes | (1 << SPTX_LOBE_OREN_NAYAR))) != 0;
        constexpr int kRowUnroll = kHeavyTerms ? 1 : 8;
        Isect it;
        SurfParam par;
        Bsdf bs;
        FSumN<MAXT> f;
        Ray ray{mk(0, 0, 0), mk(0, 0, 1), kInf};
        // ---- stage 0: interaction, emission (path.cpp:91-112) ----
        if (alive) {
            i = iCur;
            // (first: the wait is a barrier on the scoreboard the copies share with whatever else is in
            //  flight on it, e.g. the L2 prefetch below)
            SPTX_TICK(12);
            cp_async_wait_all();
            SPTX_TICK(13);
            const int t = (int)threadIdx.x;
            const float4 ra = sState[0][t], rb = sState[1][t];
            const float4 ha = sState[2][t], hb = sState[3][t];
            const float4 sx = sState[4][t];
            if (!R.betaOnes) prefetch_l2(A.beta[betaIn] + (size_t)i * SPTX_SPECTRUM_STRIDE);
            {   // the next chunk of this block
                const int kn = k + (int)gridDim.x * kShadeBlock;
                if (kn < n) iNext = queue[kn];
            }
            ray.o = mk(ra.x, ra.y, ra.z);
            ray.d = mk(ra.w, rb.x, rb.y);
            etaScale = rb.z;
            rayWeight = rb.w;
            const uint32_t info = __float_as_uint(sx.z);
            bounces = (int)(info & 0xffu);
            const bool specularBounce = ((info >> 8) & 1u) != 0;
            st.index = (uint64_t)__float_as_uint(sx.x) | ((uint64_t)__float_as_uint(sx.y) << 32);
            st.dim = (int)(info >> 16);
            const int prim = __float_as_int(ha.x), inst = __float_as_int(ha.y);
#ifdef SPTX_TRACE_PATH  // diagnosis build: the vertices of path 0 of a sptx_debug_radiance call
            if (R.sampleList && i == 0)
                printf("TRACE b%d ray o %.9g %.9g %.9g d %.9g %.9g %.9g | prim %d inst %d t %.9g dim %d\n", bounces,
                       ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, prim, inst, hb.y, st.dim);
#endif
            bool haveIsect = false;
            if (prim >= 0) {
                haveIsect = build_interaction(S, ray, prim, inst, ha.z, ha.w, hb.x, &it, kTex ? &par : nullptr);
                if (!haveIsect) atomicOr(A.counters + 3, ERR_INTERACTION);
            }
            // isect.Le(-ray.d) -> DiffuseAreaLight::L(intr, w)
            if ((bounces == 0 || specularBounce) && haveIsect && it.light >= 0) {
                const sptx_light &Lt = S.lights[it.light];
                if (Lt.two_sided || dot(it.n, -ray.d) > 0) {
                    // (few lanes of a warp get here, and the two rows are DRAM round trips the whole
                    //  warp would wait for: the rows are asked for now and added at the end of the vertex)
                    emitRow = Lt.spectrum;
                    prefetch_l2(A.L + (size_t)i * SPTX_SPECTRUM_STRIDE);
                }
            }
            alive = haveIsect && bounces < R.maxDepth;
        }
        SPTX_STAGE(0);
        // ---- stage 1: BSDF frame, light choice (integrator.cpp:85-106) ----
        bool nee = false;
        V2 uLight{0, 0}, uScattering{0, 0};
        EnvRgb texv;
        if (alive) {
            bsdf_init(S, it, &bs);
            if (kTex && bs.n > 0) {
                // Kd->Evaluate(*si) of MatteMaterial / PlasticMaterial::ComputeScatteringFunctions
                // (materials/matte.cpp:49-57, plastic.cpp:49-56): the diffuse lobe is the material's first,
                // and it is there only when the texture's value is not black
                const int R0 = S.lobes[bs.first].R;
                if (R0 <= -2) {
                    RayDiff rd;
                    if (bounces == 0) {
                        const float2 pf = A.pfilm[i];
                        camera_differentials(S, H, st.index, V2{pf.x, pf.y}, R.wavelength, ray, &rd);
                    }
                    texv = tex_evaluate(S, -2 - R0, it, par, bounces == 0 ? &rd : nullptr);
#ifdef SPTX_TRACE_PATH
                    if (R.sampleList && bounces == 0) {
                        V2 ddx, ddy;
                        uv_differentials(it, par, &rd, &ddx, &ddy);
                        printf("DEVTRACE i %d p %.9g %.9g %.9g n %.9g %.9g %.9g uv %.9g %.9g\n", i, it.p.x, it.p.y, it.p.z,
                               it.n.x, it.n.y, it.n.z, par.uv.x, par.uv.y);
                        printf("DEVTRACE   rxo %.9g %.9g %.9g rxd %.9g %.9g %.9g ryo %.9g %.9g %.9g ryd %.9g %.9g %.9g\n",
                               rd.rxo.x, rd.rxo.y, rd.rxo.z, rd.rxd.x, rd.rxd.y, rd.rxd.z, rd.ryo.x, rd.ryo.y, rd.ryo.z,
                               rd.ryd.x, rd.ryd.y, rd.ryd.z);
                        printf("DEVTRACE   dpdu %.9g %.9g %.9g dpdv %.9g %.9g %.9g dudx %.9g dvdx %.9g dudy %.9g dvdy %.9g\n",
                               par.dpdu.x, par.dpdu.y, par.dpdu.z, par.dpdv.x, par.dpdv.y, par.dpdv.z, ddx.x, ddx.y, ddy.x, ddy.y);
                        const float4 a = tex_bin4(S, texv.row[0], texv.row[1], texv.row[2], texv.w[0], texv.w[1], texv.w[2], 0);
                        const float4 b = tex_bin4(S, texv.row[0], texv.row[1], texv.row[2], texv.w[0], texv.w[1], texv.w[2], 2);
                        const float4 c = tex_bin4(S, texv.row[0], texv.row[1], texv.row[2], texv.w[0], texv.w[1], texv.w[2], 4);
                        const float4 d = tex_bin4(S, texv.row[0], texv.row[1], texv.row[2], texv.w[0], texv.w[1], texv.w[2], 6);
                        printf("DEVTRACE   Kd %.9g %.9g %.9g %.9g\n", a.z, b.z, c.z, d.z);
                    }
#endif
                    if (tex_is_black(S, texv)) {
                        bs.first += 1;
                        bs.n -= 1;
                    }
                }
            }
            if (bsdf_num_components<LOBES>(S, bs, noSpec) > 0 && S.n_lights > 0) {
                // the five values next-event estimation consumes (Get1D, Get2D, Get2D) are pure
                // functions of (index, dimension): evaluated together, consumed as the reference does
                float u5[5];
                halton_n<5>(H, st.index, st.dim, u5);
                st.dim += 1;
                lightNum = select_light(S, it.p, u5[0], &lightPdfSel);
                if (lightPdfSel != 0) {
                    uLight = V2{u5[1], u5[2]};
                    uScattering = V2{u5[3], u5[4]};
                    st.dim += 4;
                    nee = true;
                }
            }
        }
        if (iNext >= 0) stage_state(iNext);  // (this thread's slots were read in stage 0)
        iCur = iNext;
        SPTX_STAGE(2);
        // ---- stage 2: light sampling strategy (integrator.cpp:119-163) ----
        bool delta = false;
        V3 shO = mk(0, 0, 0), shD = mk(0, 0, 0), misO = mk(0, 0, 0), misD = mk(0, 0, 0);
        LiSample ls;
        ls.pdf = 0.f;
        ls.black = true;
        if (nee) {
            const sptx_light &Lt = S.lights[lightNum];
            delta = Lt.type == SPTX_LIGHT_POINT;
            RefPoint ref{it.p, it.pError, it.n};
            ls = light_sample_li<kShadeEnv>(S, Lt, ref, uLight);
        }
        SPTX_STAGE(4);
        if (nee && ls.pdf > 0 && !ls.black) {
            bsdf_f<LOBES>(S, bs, it.wo, ls.wi, noSpec, &f);
            if (kTex) tex_patch(&f, texv);
            const float ad = absdot(ls.wi, it.ns);
            const float scatteringPdf = bsdf_pdf<LOBES>(S, bs, it.wo, ls.wi, noSpec);
            const float weight = delta ? 1.f : power_heuristic(ls.pdf, scatteringPdf);
            const Row4 fa = row4(A.fa, i);
            bool fNon = false, liNon = false;
            auto faChunk = [&](int c) {
                float4 fv = (kHeavyTerms ? fsum4_call<LOBES, MAXT>(S, f, c) : fsum4<LOBES>(S, f, c)), li = li4<kShadeEnv>(S, ls, c);
                fv.x *= ad;
                fv.y *= ad;
                fv.z *= ad;
                fv.w *= ad;
                float4 o = make_float4(fv.x * li.x, fv.y * li.y, fv.z * li.z, fv.w * li.w);
                if (!delta) {
                    o.x = o.x * weight;
                    o.y = o.y * weight;
                    o.z = o.z * weight;
                    o.w = o.w * weight;
                }
                fa.st(c, make_float4(o.x / ls.pdf, o.y / ls.pdf, o.z / ls.pdf, padless(c, o.w, ls.pdf)));
                fNon |= any_nonzero4(fv, c);
                liNon |= any_nonzero4(li, c);
            };
            if (kHeavyTerms) {
#pragma unroll 1
                for (int c = 0; c < 8; ++c) faChunk(c);
            } else {
                for (int c = 0; c < 8; ++c) faChunk(c);
            }
            if (fNon && liNon) {
                // VisibilityTester(ref, pShape) -> SpawnRayTo(Interaction)
                shO = offset_ray_origin(it.p, it.pError, it.n, ls.lp.p - it.p);
                V3 target = offset_ray_origin(ls.lp.p, ls.lp.pError, ls.lp.n, shO - ls.lp.p);
                shD = target - shO;
                neeFlags |= NEE_HAS_SHADOW;
            }
        }
        SPTX_STAGE(6);
        // ---- stage 3: BSDF sampling strategy (integrator.cpp:166-213) ----
        if (nee && !delta) {
            float sPdf;
            int sType;
            bsdf_sample_f<LOBES>(S, bs, it.wo, uScattering, noSpec, &misD, &sPdf, &sType, &f);
            if (kTex) tex_patch(&f, texv);
            if (f.n > 0 && sPdf > 0) {
                float weight = 1.f;
                bool skip = false;
                if (!(sType & BSDF_SPECULAR)) {
                    RefPoint ref{it.p, it.pError, it.n};
                    float lightPdf = light_pdf_li<kShadeEnv>(S, S.lights[lightNum], ref, misD);
                    if (lightPdf == 0)
                        skip = true;
                    else
                        weight = power_heuristic(sPdf, lightPdf);
                }
                if (!skip) {
                    const float ad = absdot(misD, it.ns);
                    const sptx_light &Lm = S.lights[lightNum];
                    const int leRow = Lm.spectrum;
                    // an infinite light: Li = Le(ray) if the sampled ray leaves the scene (K4a decides)
                    const bool envLight = kShadeEnv && Lm.type == SPTX_LIGHT_INFINITE;
                    EnvRgb envLe;
                    if (envLight) envLe = env_le(S, Lm, misD);
                    const Row4 fb = row4(A.fb, i);
                    bool fNon = false;
                    auto fbChunk = [&](int c) {
                        float4 fv = (kHeavyTerms ? fsum4_call<LOBES, MAXT>(S, f, c) : fsum4<LOBES>(S, f, c));
                        const float4 le = envLight ? env_bin4(S, envLe, c) : spec4(S, leRow, c);
                        fv.x *= ad;
                        fv.y *= ad;
                        fv.z *= ad;
                        fv.w *= ad;
                        fb.st(c, make_float4(fv.x * le.x * weight / sPdf, fv.y * le.y * weight / sPdf,
                                             fv.z * le.z * weight / sPdf,
                                             padless(c, fv.w * le.w * weight, sPdf)));
                        fNon |= any_nonzero4(fv, c);
                    };
                    if (kHeavyTerms) {
#pragma unroll 1
                        for (int c = 0; c < 8; ++c) fbChunk(c);
                    } else {
                        for (int c = 0; c < 8; ++c) fbChunk(c);
                    }
                    if (fNon) {
                        misO = offset_ray_origin(it.p, it.pError, it.n, misD);
                        neeFlags |= NEE_HAS_MIS;
                    }
                }
            }
        }
        if (neeFlags) {
            A.shA[i] = make_float4(shO.x, shO.y, shO.z, shD.x);
            A.shB[i] = make_float4(shD.y, shD.z, 1.f - kShadowEpsilon, lightPdfSel);
            A.misA[i] = make_float4(misO.x, misO.y, misO.z, misD.x);
            A.misB[i] = make_float4(misD.y, misD.z, __int_as_float(lightNum), __int_as_float(neeFlags));
        }
        SPTX_STAGE(8);
        // The segment after the last scattering event (bounces + 1 == maxDepth) is traced by the reference
        // only to add the emission it finds, and that is added after a specular bounce only (path.cpp:86-99,
        // then the break at :102): after any other event the ray cannot change L, and the path ends here.
        // A variant without specular lobes then has no use for the BSDF sample either -- unless the sample's
        // next band continues the sampler stream where this path leaves it (nBands > 1).
#ifndef SPTX_TRACE_DEAD_SEGMENTS
        const bool lastScatter = bounces + 1 >= R.maxDepth;
#else
        const bool lastScatter = false;
#endif
        if (lastScatter && (LOBES & kSpecularLobes) == 0 && R.nBands == 1) alive = false;
        // ---- stage 4: BSDF sample for the next path segment (path.cpp:131-150) ----
        if (alive) {
            V3 wi;
            float pdf;
            int flags;
            bsdf_sample_f<LOBES>(S, bs, it.wo, get2d(H, st), BSDF_ALL, &wi, &pdf, &flags, &f);
            if (kTex) tex_patch(&f, texv);
            bool fNon = false;
            float maxBeta = 0.f;
            const Row4 bIn = row4(A.beta[betaIn], i), bOut = row4(A.beta[betaIn ^ 1], i);
            if (f.n > 0 && pdf != 0.f) {
                if ((flags & BSDF_SPECULAR) && (flags & BSDF_TRANSMISSION)) {
                    float eta = bs.eta;
                    etaScale *= (dot(it.wo, it.n) > 0) ? (eta * eta) : 1.f / (eta * eta);
                }
                const float ad = absdot(wi, it.ns);
                float4 bRow[kHeavyTerms ? 1 : 8];
                if (!kHeavyTerms) {
#pragma unroll
                    for (int c = 0; c < 8; ++c) bRow[c] = R.betaOnes ? make_float4(1.f, 1.f, 1.f, 1.f) : bIn.ldro(c);
                }
#pragma unroll(kRowUnroll)
                for (int c = 0; c < 8; ++c) {
                    const float4 fv = (kHeavyTerms ? fsum4_call<LOBES, MAXT>(S, f, c) : fsum4<LOBES>(S, f, c));
                    const float4 b = kHeavyTerms ? (R.betaOnes ? make_float4(1.f, 1.f, 1.f, 1.f) : bIn.ldro(c)) : bRow[kHeavyTerms ? 0 : c];
                    const float4 nb = make_float4(b.x * (fv.x * ad / pdf), b.y * (fv.y * ad / pdf),
                                                  b.z * (fv.z * ad / pdf), b.w * padless(c, fv.w * ad, pdf));
                    bOut.st(c, nb);
                    fNon |= any_nonzero4(fv, c);
                    // MaxComponentValue(beta * etaScale), std::max order (path.cpp:176-177)
                    const float v0 = nb.x * etaScale, v1 = nb.y * etaScale, v2 = nb.z * etaScale,
                                v3 = nb.w * etaScale;
                    maxBeta = (c == 0) ? v0 : ((maxBeta < v0) ? v0 : maxBeta);
                    maxBeta = (maxBeta < v1) ? v1 : maxBeta;
                    maxBeta = (maxBeta < v2) ? v2 : maxBeta;
                    if (c < 7) maxBeta = (maxBeta < v3) ? v3 : maxBeta;
                }
            }
            if (!fNon || pdf == 0.f)
                alive = false;
            else {
                // Russian roulette (path.cpp:176-184)
                if (maxBeta < R.rrThreshold && bounces > 3) {
                    float q = fmaxf(.05f, 1.f - maxBeta);
                    if (get1d(H, st) < q)
                        alive = false;
                    else {
                        const float inv = 1.f - q;
                        for (int c = 0; c < 8; ++c) {
                            float4 b = bOut.ld(c);
                            bOut.st(c, make_float4(b.x / inv, b.y / inv, b.z / inv, padless(c, b.w, inv)));
                        }
                    }
                }
                if (lastScatter && !(flags & BSDF_SPECULAR)) alive = false;
#ifdef SPTX_TRACE_PATH
                if (R.sampleList && i == 0)
                    printf("TRACE b%d sampled flags %d pdf %.9g wi %.9g %.9g %.9g alive %d p %.9g %.9g %.9g n %.9g %.9g %.9g\n",
                           bounces, flags, pdf, wi.x, wi.y, wi.z, (int)alive, it.p.x, it.p.y, it.p.z, it.n.x,
                           it.n.y, it.n.z);
#endif
                if (alive) {
                    V3 o = offset_ray_origin(it.p, it.pError, it.n, wi);
                    A.rayA[i] = make_float4(o.x, o.y, o.z, wi.x);
                    A.rayB[i] = make_float4(wi.y, wi.z, etaScale, rayWeight);
                    if (st.dim >= H.s.n_dims) atomicOr(A.counters + 3, ERR_DIM_OVERFLOW);
                    A.info[i] = pack_info(bounces + 1, (flags & BSDF_SPECULAR) ? 1 : 0, st.dim);
                }
            }
        }
        if (emitRow >= 0) {
            // L += beta * isect.Le(-ray.d) (path.cpp:91-95; beta is the vertex's incoming throughput)
            const Row4 Lr = row4(A.L, i), br = row4(A.beta[betaIn], i);
            for (int c = 0; c < 8; ++c) {
                float4 l = Lr.ld(c);
                const float4 b = R.betaOnes ? make_float4(1.f, 1.f, 1.f, 1.f) : br.ldro(c), e = spec4(S, emitRow, c);
                l.x += b.x * e.x;
                l.y += b.y * e.y;
                l.z += b.z * e.z;
                l.w += b.w * e.w;
                Lr.st(c, l);
            }
        }
        // the next band of the sample continues the sampler stream where this path stopped
        if (R.nBands > 1 && valid && !alive) A.info[i] = pack_info(bounces, 0, st.dim);
        SPTX_TICK(10);
        queue_push(A.qNee, A.counters + 2, i, neeFlags != 0);
        queue_push(A.qMis, A.counters + 4, i, (neeFlags & NEE_HAS_MIS) != 0);
        queue_push(A.q[qi ^ 1], A.counters + (qi ^ 1), i, alive);
        SPTX_TICK(11);
    }
#ifdef SPTX_STAGE_CLOCKS
    if ((threadIdx.x & 31) == 0)
        for (int c = 0; c < 24; ++c) atomicAdd(&g_stageClocks[c], (unsigned long long)stageClk[c]);
#endif
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->vertices, (unsigned long long)n);
}

// Class 0: paths whose ray left the scene.  PathIntegrator::Li adds the infinite lights' emission
// (camera rays and rays leaving a specular bounce) and ends them; with SpectralPathIntegrator bands
// the next band still needs the sampler dimension the path stopped at.
__global__ void k_shade_miss(const __grid_constant__ DScene S, PathArrays A, RenderParams R, int betaIn,
                             DeviceCounters *cnt) {
    const int n = A.counters[kClassCounter0];
    if (R.nBands > 1 || (kShadeEnv && S.n_infinite > 0))
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
            const int i = A.qc[k];
            const uint32_t info = A.info[i];
            const int bounces = (int)(info & 0xffu);
            const bool specularBounce = ((info >> 8) & 1u) != 0;
            if (kShadeEnv && S.n_infinite > 0 && (bounces == 0 || specularBounce)) {
                // L += beta * light->Le(ray) over scene.infiniteLights (integrators/path.cpp:96-99)
                const float4 ra = A.rayA[i], rb = A.rayB[i];
                const V3 d = mk(ra.w, rb.x, rb.y);
                const Row4 Lr = row4(A.L, i), br = row4(A.beta[betaIn], i);
                for (int j = 0; j < S.n_infinite; ++j) {
                    const EnvRgb e = env_le(S, S.lights[S.infinite[j]], d);
                    for (int c = 0; c < 8; ++c) {
                        float4 l = Lr.ld(c);
                        const float4 b = R.betaOnes ? make_float4(1.f, 1.f, 1.f, 1.f) : br.ldro(c), le = env_bin4(S, e, c);
                        l.x += b.x * le.x;
                        l.y += b.y * le.y;
                        l.z += b.z * le.z;
                        l.w += b.w * le.w;
                        Lr.st(c, l);
                    }
                }
            }
            if (R.nBands > 1) A.info[i] = pack_info(bounces, 0, (int)(info >> 16));
        }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->vertices, (unsigned long long)n);
}

#ifdef SPTX_SHADE_NS
}  // namespace SPTX_SHADE_NS
#endif

}  // namespace sptx
