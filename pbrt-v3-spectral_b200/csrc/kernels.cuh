// kernels.cuh -- the wavefront kernels (K1 generate, K2 intersect, K3 shade,
// K4 resolve, K5/K6 film) and the stand-alone ray-batch kernels.
#pragma once
#include "wavefront.cuh"

namespace sptx {

constexpr int kBlock = 128;
// resident blocks per SM the traversal kernels are compiled for (register budget = 65536 / (128 * n))
#ifndef SPTX_TRAV_BLOCKS
#define SPTX_TRAV_BLOCKS 6
#endif
// (the stand-alone batches walk deep BVHs and live on memory latency: measured on C4, 2^24 rays,
//  coherent closest / any / incoherent closest / any in Mrays/s -- unbounded (137 registers, 3 blocks)
//  1480 / 2319 / 558 / 817, 6 blocks 1850 / 2301 / 677 / 816, 8 blocks 1886 / 2451 / 761 / 966,
//  10 blocks 1669 / 2411 / 618 / 876)
#ifndef SPTX_BATCH_BLOCKS
#define SPTX_BATCH_BLOCKS 8
#endif
#ifndef SPTX_ANYHIT_BLOCKS
#define SPTX_ANYHIT_BLOCKS 8
#endif

// ---- K1: camera samples -> path state ---------------------------------------
// Sampler::GetCameraSample (core/sampler.cpp:46-52) + GenerateRayDifferential +
// the `if (rayWeight > 0) Li(...)` gate of SamplerIntegrator::Render
// (core/integrator.cpp:279-292).
__global__ void k_generate(DScene S, DCamera C, DSampler H, PathArrays A, RenderParams R,
                           DeviceCounters *cnt) {
    const int64_t nPaths = (int64_t)R.nPix * R.samplesInPass;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nPaths;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int pix = (int)(i % R.nPix);
        const int sampleNum = R.sampleList ? R.sampleList[i] : R.sampleBegin + (int)(i / R.nPix);
        const int2 pxy = R.pixels[pix];
        SampleStream st;
        st.index = sample_index(H, pxy.x, pxy.y, sampleNum);
        st.dim = 0;
        V2 u = get2d(H, st);
        V2 pFilm{(float)pxy.x + u.x, (float)pxy.y + u.y};
        (void)get1d(H, st);  // time
        V2 pLens = get2d(H, st);
        Ray ray;
        ray.o = mk(0, 0, 0);
        ray.d = mk(0, 0, 1);
        ray.tMax = kInf;
        float w = camera_ray(C, pFilm, pLens, R.wavelength, &ray);
        A.rayA[i] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.d.x);
        A.rayB[i] = make_float4(ray.d.y, ray.d.z, 1.f, R.ignoreRayWeight ? 1.f : w);
        A.sindex[i] = st.index;
        // SpectralPathIntegrator (integrators/spectralpath.cpp:262-283): the bands of a sample share
        // its CameraSample and draw from ONE sampler stream, band after band -- a band starts at
        // the dimension where the previous band's path stopped (left in `info` by K3)
        const int startDim = R.band > 0 ? (int)(A.info[i] >> 16) : st.dim;
        A.info[i] = pack_info(0, 0, startDim);
        A.pfilm[i] = make_float2(pFilm.x, pFilm.y);
        A.spill[i] = 0;
        // path-major spectra: 32 floats (one 128-byte line) per path.  (beta = 1 at the camera vertex is
        // not stored: RenderParams::betaOnes tells K3 / K4b of bounce 0)
        float4 *l4 = reinterpret_cast<float4 *>(A.L + (size_t)i * SPTX_SPECTRUM_STRIDE);
        for (int q = 0; q < SPTX_SPECTRUM_STRIDE / 4; ++q) l4[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        queue_push(A.q[0], A.counters + 0, (int)i, w > 0);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->cameraRays, (unsigned long long)nPaths);
}

// ---- K2: closest hit for the active queue ------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, SPTX_TRAV_BLOCKS)
k_intersect(DScene S, PathArrays A, int qi, DeviceCounters *cnt) {
    const int n = A.counters[qi];
    const int *queue = A.q[qi];
    const int lane = threadIdx.x & 31;
    TravCounters tc{0, 0};
    for (int k0 = blockIdx.x * blockDim.x; k0 < n; k0 += gridDim.x * blockDim.x) {
        const int k = k0 + threadIdx.x;
        const bool valid = k < n;
        int i = 0, cls = 0;
        if (valid) {
            i = queue[k];
            float4 ra = A.rayA[i], rb = A.rayB[i];
            Ray ray{mk(ra.x, ra.y, ra.z), mk(ra.w, rb.x, rb.y), kInf};
            HitRec h;
            scene_intersect<COUNT>(S, ray, &h, &tc);
            A.hitA[i] = make_float4(__int_as_float(h.prim), __int_as_float(h.inst), h.b0, h.b1);
            A.hitB[i] = make_float4(h.b2, ray.tMax, 0.f, 0.f);
            if (h.prim >= 0) cls = S.prim_class[h.prim];
        }
        // append to the queue of the hit's shading class: one atomic per class per warp
        const unsigned vmask = __ballot_sync(0xffffffffu, valid);
        if (valid) {
            const unsigned peers = __match_any_sync(vmask, cls);
            const int leader = __ffs(peers) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(A.counters + kClassCounter0 + cls, __popc(peers));
            base = __shfl_sync(peers, base, leader);
            A.qc[(size_t)cls * A.capacity + base + __popc(peers & ((1u << lane) - 1u))] = i;
        }
    }
    if (COUNT) {
        atomicAdd(&cnt->nodes[0], (unsigned long long)tc.nodes);
        atomicAdd(&cnt->tests[0], (unsigned long long)tc.tests);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->closestRays, (unsigned long long)n);
}

// ---- stand-alone ray batches (sptx_intersect_batch) ---------------------------
template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(kBlock, SPTX_BATCH_BLOCKS)
k_ray_batch(DScene S, const sptx_ray *rays, const int *order, int64_t n, sptx_hit *hits) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x) {
        // `order` (or NULL): the walk order of a binned batch; results land at the ray's own index
        const int64_t i = order ? (int64_t)order[k] : k;
        const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
        float4 a = __ldg(rp), b = __ldg(rp + 1);
        Ray ray{mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), a.w};
        TravCounters tc{0, 0};
        sptx_hit out;
        out.prim = -1;
        out.instance = -1;
        out.b0 = out.b1 = out.b2 = 0.f;
        if (ANY) {
            bool hit = scene_intersect_p<COUNT>(S, ray, &tc);
            out.prim = hit ? 0 : -1;
            out.t = ray.tMax;
        } else {
            HitRec h;
            scene_intersect<COUNT>(S, ray, &h, &tc);
            // (a BVH built on the device orders the primitives its own way: callers see their ids)
            out.prim = (h.prim >= 0 && S.prim_ids) ? S.prim_ids[h.prim] : h.prim;
            out.instance = h.inst;
            out.b0 = h.b0;
            out.b1 = h.b1;
            out.b2 = h.b2;
            out.t = ray.tMax;
        }
        out.n_nodes = tc.nodes;
        out.n_tests = tc.tests;
        float4 *hp = reinterpret_cast<float4 *>(hits + i);
        hp[0] = make_float4(__int_as_float(out.prim), __int_as_float(out.instance), out.t, out.b0);
        hp[1] = make_float4(out.b1, out.b2, __int_as_float(out.n_nodes), __int_as_float(out.n_tests));
    }
}

}  // namespace sptx
