// shade.cu -- the translation unit of K3 (k_shade).
//
// K3 is compiled on its own so that its arithmetic mode can differ from the rest of the
// library: traversal, the sampler and the camera (sptx_api.cu) must round like the
// reference's x86-64 build (hit ids, t, Halton values and camera rays are bit-exact), while
// per-sample radiance only has to meet north_star's tolerance (see the Makefile: SHADE_FLAGS).
#include <algorithm>

#include "shade.cuh"

namespace sptx {

#ifdef SPTX_STAGE_CLOCKS
__global__ void k_stage_clocks_print() {
    unsigned long long tot = 0;
    for (int c = 0; c < 24; ++c) tot += g_stageClocks[c];
    for (int c = 0; c < 17; ++c) {
        printf("S%02d %5.1f%%  ", c, 100.0 * (double)g_stageClocks[c] / (double)tot);
        g_stageClocks[c] = 0;
    }
    printf("\n");
}
#endif

// One launch of K3 over the class queues K2 filled.  One resident block per SM; chunks =
// class queues in block-sized pieces.
void launch_shade(cudaStream_t stream, int smCount, int64_t nPaths, const DScene *dS, const DSampler *dH,
                  const PathArrays &A, const RenderParams &R, int qi, int betaIn, DeviceCounters *cnt) {
    const int grid = (int)std::min<int64_t>((nPaths + kShadeBlock - 1) / kShadeBlock + kShadeClasses,
                                            (int64_t)smCount * kShadeBlocksPerSM);
    k_shade<<<grid, kShadeBlock, 0, stream>>>(dS, dH, A, R, qi, betaIn, cnt);
#ifdef SPTX_STAGE_CLOCKS
    k_stage_clocks_print<<<1, 1, 0, stream>>>();
#endif
}

// "exact" (the reference's roundings: no FMA contraction, IEEE division and square root) or
// "fast" (FMA contraction, approximate division / square root / reciprocal in K3 only)
const char *shade_arithmetic() {
#ifdef SPTX_SHADE_FAST
    return "fast";
#else
    return "exact";
#endif
}

}  // namespace sptx
