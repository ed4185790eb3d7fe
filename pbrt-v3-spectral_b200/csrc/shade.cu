// shade.cu -- the translation unit of K3 (k_shade), compiled FOUR times into libsptx.so: two
// arithmetic modes, each with (-DSPTX_SHADE_ENV=1, namespaces *_env) and without (=0) the code for
// Light "infinite"; sptx_upload_scene knows whether the scene has one.
//
//   SPTX_SHADE_NS=shade_exact   the reference's roundings (no FMA contraction, IEEE division and
//                               square root), like the rest of the library: per-sample radiance
//                               agrees with the reference to ~1e-7;
//   SPTX_SHADE_NS=shade_fast    -fmad=true -prec-div=false -prec-sqrt=false (-DSPTX_SHADE_FAST):
//                               FMA contraction and approximate division / square root in K3 only.
//
// Traversal, the sampler and the camera live in sptx_api.cu and are exact in both: hit ids, t,
// Halton values and camera rays stay bit-identical to the reference; only per-sample radiance
// moves (by ~1e-7 relative per sample, unbiased; north_star's bar for it is 0.5 % relative RMSE).
// The context picks the mode at run time (sptx_set_option "shade_exact").
#include <algorithm>
#include <cstdio>

#include "shade.cuh"

#ifndef SPTX_SHADE_NS
#error "compile with -DSPTX_SHADE_NS=shade_exact | shade_fast | shade_exact_env | shade_fast_env"
#endif

namespace sptx {
namespace SPTX_SHADE_NS {

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

template <int V>
static void launch_variant(cudaStream_t stream, int grid, const DScene &hS, const DSampler &hH, const PathArrays &A,
                           const RenderParams &R, int qi, int betaIn, int cls, DeviceCounters *cnt) {
    k_shade<kVariantLobes[V], kVariantTerms[V]><<<grid, kShadeBlock, 0, stream>>>(hS, hH, A, R, qi, betaIn, cls, cnt);
}

// K3 for one bounce: one launch per shading class in use, each with the variant compiled for
// the class's lobe types; one resident block per SM, a class queue is walked in block-sized
// chunks.  Returns the number of kernels launched.
int launch_shade(cudaStream_t stream, int smCount, int64_t nPaths, const DScene &hS, const DSampler &hH,
                 const PathArrays &A, const RenderParams &R, int qi, int betaIn, const ShadePlan &plan,
                 DeviceCounters *cnt) {
    const int grid = (int)std::min<int64_t>((nPaths + kShadeBlock - 1) / kShadeBlock,
                                            (int64_t)smCount * kShadeBlocksPerSM);
    int launches = 1;
    // (class 0 has work only with bands or infinite lights; the scene says which, the grid is sized for either)
    k_shade_miss<<<std::max(1, std::min(grid, smCount * 4)), 256, 0, stream>>>(hS, A, R, betaIn, cnt);
    for (int cls = 1; cls < kShadeClasses; ++cls) {
        switch (plan.variant[cls]) {
        case 0: launch_variant<0>(stream, grid, hS, hH, A, R, qi, betaIn, cls, cnt); break;
        case 1: launch_variant<1>(stream, grid, hS, hH, A, R, qi, betaIn, cls, cnt); break;
        case 2: launch_variant<2>(stream, grid, hS, hH, A, R, qi, betaIn, cls, cnt); break;
        case 3: launch_variant<3>(stream, grid, hS, hH, A, R, qi, betaIn, cls, cnt); break;
        case 4: launch_variant<4>(stream, grid, hS, hH, A, R, qi, betaIn, cls, cnt); break;
        default: continue;  // no primitive of the scene is in this class
        }
        ++launches;
    }
#ifdef SPTX_STAGE_CLOCKS
    k_stage_clocks_print<<<1, 1, 0, stream>>>();
#endif
    return launches;
}

}  // namespace SPTX_SHADE_NS
}  // namespace sptx
