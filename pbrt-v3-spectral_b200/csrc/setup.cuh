// setup.cuh -- one-time kernels run by sptx_upload_scene and the test hooks.
#pragma once
#include "resolve.cuh"

namespace sptx {

// Per-primitive 48-byte leaf block (see scene.cuh).
__global__ void k_build_leafv(const sptx_prim *prims, int64_t nPrims, const sptx_triangle *tris,
                              const sptx_mesh *meshes, const float *positions, float4 *leafv) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nPrims;
         i += (int64_t)gridDim.x * blockDim.x) {
        const sptx_prim P = prims[i];
        float4 v0 = make_float4(0, 0, 0, 0), v1 = v0, v2 = v0;
        if (P.kind == SPTX_PRIM_TRIANGLE) {
            const sptx_triangle T = tris[P.shape];
            const sptx_mesh M = meshes[T.mesh];
            const float *p0 = positions + 3 * (size_t)(M.p_base + T.v[0]);
            const float *p1 = positions + 3 * (size_t)(M.p_base + T.v[1]);
            const float *p2 = positions + 3 * (size_t)(M.p_base + T.v[2]);
            v0 = make_float4(p0[0], p0[1], p0[2], p1[0]);
            v1 = make_float4(p1[1], p1[2], p2[0], p2[1]);
            v2.x = p2[2];
        }
        v2.y = __int_as_float(P.shape);
        v2.w = __int_as_float(P.kind);
        leafv[3 * i] = v0;
        leafv[3 * i + 1] = v1;
        leafv[3 * i + 2] = v2;
    }
}

// HaltonSampler::GetIndexForSample's per-pixel offset for every pixel of the
// sample bounds (samplers/halton.cpp:96-112).
__global__ void k_pixel_offsets(sptx_sampler s, int x0, int y0, int w, int h, uint64_t *out) {
    const int64_t n = (int64_t)w * h;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        int px = x0 + (int)(i % w), py = y0 + (int)(i / w);
        out[i] = halton_pixel_offset(s, px, py);
    }
}

// SpatialLightDistribution::ComputeDistribution for every voxel, eagerly
// (core/lightdistrib.cpp:221-300).  halton5 holds RadicalInverse(d, i) for
// d = 0..4, i = 0..127.  Row layout: func[n] | cdf[n+1] | funcInt.
// voxList (sparse tables): the voxels that get a row, row r for voxList[r]; NULL: every voxel, row = voxel.
__global__ void k_voxel_tables(DScene S, const float *halton5, float *tables, const int *voxList, int64_t nRows) {
    const int n = S.n_lights;
    const V3 wmin = mk(S.wb_min[0], S.wb_min[1], S.wb_min[2]);
    const V3 wmax = mk(S.wb_max[0], S.wb_max[1], S.wb_max[2]);
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nRows;
         r += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = voxList ? (int64_t)voxList[r] : r;
        int pz = (int)(v % S.n_voxels[2]);
        int py = (int)((v / S.n_voxels[2]) % S.n_voxels[1]);
        int px = (int)(v / ((int64_t)S.n_voxels[2] * S.n_voxels[1]));
        V3 p0 = mk((float)px / (float)S.n_voxels[0], (float)py / (float)S.n_voxels[1],
                   (float)pz / (float)S.n_voxels[2]);
        V3 p1 = mk((float)(px + 1) / (float)S.n_voxels[0], (float)(py + 1) / (float)S.n_voxels[1],
                   (float)(pz + 1) / (float)S.n_voxels[2]);
        V3 a = mk(lerpf(p0.x, wmin.x, wmax.x), lerpf(p0.y, wmin.y, wmax.y), lerpf(p0.z, wmin.z, wmax.z));
        V3 c = mk(lerpf(p1.x, wmin.x, wmax.x), lerpf(p1.y, wmin.y, wmax.y), lerpf(p1.z, wmin.z, wmax.z));
        V3 vmin = mk(fminf(a.x, c.x), fminf(a.y, c.y), fminf(a.z, c.z));
        V3 vmax = mk(fmaxf(a.x, c.x), fmaxf(a.y, c.y), fmaxf(a.z, c.z));
        float *row = tables + r * (int64_t)(2 * n + 2);
        const int nSamples = 128;
        // lightContrib[j] += Li.y() / pdf over the sample points in ascending order, as the reference adds them
        // (lightdistrib.cpp:258-274); the light loop is the outer one here so that the sum lives in a register.
        // Li.y() of a plain spectrum row (every area light: divisor 0) is the same number for every sample
        // point: it is evaluated once per row, with the same operations.
        int yRow = -1;
        float yOfRow = 0.f;
        for (int j = 0; j < n; ++j) {
            const sptx_light L = S.lights[j];
            float acc = 0.f;
            for (int i = 0; i < nSamples; ++i) {
                const float *h = halton5 + 5 * i;
                RefPoint ref;
                ref.p = mk(lerpf(__ldg(h), vmin.x, vmax.x), lerpf(__ldg(h + 1), vmin.y, vmax.y),
                           lerpf(__ldg(h + 2), vmin.z, vmax.z));
                ref.pError = mk(0, 0, 0);
                ref.n = mk(0, 0, 0);
                V2 u{__ldg(h + 3), __ldg(h + 4)};
                LiSample ls = light_sample_li(S, L, ref, u);
                if (ls.pdf > 0) {
                    float yy = 0.f;
                    if (!ls.black) {
                        if (!ls.isEnv && ls.divisor == 0.f) {
                            if (ls.spectrum != yRow) {
                                float t = 0.f;
                                for (int b = 0; b < SPTX_NBINS; ++b) t += S.cie_y[b] * spec(S, ls.spectrum, b);
                                yRow = ls.spectrum;
                                yOfRow = t;
                            }
                            yy = yOfRow;
                        } else {
                            for (int b = 0; b < SPTX_NBINS; ++b) yy += S.cie_y[b] * li_bin(S, ls, b);
                        }
                    }
                    acc += y_finish(S, yy) / ls.pdf;
                }
            }
            row[j] = acc;
        }
        float sum = 0.f;
        for (int j = 0; j < n; ++j) sum += row[j];
        float avg = sum / (float)((size_t)nSamples * (size_t)n);
        float minContrib = (avg > 0) ? (float)(.001 * (double)avg) : 1.f;
        for (int j = 0; j < n; ++j) row[j] = (row[j] < minContrib) ? minContrib : row[j];
        // Distribution1D, core/sampling.h:55-70
        float *cdf = row + n;
        cdf[0] = 0.f;
        for (int j = 1; j < n + 1; ++j) cdf[j] = cdf[j - 1] + row[j - 1] / (float)n;
        float funcInt = cdf[n];
        if (funcInt == 0) {
            for (int j = 1; j < n + 1; ++j) cdf[j] = (float)j / (float)n;
        } else {
            for (int j = 1; j < n + 1; ++j) cdf[j] /= funcInt;
        }
        row[2 * n + 1] = funcInt;
    }
}

// Sparse spatial light tables: the voxels a surface point can fall into.  Every leaf of the scene's
// top-level BVH marks the voxels its bounds overlap, widened by one voxel each way (a hit point is
// inside its primitive's bounds only up to rounding); select_light's own index arithmetic is used, so
// a point inside the bounds maps into the marked range.
SPTX_DEV int voxel_coord(const DScene &S, float p, int a) {
    float o = p - S.wb_min[a];
    if (S.wb_max[a] > S.wb_min[a]) o /= S.wb_max[a] - S.wb_min[a];
    return min(max((int)(o * (float)S.n_voxels[a]), 0), S.n_voxels[a] - 1);
}
__global__ void k_voxel_mark(DScene S, int nodeBase, int nNodes, int *flags) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nNodes; i += gridDim.x * blockDim.x) {
        const float4 n0 = __ldg(S.nodes + 2 * ((size_t)nodeBase + i)), n1 = __ldg(S.nodes + 2 * ((size_t)nodeBase + i) + 1);
        if (((unsigned)__float_as_int(n1.w) & 0xffffu) == 0) continue;  // interior node
        const float lo[3] = {n0.x, n0.y, n0.z}, hi[3] = {n0.w, n1.x, n1.y};
        int a0[3], a1[3];
        for (int a = 0; a < 3; ++a) {
            a0[a] = max(voxel_coord(S, lo[a], a) - 1, 0);
            a1[a] = min(voxel_coord(S, hi[a], a) + 1, S.n_voxels[a] - 1);
        }
        for (int x = a0[0]; x <= a1[0]; ++x)
            for (int y = a0[1]; y <= a1[1]; ++y)
                for (int z = a0[2]; z <= a1[2]; ++z) flags[((size_t)x * S.n_voxels[1] + y) * S.n_voxels[2] + z] = 1;
    }
}
// rows[v] = index of voxel v among the marked ones (or -1); list[row] = v.
__global__ void k_voxel_rows(const int *flags, const int *scan, int64_t nVox, int *rows, int *list) {
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nVox; v += (int64_t)gridDim.x * blockDim.x) {
        const bool on = flags[v] != 0;
        rows[v] = on ? scan[v] : -1;
        if (on) list[scan[v]] = (int)v;
    }
}

// Test hooks ---------------------------------------------------------------------
// Per-dimension constants of the scrambled radical inverse (core/lowdiscrepancy.cpp:405-424),
// computed with the device's own float operations.
__global__ void k_halton_dims(const int32_t *primes, const int32_t *prime_sums, const uint16_t *perms,
                              int n, uint4 *tab, float *tail) {
    int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= n) return;
    const uint32_t base = (uint32_t)primes[d];
    const float invBase = 1.f / (float)base;
    tab[d] = make_uint4(base, 0xffffffffu / base, (uint32_t)prime_sums[d], __float_as_uint(invBase));
    tail[d] = invBase * (float)(int)perms[prime_sums[d]] / (1.f - invBase);
}

// Film rows [P][32] -> contrib [P][31] followed by weight [P] (the layout sptx_film_download
// hands to the host), so that the host copy is two plain transfers.
__global__ void k_film_split(const float *pixels, int64_t nPix, float *out) {
    const int64_t n = nPix * SPTX_SPECTRUM_STRIDE;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p = k / SPTX_SPECTRUM_STRIDE;
        const int b = (int)(k % SPTX_SPECTRUM_STRIDE);
        const float v = pixels[k];
        if (b < SPTX_NBINS)
            out[p * SPTX_NBINS + b] = v;
        else
            out[nPix * SPTX_NBINS + p] = v;
    }
}

// Film merge across GPUs fused with the .dat pack: film[self] = sum over ranks (rank order) of
// film[r], peers read through their NVLink-mapped pointers; one warp per 32 pixels: lane = bin
// for the sums (one 128-byte line per pixel and rank), lane = pixel for the transposed float64
// planes (256-byte runs), through a padded shared-memory tile.
struct PeerFilms {
    const float *film[8];
    int n;
};
__global__ void __launch_bounds__(256)
k_film_merge_pack(PeerFilms P, float *own, int64_t nPix, float scale, double *planes) {
    __shared__ float tile[8][32][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t nGroups = (nPix + 31) / 32;
    for (int64_t g = (int64_t)blockIdx.x * 8 + warp; g < nGroups; g += (int64_t)gridDim.x * 8) {
        const int64_t p0 = g * 32;
        for (int k = 0; k < 32; ++k) {
            const int64_t p = p0 + k;
            float acc = 0.f;
            if (p < nPix) {
                const int64_t at = p * SPTX_SPECTRUM_STRIDE + lane;
                acc = P.film[0][at];
                for (int r = 1; r < P.n; ++r) acc += P.film[r][at];
                own[at] = acc;
            }
            tile[warp][k][lane] = acc;
        }
        __syncwarp();
        if (planes && p0 + lane < nPix)
            for (int b = 0; b < SPTX_NBINS; ++b) {
                    float v = tile[warp][lane][b] + 0.f;  // as k_pack_dat
                    v = v * scale;
                    planes[(int64_t)b * nPix + p0 + lane] = (double)v;
                }
        __syncwarp();
    }
}

// The same sum shaped as a reduce-scatter: every rank runs this over ITS slice of the film
// (float4 index range [q0, q1)), reads the slice from all ranks (its own film locally, the peers
// through NVLink), adds them in rank order -- the order of k_film_merge_pack, so the result is the
// same bit for bit -- and writes the sum into the destination rank's film.  Each NVLink carries
// 1/N of a film per peer instead of rank 0 pulling N - 1 whole films through its own ingress.
__global__ void __launch_bounds__(256)
k_film_reduce_slice(PeerFilms P, float *dst, int64_t q0, int64_t q1) {
    for (int64_t q = q0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1;
         q += (int64_t)gridDim.x * blockDim.x) {
        float4 v[8];
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (r < P.n) v[r] = reinterpret_cast<const float4 *>(P.film[r])[q];
        float4 acc = v[0];
#pragma unroll
        for (int r = 1; r < 8; ++r)
            if (r < P.n) {
                acc.x += v[r].x;
                acc.y += v[r].y;
                acc.z += v[r].z;
                acc.w += v[r].w;
            }
        reinterpret_cast<float4 *>(dst)[q] = acc;
    }
}

__global__ void k_sampler_eval(DSampler H, const int *px, const int *py, const int *sample,
                               const int *dim, int64_t n, float *out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t index;
        if (H.s.type == SPTX_SAMPLER_SOBOL)
            index = sobol_interval_to_index(H, (uint64_t)sample[i], px[i] - H.sb[0], py[i] - H.sb[1]);
        else
            index = halton_pixel_offset(H.s, px[i], py[i]) + (uint64_t)sample[i] * (uint64_t)H.s.sample_stride;
        const float v = halton_sample(H, index, dim[i]);
        // the interleaved form the shading kernel uses must give the same bits
        float v5[5], v2[2];
        const int d5 = min(dim[i], H.s.n_dims - 5), d2 = min(dim[i], H.s.n_dims - 2);
        bool same = true;
        if (d5 >= 0) {
            halton_n<5>(H, index, d5, v5);
            same = same && __float_as_uint(v5[dim[i] - d5]) == __float_as_uint(v);
        }
        if (d2 >= 0) {
            halton_n<2>(H, index, d2, v2);
            same = same && __float_as_uint(v2[dim[i] - d2]) == __float_as_uint(v);
        }
        out[i] = same ? v : __int_as_float(0x7fc00000);
    }
}

__global__ void k_camera_rays(DCamera C, const float *cs, int64_t n, float wavelength,
                              sptx_ray *rays, float *weights) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        V2 pFilm{cs[5 * i], cs[5 * i + 1]};
        V2 pLens{cs[5 * i + 3], cs[5 * i + 4]};
        Ray r;
        r.o = mk(0, 0, 0);
        r.d = mk(0, 0, 1);
        r.tMax = kInf;
        float w = camera_ray(C, pFilm, pLens, wavelength, &r);
        sptx_ray o;
        o.o[0] = r.o.x; o.o[1] = r.o.y; o.o[2] = r.o.z;
        o.d[0] = r.d.x; o.d[1] = r.d.y; o.d[2] = r.d.z;
        o.t_max = r.tMax;
        o.time = lerpf(cs[5 * i + 2], C.c.shutter_open, C.c.shutter_close);
        rays[i] = o;
        weights[i] = w;
    }
}

// Film::WriteImage's spectral branch up to the file (core/film.cpp:226-306):
// value * scale as float64, bin-major planes.
__global__ void k_pack_dat(DFilm F, double *planes) {
    const int64_t nPix = (int64_t)F.width * F.height;
    const int64_t n = nPix * SPTX_NBINS;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t b = i / nPix, p = i % nPix;
        float v = F.pixels[p * SPTX_SPECTRUM_STRIDE + b];
        // spectralData = L; += splatScale * splat (0); *= scale   (all float32)
        v = v + 0.f;
        v = v * F.f.scale;
        planes[i] = (double)v;
    }
}

}  // namespace sptx
