// lbvh.cuh -- BVH construction on the device.
//
// What the reference does on the host, serially or over a few threads, before a single ray is
// traced: BVHAccel::BVHAccel (accelerators/bvh.cpp:183-225), for "hlbvh" HLBVHBuild (:404-638):
// Morton codes of the primitive centroids in the centroid bounds (10 bits per axis, :420-432),
// a radix sort (:139-180), LBVH treelets split on Morton bits (emitLBVH :474-538), an upper SAH
// tree over the treelets, then a depth-first flattening into LinearBVHNode (:640-660).
//
// The device version keeps the reference's inputs and outputs -- primitive world bounds in, a
// LinearBVHNode array in depth-first order out (first child at i + 1, second child at `offset`,
// split `axis` for the near / far order), so the traversal kernels are the same -- and replaces the
// serial middle by the fully parallel LBVH of Karras (2012): every internal node finds its range of
// the sorted Morton codes and its split from common-prefix lengths alone, bounds are fitted bottom
// up with one atomic counter per node, and the depth-first numbering comes from the subtree sizes.
// One primitive per leaf.  The tree is NOT the reference's tree (no SAH step, other node order),
// so parity for scenes built here is t-only (hit / miss identical, |dt|/t <= 1e-5) plus the id of
// the hit primitive wherever the closest hit is unique; tests/test_gpu_lbvh.py.
#pragma once
#include <cub/cub.cuh>

#include "scene.cuh"

namespace sptx {

// float <-> ordered int, so that atomicMin / atomicMax on ints order floats
SPTX_DEV int f2ord(float f) {
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
SPTX_DEV float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// Primitive world bounds: Triangle::WorldBound (shapes/triangle.cpp:180-186), Sphere: the
// transformed object bound (shapes/sphere.cpp:45-48, core/transform.cpp Bounds3f operator()),
// instances: the instanced aggregate's root bound through PrimitiveToWorld (core/primitive.cpp:68-70).
__global__ void k_lbvh_prim_bounds(DScene S, int primBase, int n, const float *instBounds, float *bmin, float *bmax,
                                   int *cmin, int *cmax) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int gp = primBase + i;
    const float4 v0 = S.leafv[3 * (size_t)gp], v1 = S.leafv[3 * (size_t)gp + 1], v2 = S.leafv[3 * (size_t)gp + 2];
    const int kind = __float_as_int(v2.w), shape = __float_as_int(v2.y);
    float lo[3], hi[3];
    if (kind == SPTX_PRIM_TRIANGLE) {
        const float p[3][3] = {{v0.x, v0.y, v0.z}, {v0.w, v1.x, v1.y}, {v1.z, v1.w, v2.x}};
        for (int a = 0; a < 3; ++a) {
            lo[a] = fminf(fminf(p[0][a], p[1][a]), p[2][a]);
            hi[a] = fmaxf(fmaxf(p[0][a], p[1][a]), p[2][a]);
        }
    } else {
        float ob[2][3];
        const float *m;
        if (kind == SPTX_PRIM_SPHERE) {
            const sptx_sphere &sp = S.spheres[shape];
            ob[0][0] = ob[0][1] = -sp.radius;
            ob[1][0] = ob[1][1] = sp.radius;
            ob[0][2] = sp.z_min;
            ob[1][2] = sp.z_max;
            m = sp.o2w;
        } else {
            const sptx_instance &I = S.instances[shape];
            for (int a = 0; a < 3; ++a) {
                ob[0][a] = instBounds[6 * I.bvh + a];
                ob[1][a] = instBounds[6 * I.bvh + 3 + a];
            }
            m = I.p2w;
        }
        for (int a = 0; a < 3; ++a) {
            lo[a] = kInf;
            hi[a] = -kInf;
        }
        for (int c = 0; c < 8; ++c) {
            const V3 q = xf_point(m, mk(ob[c & 1][0], ob[(c >> 1) & 1][1], ob[(c >> 2) & 1][2]));
            const float qa[3] = {q.x, q.y, q.z};
            for (int a = 0; a < 3; ++a) {
                lo[a] = fminf(lo[a], qa[a]);
                hi[a] = fmaxf(hi[a], qa[a]);
            }
        }
    }
    for (int a = 0; a < 3; ++a) {
        bmin[3 * (size_t)i + a] = lo[a];
        bmax[3 * (size_t)i + a] = hi[a];
        const float c = .5f * lo[a] + .5f * hi[a];  // Bounds3::Centroid... (accelerators/bvh.cpp:63)
        atomicMin(cmin + a, f2ord(c));
        atomicMax(cmax + a, f2ord(c));
    }
}

// LeftShift3 / EncodeMorton3, accelerators/bvh.cpp:108-137.
SPTX_DEV uint32_t left_shift3(uint32_t x) {
    if (x == (1u << 10)) --x;
    x = (x | (x << 16)) & 0x30000ffu;
    x = (x | (x << 8)) & 0x300f00fu;
    x = (x | (x << 4)) & 0x30c30c3u;
    x = (x | (x << 2)) & 0x9249249u;
    return x;
}
__global__ void k_lbvh_morton(int n, const float *bmin, const float *bmax, const int *cmin, const int *cmax,
                              uint32_t *codes, int *order) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[3];
    for (int a = 0; a < 3; ++a) {
        const float lo = ord2f(cmin[a]), hi = ord2f(cmax[a]);
        const float c = .5f * bmin[3 * (size_t)i + a] + .5f * bmax[3 * (size_t)i + a];
        float o = c - lo;                    // Bounds3::Offset
        if (hi > lo) o /= hi - lo;
        const float s = o * 1024.f;          // mortonScale = 1 << mortonBits
        q[a] = (uint32_t)fminf(fmaxf(s, 0.f), 1024.f);
    }
    codes[i] = (left_shift3(q[2]) << 2) | (left_shift3(q[1]) << 1) | left_shift3(q[0]);
    order[i] = i;
}

// Common-prefix length of the keys (code, position) at sorted positions i and j; -1 outside [0, n).
SPTX_DEV int lbvh_delta(const uint32_t *codes, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = codes[i], b = codes[j];
    if (a == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);  // equal codes: the position breaks the tie
    return __clz(a ^ b);
}

// Karras 2012, "Maximizing parallelism in the construction of BVHs, octrees, and k-d trees", alg. 1.
__global__ void k_lbvh_tree(int n, const uint32_t *codes, int *left, int *right, int *parent, int *axis) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (lbvh_delta(codes, n, i, i + 1) - lbvh_delta(codes, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(codes, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(codes, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(codes, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (lbvh_delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int lc = (lo == gamma) ? ~gamma : gamma;            // leaf ids are stored as ~id
    const int rc = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    left[i] = lc;
    right[i] = rc;
    parent[lc >= 0 ? lc : n - 1 + ~lc] = i;
    parent[rc >= 0 ? rc : n - 1 + ~rc] = i;
    if (i == 0) parent[0] = -1;
    // the Morton bit the two children differ in names the split axis (bit 3k + a <-> axis a:
    // EncodeMorton3 puts x in bit 0); ranges of equal codes are split by position: axis 0
    axis[i] = dnode < 32 ? (31 - dnode) % 3 : 0;
}

// Bottom-up: leaf bounds = primitive bounds; an internal node is finished by the second child to
// arrive (atomic counter): bounds = union, size = 1 + sizes of the children.
__global__ void k_lbvh_fit(int n, const int *orderSorted, const float *bmin, const float *bmax, const int *left,
                           const int *right, const int *parent, int *flag, float *nbmin, float *nbmax, int *size) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int src = orderSorted[j];
    int node = n - 1 + j;
    for (int a = 0; a < 3; ++a) {
        nbmin[3 * (size_t)node + a] = bmin[3 * (size_t)src + a];
        nbmax[3 * (size_t)node + a] = bmax[3 * (size_t)src + a];
    }
    size[node] = 1;
    __threadfence();
    int p = parent[node];
    while (p >= 0) {
        if (atomicAdd(flag + p, 1) == 0) return;  // the sibling finishes this node
        __threadfence();
        const int lc = left[p], rc = right[p];
        const int li = lc >= 0 ? lc : n - 1 + ~lc, ri = rc >= 0 ? rc : n - 1 + ~rc;
        // (the children were written by other SMs: read them at L2, past this SM's L1)
        for (int a = 0; a < 3; ++a) {
            nbmin[3 * (size_t)p + a] = fminf(__ldcg(nbmin + 3 * (size_t)li + a), __ldcg(nbmin + 3 * (size_t)ri + a));
            nbmax[3 * (size_t)p + a] = fmaxf(__ldcg(nbmax + 3 * (size_t)li + a), __ldcg(nbmax + 3 * (size_t)ri + a));
        }
        size[p] = 1 + __ldcg(size + li) + __ldcg(size + ri);
        __threadfence();
        p = parent[p];
    }
}

// Depth-first index of every node, top down, one launch per tree level: a node at index d has its
// first child at d + 1 and its second child at d + 1 + size(first child) (flattenBVHTree,
// accelerators/bvh.cpp:640-660).  `frontier` holds the internal nodes of the current level.
__global__ void k_lbvh_number(int n, int nFrontier, const int *frontier, int *next, int *nNext, const int *left,
                              const int *right, const int *size, int *dfs) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nFrontier) return;
    const int p = frontier[k];
    const int lc = left[p], rc = right[p];
    const int li = lc >= 0 ? lc : n - 1 + ~lc, ri = rc >= 0 ? rc : n - 1 + ~rc;
    const int d = dfs[p];
    dfs[li] = d + 1;
    dfs[ri] = d + 1 + size[li];
    if (lc >= 0) next[atomicAdd(nNext, 1)] = lc;
    if (rc >= 0) next[atomicAdd(nNext, 1)] = rc;
}

// LinearBVHNode records in depth-first order; leaf j holds the primitive at sorted position j.
__global__ void k_lbvh_emit(int n, const int *left, const int *right, const int *axis, const int *dfs,
                            const float *nbmin, const float *nbmax, sptx_bvh_node *nodes) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * n - 1) return;
    sptx_bvh_node out;
    for (int a = 0; a < 3; ++a) {
        out.bmin[a] = nbmin[3 * (size_t)t + a];
        out.bmax[a] = nbmax[3 * (size_t)t + a];
    }
    if (t < n - 1) {  // interior: secondChildOffset
        const int rc = right[t];
        out.offset = dfs[rc >= 0 ? rc : n - 1 + ~rc];
        out.n_prims = 0;
        out.axis = (uint8_t)axis[t];
    } else {          // leaf: primitivesOffset
        out.offset = t - (n - 1);
        out.n_prims = 1;
        out.axis = 0;
    }
    out.pad = 0;
    nodes[dfs[t]] = out;
}

// Gathers the per-primitive tables into sorted order.
__global__ void k_lbvh_permute(int n, int primBase, const int *orderSorted, const sptx_prim *primsIn,
                               const float4 *leafvIn, const uint8_t *classIn, const int *idIn, sptx_prim *primsOut,
                               float4 *leafvOut, uint8_t *classOut, int *idOut) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int src = primBase + orderSorted[j], dst = primBase + j;
    primsOut[dst] = primsIn[src];
    leafvOut[3 * (size_t)dst] = leafvIn[3 * (size_t)src];
    leafvOut[3 * (size_t)dst + 1] = leafvIn[3 * (size_t)src + 1];
    leafvOut[3 * (size_t)dst + 2] = leafvIn[3 * (size_t)src + 2];
    classOut[dst] = classIn[src];
    idOut[dst] = idIn[src];
}

}  // namespace sptx
