// interaction.cuh -- hit record -> surface interaction.
//
// The traversal kernels emit a 32-byte hit (primitive, instance, t,
// barycentrics); the shading kernel rebuilds what SurfaceInteraction would
// hold, restricted to what the path uses with constant textures: p, pError,
// wo, geometric normal n, shading normal ns and shading dpdu (BSDF frame).
// Follows Triangle::Intersect (shapes/triangle.cpp:293-425), Sphere::Intersect
// (shapes/sphere.cpp:109-156), the SurfaceInteraction constructor and
// SetShadingGeometry (core/interaction.cpp:43-91) and
// Transform::operator()(SurfaceInteraction) (core/transform.cpp:262-297).
#pragma once
#include "intersect.cuh"

namespace sptx {

struct Isect {
    V3 p, pError, wo, n;
    V3 ns, sdpdu;  // shading.n, shading.dpdu
    int material, light, prim;
};
// The surface parameterisation at the hit (SurfaceInteraction::uv, dpdu, dpdv): image textures only,
// so only the textured shading variant asks for it (the others pass NULL).
struct SurfParam {
    V3 dpdu, dpdv;
    V2 uv;
};

SPTX_DEV V3 ld3(const float *base, size_t i) {
    return mk(__ldg(base + 3 * i), __ldg(base + 3 * i + 1), __ldg(base + 3 * i + 2));
}

// Interaction in the space the shape lives in (world for top-level triangles).
SPTX_NOINL void triangle_interaction(const DScene &S, int shape, V3 rayD, float b0, float b1,
                                   float b2, Isect *it, SurfParam *sp = nullptr) {
    const sptx_triangle T = S.triangles[shape];
    const sptx_mesh M = S.meshes[T.mesh];
    V3 p0 = ld3(S.positions, (size_t)M.p_base + T.v[0]);
    V3 p1 = ld3(S.positions, (size_t)M.p_base + T.v[1]);
    V3 p2 = ld3(S.positions, (size_t)M.p_base + T.v[2]);
    V2 uv[3];
    tri_uvs(S, T, M, uv);
    V3 dpdu = mk(0, 0, 0), dpdv = mk(0, 0, 0);
    tri_partials(uv, p0, p1, p2, &dpdu, &dpdv);
    V3 dp02 = p0 - p2, dp12 = p1 - p2;
    float xAbsSum = (fabsf(b0 * p0.x) + fabsf(b1 * p1.x) + fabsf(b2 * p2.x));
    float yAbsSum = (fabsf(b0 * p0.y) + fabsf(b1 * p1.y) + fabsf(b2 * p2.y));
    float zAbsSum = (fabsf(b0 * p0.z) + fabsf(b1 * p1.z) + fabsf(b2 * p2.z));
    it->pError = gamma_n(7) * mk(xAbsSum, yAbsSum, zAbsSum);
    it->p = b0 * p0 + b1 * p1 + b2 * p2;
    if (sp) {
        sp->dpdu = dpdu;
        sp->dpdv = dpdv;
        sp->uv = V2{b0 * uv[0].x + b1 * uv[1].x + b2 * uv[2].x, b0 * uv[0].y + b1 * uv[1].y + b2 * uv[2].y};
    }
    it->wo = normalize(-rayD);
    V3 n = normalize(cross(dp02, dp12));
    V3 ns = n;
    V3 sdpdu = dpdu;
    const bool flip = M.flip != 0;
    if (M.n_base >= 0 || M.s_base >= 0) {
        if (M.n_base >= 0) {
            V3 n0 = ld3(S.normals, (size_t)M.n_base + T.v[0]);
            V3 n1 = ld3(S.normals, (size_t)M.n_base + T.v[1]);
            V3 n2 = ld3(S.normals, (size_t)M.n_base + T.v[2]);
            ns = (b0 * n0 + b1 * n1 + b2 * n2);
            if (length_sq(ns) > 0)
                ns = normalize(ns);
            else
                ns = n;
        } else
            ns = n;
        V3 ss;
        if (M.s_base >= 0) {
            V3 s0 = ld3(S.tangents, (size_t)M.s_base + T.v[0]);
            V3 s1 = ld3(S.tangents, (size_t)M.s_base + T.v[1]);
            V3 s2 = ld3(S.tangents, (size_t)M.s_base + T.v[2]);
            ss = (b0 * s0 + b1 * s1 + b2 * s2);
            if (length_sq(ss) > 0)
                ss = normalize(ss);
            else
                ss = normalize(dpdu);
        } else
            ss = normalize(dpdu);
        V3 ts = cross(ss, ns);
        if (length_sq(ts) > 0.f) {
            ts = normalize(ts);
            ss = cross(ts, ns);
        } else
            coordinate_system(ns, &ss, &ts);
        // SetShadingGeometry(ss, ts, ..., orientationIsAuthoritative = true)
        V3 sn = normalize(cross(ss, ts));
        if (flip) sn = -sn;
        n = faceforward(n, sn);
        ns = sn;
        sdpdu = ss;
    }
    if (M.n_base >= 0)
        n = faceforward(n, ns);
    else if (flip)
        n = ns = -n;
    it->n = n;
    it->ns = ns;
    it->sdpdu = sdpdu;
}

// Object-space sphere interaction moved to the sphere's parent space.  `pHit` is the refined
// object-space hit point Sphere::Intersect computed (shapes/sphere.cpp:84-90), carried by the hit
// record; `rayD` is the direction of the ray in the sphere's parent space.
SPTX_NOINL bool sphere_interaction_at(const DScene &S, int shape, V3 rayD, V3 pHit, Isect *it,
                                      SurfParam *par = nullptr) {
    const sptx_sphere &sp = S.spheres[shape];
    SphereHit sh;
    sh.pHit = pHit;
    sh.oray.d = xf_vec(sp.w2o, rayD);
    const float radius = sp.radius;
    float theta = acosf(clampf(pHit.z / radius, -1.f, 1.f));
    float zRadius = sqrtf(pHit.x * pHit.x + pHit.y * pHit.y);
    float invZRadius = 1.f / zRadius;
    float cosPhi = pHit.x * invZRadius;
    float sinPhi = pHit.y * invZRadius;
    V3 dpdu = mk(-sp.phi_max * pHit.y, sp.phi_max * pHit.x, 0.f);
    V3 dpdv = (sp.theta_max - sp.theta_min) *
              mk(pHit.z * cosPhi, pHit.z * sinPhi, -radius * sinf(theta));
    V3 pError = gamma_n(5) * vabs(pHit);
    // SurfaceInteraction constructor in object space
    V3 n = normalize(cross(dpdu, dpdv));
    V3 wo = normalize(-sh.oray.d);
    if ((sp.reverse_orientation != 0) ^ (sp.swaps_handedness != 0)) n = -n;
    V3 ns = n;
    // (*ObjectToWorld)(SurfaceInteraction)
    it->p = xf_point_err2(sp.o2w, pHit, pError, &it->pError);
    it->n = normalize(xf_normal(sp.w2o, n));
    it->wo = normalize(xf_vec(sp.o2w, wo));
    ns = normalize(xf_normal(sp.w2o, ns));
    it->sdpdu = xf_vec(sp.o2w, dpdu);
    it->ns = faceforward(ns, it->n);
    if (par) {
        par->dpdu = it->sdpdu;
        par->dpdv = xf_vec(sp.o2w, dpdv);
        // (u, v) = (phi / phiMax, (theta - thetaMin) / (thetaMax - thetaMin)), shapes/sphere.cpp:112-114
        float phi = atan2f(pHit.y, pHit.x);
        if (phi < 0) phi += 2 * kPi;
        par->uv = V2{phi / sp.phi_max, (theta - sp.theta_min) / (sp.theta_max - sp.theta_min)};
    }
    return true;
}

// The same from a ray (Shape::Pdf intersects the emitter itself, core/shape.cpp:72-87).
SPTX_NOINL bool sphere_interaction(const DScene &S, int shape, const Ray &ray, Isect *it) {
    Ray r = ray;
    r.tMax = kInf;
    SphereHit sh;
    if (!sphere_test(S.spheres[shape], r, &sh)) return false;
    return sphere_interaction_at(S, shape, ray.d, sh.pHit, it);
}

// Scene-level interaction for a closest-hit record of world ray `ray`.
SPTX_NOINL bool build_interaction(const DScene &S, const Ray &ray, int prim, int inst, float b0,
                                float b1, float b2, Isect *it, SurfParam *par = nullptr) {
    const sptx_prim P = S.prims[prim];
    it->material = P.material;
    it->light = P.light;
    it->prim = prim;
    Ray lr = ray;
    if (inst >= 0) lr = xf_ray(S.instances[inst].w2p, ray);
    if (P.kind == SPTX_PRIM_TRIANGLE)
        triangle_interaction(S, P.shape, lr.d, b0, b1, b2, it, par);
    else  // (b0, b1, b2) of a sphere hit = its object-space hit point
        sphere_interaction_at(S, P.shape, lr.d, mk(b0, b1, b2), it, par);
    if (inst >= 0 && !S.instances[inst].is_identity) {
        // InterpolatedPrimToWorld(*isect), core/primitive.cpp:88-89
        const sptx_instance &I = S.instances[inst];
        V3 pe;
        V3 p = xf_point_err2(I.p2w, it->p, it->pError, &pe);
        it->p = p;
        it->pError = pe;
        it->n = normalize(xf_normal(I.w2p, it->n));
        it->wo = normalize(xf_vec(I.p2w, it->wo));
        V3 ns = normalize(xf_normal(I.w2p, it->ns));
        it->sdpdu = xf_vec(I.p2w, it->sdpdu);
        it->ns = faceforward(ns, it->n);
        if (par) {
            par->dpdu = xf_vec(I.p2w, par->dpdu);
            par->dpdv = xf_vec(I.p2w, par->dpdv);
        }
    }
    return true;
}

}  // namespace sptx
