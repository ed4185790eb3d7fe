// intersect.cuh -- ray/box, ray/triangle, ray/sphere and BVH traversal.
//
// Arithmetic follows the reference operation by operation so that the set of
// accepted hits, the tMax shrink sequence and therefore the winning primitive
// are bit-identical to BVHAccel::Intersect / IntersectP.
#pragma once
#include "scene.cuh"

namespace sptx {

struct HitRec {
    int prim;  // global primitive index, -1 = miss
    int inst;  // instance index, -1 = none
    float b0, b1, b2;
};

struct TravCounters {
    int nodes, tests;
};

// ---- triangle -------------------------------------------------------------

SPTX_DEV void tri_uvs(const DScene &S, const sptx_triangle &T, const sptx_mesh &M, V2 uv[3]) {
    // Triangle::GetUVs, shapes/triangle.h:98-108
    if (M.uv_base >= 0) {
        for (int k = 0; k < 3; ++k) {
            const float *p = S.uvs + 2 * (size_t)(M.uv_base + T.v[k]);
            uv[k] = V2{p[0], p[1]};
        }
    } else {
        uv[0] = V2{0.f, 0.f};
        uv[1] = V2{1.f, 0.f};
        uv[2] = V2{1.f, 1.f};
    }
}

// dpdu/dpdv of a triangle (shapes/triangle.cpp:296-318).  Returns false when
// the reference rejects the hit as degenerate (ng == 0).
SPTX_DEV bool tri_partials(const V2 uv[3], V3 p0, V3 p1, V3 p2, V3 *dpdu, V3 *dpdv) {
    V2 duv02{uv[0].x - uv[2].x, uv[0].y - uv[2].y};
    V2 duv12{uv[1].x - uv[2].x, uv[1].y - uv[2].y};
    V3 dp02 = p0 - p2, dp12 = p1 - p2;
    float determinant = duv02.x * duv12.y - duv02.y * duv12.x;
    bool degenerateUV = (double)fabsf(determinant) < 1e-8;
    if (!degenerateUV) {
        float invdet = 1.f / determinant;
        *dpdu = (duv12.y * dp02 - duv02.y * dp12) * invdet;
        *dpdv = ((-duv12.x) * dp02 + duv02.x * dp12) * invdet;
    }
    if (degenerateUV || length_sq(cross(*dpdu, *dpdv)) == 0) {
        V3 ng = cross(p2 - p0, p1 - p0);
        if (length_sq(ng) == 0) return false;
        coordinate_system(normalize(ng), dpdu, dpdv);
    }
    return true;
}

// The per-ray part of the watertight test: the permutation that makes |d| largest in z and the
// shear constants (shapes/triangle.cpp:202-215 depend on the ray alone).  HOIST = true computes it
// once per ray and BVH level instead of once per triangle.  Measured twice and off everywhere: the
// four registers it keeps live cost more than the three divisions it saves (round 1: C4
// -17..-26 %; round 2, bench scene: K2 53.1 -> 57.1 ms, K4 33.7 -> 34.5 ms per step).
struct TriPre {
    int kz;
    float Sx, Sy, Sz;
};
// Permute(v, kx, ky, kz) for the test's (kx, ky, kz) = (kz + 1, kz + 2, kz) mod 3, a rotation of the
// components: kz = 2 -> (x, y, z), kz = 0 -> (y, z, x), kz = 1 -> (z, x, y).  Written with selp:
// the compiler turned the nested conditional of comp() into branches (BSSY / BRA / BSYNC around
// moves, 13 % of K2's stall samples and instructions on the bench scene, profiles/r02x).
SPTX_DEV float selp_f(float a, float b, int pick_a) {
    float r;
    asm("{ .reg .pred p; setp.ne.s32 p, %3, 0; selp.f32 %0, %1, %2, p; }" : "=f"(r) : "f"(a), "f"(b), "r"(pick_a));
    return r;
}
SPTX_DEV V3 rot3(V3 v, int kz) {
    const int r0 = kz == 0, r1 = kz == 1;
    return mk(selp_f(v.y, selp_f(v.z, v.x, r1), r0), selp_f(v.z, selp_f(v.x, v.y, r1), r0),
              selp_f(v.x, selp_f(v.y, v.z, r1), r0));
}
SPTX_DEV TriPre tri_pre(V3 rd) {
    // MaxDimension(Abs(d)), core/geometry.h: (x > y) ? ((x > z) ? 0 : 2) : ((y > z) ? 1 : 2)
    float ax = fabsf(rd.x), ay = fabsf(rd.y), az = fabsf(rd.z);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    V3 d = rot3(rd, kz);
    TriPre pre;
    pre.kz = kz;
    pre.Sx = -d.x / d.z;
    pre.Sy = -d.y / d.z;
    pre.Sz = 1.f / d.z;
    return pre;
}

// Watertight ray/triangle test, shapes/triangle.cpp:188-291 (== :427-529).
SPTX_DEV bool tri_test_pre(V3 ro, const TriPre &pre, float tMax, V3 p0, V3 p1, V3 p2, float *tOut,
                           float *b0o, float *b1o, float *b2o) {
    V3 p0t = p0 - ro, p1t = p1 - ro, p2t = p2 - ro;
    const int kz = pre.kz;
    p0t = rot3(p0t, kz);
    p1t = rot3(p1t, kz);
    p2t = rot3(p2t, kz);
    const float Sx = pre.Sx, Sy = pre.Sy, Sz = pre.Sz;
    p0t.x += Sx * p0t.z;
    p0t.y += Sy * p0t.z;
    p1t.x += Sx * p1t.z;
    p1t.y += Sy * p1t.z;
    p2t.x += Sx * p2t.z;
    p2t.y += Sy * p2t.z;
    float e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    float e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    float e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = (double)p2t.x * (double)p1t.y;
        double p2typ1tx = (double)p2t.y * (double)p1t.x;
        e0 = (float)(p2typ1tx - p2txp1ty);
        double p0txp2ty = (double)p0t.x * (double)p2t.y;
        double p0typ2tx = (double)p0t.y * (double)p2t.x;
        e1 = (float)(p0typ2tx - p0txp2ty);
        double p1txp0ty = (double)p1t.x * (double)p0t.y;
        double p1typ0tx = (double)p1t.y * (double)p0t.x;
        e2 = (float)(p1typ0tx - p1txp0ty);
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    p0t.z *= Sz;
    p1t.z *= Sz;
    p2t.z *= Sz;
    float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0 && (tScaled >= 0 || tScaled < tMax * det))
        return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > tMax * det))
        return false;
    float invDet = 1.f / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;
    float maxZt = fmaxf(fmaxf(fabsf(p0t.z), fabsf(p1t.z)), fabsf(p2t.z));
    float deltaZ = gamma_n(3) * maxZt;
    float maxXt = fmaxf(fmaxf(fabsf(p0t.x), fabsf(p1t.x)), fabsf(p2t.x));
    float maxYt = fmaxf(fmaxf(fabsf(p0t.y), fabsf(p1t.y)), fabsf(p2t.y));
    float deltaX = gamma_n(5) * (maxXt + maxZt);
    float deltaY = gamma_n(5) * (maxYt + maxZt);
    float deltaE = 2.f * (gamma_n(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = fmaxf(fmaxf(fabsf(e0), fabsf(e1)), fabsf(e2));
    float deltaT = 3.f * (gamma_n(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) *
                   fabsf(invDet);
    if (t <= deltaT) return false;
    *tOut = t;
    *b0o = b0;
    *b1o = b1;
    *b2o = b2;
    return true;
}
SPTX_DEV bool tri_test(V3 ro, V3 rd, float tMax, V3 p0, V3 p1, V3 p2, float *tOut, float *b0o,
                       float *b1o, float *b2o) {
    return tri_test_pre(ro, tri_pre(rd), tMax, p0, p1, p2, tOut, b0o, b1o, b2o);
}

// ---- sphere (EFloat interval arithmetic, core/efloat.h) -------------------

struct EF {
    float v, lo, hi;
};
SPTX_DEV EF ef(float v) { return EF{v, v, v}; }
SPTX_DEV EF ef(float v, float err) {
    if (err == 0.f) return EF{v, v, v};
    return EF{v, next_down(v - err), next_up(v + err)};
}
SPTX_DEV EF operator+(EF a, EF b) {
    return EF{a.v + b.v, next_down(a.lo + b.lo), next_up(a.hi + b.hi)};
}
SPTX_DEV EF operator-(EF a, EF b) {
    return EF{a.v - b.v, next_down(a.lo - b.hi), next_up(a.hi - b.lo)};
}
// std::min / std::max semantics (NaN-order sensitive): min(a,b) = (b < a) ? b : a
SPTX_DEV float smin(float a, float b) { return (b < a) ? b : a; }
SPTX_DEV float smax(float a, float b) { return (a < b) ? b : a; }
SPTX_DEV EF operator*(EF a, EF b) {
    float p0 = a.lo * b.lo, p1 = a.hi * b.lo, p2 = a.lo * b.hi, p3 = a.hi * b.hi;
    return EF{a.v * b.v, next_down(smin(smin(p0, p1), smin(p2, p3))),
              next_up(smax(smax(p0, p1), smax(p2, p3)))};
}
SPTX_DEV EF operator/(EF a, EF b) {
    EF r;
    r.v = a.v / b.v;
    if (b.lo < 0 && b.hi > 0) {
        r.lo = -kInf;
        r.hi = kInf;
    } else {
        float d0 = a.lo / b.lo, d1 = a.hi / b.lo, d2 = a.lo / b.hi, d3 = a.hi / b.hi;
        r.lo = next_down(smin(smin(d0, d1), smin(d2, d3)));
        r.hi = next_up(smax(smax(d0, d1), smax(d2, d3)));
    }
    return r;
}
// Quadratic(EFloat...), core/efloat.h:266-286.
SPTX_DEV bool ef_quadratic(EF A, EF B, EF C, EF *t0, EF *t1) {
    double discrim = (double)B.v * (double)B.v - 4. * (double)A.v * (double)C.v;
    if (discrim < 0.) return false;
    double rootDiscrim = sqrt(discrim);
    EF frd = ef((float)rootDiscrim, (float)((double)kMachineEpsilon * rootDiscrim));
    EF q;
    if (B.v < 0)
        q = ef(-.5f) * (B - frd);
    else
        q = ef(-.5f) * (B + frd);
    *t0 = q / A;
    *t1 = C / q;
    if (t0->v > t1->v) {
        EF tmp = *t0;
        *t0 = *t1;
        *t1 = tmp;
    }
    return true;
}

struct SphereHit {
    V3 pHit;     // object space, refined
    float phi;
    float tHit;
    Ray oray;    // object-space ray
};

// A value with a bound on how far the reference's EFloat interval reaches from it:
// lo >= v - r and hi <= v + r for the interval [lo, hi] core/efloat.h computes for the same
// expression.  The bound is propagated with plain float operations (no directed rounding): every
// EFloat operation widens its exact interval result by at most one rounding (half an ulp of the
// endpoint) plus one NextFloatUp / NextFloatDown step (one ulp), i.e. by less than 3 * 2^-24 of the
// endpoint's magnitude; kRbU = 2^-21 covers that five times over and also absorbs the roundings of
// the bound's own arithmetic.  Used to decide Sphere::Intersect's four interval comparisons from
// the value parts; the EFloat arithmetic itself runs only when a value lies within its bound of
// the threshold (a ray that starts on the sphere, a hit exactly at tMax).
struct RB {
    float v, r;
};
constexpr float kRbU = 0x1p-21f;
SPTX_DEV RB rb_make(float v, float err) { return RB{v, err + kRbU * (fabsf(v) + err)}; }
SPTX_DEV RB rb_add(RB a, RB b, float v) {
    const float w = a.r + b.r;
    return RB{v, w + kRbU * (fabsf(a.v) + fabsf(b.v) + w)};
}
SPTX_DEV RB rb_mul(RB a, RB b, float v) {
    const float aa = fabsf(a.v), ab = fabsf(b.v);
    return RB{v, aa * b.r + ab * a.r + a.r * b.r + kRbU * ((aa + a.r) * (ab + b.r))};
}
SPTX_DEV RB rb_div(RB a, RB b, float v) {
    const float den = fabsf(b.v) - b.r;
    // the divisor's interval may come close to zero (or contain it): no usable bound
    if (!(den > .5f * fabsf(b.v))) return RB{v, kInf};
    const float q = fabsf(v);
    const float w = (a.r + q * b.r) / den;
    return RB{v, w + kRbU * (q + w)};
}

// Sphere::Intersect / IntersectP up to the clipping tests,
// shapes/sphere.cpp:49-107 (== :158-215).
SPTX_NOINL bool sphere_test(const sptx_sphere &sp, const Ray &wr, SphereHit *h) {
    V3 oErr, dErr;
    Ray ray = xf_ray_err(sp.w2o, wr, &oErr, &dErr);
    // The value parts of a, b, c and of the roots, with the float operations of the EFloat
    // operators and of Quadratic(EFloat...) (core/efloat.h:266-286).
    const V3 o = ray.o, d = ray.d;
    const float dxx = d.x * d.x, dyy = d.y * d.y, dzz = d.z * d.z;
    const float av = dxx + dyy + dzz;
    const float dox = d.x * o.x, doy = d.y * o.y, doz = d.z * o.z;
    const float bh = dox + doy + doz;
    const float bv = 2.f * bh;
    const float oxx = o.x * o.x, oyy = o.y * o.y, ozz = o.z * o.z;
    const float oo = oxx + oyy + ozz, rr = sp.radius * sp.radius;
    const float cv = oo - rr;
    // Quadratic decides "no real root" from the value parts alone, in double.
    const double discrim = (double)bv * (double)bv - 4. * (double)av * (double)cv;
    if (discrim < 0.) return false;
    const double rootDiscrim = sqrt(discrim);
    const float rd = (float)rootDiscrim;
    const float qs = (bv < 0) ? bv - rd : bv + rd;
    const float q = -.5f * qs;
    float t0v = q / av, t1v = cv / q;
    const bool swapped = t0v > t1v;
    if (swapped) {
        const float t = t0v;
        t0v = t1v;
        t1v = t;
    }
    // An interval contains its value (lo <= v <= hi): "t0.hi > tMax" holds whenever t0.v > tMax and
    // "t1.lo <= 0" whenever t1.v <= 0 (every shadow ray aimed at a sphere light ends just short of
    // it: t0.v > tMax).
    if (t0v > ray.tMax || t1v <= 0) return false;
    // The comparisons left: t0.hi > tMax, t1.lo <= 0, t0.lo <= 0, t1.hi > tMax.
    bool useT1, t1Over;
    bool decided = false;
    {
        const RB ox = rb_make(o.x, oErr.x), oy = rb_make(o.y, oErr.y), oz = rb_make(o.z, oErr.z);
        const RB dx = rb_make(d.x, dErr.x), dy = rb_make(d.y, dErr.y), dz = rb_make(d.z, dErr.z);
        const RB a = rb_add(rb_add(rb_mul(dx, dx, dxx), rb_mul(dy, dy, dyy), dxx + dyy), rb_mul(dz, dz, dzz), av);
        const RB bs = rb_add(rb_add(rb_mul(dx, ox, dox), rb_mul(dy, oy, doy), dox + doy), rb_mul(dz, oz, doz), bh);
        const RB b = rb_mul(RB{2.f, 0.f}, bs, bv);
        const RB cs = rb_add(rb_add(rb_mul(ox, ox, oxx), rb_mul(oy, oy, oyy), oxx + oyy), rb_mul(oz, oz, ozz), oo);
        const RB c = rb_add(cs, rb_mul(RB{sp.radius, 0.f}, RB{sp.radius, 0.f}, rr), cv);
        const RB frd = rb_make(rd, (float)((double)kMachineEpsilon * rootDiscrim));
        const RB qq = rb_mul(RB{-.5f, 0.f}, rb_add(b, frd, qs), q);
        RB ta = rb_div(qq, a, q / av), tb = rb_div(c, qq, cv / q);
        if (swapped) {
            const RB t = ta;
            ta = tb;
            tb = t;
        }
        // slack for the bound's own roundings and for results in the denormal range
        const float r0 = ta.r * 1.01f + 1e-30f, r1 = tb.r * 1.01f + 1e-30f;
        const bool d1 = t0v + r0 <= ray.tMax;        // then t0.hi <= tMax for certain
        const bool d2 = t1v - r1 > 0.f;              // then t1.lo > 0 for certain
        const bool d3 = t0v <= 0.f || t0v - r0 > 0.f;  // t0.lo <= 0 known either way
        const bool d4 = t1v > ray.tMax || t1v + r1 <= ray.tMax;
        if (d1 && d2 && d3 && d4) {
            decided = true;
            useT1 = t0v <= 0.f;
            t1Over = t1v > ray.tMax;
        }
    }
    if (!decided) {
        EF ox = ef(ray.o.x, oErr.x), oy = ef(ray.o.y, oErr.y), oz = ef(ray.o.z, oErr.z);
        EF dx = ef(ray.d.x, dErr.x), dy = ef(ray.d.y, dErr.y), dz = ef(ray.d.z, dErr.z);
        EF a = dx * dx + dy * dy + dz * dz;
        EF b = ef(2.f) * (dx * ox + dy * oy + dz * oz);
        EF c = ox * ox + oy * oy + oz * oz - ef(sp.radius) * ef(sp.radius);
        EF t0, t1;
        if (!ef_quadratic(a, b, c, &t0, &t1)) return false;
        if (t0.hi > ray.tMax || t1.lo <= 0) return false;
        useT1 = t0.lo <= 0;
        t1Over = t1.hi > ray.tMax;
        t0v = t0.v;
        t1v = t1.v;
    }
    float tsv = t0v;
    if (useT1) {
        if (t1Over) return false;
        tsv = t1v;
    }
    const float radius = sp.radius;
    V3 pHit = ray.o + ray.d * tsv;
    pHit = pHit * (radius / length(pHit));
    if (pHit.x == 0 && pHit.y == 0) pHit.x = 1e-5f * radius;
    // phi = atan2(y, x), + 2 Pi when negative, lies in [0, float(2 Pi)]: a sphere whose phiMax is
    // float(2 Pi) or more (every full sphere: Radians(360) is that float) never clips on phi, and the
    // atan2 is needed only for the others (shapes/sphere.cpp:84-104)
    const bool clipsPhi = sp.phi_max < 2 * kPi;
    float phi = 0.f;
    if (clipsPhi) {
        phi = atan2f(pHit.y, pHit.x);
        if (phi < 0) phi += 2 * kPi;
    }
    if ((sp.z_min > -radius && pHit.z < sp.z_min) || (sp.z_max < radius && pHit.z > sp.z_max) ||
        phi > sp.phi_max) {
        if (tsv == t1v) return false;
        if (t1Over) return false;
        tsv = t1v;
        pHit = ray.o + ray.d * tsv;
        pHit = pHit * (radius / length(pHit));
        if (pHit.x == 0 && pHit.y == 0) pHit.x = 1e-5f * radius;
        if (clipsPhi) {
            phi = atan2f(pHit.y, pHit.x);
            if (phi < 0) phi += 2 * kPi;
        }
        if ((sp.z_min > -radius && pHit.z < sp.z_min) ||
            (sp.z_max < radius && pHit.z > sp.z_max) || phi > sp.phi_max)
            return false;
    }
    h->pHit = pHit;
    h->tHit = tsv;
    h->oray = ray;
    return true;
}

// ---- BVH ------------------------------------------------------------------

// Bounds3::IntersectP(ray, invDir, dirIsNeg), core/geometry.h:1420-1447.
SPTX_DEV bool slab_test(float4 n0, float4 n1, V3 o, V3 invDir, const int neg[3], float tMaxRay) {
    // n0 = {min.x, min.y, min.z, max.x}, n1 = {max.y, max.z, offset, meta}
    const float bminx = n0.x, bminy = n0.y, bminz = n0.z, bmaxx = n0.w, bmaxy = n1.x,
                bmaxz = n1.y;
    const float k = 1.f + 2.f * gamma_n(3);
    float tMin = ((neg[0] ? bmaxx : bminx) - o.x) * invDir.x;
    float tMax = ((neg[0] ? bminx : bmaxx) - o.x) * invDir.x;
    float tyMin = ((neg[1] ? bmaxy : bminy) - o.y) * invDir.y;
    float tyMax = ((neg[1] ? bminy : bmaxy) - o.y) * invDir.y;
    tMax *= k;
    tyMax *= k;
    if (tMin > tyMax || tyMin > tMax) return false;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = ((neg[2] ? bmaxz : bminz) - o.z) * invDir.z;
    float tzMax = ((neg[2] ? bminz : bmaxz) - o.z) * invDir.z;
    tzMax *= k;
    if (tMin > tzMax || tzMin > tMax) return false;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    return (tMin < tMaxRay) && (tMax > 0);
}

// Tests primitive `gp` (global index).  Instances are handled by the caller.
template <bool ANY, bool HOIST>
SPTX_DEV bool prim_test(const DScene &S, int gp, Ray &ray, HitRec *hit, int kind, int shape,
                        float4 v0, float4 v1, float4 v2, const TriPre &pre) {
    if (kind == SPTX_PRIM_TRIANGLE) {
        V3 p0 = mk(v0.x, v0.y, v0.z), p1 = mk(v0.w, v1.x, v1.y), p2 = mk(v1.z, v1.w, v2.x);
        float t, b0, b1, b2;
        if (HOIST) {
            if (!tri_test_pre(ray.o, pre, ray.tMax, p0, p1, p2, &t, &b0, &b1, &b2)) return false;
        } else if (!tri_test(ray.o, ray.d, ray.tMax, p0, p1, p2, &t, &b0, &b1, &b2))
            return false;
        if (ANY) return true;
        // Degenerate-triangle rejection of Triangle::Intersect (:296-318).
        const sptx_triangle T = S.triangles[shape];
        const sptx_mesh M = S.meshes[T.mesh];
        V2 uv[3];
        tri_uvs(S, T, M, uv);
        V3 dpdu, dpdv;
        if (!tri_partials(uv, p0, p1, p2, &dpdu, &dpdv)) return false;
        ray.tMax = t;
        hit->prim = gp;
        hit->b0 = b0;
        hit->b1 = b1;
        hit->b2 = b2;
        return true;
    } else {  // sphere
        SphereHit sh;
        if (!sphere_test(S.spheres[shape], ray, &sh)) return false;
        if (ANY) return true;
        ray.tMax = sh.tHit;
        hit->prim = gp;
        // a sphere hit carries its refined object-space hit point where a triangle hit carries
        // its barycentrics: the interaction is built from it without intersecting again
        hit->b0 = sh.pHit.x;
        hit->b1 = sh.pHit.y;
        hit->b2 = sh.pHit.z;
        return true;
    }
}

// One aggregate: BVHAccel::Intersect (accelerators/bvh.cpp:662-700) when
// !ANY, BVHAccel::IntersectP (:702-738) when ANY.  LEVEL 0 = scene aggregate
// (may contain instances), LEVEL 1 = an instance's aggregate.
template <bool ANY, bool COUNT, int LEVEL, bool HOIST>
__device__ bool traverse_bvh(const DScene &S, int bvhIndex, Ray &ray, HitRec *hit,
                             TravCounters *cnt) {
    const sptx_bvh B = S.bvhs[bvhIndex];
    bool found = false;
    TriPre pre{0, 0.f, 0.f, 0.f};
    if (HOIST) pre = tri_pre(ray.d);
    if (B.n_nodes == 0) {
        // node-less aggregate: a single bare primitive (or an empty BVH)
        for (int i = 0; i < B.n_prims; ++i) {
            int gp = B.prim_base + i;
            float4 v0 = __ldg(S.leafv + 3 * (size_t)gp), v1 = __ldg(S.leafv + 3 * (size_t)gp + 1),
                   v2 = __ldg(S.leafv + 3 * (size_t)gp + 2);
            if (COUNT) cnt->tests++;
            if (prim_test<ANY, HOIST>(S, gp, ray, hit, __float_as_int(v2.w), __float_as_int(v2.y), v0, v1,
                                      v2, pre)) {
                if (ANY) return true;
                found = true;
            }
        }
        return found;
    }
    const float4 *nodes = S.nodes + 2 * (size_t)B.node_base;
    V3 invDir = mk(1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z);
    int neg[3] = {invDir.x < 0, invDir.y < 0, invDir.z < 0};
    const unsigned negMask = (unsigned)neg[0] | ((unsigned)neg[1] << 1) | ((unsigned)neg[2] << 2);
    int stack[64];
    int sp = 0, cur = 0;
    // "while-while": every lane first walks interior nodes until it holds a
    // leaf (or runs out of nodes); the warp reconverges, then the lanes that
    // hold a leaf run the (long) primitive tests together.  Per ray the node
    // and primitive order is exactly the reference's.
    bool done = false;
    while (!done) {
        int leafOffset = 0, leafN = 0;
        while (true) {
            float4 n0 = __ldg(nodes + 2 * (size_t)cur), n1 = __ldg(nodes + 2 * (size_t)cur + 1);
            if (COUNT) cnt->nodes++;
            if (slab_test(n0, n1, ray.o, invDir, neg, ray.tMax)) {
                int offset = __float_as_int(n1.z);
                unsigned meta = (unsigned)__float_as_int(n1.w);
                int nPrims = (int)(meta & 0xffffu);
                if (nPrims > 0) {
                    leafOffset = offset;
                    leafN = nPrims;
                    break;
                }
                int axis = (int)((meta >> 16) & 0xffu);
                int far;
                // dirIsNeg[node->axis]: near child first (accelerators/bvh.cpp:686-692)
                const bool negAxis = ((negMask >> axis) & 1u) != 0;
                far = negAxis ? cur + 1 : offset;
                cur = negAxis ? offset : cur + 1;
                stack[sp++] = far;
                // the far child is visited some iterations from now: have its node in L1 by then
                asm volatile("prefetch.global.L1 [%0];" ::"l"(nodes + 2 * (size_t)far));
            } else {
                if (sp == 0) {
                    done = true;
                    break;
                }
                cur = stack[--sp];
            }
        }
        if (leafN > 0) {
            // (any-hit leaves through the flags, not a return: the loops stay structured and
            //  the warp keeps reconverging after the node phase)
            bool stop = false;
            for (int i = 0; i < leafN && !stop; ++i) {
                int gp = B.prim_base + leafOffset + i;
                const float4 *lv = S.leafv + 3 * (size_t)gp;
                float4 v0 = __ldg(lv), v1 = __ldg(lv + 1), v2 = __ldg(lv + 2);
                int kind = __float_as_int(v2.w), shape = __float_as_int(v2.y);
                if (COUNT) cnt->tests++;
                if (LEVEL == 0 && kind == SPTX_PRIM_INSTANCE) {
                    // TransformedPrimitive::Intersect[P], core/primitive.cpp:78-99
                    const sptx_instance &I = S.instances[shape];
                    Ray r2 = xf_ray(I.w2p, ray);
                    HitRec h2 = *hit;
                    if (traverse_bvh<ANY, COUNT, 1, HOIST>(S, I.bvh, r2, &h2, cnt)) {
                        found = true;
                        if (ANY)
                            stop = true;
                        else {
                            ray.tMax = r2.tMax;
                            *hit = h2;
                            hit->inst = shape;
                        }
                    }
                } else if (prim_test<ANY, HOIST>(S, gp, ray, hit, kind, shape, v0, v1, v2, pre)) {
                    found = true;
                    if (ANY)
                        stop = true;
                    else if (LEVEL == 0)
                        hit->inst = -1;
                }
            }
            if (stop || sp == 0)
                done = true;
            else
                cur = stack[--sp];
        }
    }
    return found;
}

// Scene::Intersect (core/scene.cpp:45-50): closest hit; ray.tMax is updated.
template <bool COUNT, bool HOIST = false>
SPTX_DEV bool scene_intersect(const DScene &S, Ray &ray, HitRec *hit, TravCounters *cnt) {
    hit->prim = -1;
    hit->inst = -1;
    hit->b0 = hit->b1 = hit->b2 = 0.f;
    return traverse_bvh<false, COUNT, 0, HOIST>(S, 0, ray, hit, cnt);
}
// Scene::IntersectP (core/scene.cpp:52-55).
template <bool COUNT, bool HOIST = false>
SPTX_DEV bool scene_intersect_p(const DScene &S, const Ray &ray, TravCounters *cnt) {
    Ray r = ray;
    HitRec h;
    return traverse_bvh<true, COUNT, 0, HOIST>(S, 0, r, &h, cnt);
}

}  // namespace sptx
