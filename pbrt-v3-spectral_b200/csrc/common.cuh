// common.cuh -- float32 vector math with the reference's rounding behaviour.
//
// Everything here restates arithmetic from the reference so that device
// results round the same way as its x86-64 build (no FMA contraction; the
// library is compiled with -fmad=false).  Citations are to /root/reference/src.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "sptx.h"

#define SPTX_DEV __device__ __forceinline__
#define SPTX_HD __host__ __device__ __forceinline__
// (was __noinline__ while the shade kernel thrashed the instruction cache; with staged,
//  class-sorted warps inlining wins again: structs stay in registers instead of the stack)
#ifdef SPTX_NO_INLINE
#define SPTX_NOINL __device__ __noinline__
#else
#define SPTX_NOINL __device__ __forceinline__
#endif

namespace sptx {

// core/pbrt.h:196-215
constexpr float kInf = __builtin_huge_valf();
constexpr float kMachineEpsilon = 5.9604644775390625e-08f;  // 2^-24
constexpr float kShadowEpsilon = 0.0001f;
constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInv2Pi = 0.15915494309189533577f;
constexpr float kInv4Pi = 0.07957747154594766788f;
constexpr float kPiOver2 = 1.57079632679489661923f;
constexpr float kPiOver4 = 0.78539816339744830961f;
constexpr float kOneMinusEpsilon = 0x1.fffffep-1f;  // core/rng.h:53

// gamma(n), core/pbrt.h:292-294: (n * eps) / (1 - n * eps), all in float.
SPTX_HD constexpr float gamma_n(int n) {
    return ((float)n * kMachineEpsilon) / (1.f - (float)n * kMachineEpsilon);
}

struct V3 {
    float x, y, z;
};
struct V2 {
    float x, y;
};

SPTX_HD V3 mk(float x, float y, float z) { return V3{x, y, z}; }
SPTX_HD V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
SPTX_HD V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
SPTX_HD V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
SPTX_HD V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
SPTX_HD V3 operator*(float s, V3 a) { return mk(a.x * s, a.y * s, a.z * s); }
// Vector3::operator/ multiplies by the reciprocal (core/geometry.h:245-249).
SPTX_HD V3 div_by(V3 a, float f) {
    float inv = 1.f / f;
    return mk(a.x * inv, a.y * inv, a.z * inv);
}
SPTX_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
SPTX_HD float absdot(V3 a, V3 b) { return fabsf(dot(a, b)); }
SPTX_HD float length_sq(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
SPTX_HD float length(V3 a) { return sqrtf(length_sq(a)); }
SPTX_HD V3 normalize(V3 a) { return div_by(a, length(a)); }
SPTX_HD V3 vabs(V3 a) { return mk(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
// Cross is evaluated in double and rounded once (core/geometry.h:966-972).
SPTX_HD V3 cross(V3 a, V3 b) {
    double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return mk((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)),
              (float)((ax * by) - (ay * bx)));
}
SPTX_HD float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
SPTX_HD V3 faceforward(V3 n, V3 v) { return (dot(n, v) < 0.f) ? -n : n; }
SPTX_HD float dist_sq(V3 a, V3 b) { return length_sq(a - b); }
SPTX_HD float dist(V3 a, V3 b) { return length(a - b); }

// CoordinateSystem, core/geometry.h:1029-1036.
SPTX_HD void coordinate_system(V3 v1, V3 *v2, V3 *v3) {
    if (fabsf(v1.x) > fabsf(v1.y))
        *v2 = div_by(mk(-v1.z, 0.f, v1.x), sqrtf(v1.x * v1.x + v1.z * v1.z));
    else
        *v2 = div_by(mk(0.f, v1.z, -v1.y), sqrtf(v1.y * v1.y + v1.z * v1.z));
    *v3 = cross(v1, *v2);
}

// NextFloatUp / NextFloatDown, core/pbrt.h:241-265.
SPTX_HD uint32_t f2b(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}
SPTX_HD float b2f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
SPTX_HD float next_up(float v) {
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = f2b(v);
    if (v >= 0) ++ui; else --ui;
    return b2f(ui);
}
SPTX_HD float next_down(float v) {
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = f2b(v);
    if (v > 0) --ui; else ++ui;
    return b2f(ui);
}

SPTX_HD float clampf(float v, float lo, float hi) {
    // Clamp, core/pbrt.h: if (val < low) low; else if (val > high) high; else val
    if (v < lo) return lo;
    if (v > hi) return hi;
    return v;
}
SPTX_HD float lerpf(float t, float a, float b) { return (1.f - t) * a + t * b; }

// OffsetRayOrigin, core/geometry.h:1449-1469.
SPTX_HD V3 offset_ray_origin(V3 p, V3 pError, V3 n, V3 w) {
    float d = dot(vabs(n), pError);
    V3 offset = d * n;
    if (dot(w, n) < 0) offset = -offset;
    V3 po = p + offset;
    if (offset.x > 0) po.x = next_up(po.x); else if (offset.x < 0) po.x = next_down(po.x);
    if (offset.y > 0) po.y = next_up(po.y); else if (offset.y < 0) po.y = next_down(po.y);
    if (offset.z > 0) po.z = next_up(po.z); else if (offset.z < 0) po.z = next_down(po.z);
    return po;
}

// 4x4 transforms, row-major (core/transform.h:213-266).
struct M44 {
    float m[16];
};
SPTX_HD V3 xf_point(const float *m, V3 p) {
    float x = p.x, y = p.y, z = p.z;
    float xp = m[0] * x + m[1] * y + m[2] * z + m[3];
    float yp = m[4] * x + m[5] * y + m[6] * z + m[7];
    float zp = m[8] * x + m[9] * y + m[10] * z + m[11];
    float wp = m[12] * x + m[13] * y + m[14] * z + m[15];
    if (wp == 1.f) return mk(xp, yp, zp);
    float inv = 1.f / wp;  // Point3::operator/, core/geometry.h:500-504
    return mk(inv * xp, inv * yp, inv * zp);
}
SPTX_HD V3 xf_point_err(const float *m, V3 p, V3 *err) {
    float x = p.x, y = p.y, z = p.z;
    float xp = m[0] * x + m[1] * y + m[2] * z + m[3];
    float yp = m[4] * x + m[5] * y + m[6] * z + m[7];
    float zp = m[8] * x + m[9] * y + m[10] * z + m[11];
    float wp = m[12] * x + m[13] * y + m[14] * z + m[15];
    float xs = (fabsf(m[0] * x) + fabsf(m[1] * y) + fabsf(m[2] * z) + fabsf(m[3]));
    float ys = (fabsf(m[4] * x) + fabsf(m[5] * y) + fabsf(m[6] * z) + fabsf(m[7]));
    float zs = (fabsf(m[8] * x) + fabsf(m[9] * y) + fabsf(m[10] * z) + fabsf(m[11]));
    *err = gamma_n(3) * mk(xs, ys, zs);
    if (wp == 1.f) return mk(xp, yp, zp);
    float inv = 1.f / wp;
    return mk(inv * xp, inv * yp, inv * zp);
}
// Point with incoming error, core/transform.h:305-333.
SPTX_HD V3 xf_point_err2(const float *m, V3 p, V3 e, V3 *err) {
    float x = p.x, y = p.y, z = p.z;
    float xp = m[0] * x + m[1] * y + m[2] * z + m[3];
    float yp = m[4] * x + m[5] * y + m[6] * z + m[7];
    float zp = m[8] * x + m[9] * y + m[10] * z + m[11];
    float wp = m[12] * x + m[13] * y + m[14] * z + m[15];
    const float g3 = gamma_n(3);
    err->x = (g3 + 1.f) * (fabsf(m[0]) * e.x + fabsf(m[1]) * e.y + fabsf(m[2]) * e.z) +
             g3 * (fabsf(m[0] * x) + fabsf(m[1] * y) + fabsf(m[2] * z) + fabsf(m[3]));
    err->y = (g3 + 1.f) * (fabsf(m[4]) * e.x + fabsf(m[5]) * e.y + fabsf(m[6]) * e.z) +
             g3 * (fabsf(m[4] * x) + fabsf(m[5] * y) + fabsf(m[6] * z) + fabsf(m[7]));
    err->z = (g3 + 1.f) * (fabsf(m[8]) * e.x + fabsf(m[9]) * e.y + fabsf(m[10]) * e.z) +
             g3 * (fabsf(m[8] * x) + fabsf(m[9] * y) + fabsf(m[10] * z) + fabsf(m[11]));
    if (wp == 1.f) return mk(xp, yp, zp);
    float inv = 1.f / wp;
    return mk(inv * xp, inv * yp, inv * zp);
}
SPTX_HD V3 xf_vec(const float *m, V3 v) {
    float x = v.x, y = v.y, z = v.z;
    return mk(m[0] * x + m[1] * y + m[2] * z, m[4] * x + m[5] * y + m[6] * z,
              m[8] * x + m[9] * y + m[10] * z);
}
SPTX_HD V3 xf_vec_err(const float *m, V3 v, V3 *err) {
    float x = v.x, y = v.y, z = v.z;
    err->x = gamma_n(3) * (fabsf(m[0] * x) + fabsf(m[1] * y) + fabsf(m[2] * z));
    err->y = gamma_n(3) * (fabsf(m[4] * x) + fabsf(m[5] * y) + fabsf(m[6] * z));
    err->z = gamma_n(3) * (fabsf(m[8] * x) + fabsf(m[9] * y) + fabsf(m[10] * z));
    return mk(m[0] * x + m[1] * y + m[2] * z, m[4] * x + m[5] * y + m[6] * z,
              m[8] * x + m[9] * y + m[10] * z);
}
// Normals use the transposed inverse (core/transform.h:243-249); pass mInv.
SPTX_HD V3 xf_normal(const float *mi, V3 n) {
    float x = n.x, y = n.y, z = n.z;
    return mk(mi[0] * x + mi[4] * y + mi[8] * z, mi[1] * x + mi[5] * y + mi[9] * z,
              mi[2] * x + mi[6] * y + mi[10] * z);
}

struct Ray {
    V3 o, d;
    float tMax;
};

// Transform::operator()(const Ray&), core/transform.h:251-266 (tMax shrinks).
SPTX_HD Ray xf_ray(const float *m, const Ray &r) {
    V3 oErr;
    V3 o = xf_point_err(m, r.o, &oErr);
    V3 d = xf_vec(m, r.d);
    float l2 = length_sq(d);
    float tMax = r.tMax;
    if (l2 > 0) {
        float dt = dot(vabs(d), oErr) / l2;
        o = o + d * dt;
        tMax -= dt;
    }
    return Ray{o, d, tMax};
}
// Transform::operator()(const Ray&, oErr*, dErr*), core/transform.h:387-399
// (tMax is NOT shrunk in this overload).
SPTX_HD Ray xf_ray_err(const float *m, const Ray &r, V3 *oErr, V3 *dErr) {
    V3 o = xf_point_err(m, r.o, oErr);
    V3 d = xf_vec_err(m, r.d, dErr);
    float l2 = length_sq(d);
    if (l2 > 0) {
        float dt = dot(vabs(d), *oErr) / l2;
        o = o + d * dt;
    }
    return Ray{o, d, r.tMax};
}

}  // namespace sptx
