// camera.cuh -- camera-ray generation (kernel K1 of the wavefront).
//
// PerspectiveCamera::GenerateRayDifferential (cameras/perspective.cpp:95-146)
// and RealisticCamera::GenerateRay + Camera::GenerateRayDifferential
// (cameras/realistic.cpp:899-932, core/camera.cpp:60-99).  Only the main ray
// and the weight are produced: the differentials exist to filter textures,
// and only constant textures are on the hot path -- but the "both shifted rays
// vignetted -> weight 0" rule of the generic differential code is kept because
// it changes which samples contribute.
#pragma once
#include "scene.cuh"

namespace sptx {

// ConcentricSampleDisk, core/sampling.cpp:113-130.
SPTX_DEV V2 concentric_sample_disk(V2 u) {
    V2 o{2.f * u.x - 1.f, 2.f * u.y - 1.f};
    if (o.x == 0 && o.y == 0) return V2{0.f, 0.f};
    float theta, r;
    if (fabsf(o.x) > fabsf(o.y)) {
        r = o.x;
        theta = kPiOver4 * (o.y / o.x);
    } else {
        r = o.y;
        theta = kPiOver2 - kPiOver4 * (o.x / o.y);
    }
#ifdef SPTX_SHADE_FAST   // K3's fast build: hardware sine / cosine (|theta| <= 3 pi / 4, abs. error < 5e-7)
    float sn, cs;
    __sincosf(theta, &sn, &cs);
    return V2{r * cs, r * sn};
#else
    return V2{r * cosf(theta), r * sinf(theta)};
#endif
}

// Offset rays of a camera sample (RayDifferential, core/geometry.h:894-930), world space.
struct RayDiff {
    V3 rxo, rxd, ryo, ryd;
};

SPTX_DEV float perspective_ray(const DCamera &C, V2 pFilm, V2 pLensU, Ray *out, RayDiff *rd = nullptr) {
    V3 pCamera = xf_point(C.c.raster_to_camera, mk(pFilm.x, pFilm.y, 0.f));
    Ray r;
    r.o = mk(0.f, 0.f, 0.f);
    r.d = normalize(pCamera);
    r.tMax = kInf;
    if (C.c.lens_radius > 0) {
        V2 d = concentric_sample_disk(pLensU);
        V2 pLens{C.c.lens_radius * d.x, C.c.lens_radius * d.y};
        float ft = C.c.focal_distance / r.d.z;
        V3 pFocus = r.o + r.d * ft;
        r.o = mk(pLens.x, pLens.y, 0.f);
        r.d = normalize(pFocus - r.o);
    }
    if (rd) {
        // PerspectiveCamera::GenerateRayDifferential, cameras/perspective.cpp:119-141; dxCamera /
        // dyCamera as the constructor computes them (:58-61)
        const V3 c0 = xf_point(C.c.raster_to_camera, mk(0.f, 0.f, 0.f));
        const V3 dxCamera = xf_point(C.c.raster_to_camera, mk(1.f, 0.f, 0.f)) - c0;
        const V3 dyCamera = xf_point(C.c.raster_to_camera, mk(0.f, 1.f, 0.f)) - c0;
        V3 rxo, rxd, ryo, ryd;
        if (C.c.lens_radius > 0) {
            V3 dx = normalize(pCamera + dxCamera);
            float ft = C.c.focal_distance / dx.z;
            V3 pFocus = ft * dx;
            rxo = r.o;
            rxd = normalize(pFocus - rxo);
            V3 dy = normalize(pCamera + dyCamera);
            ft = C.c.focal_distance / dy.z;
            pFocus = ft * dy;
            ryo = r.o;
            ryd = normalize(pFocus - ryo);
        } else {
            rxo = ryo = r.o;
            rxd = normalize(pCamera + dxCamera);
            ryd = normalize(pCamera + dyCamera);
        }
        // CameraToWorld(RayDifferential), core/transform.h:268-277
        rd->rxo = xf_point(C.c.camera_to_world, rxo);
        rd->ryo = xf_point(C.c.camera_to_world, ryo);
        rd->rxd = xf_vec(C.c.camera_to_world, rxd);
        rd->ryd = xf_vec(C.c.camera_to_world, ryd);
    }
    *out = xf_ray(C.c.camera_to_world, r);
    return 1.f;
}

// Quadratic(Float...), core/pbrt.h:422-438 (roots found in double).
SPTX_DEV bool quadratic_f(float a, float b, float c, float *t0, float *t1) {
    double discrim = (double)b * (double)b - 4 * (double)a * (double)c;
    if (discrim < 0) return false;
    double rootDiscrim = sqrt(discrim);
    double q;
    if (b < 0)
        q = -.5 * ((double)b - rootDiscrim);
    else
        q = -.5 * ((double)b + rootDiscrim);
    *t0 = (float)(q / (double)a);
    *t1 = (float)((double)c / q);
    if (*t0 > *t1) {
        float t = *t0;
        *t0 = *t1;
        *t1 = t;
    }
    return true;
}

// Refract, core/reflection.h:96-108.
SPTX_DEV bool refract(V3 wi, V3 n, float eta, V3 *wt) {
    float cosThetaI = dot(n, wi);
    float sin2ThetaI = fmaxf(0.f, 1.f - cosThetaI * cosThetaI);
    float sin2ThetaT = eta * eta * sin2ThetaI;
    if (sin2ThetaT >= 1) return false;
    float cosThetaT = sqrtf(1.f - sin2ThetaT);
    *wt = eta * (-wi) + (eta * cosThetaI - cosThetaT) * n;
    return true;
}

// RealisticCamera::IntersectSphericalElement, cameras/realistic.cpp:372-392.
SPTX_DEV bool lens_sphere(float radius, float zCenter, const Ray &ray, float *t, V3 *n) {
    V3 o = ray.o - mk(0.f, 0.f, zCenter);
    float A = ray.d.x * ray.d.x + ray.d.y * ray.d.y + ray.d.z * ray.d.z;
    float B = 2.f * (ray.d.x * o.x + ray.d.y * o.y + ray.d.z * o.z);
    float Cc = o.x * o.x + o.y * o.y + o.z * o.z - radius * radius;
    float t0, t1;
    if (!quadratic_f(A, B, Cc, &t0, &t1)) return false;
    bool useCloserT = (ray.d.z > 0) ^ (radius < 0);
    *t = useCloserT ? fminf(t0, t1) : fmaxf(t0, t1);
    if (*t < 0) return false;
    V3 nn = o + (*t) * ray.d;
    *n = faceforward(normalize(nn), -ray.d);
    return true;
}

// RealisticCamera::TraceLensesFromFilm, cameras/realistic.cpp:302-370.
SPTX_DEV bool trace_lenses_from_film(const DCamera &C, const Ray &rCamera, float wavelength,
                                     Ray *rOut) {
    const float flipZ[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 1};
    float elementZ = 0;
    Ray rLens = xf_ray(flipZ, rCamera);
    for (int i = C.c.n_elements - 1; i >= 0; --i) {
        const float curvatureRadius = __ldg(C.lens + 4 * i + 0);
        const float thickness = __ldg(C.lens + 4 * i + 1);
        const float eta = __ldg(C.lens + 4 * i + 2);
        const float apertureRadius = __ldg(C.lens + 4 * i + 3);
        elementZ -= thickness;
        float t;
        V3 n = mk(0.f, 0.f, 0.f);
        bool isStop = (curvatureRadius == 0);
        if (isStop) {
            if (rLens.d.z >= 0.0) return false;
            t = (elementZ - rLens.o.z) / rLens.d.z;
        } else {
            float zCenter = elementZ + curvatureRadius;
            if (!lens_sphere(curvatureRadius, zCenter, rLens, &t, &n)) return false;
        }
        V3 pHit = rLens.o + rLens.d * t;
        float r2 = pHit.x * pHit.x + pHit.y * pHit.y;
        if (r2 > apertureRadius * apertureRadius) return false;
        rLens.o = pHit;
        if (!isStop) {
            V3 w;
            float etaI = eta;
            float etaPrev = (i > 0) ? __ldg(C.lens + 4 * (i - 1) + 2) : 0.f;
            float etaT = (i > 0 && etaPrev != 0) ? etaPrev : 1.f;
            if (C.c.ca_flag && (wavelength >= 400) && (wavelength <= 700)) {
                // (wavelength - 550) * -.04 / 300 + eta, evaluated in double
                if (etaI != 1) etaI = (float)((double)(wavelength - 550.f) * -.04 / 300 + (double)etaI);
                if (etaT != 1) etaT = (float)((double)(wavelength - 550.f) * -.04 / 300 + (double)etaT);
            }
            if (!refract(normalize(-rLens.d), n, etaI / etaT, &w)) return false;
            rLens.d = w;
        }
    }
    if (rOut) *rOut = xf_ray(flipZ, rLens);
    return true;
}

// RealisticCamera::GenerateRay, cameras/realistic.cpp:899-932 (+ SampleExitPupil :832-851).
SPTX_DEV float realistic_generate_ray(const DCamera &C, V2 pFilmRaster, V2 pLensU,
                                      float wavelength, Ray *out) {
    V2 s{pFilmRaster.x / (float)C.full_res[0], pFilmRaster.y / (float)C.full_res[1]};
    V2 pFilm2{lerpf(s.x, C.film_extent[0], C.film_extent[2]),
              lerpf(s.y, C.film_extent[1], C.film_extent[3])};
    V3 pFilm = mk(-pFilm2.x, pFilm2.y, 0.f);
    // SampleExitPupil
    float rFilm = sqrtf(pFilm.x * pFilm.x + pFilm.y * pFilm.y);
    int rIndex = (int)(rFilm / (C.c.film_diagonal / 2.f) * (float)C.c.n_pupil_bounds);
    rIndex = min(C.c.n_pupil_bounds - 1, rIndex);
    const float bx0 = __ldg(C.pupil + 4 * rIndex), by0 = __ldg(C.pupil + 4 * rIndex + 1),
                bx1 = __ldg(C.pupil + 4 * rIndex + 2), by1 = __ldg(C.pupil + 4 * rIndex + 3);
    float boundsArea = (bx1 - bx0) * (by1 - by0);
    V2 pLens{lerpf(pLensU.x, bx0, bx1), lerpf(pLensU.y, by0, by1)};
    float sinTheta = (rFilm != 0) ? pFilm.y / rFilm : 0.f;
    float cosTheta = (rFilm != 0) ? pFilm.x / rFilm : 1.f;
    const float lensRearZ = __ldg(C.lens + 4 * (C.c.n_elements - 1) + 1);
    V3 pRear = mk(cosTheta * pLens.x - sinTheta * pLens.y, sinTheta * pLens.x + cosTheta * pLens.y,
                  lensRearZ);
    Ray rFilmRay{pFilm, pRear - pFilm, kInf};
    Ray r;
    if (!trace_lenses_from_film(C, rFilmRay, wavelength, &r)) return 0.f;
    r = xf_ray(C.c.camera_to_world, r);
    r.d = normalize(r.d);
    *out = r;
    float cosT = normalize(rFilmRay.d).z;
    float cos4 = (cosT * cosT) * (cosT * cosT);
    if (C.c.simple_weighting) {
        const float a0 = (__ldg(C.pupil + 2) - __ldg(C.pupil + 0)) *
                         (__ldg(C.pupil + 3) - __ldg(C.pupil + 1));
        return cos4 * boundsArea / a0;
    }
    return (C.c.shutter_close - C.c.shutter_open) * (cos4 * boundsArea) / (lensRearZ * lensRearZ);
}

// Camera::GenerateRayDifferential for the realistic camera (core/camera.cpp:60-99):
// weight 0 if the main ray or both +-0.05 px shifted rays in x or in y are vignetted.
SPTX_DEV float realistic_ray(const DCamera &C, V2 pFilm, V2 pLensU, float wavelength, Ray *out,
                             RayDiff *rd = nullptr) {
    float wt = realistic_generate_ray(C, pFilm, pLensU, wavelength, out);
    if (wt == 0) return 0.f;
    Ray tmp{out->o, out->d, kInf};
    const float eps[2] = {.05f, -.05f};
    float wtx = 0.f;
    for (int k = 0; k < 2; ++k) {
        V2 p = pFilm;
        p.x += eps[k];
        wtx = realistic_generate_ray(C, p, pLensU, wavelength, &tmp);
        if (rd) {  // rxOrigin = o + (rx.o - o) / eps, likewise the direction (a failed trace leaves rx as it was)
            rd->rxo = out->o + div_by(tmp.o - out->o, eps[k]);
            rd->rxd = out->d + div_by(tmp.d - out->d, eps[k]);
        }
        if (wtx != 0) break;
    }
    if (wtx == 0) return 0.f;
    float wty = 0.f;
    tmp = Ray{out->o, out->d, kInf};
    for (int k = 0; k < 2; ++k) {
        V2 p = pFilm;
        p.y += eps[k];
        wty = realistic_generate_ray(C, p, pLensU, wavelength, &tmp);
        if (rd) {
            rd->ryo = out->o + div_by(tmp.o - out->o, eps[k]);
            rd->ryd = out->d + div_by(tmp.d - out->d, eps[k]);
        }
        if (wty != 0) break;
    }
    if (wty == 0) return 0.f;
    return wt;
}

SPTX_DEV float camera_ray(const DCamera &C, V2 pFilm, V2 pLensU, float wavelength, Ray *out,
                          RayDiff *rd = nullptr) {
    if (C.c.type == SPTX_CAMERA_PERSPECTIVE) return perspective_ray(C, pFilm, pLensU, out, rd);
    return realistic_ray(C, pFilm, pLensU, wavelength, out, rd);
}

}  // namespace sptx
