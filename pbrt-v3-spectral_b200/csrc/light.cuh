// light.cuh -- DiffuseAreaLight / PointLight sampling and light selection.
//
// lights/diffuse.cpp:64-87, lights/point.cpp:44-54, core/shape.cpp:56-87,
// shapes/triangle.cpp:575-608, shapes/sphere.cpp:219-306,
// core/sampling.h:55-109 (Distribution1D), core/lightdistrib.cpp.
#pragma once
#include "bsdf_eval.cuh"
#include "sampler.cuh"  // mod_i

namespace sptx {

// Reference point of a light query: a surface point, or a bare point with
// n = 0 (what SpatialLightDistribution::ComputeDistribution passes).
struct RefPoint {
    V3 p, pError, n;
};

struct LightPoint {  // Interaction returned by Shape::Sample
    V3 p, pError, n;
};

// SphericalTheta / SphericalPhi, core/geometry.h:1485-1492.
SPTX_DEV float spherical_theta(V3 v) { return acosf(clampf(v.z, -1.f, 1.f)); }
SPTX_DEV float spherical_phi(V3 v) {
    const float p = atan2f(v.y, v.x);
    return (p < 0) ? (p + 2 * kPi) : p;
}

SPTX_DEV V3 uniform_sample_sphere(V2 u) {
    float z = 1.f - 2.f * u.x;
    float r = sqrtf(fmaxf(0.f, 1.f - z * z));
    float phi = 2.f * kPi * u.y;
    return mk(r * cosf(phi), r * sinf(phi), z);
}

SPTX_DEV void tri_vertices(const DScene &S, int shape, V3 *p0, V3 *p1, V3 *p2, sptx_triangle *T,
                           sptx_mesh *M) {
    *T = S.triangles[shape];
    *M = S.meshes[T->mesh];
    *p0 = ld3(S.positions, (size_t)M->p_base + T->v[0]);
    *p1 = ld3(S.positions, (size_t)M->p_base + T->v[1]);
    *p2 = ld3(S.positions, (size_t)M->p_base + T->v[2]);
}

// Triangle::Sample(u, pdf), shapes/triangle.cpp:583-608.
SPTX_DEV LightPoint triangle_sample(const DScene &S, int shape, V2 u, float area, float *pdf) {
    float su0 = sqrtf(u.x);
    float b0 = 1.f - su0, b1 = u.y * su0;
    V3 p0, p1, p2;
    sptx_triangle T;
    sptx_mesh M;
    tri_vertices(S, shape, &p0, &p1, &p2, &T, &M);
    LightPoint it;
    float b2 = (1.f - b0 - b1);
    it.p = b0 * p0 + b1 * p1 + b2 * p2;
    it.n = normalize(cross(p1 - p0, p2 - p0));
    if (M.n_base >= 0) {
        V3 n0 = ld3(S.normals, (size_t)M.n_base + T.v[0]);
        V3 n1 = ld3(S.normals, (size_t)M.n_base + T.v[1]);
        V3 n2 = ld3(S.normals, (size_t)M.n_base + T.v[2]);
        V3 ns = b0 * n0 + b1 * n1 + b2 * n2;
        it.n = faceforward(it.n, ns);
    } else if (M.flip)
        it.n = it.n * -1.f;
    V3 pAbsSum = vabs(b0 * p0) + vabs(b1 * p1) + vabs(b2 * p2);
    it.pError = gamma_n(6) * pAbsSum;
    *pdf = 1.f / area;
    return it;
}

// Sphere::Sample(u, pdf), shapes/sphere.cpp:219-230.
SPTX_DEV LightPoint sphere_sample_area(const sptx_sphere &sp, V2 u, float area, float *pdf) {
    V3 pObj = mk(0, 0, 0) + sp.radius * uniform_sample_sphere(u);
    LightPoint it;
    it.n = normalize(xf_normal(sp.w2o, pObj));
    if (sp.reverse_orientation) it.n = it.n * -1.f;
    pObj = pObj * (sp.radius / length(pObj));
    V3 pObjError = gamma_n(5) * vabs(pObj);
    it.p = xf_point_err2(sp.o2w, pObj, pObjError, &it.pError);
    *pdf = 1.f / area;
    return it;
}

// Shape::Sample(ref, u, pdf), core/shape.cpp:56-70 (area -> solid angle).
SPTX_DEV void area_to_solid_angle(const RefPoint &ref, const LightPoint &lp, float *pdf,
                                  bool checkInfAlways) {
    V3 wi = lp.p - ref.p;
    if (length_sq(wi) == 0)
        *pdf = 0.f;
    else {
        wi = normalize(wi);
        *pdf *= dist_sq(ref.p, lp.p) / absdot(lp.n, -wi);
        if (!checkInfAlways && isinf(*pdf)) *pdf = 0.f;
    }
    if (checkInfAlways && isinf(*pdf)) *pdf = 0.f;
}

// Sphere::Sample(ref, u, pdf), shapes/sphere.cpp:232-292.
SPTX_DEV LightPoint sphere_sample(const sptx_sphere &sp, const RefPoint &ref, V2 u, float area,
                                  float *pdf) {
    V3 pCenter = xf_point(sp.o2w, mk(0, 0, 0));
    V3 pOrigin = offset_ray_origin(ref.p, ref.pError, ref.n, pCenter - ref.p);
    const float radius = sp.radius;
    if (dist_sq(pOrigin, pCenter) <= radius * radius) {
        LightPoint lp = sphere_sample_area(sp, u, area, pdf);
        area_to_solid_angle(ref, lp, pdf, true);
        return lp;
    }
    V3 wc = normalize(pCenter - ref.p);
    V3 wcX, wcY;
    coordinate_system(wc, &wcX, &wcY);
    float sinThetaMax2 = radius * radius / dist_sq(ref.p, pCenter);
    float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
    float cosTheta = (1.f - u.x) + u.x * cosThetaMax;
    float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
    float phi = u.y * 2.f * kPi;
    float dc = dist(ref.p, pCenter);
    float ds = dc * cosTheta - sqrtf(fmaxf(0.f, radius * radius - dc * dc * sinTheta * sinTheta));
    float cosAlpha = (dc * dc + radius * radius - ds * ds) / (2.f * dc * radius);
    float sinAlpha = sqrtf(fmaxf(0.f, 1.f - cosAlpha * cosAlpha));
    // SphericalDirection(sinAlpha, cosAlpha, phi, -wcX, -wcY, -wc), core/geometry.h:1478-1483
    V3 nWorld = sinAlpha * cosf(phi) * (-wcX) + sinAlpha * sinf(phi) * (-wcY) + cosAlpha * (-wc);
    V3 pWorld = pCenter + radius * nWorld;
    LightPoint it;
    it.p = pWorld;
    it.pError = gamma_n(5) * vabs(pWorld);
    it.n = nWorld;
    if (sp.reverse_orientation) it.n = it.n * -1.f;
    *pdf = 1.f / (2.f * kPi * (1.f - cosThetaMax));
    return it;
}

// Shape::Pdf(ref, wi), core/shape.cpp:72-87, for one triangle / one sphere.
SPTX_NOINL float shape_pdf_generic(const DScene &S, int kind, int shape, const RefPoint &ref, V3 wi,
                                 float area) {
    Ray ray{offset_ray_origin(ref.p, ref.pError, ref.n, wi), wi, kInf};
    V3 pHit, nHit;
    if (kind == SPTX_PRIM_TRIANGLE) {
        V3 p0, p1, p2;
        sptx_triangle T;
        sptx_mesh M;
        tri_vertices(S, shape, &p0, &p1, &p2, &T, &M);
        float t, b0, b1, b2;
        if (!tri_test(ray.o, ray.d, ray.tMax, p0, p1, p2, &t, &b0, &b1, &b2)) return 0.f;
        V2 uv[3];
        tri_uvs(S, T, M, uv);
        V3 du, dv;
        if (!tri_partials(uv, p0, p1, p2, &du, &dv)) return 0.f;
        // of the interaction Triangle::Intersect builds only p and |n . wi| are used: the shading
        // normal and the orientation flips change the sign of n at most (shapes/triangle.cpp:320-415)
        pHit = b0 * p0 + b1 * p1 + b2 * p2;
        nHit = normalize(cross(p0 - p2, p1 - p2));
    } else {
        Isect it;
        if (!sphere_interaction(S, shape, ray, &it)) return 0.f;
        pHit = it.p;
        nHit = it.n;
    }
    float pdf = dist_sq(ref.p, pHit) / (absdot(nHit, -wi) * area);
    if (isinf(pdf)) pdf = 0.f;
    return pdf;
}

// Sphere::Pdf(ref, wi), shapes/sphere.cpp:294-306.
SPTX_DEV float sphere_pdf(const DScene &S, int shape, const RefPoint &ref, V3 wi, float area) {
    const sptx_sphere &sp = S.spheres[shape];
    V3 pCenter = xf_point(sp.o2w, mk(0, 0, 0));
    V3 pOrigin = offset_ray_origin(ref.p, ref.pError, ref.n, pCenter - ref.p);
    if (dist_sq(pOrigin, pCenter) <= sp.radius * sp.radius)
        return shape_pdf_generic(S, SPTX_PRIM_SPHERE, shape, ref, wi, area);
    float sinThetaMax2 = sp.radius * sp.radius / dist_sq(ref.p, pCenter);
    float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
    return 1.f / (2.f * kPi * (1.f - cosThetaMax));
}

// Light::Pdf_Li (DiffuseAreaLight: shape->Pdf; PointLight: 0).
// (ENV = false: compiled for scenes without infinite lights, see shade.cu)
template <bool ENV = true>
SPTX_NOINL float light_pdf_li(const DScene &S, const sptx_light &L, const RefPoint &ref, V3 wi) {
    if (ENV && L.type == SPTX_LIGHT_INFINITE) {
        // InfiniteAreaLight::Pdf_Li, lights/infinite.cpp:125-133; Distribution2D::Pdf, sampling.h:135-142
        const float *base = S.env + L.shape;
        const int W = (int)__ldg(base), H = (int)__ldg(base + 1), nu = (int)__ldg(base + 2), nv = (int)__ldg(base + 3);
        const V3 w = xf_vec(base + 24, wi);
        const float theta = spherical_theta(w), phi = spherical_phi(w);
        const float sinTheta = sinf(theta);
        if (sinTheta == 0) return 0.f;
        const float *marginal = base + 40 + 3 * (size_t)W * H;
        const int iu = min(max((int)(phi * kInv2Pi * (float)nu), 0), nu - 1);
        const int iv = min(max((int)(theta * kInvPi * (float)nv), 0), nv - 1);
        const float *row = marginal + (2 * nv + 2) + (size_t)iv * (2 * nu + 2);
        return (__ldg(row + iu) / __ldg(marginal + 2 * nv + 1)) / (2 * kPi * kPi * sinTheta);
    }
    if (L.type != SPTX_LIGHT_DIFFUSE_AREA) return 0.f;
    if (L.shape_kind == SPTX_PRIM_SPHERE) return sphere_pdf(S, L.shape, ref, wi, L.area);
    return shape_pdf_generic(S, SPTX_PRIM_TRIANGLE, L.shape, ref, wi, L.area);
}

// ---- InfiniteAreaLight (lights/infinite.cpp) ---------------------------------------------------
// Radiance in a direction = the RGB map looked up bilinearly at its finest level
// (MIPMap::Lookup(st, width = 0) -> triangle(0, st), core/mipmap.h:247-287, wrap mode Repeat),
// converted by SampledSpectrum::FromRGB(rgb, Illuminant) (core/spectrum.cpp:135-168): up to three
// of the seven basis spectra with scalar weights, times .86445, clamped at zero.  A radiance value
// is therefore carried as three (row, weight) pairs and evaluated per bin where it is consumed.
struct EnvRgb {
    int row[3];
    float w[3];
};
struct EnvMap {  // view of one light's block in DScene::env
    const float *base;
    int W, H, nu, nv;
    SPTX_DEV const float *l2w() const { return base + 8; }
    SPTX_DEV const float *w2l() const { return base + 24; }
    SPTX_DEV const float *texels() const { return base + 40; }
    SPTX_DEV const float *marginal() const { return texels() + 3 * (size_t)W * H; }
    SPTX_DEV const float *row(int v) const { return marginal() + (2 * nv + 2) + (size_t)v * (2 * nu + 2); }
};
SPTX_DEV EnvMap env_map(const DScene &S, const sptx_light &L) {
    EnvMap m;
    m.base = S.env + L.shape;
    m.W = (int)__ldg(m.base);
    m.H = (int)__ldg(m.base + 1);
    m.nu = (int)__ldg(m.base + 2);
    m.nv = (int)__ldg(m.base + 3);
    return m;
}
SPTX_DEV void env_texel(const EnvMap &m, int s, int t, float rgb[3]) {
    s = mod_i(s, m.W);  // ImageWrap::Repeat
    t = mod_i(t, m.H);
    const float *p = m.texels() + 3 * ((size_t)t * m.W + s);
    rgb[0] = __ldg(p);
    rgb[1] = __ldg(p + 1);
    rgb[2] = __ldg(p + 2);
}
// FromRGB(rgb, Illuminant) as (row, weight) pairs; basis rows: White, Cyan, Magenta, Yellow, Red, Green, Blue.
SPTX_DEV EnvRgb env_from_rgb(const float rgb[3], int basis) {
    enum { W = 0, C = 1, M = 2, Y = 3, R = 4, G = 5, B = 6 };
    EnvRgb e;
    const float r = rgb[0], g = rgb[1], b = rgb[2];
    if (r <= g && r <= b) {
        e.row[0] = W, e.w[0] = r;
        if (g <= b) {
            e.row[1] = C, e.w[1] = g - r;
            e.row[2] = B, e.w[2] = b - g;
        } else {
            e.row[1] = C, e.w[1] = b - r;
            e.row[2] = G, e.w[2] = g - b;
        }
    } else if (g <= r && g <= b) {
        e.row[0] = W, e.w[0] = g;
        if (r <= b) {
            e.row[1] = M, e.w[1] = r - g;
            e.row[2] = B, e.w[2] = b - r;
        } else {
            e.row[1] = M, e.w[1] = b - g;
            e.row[2] = R, e.w[2] = r - b;
        }
    } else {
        e.row[0] = W, e.w[0] = b;
        if (r <= g) {
            e.row[1] = Y, e.w[1] = r - b;
            e.row[2] = G, e.w[2] = g - r;
        } else {
            e.row[1] = Y, e.w[1] = g - b;
            e.row[2] = R, e.w[2] = r - g;
        }
    }
    for (int k = 0; k < 3; ++k) e.row[k] += basis;
    return e;
}
// One bin of FromRGB: r = 0 + w0 B0; r += w1 B1; r += w2 B2; r *= .86445f; Clamp(0, inf).
SPTX_DEV float env_bin1(float b0, float b1, float b2, const EnvRgb &e) {
    float r = e.w[0] * b0;
    r = r + e.w[1] * b1;
    r = r + e.w[2] * b2;
    r = r * .86445f;
    return (r < 0.f) ? 0.f : r;  // Clamp(val, 0, Infinity)
}
SPTX_DEV float env_bin(const DScene &S, const EnvRgb &e, int b) {
    return env_bin1(spec(S, e.row[0], b), spec(S, e.row[1], b), spec(S, e.row[2], b), e);
}
SPTX_DEV float4 env_bin4(const DScene &S, const EnvRgb &e, int c) {
    const float4 b0 = spec4(S, e.row[0], c), b1 = spec4(S, e.row[1], c), b2 = spec4(S, e.row[2], c);
    return make_float4(env_bin1(b0.x, b1.x, b2.x, e), env_bin1(b0.y, b1.y, b2.y, e), env_bin1(b0.z, b1.z, b2.z, e),
                       c < 7 ? env_bin1(b0.w, b1.w, b2.w, e) : 0.f);
}
// Lmap->Lookup(st) with width 0: bilinear at level 0.
SPTX_DEV EnvRgb env_lookup(const DScene &S, const sptx_light &L, const EnvMap &m, float s01, float t01) {
    const float s = s01 * (float)m.W - 0.5f, t = t01 * (float)m.H - 0.5f;
    const float fs = floorf(s), ft = floorf(t);
    const int s0 = (int)fs, t0 = (int)ft;
    const float ds = s - (float)s0, dt = t - (float)t0;
    float a[3], b[3], c[3], d[3], rgb[3];
    env_texel(m, s0, t0, a);
    env_texel(m, s0, t0 + 1, b);
    env_texel(m, s0 + 1, t0, c);
    env_texel(m, s0 + 1, t0 + 1, d);
    const float wa = (1 - ds) * (1 - dt), wb = (1 - ds) * dt, wc = ds * (1 - dt), wd = ds * dt;
    for (int k = 0; k < 3; ++k) rgb[k] = wa * a[k] + wb * b[k] + wc * c[k] + wd * d[k];
    return env_from_rgb(rgb, L.spectrum);
}
// InfiniteAreaLight::Le(ray), lights/infinite.cpp:92-96.
SPTX_DEV EnvRgb env_le(const DScene &S, const sptx_light &L, V3 dir) {
    const EnvMap m = env_map(S, L);
    const V3 w = normalize(xf_vec(m.w2l(), dir));
    return env_lookup(S, L, m, spherical_phi(w) * kInv2Pi, spherical_theta(w) * kInvPi);
}
// Distribution1D::SampleContinuous, core/sampling.h:72-89, on a {func[n], cdf[n+1], funcInt} block.
SPTX_DEV float dist1d_sample(const float *d, int n, float u, float *pdf, int *off) {
    const float *func = d, *cdf = d + n;
    int first = 0, len = n + 1;  // FindInterval(cdf.size(), cdf[i] <= u)
    while (len > 0) {
        int half = len >> 1, middle = first + half;
        if (__ldg(cdf + middle) <= u) {
            first = middle + 1;
            len -= half + 1;
        } else
            len = half;
    }
    const int offset = min(max(first - 1, 0), n - 1);
    if (off) *off = offset;
    float du = u - __ldg(cdf + offset);
    const float span = __ldg(cdf + offset + 1) - __ldg(cdf + offset);
    if (span > 0) du /= span;
    const float funcInt = __ldg(d + 2 * n + 1);
    if (pdf) *pdf = (funcInt > 0) ? __ldg(func + offset) / funcInt : 0.f;
    return ((float)offset + du) / (float)n;
}

// Result of Light::Sample_Li.  The radiance is row `spectrum` scaled per bin:
// area lights: Lemit[b] (emit != 0) ; point lights: I[b] / dist2.
struct LiSample {
    V3 wi;
    float pdf;
    bool black;     // Li is black (or pdf == 0)
    int spectrum;   // row
    float divisor;  // 0: Li[b] = row[b]; else Li[b] = row[b] / divisor
    LightPoint lp;  // sampled point (shadow-ray target)
    bool isEnv;     // infinite light: Li = env (three basis rows with weights) instead of a row
    EnvRgb env;
};

SPTX_DEV float li_bin(const DScene &S, const LiSample &ls, int b) {
    if (ls.isEnv) return env_bin(S, ls.env, b);
    float v = spec(S, ls.spectrum, b);
    return ls.divisor != 0.f ? v / ls.divisor : v;
}

// Four consecutive bins of the sampled radiance (the thread-per-path form of li_bin).
template <bool ENV = true>
SPTX_DEV float4 li4(const DScene &S, const LiSample &ls, int c) {
    if (ENV && ls.isEnv) return env_bin4(S, ls.env, c);
    float4 li = spec4(S, ls.spectrum, c);
    if (ls.divisor != 0.f) {
        li.x = li.x / ls.divisor;
        li.y = li.y / ls.divisor;
        li.z = li.z / ls.divisor;
        li.w = padless(c, li.w, ls.divisor);
    }
    return li;
}

template <bool ENV = true>
SPTX_NOINL LiSample light_sample_li(const DScene &S, const sptx_light &L, const RefPoint &ref, V2 u) {
    LiSample r;
    r.spectrum = L.spectrum;
    r.divisor = 0.f;
    r.black = true;
    r.pdf = 0.f;
    r.wi = mk(0, 0, 0);
    if (ENV) r.isEnv = false;
    if (ENV && L.type == SPTX_LIGHT_INFINITE) {
        // InfiniteAreaLight::Sample_Li, lights/infinite.cpp:98-123
        const EnvMap m = env_map(S, L);
        float pdfs[2];
        int v;
        const float d1 = dist1d_sample(m.marginal(), m.nv, u.y, &pdfs[1], &v);
        const float d0 = dist1d_sample(m.row(v), m.nu, u.x, &pdfs[0], nullptr);
        const float mapPdf = pdfs[0] * pdfs[1];
        if (mapPdf == 0) return r;
        const float theta = d1 * kPi, phi = d0 * 2 * kPi;
        const float cosTheta = cosf(theta), sinTheta = sinf(theta);
        const float sinPhi = sinf(phi), cosPhi = cosf(phi);
        r.wi = xf_vec(m.l2w(), mk(sinTheta * cosPhi, sinTheta * sinPhi, cosTheta));
        r.pdf = mapPdf / (2 * kPi * kPi * sinTheta);
        if (sinTheta == 0) r.pdf = 0.f;
        // VisibilityTester(ref, Interaction(ref.p + wi * (2 * worldRadius)))
        r.lp.p = ref.p + r.wi * (2 * L.area);
        r.lp.pError = mk(0, 0, 0);
        r.lp.n = mk(0, 0, 0);
        r.isEnv = true;
        r.env = env_lookup(S, L, m, d0, d1);
        r.black = false;  // blackness is checked per bin by the caller
        return r;
    }
    if (L.type == SPTX_LIGHT_POINT) {
        V3 pl = mk(L.pos[0], L.pos[1], L.pos[2]);
        r.wi = normalize(pl - ref.p);
        r.pdf = 1.f;
        r.lp.p = pl;
        r.lp.pError = mk(0, 0, 0);
        r.lp.n = mk(0, 0, 0);
        r.divisor = dist_sq(pl, ref.p);
        r.black = false;  // blackness of I/d^2 is checked per bin by the caller
        return r;
    }
    float pdf;
    LightPoint lp;
    if (L.shape_kind == SPTX_PRIM_SPHERE)
        lp = sphere_sample(S.spheres[L.shape], ref, u, L.area, &pdf);
    else {
        lp = triangle_sample(S, L.shape, u, L.area, &pdf);
        area_to_solid_angle(ref, lp, &pdf, false);
    }
    r.lp = lp;
    if (pdf == 0 || length_sq(lp.p - ref.p) == 0) return r;  // pdf = 0, black
    r.wi = normalize(lp.p - ref.p);
    r.pdf = pdf;
    // DiffuseAreaLight::L(pShape, -wi), lights/diffuse.h:56-58
    r.black = !(L.two_sided || dot(lp.n, -r.wi) > 0);
    return r;
}

// Distribution1D::SampleDiscrete, core/sampling.h:90-100, on func[n] cdf[n+1].
SPTX_DEV int sample_discrete(const float *func, const float *cdf, int n, float funcInt, float u,
                             float *pdf) {
    // FindInterval(size = n + 1, cdf[i] <= u), core/pbrt.h:398-416
    int size = n + 1;
    int first = 0, len = size;
    while (len > 0) {
        int half = len >> 1, middle = first + half;
        if (__ldg(cdf + middle) <= u) {
            first = middle + 1;
            len -= half + 1;
        } else
            len = half;
    }
    int offset = min(max(first - 1, 0), size - 2);
    *pdf = (funcInt > 0) ? __ldg(func + offset) / (funcInt * (float)n) : 0.f;
    return offset;
}

// LightDistribution::Lookup + Distribution1D::SampleDiscrete.
SPTX_DEV int select_light(const DScene &S, V3 p, float u, float *pdf) {
    const int n = S.n_lights;
    if (S.light_strategy != SPTX_LIGHTS_SPATIAL)
        return sample_discrete(S.light_func, S.light_func + n, n, S.light_func_int, u, pdf);
    // SpatialLightDistribution::Lookup, core/lightdistrib.cpp:135-147
    V3 pmin = mk(S.wb_min[0], S.wb_min[1], S.wb_min[2]);
    V3 pmax = mk(S.wb_max[0], S.wb_max[1], S.wb_max[2]);
    V3 o = p - pmin;  // Bounds3::Offset
    if (pmax.x > pmin.x) o.x /= pmax.x - pmin.x;
    if (pmax.y > pmin.y) o.y /= pmax.y - pmin.y;
    if (pmax.z > pmin.z) o.z /= pmax.z - pmin.z;
    int pi0 = min(max((int)(o.x * (float)S.n_voxels[0]), 0), S.n_voxels[0] - 1);
    int pi1 = min(max((int)(o.y * (float)S.n_voxels[1]), 0), S.n_voxels[1] - 1);
    int pi2 = min(max((int)(o.z * (float)S.n_voxels[2]), 0), S.n_voxels[2] - 1);
    size_t voxel = ((size_t)pi0 * S.n_voxels[1] + pi1) * S.n_voxels[2] + pi2;
    if (S.voxel_rows) {
        // sparse tables (many lights): rows exist for the voxels that hold a surface.  A point in a
        // voxel without one (none should: the marking is conservative) picks a light uniformly, which
        // is still an unbiased choice.
        const int r = __ldg(S.voxel_rows + voxel);
        if (r < 0) {
            *pdf = 1.f / (float)n;
            return min((int)(u * (float)n), n - 1);
        }
        voxel = (size_t)r;
    }
    const float *row = S.voxel_tables + voxel * (size_t)(2 * n + 2);
    return sample_discrete(row, row + n, n, __ldg(row + 2 * n + 1), u, pdf);
}

SPTX_DEV float power_heuristic(float fPdf, float gPdf) {
    float f = 1.f * fPdf, g = 1.f * gPdf;
    return (f * f) / (f * f + g * g);
}

}  // namespace sptx
