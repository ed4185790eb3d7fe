// tex.cuh -- ImageTexture<RGBSpectrum, Spectrum>::Evaluate on the device.
//
// textures/imagemap.h:89-101 (Evaluate), core/texture.cpp:93-99 (UVMapping2D::Map),
// core/mipmap.h:222-365 (Texel, Lookup, triangle, EWA), core/interaction.cpp:103-149
// (SurfaceInteraction::ComputeDifferentials), core/spectrum.cpp:135-168 (FromRGB; convertOut passes no
// SpectrumType, so it is the Illuminant expansion, core/spectrum.h:428-429).
// The texture's MIP pyramid is the one the reference built on the host (resampled to a power of
// two, gamma / scale applied); it arrives as a block of DScene::tex (layout: include/sptx.h).
// The value of the texture at a hit is carried like an infinite light's radiance: three
// (basis row, weight) pairs (EnvRgb) that the per-bin loops expand (tex_bin4, bsdf.cuh).
#pragma once
#include "interaction.cuh"
#include "light.cuh"

namespace sptx {

struct TexMap {
    const float *base;
    int levels, wrap, trilinear;
    float maxAniso;
    float su, sv, du, dv;
    int basis;
    const float *lut;
    SPTX_DEV int geti(int k) const { return __float_as_int(__ldg(base + k)); }
    SPTX_DEV int usize(int level) const { return geti(16 + 3 * level); }
    SPTX_DEV int vsize(int level) const { return geti(16 + 3 * level + 1); }
    SPTX_DEV const float *texels(int level) const { return base + geti(16 + 3 * level + 2); }
};
SPTX_DEV TexMap tex_map(const DScene &S, int offset) {
    TexMap m;
    m.base = S.tex + offset;
    m.levels = m.geti(0);
    m.wrap = m.geti(1);
    m.trilinear = m.geti(2);
    m.maxAniso = __ldg(m.base + 3);
    m.su = __ldg(m.base + 4);
    m.sv = __ldg(m.base + 5);
    m.du = __ldg(m.base + 6);
    m.dv = __ldg(m.base + 7);
    m.basis = m.geti(8);
    m.lut = S.tex + m.geti(9);
    return m;
}

struct Rgb {
    float r, g, b;
};
SPTX_DEV Rgb operator*(float a, Rgb v) { return Rgb{v.r * a, v.g * a, v.b * a}; }  // Spectrum * Float
SPTX_DEV Rgb operator+(Rgb a, Rgb b) { return Rgb{a.r + b.r, a.g + b.g, a.b + b.b}; }

// MIPMap::Texel, core/mipmap.h:222-244.
SPTX_DEV Rgb tex_texel(const TexMap &m, int level, int s, int t) {
    const int us = m.usize(level), vs = m.vsize(level);
    if (m.wrap == 0) {  // Repeat
        s = mod_i(s, us);
        t = mod_i(t, vs);
    } else if (m.wrap == 2) {  // Clamp
        s = min(max(s, 0), us - 1);
        t = min(max(t, 0), vs - 1);
    } else if (s < 0 || s >= us || t < 0 || t >= vs)  // Black
        return Rgb{0.f, 0.f, 0.f};
    const float *p = m.texels(level) + 3 * ((size_t)t * us + s);
    return Rgb{__ldg(p), __ldg(p + 1), __ldg(p + 2)};
}
// MIPMap::triangle, core/mipmap.h:276-287.
SPTX_DEV Rgb tex_triangle(const TexMap &m, int level, V2 st) {
    level = min(max(level, 0), m.levels - 1);
    const float s = st.x * (float)m.usize(level) - 0.5f, t = st.y * (float)m.vsize(level) - 0.5f;
    const int s0 = (int)floorf(s), t0 = (int)floorf(t);
    const float ds = s - (float)s0, dt = t - (float)t0;
    return (1 - ds) * (1 - dt) * tex_texel(m, level, s0, t0) + (1 - ds) * dt * tex_texel(m, level, s0, t0 + 1) +
           ds * (1 - dt) * tex_texel(m, level, s0 + 1, t0) + ds * dt * tex_texel(m, level, s0 + 1, t0 + 1);
}
SPTX_DEV Rgb tex_lerp(float t, Rgb a, Rgb b) { return (1 - t) * a + t * b; }
SPTX_DEV float tex_log2(float x) { return logf(x) * 1.442695040888963387004650940071f; }
// MIPMap::Lookup(st, width), core/mipmap.h:246-274 (filtered textures only).
SPTX_DEV Rgb tex_lookup_width(const TexMap &m, V2 st, float width) {
    const float level = (float)(m.levels - 1) + tex_log2(fmaxf(width, 1e-8f));
    if (level < 0) return tex_triangle(m, 0, st);
    if (level >= (float)(m.levels - 1)) return tex_texel(m, m.levels - 1, 0, 0);
    const int iLevel = (int)floorf(level);
    const float delta = level - (float)iLevel;
    return tex_lerp(delta, tex_triangle(m, iLevel, st), tex_triangle(m, iLevel + 1, st));
}
// MIPMap::EWA, core/mipmap.h:318-362.
SPTX_NOINL Rgb tex_ewa(const TexMap &m, int level, V2 st, V2 dst0, V2 dst1) {
    if (level >= m.levels) return tex_texel(m, m.levels - 1, 0, 0);
    const float us = (float)m.usize(level), vs = (float)m.vsize(level);
    st.x = st.x * us - 0.5f;
    st.y = st.y * vs - 0.5f;
    dst0.x *= us;
    dst0.y *= vs;
    dst1.x *= us;
    dst1.y *= vs;
    float A = dst0.y * dst0.y + dst1.y * dst1.y + 1;
    float B = -2 * (dst0.x * dst0.y + dst1.x * dst1.y);
    float C = dst0.x * dst0.x + dst1.x * dst1.x + 1;
    const float invF = 1 / (A * C - B * B * 0.25f);
    A *= invF;
    B *= invF;
    C *= invF;
    const float det = -B * B + 4 * A * C;
    const float invDet = 1 / det;
    const float uSqrt = sqrtf(det * C), vSqrt = sqrtf(A * det);
    const int s0 = (int)ceilf(st.x - 2 * invDet * uSqrt), s1 = (int)floorf(st.x + 2 * invDet * uSqrt);
    const int t0 = (int)ceilf(st.y - 2 * invDet * vSqrt), t1 = (int)floorf(st.y + 2 * invDet * vSqrt);
    Rgb sum{0.f, 0.f, 0.f};
    float sumWts = 0;
    for (int it = t0; it <= t1; ++it) {
        const float tt = (float)it - st.y;
        for (int is = s0; is <= s1; ++is) {
            const float ss = (float)is - st.x;
            const float r2 = A * ss * ss + B * ss * tt + C * tt * tt;
            if (r2 < 1) {
                const int index = min((int)(r2 * 128.f), 127);  // WeightLUTSize = 128
                const float weight = __ldg(m.lut + index);
                sum = sum + weight * tex_texel(m, level, is, it);
                sumWts += weight;
            }
        }
    }
    return Rgb{sum.r / sumWts, sum.g / sumWts, sum.b / sumWts};
}
// MIPMap::Lookup(st, dst0, dst1), core/mipmap.h:289-316.
SPTX_DEV Rgb tex_lookup(const TexMap &m, V2 st, V2 dst0, V2 dst1) {
    if (m.trilinear) {
        const float width = fmaxf(fmaxf(fabsf(dst0.x), fabsf(dst0.y)), fmaxf(fabsf(dst1.x), fabsf(dst1.y)));
        return tex_lookup_width(m, st, 2 * width);
    }
    if (dst0.x * dst0.x + dst0.y * dst0.y < dst1.x * dst1.x + dst1.y * dst1.y) {
        const V2 t = dst0;
        dst0 = dst1;
        dst1 = t;
    }
    const float majorLength = sqrtf(dst0.x * dst0.x + dst0.y * dst0.y);
    float minorLength = sqrtf(dst1.x * dst1.x + dst1.y * dst1.y);
    if (minorLength * m.maxAniso < majorLength && minorLength > 0) {
        const float scale = majorLength / (minorLength * m.maxAniso);
        dst1.x *= scale;
        dst1.y *= scale;
        minorLength *= scale;
    }
    if (minorLength == 0) return tex_triangle(m, 0, st);
    const float lod = fmaxf(0.f, (float)m.levels - 1.f + tex_log2(minorLength));
    const int ilod = (int)floorf(lod);
    return tex_lerp(lod - (float)ilod, tex_ewa(m, ilod, st, dst0, dst1), tex_ewa(m, ilod + 1, st, dst0, dst1));
}

// SurfaceInteraction::ComputeDifferentials, core/interaction.cpp:103-149: (du/dx, dv/dx), (du/dy, dv/dy)
// from the camera ray's offset rays (rd == NULL: the ray has none -- every ray but the camera ray).
SPTX_DEV bool solve2x2(const float A[2][2], const float B[2], float *x0, float *x1) {
    const float det = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    if (fabsf(det) < 1e-10f) return false;
    *x0 = (A[1][1] * B[0] - A[0][1] * B[1]) / det;
    *x1 = (A[0][0] * B[1] - A[1][0] * B[0]) / det;
    if (isnan(*x0) || isnan(*x1)) return false;
    return true;
}
SPTX_DEV void uv_differentials(const Isect &it, const SurfParam &par, const RayDiff *rd, V2 *ddx, V2 *ddy) {
    *ddx = V2{0.f, 0.f};
    *ddy = V2{0.f, 0.f};
    if (!rd) return;
    const V3 n = it.n, p = it.p;
    const float d = dot(n, p);
    const float tx = -(dot(n, rd->rxo) - d) / dot(n, rd->rxd);
    if (isinf(tx) || isnan(tx)) return;
    const V3 px = rd->rxo + tx * rd->rxd;
    const float ty = -(dot(n, rd->ryo) - d) / dot(n, rd->ryd);
    if (isinf(ty) || isnan(ty)) return;
    const V3 py = rd->ryo + ty * rd->ryd;
    int d0, d1;
    if (fabsf(n.x) > fabsf(n.y) && fabsf(n.x) > fabsf(n.z)) {
        d0 = 1;
        d1 = 2;
    } else if (fabsf(n.y) > fabsf(n.z)) {
        d0 = 0;
        d1 = 2;
    } else {
        d0 = 0;
        d1 = 1;
    }
    const float A[2][2] = {{comp(par.dpdu, d0), comp(par.dpdv, d0)}, {comp(par.dpdu, d1), comp(par.dpdv, d1)}};
    const float Bx[2] = {comp(px, d0) - comp(p, d0), comp(px, d1) - comp(p, d1)};
    const float By[2] = {comp(py, d0) - comp(p, d0), comp(py, d1) - comp(p, d1)};
    if (!solve2x2(A, Bx, &ddx->x, &ddx->y)) *ddx = V2{0.f, 0.f};
    if (!solve2x2(A, By, &ddy->x, &ddy->y)) *ddy = V2{0.f, 0.f};
}

// RayDifferential::ScaleDifferentials, core/geometry.h:917-922.
SPTX_DEV void scale_differentials(const Ray &ray, float s, RayDiff *rd) {
    rd->rxo = ray.o + (rd->rxo - ray.o) * s;
    rd->ryo = ray.o + (rd->ryo - ray.o) * s;
    rd->rxd = ray.d + (rd->rxd - ray.d) * s;
    rd->ryd = ray.d + (rd->ryd - ray.d) * s;
}

// Kd->Evaluate(si) of an image texture at the hit: FromRGB(rgb) as (row, weight) pairs.
SPTX_NOINL EnvRgb tex_evaluate(const DScene &S, int offset, const Isect &it, const SurfParam &par,
                                const RayDiff *rd) {
    const TexMap m = tex_map(S, offset);
    V2 ddx, ddy;
    uv_differentials(it, par, rd, &ddx, &ddy);
    const V2 dstdx{m.su * ddx.x, m.sv * ddx.y}, dstdy{m.su * ddy.x, m.sv * ddy.y};
    const V2 st{m.su * par.uv.x + m.du, m.sv * par.uv.y + m.dv};
    const Rgb v = tex_lookup(m, st, dstdx, dstdy);
    const float rgb[3] = {v.r, v.g, v.b};
    return env_from_rgb(rgb, m.basis);
}
// Spectrum::IsBlack() of the texture value (materials add the diffuse lobe only for a non-black Kd).
SPTX_DEV bool tex_is_black(const DScene &S, const EnvRgb &e) {
    bool non = false;
    for (int c = 0; c < 8; ++c) {
        const float4 v = tex_bin4(S, e.row[0], e.row[1], e.row[2], e.w[0], e.w[1], e.w[2], c);
        non |= v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f;
    }
    return !non;
}

}  // namespace sptx
