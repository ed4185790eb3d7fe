// bsdf.cuh -- BSDF::f / Sample_f / Pdf over 31-bin spectra.
//
// A material's BSDF is a short list of lobes (sptx_lobe).  Every lobe value is
// "spectrum row x a few scalars", so instead of materialising 31-float arrays
// per thread a BSDF evaluation returns an FSum: up to kMaxTerms terms, each a
// formula id + spectrum rows + scalars.  Consumers stream over the 31 bins and
// evaluate fsum_bin(), which reproduces the reference's per-bin operation
// order (core/reflection.cpp, core/microfacet.cpp).
#pragma once
#include "camera.cuh"  // refract()
#include "interaction.cuh"

namespace sptx {

enum {
    BSDF_REFLECTION = 1 << 0,
    BSDF_TRANSMISSION = 1 << 1,
    BSDF_DIFFUSE = 1 << 2,
    BSDF_GLOSSY = 1 << 3,
    BSDF_SPECULAR = 1 << 4,
    BSDF_ALL = 31
};

constexpr int kMaxTerms = 4;

enum {
    TERM_R_S = 0,        // R[b] * s0                       (Lambert: s0 = InvPi)
    TERM_R_S_S = 1,      // R[b] * s0 * s1                  (Oren-Nayar)
    TERM_MICRO_S = 2,    // R[b] * s0 * s1 * s2 / s3        (microfacet, scalar Fresnel s2)
    TERM_MICRO_COND = 3, // R[b] * s0 * s1 * FrCond[b] / s3 (microfacet, conductor, cos = s2)
    TERM_S_R_DIV = 4,    // s0 * R[b] / s1                  (specular reflection / FresnelSpecular R)
    TERM_COND_R_DIV = 5, // FrCond[b](s0) * R[b] / s1       (specular reflection, conductor)
    TERM_T_S_S_DIV = 6,  // (T[b] * s0) * s1 / s2           (specular transmission)
    TERM_BLEND = 7,      // FresnelBlend::f: Rd = R, Rs = e0; diffuse and specular scalars s0..s3
    TERM_TEX_S = 8,      // Tex[b] * s0, Tex = an image texture's value at the hit: FromRGB over the
                         // basis rows e0, e1, e2 with weights s1, s2, s3 (tex.cuh); Lambert
};

// FresnelBlend::f for one bin, core/reflection.cpp:285-299 (+ SchlickFresnel, reflection.h):
// s0, s1 = 1 - pow5(1 - |cos|/2) for wi and wo, s2 = pow5(1 - wi.wh), s3 = D / (4 |wi.wh| max|cos|).
SPTX_DEV float blend_bin(float rd, float rs, float s0, float s1, float s2, float s3) {
    const float kd = 28.f / (23.f * kPi);
    const float diffuse = rd * kd * (1.f - rs) * s0 * s1;
    const float schlick = rs + s2 * (1.f - rs);
    return diffuse + schlick * s3;
}

struct FTerm {
    int kind;
    int R;
    int e0, e1, e2;  // conductor rows: etaI, etaT, k
    float s0, s1, s2, s3;
};
template <int MAXT>
struct FSumN {
    int n;
    FTerm t[MAXT];
};
using FSum = FSumN<kMaxTerms>;

// A shading variant (K3 is compiled once per variant, see shade.cuh) is the set of lobe types its
// code can evaluate: branches for other types are removed at compile time.
constexpr int kAllLobes = (1 << (SPTX_LOBE_FRESNEL_BLEND + 1)) - 1;
// not a lobe type: the variant's materials may have an image texture as their Lambert lobe's R
constexpr int kTexturedLobes = 1 << 16;
template <int LOBES>
SPTX_DEV constexpr bool has_lobe(int type) { return ((LOBES >> type) & 1) != 0; }
constexpr int kSpecularLobes = (1 << SPTX_LOBE_SPECULAR_REFL) | (1 << SPTX_LOBE_SPECULAR_TRANS) |
                               (1 << SPTX_LOBE_FRESNEL_SPECULAR);
constexpr int kMicrofacetLobes = (1 << SPTX_LOBE_MICROFACET_REFL) | (1 << SPTX_LOBE_MICROFACET_TRANS) |
                                 (1 << SPTX_LOBE_FRESNEL_BLEND);

SPTX_DEV float spec(const DScene &S, int row, int b) {
    return __ldg(S.spectra + (size_t)row * SPTX_SPECTRUM_STRIDE + b);
}

// FrDielectric, core/reflection.cpp:47-68.
SPTX_NOINL float fr_dielectric(float cosThetaI, float etaI, float etaT) {
    cosThetaI = clampf(cosThetaI, -1.f, 1.f);
    bool entering = cosThetaI > 0.f;
    if (!entering) {
        float t = etaI;
        etaI = etaT;
        etaT = t;
        cosThetaI = fabsf(cosThetaI);
    }
    float sinThetaI = sqrtf(fmaxf(0.f, 1.f - cosThetaI * cosThetaI));
    float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1) return 1.f;
    float cosThetaT = sqrtf(fmaxf(0.f, 1.f - sinThetaT * sinThetaT));
    float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (Rparl * Rparl + Rperp * Rperp) / 2.f;
}

// FrConductor for one bin, core/reflection.cpp:71-94.
SPTX_DEV float fr_conductor_bin(float cosThetaI, float etai, float etat, float k) {
    cosThetaI = clampf(cosThetaI, -1.f, 1.f);
    float eta = etat / etai;
    float etak = k / etai;
    float cosThetaI2 = cosThetaI * cosThetaI;
    float sinThetaI2 = (float)(1. - (double)cosThetaI2);
    float eta2 = eta * eta;
    float etak2 = etak * etak;
    float t0 = eta2 - etak2 - sinThetaI2;
    float a2plusb2 = sqrtf(t0 * t0 + 4.f * eta2 * etak2);
    float t1 = a2plusb2 + cosThetaI2;
    float a = sqrtf(0.5f * (a2plusb2 + t0));
    float t2 = 2.f * cosThetaI * a;
    float Rs = (t1 - t2) / (t1 + t2);
    float t3 = cosThetaI2 * a2plusb2 + sinThetaI2 * sinThetaI2;
    float t4 = t2 * sinThetaI2;
    float Rp = Rs * (t3 - t4) / (t3 + t4);
    return 0.5f * (Rp + Rs);
}

SPTX_DEV float fterm_bin(const DScene &S, const FTerm &t, int b) {
    switch (t.kind) {
    case TERM_R_S:
        return spec(S, t.R, b) * t.s0;
    case TERM_R_S_S:
        return spec(S, t.R, b) * t.s0 * t.s1;
    case TERM_MICRO_S:
        return spec(S, t.R, b) * t.s0 * t.s1 * t.s2 / t.s3;
    case TERM_MICRO_COND: {
        float F = fr_conductor_bin(t.s2, spec(S, t.e0, b), spec(S, t.e1, b), spec(S, t.e2, b));
        return spec(S, t.R, b) * t.s0 * t.s1 * F / t.s3;
    }
    case TERM_S_R_DIV:
        return t.s0 * spec(S, t.R, b) / t.s1;
    case TERM_COND_R_DIV: {
        float F = fr_conductor_bin(t.s0, spec(S, t.e0, b), spec(S, t.e1, b), spec(S, t.e2, b));
        return F * spec(S, t.R, b) / t.s1;
    }
    case TERM_BLEND:
        return blend_bin(spec(S, t.R, b), spec(S, t.e0, b), t.s0, t.s1, t.s2, t.s3);
    default:  // TERM_T_S_S_DIV
        return (spec(S, t.R, b) * t.s0) * t.s1 / t.s2;
    }
}
// Sum of terms in lobe order, starting from Spectrum(0) as BSDF::f does.
template <int MAXT>
SPTX_DEV float fsum_bin(const DScene &S, const FSumN<MAXT> &f, int b) {
    float v = 0.f;
    for (int i = 0; i < f.n; ++i) v += fterm_bin(S, f.t[i], b);
    return v;
}

// Four consecutive bins (4c .. 4c+3) of a term / of the sum: the thread-per-path
// form.  Same per-bin operation order as fterm_bin.
SPTX_DEV float4 spec4(const DScene &S, int row, int c) {
    return __ldg(reinterpret_cast<const float4 *>(S.spectra + (size_t)row * SPTX_SPECTRUM_STRIDE) + c);
}
// x / d for the .w word of chunk c: the last chunk's .w is the row's pad word (always 0), and
// 0 / d (or a conductor Fresnel term of an all-zero bin) would run the IEEE division and
// square-root slow paths for nothing
SPTX_DEV float padless(int c, float x, float d) { return c < 7 ? x / d : 0.f; }
// (term kinds a variant can meet: R_S Lambert / Lambert transmission; R_S_S Oren-Nayar /
//  microfacet transmission; MICRO_* microfacet reflection; *_DIV the specular lobes; BLEND)
// One bin of an image texture's value: ImageTexture::convertOut calls Spectrum::FromRGB(rgb) WITHOUT a
// SpectrumType, and the default is Illuminant (core/spectrum.h:428-429, textures/imagemap.h:113-117):
// r = 0 + w0 B0; r += w1 B1; r += w2 B2 over the rgbIllum2Spect* rows, r *= .86445f, Clamp(0, inf)
// (core/spectrum.cpp:135-168, :250) -- the same expansion as an infinite light's map (light.cuh).
SPTX_DEV float tex_bin1(float b0, float b1, float b2, float w0, float w1, float w2) {
    float r = w0 * b0;
    r = r + w1 * b1;
    r = r + w2 * b2;
    r = r * .86445f;
    return (r < 0.f) ? 0.f : r;
}
SPTX_DEV float4 tex_bin4(const DScene &S, int r0, int r1, int r2, float w0, float w1, float w2, int c) {
    const float4 b0 = spec4(S, r0, c), b1 = spec4(S, r1, c), b2 = spec4(S, r2, c);
    return make_float4(tex_bin1(b0.x, b1.x, b2.x, w0, w1, w2), tex_bin1(b0.y, b1.y, b2.y, w0, w1, w2),
                       tex_bin1(b0.z, b1.z, b2.z, w0, w1, w2), c < 7 ? tex_bin1(b0.w, b1.w, b2.w, w0, w1, w2) : 0.f);
}
template <int LOBES>
SPTX_DEV float4 fterm4(const DScene &S, const FTerm &t, int c) {
    if ((LOBES & kTexturedLobes) && t.kind == TERM_TEX_S) {
        const float4 R = tex_bin4(S, t.e0, t.e1, t.e2, t.s1, t.s2, t.s3, c);
        return make_float4(R.x * t.s0, R.y * t.s0, R.z * t.s0, R.w * t.s0);
    }
    const float4 R = spec4(S, t.R, c);
    constexpr bool kRS = has_lobe<LOBES>(SPTX_LOBE_LAMBERT) || has_lobe<LOBES>(SPTX_LOBE_LAMBERT_TRANS);
    constexpr bool kRSS = has_lobe<LOBES>(SPTX_LOBE_OREN_NAYAR) || has_lobe<LOBES>(SPTX_LOBE_MICROFACET_TRANS);
    constexpr bool kMicro = has_lobe<LOBES>(SPTX_LOBE_MICROFACET_REFL);
    constexpr bool kSpecR = has_lobe<LOBES>(SPTX_LOBE_SPECULAR_REFL) || has_lobe<LOBES>(SPTX_LOBE_FRESNEL_SPECULAR);
    constexpr bool kSpecRC = has_lobe<LOBES>(SPTX_LOBE_SPECULAR_REFL);
    constexpr bool kSpecT = has_lobe<LOBES>(SPTX_LOBE_SPECULAR_TRANS) || has_lobe<LOBES>(SPTX_LOBE_FRESNEL_SPECULAR);
    constexpr bool kBlend = has_lobe<LOBES>(SPTX_LOBE_FRESNEL_BLEND);
    if (kRS && (t.kind == TERM_R_S || !(kRSS || kMicro || kSpecR || kSpecT || kBlend)))
        return make_float4(R.x * t.s0, R.y * t.s0, R.z * t.s0, R.w * t.s0);
    if (kRSS && t.kind == TERM_R_S_S)
        return make_float4(R.x * t.s0 * t.s1, R.y * t.s0 * t.s1, R.z * t.s0 * t.s1, R.w * t.s0 * t.s1);
    if (kMicro && t.kind == TERM_MICRO_S)
        return make_float4(R.x * t.s0 * t.s1 * t.s2 / t.s3, R.y * t.s0 * t.s1 * t.s2 / t.s3,
                           R.z * t.s0 * t.s1 * t.s2 / t.s3, padless(c, R.w * t.s0 * t.s1 * t.s2, t.s3));
    if (kMicro && t.kind == TERM_MICRO_COND) {
        const float4 ei = spec4(S, t.e0, c), et = spec4(S, t.e1, c), k = spec4(S, t.e2, c);
        return make_float4(R.x * t.s0 * t.s1 * fr_conductor_bin(t.s2, ei.x, et.x, k.x) / t.s3,
                           R.y * t.s0 * t.s1 * fr_conductor_bin(t.s2, ei.y, et.y, k.y) / t.s3,
                           R.z * t.s0 * t.s1 * fr_conductor_bin(t.s2, ei.z, et.z, k.z) / t.s3,
                           c < 7 ? R.w * t.s0 * t.s1 * fr_conductor_bin(t.s2, ei.w, et.w, k.w) / t.s3 : 0.f);
    }
    if (kSpecR && t.kind == TERM_S_R_DIV)
        return make_float4(t.s0 * R.x / t.s1, t.s0 * R.y / t.s1, t.s0 * R.z / t.s1,
                           padless(c, t.s0 * R.w, t.s1));
    if (kSpecRC && t.kind == TERM_COND_R_DIV) {
        const float4 ei = spec4(S, t.e0, c), et = spec4(S, t.e1, c), k = spec4(S, t.e2, c);
        return make_float4(fr_conductor_bin(t.s0, ei.x, et.x, k.x) * R.x / t.s1,
                           fr_conductor_bin(t.s0, ei.y, et.y, k.y) * R.y / t.s1,
                           fr_conductor_bin(t.s0, ei.z, et.z, k.z) * R.z / t.s1,
                           c < 7 ? fr_conductor_bin(t.s0, ei.w, et.w, k.w) * R.w / t.s1 : 0.f);
    }
    if (kBlend && t.kind == TERM_BLEND) {
        const float4 rs = spec4(S, t.e0, c);
        return make_float4(blend_bin(R.x, rs.x, t.s0, t.s1, t.s2, t.s3), blend_bin(R.y, rs.y, t.s0, t.s1, t.s2, t.s3),
                           blend_bin(R.z, rs.z, t.s0, t.s1, t.s2, t.s3),
                           c < 7 ? blend_bin(R.w, rs.w, t.s0, t.s1, t.s2, t.s3) : 0.f);
    }
    if (kSpecT)  // TERM_T_S_S_DIV
        return make_float4((R.x * t.s0) * t.s1 / t.s2, (R.y * t.s0) * t.s1 / t.s2,
                           (R.z * t.s0) * t.s1 / t.s2, padless(c, (R.w * t.s0) * t.s1, t.s2));
    return make_float4(0.f, 0.f, 0.f, 0.f);
}
template <int LOBES = kAllLobes, int MAXT = kMaxTerms>
SPTX_DEV float4 fsum4(const DScene &S, const FSumN<MAXT> &f, int c) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < MAXT; ++i) {
        if (i < f.n) {
            const float4 t = fterm4<LOBES>(S, f.t[i], c);
            v.x += t.x;
            v.y += t.y;
            v.z += t.z;
            v.w += t.w;
        }
    }
    return v;
}

// The same as ONE function per kernel instead of a copy per row loop: for the variants whose terms
// carry per-bin Fresnel / microfacet code (three row loops per vertex call it; shade.cuh).
template <int LOBES, int MAXT>
static __device__ __noinline__ float4 fsum4_call(const DScene &S, const FSumN<MAXT> &f, int c) {
    return fsum4<LOBES, MAXT>(S, f, c);
}

// ---- local shading frame trigonometry, core/reflection.h:55-100 ------------
SPTX_DEV float cos_theta(V3 w) { return w.z; }
SPTX_DEV float cos2_theta(V3 w) { return w.z * w.z; }
SPTX_DEV float abs_cos_theta(V3 w) { return fabsf(w.z); }
SPTX_DEV float sin2_theta(V3 w) { return fmaxf(0.f, 1.f - cos2_theta(w)); }
SPTX_DEV float sin_theta(V3 w) { return sqrtf(sin2_theta(w)); }
SPTX_DEV float tan_theta(V3 w) { return sin_theta(w) / cos_theta(w); }
SPTX_DEV float tan2_theta(V3 w) { return sin2_theta(w) / cos2_theta(w); }
SPTX_DEV float cos_phi(V3 w) {
    float s = sin_theta(w);
    return (s == 0) ? 1.f : clampf(w.x / s, -1.f, 1.f);
}
SPTX_DEV float sin_phi(V3 w) {
    float s = sin_theta(w);
    return (s == 0) ? 0.f : clampf(w.y / s, -1.f, 1.f);
}
SPTX_DEV float cos2_phi(V3 w) { return cos_phi(w) * cos_phi(w); }
SPTX_DEV float sin2_phi(V3 w) { return sin_phi(w) * sin_phi(w); }
SPTX_DEV bool same_hemisphere(V3 a, V3 b) { return a.z * b.z > 0; }

// ---- Trowbridge-Reitz, core/microfacet.cpp:155-184, 238-344 -----------------
// Real functions (scalar / by-value arguments, no per-thread structs behind pointers): inlined at every
// use they were a third of the plastic / metal variant's 25 000 instructions, a kernel whose top stall
// was `no_instruction` (profiles/r02zz_step_ncu.txt).
#define SPTX_CALL static __device__ __noinline__
SPTX_CALL float tr_D(float ax, float ay, V3 wh) {
    float tan2Theta = tan2_theta(wh);
    if (isinf(tan2Theta)) return 0.f;
    const float cos4Theta = cos2_theta(wh) * cos2_theta(wh);
    float e = (cos2_phi(wh) / (ax * ax) + sin2_phi(wh) / (ay * ay)) * tan2Theta;
    return 1.f / (kPi * ax * ay * cos4Theta * (1.f + e) * (1.f + e));
}
SPTX_CALL float tr_lambda(float ax, float ay, V3 w) {
    float absTanTheta = fabsf(tan_theta(w));
    if (isinf(absTanTheta)) return 0.f;
    float alpha = sqrtf(cos2_phi(w) * ax * ax + sin2_phi(w) * ay * ay);
    float alpha2Tan2Theta = (alpha * absTanTheta) * (alpha * absTanTheta);
    return (-1.f + sqrtf(1.f + alpha2Tan2Theta)) / 2.f;
}
SPTX_DEV float tr_G1(float ax, float ay, V3 w) { return 1.f / (1.f + tr_lambda(ax, ay, w)); }
SPTX_DEV float tr_G(float ax, float ay, V3 wo, V3 wi) {
    return 1.f / (1.f + tr_lambda(ax, ay, wo) + tr_lambda(ax, ay, wi));
}
SPTX_CALL float tr_pdf(const sptx_lobe &L, V3 wo, V3 wh) {
    if (L.sample_visible)
        return tr_D(L.alpha_x, L.alpha_y, wh) * tr_G1(L.alpha_x, L.alpha_y, wo) * absdot(wo, wh) /
               abs_cos_theta(wo);
    return tr_D(L.alpha_x, L.alpha_y, wh) * abs_cos_theta(wh);
}
SPTX_DEV void tr_sample11(float cosTheta, float U1, float U2, float *slope_x, float *slope_y) {
    if ((double)cosTheta > .9999) {
        float r = sqrtf(U1 / (1.f - U1));
        float phi = (float)(6.28318530718 * (double)U2);
        *slope_x = r * cosf(phi);
        *slope_y = r * sinf(phi);
        return;
    }
    float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
    float tanTheta = sinTheta / cosTheta;
    float a = 1.f / tanTheta;
    float G1 = 2.f / (1.f + sqrtf(1.f + 1.f / (a * a)));
    float A = 2.f * U1 / G1 - 1.f;
    float tmp = 1.f / (A * A - 1.f);
    if (tmp > 1e10f) tmp = 1e10f;
    float B = tanTheta;
    float D = sqrtf(fmaxf(B * B * tmp * tmp - (A * A - B * B) * tmp, 0.f));
    float slope_x_1 = B * tmp - D;
    float slope_x_2 = B * tmp + D;
    *slope_x = (A < 0 || slope_x_2 > 1.f / tanTheta) ? slope_x_1 : slope_x_2;
    float Sg;
    if (U2 > 0.5f) {
        Sg = 1.f;
        U2 = 2.f * (U2 - .5f);
    } else {
        Sg = -1.f;
        U2 = 2.f * (.5f - U2);
    }
    float z = (U2 * (U2 * (U2 * 0.27385f - 0.73369f) + 0.46341f)) /
              (U2 * (U2 * (U2 * 0.093073f + 0.309420f) - 1.000000f) + 0.597999f);
    *slope_y = Sg * z * sqrtf(1.f + *slope_x * *slope_x);
}
SPTX_CALL V3 tr_sample_wh(const sptx_lobe &L, V3 wo, V2 u) {
    const float ax = L.alpha_x, ay = L.alpha_y;
    V3 wh;
    if (!L.sample_visible) {
        float cosTheta = 0, phi = (2 * kPi) * u.y;
        if (ax == ay) {
            float tanTheta2 = ax * ax * u.x / (1.0f - u.x);
            cosTheta = 1.f / sqrtf(1.f + tanTheta2);
        } else {
            phi = atanf(ay / ax * tanf(2 * kPi * u.y + .5f * kPi));
            if (u.y > .5f) phi += kPi;
            float sinPhi = sinf(phi), cosPhi = cosf(phi);
            const float ax2 = ax * ax, ay2 = ay * ay;
            const float alpha2 = 1.f / (cosPhi * cosPhi / ax2 + sinPhi * sinPhi / ay2);
            float tanTheta2 = alpha2 * u.x / (1.f - u.x);
            cosTheta = 1.f / sqrtf(1.f + tanTheta2);
        }
        float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
        wh = mk(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
        if (!same_hemisphere(wo, wh)) wh = -wh;
    } else {
        bool flip = wo.z < 0;
        V3 wi = flip ? -wo : wo;
        V3 ws = normalize(mk(ax * wi.x, ay * wi.y, wi.z));
        float sx, sy;
        tr_sample11(cos_theta(ws), u.x, u.y, &sx, &sy);
        float tmp = cos_phi(ws) * sx - sin_phi(ws) * sy;
        sy = sin_phi(ws) * sx + cos_phi(ws) * sy;
        sx = tmp;
        sx = ax * sx;
        sy = ay * sy;
        wh = normalize(mk(-sx, -sy, 1.f));
        if (flip) wh = -wh;
    }
    return wh;
}

}  // namespace sptx
