// bsdf_eval.cuh -- BSDF::f, BSDF::Pdf, BSDF::Sample_f (core/reflection.cpp:670-785).
#pragma once
#include "bsdf.cuh"

namespace sptx {

struct Bsdf {
    V3 ns, ng, ss, ts;
    float eta;
    int first, n;  // lobes [first, first+n)
};

SPTX_DEV void bsdf_init(const DScene &S, const Isect &it, Bsdf *b) {
    // BSDF::BSDF, core/reflection.h:156-161
    const sptx_material M = S.materials[it.material];
    b->eta = M.eta;
    b->first = M.first_lobe;
    b->n = M.n_lobes;
    b->ns = it.ns;
    b->ng = it.n;
    b->ss = normalize(it.sdpdu);
    b->ts = cross(b->ns, b->ss);
}
SPTX_DEV V3 to_local(const Bsdf &b, V3 v) { return mk(dot(v, b.ss), dot(v, b.ts), dot(v, b.ns)); }
SPTX_DEV V3 to_world(const Bsdf &b, V3 v) {
    return mk(b.ss.x * v.x + b.ts.x * v.y + b.ns.x * v.z, b.ss.y * v.x + b.ts.y * v.y + b.ns.y * v.z,
              b.ss.z * v.x + b.ts.z * v.y + b.ns.z * v.z);
}

SPTX_DEV int lobe_flags(int type) {
    switch (type) {
    case SPTX_LOBE_LAMBERT:
    case SPTX_LOBE_OREN_NAYAR:
        return BSDF_REFLECTION | BSDF_DIFFUSE;
    case SPTX_LOBE_SPECULAR_REFL:
        return BSDF_REFLECTION | BSDF_SPECULAR;
    case SPTX_LOBE_SPECULAR_TRANS:
        return BSDF_TRANSMISSION | BSDF_SPECULAR;
    case SPTX_LOBE_FRESNEL_SPECULAR:
        return BSDF_REFLECTION | BSDF_TRANSMISSION | BSDF_SPECULAR;
    case SPTX_LOBE_LAMBERT_TRANS:
        return BSDF_TRANSMISSION | BSDF_DIFFUSE;
    case SPTX_LOBE_MICROFACET_TRANS:
        return BSDF_TRANSMISSION | BSDF_GLOSSY;
    default:
        return BSDF_REFLECTION | BSDF_GLOSSY;
    }
}
SPTX_DEV bool matches(int lobeFlags, int flags) { return (lobeFlags & flags) == lobeFlags; }

template <int LOBES = kAllLobes>
SPTX_DEV int bsdf_num_components(const DScene &S, const Bsdf &b, int flags) {
    // a variant made of specular lobes only has no component outside BSDF_SPECULAR
    if ((LOBES & ~kSpecularLobes) == 0 && !(flags & BSDF_SPECULAR)) return 0;
    if (LOBES == (1 << SPTX_LOBE_LAMBERT)) return b.n;  // 0 or 1 diffuse reflection lobe
    int num = 0;
    for (int i = 0; i < b.n; ++i)
        if (matches(lobe_flags(S.lobes[b.first + i].type), flags)) ++num;
    return num;
}

// BxDF::f for the non-specular lobes; appends one term (or nothing when the
// lobe evaluates to zero).
template <int LOBES = kAllLobes, int MAXT = kMaxTerms>
SPTX_NOINL void lobe_f(const sptx_lobe &L, V3 wo, V3 wi, FSumN<MAXT> *out) {
    FTerm t;
    t.R = L.R;
    t.e0 = L.fr_eta_i;
    t.e1 = L.fr_eta_t;
    t.e2 = L.fr_k;
    t.s0 = t.s1 = t.s2 = t.s3 = 0.f;
    if (has_lobe<LOBES>(SPTX_LOBE_LAMBERT) && L.type == SPTX_LOBE_LAMBERT) {
        t.kind = TERM_R_S;
        t.s0 = kInvPi;
    } else if (has_lobe<LOBES>(SPTX_LOBE_OREN_NAYAR) && L.type == SPTX_LOBE_OREN_NAYAR) {
        // OrenNayar::f, core/reflection.cpp:197-219
        float sinThetaI = sin_theta(wi), sinThetaO = sin_theta(wo);
        float maxCos = 0;
        if ((double)sinThetaI > 1e-4 && (double)sinThetaO > 1e-4) {
            float sinPhiI = sin_phi(wi), cosPhiI = cos_phi(wi);
            float sinPhiO = sin_phi(wo), cosPhiO = cos_phi(wo);
            float dCos = cosPhiI * cosPhiO + sinPhiI * sinPhiO;
            maxCos = fmaxf(0.f, dCos);
        }
        float sinAlpha, tanBeta;
        if (abs_cos_theta(wi) > abs_cos_theta(wo)) {
            sinAlpha = sinThetaO;
            tanBeta = sinThetaI / abs_cos_theta(wi);
        } else {
            sinAlpha = sinThetaI;
            tanBeta = sinThetaO / abs_cos_theta(wo);
        }
        t.kind = TERM_R_S_S;
        t.s0 = kInvPi;
        t.s1 = (L.A + L.B * maxCos * sinAlpha * tanBeta);
    } else if (has_lobe<LOBES>(SPTX_LOBE_MICROFACET_REFL) && L.type == SPTX_LOBE_MICROFACET_REFL) {
        // MicrofacetReflection::f, core/reflection.cpp:226-236
        float cosThetaO = abs_cos_theta(wo), cosThetaI = abs_cos_theta(wi);
        V3 wh = wi + wo;
        if (cosThetaI == 0 || cosThetaO == 0) return;
        if (wh.x == 0 && wh.y == 0 && wh.z == 0) return;
        wh = normalize(wh);
        float c = dot(wi, wh);
        t.s0 = tr_D(L.alpha_x, L.alpha_y, wh);
        t.s1 = tr_G(L.alpha_x, L.alpha_y, wo, wi);
        t.s3 = 4.f * cosThetaI * cosThetaO;
        if (L.fresnel == SPTX_FRESNEL_CONDUCTOR) {
            t.kind = TERM_MICRO_COND;
            t.s2 = fabsf(c);
        } else {
            t.kind = TERM_MICRO_S;
            t.s2 = (L.fresnel == SPTX_FRESNEL_DIELECTRIC) ? fr_dielectric(c, L.eta_a, L.eta_b) : 1.f;
        }
    } else if (has_lobe<LOBES>(SPTX_LOBE_LAMBERT_TRANS) && L.type == SPTX_LOBE_LAMBERT_TRANS) {
        // LambertianTransmission::f, core/reflection.cpp:187-190
        t.kind = TERM_R_S;
        t.R = L.T;
        t.s0 = kInvPi;
    } else if (has_lobe<LOBES>(SPTX_LOBE_MICROFACET_TRANS) && L.type == SPTX_LOBE_MICROFACET_TRANS) {
        // MicrofacetTransmission::f, core/reflection.cpp:244-266
        if (same_hemisphere(wo, wi)) return;
        float cosThetaO = cos_theta(wo), cosThetaI = cos_theta(wi);
        if (cosThetaI == 0 || cosThetaO == 0) return;
        float eta = cos_theta(wo) > 0 ? (L.eta_b / L.eta_a) : (L.eta_a / L.eta_b);
        V3 wh = normalize(wo + wi * eta);
        if (wh.z < 0) wh = -wh;
        float F = fr_dielectric(dot(wo, wh), L.eta_a, L.eta_b);
        float sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
        float factor = 1.f / eta;  // TransportMode::Radiance
        t.kind = TERM_R_S_S;  // ((1 - F) * T[b]) * |...|
        t.R = L.T;
        t.s0 = 1.f - F;
        t.s1 = fabsf(tr_D(L.alpha_x, L.alpha_y, wh) * tr_G(L.alpha_x, L.alpha_y, wo, wi) * eta * eta *
                     absdot(wi, wh) * absdot(wo, wh) * factor * factor /
                     (cosThetaI * cosThetaO * sqrtDenom * sqrtDenom));
    } else if (has_lobe<LOBES>(SPTX_LOBE_FRESNEL_BLEND) && L.type == SPTX_LOBE_FRESNEL_BLEND) {
        // FresnelBlend::f, core/reflection.cpp:285-299
        V3 wh = wi + wo;
        if (wh.x == 0 && wh.y == 0 && wh.z == 0) return;
        wh = normalize(wh);
        const float a = 1.f - .5f * abs_cos_theta(wi), b = 1.f - .5f * abs_cos_theta(wo);
        const float c = 1.f - dot(wi, wh);
        t.kind = TERM_BLEND;
        t.e0 = L.T;  // Rs
        t.s0 = 1.f - (a * a) * (a * a) * a;
        t.s1 = 1.f - (b * b) * (b * b) * b;
        t.s2 = (c * c) * (c * c) * c;
        t.s3 = tr_D(L.alpha_x, L.alpha_y, wh) /
               (4.f * absdot(wi, wh) * fmaxf(abs_cos_theta(wi), abs_cos_theta(wo)));
    } else
        return;  // specular lobes: f() == 0
    if (out->n < MAXT) out->t[out->n++] = t;
}

// BxDF::Pdf, core/reflection.cpp:386-388 and MicrofacetReflection::Pdf :418-422.
template <int LOBES = kAllLobes>
SPTX_NOINL float lobe_pdf(const sptx_lobe &L, V3 wo, V3 wi) {
    if ((has_lobe<LOBES>(SPTX_LOBE_LAMBERT) && L.type == SPTX_LOBE_LAMBERT) ||
        (has_lobe<LOBES>(SPTX_LOBE_OREN_NAYAR) && L.type == SPTX_LOBE_OREN_NAYAR))
        return same_hemisphere(wo, wi) ? abs_cos_theta(wi) * kInvPi : 0.f;
    if (has_lobe<LOBES>(SPTX_LOBE_MICROFACET_REFL) && L.type == SPTX_LOBE_MICROFACET_REFL) {
        if (!same_hemisphere(wo, wi)) return 0.f;
        V3 wh = normalize(wo + wi);
        return tr_pdf(L, wo, wh) / (4.f * dot(wo, wh));
    }
    if (has_lobe<LOBES>(SPTX_LOBE_LAMBERT_TRANS) && L.type == SPTX_LOBE_LAMBERT_TRANS)  // core/reflection.cpp:400-403
        return !same_hemisphere(wo, wi) ? abs_cos_theta(wi) * kInvPi : 0.f;
    if (has_lobe<LOBES>(SPTX_LOBE_MICROFACET_TRANS) && L.type == SPTX_LOBE_MICROFACET_TRANS) {
        // MicrofacetTransmission::Pdf, core/reflection.cpp:436-448
        if (same_hemisphere(wo, wi)) return 0.f;
        float eta = cos_theta(wo) > 0 ? (L.eta_b / L.eta_a) : (L.eta_a / L.eta_b);
        V3 wh = normalize(wo + wi * eta);
        float sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
        float dwh_dwi = fabsf((eta * eta * dot(wi, wh)) / (sqrtDenom * sqrtDenom));
        return tr_pdf(L, wo, wh) * dwh_dwi;
    }
    if (has_lobe<LOBES>(SPTX_LOBE_FRESNEL_BLEND) && L.type == SPTX_LOBE_FRESNEL_BLEND) {
        // FresnelBlend::Pdf, core/reflection.cpp:470-475
        if (!same_hemisphere(wo, wi)) return 0.f;
        V3 wh = normalize(wo + wi);
        float pdf_wh = tr_pdf(L, wo, wh);
        return .5f * (abs_cos_theta(wi) * kInvPi + pdf_wh / (4.f * dot(wo, wh)));
    }
    return 0.f;  // specular lobes
}

// BSDF::f, core/reflection.cpp:670-683.
// The matte variant: a material is one LambertianReflection lobe (or none, Kd black), which
// matches every `flags` the path asks with; the generic lobe loops below reduce to this.
constexpr int kLambertOnly = 1 << SPTX_LOBE_LAMBERT;
template <int MAXT>
SPTX_DEV void lambert_term(const DScene &S, const Bsdf &b, FSumN<MAXT> *f) {
    FTerm t;
    t.kind = TERM_R_S;
    t.R = S.lobes[b.first].R;
    t.e0 = t.e1 = t.e2 = -1;
    t.s0 = kInvPi;
    t.s1 = t.s2 = t.s3 = 0.f;
    f->t[0] = t;
    f->n = 1;
}

template <int LOBES = kAllLobes, int MAXT = kMaxTerms>
SPTX_NOINL void bsdf_f(const DScene &S, const Bsdf &b, V3 woW, V3 wiW, int flags, FSumN<MAXT> *f) {
    f->n = 0;
    V3 wi = to_local(b, wiW), wo = to_local(b, woW);
    if (wo.z == 0) return;
    bool reflect = dot(wiW, b.ng) * dot(woW, b.ng) > 0;
    if (LOBES == kLambertOnly) {
        if (b.n > 0 && reflect) lambert_term(S, b, f);
        return;
    }
    for (int i = 0; i < b.n; ++i) {
        const sptx_lobe &L = S.lobes[b.first + i];
        int lf = lobe_flags(L.type);
        if (matches(lf, flags) &&
            ((reflect && (lf & BSDF_REFLECTION)) || (!reflect && (lf & BSDF_TRANSMISSION))))
            lobe_f<LOBES>(L, wo, wi, f);
    }
}

// BSDF::Pdf, core/reflection.cpp:770-785.
template <int LOBES = kAllLobes>
SPTX_NOINL float bsdf_pdf(const DScene &S, const Bsdf &b, V3 woW, V3 wiW, int flags) {
    if (b.n == 0) return 0.f;
    V3 wo = to_local(b, woW), wi = to_local(b, wiW);
    if (wo.z == 0) return 0.f;
    if (LOBES == kLambertOnly) return same_hemisphere(wo, wi) ? abs_cos_theta(wi) * kInvPi : 0.f;
    float pdf = 0.f;
    int matching = 0;
    for (int i = 0; i < b.n; ++i) {
        const sptx_lobe &L = S.lobes[b.first + i];
        if (matches(lobe_flags(L.type), flags)) {
            ++matching;
            pdf += lobe_pdf<LOBES>(L, wo, wi);
        }
    }
    return matching > 0 ? pdf / (float)matching : 0.f;
}

// CosineSampleHemisphere, core/sampling.h:155-159.
SPTX_DEV V3 cosine_sample_hemisphere(V2 u) {
    V2 d = concentric_sample_disk(u);
    float z = sqrtf(fmaxf(0.f, 1.f - d.x * d.x - d.y * d.y));
    return mk(d.x, d.y, z);
}

// BSDF::Sample_f, core/reflection.cpp:703-768.  Returns false when the
// reference returns a black spectrum with pdf == 0 (f->n == 0, *pdf == 0).
template <int LOBES = kAllLobes, int MAXT = kMaxTerms>
SPTX_NOINL void bsdf_sample_f(const DScene &S, const Bsdf &b, V3 woW, V2 u, int type, V3 *wiW,
                            float *pdf, int *sampledType, FSumN<MAXT> *f) {
    f->n = 0;
    *pdf = 0.f;
    *sampledType = 0;
    if (LOBES == kLambertOnly) {
        // matchingComps = 1, comp = 0, uRemapped = u (core/reflection.cpp:707-722), then
        // BxDF::Sample_f (:378-385) and the f of the one lobe
        if (b.n == 0) return;
        const V3 wo = to_local(b, woW);
        if (wo.z == 0) return;
        *sampledType = BSDF_REFLECTION | BSDF_DIFFUSE;
        V3 wi = cosine_sample_hemisphere(V2{fminf(u.x * 1.f - 0.f, kOneMinusEpsilon), u.y});
        if (wo.z < 0) wi.z *= -1.f;
        *pdf = same_hemisphere(wo, wi) ? abs_cos_theta(wi) * kInvPi : 0.f;
        if (*pdf == 0) {
            *sampledType = 0;
            return;
        }
        *wiW = to_world(b, wi);
        if (dot(*wiW, b.ng) * dot(woW, b.ng) > 0) lambert_term(S, b, f);
        return;
    }
    int matching = bsdf_num_components<LOBES>(S, b, type);
    if (matching == 0) return;
    int comp = min((int)floorf(u.x * (float)matching), matching - 1);
    int chosen = -1, count = comp;
    for (int i = 0; i < b.n; ++i)
        if (matches(lobe_flags(S.lobes[b.first + i].type), type) && count-- == 0) {
            chosen = i;
            break;
        }
    V2 ur{fminf(u.x * (float)matching - (float)comp, kOneMinusEpsilon), u.y};
    V3 wo = to_local(b, woW), wi = mk(0, 0, 0);
    if (wo.z == 0) return;
    const sptx_lobe &L = S.lobes[b.first + chosen];
    const int lf = lobe_flags(L.type);
    *sampledType = lf;
    FTerm t;
    t.R = L.R;
    t.e0 = L.fr_eta_i;
    t.e1 = L.fr_eta_t;
    t.e2 = L.fr_k;
    t.s0 = t.s1 = t.s2 = t.s3 = 0.f;
    bool haveTerm = false;
    if ((has_lobe<LOBES>(SPTX_LOBE_LAMBERT) && L.type == SPTX_LOBE_LAMBERT) ||
        (has_lobe<LOBES>(SPTX_LOBE_OREN_NAYAR) && L.type == SPTX_LOBE_OREN_NAYAR)) {
        // BxDF::Sample_f, core/reflection.cpp:378-385
        wi = cosine_sample_hemisphere(ur);
        if (wo.z < 0) wi.z *= -1.f;
        *pdf = lobe_pdf<LOBES>(L, wo, wi);
    } else if (has_lobe<LOBES>(SPTX_LOBE_MICROFACET_REFL) && L.type == SPTX_LOBE_MICROFACET_REFL) {
        // MicrofacetReflection::Sample_f, core/reflection.cpp:405-416
        V3 wh = tr_sample_wh(L, wo, ur);
        wi = -wo + 2.f * dot(wo, wh) * wh;  // Reflect(wo, wh)
        if (same_hemisphere(wo, wi)) *pdf = tr_pdf(L, wo, wh) / (4.f * dot(wo, wh));
    } else if (has_lobe<LOBES>(SPTX_LOBE_LAMBERT_TRANS) && L.type == SPTX_LOBE_LAMBERT_TRANS) {
        // LambertianTransmission::Sample_f, core/reflection.cpp:391-398
        wi = cosine_sample_hemisphere(ur);
        if (wo.z > 0) wi.z *= -1.f;
        *pdf = lobe_pdf<LOBES>(L, wo, wi);
    } else if (has_lobe<LOBES>(SPTX_LOBE_MICROFACET_TRANS) && L.type == SPTX_LOBE_MICROFACET_TRANS) {
        // MicrofacetTransmission::Sample_f, core/reflection.cpp:425-434
        V3 wh = tr_sample_wh(L, wo, ur);
        float eta = cos_theta(wo) > 0 ? (L.eta_a / L.eta_b) : (L.eta_b / L.eta_a);
        if (refract(wo, wh, eta, &wi)) *pdf = lobe_pdf<LOBES>(L, wo, wi);
    } else if (has_lobe<LOBES>(SPTX_LOBE_FRESNEL_BLEND) && L.type == SPTX_LOBE_FRESNEL_BLEND) {
        // FresnelBlend::Sample_f, core/reflection.cpp:450-468
        if (ur.x < .5f) {
            V2 u2{fminf(2.f * ur.x, kOneMinusEpsilon), ur.y};
            wi = cosine_sample_hemisphere(u2);
            if (wo.z < 0) wi.z *= -1.f;
            *pdf = lobe_pdf<LOBES>(L, wo, wi);
        } else {
            V2 u2{fminf(2.f * (ur.x - .5f), kOneMinusEpsilon), ur.y};
            V3 wh = tr_sample_wh(L, wo, u2);
            wi = -wo + 2.f * dot(wo, wh) * wh;  // Reflect(wo, wh)
            if (same_hemisphere(wo, wi)) *pdf = lobe_pdf<LOBES>(L, wo, wi);
        }
    } else if (has_lobe<LOBES>(SPTX_LOBE_SPECULAR_REFL) && L.type == SPTX_LOBE_SPECULAR_REFL) {
        // SpecularReflection::Sample_f, core/reflection.cpp:136-143
        wi = mk(-wo.x, -wo.y, wo.z);
        *pdf = 1.f;
        t.s1 = abs_cos_theta(wi);
        if (L.fresnel == SPTX_FRESNEL_CONDUCTOR) {
            t.kind = TERM_COND_R_DIV;
            t.s0 = fabsf(cos_theta(wi));
        } else {
            t.kind = TERM_S_R_DIV;
            t.s0 = (L.fresnel == SPTX_FRESNEL_DIELECTRIC) ? fr_dielectric(cos_theta(wi), L.eta_a, L.eta_b)
                                                          : 1.f;
        }
        haveTerm = true;
    } else if (has_lobe<LOBES>(SPTX_LOBE_SPECULAR_TRANS) && L.type == SPTX_LOBE_SPECULAR_TRANS) {
        // SpecularTransmission::Sample_f, core/reflection.cpp:150-166
        bool entering = cos_theta(wo) > 0;
        float etaI = entering ? L.eta_a : L.eta_b;
        float etaT = entering ? L.eta_b : L.eta_a;
        V3 nf = faceforward(mk(0, 0, 1), wo);
        if (refract(wo, nf, etaI / etaT, &wi)) {
            *pdf = 1.f;
            t.kind = TERM_T_S_S_DIV;
            t.R = L.T;
            t.s0 = 1.f - fr_dielectric(cos_theta(wi), L.eta_a, L.eta_b);
            t.s1 = (etaI * etaI) / (etaT * etaT);
            t.s2 = abs_cos_theta(wi);
            haveTerm = true;
        }
    } else if (has_lobe<LOBES>(SPTX_LOBE_FRESNEL_SPECULAR)) {  // core/reflection.cpp:477-511
        float F = fr_dielectric(cos_theta(wo), L.eta_a, L.eta_b);
        if (ur.x < F) {
            wi = mk(-wo.x, -wo.y, wo.z);
            *sampledType = BSDF_SPECULAR | BSDF_REFLECTION;
            *pdf = F;
            t.kind = TERM_S_R_DIV;
            t.s0 = F;
            t.s1 = abs_cos_theta(wi);
            haveTerm = true;
        } else {
            bool entering = cos_theta(wo) > 0;
            float etaI = entering ? L.eta_a : L.eta_b;
            float etaT = entering ? L.eta_b : L.eta_a;
            V3 nf = faceforward(mk(0, 0, 1), wo);
            if (refract(wo, nf, etaI / etaT, &wi)) {
                *sampledType = BSDF_SPECULAR | BSDF_TRANSMISSION;
                *pdf = 1.f - F;
                t.kind = TERM_T_S_S_DIV;
                t.R = L.T;
                t.s0 = 1.f - F;
                t.s1 = (etaI * etaI) / (etaT * etaT);
                t.s2 = abs_cos_theta(wi);
                haveTerm = true;
            }
        }
    }
    if (*pdf == 0) {
        *sampledType = 0;
        return;
    }
    *wiW = to_world(b, wi);
    if (!(lf & BSDF_SPECULAR) && matching > 1)
        for (int i = 0; i < b.n; ++i) {
            const sptx_lobe &Lo = S.lobes[b.first + i];
            if (i != chosen && matches(lobe_flags(Lo.type), type)) *pdf += lobe_pdf<LOBES>(Lo, wo, wi);
        }
    if (matching > 1) *pdf /= (float)matching;
    if (!(lf & BSDF_SPECULAR)) {
        bool reflect = dot(*wiW, b.ng) * dot(woW, b.ng) > 0;
        for (int i = 0; i < b.n; ++i) {
            const sptx_lobe &Lo = S.lobes[b.first + i];
            int f2 = lobe_flags(Lo.type);
            if (matches(f2, type) &&
                ((reflect && (f2 & BSDF_REFLECTION)) || (!reflect && (f2 & BSDF_TRANSMISSION))))
                lobe_f<LOBES>(Lo, wo, wi, f);
        }
    } else if (haveTerm) {
        f->t[0] = t;
        f->n = 1;
    }
}

}  // namespace sptx
