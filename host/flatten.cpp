// host/flatten.cpp -- reference scene graph -> POD tables (include/sptx.h).
//
// Product host code.  The reference keeps the data this needs in private
// members and has no accessor for most of it (BVHAccel::nodes, Triangle::mesh,
// GeometricPrimitive::shape, Film::filterTable ...).  A maintainer would add
// `friend` declarations (INTEGRATION.md lists them); this in-place build opens
// the same members by re-defining the access keywords for the reference's
// headers ONLY, after every standard header has been included.
#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#include <glog/logging.h>

#define private public
#define protected public
#include "accelerators/bvh.h"
#include "camera.h"
#include "cameras/perspective.h"
#include "cameras/realistic.h"
#include "film.h"
#include "filter.h"
#include "interaction.h"
#include "light.h"
#include "lights/diffuse.h"
#include "lights/infinite.h"
#include "mipmap.h"
#include "sampling.h"
#include "lights/point.h"
#include "lowdiscrepancy.h"
#include "material.h"
#include "materials/glass.h"
#include "materials/matte.h"
#include "materials/metal.h"
#include "materials/mirror.h"
#include "materials/plastic.h"
#include "materials/substrate.h"
#include "materials/translucent.h"
#include "materials/uber.h"
#include "memory.h"
#include "microfacet.h"
#include "primitive.h"
#include "reflection.h"
#include "sampler.h"
#include "samplers/halton.h"
#include "samplers/random.h"
#include "samplers/sobol.h"
#include "samplers/zerotwosequence.h"
#include "sobolmatrices.h"
#include "scene.h"
#include "shapes/sphere.h"
#include "shapes/triangle.h"
#include "spectrum.h"
#include "texture.h"
#include "textures/constant.h"
#include "textures/imagemap.h"
#include "transform.h"
#undef private
#undef protected

#include "flatten.h"

using namespace pbrt;

namespace sptx_host {

namespace {

// Same bytes as the reference's file-local LinearBVHNode
// (src/accelerators/bvh.cpp:95-104); sptx_bvh_node mirrors it field by field.
static_assert(sizeof(sptx_bvh_node) == 32, "LinearBVHNode is 32 bytes");

struct Builder {
    FlatScene *fs;
    std::string err;
    std::map<const TriangleMesh *, int> meshIndex;
    std::map<const Material *, int> materialIndex;
    std::map<const Light *, int> lightIndex;
    std::map<const Primitive *, int> bvhIndex;     // nested aggregates
    std::map<const Shape *, int> triangleOfShape;  // for area lights
    std::map<const Shape *, int> sphereOfShape;
    std::vector<const Primitive *> pending;        // aggregates to copy

    bool Fail(const std::string &m) {
        if (err.empty()) err = m;
        return false;
    }

    int AddSpectrum(const Spectrum &s) {
        int row = (int)(fs->spectra.size() / SPTX_SPECTRUM_STRIDE);
        for (int i = 0; i < SPTX_NBINS; ++i) fs->spectra.push_back(s[i]);
        for (int i = SPTX_NBINS; i < SPTX_SPECTRUM_STRIDE; ++i)
            fs->spectra.push_back(0.f);
        return row;
    }

    static void CopyMatrix(const Matrix4x4 &m, float *dst) {
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) dst[4 * r + c] = m.m[r][c];
    }

    template <typename T>
    static bool IsConstant(const std::shared_ptr<Texture<T>> &t) {
        return !t || dynamic_cast<const ConstantTexture<T> *>(t.get());
    }

    typedef ImageTexture<RGBSpectrum, Spectrum> SpectrumImage;
    static const SpectrumImage *AsImage(const std::shared_ptr<Texture<Spectrum>> &t) {
        return dynamic_cast<const SpectrumImage *>(t.get());
    }
    std::map<const void *, int> textureOffset;
    int weightLutOffset = -1;
    // The seven rgbIllum2Spect* rows: what SampledSpectrum::FromRGB expands an RGB value over when
    // called without a SpectrumType (core/spectrum.h:428-429: the default is Illuminant) -- that is
    // how InfiniteAreaLight reads its map AND how ImageTexture::convertOut turns a texel into a
    // Spectrum (textures/imagemap.h:113-117).
    int illumBasisRow = -1;
    int IlluminantBasis() {
        if (illumBasisRow < 0) {
            illumBasisRow = AddSpectrum(SampledSpectrum::rgbIllum2SpectWhite);
            AddSpectrum(SampledSpectrum::rgbIllum2SpectCyan);
            AddSpectrum(SampledSpectrum::rgbIllum2SpectMagenta);
            AddSpectrum(SampledSpectrum::rgbIllum2SpectYellow);
            AddSpectrum(SampledSpectrum::rgbIllum2SpectRed);
            AddSpectrum(SampledSpectrum::rgbIllum2SpectGreen);
            AddSpectrum(SampledSpectrum::rgbIllum2SpectBlue);
        }
        return illumBasisRow;
    }
    int AddImageTexture(const SpectrumImage *t);
    bool CheckMaterialTextures(const Material *m);
    int AddMaterial(const Material *m);
    bool AddLobe(const BxDF *b);
    int AddMesh(const TriangleMesh *mesh, bool flip);
    bool AddShape(const Shape *shape, sptx_prim *out);
    bool AddPrimitive(const Primitive *p, sptx_prim *out);
    bool AddAggregate(const Primitive *agg, bool topLevel);
};

bool Builder::CheckMaterialTextures(const Material *m) {
    // Only constant textures are on the hot path (SURVEY.md 2.1 "Textures");
    // bump maps perturb the shading frame per hit and are out as well.
    if (auto *x = dynamic_cast<const MatteMaterial *>(m)) {
        if (x->bumpMap) return Fail("matte: bumpmap is not supported");
        if ((!IsConstant(x->Kd) && !AsImage(x->Kd)) || !IsConstant(x->sigma))
            return Fail("matte: only constant textures (and an imagemap Kd) are supported");
        return true;
    }
    if (auto *x = dynamic_cast<const PlasticMaterial *>(m)) {
        if (x->bumpMap) return Fail("plastic: bumpmap is not supported");
        if ((!IsConstant(x->Kd) && !AsImage(x->Kd)) || !IsConstant(x->Ks) ||
            !IsConstant(x->roughness))
            return Fail("plastic: only constant textures (and an imagemap Kd) are supported");
        return true;
    }
    if (auto *x = dynamic_cast<const GlassMaterial *>(m)) {
        if (x->bumpMap) return Fail("glass: bumpmap is not supported");
        if (!IsConstant(x->Kr) || !IsConstant(x->Kt) ||
            !IsConstant(x->uRoughness) || !IsConstant(x->vRoughness) ||
            !IsConstant(x->index))
            return Fail("glass: only constant textures are supported");
        return true;
    }
    if (auto *x = dynamic_cast<const MetalMaterial *>(m)) {
        if (x->bumpMap) return Fail("metal: bumpmap is not supported");
        if (!IsConstant(x->eta) || !IsConstant(x->k) ||
            !IsConstant(x->roughness) || !IsConstant(x->uRoughness) ||
            !IsConstant(x->vRoughness))
            return Fail("metal: only constant textures are supported");
        return true;
    }
    if (auto *x = dynamic_cast<const MirrorMaterial *>(m)) {
        if (x->bumpMap) return Fail("mirror: bumpmap is not supported");
        if (!IsConstant(x->Kr))
            return Fail("mirror: only constant textures are supported");
        return true;
    }
    if (auto *x = dynamic_cast<const UberMaterial *>(m)) {
        // src/materials/uber.cpp:44-92: up to five lobes, all of types the
        // device already evaluates (SpecularTransmission for 1 - opacity,
        // Lambertian, Trowbridge-Reitz microfacet and specular reflection with
        // a dielectric Fresnel term, SpecularTransmission for Kt)
        if (x->bumpMap) return Fail("uber: bumpmap is not supported");
        if ((!IsConstant(x->Kd) && !AsImage(x->Kd)) || !IsConstant(x->Ks) || !IsConstant(x->Kr) ||
            !IsConstant(x->Kt) || !IsConstant(x->opacity) ||
            !IsConstant(x->roughness) || !IsConstant(x->roughnessu) ||
            !IsConstant(x->roughnessv) || !IsConstant(x->eta))
            return Fail("uber: only constant textures (and an imagemap Kd) are supported");
        if (AsImage(x->Kd)) {
            // kd = opacity * Kd (uber.cpp:52-57): the device takes the texture's value as it is
            SurfaceInteraction si;
            const Spectrum op = x->opacity->Evaluate(si);
            for (int i = 0; i < nSpectralSamples; ++i)
                if (op[i] != 1.f) return Fail("uber: an imagemap Kd needs opacity 1");
        }
        return true;
    }
    if (auto *x = dynamic_cast<const TranslucentMaterial *>(m)) {
        // src/materials/translucent.cpp:45-83: Lambertian reflection and
        // transmission, microfacet reflection and transmission
        if (x->bumpMap) return Fail("translucent: bumpmap is not supported");
        if (!IsConstant(x->Kd) || !IsConstant(x->Ks) ||
            !IsConstant(x->roughness) || !IsConstant(x->reflect) ||
            !IsConstant(x->transmit))
            return Fail("translucent: only constant textures are supported");
        return true;
    }
    if (auto *x = dynamic_cast<const SubstrateMaterial *>(m)) {
        // src/materials/substrate.cpp:45-65: one FresnelBlend lobe
        if (x->bumpMap) return Fail("substrate: bumpmap is not supported");
        if (!IsConstant(x->Kd) || !IsConstant(x->Ks) || !IsConstant(x->nu) ||
            !IsConstant(x->nv))
            return Fail("substrate: only constant textures are supported");
        return true;
    }
    return Fail(
        "material type is outside the GPU hot path (supported: matte, "
        "plastic, glass, metal, mirror, uber, translucent, substrate)");
}

bool Builder::AddLobe(const BxDF *b) {
    sptx_lobe l;
    std::memset(&l, 0, sizeof(l));
    l.R = l.T = l.fr_eta_i = l.fr_eta_t = l.fr_k = -1;
    auto setFresnel = [&](const Fresnel *f) -> bool {
        if (dynamic_cast<const FresnelNoOp *>(f)) {
            l.fresnel = SPTX_FRESNEL_NOOP;
        } else if (auto *d = dynamic_cast<const FresnelDielectric *>(f)) {
            l.fresnel = SPTX_FRESNEL_DIELECTRIC;
            l.eta_a = d->etaI;
            l.eta_b = d->etaT;
        } else if (auto *c = dynamic_cast<const FresnelConductor *>(f)) {
            l.fresnel = SPTX_FRESNEL_CONDUCTOR;
            l.fr_eta_i = AddSpectrum(c->etaI);
            l.fr_eta_t = AddSpectrum(c->etaT);
            l.fr_k = AddSpectrum(c->k);
        } else
            return Fail("unknown Fresnel type");
        return true;
    };
    if (auto *x = dynamic_cast<const LambertianReflection *>(b)) {
        l.type = SPTX_LOBE_LAMBERT;
        l.R = AddSpectrum(x->R);
    } else if (auto *x = dynamic_cast<const OrenNayar *>(b)) {
        l.type = SPTX_LOBE_OREN_NAYAR;
        l.R = AddSpectrum(x->R);
        l.A = x->A;
        l.B = x->B;
    } else if (auto *x = dynamic_cast<const SpecularReflection *>(b)) {
        l.type = SPTX_LOBE_SPECULAR_REFL;
        l.R = AddSpectrum(x->R);
        if (!setFresnel(x->fresnel)) return false;
    } else if (auto *x = dynamic_cast<const SpecularTransmission *>(b)) {
        l.type = SPTX_LOBE_SPECULAR_TRANS;
        l.T = AddSpectrum(x->T);
        l.eta_a = x->etaA;
        l.eta_b = x->etaB;
        if (x->mode != TransportMode::Radiance)
            return Fail("importance transport is not on the path");
    } else if (auto *x = dynamic_cast<const FresnelSpecular *>(b)) {
        l.type = SPTX_LOBE_FRESNEL_SPECULAR;
        l.R = AddSpectrum(x->R);
        l.T = AddSpectrum(x->T);
        l.eta_a = x->etaA;
        l.eta_b = x->etaB;
        if (x->mode != TransportMode::Radiance)
            return Fail("importance transport is not on the path");
    } else if (auto *x = dynamic_cast<const MicrofacetReflection *>(b)) {
        l.type = SPTX_LOBE_MICROFACET_REFL;
        l.R = AddSpectrum(x->R);
        auto *tr = dynamic_cast<const TrowbridgeReitzDistribution *>(
            x->distribution);
        if (!tr)
            return Fail("only the Trowbridge-Reitz distribution is supported");
        l.alpha_x = tr->alphax;
        l.alpha_y = tr->alphay;
        l.sample_visible = tr->sampleVisibleArea ? 1 : 0;
        if (!setFresnel(x->fresnel)) return false;
    } else if (auto *x = dynamic_cast<const LambertianTransmission *>(b)) {
        l.type = SPTX_LOBE_LAMBERT_TRANS;
        l.T = AddSpectrum(x->T);
    } else if (auto *x = dynamic_cast<const MicrofacetTransmission *>(b)) {
        l.type = SPTX_LOBE_MICROFACET_TRANS;
        l.T = AddSpectrum(x->T);
        l.eta_a = x->etaA;
        l.eta_b = x->etaB;
        auto *tr = dynamic_cast<const TrowbridgeReitzDistribution *>(
            x->distribution);
        if (!tr)
            return Fail("only the Trowbridge-Reitz distribution is supported");
        l.alpha_x = tr->alphax;
        l.alpha_y = tr->alphay;
        l.sample_visible = tr->sampleVisibleArea ? 1 : 0;
        if (x->mode != TransportMode::Radiance)
            return Fail("importance transport is not on the path");
    } else if (auto *x = dynamic_cast<const FresnelBlend *>(b)) {
        l.type = SPTX_LOBE_FRESNEL_BLEND;
        l.R = AddSpectrum(x->Rd);
        l.T = AddSpectrum(x->Rs);
        auto *tr = dynamic_cast<const TrowbridgeReitzDistribution *>(
            x->distribution);
        if (!tr)
            return Fail("only the Trowbridge-Reitz distribution is supported");
        l.alpha_x = tr->alphax;
        l.alpha_y = tr->alphay;
        l.sample_visible = tr->sampleVisibleArea ? 1 : 0;
    } else
        return Fail("BxDF outside the GPU hot path: " + b->ToString());
    fs->lobes.push_back(l);
    return true;
}

// One ImageTexture<RGBSpectrum, Spectrum> as a block of tex_floats (layout: include/sptx.h).
// Returns the block's offset, or -1 (err set) for what the device does not evaluate.
int Builder::AddImageTexture(const SpectrumImage *t) {
    auto it = textureOffset.find(t);
    if (it != textureOffset.end()) return it->second;
    auto *uv = dynamic_cast<const UVMapping2D *>(t->mapping.get());
    if (!uv) {
        Fail("imagemap: only the \"uv\" mapping is on the GPU hot path");
        return -1;
    }
    const MIPMap<RGBSpectrum> &mip = *t->mipmap;
    if (mip.noFiltering || t->spdFlag) {
        Fail("imagemap: noFiltering / display-primary textures are outside the GPU hot path");
        return -1;
    }
    std::vector<float> &T = fs->texFloats;
    auto putInt = [&](int v) {
        float f;
        std::memcpy(&f, &v, 4);
        T.push_back(f);
    };
    if (weightLutOffset < 0) {  // MIPMap::weightLut, filled by the first MIPMap constructed
        weightLutOffset = (int)T.size();
        for (int i = 0; i < 128; ++i) T.push_back(MIPMap<RGBSpectrum>::weightLut[i]);
    }
    const int basisRow = IlluminantBasis();
    const int block = (int)T.size();
    const int nLevels = mip.Levels();
    putInt(nLevels);
    putInt(mip.wrapMode == ImageWrap::Repeat ? 0 : (mip.wrapMode == ImageWrap::Black ? 1 : 2));
    putInt(mip.doTrilinear ? 1 : 0);
    T.push_back(mip.maxAnisotropy);
    T.push_back(uv->su);
    T.push_back(uv->sv);
    T.push_back(uv->du);
    T.push_back(uv->dv);
    putInt(basisRow);
    putInt(weightLutOffset);
    for (int i = 10; i < 16; ++i) T.push_back(0.f);
    const size_t levelTable = T.size();
    for (int l = 0; l < nLevels; ++l) {
        putInt(mip.pyramid[l]->uSize());
        putInt(mip.pyramid[l]->vSize());
        putInt(0);
    }
    for (int l = 0; l < nLevels; ++l) {
        const BlockedArray<RGBSpectrum> &lvl = *mip.pyramid[l];
        const int off = (int)T.size() - block;
        std::memcpy(&T[levelTable + 3 * l + 2], &off, 4);
        for (int v = 0; v < lvl.vSize(); ++v)
            for (int u = 0; u < lvl.uSize(); ++u) {
                Float rgb[3];
                lvl(u, v).ToRGB(rgb);
                T.push_back(rgb[0]);
                T.push_back(rgb[1]);
                T.push_back(rgb[2]);
            }
    }
    textureOffset[t] = block;
    return block;
}

int Builder::AddMaterial(const Material *m) {
    if (!m) return -1;
    auto it = materialIndex.find(m);
    if (it != materialIndex.end()) return it->second;
    if (!CheckMaterialTextures(m)) return -2;
    // An imagemap Kd (matte, plastic): the material's lobes are taken with a stand-in constant Kd
    // (so that the diffuse lobe, which the material adds first and only for a non-black value, is
    // there), then that lobe's R is pointed at the texture, which the device evaluates per hit.
    std::shared_ptr<Texture<Spectrum>> *kdSlot = nullptr;
    if (auto *x = dynamic_cast<const MatteMaterial *>(m))
        kdSlot = const_cast<std::shared_ptr<Texture<Spectrum>> *>(&x->Kd);
    else if (auto *x = dynamic_cast<const PlasticMaterial *>(m))
        kdSlot = const_cast<std::shared_ptr<Texture<Spectrum>> *>(&x->Kd);
    else if (auto *x = dynamic_cast<const UberMaterial *>(m))
        kdSlot = const_cast<std::shared_ptr<Texture<Spectrum>> *>(&x->Kd);
    int texBlock = -1;
    std::shared_ptr<Texture<Spectrum>> savedKd;
    if (kdSlot && AsImage(*kdSlot)) {
        texBlock = AddImageTexture(AsImage(*kdSlot));
        if (texBlock < 0) return -2;
        savedKd = *kdSlot;
        *kdSlot = std::make_shared<ConstantTexture<Spectrum>>(Spectrum(.5f));
    }
    // With constant textures the BSDF a material builds does not depend on
    // the hit, so build it once through the material's own code
    // (Material::ComputeScatteringFunctions, allowMultipleLobes = true as
    // src/integrators/path.cpp:107 passes) and record its lobes.
    MemoryArena arena;
    SurfaceInteraction si(Point3f(0, 0, 0), Vector3f(0, 0, 0), Point2f(0, 0),
                          Vector3f(0, 0, 1), Vector3f(1, 0, 0),
                          Vector3f(0, 1, 0), Normal3f(0, 0, 0),
                          Normal3f(0, 0, 0), 0, nullptr);
    m->ComputeScatteringFunctions(&si, arena, TransportMode::Radiance, true);
    if (savedKd) *kdSlot = savedKd;
    if (!si.bsdf) {
        Fail("material produced no BSDF");
        return -2;
    }
    if (si.bssrdf) {
        Fail("subsurface scattering is outside the GPU hot path");
        return -2;
    }
    sptx_material fm;
    fm.first_lobe = (int)fs->lobes.size();
    fm.n_lobes = si.bsdf->nBxDFs;
    fm.eta = si.bsdf->eta;
    fm.material_id = m->materialId;
    for (int i = 0; i < si.bsdf->nBxDFs; ++i)
        if (!AddLobe(si.bsdf->bxdfs[i])) return -2;
    if (texBlock >= 0) {
        sptx_lobe &l0 = fs->lobes[(size_t)fm.first_lobe];
        if (fm.n_lobes < 1 || l0.type != SPTX_LOBE_LAMBERT) {
            Fail("textured Kd: the material's first lobe is not a Lambert lobe (matte with sigma != 0?)");
            return -2;
        }
        l0.R = SPTX_TEX_ROW(texBlock);
        // the device's textured shading variant evaluates Lambert and microfacet-reflection lobes
        for (int i = 0; i < fm.n_lobes; ++i) {
            const int ty = fs->lobes[(size_t)fm.first_lobe + i].type;
            if (ty != SPTX_LOBE_LAMBERT && ty != SPTX_LOBE_MICROFACET_REFL) {
                Fail("textured Kd: the material has lobes besides diffuse and glossy reflection (uber with Kr / Kt?)");
                return -2;
            }
        }
    }
    int idx = (int)fs->materials.size();
    fs->materials.push_back(fm);
    materialIndex[m] = idx;
    return idx;
}

int Builder::AddMesh(const TriangleMesh *mesh, bool flip) {
    auto it = meshIndex.find(mesh);
    if (it != meshIndex.end()) {
        if ((fs->meshes[it->second].flip != 0) != flip) {
            Fail("triangles of one mesh disagree on orientation");
            return -1;
        }
        return it->second;
    }
    if (mesh->alphaMask || mesh->shadowAlphaMask) {
        Fail("alpha-masked meshes are outside the GPU hot path");
        return -1;
    }
    sptx_mesh fm;
    fm.n_vertices = mesh->nVertices;
    fm.flip = flip ? 1 : 0;
    fm.p_base = (int)(fs->positions.size() / 3);
    for (int i = 0; i < mesh->nVertices; ++i) {
        fs->positions.push_back(mesh->p[i].x);
        fs->positions.push_back(mesh->p[i].y);
        fs->positions.push_back(mesh->p[i].z);
    }
    fm.n_base = fm.s_base = fm.uv_base = -1;
    if (mesh->n) {
        fm.n_base = (int)(fs->normals.size() / 3);
        for (int i = 0; i < mesh->nVertices; ++i) {
            fs->normals.push_back(mesh->n[i].x);
            fs->normals.push_back(mesh->n[i].y);
            fs->normals.push_back(mesh->n[i].z);
        }
    }
    if (mesh->s) {
        fm.s_base = (int)(fs->tangents.size() / 3);
        for (int i = 0; i < mesh->nVertices; ++i) {
            fs->tangents.push_back(mesh->s[i].x);
            fs->tangents.push_back(mesh->s[i].y);
            fs->tangents.push_back(mesh->s[i].z);
        }
    }
    if (mesh->uv) {
        fm.uv_base = (int)(fs->uvs.size() / 2);
        for (int i = 0; i < mesh->nVertices; ++i) {
            fs->uvs.push_back(mesh->uv[i].x);
            fs->uvs.push_back(mesh->uv[i].y);
        }
    }
    int idx = (int)fs->meshes.size();
    fs->meshes.push_back(fm);
    meshIndex[mesh] = idx;
    return idx;
}

bool Builder::AddShape(const Shape *shape, sptx_prim *out) {
    if (auto *tri = dynamic_cast<const Triangle *>(shape)) {
        auto known = triangleOfShape.find(shape);
        if (known != triangleOfShape.end()) {
            out->kind = SPTX_PRIM_TRIANGLE;
            out->shape = known->second;
            return true;
        }
        bool flip = tri->reverseOrientation ^ tri->transformSwapsHandedness;
        int mesh = AddMesh(tri->mesh.get(), flip);
        if (mesh < 0) return false;
        sptx_triangle t;
        t.v[0] = tri->v[0];
        t.v[1] = tri->v[1];
        t.v[2] = tri->v[2];
        t.mesh = mesh;
        out->kind = SPTX_PRIM_TRIANGLE;
        out->shape = (int)fs->triangles.size();
        fs->triangles.push_back(t);
        triangleOfShape[shape] = out->shape;
        return true;
    }
    if (auto *sph = dynamic_cast<const Sphere *>(shape)) {
        auto known = sphereOfShape.find(shape);
        if (known != sphereOfShape.end()) {
            out->kind = SPTX_PRIM_SPHERE;
            out->shape = known->second;
            return true;
        }
        sptx_sphere s;
        std::memset(&s, 0, sizeof(s));
        CopyMatrix(sph->ObjectToWorld->m, s.o2w);
        CopyMatrix(sph->ObjectToWorld->mInv, s.w2o);
        // WorldToObject is looked up separately (src/core/api.cpp:1367); it
        // must be the exact inverse pair for the device to use one copy.
        if (std::memcmp(&sph->WorldToObject->m, &sph->ObjectToWorld->mInv,
                        sizeof(Matrix4x4)) != 0)
            return Fail("sphere: WorldToObject is not ObjectToWorld's inverse");
        s.radius = sph->radius;
        s.z_min = sph->zMin;
        s.z_max = sph->zMax;
        s.theta_min = sph->thetaMin;
        s.theta_max = sph->thetaMax;
        s.phi_max = sph->phiMax;
        s.reverse_orientation = sph->reverseOrientation;
        s.swaps_handedness = sph->transformSwapsHandedness;
        out->kind = SPTX_PRIM_SPHERE;
        out->shape = (int)fs->spheres.size();
        fs->spheres.push_back(s);
        sphereOfShape[shape] = out->shape;
        return true;
    }
    return Fail(
        "shape type is outside the GPU hot path (supported: trianglemesh and "
        "the generators that emit one, sphere)");
}

bool Builder::AddPrimitive(const Primitive *p, sptx_prim *out) {
    out->kind = -1;
    out->shape = -1;
    out->material = -1;
    out->light = -1;
    if (auto *g = dynamic_cast<const GeometricPrimitive *>(p)) {
        if (g->mediumInterface.inside || g->mediumInterface.outside)
            return Fail("participating media are outside the GPU hot path");
        if (!AddShape(g->shape.get(), out)) return false;
        out->material = AddMaterial(g->material.get());
        if (out->material == -2) return false;
        if (out->material == -1)
            return Fail("primitives without a material (medium boundaries) "
                        "are outside the GPU hot path");
        if (g->areaLight) {
            auto it = lightIndex.find(g->areaLight.get());
            if (it == lightIndex.end())
                return Fail("area light of a primitive is not in scene.lights");
            out->light = it->second;
        }
        return true;
    }
    if (auto *t = dynamic_cast<const TransformedPrimitive *>(p)) {
        const AnimatedTransform &at = t->PrimitiveToWorld;
        if (at.actuallyAnimated)
            return Fail("animated instance transforms are not supported");
        sptx_instance inst;
        std::memset(&inst, 0, sizeof(inst));
        CopyMatrix(at.startTransform->m, inst.p2w);
        CopyMatrix(at.startTransform->mInv, inst.w2p);
        inst.instance_id = (int)t->instanceId;
        inst.is_identity = at.startTransform->IsIdentity() ? 1 : 0;
        const Primitive *inner = t->primitive.get();
        if (dynamic_cast<const TransformedPrimitive *>(inner))
            return Fail("nested object instances are not supported");
        auto known = bvhIndex.find(inner);
        if (known != bvhIndex.end())
            inst.bvh = known->second;
        else {
            inst.bvh = (int)fs->bvhs.size() + (int)pending.size();
            bvhIndex[inner] = inst.bvh;
            pending.push_back(inner);
        }
        out->kind = SPTX_PRIM_INSTANCE;
        out->shape = (int)fs->instances.size();
        fs->instances.push_back(inst);
        return true;
    }
    return Fail("primitive type is outside the GPU hot path");
}

// Copies one aggregate into slot fs->bvhs.size().  `agg` is either a BVHAccel
// or (for a one-primitive instance, src/core/api.cpp:1590) a bare
// GeometricPrimitive, stored as a node-less BVH of one primitive.
bool Builder::AddAggregate(const Primitive *agg, bool topLevel) {
    sptx_bvh fb;
    fb.node_base = (int)fs->nodes.size();
    fb.prim_base = (int)fs->prims.size();
    fb.n_nodes = 0;
    fb.n_prims = 0;
    std::vector<const Primitive *> members;
    if (auto *bvh = dynamic_cast<const BVHAccel *>(agg)) {
        fb.n_prims = (int)bvh->primitives.size();
        for (auto &sp : bvh->primitives) members.push_back(sp.get());
        if (bvh->nodes) {
            // The node count is a local of the constructor; recover it by
            // walking the depth-first layout (src/accelerators/bvh.cpp:640-658).
            const sptx_bvh_node *n =
                reinterpret_cast<const sptx_bvh_node *>(bvh->nodes);
            int maxIndex = 0;
            std::vector<int> stack{0};
            while (!stack.empty()) {
                int i = stack.back();
                stack.pop_back();
                maxIndex = std::max(maxIndex, i);
                if (n[i].n_prims == 0) {
                    stack.push_back(i + 1);
                    stack.push_back(n[i].offset);
                }
            }
            fb.n_nodes = maxIndex + 1;
            fs->nodes.insert(fs->nodes.end(), n, n + fb.n_nodes);
        }
    } else if (dynamic_cast<const GeometricPrimitive *>(agg) && !topLevel) {
        fb.n_prims = 1;
        members.push_back(agg);
    } else
        return Fail(
            "aggregate is not a BVHAccel (Accelerator \"bvh\" is the only "
            "one on the GPU hot path)");
    int slot = (int)fs->bvhs.size();
    fs->bvhs.push_back(fb);
    fs->prims.resize(fb.prim_base + fb.n_prims);
    for (int i = 0; i < fb.n_prims; ++i) {
        sptx_prim fp;
        if (!AddPrimitive(members[i], &fp)) return false;
        if (!topLevel && fp.kind == SPTX_PRIM_INSTANCE)
            return Fail("nested object instances are not supported");
        fs->prims[fb.prim_base + i] = fp;
    }
    (void)slot;
    return true;
}

}  // namespace

void FlatScene::Finish() {
    desc.abi_version = SPTX_ABI_VERSION;
#define SPTX_SET(field, count, vec, div) \
    desc.count = (int64_t)(vec.size() / (div)); \
    desc.field = vec.empty() ? nullptr : vec.data();
    SPTX_SET(nodes, n_nodes, nodes, 1)
    SPTX_SET(bvhs, n_bvhs, bvhs, 1)
    SPTX_SET(prims, n_prims, prims, 1)
    SPTX_SET(meshes, n_meshes, meshes, 1)
    SPTX_SET(triangles, n_triangles, triangles, 1)
    SPTX_SET(positions, n_positions, positions, 3)
    SPTX_SET(normals, n_normals, normals, 3)
    SPTX_SET(tangents, n_tangents, tangents, 3)
    SPTX_SET(uvs, n_uvs, uvs, 2)
    SPTX_SET(spheres, n_spheres, spheres, 1)
    SPTX_SET(instances, n_instances, instances, 1)
    SPTX_SET(materials, n_materials, materials, 1)
    SPTX_SET(lobes, n_lobes, lobes, 1)
    SPTX_SET(spectra, n_spectra, spectra, SPTX_SPECTRUM_STRIDE)
    SPTX_SET(lights, n_lights, lights, 1)
    SPTX_SET(lens_elements, n_lens_floats, lensElements, 1)
    SPTX_SET(exit_pupil_bounds, n_pupil_floats, exitPupilBounds, 1)
    SPTX_SET(primes, n_primes, primes, 1)
    SPTX_SET(prime_sums, n_prime_sums, primeSums, 1)
    SPTX_SET(perms, n_perms, perms, 1)
    SPTX_SET(env_floats, n_env_floats, envFloats, 1)
    SPTX_SET(tex_floats, n_tex_floats, texFloats, 1)
#undef SPTX_SET
}

bool FlattenScene(const Scene &scene, const Camera &camera, Sampler &sampler,
                  const PathParams &params, FlatScene *out, std::string *err) {
    Builder b;
    b.fs = out;
    std::memset(&out->desc, 0, sizeof(out->desc));
    auto fail = [&](const std::string &m) {
        *err = m;
        return false;
    };

    // ---- lights: index = position in scene.lights (light selection and
    // GetAreaLight() identity both depend on it, src/core/integrator.cpp:85-106)
    for (size_t i = 0; i < scene.lights.size(); ++i)
        b.lightIndex[scene.lights[i].get()] = (int)i;

    // ---- aggregates and everything hanging off them
    if (!b.AddAggregate(scene.aggregate.get(), true)) return fail(b.err);
    for (size_t i = 0; i < b.pending.size(); ++i)
        if (!b.AddAggregate(b.pending[i], false)) return fail(b.err);

    for (size_t i = 0; i < scene.lights.size(); ++i) {
        const Light *l = scene.lights[i].get();
        sptx_light fl;
        std::memset(&fl, 0, sizeof(fl));
        fl.shape = -1;
        if (l->mediumInterface.inside || l->mediumInterface.outside)
            return fail("participating media are outside the GPU hot path");
        if (auto *d = dynamic_cast<const DiffuseAreaLight *>(l)) {
            fl.type = SPTX_LIGHT_DIFFUSE_AREA;
            sptx_prim tmp;
            if (!b.AddShape(d->shape.get(), &tmp)) return fail(b.err);
            fl.shape_kind = tmp.kind;
            fl.shape = tmp.shape;
            fl.two_sided = d->twoSided ? 1 : 0;
            fl.spectrum = b.AddSpectrum(d->Lemit);
            fl.area = d->area;
        } else if (auto *p = dynamic_cast<const PointLight *>(l)) {
            fl.type = SPTX_LIGHT_POINT;
            fl.spectrum = b.AddSpectrum(p->I);
            fl.pos[0] = p->pLight.x;
            fl.pos[1] = p->pLight.y;
            fl.pos[2] = p->pLight.z;
        } else if (auto *inf = dynamic_cast<const InfiniteAreaLight *>(l)) {
            // InfiniteAreaLight (src/lights/infinite.cpp): the radiance map's finest level as
            // the MIPMap holds it (resampled to powers of two, already scaled by L), the
            // Distribution2D the constructor built, and the bounding sphere of Preprocess().
            fl.type = SPTX_LIGHT_INFINITE;
            fl.shape = (int)out->envFloats.size();
            fl.area = inf->worldRadius;
            fl.pos[0] = inf->worldCenter.x;
            fl.pos[1] = inf->worldCenter.y;
            fl.pos[2] = inf->worldCenter.z;
            fl.spectrum = b.IlluminantBasis();
            const MIPMap<RGBSpectrum> &mip = *inf->Lmap;
            if (mip.wrapMode != ImageWrap::Repeat || mip.noFiltering)
                return fail("infinite light map with a wrap mode / filter outside the GPU hot path");
            const BlockedArray<RGBSpectrum> &lvl = *mip.pyramid[0];
            const int W = lvl.uSize(), H = lvl.vSize();
            const Distribution2D &d2 = *inf->distribution;
            const int nv = d2.pMarginal->Count(), nu = d2.pConditionalV[0]->Count();
            std::vector<float> &E = out->envFloats;
            E.push_back((float)W);
            E.push_back((float)H);
            E.push_back((float)nu);
            E.push_back((float)nv);
            E.push_back(inf->Power().y());
            E.push_back(0.f);
            E.push_back(0.f);
            E.push_back(0.f);
            float m[16];
            Builder::CopyMatrix(inf->LightToWorld.m, m);
            E.insert(E.end(), m, m + 16);
            Builder::CopyMatrix(inf->WorldToLight.m, m);
            E.insert(E.end(), m, m + 16);
            for (int v = 0; v < H; ++v)
                for (int u = 0; u < W; ++u) {
                    Float rgb[3];
                    lvl(u, v).ToRGB(rgb);
                    E.push_back(rgb[0]);
                    E.push_back(rgb[1]);
                    E.push_back(rgb[2]);
                }
            auto addDist = [&](const Distribution1D &d) {
                E.insert(E.end(), d.func.begin(), d.func.end());
                E.insert(E.end(), d.cdf.begin(), d.cdf.end());
                E.push_back(d.funcInt);
            };
            addDist(*d2.pMarginal);
            for (int v = 0; v < nv; ++v) addDist(*d2.pConditionalV[v]);
        } else
            return fail(
                "light type is outside the GPU hot path (supported: "
                "diffuse area lights on triangles/spheres, point, infinite)");
        out->lights.push_back(fl);
    }

    // ---- camera
    sptx_camera &fc = out->desc.camera;
    if (camera.CameraToWorld.actuallyAnimated)
        return fail("animated cameras are not supported");
    if (camera.medium) return fail("camera media are not supported");
    Builder::CopyMatrix(camera.CameraToWorld.startTransform->m,
                        fc.camera_to_world);
    fc.shutter_open = camera.shutterOpen;
    fc.shutter_close = camera.shutterClose;
    fc.film_diagonal = camera.film->diagonal;
    if (auto *pc = dynamic_cast<const PerspectiveCamera *>(&camera)) {
        fc.type = SPTX_CAMERA_PERSPECTIVE;
        Builder::CopyMatrix(pc->RasterToCamera.m, fc.raster_to_camera);
        fc.lens_radius = pc->lensRadius;
        fc.focal_distance = pc->focalDistance;
    } else if (auto *rc = dynamic_cast<const RealisticCamera *>(&camera)) {
        fc.type = SPTX_CAMERA_REALISTIC;
        fc.simple_weighting = rc->simpleWeighting ? 1 : 0;
        if (rc->noWeighting) fc.simple_weighting = 2;
        fc.ca_flag = rc->caFlag ? 1 : 0;
        for (const auto &e : rc->elementInterfaces) {
            out->lensElements.push_back(e.curvatureRadius);
            out->lensElements.push_back(e.thickness);
            out->lensElements.push_back(e.eta);
            out->lensElements.push_back(e.apertureRadius);
        }
        for (const auto &bd : rc->exitPupilBounds) {
            out->exitPupilBounds.push_back(bd.pMin.x);
            out->exitPupilBounds.push_back(bd.pMin.y);
            out->exitPupilBounds.push_back(bd.pMax.x);
            out->exitPupilBounds.push_back(bd.pMax.y);
        }
        fc.n_elements = (int)rc->elementInterfaces.size();
        fc.n_pupil_bounds = (int)rc->exitPupilBounds.size();
    } else
        return fail(
            "camera type is outside the GPU hot path (supported: "
            "perspective, realistic)");

    // ---- film
    const Film &film = *camera.film;
    sptx_film &ff = out->desc.film;
    ff.full_res[0] = film.fullResolution.x;
    ff.full_res[1] = film.fullResolution.y;
    ff.cropped[0] = film.croppedPixelBounds.pMin.x;
    ff.cropped[1] = film.croppedPixelBounds.pMin.y;
    ff.cropped[2] = film.croppedPixelBounds.pMax.x;
    ff.cropped[3] = film.croppedPixelBounds.pMax.y;
    Bounds2i sb = film.GetSampleBounds();
    ff.sample_bounds[0] = sb.pMin.x;
    ff.sample_bounds[1] = sb.pMin.y;
    ff.sample_bounds[2] = sb.pMax.x;
    ff.sample_bounds[3] = sb.pMax.y;
    ff.filter_radius[0] = film.filter->radius.x;
    ff.filter_radius[1] = film.filter->radius.y;
    static_assert(Film::filterTableWidth == SPTX_FILTER_TABLE_WIDTH, "");
    std::memcpy(ff.filter_table, film.filterTable, sizeof(ff.filter_table));
    ff.max_sample_luminance = film.maxSampleLuminance;
    ff.scale = film.scale;
    ff.diagonal = film.diagonal;
    ff.spectral_flag = film.spectralFlag ? 1 : 0;

    // ---- sampler
    sptx_sampler &fsm = out->desc.sampler;
    fsm.spp = (int)sampler.samplesPerPixel;
    if (auto *h = dynamic_cast<HaltonSampler *>(&sampler)) {
        fsm.type = SPTX_SAMPLER_HALTON;
        fsm.sample_at_center = h->sampleAtPixelCenter ? 1 : 0;
        fsm.base_scales[0] = h->baseScales[0];
        fsm.base_scales[1] = h->baseScales[1];
        fsm.base_exponents[0] = h->baseExponents[0];
        fsm.base_exponents[1] = h->baseExponents[1];
        fsm.sample_stride = h->sampleStride;
        fsm.mult_inverse[0] = h->multInverse[0];
        fsm.mult_inverse[1] = h->multInverse[1];
        if (!h->samples1DArraySizes.empty() || !h->samples2DArraySizes.empty())
            return fail("sampler array requests are not on the path");
        // Dimensions a path can consume: 5 camera dims + per bounce at most
        // 1+2+2 (light pick, uLight, uScattering) + 2 (BSDF) + 1 (RR).
        // SpectralPathIntegrator runs one path per band off the same stream.
        int dims = 5 + 8 * (std::max(params.maxDepth, 0) + 2) *
                           std::max(params.numCABands, 1);
        dims = std::min(std::max(dims, 64), (int)PrimeTableSize);
        fsm.n_dims = dims;
        out->primes.assign(Primes, Primes + dims);
        out->primeSums.assign(PrimeSums, PrimeSums + dims);
        size_t nperm = (size_t)PrimeSums[dims - 1] + Primes[dims - 1];
        const std::vector<uint16_t> &perms =
            HaltonSampler::radicalInversePermutations;
        if (perms.size() < nperm) return fail("Halton permutations missing");
        out->perms.assign(perms.begin(), perms.begin() + nperm);
    } else if (auto *sb = dynamic_cast<SobolSampler *>(&sampler)) {
        // SobolSampler (src/samplers/sobol.cpp): index from the pixel and the
        // sample number through the two 64-bit van der Corput / Sobol' matrices
        // of this resolution, values from SobolMatrices32.
        fsm.type = SPTX_SAMPLER_SOBOL;
        fsm.log2_resolution = sb->log2Resolution;
        if (!sb->samples1DArraySizes.empty() || !sb->samples2DArraySizes.empty())
            return fail("sampler array requests are not on the path");
        int dims = 5 + 8 * (std::max(params.maxDepth, 0) + 2) *
                           std::max(params.numCABands, 1);
        dims = std::min(std::max(dims, 64), (int)NumSobolDimensions);
        fsm.n_dims = dims;
        out->primes.resize((size_t)dims * SobolMatrixSize);
        std::memcpy(out->primes.data(), SobolMatrices32,
                    out->primes.size() * sizeof(int32_t));
        out->primeSums.assign((size_t)4 * SobolMatrixSize, 0);
        if (sb->log2Resolution > 0) {
            std::memcpy(out->primeSums.data(),
                        VdCSobolMatrices[sb->log2Resolution - 1],
                        SobolMatrixSize * sizeof(uint64_t));
            std::memcpy(out->primeSums.data() + 2 * SobolMatrixSize,
                        VdCSobolMatricesInv[sb->log2Resolution - 1],
                        SobolMatrixSize * sizeof(uint64_t));
        }
    } else if (dynamic_cast<RandomSampler *>(&sampler)) {
        fsm.type = SPTX_SAMPLER_RANDOM;
    } else if (auto *zt = dynamic_cast<ZeroTwoSequenceSampler *>(&sampler)) {
        // ZeroTwoSequenceSampler (src/samplers/zerotwosequence.cpp): samplesPerPixel is already
        // rounded up to a power of two; the first nSampledDimensions 1D / 2D dimensions of a
        // pixel's samples are (0,2)-sequence tables, the rest comes from the tile's RNG
        if (!zt->samples1DArraySizes.empty() || !zt->samples2DArraySizes.empty())
            return fail("sampler array requests are not on the path");
        fsm.type = SPTX_SAMPLER_ZEROTWO;
        fsm.n_dims = (int)zt->samples1D.size();
    } else
        return fail(
            "sampler type is outside the GPU hot path (supported: halton, "
            "sobol, random, 02sequence)");

    // ---- integrator
    sptx_integrator &fi = out->desc.integrator;
    fi.max_depth = params.maxDepth;
    fi.rr_threshold = params.rrThreshold;
    for (int i = 0; i < 4; ++i) fi.pixel_bounds[i] = params.pixelBounds[i];
    fi.num_ca_bands = params.numCABands;
    out->desc.metadata_strategy = (uint32_t)params.metadataStrategy;
    const bool tileSerial = out->desc.sampler.type == SPTX_SAMPLER_RANDOM ||
                            out->desc.sampler.type == SPTX_SAMPLER_ZEROTWO;
    if (params.metadataStrategy && tileSerial)
        return fail(
            "Integrator \"metadata\" is on the GPU hot path for the halton "
            "and sobol samplers only");
    if (params.numCABands > 1 && tileSerial)
        return fail(
            "spectralpath with numCABands > 1 is on the GPU hot path for the "
            "halton and sobol samplers only");
    if (params.lightSampleStrategy == "uniform")
        fi.light_strategy = SPTX_LIGHTS_UNIFORM;
    else if (params.lightSampleStrategy == "power")
        fi.light_strategy = SPTX_LIGHTS_POWER;
    else  // "spatial", and unknown names (src/core/lightdistrib.cpp:60-66)
        fi.light_strategy = SPTX_LIGHTS_SPATIAL;

    // ---- colour matching (SampledSpectrum::y, src/core/spectrum.h:415-421)
    for (int i = 0; i < SPTX_NBINS; ++i)
        out->desc.cie_y[i] = SampledSpectrum::Y[i];
    out->desc.cie_y_integral = CIE_Y_integral;

    out->Finish();
    return true;
}

}  // namespace sptx_host
