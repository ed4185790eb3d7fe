// sceneio.cpp -- flattened-scene files: the sptx_scene_desc header followed by
// its arrays, each padded to 16 bytes.  Written by the host layer
// (SPTX_DUMP_SCENE) and read by the Python harness / benchmarks.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "sptx.h"

namespace {

const char kMagic[8] = {'S', 'P', 'T', 'X', 'S', 'C', 'N', '1'};

struct Section {
    const void **ptr;
    size_t bytes;
};

// The arrays of a descriptor in file order, with their byte sizes.
std::vector<Section> sections(sptx_scene_desc *d) {
    std::vector<Section> s;
#define SEC(field, count, elem) s.push_back(Section{(const void **)&d->field, (size_t)(count) * (elem)})
    SEC(nodes, d->n_nodes, sizeof(sptx_bvh_node));
    SEC(bvhs, d->n_bvhs, sizeof(sptx_bvh));
    SEC(prims, d->n_prims, sizeof(sptx_prim));
    SEC(meshes, d->n_meshes, sizeof(sptx_mesh));
    SEC(triangles, d->n_triangles, sizeof(sptx_triangle));
    SEC(positions, d->n_positions, 3 * sizeof(float));
    SEC(normals, d->n_normals, 3 * sizeof(float));
    SEC(tangents, d->n_tangents, 3 * sizeof(float));
    SEC(uvs, d->n_uvs, 2 * sizeof(float));
    SEC(spheres, d->n_spheres, sizeof(sptx_sphere));
    SEC(instances, d->n_instances, sizeof(sptx_instance));
    SEC(materials, d->n_materials, sizeof(sptx_material));
    SEC(lobes, d->n_lobes, sizeof(sptx_lobe));
    SEC(spectra, d->n_spectra, SPTX_SPECTRUM_STRIDE * sizeof(float));
    SEC(lights, d->n_lights, sizeof(sptx_light));
    SEC(lens_elements, d->n_lens_floats, sizeof(float));
    SEC(exit_pupil_bounds, d->n_pupil_floats, sizeof(float));
    SEC(primes, d->n_primes, sizeof(int32_t));
    SEC(prime_sums, d->n_prime_sums, sizeof(int32_t));
    SEC(perms, d->n_perms, sizeof(uint16_t));
    SEC(env_floats, d->n_env_floats, sizeof(float));
    SEC(tex_floats, d->n_tex_floats, sizeof(float));
#undef SEC
    return s;
}

std::string g_io_error;

}  // namespace

extern "C" const char *sptx_io_last_error(void) { return g_io_error.c_str(); }

extern "C" int sptx_scene_save(const char *path, const sptx_scene_desc *scene) {
    if (!path || !scene) return SPTX_EINVAL;
    FILE *f = fopen(path, "wb");
    if (!f) {
        g_io_error = std::string("cannot open ") + path;
        return SPTX_EIO;
    }
    sptx_scene_desc copy = *scene;
    std::vector<Section> src = sections(&copy);
    std::vector<const void *> data;
    for (Section &s : src) {
        data.push_back(*s.ptr);
        *s.ptr = nullptr;  // no host addresses in the file
    }
    uint64_t headerBytes = sizeof(sptx_scene_desc);
    bool ok = fwrite(kMagic, 1, 8, f) == 8 && fwrite(&headerBytes, 8, 1, f) == 1 &&
              fwrite(&copy, sizeof(copy), 1, f) == 1;
    static const char pad[16] = {0};
    size_t pos = 16 + sizeof(copy);
    if (ok && pos % 16) {
        ok = fwrite(pad, 1, 16 - pos % 16, f) == 16 - pos % 16;
        pos += 16 - pos % 16;
    }
    for (size_t i = 0; ok && i < src.size(); ++i) {
        if (src[i].bytes && data[i]) ok = fwrite(data[i], 1, src[i].bytes, f) == src[i].bytes;
        size_t r = src[i].bytes % 16;
        if (ok && r) ok = fwrite(pad, 1, 16 - r, f) == 16 - r;
    }
    fclose(f);
    if (!ok) {
        g_io_error = std::string("short write to ") + path;
        return SPTX_EIO;
    }
    return SPTX_OK;
}

extern "C" int sptx_scene_load(const char *path, sptx_scene_desc **out) {
    if (!path || !out) return SPTX_EINVAL;
    *out = nullptr;
    FILE *f = fopen(path, "rb");
    if (!f) {
        g_io_error = std::string("cannot open ") + path;
        return SPTX_EIO;
    }
    fseek(f, 0, SEEK_END);
    long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    // One allocation: descriptor first, the file image (arrays) behind it; sptx_scene_free frees it.
    // The file is read kFront bytes in, so that the full-size descriptor at the front never
    // overlaps the arrays of a file whose own descriptor was shorter.
    const size_t kFront = 64;
    char *base = (char *)malloc((size_t)size + kFront + 16);
    if (!base) {
        fclose(f);
        return SPTX_ENOMEM;
    }
    char *blob = base + kFront;
    bool ok = fread(blob, 1, (size_t)size, f) == (size_t)size;
    fclose(f);
    uint64_t headerBytes = 0;
    if (ok && size >= 16) memcpy(&headerBytes, blob + 8, 8);
    // (files written before the trailing env_floats / tex_floats members existed have a descriptor
    //  16 or 32 bytes shorter)
    const uint64_t kOldHeader = sizeof(sptx_scene_desc) - 32;
    if (!ok || size < (long)(16 + kOldHeader) || memcmp(blob, kMagic, 8) != 0 ||
        (headerBytes != sizeof(sptx_scene_desc) && headerBytes != kOldHeader && headerBytes != kOldHeader + 16)) {
        free(base);
        g_io_error = std::string("not a flattened-scene file of this ABI: ") + path;
        return SPTX_EINVAL;
    }
    // The descriptor goes to the front of the allocation.
    sptx_scene_desc *d = (sptx_scene_desc *)base;
    sptx_scene_desc tmp;
    memset(&tmp, 0, sizeof(tmp));
    memcpy(&tmp, blob + 16, (size_t)headerBytes);
    size_t pos = 16 + (size_t)headerBytes;
    if (pos % 16) pos += 16 - pos % 16;
    // the counts come from the file: none may be negative (a negative count would wrap the size
    // arithmetic below past the truncation check)
    const int64_t counts[] = {tmp.n_nodes, tmp.n_bvhs, tmp.n_prims, tmp.n_meshes, tmp.n_triangles,
                              tmp.n_positions, tmp.n_normals, tmp.n_tangents, tmp.n_uvs, tmp.n_spheres,
                              tmp.n_instances, tmp.n_materials, tmp.n_lobes, tmp.n_spectra, tmp.n_lights,
                              tmp.n_lens_floats, tmp.n_pupil_floats, tmp.n_primes, tmp.n_prime_sums,
                              tmp.n_perms, tmp.n_env_floats, tmp.n_tex_floats};
    for (int64_t c : counts)
        if (c < 0 || c > (int64_t)size) {  // every element is at least one byte
            free(base);
            g_io_error = std::string("corrupt scene file (array count out of range): ") + path;
            return SPTX_EINVAL;
        }
    std::vector<Section> secs = sections(&tmp);
    for (Section &s : secs) {
        if (s.bytes > (size_t)size || pos > (size_t)size - s.bytes) {
            free(base);
            g_io_error = std::string("truncated scene file: ") + path;
            return SPTX_EINVAL;
        }
        *s.ptr = s.bytes ? (const void *)(blob + pos) : nullptr;
        pos += s.bytes;
        if (s.bytes % 16) pos += 16 - s.bytes % 16;
    }
    static_assert(sizeof(sptx_scene_desc) <= 64 + 16 + sizeof(sptx_scene_desc) - 32, "descriptor overlaps the arrays");
    memcpy(d, &tmp, sizeof(tmp));
    *out = d;
    return SPTX_OK;
}

extern "C" void sptx_scene_free(sptx_scene_desc *scene) { free(scene); }
