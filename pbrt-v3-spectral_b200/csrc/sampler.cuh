// sampler.cuh -- HaltonSampler reproduced on the device.
//
// A sample value is a pure function of (pixel, sample number, dimension):
//   index = offsetForPixel + sampleNum * sampleStride   (samplers/halton.cpp:96-115)
//   value = SampleDimension(index, dim)                 (samplers/halton.cpp:117-127)
// so every path owns its own stream and streams match the reference bit for bit.
#pragma once
#include "scene.cuh"

namespace sptx {

SPTX_HD uint32_t reverse_bits32(uint32_t n) {
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ffu) << 8) | ((n & 0xff00ff00u) >> 8);
    n = ((n & 0x0f0f0f0fu) << 4) | ((n & 0xf0f0f0f0u) >> 4);
    n = ((n & 0x33333333u) << 2) | ((n & 0xccccccccu) >> 2);
    n = ((n & 0x55555555u) << 1) | ((n & 0xaaaaaaaau) >> 1);
    return n;
}
SPTX_HD uint64_t reverse_bits64(uint64_t n) {
    uint64_t n0 = reverse_bits32((uint32_t)n);
    uint64_t n1 = reverse_bits32((uint32_t)(n >> 32));
    return (n0 << 32) | n1;
}

// InverseRadicalInverse<base>, core/lowdiscrepancy.h:82-91.
SPTX_HD uint64_t inverse_radical_inverse(uint64_t inverse, int base, int nDigits) {
    uint64_t index = 0;
    for (int i = 0; i < nDigits; ++i) {
        uint64_t digit = inverse % (uint64_t)base;
        inverse /= (uint64_t)base;
        index = index * (uint64_t)base + digit;
    }
    return index;
}

// Mod for ints, core/pbrt.h:316-320.
SPTX_HD int mod_i(int a, int b) {
    int r = a - (a / b) * b;
    return (r < 0) ? r + b : r;
}

// HaltonSampler::GetIndexForSample's per-pixel part (samplers/halton.cpp:96-112).
SPTX_HD uint64_t halton_pixel_offset(const sptx_sampler &s, int px, int py) {
    const int kMaxResolution = 128;
    uint64_t off = 0;
    if (s.sample_stride > 1) {
        int pm0 = mod_i(px, kMaxResolution), pm1 = mod_i(py, kMaxResolution);
        uint64_t d0 = inverse_radical_inverse((uint64_t)pm0, 2, s.base_exponents[0]);
        off += d0 * (uint64_t)(s.sample_stride / s.base_scales[0]) * (uint64_t)s.mult_inverse[0];
        uint64_t d1 = inverse_radical_inverse((uint64_t)pm1, 3, s.base_exponents[1]);
        off += d1 * (uint64_t)(s.sample_stride / s.base_scales[1]) * (uint64_t)s.mult_inverse[1];
        off %= (uint64_t)s.sample_stride;
    }
    return off;
}

// RadicalInverseSpecialized<base>, core/lowdiscrepancy.cpp:389-403.
SPTX_HD float radical_inverse_base(uint64_t a, uint32_t base) {
    const float invBase = 1.f / (float)base;
    uint64_t reversed = 0;
    float invBaseN = 1.f;
    if (a >> 32) {
        while (a >> 32) {
            uint64_t next = a / base;
            uint64_t digit = a - next * base;
            reversed = reversed * base + digit;
            invBaseN *= invBase;
            a = next;
        }
    }
    uint32_t a32 = (uint32_t)a;
    while (a32) {
        uint32_t next = a32 / base;
        uint32_t digit = a32 - next * base;
        reversed = reversed * base + digit;
        invBaseN *= invBase;
        a32 = next;
    }
    float r = (float)reversed * invBaseN;
    return (kOneMinusEpsilon < r) ? kOneMinusEpsilon : r;
}

// ScrambledRadicalInverseSpecialized<base>, core/lowdiscrepancy.cpp:405-424.
// Same integer digits and the same float operations; the 32-bit quotient is
// found with a reciprocal multiply (magic = floor(2^32 / base), one fix-up)
// instead of a hardware-emulated division.
SPTX_DEV float scrambled_radical_inverse(const uint16_t *perm, uint64_t a, uint32_t base) {
    const float invBase = 1.f / (float)base;
    uint64_t reversed = 0;
    float invBaseN = 1.f;
    while (a >> 32) {
        uint64_t next = a / base;
        uint64_t digit = a - next * base;
        reversed = reversed * base + __ldg(perm + digit);
        invBaseN *= invBase;
        a = next;
    }
    uint32_t a32 = (uint32_t)a;
    const uint32_t magic = 0xffffffffu / base;  // floor((2^32 - 1) / base) <= 2^32 / base
    while (a32) {
        uint32_t next = __umulhi(a32, magic);    // in {q - 1, q}  (a32 * (1/base - magic/2^32) < 2)
        uint32_t digit = a32 - next * base;
        if (digit >= base) {
            ++next;
            digit -= base;
        }
        if (digit >= base) {  // magic may undershoot twice when base divides 2^32 - 1 oddly
            ++next;
            digit -= base;
        }
        reversed = reversed * base + __ldg(perm + digit);
        invBaseN *= invBase;
        a32 = next;
    }
    float r = invBaseN * ((float)reversed + invBase * (float)(int)__ldg(perm) / (1.f - invBase));
    return (kOneMinusEpsilon < r) ? kOneMinusEpsilon : r;
}

// ---- SobolSampler (samplers/sobol.cpp, core/lowdiscrepancy.h:229-274) -------------------
constexpr int kSobolMatrixSize = 52;
// SobolIntervalToIndex: the index of sample number `frame` of pixel p (relative to the sample
// bounds) in the global Sobol' sequence.
SPTX_DEV uint64_t sobol_interval_to_index(const DSampler &H, uint64_t frame, int px, int py) {
    const uint32_t m = (uint32_t)H.s.log2_resolution;
    if (m == 0) return 0;
    const uint64_t *vdc = reinterpret_cast<const uint64_t *>(H.prime_sums);
    const uint64_t *vdcInv = vdc + kSobolMatrixSize;
    const uint32_t m2 = m << 1;
    uint64_t index = frame << m2;
    uint64_t delta = 0;
    for (int c = 0; frame; frame >>= 1, ++c)
        if (frame & 1) delta ^= __ldg(vdc + c);
    uint64_t b = (((uint64_t)((uint32_t)px) << m) | ((uint32_t)py)) ^ delta;
    for (int c = 0; b; b >>= 1, ++c)
        if (b & 1) index ^= __ldg(vdcInv + c);
    return index;
}
// SobolSampler::SampleDimension = SobolSampleFloat + the pixel remap of dimensions 0 and 1.
// The pixel the index belongs to is read off the sample's own top m bits (that is what
// SobolIntervalToIndex arranges), so a path needs to carry nothing but its index.
SPTX_NOINL float sobol_sample(const DSampler &H, uint64_t index, int dim) {
    const int d = dim < H.s.n_dims ? dim : H.s.n_dims - 1;  // callers flag the overflow
    const uint32_t *M = reinterpret_cast<const uint32_t *>(H.primes) + (size_t)d * kSobolMatrixSize;
    uint32_t v = 0;
    for (int i = 0; index != 0; index >>= 1, ++i)
        if (index & 1) v ^= __ldg(M + i);
    float s = (float)v * 0x1p-32f;
    s = (kOneMinusEpsilon < s) ? kOneMinusEpsilon : s;
    if (d == 0 || d == 1) {
        const int m = H.s.log2_resolution;
        const int resolution = 1 << m;
        s = s * (float)resolution + (float)H.sb[d];
        const int cur = H.sb[d] + (m > 0 ? (int)(v >> (32 - m)) : 0);
        s = s - (float)cur;
        s = (s < 0.f) ? 0.f : ((s > kOneMinusEpsilon) ? kOneMinusEpsilon : s);
    }
    return s;
}

// HaltonSampler::SampleDimension, samplers/halton.cpp:117-127.
SPTX_NOINL float halton_sample(const DSampler &H, uint64_t index, int dim) {
    if (H.s.type == SPTX_SAMPLER_SOBOL) return sobol_sample(H, index, dim);
    if (H.s.sample_at_center && (dim == 0 || dim == 1)) return 0.5f;
    if (dim == 0) {
        // RadicalInverse(0, a) = ReverseBits64(a) * 0x1p-64 in DOUBLE, rounded
        // to float on return (core/lowdiscrepancy.cpp:430-437).
        uint64_t a = index >> H.s.base_exponents[0];
        return (float)((double)reverse_bits64(a) * 0x1p-64);
    } else if (dim == 1) {
        return radical_inverse_base(index / (uint64_t)H.s.base_scales[1], 3u);
    }
    int d = dim < H.s.n_dims ? dim : H.s.n_dims - 1;  // callers flag the overflow
    return scrambled_radical_inverse(H.perms + __ldg(H.prime_sums + d), index,
                                     (uint32_t)__ldg(H.primes + d));
}

// N consecutive scrambled dimensions of one index at once.  The digit loops are serial
// chains (multiply-high, permutation lookup, multiply-add); interleaving the chains of the
// dimensions a vertex consumes together overlaps their latencies.  Same integer digits and the
// same float operations per dimension as scrambled_radical_inverse; the reversed digits stay in
// 32 bits, which holds while index * base < 2^32 (checked, else the scalar path).
template <int N>
SPTX_DEV void halton_n(const DSampler &H, uint64_t index, int dim, float *out) {
    bool fast = H.s.type == SPTX_SAMPLER_HALTON && dim >= 2 && dim + N <= H.s.n_dims && (index >> 32) == 0;
    uint4 tab[N];
    if (fast) {
#pragma unroll
        for (int j = 0; j < N; ++j) {
            tab[j] = __ldg(H.dim_tab + dim + j);
            if ((uint32_t)index > tab[j].y) fast = false;  // index * base would leave 32 bits
        }
    }
    if (!fast) {
#pragma unroll
        for (int j = 0; j < N; ++j) out[j] = halton_sample(H, index, dim + j);
        return;
    }
    // (fast path: index * base < 2^32, so a < 2^31 and the round-down magic undershoots the
    //  quotient by at most one: a * (1/base - magic/2^32) <= a / 2^32 < 1/2 -- one fix-up.  The
    //  bases of consecutive dimensions ascend, so dimension `dim` has the most digits: the loop
    //  runs while ITS digits last, the others drop out through their predicate.)
    uint32_t a[N], rev[N];
    float invN[N];
    const uint16_t *perm[N];
#pragma unroll
    for (int j = 0; j < N; ++j) {
        a[j] = (uint32_t)index;
        rev[j] = 0;
        invN[j] = 1.f;
        perm[j] = H.perms + tab[j].z;
    }
    while (a[0]) {
#pragma unroll
        for (int j = 0; j < N; ++j) {
            if (j == 0 || a[j]) {
                const uint32_t base = tab[j].x;
                uint32_t next = __umulhi(a[j], tab[j].y);
                uint32_t digit = a[j] - next * base;
                if (digit >= base) {
                    ++next;
                    digit -= base;
                }
                rev[j] = rev[j] * base + __ldg(perm[j] + digit);
                invN[j] *= __uint_as_float(tab[j].w);
                a[j] = next;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < N; ++j) {
        float r = invN[j] * ((float)rev[j] + __ldg(H.dim_tail + dim + j));
        out[j] = (kOneMinusEpsilon < r) ? kOneMinusEpsilon : r;
    }
}

// Index of sample number `sampleNum` of pixel (px, py): GlobalSampler::GetIndexForSample.
SPTX_DEV uint64_t sample_index(const DSampler &H, int px, int py, int sampleNum) {
    if (H.s.type == SPTX_SAMPLER_SOBOL)
        return sobol_interval_to_index(H, (uint64_t)sampleNum, px - H.sb[0], py - H.sb[1]);
    const uint64_t off = H.pixel_offset[(size_t)(py - H.sb[1]) * (H.sb[2] - H.sb[0]) + (px - H.sb[0])];
    return off + (uint64_t)sampleNum * (uint64_t)H.s.sample_stride;
}

// Per-path view of the sampler: GlobalSampler::Get1D/Get2D with no array
// requests (core/sampler.cpp:180-195): dimensions are consumed sequentially.
struct SampleStream {
    uint64_t index;
    int dim;
};
SPTX_DEV float get1d(const DSampler &H, SampleStream &st) { return halton_sample(H, st.index, st.dim++); }
SPTX_DEV V2 get2d(const DSampler &H, SampleStream &st) {
    float v[2];
    halton_n<2>(H, st.index, st.dim, v);
    st.dim += 2;
    return V2{v[0], v[1]};
}

}  // namespace sptx
