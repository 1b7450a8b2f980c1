// Halton sampler on the device: a pure function of (pixel, sample number,
// dimension), exactly the stream the reference's HaltonSampler hands to
// PathIntegrator::Li.
//   HaltonSampler ctor / GetIndexForSample / SampleDimension   samplers/halton.cpp:65-127
//   GlobalSampler::Get1D / Get2D                                core/sampler.cpp:181-196
//   RadicalInverse / ScrambledRadicalInverse                    core/lowdiscrepancy.cpp:390-433, 2506-
//   InverseRadicalInverse                                       core/lowdiscrepancy.h:83-91
// The reference unrolls one radical inverse per prime base with a compile-time
// divisor; here the base comes from a table and the quotient is taken with a
// 32-bit divide whenever the running index fits (it does for every BASELINE
// config: stride 31104 x 4096 spp < 2^32), which yields the same digits.
#ifndef PG_DEVICE_SAMPLER_H
#define PG_DEVICE_SAMPLER_H

#include "../common/pg_math.h"

namespace pg {

constexpr int kHaltonMaxResolution = 128;  // halton.cpp:42

constexpr int kHaltonPrimes = 1000;  // entries of the prime tables (= kPrimeTableSize, common/halton_tables.h)
constexpr int kArrayStartDim = 5;  // GlobalSampler::arrayStartDim (sampler.h:122) -- although this fork's camera sample uses 6 dimensions

// Everything SampleDimension needs about one dimension >= 2, in one 32-byte record (two 16-byte loads instead of five
// scattered ones, and no division or bit scan per sample): the prime, its 32-bit division magic and shift (below), where
// its digit permutation starts, and the two per-base constants of ScrambledRadicalInverseSpecialized
// (lowdiscrepancy.cpp:406-423) -- invBase = 1 / base and the tail invBase * perm[0] / (1 - invBase), both computed on the
// host by the same correctly rounded float operations in the same order, so their bits are the device's.
struct HaltonDim {
    unsigned base, magic32;
    int shift, permOffset;
    float invBase, tail;
    unsigned pad[2];
};

struct HaltonParams {
    int baseScales[2], baseExponents[2];
    int sampleStride;
    int multInverse[2];
    int pixelOffsetMul[2];         // ((sampleStride / baseScales[i]) * multInverse[i]) mod sampleStride
    int sampleAtPixelCenter;
    int arrayEndDim;               // GlobalSampler: dimensions [kArrayStartDim, arrayEndDim) hold the sample arrays (sampler.cpp:126-148)
    const int *primes;             // 1000 entries
    const int *primeSums;          // 1000 entries
    const unsigned long long *primeMagic;  // 2 x 1000 entries (halton_tables.h:primeMagics): 64-bit magics, then 32-bit ones
    const unsigned short *perms;   // 3,682,913 entries
    const HaltonDim *dims;         // 1000 entries (fillHaltonDims)
};

PG_HD int modPos(int a, int b) {  // Mod(), pbrt.h
    int r = a - (a / b) * b;
    return r < 0 ? r + b : r;
}

// GetIndexForSample(0) for a pixel: the CRT offset of the pixel inside the
// 128-pixel Halton tile (halton.cpp:96-116).
PG_HD long long haltonPixelOffset(const HaltonParams &h, int px, int py) {
    unsigned long long offset = 0;
    if (h.sampleStride > 1) {
        int pm[2] = {modPos(px, kHaltonMaxResolution), modPos(py, kHaltonMaxResolution)};
        for (int i = 0; i < 2; ++i) {
            unsigned base = i == 0 ? 2u : 3u;
            unsigned inverse = (unsigned)pm[i], index = 0;
            for (int k = 0; k < h.baseExponents[i]; ++k) {
                unsigned digit = inverse % base;
                inverse /= base;
                index = index * base + digit;
            }
            // halton.cpp:110-112 adds index * (stride / scale_i) * multInverse_i and reduces mod stride at the end; the
            // factor is reduced mod stride beforehand (makeHaltonParams), which leaves the residue unchanged and keeps
            // every product below 2^32 -- a 32-bit remainder per sample instead of two divisions and a 64-bit one.
            offset += (unsigned long long)(index * (unsigned)h.pixelOffsetMul[i]);
        }
        offset = (unsigned long long)((unsigned)offset % (unsigned)h.sampleStride);
    }
    return (long long)offset;
}

PG_HD unsigned long long reverseBits64(unsigned long long n) {
#if defined(__CUDA_ARCH__)
    return __brevll(n);
#else
    unsigned long long r = 0;
    for (int i = 0; i < 64; ++i) r |= ((n >> i) & 1ull) << (63 - i);
    return r;
#endif
}

// RadicalInverseSpecialized<base>, lowdiscrepancy.cpp:390-404
PG_HD float radicalInverseBase(unsigned base, unsigned long long a) {
    const float invBase = (float)1 / (float)base;
    unsigned long long reversedDigits = 0;
    float invBaseN = 1;
    while (a > 0xffffffffull) {
        unsigned long long next = a / base;
        unsigned long long digit = a - next * base;
        reversedDigits = reversedDigits * base + digit;
        invBaseN *= invBase;
        a = next;
    }
    unsigned a32 = (unsigned)a;
    while (a32) {
        unsigned next = a32 / base;
        unsigned digit = a32 - next * base;
        reversedDigits = reversedDigits * base + digit;
        invBaseN *= invBase;
        a32 = next;
    }
    return minf((float)reversedDigits * invBaseN, kOneMinusEpsilon);
}

PG_HD unsigned long long mulhi64(unsigned long long a, unsigned long long b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (unsigned long long)(((unsigned __int128)a * (unsigned __int128)b) >> 64);
#endif
}

// ScrambledRadicalInverseSpecialized<base>, lowdiscrepancy.cpp:406-423.  The
// quotient a / base is taken as mulhi64(a, magic) once a fits 32 bits: with
// magic = floor(2^64 / base) + 1 the product overshoots a / base by less than
// 2^-32, which cannot carry past the fraction's distance (>= 1 / base) to the
// next integer -- so the digits are the reference's.
//
// Indices below 2^31 (every BASELINE configuration: stride 31104 x 4096 spp = 1.3e8) take a shorter loop:
// with s = floor(log2 base) and m = ceil(2^(32+s) / base) < 2^32, e = m * base - 2^(32+s) lies in (0, base),
// so a * m / 2^(32+s) = a / base + a * e / (base * 2^(32+s)), and the excess is below 1 / base whenever
// a * e < 2^(32+s) -- true for a < 2^31 because e < base < 2^(s+1).  Hence mulhi32(a, m) >> s = a / base.
// All digits but the most significant one are accumulated in 32 bits (before the last step the reversed
// digits are below base^(k-1) <= a < 2^31); the last step widens to the reference's 64-bit value.
PG_HD unsigned mulhi32(unsigned a, unsigned b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (unsigned)(((unsigned long long)a * (unsigned long long)b) >> 32);
#endif
}
PG_HD int floorLog2u(unsigned v) {
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)v);
#else
    return 31 - __builtin_clz(v);
#endif
}
// the short loop with the dimension's constants from its record
PG_HD float scrambledRadicalInverseSmall(const HaltonDim &d, const unsigned short *perm, unsigned a32) {
    const unsigned base = d.base, magic32 = d.magic32;
    const int s = d.shift;
    const float invBase = d.invBase;
    unsigned long long reversedDigits = 0;
    float invBaseN = 1;
    unsigned rev = 0;
    if (a32) {
        for (;;) {
            unsigned next = mulhi32(a32, magic32) >> s;
            unsigned digit = a32 - next * base;
            unsigned pd = PG_LD(perm + digit);
            invBaseN *= invBase;
            if (next == 0) {
                reversedDigits = (unsigned long long)rev * base + pd;
                break;
            }
            rev = rev * base + pd;
            a32 = next;
        }
    }
    return minf(invBaseN * ((float)reversedDigits + d.tail), kOneMinusEpsilon);
}
PG_HD float scrambledRadicalInverseBase(unsigned base, unsigned long long magic, unsigned magic32, const unsigned short *perm,
                                        unsigned long long a) {
    const float invBase = (float)1 / (float)base;
    unsigned long long reversedDigits = 0;
    float invBaseN = 1;
    if (a <= 0x7fffffffull) {
        unsigned a32 = (unsigned)a, rev = 0;
        if (a32) {
            const int s = floorLog2u(base);
            for (;;) {
                unsigned next = mulhi32(a32, magic32) >> s;
                unsigned digit = a32 - next * base;
                unsigned pd = PG_LD(perm + digit);
                invBaseN *= invBase;
                if (next == 0) {
                    reversedDigits = (unsigned long long)rev * base + pd;
                    break;
                }
                rev = rev * base + pd;
                a32 = next;
            }
        }
        return minf(invBaseN * ((float)reversedDigits + invBase * (float)PG_LD(perm) / (1 - invBase)), kOneMinusEpsilon);
    }
    while (a > 0xffffffffull) {
        unsigned long long next = a / base;
        unsigned long long digit = a - next * base;
        reversedDigits = reversedDigits * base + PG_LD(perm + digit);
        invBaseN *= invBase;
        a = next;
    }
    unsigned a32 = (unsigned)a;
    while (a32) {
        unsigned next = (unsigned)mulhi64((unsigned long long)a32, magic);
        unsigned digit = a32 - next * base;
        reversedDigits = reversedDigits * base + PG_LD(perm + digit);
        invBaseN *= invBase;
        a32 = next;
    }
    return minf(invBaseN * ((float)reversedDigits + invBase * (float)PG_LD(perm) / (1 - invBase)), kOneMinusEpsilon);
}

// HaltonSampler::SampleDimension, halton.cpp:118-127
PG_NI_SAMPLER float haltonSample(const HaltonParams &h, long long index, int dim) {
    if (h.sampleAtPixelCenter && (dim == 0 || dim == 1)) return 0.5f;
    if (dim == 0) {
        unsigned long long a = (unsigned long long)(index >> h.baseExponents[0]);
        return (float)((double)reverseBits64(a) * 0x1p-64);  // RadicalInverse case 0
    }
    if (dim == 1) return radicalInverseBase(3u, (unsigned long long)(index / h.baseScales[1]));
    if ((unsigned long long)index <= 0x7fffffffull) {  // every BASELINE configuration
#if defined(__CUDA_ARCH__)
        const uint4 q0 = __ldg(reinterpret_cast<const uint4 *>(h.dims + dim));
        const uint2 q1 = __ldg(reinterpret_cast<const uint2 *>(h.dims + dim) + 2);
        HaltonDim d;
        d.base = q0.x, d.magic32 = q0.y, d.shift = (int)q0.z, d.permOffset = (int)q0.w;
        d.invBase = __uint_as_float(q1.x), d.tail = __uint_as_float(q1.y);
#else
        const HaltonDim d = h.dims[dim];
#endif
        return scrambledRadicalInverseSmall(d, h.perms + d.permOffset, (unsigned)index);
    }
    unsigned base = (unsigned)PG_LD(h.primes + dim);
    return scrambledRadicalInverseBase(base, PG_LD(h.primeMagic + dim), (unsigned)PG_LD(h.primeMagic + kHaltonPrimes + dim),
                                       h.perms + PG_LD(h.primeSums + dim), (unsigned long long)index);
}

// HaltonSampler's constructor (halton.cpp:65-93): scales/exponents that cover
// min(resolution, 128) in bases 2 and 3, the stride and the CRT inverses.
// Host side; the table pointers are filled in by the caller.
inline void haltonExtendedGCD(unsigned long long a, unsigned long long b, long long *x, long long *y) {
    if (b == 0) {
        *x = 1;
        *y = 0;
        return;
    }
    long long d = (long long)(a / b), xp, yp;
    haltonExtendedGCD(b, a % b, &xp, &yp);
    *x = yp;
    *y = xp - (d * yp);
}
inline HaltonParams makeHaltonParams(const int sampleBounds[4], int sampleAtPixelCenter) {
    HaltonParams h;
    int res[2] = {sampleBounds[2] - sampleBounds[0], sampleBounds[3] - sampleBounds[1]};
    for (int i = 0; i < 2; ++i) {
        int base = (i == 0) ? 2 : 3;
        int scale = 1, exp = 0;
        int lim = res[i] < kHaltonMaxResolution ? res[i] : kHaltonMaxResolution;
        while (scale < lim) {
            scale *= base;
            ++exp;
        }
        h.baseScales[i] = scale;
        h.baseExponents[i] = exp;
    }
    h.sampleStride = h.baseScales[0] * h.baseScales[1];
    for (int i = 0; i < 2; ++i) {
        long long x, y, n = h.baseScales[i];
        haltonExtendedGCD((unsigned long long)h.baseScales[1 - i], (unsigned long long)n, &x, &y);
        long long m = x - (x / n) * n;  // Mod(x, n)
        h.multInverse[i] = (int)(m < 0 ? m + n : m);
    }
    for (int i = 0; i < 2; ++i)
        h.pixelOffsetMul[i] = (int)(((unsigned long long)(h.sampleStride / h.baseScales[i]) * (unsigned long long)h.multInverse[i]) %
                                    (unsigned long long)h.sampleStride);
    h.sampleAtPixelCenter = sampleAtPixelCenter;
    h.arrayEndDim = kArrayStartDim;  // no sample arrays requested
    h.primes = nullptr;
    h.primeSums = nullptr;
    h.primeMagic = nullptr;
    h.perms = nullptr;
    h.dims = nullptr;
    return h;
}

// the per-dimension records; primeSums[i] = sum of the primes before prime i, perms as radicalInversePermutations lays them out
inline void fillHaltonDims(const int *primes, const int *primeSums, const unsigned short *perms, int n, HaltonDim *out) {
    for (int i = 0; i < n; ++i) {
        HaltonDim &d = out[i];
        d.base = (unsigned)primes[i];
        int s = 0;
        while ((2u << s) <= d.base) ++s;  // floor(log2 base)
        d.shift = s;
        d.magic32 = (unsigned)(((1ull << (32 + s)) + (unsigned long long)d.base - 1) / (unsigned long long)d.base);
        d.permOffset = primeSums[i];
        const float invBase = (float)1 / (float)d.base;
        d.invBase = invBase;
        d.tail = invBase * (float)perms[primeSums[i]] / (1 - invBase);
        d.pad[0] = d.pad[1] = 0;
    }
}

// The per-path view of GlobalSampler: interval sample index + running dimension.
struct PathSampler {
    long long index;
    int dim;
};
// GlobalSampler::Get1D / Get2D (sampler.cpp:181-196), including the jump over the array dimensions.
PG_HD float get1D(const HaltonParams &h, PathSampler &s) {
    if (s.dim >= kArrayStartDim && s.dim < h.arrayEndDim) s.dim = h.arrayEndDim;
    return haltonSample(h, s.index, s.dim++);
}
PG_HD void get2D(const HaltonParams &h, PathSampler &s, float *u0, float *u1) {
    if (s.dim + 1 >= kArrayStartDim && s.dim < h.arrayEndDim) s.dim = h.arrayEndDim;
    *u0 = haltonSample(h, s.index, s.dim);
    *u1 = haltonSample(h, s.index, s.dim + 1);
    s.dim += 2;
}

// Element k of 2D sample array `a` for pixel sample `sampleNum` (GlobalSampler::StartPixel, sampler.cpp:138-148,
// + Sampler::Get2DArray, sampler.cpp:113-118): array a lives in dimensions kArrayStartDim + 2a, +1 and is filled
// from the pixel's samples j = 0 .. n * spp - 1; pixel sample s reads elements s * n .. s * n + n - 1.
// `index` is the Halton index of pixel sample `sampleNum` (GetIndexForSample: offset + sampleNum * stride).
PG_HD void arraySample2D(const HaltonParams &h, long long index, int sampleNum, int a, int n, int k, float *u0, float *u1) {
    long long j = (long long)sampleNum * n + k;
    long long idx = index + (j - sampleNum) * h.sampleStride;
    int dim = kArrayStartDim + 2 * a;
    *u0 = haltonSample(h, idx, dim);
    *u1 = haltonSample(h, idx, dim + 1);
}

}  // namespace pg
#endif
