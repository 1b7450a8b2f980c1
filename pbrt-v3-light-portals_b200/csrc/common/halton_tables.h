// Tables behind the Halton sampler, generated rather than stored:
//   * the first 1000 primes and their running sums (Primes / PrimeSums,
//     src/core/lowdiscrepancy.cpp:40-230),
//   * the digit permutations for scrambled radical inverses, produced by
//     shuffling 0..p-1 for every prime p with the default-seeded PCG32
//     (ComputeRadicalInversePermutations lowdiscrepancy.cpp:2490-2504,
//      Shuffle sampling.h:152-158, RNG rng.h:56-148; HaltonSampler seeds
//      `RNG rng;` halton.cpp:69-72).
#ifndef PG_HALTON_TABLES_H
#define PG_HALTON_TABLES_H

#include <cstdint>
#include <utility>
#include <vector>

namespace pg {

constexpr int kPrimeTableSize = 1000;

inline std::vector<int> firstPrimes(int n = kPrimeTableSize) {
    std::vector<int> primes;
    for (int c = 2; (int)primes.size() < n; ++c) {
        bool isPrime = true;
        for (int p : primes) {
            if (p * p > c) break;
            if (c % p == 0) { isPrime = false; break; }
        }
        if (isPrime) primes.push_back(c);
    }
    return primes;
}

// Division magics, two tables of primes.size() entries back to back (see sampler.h):
//   [i]      floor(2^64 / p) + 1                     exact a / p = mulhi64(a, .) for a < 2^32
//   [n + i]  ceil(2^(32 + s) / p), s = floor(log2 p)  exact a / p = mulhi32(a, .) >> s for a < 2^31 (fits 32 bits)
inline std::vector<unsigned long long> primeMagics(const std::vector<int> &primes) {
    std::vector<unsigned long long> m;
    for (int p : primes) m.push_back(~0ull / (unsigned long long)p + 1);
    for (int p : primes) {
        int s = 0;
        while ((2u << s) <= (unsigned)p) ++s;
        m.push_back(((1ull << (32 + s)) + (unsigned long long)p - 1) / (unsigned long long)p);
    }
    return m;
}

struct Pcg32 {  // rng.h:56-148
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    uint32_t uniformUInt32() {
        uint64_t oldstate = state;
        state = oldstate * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t)(((oldstate >> 18u) ^ oldstate) >> 27u);
        uint32_t rot = (uint32_t)(oldstate >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    uint32_t uniformUInt32(uint32_t b) {
        uint32_t threshold = (~b + 1u) % b;
        while (true) {
            uint32_t r = uniformUInt32();
            if (r >= threshold) return r % b;
        }
    }
};

// perms[primeSums[i] + d] = permuted digit d in base primes[i]
inline std::vector<uint16_t> radicalInversePermutations(const std::vector<int> &primes) {
    size_t total = 0;
    for (int p : primes) total += p;
    std::vector<uint16_t> perms(total);
    Pcg32 rng;
    uint16_t *p = perms.data();
    for (int prime : primes) {
        for (int j = 0; j < prime; ++j) p[j] = (uint16_t)j;
        for (int i = 0; i < prime; ++i) {  // Shuffle(p, prime, 1, rng)
            int other = i + (int)rng.uniformUInt32((uint32_t)(prime - i));
            std::swap(p[i], p[other]);
        }
        p += prime;
    }
    return perms;
}

}  // namespace pg
#endif
