// JVM random streams that leak into MLlib's results, usable on host and device.
//   XorShift  org.apache.spark.util.random.XORShiftRandom (hashSeed over a 64-BYTE buffer)
//   JavaRand  java.util.Random
//   Well19937c + Poisson: org.apache.commons.math3 3.4.1 (restated from the published
//   algorithm, not validated against a JVM)
#pragma once
#include <cstdint>

namespace har {

__host__ __device__ inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

__host__ __device__ inline uint32_t murmur3_64byte_seedbuf(int64_t seed, uint32_t h) {
    // MurmurHash3_x86_32 over 8 big-endian seed bytes followed by 56 zero bytes
    uint8_t b[8];
    for (int i = 0; i < 8; i++) b[i] = (uint8_t)((uint64_t)seed >> (56 - 8 * i));
    for (int blk = 0; blk < 16; blk++) {
        uint32_t k = 0;
        if (blk < 2) k = (uint32_t)b[4 * blk] | ((uint32_t)b[4 * blk + 1] << 8) | ((uint32_t)b[4 * blk + 2] << 16) |
                         ((uint32_t)b[4 * blk + 3] << 24);
        k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u;
        h ^= k; h = rotl32(h, 13); h = h * 5 + 0xe6546b64u;
    }
    h ^= 64u;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}

struct XorShift {
    uint64_t s;
    __host__ __device__ explicit XorShift(int64_t seed) {
        uint32_t lo = murmur3_64byte_seedbuf(seed, 0x3c074a61u);
        uint32_t hi = murmur3_64byte_seedbuf(seed, lo);
        s = ((uint64_t)hi << 32) | lo;
    }
    __host__ __device__ uint32_t next(int bits) {
        s ^= s << 21; s ^= s >> 35; s ^= s << 4;
        return (uint32_t)(s & ((1ULL << bits) - 1));
    }
    __host__ __device__ double next_double() {
        uint64_t a = next(26), b = next(27);
        return (double)((a << 27) + b) * (1.0 / 9007199254740992.0);
    }
    __host__ __device__ int32_t next_int() { return (int32_t)next(32); }
};

struct JavaRand {
    uint64_t s;
    explicit JavaRand(int64_t seed) : s(((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1)) {}
    int32_t next(int bits) {
        s = (s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
        return (int32_t)(int64_t)(s >> (48 - bits));
    }
    int64_t next_long() {
        int64_t hi = next(32), lo = next(32);
        return (int64_t)((uint64_t)hi << 32) + lo;
    }
};

struct Well19937c {
    uint32_t v[624];
    int index;
    __host__ __device__ void seed(int64_t sd) {
        v[0] = (uint32_t)((uint64_t)sd >> 32);
        v[1] = (uint32_t)((uint64_t)sd & 0xffffffffu);
        for (int i = 2; i < 624; i++) {
            int64_t l = (int64_t)(int32_t)v[i - 2];
            v[i] = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
        }
        index = 0;
    }
    __host__ __device__ uint32_t next(int bits) {
        int idx = index;
        int iRm1 = idx + 623; if (iRm1 >= 624) iRm1 -= 624;
        int iRm2 = idx + 622; if (iRm2 >= 624) iRm2 -= 624;
        int i1 = idx + 70; if (i1 >= 624) i1 -= 624;
        int i2 = idx + 179; if (i2 >= 624) i2 -= 624;
        int i3 = idx + 449; if (i3 >= 624) i3 -= 624;
        uint32_t v0 = v[idx], vM1 = v[i1], vM2 = v[i2], vM3 = v[i3];
        uint32_t z0 = (0x80000000u & v[iRm1]) ^ (0x7FFFFFFFu & v[iRm2]);
        uint32_t z1 = (v0 ^ (v0 << 25)) ^ (vM1 ^ (vM1 >> 27));
        uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
        uint32_t z3 = z1 ^ z2;
        uint32_t z4 = z0 ^ (z1 ^ (z1 << 9)) ^ (z2 ^ (z2 << 21)) ^ (z3 ^ (z3 >> 21));
        v[idx] = z3;
        v[iRm1] = z4;
        v[iRm2] &= 0x80000000u;
        index = iRm1;
        z4 ^= (z4 << 7) & 0xe46e1700u;
        z4 ^= (z4 << 15) & 0x9b868000u;
        return z4 >> (32 - bits);
    }
    __host__ __device__ double next_double() {
        uint64_t hi = (uint64_t)next(26) << 26;
        uint64_t lo = next(26);
        return (double)(hi | lo) * (1.0 / 4503599627370496.0);
    }
    // PoissonDistribution.sample() for mean < 40
    __host__ __device__ int poisson(double mean, double p) {
        long n = 0;
        double r = 1.0;
        while ((double)n < 1000.0 * mean) {
            r *= next_double();
            if (r >= p) n++; else return (int)n;
        }
        return (int)n;
    }
};

}  // namespace har
