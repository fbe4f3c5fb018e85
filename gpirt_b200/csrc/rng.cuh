// Counter-based random streams for the sweep: Philox4x32-10 (Salmon et al., SC'11) keyed by the chain
// seed, countered by (draw block, unit, sweep, stream|chain). Replaces the reference's per-thread
// mt19937 ThreadRNG (src/gpirt.h:13-41), whose streams depend on the OpenMP thread that happens to
// pick an item up; here a unit's stream is a pure function of (seed, chain, sweep, block, unit), so
// results do not depend on the launch geometry or on how items are sharded across GPUs.
#pragma once
#include <cstdint>

namespace gpirt {

enum Stream : uint32_t {
    ST_F_Z = 0, ST_F_U = 1, ST_FSTAR_Z = 2, ST_THETA_Z = 3, ST_THETA_U = 4,
    ST_BETA_Z = 5, ST_BETA_U = 6, ST_THR_Z = 7, ST_THR_U = 8, ST_INIT_BETA = 9, ST_INIT_F = 10
};

struct RngKey {
    uint32_t k0, k1;   // seed
    uint32_t tag, iter;  // tag = session << 8 | chain << 16: ORed into the stream id (streams are < 256; session 0 of chain c is the plain chain)
};

__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
        uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
#else
        uint64_t p0 = uint64_t(0xD2511F53u) * c[0], p1 = uint64_t(0xCD9E8D57u) * c[2];
        uint32_t hi0 = uint32_t(p0 >> 32), lo0 = uint32_t(p0), hi1 = uint32_t(p1 >> 32), lo1 = uint32_t(p1);
#endif
        uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

// 52 random bits + 1/2, scaled by 2^-52: strictly inside (0,1), every value exact in fp64.
__host__ __device__ inline double u01_from_bits(uint32_t a, uint32_t b) {
    uint64_t k = (uint64_t(a >> 6) << 26) | uint64_t(b >> 6);
    return (double(k) + 0.5) * (1.0 / 4503599627370496.0);
}

__host__ __device__ inline void philox_uniform2(const RngKey& key, uint32_t stream, uint32_t unit, uint32_t block,
                                                double& a, double& b) {
    uint32_t c[4] = {block, unit, key.iter, stream | key.tag};
    philox4x32_10(c, key.k0, key.k1);
    a = u01_from_bits(c[0], c[1]);
    b = u01_from_bits(c[2], c[3]);
}

// k-th uniform of a unit's stream
__device__ inline double philox_uniform(const RngKey& key, uint32_t stream, uint32_t unit, uint32_t k) {
    double a, b;
    philox_uniform2(key, stream, unit, k >> 1, a, b);
    return (k & 1) ? b : a;
}

// Box-Muller pair: normals 2*block and 2*block+1 of a unit's stream
__device__ inline void philox_normal2(const RngKey& key, uint32_t stream, uint32_t unit, uint32_t block,
                                      double& z0, double& z1) {
    double u1, u2;
    philox_uniform2(key, stream, unit, block, u1, u2);
    double r = sqrt(-2.0 * log(u1));
    double s, c;
    sincos(6.283185307179586476925286766559 * u2, &s, &c);
    z0 = r * c;
    z1 = r * s;
}

__device__ inline double philox_normal(const RngKey& key, uint32_t stream, uint32_t unit, uint32_t i) {
    double a, b;
    philox_normal2(key, stream, unit, i >> 1, a, b);
    return (i & 1) ? b : a;
}

// Uniform source of one ESS unit: supplied array (parity mode) or Philox.
struct UniformSrc {
    const double* u;      // device pointer or nullptr
    long long stride;
    __device__ inline double get(const RngKey& key, uint32_t stream, uint32_t unit, uint32_t k, int* overrun) const {
        if (u) {
            if ((long long)k >= stride) { *overrun = 1; return 0.5; }
            return u[(long long)unit * stride + k];
        }
        return philox_uniform(key, stream, unit, k);
    }
};

}  // namespace gpirt
