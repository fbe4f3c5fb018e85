// TEST INFRASTRUCTURE — random-stream sources for the CPU oracle (see orc_math.h header).
//
// Three interchangeable sources, each handing out, per sampling unit (an item or a respondent),
// a sequence of standard normals and a sequence of U(0,1) uniforms in the order the reference
// consumes them (normals first, then u, eps0, one uniform per rejection):
//   SuppliedRng  caller-provided arrays (parity mode: the GPU path gets the very same arrays)
//   PhiloxRng    the counter-based generator the CUDA path uses in production mode
//                (Philox4x32-10, Salmon et al. SC'11; spec restated, not shared, from
//                gpirt_b200/csrc/rng.cuh); uniforms are bit-exact vs the GPU, normals agree
//                to libm rounding
//   MtRng        the reference's ThreadRNG (src/gpirt.h:13-41): std::mt19937 +
//                std::normal_distribution + std::uniform_real_distribution, reseeded per sweep
//                with iter*10000 + tid*1000 (src/gpirt.h:178-182), distribution objects
//                persisting across reseeds exactly as in the reference.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <random>
#include <vector>

namespace orc {

enum Stream : uint32_t {
    ST_F_Z = 0, ST_F_U = 1, ST_FSTAR_Z = 2, ST_THETA_Z = 3, ST_THETA_U = 4,
    ST_BETA_Z = 5, ST_BETA_U = 6, ST_THR_Z = 7, ST_THR_U = 8, ST_INIT_BETA = 9, ST_INIT_F = 10
};

struct Philox {
    static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
        uint64_t p = uint64_t(a) * uint64_t(b);
        hi = uint32_t(p >> 32);
        lo = uint32_t(p);
    }
    static inline void run(uint32_t c[4], uint32_t k0, uint32_t k1) {
        for (int r = 0; r < 10; ++r) {
            uint32_t hi0, lo0, hi1, lo1;
            mulhilo(0xD2511F53u, c[0], hi0, lo0);
            mulhilo(0xCD9E8D57u, c[2], hi1, lo1);
            uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
            c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
            k0 += 0x9E3779B9u;
            k1 += 0xBB67AE85u;
        }
    }
    // 52 random bits + 1/2, scaled: strictly inside (0,1), every value exactly representable
    static inline double u01(uint32_t a, uint32_t b) {
        uint64_t k = (uint64_t(a >> 6) << 26) | uint64_t(b >> 6);
        return (double(k) + 0.5) * (1.0 / 4503599627370496.0);
    }
    // pair of uniforms for (block, unit, iter, stream|chain<<16)
    static inline void uniform2(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stream,
                                uint32_t unit, uint32_t block, double& a, double& b) {
        uint32_t c[4] = {block, unit, iter, stream | (chain << 16)};
        run(c, uint32_t(seed), uint32_t(seed >> 32));
        a = u01(c[0], c[1]);
        b = u01(c[2], c[3]);
    }
    static inline void normal2(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stream,
                               uint32_t unit, uint32_t block, double& z0, double& z1) {
        double u1, u2;
        uniform2(seed, chain, iter, stream, unit, block, u1, u2);
        double r = std::sqrt(-2.0 * std::log(u1));
        double t = 6.283185307179586476925286766559 * u2;
        z0 = r * std::cos(t);
        z1 = r * std::sin(t);
    }
};

// Sequential per-unit view used by the samplers.
struct UnitRng {
    virtual double norm() = 0;
    virtual double unif() = 0;
    virtual ~UnitRng() {}
};

struct SuppliedRng : UnitRng {
    const double* z; const double* u; size_t nz, nu, iz = 0, iu = 0;
    bool overrun = false;
    SuppliedRng(const double* z_, size_t nz_, const double* u_, size_t nu_) : z(z_), u(u_), nz(nz_), nu(nu_) {}
    double norm() override { if (iz >= nz) { overrun = true; return 0.0; } return z[iz++]; }
    double unif() override { if (iu >= nu) { overrun = true; return 0.5; } return u[iu++]; }
};

struct PhiloxRng : UnitRng {
    uint64_t seed; uint32_t chain, iter, zs, us, unit; uint32_t iz = 0, iu = 0;
    PhiloxRng(uint64_t seed_, uint32_t chain_, uint32_t iter_, uint32_t zs_, uint32_t us_, uint32_t unit_)
        : seed(seed_), chain(chain_), iter(iter_), zs(zs_), us(us_), unit(unit_) {}
    double norm() override {
        double a, b;
        Philox::normal2(seed, chain, iter, zs, unit, iz >> 1, a, b);
        return (iz++ & 1) ? b : a;
    }
    double unif() override {
        double a, b;
        Philox::uniform2(seed, chain, iter, us, unit, iu >> 1, a, b);
        return (iu++ & 1) ? b : a;
    }
};

// The reference's ThreadRNG; shared sequentially by every unit a thread handles.
struct MtRng : UnitRng {
    std::mt19937 gen;
    std::normal_distribution<double> nd{0.0, 1.0};
    std::uniform_real_distribution<double> ud{0.0, 1.0};
    void seed(unsigned s) { gen.seed(s); }
    double norm() override { return nd(gen); }
    double unif() override { return ud(gen); }
};

// Wraps any source and keeps what a unit drew: tests replay an mt19937 run as supplied streams on the GPU path.
struct RecordingRng : UnitRng {
    UnitRng& base; std::vector<double>& z; std::vector<double>& u;
    RecordingRng(UnitRng& b, std::vector<double>& z_, std::vector<double>& u_) : base(b), z(z_), u(u_) {}
    double norm() override { double v = base.norm(); z.push_back(v); return v; }
    double unif() override { double v = base.unif(); u.push_back(v); return v; }
};

}  // namespace orc
