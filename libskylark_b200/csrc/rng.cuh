// Device-side counter RNG and variate generation for the sketch kernels.
//
// What the reference does (and what these functions reproduce bit for bit on
// the integer side):
//   * one sample per global stream index g = base + i: a fresh
//     r123::MicroURNG<Threefry2x64_R<13>> with ctr = {g, 0}, key = {seed, 0}
//     (base/randgen.hpp:27-30, 98-115);
//   * the distribution object (Boost.Random) pulls 64-bit draws from it; block n
//     of the MicroURNG has n in the high 32 bits of ctr[1] and hands out X1
//     before X0.
// Distribution semantics follow Boost 1.60 (the conda pin); the Rademacher draw
// order of Boost <= 1.55 is selectable (SKB_ORDER_155).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace skb {

#define SKB_ORDER_160 160
#define SKB_ORDER_155 155

#define SKB_BOOST_PI 3.14159265358979323846

__host__ __device__ __forceinline__ uint64_t rotl64(uint64_t x, int r) {
    return (x << r) | (x >> (64 - r));
}

// Threefry2x64, 13 rounds, key = {k0, 0}.  Fully unrolled; ks2 folds the parity
// constant.  Returns X0 in .x and X1 in .y.
struct u64x2 { uint64_t x, y; };

__host__ __device__ __forceinline__ u64x2 threefry2x64_13(uint64_t c0, uint64_t c1,
                                                          uint64_t k0) {
    const uint64_t ks0 = k0, ks1 = 0, ks2 = 0x1BD11BDAA9FC1A22ULL ^ k0;
    uint64_t x0 = c0 + ks0, x1 = c1 + ks1;
#define SKB_MIX(R) x0 += x1; x1 = rotl64(x1, R); x1 ^= x0;
    SKB_MIX(16) SKB_MIX(42) SKB_MIX(12) SKB_MIX(31)
    x0 += ks1; x1 += ks2 + 1;
    SKB_MIX(16) SKB_MIX(32) SKB_MIX(24) SKB_MIX(21)
    x0 += ks2; x1 += ks0 + 2;
    SKB_MIX(16) SKB_MIX(42) SKB_MIX(12) SKB_MIX(31)
    x0 += ks0; x1 += ks1 + 3;
    SKB_MIX(16)
#undef SKB_MIX
    u64x2 r; r.x = x0; r.y = x1; return r;
}

// MicroURNG over one stream index.
struct MicroUrng {
    uint64_t c0, k0, r0, r1;
    uint32_t n;
    int last;
    __host__ __device__ __forceinline__ MicroUrng(uint64_t key, uint64_t idx)
        : c0(idx), k0(key), r0(0), r1(0), n(0), last(0) {}
    __host__ __device__ __forceinline__ uint64_t next() {
        if (last == 0) {
            u64x2 b = threefry2x64_13(c0, (uint64_t)n << 32, k0);
            r0 = b.x; r1 = b.y; n++; last = 2;
        }
        --last;
        return last == 1 ? r1 : r0;
    }
};

__host__ __device__ __forceinline__ uint64_t seed_to_key(int seed) {
    return (uint64_t)(int64_t)seed;   // static_cast<uint64_t>(int), randgen.hpp:37
}

// u01 of a draw without the conversion pipe and without the final multiply: the two halves
// are planted in doubles whose exponents already carry the 2^-64 factor,
//   dhi = 2^20 + hi*2^-32,  dlo = 2^-12 + lo*2^-64,  result = (dhi - (2^20 + 2^-12)) + dlo,
// every step exact except the last add, which rounds hi*2^-32 + lo*2^-64 = x*2^-64 once
// (to nearest even) -- bit-identical to double(x) * 2^-64.
__device__ __forceinline__ double u64_to_u01_rn(uint64_t x) {
    double dhi = __hiloint2double(0x41300000, (int)(uint32_t)(x >> 32));   // 2^20 + hi*2^-32
    double dlo = __hiloint2double(0x3F300000, (int)(uint32_t)x);           // 2^-12 + lo*2^-64
    return __dadd_rn(__dsub_rn(dhi, 1048576.000244140625), dlo);           // 2^20 + 2^-12
}

// double(x) for a 64-bit unsigned x, round to nearest even, without the
// conversion pipe: hi*2^32 and lo are exact, their sum rounds once.
__device__ __forceinline__ double u64_to_double_rn(uint64_t x) {
#ifdef __CUDA_ARCH__
    double dhi = __hiloint2double(0x45300000, (int)(uint32_t)(x >> 32));   // 2^84 + hi*2^32
    double dlo = __hiloint2double(0x43300000, (int)(uint32_t)x);           // 2^52 + lo
    return __dadd_rn(__dsub_rn(dhi, 19342813118337666422669312.0), dlo);   // 2^84 + 2^52
#else
    return (double)x;
#endif
}

__host__ __device__ __forceinline__ double u64_to_double(uint64_t x) {
#ifdef __CUDA_ARCH__
    return u64_to_double_rn(x);
#else
    return (double)x;
#endif
}

// boost::random::uniform_01<double>
__host__ __device__ __forceinline__ double boost_u01(MicroUrng &u) {
    for (;;) {
        double r = u64_to_double(u.next()) * 0x1p-64;
        if (r < 1.0) return r;
    }
}

// boost::random::detail::generate_uniform_int for range < 2^64-1
__host__ __device__ __forceinline__ uint64_t boost_uniform_int(MicroUrng &u,
                                                               uint64_t range) {
    if (range == 0) return 0;
    const uint64_t brange = ~(uint64_t)0;
    uint64_t bucket = brange / (range + 1);
    if (brange % (range + 1) == range) bucket++;
    for (;;) {
        uint64_t r = u.next() / bucket;
        if (r <= range) return r;
    }
}

// utility/distributions.hpp:40-52 (discrete_distribution{.5,0,.5} - 1)
__host__ __device__ __forceinline__ double boost_rademacher(MicroUrng &u, int order) {
    uint64_t r; double t;
    if (order == SKB_ORDER_155) { t = boost_u01(u); r = boost_uniform_int(u, 2); }
    else                        { r = boost_uniform_int(u, 2); t = boost_u01(u); }
    // alias table [0]=(0.5,->2) [1]=(0,->0) [2]=(1,-)
    bool minus = (r == 1) || (r == 0 && t < 0.5);
    return minus ? -1.0 : 1.0;
}

#ifdef __CUDACC__
}  // namespace skb
#include "cr_math.cuh"
namespace skb {
// SKB_LIBM_FAST (compile-time, for A/B timing only): the previous generators -- libdevice log / tan
// (1-2 ulp) and the 0.6-ulp cosine below -- instead of the nearly correctly rounded cr_math.cuh.

// cos(x) for x in [0, 2*pi]: Cody-Waite reduction by pi/2 in 33-bit pieces (exact products,
// with the extra rounds when x sits next to a multiple of pi/2) and the classic degree-13/14
// minimax kernels evaluated on the (head, tail) pair -- about 0.6 ulp, against the 1-2 ulp of
// the general-purpose routine; the Box-Muller product then stays within 2 ulp of a libm
// evaluation of the same expression far more often.
__device__ __forceinline__ double cos_0_2pi(double x) {
    const double invpio2 = 6.36619772367581382433e-01;
    const double pio2_1 = 1.57079632673412561417e+00, pio2_1t = 6.07710050650619224932e-11;
    const double pio2_2 = 6.07710050630396597660e-11, pio2_2t = 2.02226624879595063154e-21;
    const double pio2_3 = 2.02226624871116645580e-21, pio2_3t = 8.47842766036889956997e-32;
    const int n = (int)(__dmul_rn(x, invpio2) + 0.5);          // 0..4
    const double fn = (double)n;
    double r = __dsub_rn(x, __dmul_rn(fn, pio2_1));            // exact: 33-bit piece times n <= 4
    double w = __dmul_rn(fn, pio2_1t);
    double y0 = __dsub_rn(r, w);
    const int ex = (__double2hiint(x) >> 20) & 0x7ff;
    if (ex - ((__double2hiint(y0) >> 20) & 0x7ff) > 16) {      // heavy cancellation: 2nd piece
        double t = r;
        w = __dmul_rn(fn, pio2_2);
        r = __dsub_rn(t, w);
        w = __dsub_rn(__dmul_rn(fn, pio2_2t), __dsub_rn(__dsub_rn(t, r), w));
        y0 = __dsub_rn(r, w);
        if (ex - ((__double2hiint(y0) >> 20) & 0x7ff) > 49) {  // and the 3rd
            t = r;
            w = __dmul_rn(fn, pio2_3);
            r = __dsub_rn(t, w);
            w = __dsub_rn(__dmul_rn(fn, pio2_3t), __dsub_rn(__dsub_rn(t, r), w));
            y0 = __dsub_rn(r, w);
        }
    }
    const double y1 = __dsub_rn(__dsub_rn(r, y0), w);
    // cos(x) = +cos(y), -sin(y), -cos(y), +sin(y) for n = 0, 1, 2, 3 (4 = 0).  One Horner chain
    // with the coefficient set picked per lane, so that lanes in different quadrants do not
    // serialise two polynomial evaluations.
    const bool use_sin = n & 1;
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    const double z = y0 * y0;
    double t = use_sin ? 0.0 : C6;
    t = fma(t, z, use_sin ? S6 : C5);
    t = fma(t, z, use_sin ? S5 : C4);
    t = fma(t, z, use_sin ? S4 : C3);
    t = fma(t, z, use_sin ? S3 : C2);
    t = fma(t, z, use_sin ? S2 : C1);
    double res;
    if (use_sin) {
        const double v = z * y0;
        res = y0 - ((z * (0.5 * y1 - v * t) - y1) - v * S1);
    } else {
        const double rc = t * z, ax = fabs(y0);
        const double qx = ax < 0.3 ? 0.0
                        : (ax > 0.78125 ? 0.28125
                                        : __hiloint2double(__double2hiint(ax) - 0x00200000, 0));
        const double hz = 0.5 * z - qx, a = 1.0 - qx;
        res = a - (hz - (z * rc - y0 * y1));
    }
    return ((n + 1) & 2) ? -res : res;
}

// boost::random::normal_distribution (Box-Muller, cosine branch)
// r1, r2 = the two uniform_01 draws, in draw order
__device__ __forceinline__ double box_muller_cos(double r1, double r2) {
#ifdef SKB_LIBM_FAST
    double rho = sqrt(__dmul_rn(-2.0, log(__dsub_rn(1.0, r2))));
    return __dmul_rn(rho, cos_0_2pi(__dmul_rn(2.0 * SKB_BOOST_PI, r1)));
#else
    double rho = sqrt(__dmul_rn(-2.0, log_cr(__dsub_rn(1.0, r2))));     // 1 - r2 in [2^-53, 1]
    return __dmul_rn(rho, cos_cr_0_2pi(__dmul_rn(2.0 * SKB_BOOST_PI, r1)));
#endif
}
__device__ __forceinline__ double boost_normal(MicroUrng &u) {
    double r1 = boost_u01(u);
    double r2 = boost_u01(u);
    return box_muller_cos(r1, r2);
}

// boost::random::cauchy_distribution(0,1)
__device__ __forceinline__ double boost_cauchy(MicroUrng &u) {
#ifdef SKB_LIBM_FAST
    return tan(__dmul_rn(SKB_BOOST_PI, __dsub_rn(boost_u01(u), 0.5)));
#else
    return tan_cr_halfpi(__dmul_rn(SKB_BOOST_PI, __dsub_rn(boost_u01(u), 0.5)));
#endif
}

// boost::random::exponential_distribution(1)
__device__ __forceinline__ double boost_exponential(MicroUrng &u) {
#ifdef SKB_LIBM_FAST
    return -log(__dsub_rn(1.0, boost_u01(u)));
#else
    return -log_cr(__dsub_rn(1.0, boost_u01(u)));
#endif
}

// Fast path for the dense generators: both draws of block 0 taken directly; the
// (probability 2^-53) rejection case falls back to the general routine.
__device__ __forceinline__ double normal_at(uint64_t key, uint64_t idx) {
    u64x2 b = threefry2x64_13(idx, 0, key);
    if (b.x >= 0xFFFFFFFFFFFFFC00ULL || b.y >= 0xFFFFFFFFFFFFFC00ULL) {
        MicroUrng u(key, idx);
        return boost_normal(u);
    }
    return box_muller_cos(u64_to_u01_rn(b.y), u64_to_u01_rn(b.x));
}

__device__ __forceinline__ double cauchy_at(uint64_t key, uint64_t idx) {
    MicroUrng u(key, idx);
    return boost_cauchy(u);
}
#endif

// boost::random::uniform_real_distribution(a,b)
__host__ __device__ __forceinline__ double boost_uniform_real(MicroUrng &u, double a,
                                                              double b) {
    for (;;) {
        double num = u64_to_double(u.next());
#ifdef __CUDA_ARCH__
        double r = __dadd_rn(__dmul_rn(num * 0x1p-64, b - a), a);   // no fma contraction
#else
        double r = num / 0x1p64 * (b - a) + a;
#endif
        if (r < b) return r;
    }
}

}  // namespace skb
