// Host check of libskylark_b200/csrc/cr_math.cuh against __float128 (libquadmath).
//   g++ -O2 -ffp-contract=off -o /tmp/cr_test tools/cr_math_host_test.cpp -lquadmath && /tmp/cr_test
// The device arithmetic is IEEE double with explicit fma, so the host build of the same
// header computes bit-identical results.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <quadmath.h>
#include <random>

#define __device__
#define __forceinline__ inline
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __drcp_rn(double a) { return 1.0 / a; }
static inline long long __double_as_longlong(double x) { long long r; memcpy(&r, &x, 8); return r; }
static inline double __longlong_as_double(long long x) { double r; memcpy(&r, &x, 8); return r; }
template <typename T> static inline T __ldg(const T *p) { return *p; }
using std::fma;
#include "../libskylark_b200/csrc/cr_math.cuh"

static long long ulps(double a, double b) {
    long long x = __double_as_longlong(a), y = __double_as_longlong(b);
    if (x < 0) x = (long long)0x8000000000000000ULL - x;
    if (y < 0) y = (long long)0x8000000000000000ULL - y;
    return x > y ? x - y : y - x;
}

int main() {
    std::mt19937_64 g(12345);
    const double PI = 3.14159265358979323846;
    long long n = 4000000, bad_log = 0, bad_cos = 0, bad_tan = 0, max_log = 0, max_cos = 0, max_tan = 0;
    for (long long it = 0; it < n; it++) {
        uint64_t a = g(), b = g();
        double r1 = (double)a * 0x1p-64, r2 = (double)b * 0x1p-64;
        if (r1 >= 1.0 || r2 >= 1.0) continue;
        double x = 1.0 - r2;
        if (it % 7 == 0) x = ldexp(x, -(int)(a % 53));          // all exponents down to 2^-53
        if (it % 11 == 0) x = 1.0 - ldexp(r2, -(int)(a % 50));  // next to 1
        double want = (double)logq((__float128)x), got = skb::log_cr(x);
        long long u = ulps(want, got); if (u) bad_log++; if (u > max_log) max_log = u;
        double y = 2.0 * PI * r1;
        if (it % 13 == 0) y = (a % 5) * (PI / 2) + ldexp(r2 - 0.5, -(int)(b % 45));   // next to the zeros
        if (y < 0) y = -y;
        if (y > 2.0 * PI) y = 2.0 * PI;
        want = (double)cosq((__float128)y); got = skb::cos_cr_0_2pi(y);
        u = ulps(want, got); if (u) bad_cos++; if (u > max_cos) max_cos = u;
        double t = PI * (r1 - 0.5);
        if (it % 17 == 0) t = (a & 1 ? 1 : -1) * (PI * 0.5 - ldexp(r2, -(int)(b % 50)));   // next to the poles
        if (t > PI * 0.5) t = PI * 0.5;
        if (t < -PI * 0.5) t = -PI * 0.5;
        want = (double)tanq((__float128)t); got = skb::tan_cr_halfpi(t);
        u = ulps(want, got); if (u) bad_tan++; if (u > max_tan) max_tan = u;
    }
    printf("samples %lld\nlog: %lld not correctly rounded (max %lld ulp)\ncos: %lld (max %lld)\ntan: %lld (max %lld)\n",
           n, bad_log, max_log, bad_cos, max_cos, bad_tan, max_tan);
    // edge cases
    printf("log(1) = %g, log(2^-53) ok %d, cos(0) = %.17g, cos(2pi) = %.17g, tan(-pi/2) ok %d\n", skb::log_cr(1.0),
           skb::log_cr(0x1p-53) == (double)logq(0x1p-53Q), skb::cos_cr_0_2pi(0.0), skb::cos_cr_0_2pi(2.0 * PI),
           skb::tan_cr_halfpi(-PI * 0.5) == (double)tanq((__float128)(-PI * 0.5)));
    return (max_log > 1 || max_cos > 1 || max_tan > 1) ? 1 : 0;
}
