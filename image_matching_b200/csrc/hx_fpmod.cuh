// hx_fpmod.cuh -- exact modular arithmetic on FP64 for the 45-bit scaling primes (q < 2^46: 11 of the 12 Q limbs of approaches
// 4/5, 12..18 of 13..19 for approaches 1-3).  Residues are exact integers held in doubles, CENTRED and LAZY (|x| < 2^51); every
// function below is exact (all intermediate integers stay below 2^53, products are split error-free with one FMA), so results are
// congruent mod q to exact integer arithmetic and a final fp_normalize() gives the canonical residue OpenFHE's NativeInteger ops
// produce.  Compiles for host as well (tests/cpp/fp_arith_test.cpp checks it against __int128 on the CPU).
#pragma once
#include <cstdint>
#if defined(__CUDA_ARCH__)
#define HX_FMA(a, b, c) __fma_rn(a, b, c)
#define HX_MUL(a, b) __dmul_rn(a, b)
#define HX_ADD(a, b) __dadd_rn(a, b)
#define HX_SUB(a, b) __dsub_rn(a, b)
#else
#include <cmath>
#include <cstring>
#define HX_FMA(a, b, c) std::fma(a, b, c)
#define HX_MUL(a, b) ((a) * (b))    // host builds of this header use -ffp-contract=off
#define HX_ADD(a, b) ((a) + (b))
#define HX_SUB(a, b) ((a) - (b))
#endif
#if defined(__CUDACC__)
#define HX_HD __host__ __device__ __forceinline__
#else
#define HX_HD inline
#endif

namespace hx {

constexpr double kFpMagic52 = 4503599627370496.0;          // 2^52
constexpr double kFpMagicRint = 6755399441055744.0;        // 1.5 * 2^52: x + M - M = rint(x) for |x| < 2^51

HX_HD double fp_bits_to_double(uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    double d; std::memcpy(&d, &b, 8); return d;
#endif
}
HX_HD uint64_t fp_double_to_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t b; std::memcpy(&b, &d, 8); return b;
#endif
}
// exact for 0 <= x < 2^52
HX_HD double fp_from_u64(uint64_t x) { return HX_SUB(fp_bits_to_double(x | 0x4330000000000000ull), kFpMagic52); }
// exact for an integer-valued 0 <= d < 2^52
HX_HD uint64_t fp_to_u64(double d) { return fp_double_to_bits(HX_ADD(d, kFpMagic52)) & 0x000FFFFFFFFFFFFFull; }

// a * b mod q, centred and lazy: for |a| < 2^51, 0 <= b < q < 2^46 (all integer-valued), qinv = RN(1/q), returns an integer m with
// m = a*b (mod q) and |m| <= 0.75 q + 2^43 (|m| <= 0.51 q + 2^38 when |a| < q).
//   hi + lo = a*b exactly (error-free product);  c = rint(hi / q) up to +-0.25 (the quotient is below 2^51, so the magic-constant
//   rounding applies);  hi - c*q is an integer below 2^46 in magnitude, hence exact in the FMA;  |lo| <= ulp(hi)/2 <= 2^43.
HX_HD double fp_mulmod(double a, double b, double q, double qinv) {
    const double hi = HX_MUL(a, b);
    const double lo = HX_FMA(a, b, -hi);
    const double c = HX_SUB(HX_FMA(hi, qinv, kFpMagicRint), kFpMagicRint);
    return HX_ADD(HX_FMA(-c, q, hi), lo);
}
// any lazy value |v| < 2^51 -> canonical residue in [0, q) as a double
HX_HD double fp_normalize(double v, double q, double qinv) {
    const double c = HX_SUB(HX_FMA(v, qinv, kFpMagicRint), kFpMagicRint);
    double r = HX_FMA(-c, q, v);                // in [-q/2 - 1, q/2 + 1], exact
    if (r < 0.0) r = HX_ADD(r, q);
    return r;
}

}  // namespace hx
