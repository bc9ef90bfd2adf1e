// hx_modarith.cuh -- 64-bit RNS modular arithmetic for sm_100a.
// All moduli are < 2^60, so 4q < 2^62 and lazy [0,2q)/[0,4q) ranges fit a word.
// Three multiplication flavours, chosen by what is constant:
//   shoup_*   : one operand is a long-lived constant w with wsh = floor(w*2^64/q)  (twiddles, CRT scalars)
//   mont_*    : one operand is stored in Montgomery form (keys, database, base-conversion matrix);
//               sums of products are accumulated lazily in 128 bits and reduced once (REDC)
//   mulmod    : both operands variable (ct x ct tensor products): two REDCs via R^2
// Results are always the canonical residue in [0,q): every kernel output is bit-identical to
// exact modular arithmetic, which is what OpenFHE's NativeInteger ops produce.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hx {
typedef unsigned long long u64;   // matches CUDA's 64-bit intrinsics
typedef unsigned int u32;

struct u128d { u64 lo, hi; };

// Per-prime constants kept in device memory (one entry per limb of Q then P).
struct PrimeConst {
    u64 q;        // modulus
    u64 qinv_neg; // -q^-1 mod 2^64
    u64 r2;       // 2^128 mod q  (to Montgomery form / second REDC of mulmod)
    u64 r1;       // 2^64 mod q
    u64 mu64;     // floor(2^64 / q)  (reduce an arbitrary 64-bit word)
    u64 ninv;     // N^-1 mod q
    u64 ninv_sh;  // Shoup quotient of ninv
    double qinv_d; // 1/q rounded to double (FP64 NTT back end, primes below 2^46)
};

__device__ __forceinline__ u128d mul_wide(u64 a, u64 b) {
    u128d r;
    r.lo = a * b;
    r.hi = __umul64hi(a, b);
    return r;
}
__device__ __forceinline__ void acc_wide(u128d& acc, u64 a, u64 b) {
    u64 lo = a * b, hi = __umul64hi(a, b);
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(acc.lo), "+l"(acc.hi) : "l"(lo), "l"(hi));
}
__device__ __forceinline__ u64 csub(u64 x, u64 q) { return x >= q ? x - q : x; }
__device__ __forceinline__ u64 add_mod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
__device__ __forceinline__ u64 sub_mod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }

// a*w mod q in [0,2q) for ANY 64-bit a (w < q, wsh = floor(w*2^64/q)).
__device__ __forceinline__ u64 shoup_lazy(u64 a, u64 w, u64 wsh, u64 q) {
    u64 hi = __umul64hi(a, wsh);
    return a * w - hi * q;
}
__device__ __forceinline__ u64 shoup_mul(u64 a, u64 w, u64 wsh, u64 q) { return csub(shoup_lazy(a, w, wsh, q), q); }
// Cheaper lazy product for the NTT butterflies: the quotient estimate drops a_lo*wsh_lo and the carries
// of the two cross terms, so it is hi, hi-1 or hi-2 and the result lies in [0,4q) (still a*w mod q).
// Three IMAD.WIDE instead of four and no carry chain (measured 14.4 vs 16.4 issue slots, tools/microbench2.cu).
__device__ __forceinline__ u64 umulhi_approx(u64 a, u64 b) {
    const u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32);
    const u64 s1 = (u64)a0 * b1, s2 = (u64)a1 * b0;
    return (u64)a1 * b1 + ((s1 >> 32) + (s2 >> 32));
}
__device__ __forceinline__ u64 shoup_lazy4(u64 a, u64 w, u64 wsh, u64 q) { return a * w - umulhi_approx(a, wsh) * q; }

// Montgomery REDC: T = hi*2^64 + lo < q*2^64  ->  T * 2^-64 mod q in [0,q).
__device__ __forceinline__ u64 redc(u128d t, u64 q, u64 qinv_neg) {
    u64 m = t.lo * qinv_neg;
    u64 mh = __umul64hi(m, q);
    // (T + m*q) / 2^64 : low words cancel to 0 with a carry iff t.lo != 0
    u64 r = t.hi + mh + (t.lo != 0 ? 1ull : 0ull);
    return csub(r, q);
}
// Reduce an arbitrary 64-bit word mod q.
__device__ __forceinline__ u64 reduce64(u64 x, u64 q, u64 mu64) {
    u64 e = __umul64hi(x, mu64);
    u64 r = x - e * q;          // in [0, 2q)
    return csub(csub(r, q), q);
}
// Fold a full 128-bit accumulator below q*2^64 so redc() applies: hi <- hi mod q.
__device__ __forceinline__ u128d fold128(u128d t, u64 q, u64 mu64) {
    t.hi = reduce64(t.hi, q, mu64);
    return t;
}
// a * bm * 2^-64 mod q ; with bm = b*2^64 mod q this is a*b mod q.
__device__ __forceinline__ u64 mont_mul(u64 a, u64 bm, u64 q, u64 qinv_neg) {
    return redc(mul_wide(a, bm), q, qinv_neg);
}
// Both operands plain: a*b mod q.
__device__ __forceinline__ u64 mulmod(u64 a, u64 b, const PrimeConst& pc) {
    u64 t = redc(mul_wide(a, b), pc.q, pc.qinv_neg);     // a*b*R^-1
    return redc(mul_wide(t, pc.r2), pc.q, pc.qinv_neg);  // * R^2 * R^-1
}

// 128-bit global access (limb-coalesced: two adjacent coefficients of one limb).
__device__ __forceinline__ ulonglong2 ldg128(const u64* p) { return __ldg(reinterpret_cast<const ulonglong2*>(p)); }
__device__ __forceinline__ ulonglong2 ld_stream128(const u64* p) {
    ulonglong2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ void st128(u64* p, ulonglong2 v) { *reinterpret_cast<ulonglong2*>(p) = v; }

}  // namespace hx
