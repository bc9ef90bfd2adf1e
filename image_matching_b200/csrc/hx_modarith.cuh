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

// ---- carry-free lazy multiply-accumulate on 30-bit halves ----------------------------------------
// x = x1*2^30 + x0 with x0, x1 < 2^30 (any x < 2^60).  A product x*y is the three columns
//   lo = x0*y0,  mid = x0*y1 + x1*y0,  hi = x1*y1      (x*y = lo + mid*2^30 + hi*2^60)
// each column is one or two IMAD.WIDE.U32 with a 64-bit addend and never overflows while at most 7
// products are accumulated (mid < 2*7*2^60 < 2^64): 4 multiplier-pipe instructions per product and no
// carry chain, against ~12 for a 64x64->128 multiply-accumulate.  Long-lived operands (evaluation keys,
// conversion matrices) are stored pre-split as (x1 << 32) | x0, so only the streaming operand is split.
struct acc3 { u64 lo, mid, hi; };
constexpr u32 kMask30 = 0x3FFFFFFFu;
__device__ __host__ __forceinline__ u64 split30(u64 x) { return ((x >> 30) << 32) | (x & kMask30); }
__device__ __host__ __forceinline__ u64 unsplit30(u64 s) { return ((s >> 32) << 30) | (s & 0xFFFFFFFFull); }
__device__ __forceinline__ u64 madw(u32 a, u32 b, u64 c) {
    u64 d;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
    return d;
}
// acc += x * y ; x plain (< 2^60), ys = split30(y)
__device__ __forceinline__ void mac30(acc3& a, u64 x, u64 ys) {
    const u32 x0 = (u32)x & kMask30, x1 = (u32)(x >> 30);
    const u32 y0 = (u32)ys, y1 = (u32)(ys >> 32);
    a.lo = madw(x0, y0, a.lo);
    a.mid = madw(x0, y1, a.mid);
    a.mid = madw(x1, y0, a.mid);
    a.hi = madw(x1, y1, a.hi);
}
// acc += x * y with both operands stored split (xs = split30(x), ys = split30(y)): no unpacking instructions at all
__device__ __forceinline__ void mac30ss(acc3& a, u64 xs, u64 ys) {
    const u32 x0 = (u32)xs, x1 = (u32)(xs >> 32), y0 = (u32)ys, y1 = (u32)(ys >> 32);
    a.lo = madw(x0, y0, a.lo);
    a.mid = madw(x0, y1, a.mid);
    a.mid = madw(x1, y0, a.mid);
    a.hi = madw(x1, y1, a.hi);
}
// the accumulated value as a 128-bit integer
__device__ __forceinline__ u128d fold30(const acc3& a) {
    u128d t;
    const u64 m_lo = a.mid << 30, m_hi = a.mid >> 34, h_lo = a.hi << 60, h_hi = a.hi >> 4;
    asm("add.cc.u64 %0, %2, %3;\n\taddc.u64 %1, %4, %5;" : "=l"(t.lo), "=l"(t.hi) : "l"(a.lo), "l"(m_lo), "l"(m_hi), "l"(h_hi));
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(t.lo), "+l"(t.hi) : "l"(h_lo));
    return t;
}

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
// ---- global loads: which cache path for which data ------------------------------------------------------------------------
// MUTABLE data (ciphertexts, digits, scratch: anything a kernel of this process writes during a query) is read with ld.global.cg
// (L2 only).  The non-coherent path (ld.global.nc / __ldg, or what the compiler emits for a `const __restrict__` pointer) is kept
// for data that never changes after setup (twiddles, evaluation keys, database blocks, tables).  Measured on B200
// (tools/race_probe.py, profiles/r2_race_*.log): with the compare chain of one scenario on the engine stream and the next
// scenario's rotations running concurrently on the lane streams, ~3 % of the calls returned a ciphertext with stale 128-byte
// lines when produced data was read through LD.NC -- the pool recycles addresses between the streams and a line cached through
// the non-coherent path on an SM outlived the kernel boundary of the OTHER stream.  With .cg loads: 0 mismatches in 2560 calls.
// These words are read once per CTA, so L1 gives nothing for them anyway.
__device__ __forceinline__ ulonglong2 ldg128(const u64* p) {
    ulonglong2 v;
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ u64 ldg64(const u64* p) {
    u64 v;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ ulonglong2 ld_stream128(const u64* p) {
    ulonglong2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
// Programmatic dependent launch: a kernel launched with launch_pdl() may be scheduled while the previous kernel of its stream is
// draining; it must not touch that kernel's output before this returns (full completion + memory flush of the predecessor).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// (An early "griddepcontrol.launch_dependents" at the top of every kernel was tried and gave wrong results in chains of three
// or more dependent kernels, so the trigger is left implicit: the dependent is scheduled as this kernel's CTAs exit.)
// Asynchronous HBM -> L2 prefetch of `bytes` (multiple of 16) contiguous bytes at a 16-byte aligned address.
__device__ __forceinline__ void l2_prefetch(const void* p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// ---- TMA bulk copy global -> shared, completion on an mbarrier (sm_90+; SASS UBLKCP) -----------------
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// bytes: multiple of 16; both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, u32 bytes, u64* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(u64* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
    u32 done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void st128(u64* p, ulonglong2 v) { *reinterpret_cast<ulonglong2*>(p) = v; }

}  // namespace hx
