// hx_ntt_tile.cuh -- compile-time-geometry negacyclic NTT passes over 2048-element shared-memory
// tiles, shared by the stand-alone NTT kernels and the fused key-switch kernels.
//
// Transform (same as OpenFHE's ChineseRemainderTransformFTT, reached from every key switch and
// rescale of /root/reference/src/sender/sender_diag.cpp:25,79,80):
//   forward : Cooley-Tukey, natural order in, bit-reversed out, stage s uses tw[2^s + group]
//   inverse : Gentleman-Sande with psi^-1 twiddles, then N^-1
// A limb of N = 2^LOGN words is viewed as 2^A rows x 2^B columns (A = LOGN/2):
//   column pass = stages 0..A-1   on a tile of all 2^A rows x C = 2^(11-A) adjacent columns
//   row pass    = stages A..LOGN-1 on a tile of D = 2^(11-B) whole rows = 2048 contiguous words
// A thread owns 8 tile elements per round and runs up to 3 stages in registers (radix-8) between
// shared-memory exchanges.  Harvey lazy butterflies on Shoup twiddles with the approximate quotient
// of shoup_lazy4 (product in [0,4q)): forward values live in [0,8q), inverse values in [0,4q)
// (q < 2^60, so 8q fits a word); callers normalise on the way out.
#pragma once
#include "hx_modarith.cuh"

namespace hx {

constexpr int kLogTile = 11;
constexpr int kTileN = 1 << kLogTile;          // 2048 elements
constexpr int kTileThreads = kTileN / 8;       // 256 threads
constexpr int kTileWords = kTileN + (kTileN >> 4) + 8;   // padded: 1 word per 16 against bank conflicts
__device__ __forceinline__ int tphys(int e) { return e + (e >> 4); }
// Compile-time offsets of the padded layout (verified exhaustively): with base = insert_bits3(t, p),
//   tphys(base + (k << p)) == tphys(base) + toff(k, p)            (the 8 elements a thread owns in a round)
//   tphys((u*256 + t)*2 + j) == tphys(2t) + u*kPairStep + j        (128-bit load/store loops)
//   tphys(u*256 + t)        == tphys(t)  + u*kWordStep             (column-tile loops)
// so a thread computes one padded address per phase and the rest are immediate offsets.
__host__ __device__ constexpr int toff(int k, int p) { return (k << p) + ((k << p) >> 4); }
constexpr int kPairStep = 2 * kTileThreads + (2 * kTileThreads >> 4);   // 544
constexpr int kWordStep = kTileThreads + (kTileThreads >> 4);           // 272

template <int LOGN>
struct NttGeo {
    static_assert(LOGN >= 12 && LOGN <= 16, "tile NTT supports N = 2^12 .. 2^16");
    static constexpr int A = LOGN / 2;
    static constexpr int B = LOGN - A;
    static constexpr int LOGC = kLogTile - A;      // columns per column-pass tile
    static constexpr int LOGD = kLogTile - B;      // rows per row-pass tile
    static constexpr int TILES = 1 << (LOGN - kLogTile);
};

__device__ __forceinline__ int insert_bits3(int t, int p) { return ((t >> p) << (p + 3)) | (t & ((1 << p) - 1)); }
// Which of the 256 element groups a thread takes in a round. In the round with element stride 4 (p = 2) the natural order puts
// the four 4-lane groups of a half-warp on blocks 0,1,2,3 of 32 elements, i.e. 34 doubles apart = 4 banks of 8 bytes: their
// 4-double windows overlap pairwise and every LDS.64 / STS.64 of the round takes two wavefronts per half-warp (ncu: 2x).
// Rotating the low three bits of the group number puts a half-warp on blocks 0,2,4,6 (68 doubles apart = 8 banks): conflict free.
template <int P>
__device__ __forceinline__ int round_thread(int t) {
    if constexpr (P == 2) {
        const int g = t >> 2;
        return ((((g & ~7) | ((g & 3) << 1) | ((g >> 2) & 1))) << 2) | (t & 3);
    } else {
        return t;
    }
}

// ---- forward --------------------------------------------------------------------------------
// Two arithmetic back ends, chosen per prime (block-uniform):
//  * integer (any q < 2^60): Harvey butterflies on Shoup twiddles, values in [0,8q);
//  * FP64 (q < 2^46, i.e. the 45-bit scaling primes = 11 of the 12 Q limbs of approaches 4/5): values are
//    exact integers held in doubles, CENTRED and lazy.  y*w mod q = y*w - rint(y*(w/q))*q is evaluated
//    exactly with an error-free product (h = y*w rounded, l = fma(y,w,-h)) and two more FMAs, 8 DP
//    instructions per butterfly against ~40 integer instructions; |result| <= 0.58q, so a value grows by
//    at most 0.58q per stage: 16 stages from a canonical input stay below 10.3q < 2^50 and the pass needs
//    no comparison or conditional subtraction at all.  All intermediate integers are < 2^53, so every
//    operation is exact and the canonical output is bit-identical to the integer path.
// Between the column and the row pass the words in shared/global memory are "pass-internal": lazy
// integers (integer back end) or double bit patterns (FP64 back end).  fwd_encode() makes them from
// canonical residues, fwd_normalize() turns them back.
__device__ __forceinline__ bool fp_ok(u64 q) { return (q >> 46) == 0; }
constexpr double kMagic52 = 4503599627370496.0;          // 2^52
constexpr double kMagicRint = 6755399441055744.0;        // 1.5 * 2^52: x + M - M = rint(x) for |x| < 2^51
// exact for 0 <= x < 2^52
__device__ __forceinline__ double u64_to_double_small(u64 x) {
    return __dsub_rn(__longlong_as_double((long long)(x | 0x4330000000000000ull)), kMagic52);
}
__device__ __forceinline__ u64 fwd_encode(u64 x, bool fp) {
    return fp ? (u64)__double_as_longlong(u64_to_double_small(x)) : x;
}
// any forward-pass word -> canonical residue
__device__ __forceinline__ u64 fwd_normalize(u64 x, const PrimeConst& pc, bool fp) {
    if (fp) {
        const double v = __longlong_as_double((long long)x), qd = u64_to_double_small(pc.q);
        const double c = __dsub_rn(__fma_rn(v, pc.qinv_d, kMagicRint), kMagicRint);
        double r = __fma_rn(-c, qd, v);                   // in [-q/2 - 1, q/2 + 1], exact
        if (r < 0.0) r = __dadd_rn(r, qd);
        return (u64)__double_as_longlong(__dadd_rn(r, kMagic52)) & 0x000FFFFFFFFFFFFFull;
    }
    const u64 q = pc.q;
    return csub(csub(csub(x, q << 2), q << 1), q);
}

// One round of R consecutive forward stages.  SL0 = first local stage (local stage sl toggles
// local bit 10-sl); S0 = the global stage of SL0; SHIFT0 = S0 - (stage at which `tileoff` has
// weight 1) so that twiddle index = 2^s + (tileoff << (s - S0 + SHIFT0)) + local group.
// The R stages of a round on the 8 register-held elements of a thread (element k of `v` = tile element base + (k << p)); no
// shared-memory access: kernels that feed a round from registers (multiply-accumulate results) or consume it in registers
// (scaling, direct global stores) call these, the *_round_* wrappers below add the shared-memory load and store.
template <int SL0, int R, int S0, int SHIFT0>
__device__ __forceinline__ void fwd_regs_ct(u64 (&v)[8], int base, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    constexpr int p = kLogTile - SL0 - R;
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int kb = R - 1 - r;
        const int b = p + kb;
        const int s = S0 + r;
        const ulonglong2* twp = tw + (1 << s) + (tileoff << (SHIFT0 + r)) + (base >> (b + 1));
#pragma unroll
        for (int h = 0; h < 4; h++) {
            const int lowmask = (1 << kb) - 1;
            const int k0 = ((h & ~lowmask) << 1) | (h & lowmask);
            const int k1 = k0 | (1 << kb);
            const ulonglong2 w = __ldg(twp + (k0 >> (kb + 1)));
            u64 x = v[k0];
            x = x >= q4 ? x - q4 : x;
            const u64 y = shoup_lazy4(v[k1], w.x, w.y, q);
            v[k0] = x + y;
            v[k1] = x - y + q4;
        }
    }
}
template <int P>
__device__ __forceinline__ int round_base(int t) { return insert_bits3(round_thread<P>(t), P); }
template <int SL0, int R, int S0, int SHIFT0>
__device__ __forceinline__ void fwd_round_ct(u64* sm, int t, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    constexpr int p = kLogTile - SL0 - R;
    const int base = round_base<p>(t);
    u64* const sp = sm + tphys(base);
    u64 v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = sp[toff(k, p)];
    fwd_regs_ct<SL0, R, S0, SHIFT0>(v, base, tileoff, tw, q, q4);
#pragma unroll
    for (int k = 0; k < 8; k++) sp[toff(k, p)] = v[k];
}
// FP64 back end of the same round: tw = {w, w/q} as doubles.
template <int SL0, int R, int S0, int SHIFT0>
__device__ __forceinline__ void fwd_regs_fp(double (&v)[8], int base, int tileoff, const double2* __restrict__ tw, double q) {
    constexpr int p = kLogTile - SL0 - R;
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int kb = R - 1 - r;
        const int b = p + kb;
        const int s = S0 + r;
        const double2* twp = tw + (1 << s) + (tileoff << (SHIFT0 + r)) + (base >> (b + 1));
#pragma unroll
        for (int h = 0; h < 4; h++) {
            const int lowmask = (1 << kb) - 1;
            const int k0 = ((h & ~lowmask) << 1) | (h & lowmask);
            const int k1 = k0 | (1 << kb);
            const double2 w = __ldg(twp + (k0 >> (kb + 1)));
            const double y = v[k1];
            const double c = __dsub_rn(__fma_rn(y, w.y, kMagicRint), kMagicRint);   // rint(y * w/q)
            const double hi = __dmul_rn(y, w.x);
            const double lo = __fma_rn(y, w.x, -hi);                               // y*w = hi + lo exactly
            const double m = __dadd_rn(__fma_rn(-c, q, hi), lo);                   // y*w - c*q, |m| <= 0.58 q
            const double x = v[k0];
            v[k0] = __dadd_rn(x, m);
            v[k1] = __dsub_rn(x, m);
        }
    }
}
template <int SL0, int R, int S0, int SHIFT0>
__device__ __forceinline__ void fwd_round_fp(double* sm, int t, int tileoff, const double2* __restrict__ tw, double q) {
    constexpr int p = kLogTile - SL0 - R;
    const int base = round_base<p>(t);
    double* const sp = sm + tphys(base);
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = sp[toff(k, p)];
    fwd_regs_fp<SL0, R, S0, SHIFT0>(v, base, tileoff, tw, q);
#pragma unroll
    for (int k = 0; k < 8; k++) sp[toff(k, p)] = v[k];
}

// Local stages [SL, SL_END) forward; a __syncthreads() after every round.
template <int SL, int SL_END, int S, int SHIFT>
__device__ __forceinline__ void fwd_stages(u64* sm, int t, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    if constexpr (SL < SL_END) {
        constexpr int R = (SL_END - SL >= 3) ? 3 : (SL_END - SL);
        fwd_round_ct<SL, R, S, SHIFT>(sm, t, tileoff, tw, q, q4);
        __syncthreads();
        fwd_stages<SL + R, SL_END, S + R, SHIFT + R>(sm, t, tileoff, tw, q, q4);
    }
}
template <int SL, int SL_END, int S, int SHIFT>
__device__ __forceinline__ void fwd_stages_fp(double* sm, int t, int tileoff, const double2* __restrict__ tw, double q) {
    if constexpr (SL < SL_END) {
        constexpr int R = (SL_END - SL >= 3) ? 3 : (SL_END - SL);
        fwd_round_fp<SL, R, S, SHIFT>(sm, t, tileoff, tw, q);
        __syncthreads();
        fwd_stages_fp<SL + R, SL_END, S + R, SHIFT + R>(sm, t, tileoff, tw, q);
    }
}
// The same passes with the FIRST round fed straight from global memory: `src` is the tile's 2048 contiguous pass-internal words,
// read in the round's own ownership (a warp reads 256 contiguous bytes per element index).  Saves the staging stores, the first
// round's shared loads and one barrier; the tile in shared memory need not be initialised.
template <int SL, int SL_END, int S, int SHIFT>
__device__ __forceinline__ void fwd_stages_g(u64* sm, const u64* src, int t, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    constexpr int R = (SL_END - SL >= 3) ? 3 : (SL_END - SL);
    constexpr int p = kLogTile - SL - R;
    const int base = round_base<p>(t);
    u64 v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = ldg64(src + base + (k << p));
    fwd_regs_ct<SL, R, S, SHIFT>(v, base, tileoff, tw, q, q4);
    u64* const sp = sm + tphys(base);
#pragma unroll
    for (int k = 0; k < 8; k++) sp[toff(k, p)] = v[k];
    __syncthreads();
    fwd_stages<SL + R, SL_END, S + R, SHIFT + R>(sm, t, tileoff, tw, q, q4);
}
template <int SL, int SL_END, int S, int SHIFT>
__device__ __forceinline__ void fwd_stages_fp_g(double* sm, const u64* src, int t, int tileoff, const double2* __restrict__ tw, double q) {
    constexpr int R = (SL_END - SL >= 3) ? 3 : (SL_END - SL);
    constexpr int p = kLogTile - SL - R;
    const int base = round_base<p>(t);
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = __longlong_as_double((long long)ldg64(src + base + (k << p)));
    fwd_regs_fp<SL, R, S, SHIFT>(v, base, tileoff, tw, q);
    double* const sp = sm + tphys(base);
#pragma unroll
    for (int k = 0; k < 8; k++) sp[toff(k, p)] = v[k];
    __syncthreads();
    fwd_stages_fp<SL + R, SL_END, S + R, SHIFT + R>(sm, t, tileoff, tw, q);
}
// column pass: global stages 0..A-1 = local stages 0..A-1, twiddle = 2^s + (row >> (A-s)); no tile offset.
// Shared memory holds pass-internal words (fwd_encode of canonical residues) on entry.
template <int LOGN, class Tables>
__device__ __forceinline__ void fwd_col_pass(u64* sm, int t, const Tables& T, int prime, u64 q) {
    if (fp_ok(q)) fwd_stages_fp<0, NttGeo<LOGN>::A, 0, 0>(reinterpret_cast<double*>(sm), t, 0, T.tw_fwd_d + (size_t)prime * T.N, u64_to_double_small(q));
    else fwd_stages<0, NttGeo<LOGN>::A, 0, 0>(sm, t, 0, T.tw_fwd + (size_t)prime * T.N, q, q << 2);
}
// row pass: global stages A..LOGN-1 = local stages LOGD..10; tile `tile` starts at row tile<<LOGD
template <int LOGN, class Tables>
__device__ __forceinline__ void fwd_row_pass(u64* sm, int t, int tile, const Tables& T, int prime, u64 q) {
    using G = NttGeo<LOGN>;
    // at stage A a row index has weight 1 in the group number; local groups are numbered inside the tile
    if (fp_ok(q)) fwd_stages_fp<G::LOGD, kLogTile, G::A, G::LOGD>(reinterpret_cast<double*>(sm), t, tile, T.tw_fwd_d + (size_t)prime * T.N, u64_to_double_small(q));
    else fwd_stages<G::LOGD, kLogTile, G::A, G::LOGD>(sm, t, tile, T.tw_fwd + (size_t)prime * T.N, q, q << 2);
}

// row pass with its first round fed from global memory (see fwd_stages_g): `src` = the tile's 2048 contiguous pass-internal words
template <int LOGN, class Tables>
__device__ __forceinline__ void fwd_row_pass_g(u64* sm, const u64* src, int t, int tile, const Tables& T, int prime, u64 q) {
    using G = NttGeo<LOGN>;
    if (fp_ok(q)) fwd_stages_fp_g<G::LOGD, kLogTile, G::A, G::LOGD>(reinterpret_cast<double*>(sm), src, t, tile, T.tw_fwd_d + (size_t)prime * T.N, u64_to_double_small(q));
    else fwd_stages_g<G::LOGD, kLogTile, G::A, G::LOGD>(sm, src, t, tile, T.tw_fwd + (size_t)prime * T.N, q, q << 2);
}

// ---- inverse --------------------------------------------------------------------------------
// One round of R consecutive inverse (Gentleman-Sande) stages: local stages SH-1, SH-2, .., SH-R.
// S_TOP = global stage of local stage SH-1; SHIFT_TOP = shift of tileoff at that stage.
template <int SH, int R, int S_TOP, int SHIFT_TOP>
__device__ __forceinline__ void inv_regs_ct(u64 (&v)[8], int base, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    constexpr int p = kLogTile - SH;
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int kb = r;
        const int b = p + kb;
        const int s = S_TOP - r;
        const ulonglong2* twp = tw + (1 << s) + (tileoff << (SHIFT_TOP - r)) + (base >> (b + 1));
#pragma unroll
        for (int h = 0; h < 4; h++) {
            const int lowmask = (1 << kb) - 1;
            const int k0 = ((h & ~lowmask) << 1) | (h & lowmask);
            const int k1 = k0 | (1 << kb);
            const ulonglong2 w = __ldg(twp + (k0 >> (kb + 1)));
            const u64 x = v[k0], y = v[k1];
            u64 s_ = x + y;
            s_ = s_ >= q4 ? s_ - q4 : s_;
            v[k0] = s_;
            v[k1] = shoup_lazy4(x - y + q4, w.x, w.y, q);
        }
    }
}
template <int SH, int R, int S_TOP, int SHIFT_TOP>
__device__ __forceinline__ void inv_round_ct(u64* sm, int t, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    constexpr int p = kLogTile - SH;
    const int base = round_base<p>(t);
    u64* const sp = sm + tphys(base);
    u64 v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = sp[toff(k, p)];
    inv_regs_ct<SH, R, S_TOP, SHIFT_TOP>(v, base, tileoff, tw, q, q4);
#pragma unroll
    for (int k = 0; k < 8; k++) sp[toff(k, p)] = v[k];
}
// Local stages SH-1 .. SL (descending); remainder round first so the radix-8 rounds are aligned.
template <int SL, int SH, int S_TOP, int SHIFT_TOP>
__device__ __forceinline__ void inv_stages(u64* sm, int t, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    if constexpr (SH > SL) {
        constexpr int rem = (SH - SL) % 3;
        constexpr int R = rem == 0 ? 3 : rem;
        inv_round_ct<SH, R, S_TOP, SHIFT_TOP>(sm, t, tileoff, tw, q, q4);
        __syncthreads();
        inv_stages<SL, SH - R, S_TOP - R, SHIFT_TOP - R>(sm, t, tileoff, tw, q, q4);
    }
}
template <int LOGN>
__device__ __forceinline__ void inv_row_pass(u64* sm, int t, int tile, const ulonglong2* __restrict__ tw, u64 q) {
    using G = NttGeo<LOGN>;
    // top stage LOGN-1: shift of the tile offset there = LOGD + (LOGN-1-A) = 10
    inv_stages<G::LOGD, kLogTile, LOGN - 1, kLogTile - 1>(sm, t, tile, tw, q, q << 2);
}
// The same with the LAST round written straight to global memory (`dst` = the tile's 2048 contiguous words; a warp writes 256
// contiguous bytes per element index): no final shared stores, no re-read in another ownership, one barrier fewer.  Values [0,4q).
template <int SL, int SH, int S_TOP, int SHIFT_TOP>
__device__ __forceinline__ void inv_stages_g(u64* sm, u64* dst, int t, int tileoff, const ulonglong2* __restrict__ tw, u64 q, u64 q4) {
    static_assert(SH > SL, "at least one stage");
    constexpr int rem = (SH - SL) % 3;
    constexpr int R = rem == 0 ? 3 : rem;
    if constexpr (SH - R == SL) {
        constexpr int p = kLogTile - SH;
        const int base = round_base<p>(t);
        const u64* const sp = sm + tphys(base);
        u64 v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = sp[toff(k, p)];
        inv_regs_ct<SH, R, S_TOP, SHIFT_TOP>(v, base, tileoff, tw, q, q4);
#pragma unroll
        for (int k = 0; k < 8; k++) dst[base + (k << p)] = v[k];
    } else {
        inv_round_ct<SH, R, S_TOP, SHIFT_TOP>(sm, t, tileoff, tw, q, q4);
        __syncthreads();
        inv_stages_g<SL, SH - R, S_TOP - R, SHIFT_TOP - R>(sm, dst, t, tileoff, tw, q, q4);
    }
}
template <int LOGN>
__device__ __forceinline__ void inv_row_pass_g(u64* sm, u64* dst, int t, int tile, const ulonglong2* __restrict__ tw, u64 q) {
    using G = NttGeo<LOGN>;
    inv_stages_g<G::LOGD, kLogTile, LOGN - 1, kLogTile - 1>(sm, dst, t, tile, tw, q, q << 2);
}
template <int LOGN>
__device__ __forceinline__ void inv_col_pass(u64* sm, int t, const ulonglong2* __restrict__ tw, u64 q) {
    using G = NttGeo<LOGN>;
    inv_stages<0, G::A, G::A - 1, 16>(sm, t, 0, tw, q, q << 2);   // tile offset unused (0); 16 keeps shifts non-negative
}

// ---- tile addressing ------------------------------------------------------------------------
// Column tile: local index l = r*C + c  <->  limb offset r*2^B + tile*C + c.
template <int LOGN>
__device__ __forceinline__ size_t col_offset(int tile, int l) {
    using G = NttGeo<LOGN>;
    return ((size_t)(l >> G::LOGC) << G::B) + ((size_t)tile << G::LOGC) + (l & ((1 << G::LOGC) - 1));
}

}  // namespace hx
