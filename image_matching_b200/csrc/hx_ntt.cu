// hx_ntt.cu -- batched negacyclic NTT / iNTT over RNS limbs for sm_100a.
//
// Same transform as OpenFHE's ChineseRemainderTransformFTT (forward: Cooley-Tukey, natural order
// in, bit-reversed order out, out[i] = f(psi^(2*brev(i)+1)); inverse: Gentleman-Sande + N^-1),
// reached from every key switch / rescale on the reference path (sender_diag.cpp:25,79,80).
//
// Layout: a limb is N contiguous u64. One limb (N = 2^15: 256 KiB) exceeds shared memory, so the
// transform is split into two launches over 2048-element shared-memory tiles:
//   column pass : stages 0..A-1   (butterfly distance >= 2^B): tile = 2^A rows x C adjacent columns
//   row pass    : stages A..logN-1 (distance < 2^B)          : tile = D contiguous 2^B-element rows
// Inside a tile each thread keeps 8 elements in registers and runs up to three stages (radix-8)
// between shared-memory exchanges. Butterflies are Harvey lazy butterflies on Shoup twiddles:
// values stay in [0,4q) (forward) / [0,2q) (inverse) and are normalised to [0,q) on the way out,
// so results are the canonical residues.
#include "hx_kernels.cuh"
#include "hx_ntt_tile.cuh"

namespace hx {

constexpr int kTileLog = 11;               // 2048 elements = 16 KiB (+ padding) per tile
constexpr int kTile = 1 << kTileLog;
constexpr int kEPT = 8;                    // elements per thread
__device__ __forceinline__ int phys(int e) { return e + (e >> 4); }   // pad 1 word per 16 against bank conflicts
constexpr int kSmemWords = kTile + (kTile >> 4) + 8;

struct NttGeom {
    int logN, A, B;       // N = 2^(A+B)
    int logT;             // tile = 2^logT elements
    int logC;             // column pass: C = 2^logC columns per tile (logT = A + logC)
    int logD;             // row pass: D = 2^logD rows per tile      (logT = B + logD)
};

__device__ __forceinline__ int insert3(int t, int pp) {
    return ((t >> pp) << (pp + 3)) | (t & ((1 << pp) - 1));
}

// One forward round: R (<=3) consecutive Cooley-Tukey stages sg..sg+R-1 on the 8 elements a thread
// holds (local index bits [p, p+3), p = logT - sg - R). Local stage s toggles local bit logT-1-s;
// the twiddle group of element e at that stage is e >> (logT - s).
template <int R, bool kRowPass>
__device__ __forceinline__ void fwd_round(u64* sm, int t, int logT, int sg, int sglob0, int A, int chunk0,
                                          const ulonglong2* __restrict__ tw, u64 q, u64 q2) {
    const int p = logT - sg - R;
    const int base = insert3(t, p);
    u64 v[kEPT];
#pragma unroll
    for (int k = 0; k < kEPT; k++) v[k] = sm[phys(base + (k << p))];
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int kb = R - 1 - r;           // bit inside k toggled by this stage
        const int b = p + kb;               // local bit
        const int s = sglob0 + sg + r;      // global stage
        const int twb = (1 << s) + (kRowPass ? (chunk0 << (s - A)) : 0);
#pragma unroll
        for (int h = 0; h < kEPT / 2; h++) {
            const int lowmask = (1 << kb) - 1;
            const int k0 = ((h & ~lowmask) << 1) | (h & lowmask);
            const int k1 = k0 | (1 << kb);
            const int e = base + (k0 << p);
            ulonglong2 w = __ldg(tw + twb + (e >> (b + 1)));
            u64 x = v[k0];
            x = x >= q2 ? x - q2 : x;
            u64 y = shoup_lazy(v[k1], w.x, w.y, q);
            v[k0] = x + y;
            v[k1] = x - y + q2;
        }
    }
#pragma unroll
    for (int k = 0; k < kEPT; k++) sm[phys(base + (k << p))] = v[k];
    __syncthreads();
}

// Forward local stages [s0, s1): full radix-8 rounds first, remainder (1 or 2 stages) last.
template <bool kRowPass>
__device__ __forceinline__ void fwd_rounds(u64* sm, int t, int logT, int s0, int s1, int sglob0, int A, int chunk0,
                                           const ulonglong2* __restrict__ tw, u64 q, u64 q2) {
    int sg = s0;
    for (; sg + 3 <= s1; sg += 3) fwd_round<3, kRowPass>(sm, t, logT, sg, sglob0, A, chunk0, tw, q, q2);
    if (s1 - sg == 2) fwd_round<2, kRowPass>(sm, t, logT, sg, sglob0, A, chunk0, tw, q, q2);
    else if (s1 - sg == 1) fwd_round<1, kRowPass>(sm, t, logT, sg, sglob0, A, chunk0, tw, q, q2);
}

// One inverse round: Gentleman-Sande stages sh-1, sh-2, .., sh-R (local), elements at bits [p, p+3),
// p = logT - sh; stage sh-1-r toggles bit p + r. Values stay in [0,2q).
template <int R, bool kRowPass>
__device__ __forceinline__ void inv_round(u64* sm, int t, int logT, int sh, int sglob0, int A, int chunk0,
                                          const ulonglong2* __restrict__ tw, u64 q, u64 q2) {
    const int p = logT - sh;
    const int base = insert3(t, p);
    u64 v[kEPT];
#pragma unroll
    for (int k = 0; k < kEPT; k++) v[k] = sm[phys(base + (k << p))];
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int kb = r;
        const int b = p + kb;
        const int s = sglob0 + sh - 1 - r;
        const int twb = (1 << s) + (kRowPass ? (chunk0 << (s - A)) : 0);
#pragma unroll
        for (int h = 0; h < kEPT / 2; h++) {
            const int lowmask = (1 << kb) - 1;
            const int k0 = ((h & ~lowmask) << 1) | (h & lowmask);
            const int k1 = k0 | (1 << kb);
            const int e = base + (k0 << p);
            ulonglong2 w = __ldg(tw + twb + (e >> (b + 1)));
            u64 x = v[k0], y = v[k1];
            u64 s_ = x + y;
            s_ = s_ >= q2 ? s_ - q2 : s_;
            u64 d = x - y + q2;
            v[k0] = s_;
            v[k1] = shoup_lazy(d, w.x, w.y, q);
        }
    }
#pragma unroll
    for (int k = 0; k < kEPT; k++) sm[phys(base + (k << p))] = v[k];
    __syncthreads();
}

// Inverse local stages s1-1 .. s0: remainder round first (smallest distances), then radix-8 rounds.
template <bool kRowPass>
__device__ __forceinline__ void inv_rounds(u64* sm, int t, int logT, int s0, int s1, int sglob0, int A, int chunk0,
                                           const ulonglong2* __restrict__ tw, u64 q, u64 q2) {
    int sh = s1;
    int rem = (s1 - s0) % 3;
    if (rem == 2) { inv_round<2, kRowPass>(sm, t, logT, sh, sglob0, A, chunk0, tw, q, q2); sh -= 2; }
    else if (rem == 1) { inv_round<1, kRowPass>(sm, t, logT, sh, sglob0, A, chunk0, tw, q, q2); sh -= 1; }
    for (; sh > s0; sh -= 3) inv_round<3, kRowPass>(sm, t, logT, sh, sglob0, A, chunk0, tw, q, q2);
}

// grid: (tiles per limb, nbatch). Row pass: tile = contiguous 2^logT elements.
template <bool kInverse>
__global__ void __launch_bounds__(kTile / kEPT) k_ntt_row(DevTables T, NttGeom g, u64* __restrict__ data, LimbMap lm) {
    __shared__ u64 sm[kSmemWords];
    int prime = lm.prime[blockIdx.y % lm.period];
    if (prime < 0) return;                 // limb not part of this transform (block-uniform)
    const PrimeConst pc = T.pc[prime];
    const u64 q = pc.q, q2 = q << 1;
    const int nthr = blockDim.x, t = threadIdx.x;
    const int tile = 1 << g.logT;
    u64* limb = data + (size_t)blockIdx.y * T.N + (size_t)blockIdx.x * tile;
    for (int e = t * 2; e < tile; e += nthr * 2) {
        ulonglong2 x = ldg128(limb + e);
        sm[phys(e)] = x.x;
        sm[phys(e + 1)] = x.y;
    }
    __syncthreads();
    int chunk0 = blockIdx.x << g.logD;
    if (!kInverse) {
        const ulonglong2* tw = T.tw_fwd + (size_t)prime * T.N;
        fwd_rounds<true>(sm, t, g.logT, g.logD, g.logT, g.A - g.logD, g.A, chunk0, tw, q, q2);
        for (int e = t * 2; e < tile; e += nthr * 2) {
            ulonglong2 x;
            x.x = csub(csub(sm[phys(e)], q2), q);
            x.y = csub(csub(sm[phys(e + 1)], q2), q);
            st128(limb + e, x);
        }
    } else {
        const ulonglong2* tw = T.tw_inv + (size_t)prime * T.N;
        inv_rounds<true>(sm, t, g.logT, g.logD, g.logT, g.A - g.logD, g.A, chunk0, tw, q, q2);
        bool last = (g.A == 0);   // single-pass transform: apply N^-1 here
        for (int e = t * 2; e < tile; e += nthr * 2) {
            ulonglong2 x;
            x.x = sm[phys(e)];
            x.y = sm[phys(e + 1)];
            if (last) {
                x.x = shoup_mul(x.x, pc.ninv, pc.ninv_sh, q);
                x.y = shoup_mul(x.y, pc.ninv, pc.ninv_sh, q);
            } else {
                x.x = csub(x.x, q);
                x.y = csub(x.y, q);
            }
            st128(limb + e, x);
        }
    }
}

// Column pass: tile = 2^A rows x C columns; element (r, c) lives at global r*2^B + col0 + c, local r*C + c.
template <bool kInverse>
__global__ void __launch_bounds__(kTile / kEPT) k_ntt_col(DevTables T, NttGeom g, u64* __restrict__ data, LimbMap lm) {
    __shared__ u64 sm[kSmemWords];
    int prime = lm.prime[blockIdx.y % lm.period];
    if (prime < 0) return;                 // limb not part of this transform (block-uniform)
    const PrimeConst pc = T.pc[prime];
    const u64 q = pc.q, q2 = q << 1;
    const int nthr = blockDim.x, t = threadIdx.x;
    const int tile = 1 << g.logT;
    const int cmask = (1 << g.logC) - 1;
    u64* limb = data + (size_t)blockIdx.y * T.N + ((size_t)blockIdx.x << g.logC);
    for (int e = t; e < tile; e += nthr) {
        int r = e >> g.logC, c = e & cmask;
        sm[phys(e)] = ldg64(limb + ((size_t)r << g.B) + c);
    }
    __syncthreads();
    if (!kInverse) {
        const ulonglong2* tw = T.tw_fwd + (size_t)prime * T.N;
        fwd_rounds<false>(sm, t, g.logT, 0, g.A, 0, g.A, 0, tw, q, q2);
        for (int e = t; e < tile; e += nthr) {
            int r = e >> g.logC, c = e & cmask;
            limb[((size_t)r << g.B) + c] = csub(csub(sm[phys(e)], q2), q);
        }
    } else {
        const ulonglong2* tw = T.tw_inv + (size_t)prime * T.N;
        inv_rounds<false>(sm, t, g.logT, 0, g.A, 0, g.A, 0, tw, q, q2);
        for (int e = t; e < tile; e += nthr) {
            int r = e >> g.logC, c = e & cmask;
            limb[((size_t)r << g.B) + c] = shoup_mul(sm[phys(e)], pc.ninv, pc.ninv_sh, q);
        }
    }
}

// ---- compile-time geometry kernels (N = 2^12 .. 2^16): grid (tiles per limb, nbatch) ----------
template <int LOGN, bool kInverse>
__global__ void __launch_bounds__(kTileThreads) k_ntt_row_t(const __grid_constant__ DevTables T, u64* __restrict__ data,
                                                            const __grid_constant__ LimbMap lm) {
    __shared__ u64 sm[kTileWords];
    const int prime = lm.prime[blockIdx.y % lm.period];
    if (prime < 0) return;                 // limb not part of this transform (block-uniform)
    const PrimeConst pc = T.pc[prime];
    const u64 q = pc.q;
    const int t = threadIdx.x, tile = blockIdx.x;
    u64* limb = data + (size_t)blockIdx.y * T.N + (size_t)tile * kTileN;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        const ulonglong2 x = ldg128(limb + e);
        sm[tphys(e)] = x.x;
        sm[tphys(e + 1)] = x.y;
    }
    __syncthreads();
    if (!kInverse) {
        fwd_row_pass<LOGN>(sm, t, tile, T, prime, q);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = (u * kTileThreads + t) * 2;
            ulonglong2 x;
            x.x = fwd_normalize(sm[tphys(e)], pc, fp_ok(q));
            x.y = fwd_normalize(sm[tphys(e + 1)], pc, fp_ok(q));
            st128(limb + e, x);
        }
    } else {
        inv_row_pass<LOGN>(sm, t, tile, T.tw_inv + (size_t)prime * T.N, q);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = (u * kTileThreads + t) * 2;
            ulonglong2 x = {sm[tphys(e)], sm[tphys(e + 1)]};   // [0,2q): the column pass finishes the transform
            st128(limb + e, x);
        }
    }
}
template <int LOGN, bool kInverse>
__global__ void __launch_bounds__(kTileThreads) k_ntt_col_t(const __grid_constant__ DevTables T, u64* __restrict__ data,
                                                            const __grid_constant__ LimbMap lm) {
    __shared__ u64 sm[kTileWords];
    const int prime = lm.prime[blockIdx.y % lm.period];
    if (prime < 0) return;
    const PrimeConst pc = T.pc[prime];
    const u64 q = pc.q;
    const int t = threadIdx.x, tile = blockIdx.x;
    u64* limb = data + (size_t)blockIdx.y * T.N;
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const int l = u * kTileThreads + t;
        const u64 x = ldg64(limb + col_offset<LOGN>(tile, l));
        sm[tphys(l)] = kInverse ? x : fwd_encode(x, fp_ok(q));
    }
    __syncthreads();
    if (!kInverse) {
        fwd_col_pass<LOGN>(sm, t, T, prime, q);
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int l = u * kTileThreads + t;
            limb[col_offset<LOGN>(tile, l)] = sm[tphys(l)];   // [0,4q): the row pass normalises
        }
    } else {
        inv_col_pass<LOGN>(sm, t, T.tw_inv + (size_t)prime * T.N, q);
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int l = u * kTileThreads + t;
            limb[col_offset<LOGN>(tile, l)] = shoup_mul(sm[tphys(l)], pc.ninv, pc.ninv_sh, q);
        }
    }
}
template <int LOGN>
static void launch_ntt_t(const DevTables& T, u64* data, int nbatch, const LimbMap& lm, bool inverse, cudaStream_t st) {
    dim3 grid(NttGeo<LOGN>::TILES, nbatch);
    if (!inverse) {
        k_ntt_col_t<LOGN, false><<<grid, kTileThreads, 0, st>>>(T, data, lm);
        check_launch("k_ntt_col_t");
        k_ntt_row_t<LOGN, false><<<grid, kTileThreads, 0, st>>>(T, data, lm);
        check_launch("k_ntt_row_t");
    } else {
        k_ntt_row_t<LOGN, true><<<grid, kTileThreads, 0, st>>>(T, data, lm);
        check_launch("k_ntt_row_t");
        k_ntt_col_t<LOGN, true><<<grid, kTileThreads, 0, st>>>(T, data, lm);
        check_launch("k_ntt_col_t");
    }
}

static NttGeom make_geom(int logN) {
    NttGeom g;
    g.logN = logN;
    if (logN <= kTileLog) {
        g.A = 0; g.B = logN; g.logT = logN; g.logC = 0; g.logD = 0;
    } else {
        g.A = logN / 2; g.B = logN - g.A; g.logT = kTileLog;
        g.logC = kTileLog - g.A; g.logD = kTileLog - g.B;
    }
    return g;
}

void launch_ntt(const DevTables& T, u64* data, int nbatch, const LimbMap& lm, bool inverse, cudaStream_t st) {
    if (nbatch <= 0) return;
    switch (T.logN) {
        case 12: return launch_ntt_t<12>(T, data, nbatch, lm, inverse, st);
        case 13: return launch_ntt_t<13>(T, data, nbatch, lm, inverse, st);
        case 14: return launch_ntt_t<14>(T, data, nbatch, lm, inverse, st);
        case 15: return launch_ntt_t<15>(T, data, nbatch, lm, inverse, st);
        case 16: return launch_ntt_t<16>(T, data, nbatch, lm, inverse, st);
        default: break;
    }
    NttGeom g = make_geom(T.logN);
    int threads = (1 << g.logT) / kEPT;
    dim3 grid_row(T.N >> g.logT, nbatch), grid_col(T.N >> g.logT, nbatch);
    if (!inverse) {
        if (g.A > 0) k_ntt_col<false><<<grid_col, threads, 0, st>>>(T, g, data, lm);
        check_launch("k_ntt_col");
        k_ntt_row<false><<<grid_row, threads, 0, st>>>(T, g, data, lm);
        check_launch("k_ntt_row");
    } else {
        k_ntt_row<true><<<grid_row, threads, 0, st>>>(T, g, data, lm);
        check_launch("k_ntt_row");
        if (g.A > 0) k_ntt_col<true><<<grid_col, threads, 0, st>>>(T, g, data, lm);
        check_launch("k_ntt_col");
    }
}

}  // namespace hx
