// hx_fused.cu -- fused hybrid key-switch pipeline for sm_100a:
// EvalFastKeySwitchCoreExt + ApproxModDown + "+ c0" + automorphism of OpenFHE v1.2.3, as reached
// from cc->EvalFastRotation (/root/reference/src/sender/sender_diag.cpp:25), RelinearizeInPlace
// (:79) and EvalRotate (src/openFHE_wrapper.cpp:124), in FOUR launches per batch of keys, every
// CTA owning one 2048-element tile of one limb (18 KB of shared memory, 3-4 CTAs per SM):
//
//   K1  k_ks_p_inner   P-limb inner product  sum_j d_j (.) evk_j  fused with the inverse row pass
//   K2a k_inv_cols     inverse column pass of the P limbs, scaled by (P/p_k)^-1 * N^-1
//   K2b k_conv_fwd_col approximate base conversion to one Q limb fused with its forward column pass
//   K3  k_ks_q_finish  Q-limb inner product computed on the fly, forward row pass of the converted
//                      limb, (ip - conv) * P^-1 + addend, written through the Galois permutation
//
// The evaluation key is streamed from HBM exactly once (its P limbs in K1, its Q limbs in K3); the
// 2 x (l+K)-limb extended ciphertext of the textbook pipeline is never materialised.  The only
// intermediates are the 2 x K partially transformed P limbs and the 2 x l converted limbs.
// All results are canonical residues: bit-identical to the unfused kernels and to the CPU oracle.
#include "hx_kernels.cuh"
#include "hx_ntt_tile.cuh"
#include "hx_fpmod.cuh"

#ifndef HX_MINB
#define HX_MINB 3
#endif
namespace hx {

int g_pdl = 1;
template <class K>
static void smem_optin(K kernel, unsigned bytes) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) throw std::runtime_error(std::string("cudaFuncSetAttribute (dynamic shared memory opt-in): ") + cudaGetErrorString(e));
}

constexpr unsigned kTileBytes = kTileN * sizeof(u64);                                   // 16 KB
constexpr unsigned kStageOff = (kTileWords * sizeof(u64) + 127u) & ~127u;              // staged key tiles start here
constexpr unsigned ks_smem_bytes(int ndig) { return kStageOff + ndig * kTileBytes + 16; }

// PrecomputeAutoMap on the fly: position J (bit-reversed order) of the output reads position
// brev(((2*brev(J)+1)*g mod 2N) >> 1) of the input.
template <int LOGN>
__device__ __forceinline__ u32 automap_at(u32 J, u32 g) {
    const u32 j = __brev(J) >> (32 - LOGN);
    const u32 idx = (((2u * j + 1u) * g) & ((2u << LOGN) - 1u)) >> 1;
    return __brev(idx) >> (32 - LOGN);
}

// ---------------------------------------------------------------------------------------------
// K1: grid (TILES, 2*K, nb).  out: tmpP[b][comp][k][N], values in [0,4q), inverse row pass done.
template <int LOGN, int NDIG>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_ks_p_inner(const __grid_constant__ DevTables T, const __grid_constant__ FusedKsArgs a) {
    extern __shared__ __align__(128) unsigned char smraw[];
    u64* sm = reinterpret_cast<u64*>(smraw);
    u64* kst = reinterpret_cast<u64*>(smraw + kStageOff);
    u64* mbar = reinterpret_cast<u64*>(smraw + kStageOff + NDIG * kTileBytes);
    const int t = threadIdx.x, tile = blockIdx.x;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    const int k = blockIdx.y >> 1, comp = blockIdx.y & 1, b = blockIdx.z;
    const int prime = a.L + k, dl = a.nl + k, ext = a.nl + a.K;
    const PrimeConst pc = T.pc[prime];
    const u64* key = (comp == 0 ? a.keyB[b] : a.keyA[b]) + (size_t)prime * T.N + (size_t)tile * kTileN;
    const u64* dig = a.digits + (size_t)b * a.digits_bs + (size_t)dl * T.N + (size_t)tile * kTileN;
    const size_t kstride = (size_t)(a.L + a.K) * T.N, dstride = (size_t)ext * T.N;
    if (t == 0) {   // key tiles by TMA bulk copy (see k_ks_q_finish)
        mbar_init(mbar, 1);
        mbar_expect_tx(mbar, NDIG * kTileBytes);
#pragma unroll
        for (int j = 0; j < NDIG; j++) bulk_g2s(kst + j * kTileN, key + j * kstride, kTileBytes, mbar);
    }
    pdl_wait();      // the evaluation keys above are static; everything below reads the previous kernels' output
    __syncthreads();
    mbar_wait(mbar, 0);
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        acc3 a0 = {0, 0, 0}, a1 = {0, 0, 0};
#pragma unroll
        for (int j = 0; j < NDIG; j++) {
            const ulonglong2 d = ldg128(dig + j * dstride + e);
            const ulonglong2 kv = *reinterpret_cast<const ulonglong2*>(kst + j * kTileN + e);
            mac30ss(a0, d.x, kv.x);
            mac30ss(a1, d.y, kv.y);
        }
        smp[u * kPairStep] = redc(fold30(a0), pc.q, pc.qinv_neg);
        smp[u * kPairStep + 1] = redc(fold30(a1), pc.q, pc.qinv_neg);
    }
    __syncthreads();
    u64* o = a.tmpP + (((size_t)b * 2 + comp) * a.K + k) * T.N + (size_t)tile * kTileN;
    inv_row_pass_g<LOGN>(sm, o, t, tile, T.tw_inv + (size_t)prime * T.N, pc.q);   // last round straight to global memory
}

// ---------------------------------------------------------------------------------------------
// Conversion tables of one launch: a single table (ModDown, rescale) or the per-digit tables of a ModUp,
// whose digits then share ONE launch of each stage (grid.y runs over the limbs / targets of all digits).
struct TabSet {
    const BaseConvTable* t[kMaxDigits];
    int ndig;
    int alpha;                 // source limbs per digit (inverse-column stage: digit of limb y = min(y / alpha, ndig - 1))
    int ystart[kMaxDigits + 1]; // conversion stage: targets of digit d are grid.y in [ystart[d], ystart[d+1])
    size_t in_off[kMaxDigits];  // conversion stage: words added to the input / output pointer for digit d
    size_t out_off[kMaxDigits];
};
static TabSet single_table(const BaseConvTable* tab) {
    TabSet s{};
    s.t[0] = tab; s.ndig = 1; s.alpha = 1 << 20; s.ystart[0] = 0; s.ystart[1] = 1 << 20;
    return s;
}

// K2a: grid (TILES, n_in, nbatch), in place on [nbatch][n_in][N]: inverse column pass, then
// x * hatinv_n  (hatinv_n = (Qhat_k^-1 * N^-1) mod q_k folds the N^-1 of the inverse transform into
// the first multiplication of ApproxSwitchCRTBasis).  Output canonical, coefficient order.
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_inv_cols(const __grid_constant__ DevTables T, const __grid_constant__ TabSet ts,
                                                              u64* __restrict__ data, size_t bs) {
    pdl_wait();
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, tile = blockIdx.x;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    const int dgt = min((int)blockIdx.y / ts.alpha, ts.ndig - 1), k = blockIdx.y - dgt * ts.alpha;
    const BaseConvTable* __restrict__ tab = ts.t[dgt];
    const int prime = tab->src[k];
    const u64 q = T.pc[prime].q;
    u64* limb = data + (size_t)blockIdx.z * bs + (size_t)blockIdx.y * T.N;
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const int l = u * kTileThreads + t;
        smw[u * kWordStep] = ldg64(limb + col_offset<LOGN>(tile, l));
    }
    __syncthreads();
    inv_col_pass<LOGN>(sm, t, T.tw_inv + (size_t)prime * T.N, q);
    const u64 h = tab->hatinv_n[k], hs = tab->hatinv_n_sh[k];
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const int l = u * kTileThreads + t;
        limb[col_offset<LOGN>(tile, l)] = split30(shoup_mul(smw[u * kWordStep], h, hs, q));   // conversion source, stored split
    }
}

// K2b: grid (TILES, ceil(targets / jgroup), nbatch).  in: [nbatch][n_in][N] scaled coefficients (K2a);
// out: [nbatch][n_out][N], forward column pass done, values pass-internal (see fwd_encode).
// The n_in source column tiles are staged into shared memory once per CTA by cp.async (all of them in
// flight together, no registers held) and converted to `jgroup` target limbs from there: a large batch
// re-reads the sources from L2 n_out / jgroup times instead of n_out times.
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_conv_fwd_col(const __grid_constant__ DevTables T, const __grid_constant__ TabSet ts,
                                                                  const u64* __restrict__ in, size_t in_bs, u64* __restrict__ out, size_t out_bs,
                                                                  int ext_nl, int L, int jgroup) {
    pdl_wait();
    using G = NttGeo<LOGN>;
    extern __shared__ __align__(128) unsigned char smraw[];
    u64* sm = reinterpret_cast<u64*>(smraw);
    u64* stage = reinterpret_cast<u64*>(smraw + kStageOff);          // [n_in][2048], local index r*C + c
    const int t = threadIdx.x, tile = blockIdx.x;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    int dgt = 0;
    while (dgt + 1 < ts.ndig && (int)blockIdx.y >= ts.ystart[dgt + 1]) dgt++;
    const BaseConvTable* __restrict__ tab = ts.t[dgt];
    in += ts.in_off[dgt];
    out += ts.out_off[dgt];
    const int n_in = tab->n_in, n_out = tab->n_out;
    const int j0 = (blockIdx.y - ts.ystart[dgt]) * jgroup, j1 = min(n_out, j0 + jgroup);
    const u64* src = in + (size_t)blockIdx.z * in_bs + ((size_t)tile << G::LOGC);
    constexpr int kSegs = (1 << G::LOGC) / 2;                         // 16-byte chunks per tile row
    {
        for (int k = 0; k < n_in; k++) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int ch = i * kTileThreads + t, r = ch / kSegs, sg = ch % kSegs;
                cp_async16(stage + k * kTileN + (r << G::LOGC) + sg * 2, src + (size_t)k * T.N + ((size_t)r << G::B) + sg * 2);
            }
        }
        asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
        __syncthreads();
    }
    for (int j = j0; j < j1; j++) {
        const int prime = tab->dst[j];
        const PrimeConst pc = T.pc[prime];
        acc3 acc[8];
#pragma unroll
        for (int u = 0; u < 8; u++) acc[u] = {0, 0, 0};
        for (int k = 0; k < n_in; k++) {   // n_in <= 7 (digit width / number of P limbs)
            const u64 m = tab->mat30[k * n_out + j];
#pragma unroll
            for (int u = 0; u < 8; u++) mac30ss(acc[u], stage[k * kTileN + u * kTileThreads + t], m);
        }
#pragma unroll
        for (int u = 0; u < 8; u++) smw[u * kWordStep] = fwd_encode(redc(fold30(acc[u]), pc.q, pc.qinv_neg), fp_ok(pc.q));
        __syncthreads();
        fwd_col_pass<LOGN>(sm, t, T, prime, pc.q);
        // ext_nl >= 0 (ModUp): the output is a digit in extended-basis order (Q limbs 0..nl-1, then P); else limb j
        const int pos = ext_nl < 0 ? j : (prime < L ? prime : ext_nl + prime - L);
        u64* dst = out + (size_t)blockIdx.z * out_bs + (size_t)pos * T.N;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int l = u * kTileThreads + t;
            dst[col_offset<LOGN>(tile, l)] = smw[u * kWordStep];
        }
        __syncthreads();   // the work tile is reused by the next target limb
    }
}
// launch helper: dynamic shared memory = work tile + n_in staged source tiles
template <int LOGN>
static void launch_conv_fwd_col(const DevTables& T, const TabSet& ts, int max_n_in, int ny, int nbatch, const u64* in, size_t in_bs, u64* out,
                                size_t out_bs, int ext_nl, int L, int jgroup, cudaStream_t st) {
    const int bytes = (int)(kStageOff + max_n_in * kTileBytes);
    launch_pdl(k_conv_fwd_col<LOGN>, dim3(NttGeo<LOGN>::TILES, ny, nbatch), dim3(kTileThreads), bytes, st, T, ts, in, in_bs, out, out_bs, ext_nl, L, jgroup);
}

// ---------------------------------------------------------------------------------------------
// K3: grid (TILES, 2*nlimbs, nb) over the limbs listed in a.limbs (one launch for the FP64 limbs, one for the 60-bit ones).
//   out[b][comp][i][J] = (ip - conv + addend)[map_g(J)]     (ip, conv carry the factor P^-1: it sits in the key's Q limbs and in the
//   ModDown conversion matrix)
// FP limbs (q < 2^46): the converted tile arrives as lazy doubles from the column pass, the forward row pass runs on doubles, the
// evaluation-key inner product is computed on doubles as well (digits and the key's Q limbs are stored as doubles for these limbs:
// 6 DP instructions per product against 4 IMAD.WIDE + a 128-bit fold + REDC), so the whole epilogue is `ip - conv + c0` in
// doubles and ONE normalisation per coefficient.
// The Galois permutation inside a tile is affine in the natural index: with J = otile*2048 + e (bit-reversed order), jh = brev11(e)
// and jl = brev_h(otile), the source index is (2^h jh g + ((2 jl + 1) g >> 1)) mod N, i.e. source tile = brev_h(c mod 2^h) and
// source offset = brev11(jh*g + (c >> h)) with c = ((2 jl + 1) g) >> 1: one multiply-add and two bit reversals per PAIR of
// outputs (e, e+1 map to src, src^1 because g is odd).
template <int LOGN>
struct AutoTile {
    u32 g, c1;
    int tile;
    __device__ __forceinline__ AutoTile(u32 otile, u32 g_) : g(g_) {
        constexpr int h = LOGN - kLogTile;
        const u32 jl = h ? (__brev(otile) >> (32 - h)) : 0u;
        const u32 c = (((2u * jl + 1u) * g_) & ((2u << LOGN) - 1u)) >> 1;
        tile = h ? (int)(__brev(c << (32 - h)) ) : 0;       // brev_h(c mod 2^h)
        c1 = c >> h;
    }
    // source offset (inside the source tile) of output offset e = 2*m (m < 1024); e + 1 reads src ^ 1
    __device__ __forceinline__ int src_of_pair(u32 m) const {
        const u32 jh = __brev(m) >> 22;                       // brev11(2m) = brev10(m)
        return (int)__brev((jh * g + c1) << 21);              // brev11 of the low 11 bits
    }
};

template <int LOGN, int NDIG, bool FP>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_ks_q_finish(const __grid_constant__ DevTables T, const __grid_constant__ FusedKsArgs a) {
    extern __shared__ __align__(128) unsigned char smraw[];
    u64* sm = reinterpret_cast<u64*>(smraw);                       // work tile
    u64* kst = reinterpret_cast<u64*>(smraw + kStageOff);           // NDIG staged key tiles (TMA)
    u64* mbar = reinterpret_cast<u64*>(smraw + kStageOff + NDIG * kTileBytes);
    const int t = threadIdx.x, otile = blockIdx.x;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    const int i = a.limbs[blockIdx.y >> 1], comp = blockIdx.y & 1, b = blockIdx.z;
    const AutoTile<LOGN> am((u32)otile, a.galois[b]);
    const int tile = am.tile;              // tiles map to tiles under the automorphism
    const PrimeConst pc = T.pc[i];
    const u64 q = pc.q;
    const size_t toff = (size_t)tile * kTileN;
    const u64* cv = a.conv + (((size_t)b * 2 + comp) * a.nl + i) * T.N + toff;
    const u64* key = (comp == 0 ? a.keyB[b] : a.keyA[b]) + (size_t)i * T.N + toff;
    const u64* dig = a.digits + (size_t)b * a.digits_bs + (size_t)i * T.N + toff;
    const size_t kstride = (size_t)(a.L + a.K) * T.N, dstride = (size_t)(a.nl + a.K) * T.N;
    // the key tiles of this CTA (NDIG x 16 KB, each contiguous) are fetched by TMA bulk copies into shared memory while the
    // converted tile is loaded and transformed: 48 KB in flight per CTA without holding registers
    if (t == 0) {
        mbar_init(mbar, 1);
        mbar_expect_tx(mbar, NDIG * kTileBytes);
#pragma unroll
        for (int j = 0; j < NDIG; j++) bulk_g2s(kst + j * kTileN, key + j * kstride, kTileBytes, mbar);
    }
    pdl_wait();      // the evaluation keys above are static; everything below reads the previous kernels' output
    {   // forward row pass, one arithmetic back end per instantiation (the launch lists separate the limbs); its first round reads the
        // converted tile straight from global memory
        using G = NttGeo<LOGN>;
        if constexpr (FP) fwd_stages_fp_g<G::LOGD, kLogTile, G::A, G::LOGD>(reinterpret_cast<double*>(sm), cv, t, tile, T.tw_fwd_d + (size_t)i * T.N, fp_from_u64(q));
        else fwd_stages_g<G::LOGD, kLogTile, G::A, G::LOGD>(sm, cv, t, tile, T.tw_fwd + (size_t)i * T.N, q, q << 2);
    }
    const u64* add = (comp == 0 ? a.add0 : a.add1);
    if (add) add += (size_t)b * a.add_bs + (size_t)i * T.N + toff;
    mbar_wait(mbar, 0);
    if constexpr (FP) {
        const double qd = fp_from_u64(q), qinv = pc.qinv_d;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = (u * kTileThreads + t) * 2;
            double s0 = -fp_bits_to_double(smp[u * kPairStep]), s1 = -fp_bits_to_double(smp[u * kPairStep + 1]);
#pragma unroll
            for (int j = 0; j < NDIG; j++) {
                const ulonglong2 d = ldg128(dig + j * dstride + e);
                const ulonglong2 kv = *reinterpret_cast<const ulonglong2*>(kst + j * kTileN + e);
                s0 = __dadd_rn(s0, fp_mulmod(fp_bits_to_double(d.x), fp_bits_to_double(kv.x), qd, qinv));
                s1 = __dadd_rn(s1, fp_mulmod(fp_bits_to_double(d.y), fp_bits_to_double(kv.y), qd, qinv));
            }
            if (add) {
                const ulonglong2 ad = ldg128(add + e);
                s0 = __dadd_rn(s0, fp_from_u64(ad.x));
                s1 = __dadd_rn(s1, fp_from_u64(ad.y));
            }
            // |s| < 3 * (0.75 q + 2^43) + 10.3 q + q < 2^51: one normalisation gives the canonical residue
            smp[u * kPairStep] = fp_to_u64(fp_normalize(s0, qd, qinv));
            smp[u * kPairStep + 1] = fp_to_u64(fp_normalize(s1, qd, qinv));
        }
    } else {
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = (u * kTileThreads + t) * 2;
            acc3 a0 = {0, 0, 0}, a1 = {0, 0, 0};
#pragma unroll
            for (int j = 0; j < NDIG; j++) {
                const ulonglong2 d = ldg128(dig + j * dstride + e);
                const ulonglong2 kv = *reinterpret_cast<const ulonglong2*>(kst + j * kTileN + e);
                mac30ss(a0, d.x, kv.x);
                mac30ss(a1, d.y, kv.y);
            }
            const u64 c0 = fwd_normalize(smp[u * kPairStep], pc, false), c1 = fwd_normalize(smp[u * kPairStep + 1], pc, false);
            u64 r0 = sub_mod(redc(fold30(a0), q, pc.qinv_neg), c0, q);
            u64 r1 = sub_mod(redc(fold30(a1), q, pc.qinv_neg), c1, q);
            if (add) {
                const ulonglong2 ad = ldg128(add + e);
                r0 = add_mod(r0, ad.x, q);
                r1 = add_mod(r1, ad.y, q);
            }
            smp[u * kPairStep] = r0;
            smp[u * kPairStep + 1] = r1;
        }
    }
    __syncthreads();
    u64* o = a.out + (size_t)b * a.out_bs + ((size_t)comp * a.nl + i) * T.N + (size_t)otile * kTileN;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int m = u * kTileThreads + t;
        const int src = am.src_of_pair((u32)m), p = src & 1;
        const u64* sp = sm + tphys(src & ~1);
        ulonglong2 x;
        x.x = sp[p];
        x.y = sp[p ^ 1];
        st128(o + 2 * m, x);
    }
}

// K3 in two launches: the FP64 limbs (q < 2^46) and the 60-bit limbs (q_0) run different code; a.fp_mask has bit i set for FP limbs
template <int LOGN, int NDIG>
static void launch_q_finish(const DevTables& T, const FusedKsArgs& a0, int nb, cudaStream_t st) {
    using G = NttGeo<LOGN>;
    FusedKsArgs a = a0;
    int n = 0;
    for (int i = 0; i < a.nl; i++)
        if ((a.fp_mask >> i) & 1) a.limbs[n++] = (short)i;
    if (n) launch_pdl(k_ks_q_finish<LOGN, NDIG, true>, dim3(G::TILES, 2 * n, nb), dim3(kTileThreads), ks_smem_bytes(NDIG), st, T, a);
    n = 0;
    for (int i = 0; i < a.nl; i++)
        if (!((a.fp_mask >> i) & 1)) a.limbs[n++] = (short)i;
    if (n) launch_pdl(k_ks_q_finish<LOGN, NDIG, false>, dim3(G::TILES, 2 * n, nb), dim3(kTileThreads), ks_smem_bytes(NDIG), st, T, a);
}
template <int LOGN, int NDIG>
static void launch_fused_tn(const DevTables& T, const FusedKsArgs& a, const BaseConvTable* moddown, int nb, cudaStream_t st) {
    using G = NttGeo<LOGN>;
    launch_pdl(k_ks_p_inner<LOGN, NDIG>, dim3(G::TILES, 2 * a.K, nb), dim3(kTileThreads), ks_smem_bytes(NDIG), st, T, a);
    const TabSet ts = single_table(moddown);
    if (a.moddown_w && moddown_w_supported(LOGN, a.K)) {   // rings of 2^15 / 2^16: the whole coefficient-domain middle as warp tasks
        launch_moddown_w(T, moddown, a.K, nb * 2, a.tmpP, (size_t)a.K * T.N, a.conv, (size_t)a.nl * T.N, st);
        launch_q_finish<LOGN, NDIG>(T, a, nb, st);
        return;
    }
    launch_pdl(k_inv_cols<LOGN>, dim3(G::TILES, a.K, nb * 2), dim3(kTileThreads), 0, st, T, ts, a.tmpP, (size_t)a.K * T.N);
    // large batches: 4 target limbs per CTA share one staging of the K source tiles
    const int jg = (nb * 2 >= 16) ? 4 : 1;
    launch_conv_fwd_col<LOGN>(T, ts, a.K, (a.nl + jg - 1) / jg, nb * 2, a.tmpP, (size_t)a.K * T.N, a.conv, (size_t)a.nl * T.N, -1, 0, jg, st);
    launch_q_finish<LOGN, NDIG>(T, a, nb, st);
}
template <int LOGN>
static void launch_fused_t(const DevTables& T, const FusedKsArgs& a, const BaseConvTable* moddown, int nb, cudaStream_t st) {
    switch (a.ndig) {
        case 1: launch_fused_tn<LOGN, 1>(T, a, moddown, nb, st); break;
        case 2: launch_fused_tn<LOGN, 2>(T, a, moddown, nb, st); break;
        case 3: launch_fused_tn<LOGN, 3>(T, a, moddown, nb, st); break;
        case 4: launch_fused_tn<LOGN, 4>(T, a, moddown, nb, st); break;
        default: throw std::runtime_error("fused key switch: number of digits must be 1..4");
    }
}

// ---------------------------------------------------------------------------------------------
// Fused ModUp (EvalKeySwitchPrecomputeCore, sender_diag.cpp:22) and rescale (DropLastElementAndScale,
// sender_diag.cpp:80) are built from the same tile passes.

// Inverse row pass, out of place: grid (TILES, nlimbs, nbatch); limb y of batch z is read at
// in + z*in_bs + y*N and written to out + z*out_bs + y*N; prime index = prime0 + y.  Values [0,2q).
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_inv_rows(const __grid_constant__ DevTables T, const u64* __restrict__ in, size_t in_bs,
                                                              u64* __restrict__ out, size_t out_bs, int prime0) {
    pdl_wait();
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, tile = blockIdx.x, prime = prime0 + blockIdx.y;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    const u64 q = T.pc[prime].q;
    const u64* src = in + (size_t)blockIdx.z * in_bs + (size_t)blockIdx.y * T.N + (size_t)tile * kTileN;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        const ulonglong2 x = ldg128(src + e);
        smp[u * kPairStep] = x.x;
        smp[u * kPairStep + 1] = x.y;
    }
    __syncthreads();
    u64* dst = out + (size_t)blockIdx.z * out_bs + (size_t)blockIdx.y * T.N + (size_t)tile * kTileN;
    inv_row_pass_g<LOGN>(sm, dst, t, tile, T.tw_inv + (size_t)prime * T.N, q);   // last round straight to global memory
}

// ModUp tail: grid (TILES, ext, ndig*nbatch).  Limb y of digit d: own limbs [lo,hi) are copied from the
// input polynomial, the converted limbs get their forward row pass (in place) and are normalised.
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_modup_rows(const __grid_constant__ DevTables T, u64* __restrict__ digits, const u64* __restrict__ c,
                                                                size_t c_bs, int ndig, int nl, int L, int alpha) {
    pdl_wait();
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, tile = blockIdx.x, y = blockIdx.y;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    const int d = blockIdx.z % ndig, bz = blockIdx.z / ndig;
    const int ext = gridDim.y;
    const int lo = d * alpha, hi = (d == ndig - 1) ? nl : (d + 1) * alpha;
    u64* limb = digits + ((size_t)blockIdx.z * ext + y) * T.N + (size_t)tile * kTileN;
    // digit words as the key-switch inner products read them: doubles on the FP64 limbs (lazy there is fine: |x| < 2^51),
    // 30-bit split words on the 60-bit limbs
    const int prime = y < nl ? y : L + (y - nl);
    const PrimeConst pc = T.pc[prime];
    const u64 q = pc.q;
    const bool fp = fp_ok(q);
    if (y >= lo && y < hi) {
        const u64* src = c + (size_t)bz * c_bs + (size_t)y * T.N + (size_t)tile * kTileN;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = (u * kTileThreads + t) * 2;
            ulonglong2 x = ldg128(src + e);
            if (fp) { x.x = fp_double_to_bits(fp_from_u64(x.x)); x.y = fp_double_to_bits(fp_from_u64(x.y)); }
            else { x.x = split30(x.x); x.y = split30(x.y); }
            st128(limb + e, x);
        }
        return;
    }
    fwd_row_pass_g<LOGN>(sm, limb, t, tile, T, prime, q);   // first round straight from global memory
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        ulonglong2 x = {smp[u * kPairStep], smp[u * kPairStep + 1]};   // FP64 limbs: the row pass's lazy doubles as they are
        if (!fp) { x.x = split30(fwd_normalize(x.x, pc, false)); x.y = split30(fwd_normalize(x.y, pc, false)); }
        st128(limb + e, x);
    }
}

// Rescale, column stage: grid (TILES, nl-1, nbatch).  last: [nbatch][N] coefficients of the dropped limb
// (canonical, K2a with a 1x0 table does inverse columns + N^-1); tmp_i = lift(last) * (-q_l^-1) mod q_i
// (centred lift: v > q_l/2 means v - q_l, and q_l * (-q_l^-1) = -1, hence the "+1"), forward column pass.
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_rescale_cols(const __grid_constant__ DevTables T, const u64* __restrict__ last, size_t last_bs,
                                                                  u64* __restrict__ tmp, size_t tmp_bs, int nl, const __grid_constant__ ScalarTable negqlinv) {
    pdl_wait();
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, tile = blockIdx.x, i = blockIdx.y;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    const u64 q = T.pc[i].q, half = T.pc[nl - 1].q >> 1;
    const u64* src = last + (size_t)blockIdx.z * last_bs;
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const int l = u * kTileThreads + t;
        const u64 v = unsplit30(ldg64(src + col_offset<LOGN>(tile, l)));   // k_inv_cols stores split words
        u64 r = shoup_mul(v, negqlinv.v[i], negqlinv.sh[i], q);
        if (v > half) r = add_mod(r, 1, q);
        smw[u * kWordStep] = fwd_encode(r, fp_ok(q));
    }
    __syncthreads();
    fwd_col_pass<LOGN>(sm, t, T, i, q);
    u64* dst = tmp + (size_t)blockIdx.z * tmp_bs + (size_t)i * T.N;
#pragma unroll
    for (int u = 0; u < 8; u++) {
        const int l = u * kTileThreads + t;
        dst[col_offset<LOGN>(tile, l)] = smw[u * kWordStep];
    }
}
// Rescale, row stage: out_i = c_i * q_l^-1 + NTT(tmp_i).  grid (TILES, nl-1, nbatch).
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads, HX_MINB) k_rescale_rows(const __grid_constant__ DevTables T, u64* __restrict__ out, size_t out_bs,
                                                                  const u64* __restrict__ c, size_t c_bs, const u64* __restrict__ tmp, size_t tmp_bs,
                                                                  const __grid_constant__ ScalarTable qlinv) {
    pdl_wait();
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, tile = blockIdx.x, i = blockIdx.y;
    u64* const smp = sm + tphys(2 * t);   // this thread's element pairs: smp[u * kPairStep + {0,1}]
    u64* const smw = sm + tphys(t);       // this thread's single words:  smw[u * kWordStep]
    const PrimeConst pc = T.pc[i];
    const u64 q = pc.q;
    const size_t off = (size_t)i * T.N + (size_t)tile * kTileN;
    const u64* src = tmp + (size_t)blockIdx.z * tmp_bs + off;
    fwd_row_pass_g<LOGN>(sm, src, t, tile, T, i, q);   // first round straight from global memory
    const bool fp = fp_ok(q);
    const u64* cs = c + (size_t)blockIdx.z * c_bs + off;
    u64* dst = out + (size_t)blockIdx.z * out_bs + off;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        const ulonglong2 x = ldg128(cs + e);
        ulonglong2 r;
        r.x = add_mod(shoup_mul(x.x, qlinv.v[i], qlinv.sh[i], q), fwd_normalize(smp[u * kPairStep], pc, fp), q);
        r.y = add_mod(shoup_mul(x.y, qlinv.v[i], qlinv.sh[i], q), fwd_normalize(smp[u * kPairStep + 1], pc, fp), q);
        st128(dst + e, r);
    }
}

template <int LOGN>
static void modup_t(const DevTables& T, const ModUpArgs& m, cudaStream_t st) {
    using G = NttGeo<LOGN>;
    const int ext = m.nl + m.K;
    launch_pdl(k_inv_rows<LOGN>, dim3(G::TILES, m.nl, m.nbatch), dim3(kTileThreads), 0, st, T, m.c, m.c_bs, m.scratch, (size_t)m.nl * T.N, 0);
    if (m.warp_tasks && moddown_w_supported(LOGN, m.alpha)) {
        // rings of 2^15 / 2^16: inverse column pass, conversion and forward column pass of every digit in ONE launch of warp tasks
        launch_modup_w(T, m, st);
    } else {
        // all digits in one launch per stage: inverse columns of the nl source limbs, then every digit's
        // conversion to its complement limbs fused with their forward column pass
        TabSet ts{};
        ts.ndig = m.ndig; ts.alpha = m.alpha; ts.ystart[0] = 0;
        for (int d = 0; d < m.ndig; d++) {
            const int lo = d * m.alpha, hi = (d == m.ndig - 1) ? m.nl : (d + 1) * m.alpha, sz = hi - lo;
            ts.t[d] = m.tabs[d];
            ts.ystart[d + 1] = ts.ystart[d] + (ext - sz);
            ts.in_off[d] = (size_t)lo * T.N;
            ts.out_off[d] = (size_t)d * ext * T.N;
        }
        launch_pdl(k_inv_cols<LOGN>, dim3(G::TILES, m.nl, m.nbatch), dim3(kTileThreads), 0, st, T, ts, m.scratch, (size_t)m.nl * T.N);
        launch_conv_fwd_col<LOGN>(T, ts, m.alpha, ts.ystart[m.ndig], m.nbatch,
                                  m.scratch, (size_t)m.nl * T.N, m.digits, (size_t)m.ndig * ext * T.N, m.nl, m.L, 1, st);
    }
    launch_pdl(k_modup_rows<LOGN>, dim3(G::TILES, ext, m.ndig * m.nbatch), dim3(kTileThreads), 0, st, T, m.digits, m.c, m.c_bs, m.ndig, m.nl, m.L, m.alpha);
}
void launch_fused_modup(const DevTables& T, const ModUpArgs& m, cudaStream_t st) {
    switch (T.logN) {
        case 12: modup_t<12>(T, m, st); break;
        case 13: modup_t<13>(T, m, st); break;
        case 14: modup_t<14>(T, m, st); break;
        case 15: modup_t<15>(T, m, st); break;
        case 16: modup_t<16>(T, m, st); break;
        default: throw std::runtime_error("fused ModUp: unsupported ring dimension");
    }
}

template <int LOGN>
static void rescale_t(const DevTables& T, const RescaleArgs& r, cudaStream_t st) {
    using G = NttGeo<LOGN>;
    const int nl = r.nl;
    // dropped limb of every component -> coefficients (inverse rows, then inverse columns * N^-1 via the 1-limb table)
    launch_pdl(k_inv_rows<LOGN>, dim3(G::TILES, 1, r.ncomp), dim3(kTileThreads), 0, st, T, r.c + (size_t)(nl - 1) * T.N, r.c_bs, r.last, (size_t)T.N, nl - 1);
    launch_pdl(k_inv_cols<LOGN>, dim3(G::TILES, 1, r.ncomp), dim3(kTileThreads), 0, st, T, single_table(r.last_tab), r.last, (size_t)T.N);
    launch_pdl(k_rescale_cols<LOGN>, dim3(G::TILES, nl - 1, r.ncomp), dim3(kTileThreads), 0, st, T, r.last, (size_t)T.N, r.tmp, (size_t)(nl - 1) * T.N, nl, r.negqlinv);
    launch_pdl(k_rescale_rows<LOGN>, dim3(G::TILES, nl - 1, r.ncomp), dim3(kTileThreads), 0, st, T, r.out, r.out_bs, r.c, r.c_bs, r.tmp, (size_t)(nl - 1) * T.N, r.qlinv);
}
void launch_fused_rescale(const DevTables& T, const RescaleArgs& r, cudaStream_t st) {
    switch (T.logN) {
        case 12: rescale_t<12>(T, r, st); break;
        case 13: rescale_t<13>(T, r, st); break;
        case 14: rescale_t<14>(T, r, st); break;
        case 15: rescale_t<15>(T, r, st); break;
        case 16: rescale_t<16>(T, r, st); break;
        default: throw std::runtime_error("fused rescale: unsupported ring dimension");
    }
}

bool fused_ks_supported(int logN) { return logN >= 12 && logN <= 16; }

template <int LOGN>
static void init_fused_t() {
    const int conv_max = (int)(kStageOff + 7 * kTileBytes);   // digit width / P basis <= 7 limbs (Engine constructor)
    smem_optin(k_conv_fwd_col<LOGN>, conv_max);
    smem_optin(k_ks_p_inner<LOGN, 1>, ks_smem_bytes(1)); smem_optin(k_ks_q_finish<LOGN, 1, true>, ks_smem_bytes(1)); smem_optin(k_ks_q_finish<LOGN, 1, false>, ks_smem_bytes(1));
    smem_optin(k_ks_p_inner<LOGN, 2>, ks_smem_bytes(2)); smem_optin(k_ks_q_finish<LOGN, 2, true>, ks_smem_bytes(2)); smem_optin(k_ks_q_finish<LOGN, 2, false>, ks_smem_bytes(2));
    smem_optin(k_ks_p_inner<LOGN, 3>, ks_smem_bytes(3)); smem_optin(k_ks_q_finish<LOGN, 3, true>, ks_smem_bytes(3)); smem_optin(k_ks_q_finish<LOGN, 3, false>, ks_smem_bytes(3));
    smem_optin(k_ks_p_inner<LOGN, 4>, ks_smem_bytes(4)); smem_optin(k_ks_q_finish<LOGN, 4, true>, ks_smem_bytes(4)); smem_optin(k_ks_q_finish<LOGN, 4, false>, ks_smem_bytes(4));
}
void init_fused_kernels(int logN) {
    init_moddown_kernels(logN);
    switch (logN) {
        case 12: init_fused_t<12>(); break;
        case 13: init_fused_t<13>(); break;
        case 14: init_fused_t<14>(); break;
        case 15: init_fused_t<15>(); break;
        case 16: init_fused_t<16>(); break;
        default: break;   // smaller rings use the unfused kernels only
    }
}

void launch_fused_ks(const DevTables& T, const FusedKsArgs& a, const BaseConvTable* moddown, int nb, cudaStream_t st) {
    switch (T.logN) {
        case 12: launch_fused_t<12>(T, a, moddown, nb, st); break;
        case 13: launch_fused_t<13>(T, a, moddown, nb, st); break;
        case 14: launch_fused_t<14>(T, a, moddown, nb, st); break;
        case 15: launch_fused_t<15>(T, a, moddown, nb, st); break;
        case 16: launch_fused_t<16>(T, a, moddown, nb, st); break;
        default: throw std::runtime_error("fused key switch: unsupported ring dimension");
    }
}

}  // namespace hx
