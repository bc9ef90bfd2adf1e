// hx_fused.cu -- fused hybrid key-switch pipeline for sm_100a:
// EvalFastKeySwitchCoreExt + ApproxModDown + "+ c0" + automorphism of OpenFHE v1.2.3, as reached
// from cc->EvalFastRotation (/root/reference/src/sender/sender_diag.cpp:25), RelinearizeInPlace
// (:79) and EvalRotate (src/openFHE_wrapper.cpp:124), in THREE launches per batch of keys:
//
//   K1 k_ks_p_inner   P-limb inner product  sum_j d_j (.) evk_j  fused with the inverse row pass
//   K2 k_conv_cols    inverse column pass of the P limbs -> P^-1/N^-1 scaling -> approximate base
//                     conversion to the Q limbs -> forward column pass
//   K3 k_ks_q_finish  forward row pass of the converted limbs, Q-limb inner product computed on the
//                     fly, (ip - conv) * P^-1 + addend, written through the Galois permutation
//
// The evaluation key is streamed from HBM exactly once (its P limbs in K1, its Q limbs in K3); the
// 2 x (l+K)-limb extended ciphertext of the textbook pipeline is never materialised.  The only
// intermediates are the 2 x K partially transformed P limbs and the 2 x l converted limbs.
// All results are canonical residues: bit-identical to the unfused kernels and to the CPU oracle.
#include "hx_kernels.cuh"
#include "hx_ntt_tile.cuh"

namespace hx {

// PrecomputeAutoMap on the fly: position J (bit-reversed order) of the output reads position
// brev(((2*brev(J)+1)*g mod 2N) >> 1) of the input.
template <int LOGN>
__device__ __forceinline__ u32 automap_at(u32 J, u32 g) {
    const u32 j = __brev(J) >> (32 - LOGN);
    const u32 idx = (((2u * j + 1u) * g) & ((2u << LOGN) - 1u)) >> 1;
    return __brev(idx) >> (32 - LOGN);
}

// ---------------------------------------------------------------------------------------------
// K1: grid (TILES, 2*K, nb).  out: tmpP[b][comp][k][N], values in [0,2q), inverse row pass done.
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads) k_ks_p_inner(const __grid_constant__ DevTables T, const __grid_constant__ FusedKsArgs a) {
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, tile = blockIdx.x;
    const int k = blockIdx.y >> 1, comp = blockIdx.y & 1, b = blockIdx.z;
    const int prime = a.L + k, dl = a.nl + k, ext = a.nl + a.K;
    const PrimeConst pc = T.pc[prime];
    const u64* key = (comp == 0 ? a.keyB[b] : a.keyA[b]) + (size_t)prime * T.N + (size_t)tile * kTileN;
    const u64* dig = a.digits + (size_t)dl * T.N + (size_t)tile * kTileN;
    const size_t kstride = (size_t)(a.L + a.K) * T.N, dstride = (size_t)ext * T.N;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        u128d a0 = {0, 0}, a1 = {0, 0};
        for (int j = 0; j < a.ndig; j++) {
            const ulonglong2 d = ldg128(dig + j * dstride + e);
            const ulonglong2 kv = ld_stream128(key + j * kstride + e);
            acc_wide(a0, d.x, kv.x);
            acc_wide(a1, d.y, kv.y);
        }
        sm[tphys(e)] = redc(a0, pc.q, pc.qinv_neg);
        sm[tphys(e + 1)] = redc(a1, pc.q, pc.qinv_neg);
    }
    __syncthreads();
    inv_row_pass<LOGN>(sm, t, tile, T.tw_inv + (size_t)prime * T.N, pc.q);
    u64* o = a.tmpP + (((size_t)b * 2 + comp) * a.K + k) * T.N + (size_t)tile * kTileN;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        ulonglong2 x = {sm[tphys(e)], sm[tphys(e + 1)]};
        st128(o + e, x);
    }
}

// ---------------------------------------------------------------------------------------------
// K2: grid (TILES, nbatch).  in : [nbatch][n_in][N]  row-pass-inverted limbs in [0,2q)
//                            out: [nbatch][n_out][N] column-pass-forwarded limbs in [0,4q)
// dynamic shared memory: (n_in + 1) tiles.  hatinv_n = (Qhat_i^-1 * N^-1) mod q_i folds the N^-1 of
// the inverse transform into the first multiplication of ApproxSwitchCRTBasis.
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads) k_conv_cols(DevTables T, const BaseConvTable* __restrict__ tab, const u64* __restrict__ in,
                                                            size_t in_bs, u64* __restrict__ out, size_t out_bs) {
    extern __shared__ u64 smd[];
    const int t = threadIdx.x, tile = blockIdx.x;
    const int n_in = tab->n_in, n_out = tab->n_out;
    const u64* src = in + (size_t)blockIdx.y * in_bs;
    u64* dst = out + (size_t)blockIdx.y * out_bs;
    for (int k = 0; k < n_in; k++) {
        u64* s = smd + (size_t)k * kTileWords;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int l = u * kTileThreads + t;
            s[tphys(l)] = src[(size_t)k * T.N + col_offset<LOGN>(tile, l)];
        }
    }
    __syncthreads();
    for (int k = 0; k < n_in; k++) {
        const int prime = tab->src[k];
        const u64 q = T.pc[prime].q;
        u64* s = smd + (size_t)k * kTileWords;
        inv_col_pass<LOGN>(s, t, T.tw_inv + (size_t)prime * T.N, q);
        const u64 h = tab->hatinv_n[k], hs = tab->hatinv_n_sh[k];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int l = tphys(u * kTileThreads + t);
            s[l] = shoup_mul(s[l], h, hs, q);
        }
    }
    __syncthreads();
    u64* so = smd + (size_t)n_in * kTileWords;
    for (int j = 0; j < n_out; j++) {
        const int prime = tab->dst[j];
        const PrimeConst pc = T.pc[prime];
        u128d acc[8];
#pragma unroll
        for (int u = 0; u < 8; u++) acc[u] = {0, 0};
        for (int k = 0; k < n_in; k++) {
            const u64 m = tab->mat[k * n_out + j];
            const u64* s = smd + (size_t)k * kTileWords;
#pragma unroll
            for (int u = 0; u < 8; u++) acc_wide(acc[u], s[tphys(u * kTileThreads + t)], m);
        }
#pragma unroll
        for (int u = 0; u < 8; u++) so[tphys(u * kTileThreads + t)] = redc(acc[u], pc.q, pc.qinv_neg);
        __syncthreads();
        fwd_col_pass<LOGN>(so, t, T.tw_fwd + (size_t)prime * T.N, pc.q);
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int l = u * kTileThreads + t;
            dst[(size_t)j * T.N + col_offset<LOGN>(tile, l)] = so[tphys(l)];
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// K3: grid (TILES, 2*nl, nb).  out[b][comp][i][J] = ((ip - conv) * P^-1 + addend)[map_g(J)]
template <int LOGN>
__global__ void __launch_bounds__(kTileThreads) k_ks_q_finish(const __grid_constant__ DevTables T, const __grid_constant__ FusedKsArgs a) {
    __shared__ u64 sm[kTileWords];
    const int t = threadIdx.x, otile = blockIdx.x;
    const int i = blockIdx.y >> 1, comp = blockIdx.y & 1, b = blockIdx.z;
    const u32 g = a.galois[b];
    // tiles map to tiles under the automorphism (the low bits of the natural index map among themselves)
    const int tile = (int)(automap_at<LOGN>((u32)otile << kLogTile, g) >> kLogTile);
    const PrimeConst pc = T.pc[i];
    const u64 q = pc.q, q2 = q << 1;
    const size_t toff = (size_t)tile * kTileN;
    const u64* cv = a.conv + (((size_t)b * 2 + comp) * a.nl + i) * T.N + toff;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        const ulonglong2 x = ldg128(cv + e);
        sm[tphys(e)] = x.x;
        sm[tphys(e + 1)] = x.y;
    }
    __syncthreads();
    fwd_row_pass<LOGN>(sm, t, tile, T.tw_fwd + (size_t)i * T.N, q);
    const u64* key = (comp == 0 ? a.keyB[b] : a.keyA[b]) + (size_t)i * T.N + toff;
    const u64* dig = a.digits + (size_t)i * T.N + toff;
    const u64* add = (comp == 0 ? a.add0 : a.add1);
    if (add) add += (size_t)i * T.N + toff;
    const size_t kstride = (size_t)(a.L + a.K) * T.N, dstride = (size_t)(a.nl + a.K) * T.N;
    const u64 pinv = a.pinv.v[i], pinv_sh = a.pinv.sh[i];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        u128d a0 = {0, 0}, a1 = {0, 0};
        for (int j = 0; j < a.ndig; j++) {
            const ulonglong2 d = ldg128(dig + j * dstride + e);
            const ulonglong2 kv = ld_stream128(key + j * kstride + e);
            acc_wide(a0, d.x, kv.x);
            acc_wide(a1, d.y, kv.y);
        }
        const u64 c0 = csub(csub(sm[tphys(e)], q2), q), c1 = csub(csub(sm[tphys(e + 1)], q2), q);
        u64 r0 = shoup_mul(sub_mod(redc(a0, q, pc.qinv_neg), c0, q), pinv, pinv_sh, q);
        u64 r1 = shoup_mul(sub_mod(redc(a1, q, pc.qinv_neg), c1, q), pinv, pinv_sh, q);
        if (add) {
            const ulonglong2 ad = ldg128(add + e);
            r0 = add_mod(r0, ad.x, q);
            r1 = add_mod(r1, ad.y, q);
        }
        sm[tphys(e)] = r0;
        sm[tphys(e + 1)] = r1;
    }
    __syncthreads();
    u64* o = a.out + (size_t)b * a.out_bs + ((size_t)comp * a.nl + i) * T.N + (size_t)otile * kTileN;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int e = (u * kTileThreads + t) * 2;
        const u32 J = ((u32)otile << kLogTile) + e;
        ulonglong2 x;
        x.x = sm[tphys((int)(automap_at<LOGN>(J, g) & (kTileN - 1)))];
        x.y = sm[tphys((int)(automap_at<LOGN>(J + 1, g) & (kTileN - 1)))];
        st128(o + e, x);
    }
}

template <int LOGN>
static void launch_fused_t(const DevTables& T, const FusedKsArgs& a, const BaseConvTable* moddown, int nb, cudaStream_t st) {
    using G = NttGeo<LOGN>;
    k_ks_p_inner<LOGN><<<dim3(G::TILES, 2 * a.K, nb), kTileThreads, 0, st>>>(T, a);
    const size_t smem = (size_t)(a.K + 1) * kTileWords * sizeof(u64);
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k_conv_cols<LOGN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 9 * kTileWords * (int)sizeof(u64));
        attr_set = true;
    }
    k_conv_cols<LOGN><<<dim3(G::TILES, nb * 2), kTileThreads, smem, st>>>(T, moddown, a.tmpP, (size_t)a.K * T.N, a.conv, (size_t)a.nl * T.N);
    k_ks_q_finish<LOGN><<<dim3(G::TILES, 2 * a.nl, nb), kTileThreads, 0, st>>>(T, a);
}

bool fused_ks_supported(int logN) { return logN >= 12 && logN <= 16; }

void launch_fused_ks(const DevTables& T, const FusedKsArgs& a, const BaseConvTable* moddown, int nb, cudaStream_t st) {
    switch (T.logN) {
        case 12: launch_fused_t<12>(T, a, moddown, nb, st); break;
        case 13: launch_fused_t<13>(T, a, moddown, nb, st); break;
        case 14: launch_fused_t<14>(T, a, moddown, nb, st); break;
        case 15: launch_fused_t<15>(T, a, moddown, nb, st); break;
        case 16: launch_fused_t<16>(T, a, moddown, nb, st); break;
        default: break;
    }
}

}  // namespace hx
