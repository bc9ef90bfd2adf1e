// hx_kernels.cu -- sm_100a kernels for the CKKS hot path other than the NTT (SURVEY §8 rows
// P3-P8): approximate base conversion, key-switch inner product, ModDown epilogue with the Galois
// gather, rescale, the elementwise family and the fused diagonal multiply-accumulate.
// Every kernel is HBM/L2-streaming integer work: one thread owns two adjacent coefficients of a
// limb (128-bit loads/stores), grid.y walks limbs, grid.z walks batch entries.
#include "hx_kernels.cuh"
#include "hx_fpmod.cuh"

namespace hx {

constexpr int kThreads = 256;
static inline unsigned blocks_for(int N) { return (unsigned)((N / 2 + kThreads - 1) / kThreads); }

// ---------------------------------------------------------------------------------------------
// ApproxSwitchCRTBasis (row P4/P5): out_j = ( sum_i [x_i * hatinv_i]_{q_i} * (Qhat_i mod p_j) ) mod p_j
// hatinv as Shoup pairs; the matrix in Montgomery form so one REDC finishes each output.
// Sum of n_in <= 8 products, each < 2^60 * p_j  ->  < p_j * 2^64, REDC precondition holds.
__global__ void __launch_bounds__(kThreads) k_baseconv(DevTables T, const BaseConvTable* __restrict__ tab, int n_in, int n_out,
                                                       const u64* __restrict__ in, size_t in_bs, u64* __restrict__ out, size_t out_bs) {
    __shared__ u64 s_mat[8 * kMaxPrimes];
    __shared__ u64 s_q[kMaxPrimes], s_qi[kMaxPrimes];
    for (int i = threadIdx.x; i < n_in * n_out; i += blockDim.x) s_mat[i] = tab->mat[i];
    for (int j = threadIdx.x; j < n_out; j += blockDim.x) {
        PrimeConst pc = T.pc[tab->dst[j]];
        s_q[j] = pc.q;
        s_qi[j] = pc.qinv_neg;
    }
    __syncthreads();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const u64* src = in + (size_t)blockIdx.z * in_bs + idx;
    u64* dst = out + (size_t)blockIdx.z * out_bs + idx;
    u64 y0[8], y1[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        if (i < n_in) {
            ulonglong2 x = ldg128(src + (size_t)i * T.N);
            u64 q = T.pc[tab->src[i]].q;
            y0[i] = shoup_mul(x.x, tab->hatinv[i], tab->hatinv_sh[i], q);
            y1[i] = shoup_mul(x.y, tab->hatinv[i], tab->hatinv_sh[i], q);
        }
    }
    for (int j = 0; j < n_out; j++) {
        u128d a0 = {0, 0}, a1 = {0, 0};
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (i < n_in) {
                u64 m = s_mat[i * n_out + j];
                acc_wide(a0, y0[i], m);
                acc_wide(a1, y1[i], m);
            }
        }
        ulonglong2 r;
        r.x = redc(a0, s_q[j], s_qi[j]);
        r.y = redc(a1, s_q[j], s_qi[j]);
        st128(dst + (size_t)j * T.N, r);
    }
}

void launch_baseconv(const DevTables& T, const BaseConvTable* tab, int n_in, int n_out, const u64* in, size_t in_bs,
                     u64* out, size_t out_bs, int nbatch, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), 1, nbatch);
    k_baseconv<<<grid, kThreads, 0, st>>>(T, tab, n_in, n_out, in, in_bs, out, out_bs);
    check_launch("k_baseconv");
}

// ---------------------------------------------------------------------------------------------
// EvalFastKeySwitchCoreExt (row P5): c~0 = sum_j d_j (.) b_j ; c~1 = sum_j d_j (.) a_j over the
// nl Q-limbs and K P-limbs.  Keys are streamed once from HBM (L1 no-allocate), digits are re-read
// by every batch entry (hoisted rotations share them) and stay L2-resident.  Keys are stored in
// Montgomery form, so each output is 3 wide MACs + one REDC (3*q^2 < q*2^64).
__global__ void __launch_bounds__(kThreads) k_ks_inner(DevTables T, const u64* __restrict__ digits, int ndig, int nl, int L, int K,
                                                       const u64* const* keyA, const u64* const* keyB,
                                                       u64* __restrict__ out, size_t out_bs) {
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;                        // extended limb index 0..nl+K-1
    int kl = i < nl ? i : i - nl + L;          // limb index inside the key / prime index
    const PrimeConst pc = T.pc[kl];
    const int ext = nl + K;
    const u64* ka = reinterpret_cast<const u64*>(ldg64(reinterpret_cast<const u64*>(keyA + blockIdx.z)));   // pointer tables are pool memory
    const u64* kb = reinterpret_cast<const u64*>(ldg64(reinterpret_cast<const u64*>(keyB + blockIdx.z)));
    u128d b0 = {0, 0}, b1 = {0, 0}, a0 = {0, 0}, a1 = {0, 0};
    const bool fpkey = kl < L && (pc.q >> 46) == 0;   // engine-internal key form (launch_key_finish): doubles of plain residues on these limbs
    for (int j = 0; j < ndig; j++) {
        ulonglong2 d = ldg128(digits + ((size_t)j * ext + i) * T.N + idx);
        size_t koff = ((size_t)j * (L + K) + kl) * T.N + idx;
        ulonglong2 vb = ld_stream128(kb + koff);
        ulonglong2 va = ld_stream128(ka + koff);
        if (fpkey) {
            vb.x = fp_to_u64(fp_bits_to_double(vb.x)); vb.y = fp_to_u64(fp_bits_to_double(vb.y));
            va.x = fp_to_u64(fp_bits_to_double(va.x)); va.y = fp_to_u64(fp_bits_to_double(va.y));
        } else {
            vb.x = unsplit30(vb.x); vb.y = unsplit30(vb.y); va.x = unsplit30(va.x); va.y = unsplit30(va.y);   // Montgomery residues, stored split
        }
        acc_wide(b0, d.x, vb.x);
        acc_wide(b1, d.y, vb.y);
        acc_wide(a0, d.x, va.x);
        acc_wide(a1, d.y, va.y);
    }
    u64* o = out + (size_t)blockIdx.z * out_bs + (size_t)i * T.N + idx;
    ulonglong2 r;
    r.x = redc(b0, pc.q, pc.qinv_neg);
    r.y = redc(b1, pc.q, pc.qinv_neg);
    if (fpkey) { r.x = redc(mul_wide(r.x, pc.r2), pc.q, pc.qinv_neg); r.y = redc(mul_wide(r.y, pc.r2), pc.q, pc.qinv_neg); }   // plain key: undo the REDC's 2^-64
    st128(o, r);
    r.x = redc(a0, pc.q, pc.qinv_neg);
    r.y = redc(a1, pc.q, pc.qinv_neg);
    if (fpkey) { r.x = redc(mul_wide(r.x, pc.r2), pc.q, pc.qinv_neg); r.y = redc(mul_wide(r.y, pc.r2), pc.q, pc.qinv_neg); }
    st128(o + (size_t)ext * T.N, r);
}

void launch_ks_inner(const DevTables& T, const u64* digits, int ndig, int nl, int L, int K, const u64* const* keyA,
                     const u64* const* keyB, u64* out, size_t out_bs, int nbatch, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl + K, nbatch);
    k_ks_inner<<<grid, kThreads, 0, st>>>(T, digits, ndig, nl, L, K, keyA, keyB, out, out_bs);
    check_launch("k_ks_inner");
}

// ---------------------------------------------------------------------------------------------
// ApproxModDown epilogue (row P5) fused with "+ c0" and the automorphism gather (row P3):
//   out[i][j] = ct'[i][m] - conv'[i][m] + addend[i][m],  m = map[j]   (keys and conversion matrix carry P^-1)
// grid.y = limb (nl), grid.z = batch*2 + component; addend ([comps][nl][N]) applies to components < addend_comps.
__global__ void __launch_bounds__(kThreads) k_moddown_finish(DevTables T, const u64* ct, size_t ct_bs, const u64* conv,
                                                             size_t conv_bs, const u64* addend, int addend_comps, const u32* const* maps,
                                                             u64* __restrict__ out, size_t out_bs, int nl, int ext) {
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    int batch = blockIdx.z >> 1, comp = blockIdx.z & 1;
    const u64 q = T.pc[i].q;
    const u64* c = ct + (size_t)batch * ct_bs + ((size_t)comp * ext + i) * T.N;
    const u64* v = conv + (size_t)batch * conv_bs + ((size_t)comp * nl + i) * T.N;
    const u64* ad = (addend != nullptr && comp < addend_comps) ? addend + ((size_t)comp * nl + i) * T.N : nullptr;
    const u32* map = maps ? reinterpret_cast<const u32*>(ldg64(reinterpret_cast<const u64*>(maps + batch))) : nullptr;
    u64 r[2];
#pragma unroll
    for (int e = 0; e < 2; e++) {
        int m = map ? (int)map[idx + e] : idx + e;
        u64 d = sub_mod(ldg64(c + m), ldg64(v + m), q);
        if (ad) d = add_mod(d, ldg64(ad + m), q);
        r[e] = d;
    }
    ulonglong2 o = {r[0], r[1]};
    st128(out + (size_t)batch * out_bs + ((size_t)comp * nl + i) * T.N + idx, o);
}

void launch_moddown_finish(const DevTables& T, const u64* ct, size_t ct_bs, const u64* conv, size_t conv_bs, const u64* addend,
                           int addend_comps, const u32* const* maps, u64* out, size_t out_bs, int nl, int nbatch,
                           cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, nbatch * 2);
    int ext = (int)(ct_bs / (2 * (size_t)T.N));   // ct batch entry = [2][ext][N]
    k_moddown_finish<<<grid, kThreads, 0, st>>>(T, ct, ct_bs, conv, conv_bs, addend, addend_comps, maps, out, out_bs, nl, ext);
    check_launch("k_moddown_finish");
}

// ---------------------------------------------------------------------------------------------
// Elementwise family (row P8). grid.y = limb, grid.z = component.
__global__ void __launch_bounds__(kThreads) k_ew(DevTables T, int op, u64* __restrict__ out, const u64* __restrict__ a, const u64* __restrict__ b,
                                                 size_t cs_out, size_t cs_a, size_t cs_b) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const PrimeConst pc = T.pc[i];
    size_t off = (size_t)i * T.N + idx;
    ulonglong2 x = ldg128(a + blockIdx.z * cs_a + off);
    ulonglong2 y = ldg128(b + blockIdx.z * cs_b + off);
    ulonglong2 r;
    if (op == EW_ADD) { r.x = add_mod(x.x, y.x, pc.q); r.y = add_mod(x.y, y.y, pc.q); }
    else if (op == EW_SUB) { r.x = sub_mod(x.x, y.x, pc.q); r.y = sub_mod(x.y, y.y, pc.q); }
    else { r.x = mulmod(x.x, y.x, pc); r.y = mulmod(x.y, y.y, pc); }
    st128(out + blockIdx.z * cs_out + off, r);
}
void launch_ew(const DevTables& T, EwOp op, u64* out, const u64* a, const u64* b, int nl, int ncomp, size_t cs_out, size_t cs_a,
               size_t cs_b, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, ncomp);
    launch_pdl(k_ew, grid, dim3(kThreads), 0, st, T, (int)op, out, a, b, cs_out, cs_a, cs_b);
}

__global__ void __launch_bounds__(kThreads) k_mul_scalar(DevTables T, u64* __restrict__ out, const u64* __restrict__ a, ScalarTable s,
                                                         size_t cs_out, size_t cs_a) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const u64 q = T.pc[i].q;
    size_t off = (size_t)i * T.N + idx;
    ulonglong2 x = ldg128(a + blockIdx.z * cs_a + off);
    x.x = shoup_mul(x.x, s.v[i], s.sh[i], q);
    x.y = shoup_mul(x.y, s.v[i], s.sh[i], q);
    st128(out + blockIdx.z * cs_out + off, x);
}
void launch_mul_scalar(const DevTables& T, u64* out, const u64* a, const ScalarTable& s, int nl, int ncomp, size_t cs_out, size_t cs_a,
                       cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, ncomp);
    launch_pdl(k_mul_scalar, grid, dim3(kThreads), 0, st, T, out, a, s, cs_out, cs_a);
}

__global__ void __launch_bounds__(kThreads) k_add_scalar(DevTables T, u64* __restrict__ out, const u64* __restrict__ a, ScalarTable s, int ncomp) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const u64 q = T.pc[i].q;
    size_t off = ((size_t)blockIdx.z * gridDim.y + i) * T.N + idx;
    ulonglong2 x = ldg128(a + off);
    if (blockIdx.z % ncomp == 0) {   // the constant goes to component 0; the other components are copied
        x.x = add_mod(x.x, s.v[i], q);
        x.y = add_mod(x.y, s.v[i], q);
    }
    st128(out + off, x);
}
void launch_add_scalar(const DevTables& T, u64* out, const u64* a, const ScalarTable& s, int nl, int ncomp, int ncomps_total, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, ncomps_total);
    launch_pdl(k_add_scalar, grid, dim3(kThreads), 0, st, T, out, a, s, ncomp);
}

__global__ void __launch_bounds__(kThreads) k_wsum(DevTables T, u64* __restrict__ out, const __grid_constant__ WsumArgs a) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const int i = blockIdx.y;
    const u64 q = T.pc[i].q;
    const size_t off = ((size_t)blockIdx.z * gridDim.y + i) * T.N + idx;
    ulonglong2 acc = ldg128(a.p[0] + off);
    acc.x = shoup_mul(acc.x, a.v[0][i], a.sh[0][i], q);
    acc.y = shoup_mul(acc.y, a.v[0][i], a.sh[0][i], q);
    for (int t = 1; t < a.n; t++) {
        const ulonglong2 x = ldg128(a.p[t] + off);
        acc.x = add_mod(acc.x, shoup_mul(x.x, a.v[t][i], a.sh[t][i], q), q);
        acc.y = add_mod(acc.y, shoup_mul(x.y, a.v[t][i], a.sh[t][i], q), q);
    }
    st128(out + off, acc);
}
void launch_wsum(const DevTables& T, u64* out, const WsumArgs& a, int nl, int ncomps_total, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, ncomps_total);
    launch_pdl(k_wsum, grid, dim3(kThreads), 0, st, T, out, a);
}

// EvalMultCore (row P6, single product): out = (a0*b0, a0*b1 + a1*b0, a1*b1).
__global__ void __launch_bounds__(kThreads) k_tensor(DevTables T, u64* __restrict__ out, const u64* __restrict__ a, size_t cs_a,
                                                     const u64* __restrict__ b, size_t cs_b, int nl, int bca, int bcb) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const PrimeConst pc = T.pc[i];
    size_t off = (size_t)i * T.N + idx;
    if (!bca) a += (size_t)blockIdx.z * 2 * cs_a;
    if (!bcb) b += (size_t)blockIdx.z * 2 * cs_b;
    out += (size_t)blockIdx.z * 3 * nl * T.N;   // batch entries are contiguous ciphertexts
    ulonglong2 a0 = ldg128(a + off);
    ulonglong2 a1 = ldg128(a + cs_a + off);
    ulonglong2 b0 = ldg128(b + off);
    ulonglong2 b1 = ldg128(b + cs_b + off);
    size_t cs_o = (size_t)nl * T.N;
    ulonglong2 r;
    r.x = mulmod(a0.x, b0.x, pc); r.y = mulmod(a0.y, b0.y, pc);
    st128(out + off, r);
    r.x = add_mod(mulmod(a0.x, b1.x, pc), mulmod(a1.x, b0.x, pc), pc.q);
    r.y = add_mod(mulmod(a0.y, b1.y, pc), mulmod(a1.y, b0.y, pc), pc.q);
    st128(out + cs_o + off, r);
    r.x = mulmod(a1.x, b1.x, pc); r.y = mulmod(a1.y, b1.y, pc);
    st128(out + 2 * cs_o + off, r);
}
void launch_tensor(const DevTables& T, u64* out, const u64* a, size_t cs_a, bool bcast_a, const u64* b, size_t cs_b, bool bcast_b, int nl, int nbatch, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, nbatch);
    launch_pdl(k_tensor, grid, dim3(kThreads), 0, st, T, out, a, cs_a, b, cs_b, nl, bcast_a ? 1 : 0, bcast_b ? 1 : 0);
}

// EvalMultCore against database ciphertexts kept in Montgomery form: one REDC per output word.
__global__ void __launch_bounds__(kThreads) k_tensor_db(DevTables T, u64* __restrict__ out, const __grid_constant__ CtPtrs ap, size_t cs_a,
                                                        const u64* __restrict__ db, size_t cs_b, int nl) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const PrimeConst pc = T.pc[i];
    size_t off = (size_t)i * T.N + idx;
    const u64* a = ap.p[blockIdx.z];
    const u64* b = db + (size_t)blockIdx.z * 2 * cs_b;
    out += (size_t)blockIdx.z * 3 * nl * T.N;
    const ulonglong2 a0 = ldg128(a + off), a1 = ldg128(a + cs_a + off);
    const ulonglong2 b0 = ld_stream128(b + off), b1 = ld_stream128(b + cs_b + off);
    size_t cs_o = (size_t)nl * T.N;
    ulonglong2 r;
    r.x = mont_mul(a0.x, b0.x, pc.q, pc.qinv_neg); r.y = mont_mul(a0.y, b0.y, pc.q, pc.qinv_neg);
    st128(out + off, r);
    u128d m0 = mul_wide(a0.x, b1.x), m1 = mul_wide(a0.y, b1.y);     // 2 q^2 < q 2^64: one REDC for the cross term
    acc_wide(m0, a1.x, b0.x); acc_wide(m1, a1.y, b0.y);
    r.x = redc(m0, pc.q, pc.qinv_neg); r.y = redc(m1, pc.q, pc.qinv_neg);
    st128(out + cs_o + off, r);
    r.x = mont_mul(a1.x, b1.x, pc.q, pc.qinv_neg); r.y = mont_mul(a1.y, b1.y, pc.q, pc.qinv_neg);
    st128(out + 2 * cs_o + off, r);
}
void launch_tensor_db(const DevTables& T, u64* out, const CtPtrs& a, size_t cs_a, const u64* db, size_t cs_b, int nl, int nbatch, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl, nbatch);
    launch_pdl(k_tensor_db, grid, dim3(kThreads), 0, st, T, out, a, cs_a, db, cs_b, nl);
}
__global__ void __launch_bounds__(kThreads) k_batch_sum(DevTables T, u64* __restrict__ acc, const u64* __restrict__ in, int nl, size_t bs, int nbatch,
                                                        int accumulate) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const u64 q = T.pc[blockIdx.y % nl].q;
    size_t off = (size_t)blockIdx.y * T.N + idx;
    ulonglong2 s = accumulate ? ldg128(acc + off) : make_ulonglong2(0, 0);
    for (int z = 0; z < nbatch; z++) {
        const ulonglong2 x = ldg128(in + (size_t)z * bs + off);
        s.x = add_mod(s.x, x.x, q);
        s.y = add_mod(s.y, x.y, q);
    }
    st128(acc + off, s);
}
void launch_batch_sum(const DevTables& T, u64* acc, const u64* in, int nl, int ncomp, int nbatch, int accumulate, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl * ncomp, 1);
    launch_pdl(k_batch_sum, grid, dim3(kThreads), 0, st, T, acc, in, nl, (size_t)ncomp * nl * T.N, nbatch, accumulate);
}

// multi-GPU combine (sharded.py): the ranks' partial sums are added as plain 64-bit integers by the collective; this brings
// them back to residues
__global__ void __launch_bounds__(kThreads) k_fold_mod(DevTables T, u64* __restrict__ data, int nl) {
    pdl_wait();
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const PrimeConst pc = T.pc[blockIdx.y % nl];
    u64* p = data + (size_t)blockIdx.y * T.N + idx;
    ulonglong2 x = ldg128(p);
    x.x = reduce64(x.x, pc.q, pc.mu64);
    x.y = reduce64(x.y, pc.q, pc.mu64);
    st128(p, x);
}
void launch_fold_mod(const DevTables& T, u64* data, int nl, int ncomps, cudaStream_t st) {
    launch_pdl(k_fold_mod, dim3(blocks_for(T.N), nl * ncomps, 1), dim3(kThreads), 0, st, T, data, nl);
}
// AutomorphismTransform (row P3): out[l][j] = in[l][map[j]].
__global__ void __launch_bounds__(kThreads) k_automorph(int N, u64* __restrict__ out, const u64* __restrict__ in, const u32* __restrict__ map) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N) return;
    size_t off = (size_t)blockIdx.y * N;
    out[off + idx] = ldg64(in + off + map[idx]);
}
void launch_automorph(const DevTables& T, u64* out, const u64* in, const u32* map, int nlimbs_total, cudaStream_t st) {
    dim3 grid((T.N + kThreads - 1) / kThreads, nlimbs_total, 1);
    k_automorph<<<grid, kThreads, 0, st>>>(T.N, out, in, map);
    check_launch("k_automorph");
}

// PrecomputeAutoMap: map[brev(j)] = brev(((2j+1)*k mod 2N) >> 1).
__global__ void k_automap(int N, int logN, u64 k, u32* __restrict__ map) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    u64 m = (u64)N << 1;
    u64 idx = ((((u64)j << 1) + 1) * k & (m - 1)) >> 1;
    u32 jr = __brev((u32)j) >> (32 - logN);
    u32 ir = __brev((u32)idx) >> (32 - logN);
    map[jr] = ir;
}
void launch_automap(int N, int logN, u64 k, u32* map, cudaStream_t st) {
    k_automap<<<(N + 255) / 256, 256, 0, st>>>(N, logN, k, map);
    check_launch("k_automap");
}

// x -> x * 2^64 mod q (Montgomery form) for keys / database blocks at load time.
__global__ void __launch_bounds__(kThreads) k_to_mont(DevTables T, u64* __restrict__ data, LimbMap lm, bool inverse) {
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    size_t b = (size_t)blockIdx.y + (size_t)blockIdx.z * gridDim.y;
    const PrimeConst pc = T.pc[lm.prime[b % lm.period]];
    u64* p = data + b * T.N + idx;
    ulonglong2 x = ldg128(p);
    const u64 f = inverse ? 1ull : pc.r2;          // inverse: x * 2^-64, back to plain residues
    x.x = redc(mul_wide(x.x, f), pc.q, pc.qinv_neg);
    x.y = redc(mul_wide(x.y, f), pc.q, pc.qinv_neg);
    st128(p, x);
}
__global__ void __launch_bounds__(kThreads) k_split30(u64* __restrict__ data, size_t words) {
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i >= words) return;
    ulonglong2 x = ldg128(data + i);
    x.x = split30(x.x); x.y = split30(x.y);
    st128(data + i, x);
}
// grid (N / 512, L + K, dnum)
__global__ void __launch_bounds__(kThreads) k_key_finish(DevTables T, u64* __restrict__ key, int L, int K, ScalarTable pinv) {
    const int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const int limb = blockIdx.y;
    const PrimeConst pc = T.pc[limb];
    u64* p = key + ((size_t)blockIdx.z * (L + K) + limb) * T.N + idx;
    ulonglong2 x = ldg128(p);
    if (limb < L) {
        x.x = shoup_mul(x.x, pinv.v[limb], pinv.sh[limb], pc.q);
        x.y = shoup_mul(x.y, pinv.v[limb], pinv.sh[limb], pc.q);
    }
    if (limb < L && (pc.q >> 46) == 0) {
        x.x = fp_double_to_bits(fp_from_u64(x.x));
        x.y = fp_double_to_bits(fp_from_u64(x.y));
    } else {
        x.x = split30(redc(mul_wide(x.x, pc.r2), pc.q, pc.qinv_neg));
        x.y = split30(redc(mul_wide(x.y, pc.r2), pc.q, pc.qinv_neg));
    }
    st128(p, x);
}
void launch_key_finish(const DevTables& T, u64* key, int dnum, int L, int K, const ScalarTable& pinv, cudaStream_t st) {
    k_key_finish<<<dim3(blocks_for(T.N), L + K, dnum), kThreads, 0, st>>>(T, key, L, K, pinv);
    check_launch("k_key_finish");
}
__global__ void __launch_bounds__(kThreads) k_key_unfinish(DevTables T, u64* __restrict__ out, const u64* __restrict__ key, int L, int K, ScalarTable pmod) {
    const int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const int limb = blockIdx.y;
    const PrimeConst pc = T.pc[limb];
    const size_t off = ((size_t)blockIdx.z * (L + K) + limb) * T.N + idx;
    ulonglong2 x = ld_stream128(key + off);
    if (limb < L && (pc.q >> 46) == 0) {
        x.x = fp_to_u64(fp_bits_to_double(x.x));
        x.y = fp_to_u64(fp_bits_to_double(x.y));
    } else {
        u128d a = {unsplit30(x.x), 0}, b = {unsplit30(x.y), 0};
        x.x = redc(a, pc.q, pc.qinv_neg);      // out of Montgomery form
        x.y = redc(b, pc.q, pc.qinv_neg);
    }
    if (limb < L) {
        x.x = shoup_mul(x.x, pmod.v[limb], pmod.sh[limb], pc.q);
        x.y = shoup_mul(x.y, pmod.v[limb], pmod.sh[limb], pc.q);
    }
    st128(out + off, x);
}
void launch_key_unfinish(const DevTables& T, u64* out, const u64* key, int dnum, int L, int K, const ScalarTable& pmod, cudaStream_t st) {
    k_key_unfinish<<<dim3(blocks_for(T.N), L + K, dnum), kThreads, 0, st>>>(T, out, key, L, K, pmod);
    check_launch("k_key_unfinish");
}
void launch_split30(u64* data, size_t words, cudaStream_t st) {
    k_split30<<<(unsigned)((words / 2 + kThreads - 1) / kThreads), kThreads, 0, st>>>(data, words);
    check_launch("k_split30");
}
void launch_to_mont(const DevTables& T, u64* data, int nbatch, const LimbMap& lm, cudaStream_t st, bool inverse) {
    // nbatch must be a multiple of lm.period; split so grid.y stays a multiple of the period
    int per = lm.period;
    int groups = nbatch / per;
    dim3 grid(blocks_for(T.N), per, groups);
    k_to_mont<<<grid, kThreads, 0, st>>>(T, data, lm, inverse);
    check_launch("k_to_mont");
}

// ---------------------------------------------------------------------------------------------
// DropLastElementAndScale (row P7).
// prep: tmp_i = centred_lift(last) * (-q_l^-1) mod q_i.  The centred lift of v > q_l/2 is v - q_l,
// and q_l * (-q_l^-1) = -1, so the correction is simply "+1".
__global__ void __launch_bounds__(kThreads) k_rescale_prep(DevTables T, const u64* __restrict__ last, size_t last_bs, u64* __restrict__ tmp,
                                                           size_t tmp_bs, int nl, ScalarTable c) {
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const u64 q = T.pc[i].q;
    const u64 half = T.pc[nl - 1].q >> 1;
    ulonglong2 v = ldg128(last + blockIdx.z * last_bs + idx);
    ulonglong2 r;
    r.x = shoup_mul(v.x, c.v[i], c.sh[i], q);
    r.y = shoup_mul(v.y, c.v[i], c.sh[i], q);
    if (v.x > half) r.x = add_mod(r.x, 1, q);
    if (v.y > half) r.y = add_mod(r.y, 1, q);
    st128(tmp + blockIdx.z * tmp_bs + (size_t)i * T.N + idx, r);
}
void launch_rescale_prep(const DevTables& T, const u64* last, size_t last_bs, u64* tmp, size_t tmp_bs, int nl,
                         const ScalarTable& negqlinv, int nbatch, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl - 1, nbatch);
    k_rescale_prep<<<grid, kThreads, 0, st>>>(T, last, last_bs, tmp, tmp_bs, nl, negqlinv);
    check_launch("k_rescale_prep");
}
__global__ void __launch_bounds__(kThreads) k_rescale_finish(DevTables T, u64* __restrict__ out, size_t out_bs, const u64* __restrict__ c, size_t c_bs,
                                                             const u64* __restrict__ tmp, size_t tmp_bs, ScalarTable s) {
    int idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const u64 q = T.pc[i].q;
    size_t off = (size_t)i * T.N + idx;
    ulonglong2 x = ldg128(c + blockIdx.z * c_bs + off);
    ulonglong2 t = ldg128(tmp + blockIdx.z * tmp_bs + off);
    x.x = add_mod(shoup_mul(x.x, s.v[i], s.sh[i], q), t.x, q);
    x.y = add_mod(shoup_mul(x.y, s.v[i], s.sh[i], q), t.y, q);
    st128(out + blockIdx.z * out_bs + off, x);
}
void launch_rescale_finish(const DevTables& T, u64* out, size_t out_bs, const u64* c, size_t c_bs, const u64* tmp, size_t tmp_bs,
                           int nl, const ScalarTable& qlinv, int nbatch, cudaStream_t st) {
    dim3 grid(blocks_for(T.N), nl - 1, nbatch);
    k_rescale_finish<<<grid, kThreads, 0, st>>>(T, out, out_bs, c, c_bs, tmp, tmp_bs, qlinv);
    check_launch("k_rescale_finish");
}

// ---------------------------------------------------------------------------------------------
// Fused diagonal multiply-accumulate (row P6): sum over diagonals d of rotq[d] (x) db[d]
// (sender_diag.cpp:70-77: 512 EvalMultNoRelin + 511 EvalAddInPlace).  The database block is in
// Montgomery form; products are accumulated un-reduced in 128 bits and folded every kFold diagonals
// (2*kFold*q^2 < 2^128 for q < 2^60), so the kernel does 4 wide MACs per 32 bytes streamed.
constexpr int kFold = 64;
// For the 45-bit limbs (q < 2^46, all but q_0) the operands are cut at 23 bits: the four partial products of a pair are below
// 2^46, so three 64-bit columns take all 512 diagonals (and both products of component 1) without a carry and without a fold:
// 4 mad.wide per product and nothing else, against ~16 instructions for a 64x64->128 multiply-accumulate.
constexpr int kCut = 23;
constexpr int kMaxDiagSmall = 1 << 14;      // columns stay below 2^63: 2 * 2^46 * 2^14 = 2^61
struct cols3 { u64 lo, mid, hi; };
__device__ __forceinline__ void mac23(cols3& c, u32 x0, u32 x1, u32 y0, u32 y1) {
    c.lo = madw(x0, y0, c.lo);
    c.mid = madw(x0, y1, c.mid);
    c.mid = madw(x1, y0, c.mid);
    c.hi = madw(x1, y1, c.hi);
}
__device__ __forceinline__ u128d fold23(const cols3& c) {      // lo + mid * 2^23 + hi * 2^46 as a 128-bit integer
    u128d t;
    const u64 m_lo = c.mid << kCut, m_hi = c.mid >> (64 - kCut), h_lo = c.hi << (2 * kCut), h_hi = c.hi >> (64 - 2 * kCut);
    asm("add.cc.u64 %0, %2, %3;\n\taddc.u64 %1, %4, %5;" : "=l"(t.lo), "=l"(t.hi) : "l"(c.lo), "l"(m_lo), "l"(m_hi), "l"(h_hi));
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(t.lo), "+l"(t.hi) : "l"(h_lo));
    return t;
}
// NB database blocks per CTA: the rotated-query operands are loaded and cut once for NB blocks.
template <int NB>
__global__ void __launch_bounds__(kThreads, 2) k_diag_mac(DevTables T, const u64* __restrict__ rotq, const __grid_constant__ DbBlockPtrs dbs,
                                                          u64* __restrict__ out, int ndiag, int nl, int nblocks) {
    pdl_wait();
    int idx = (blockIdx.y * blockDim.x + threadIdx.x) * 2;
    if (idx >= T.N) return;
    const int i = blockIdx.z, g0 = blockIdx.x * NB;
    const PrimeConst pc = T.pc[i];
    const size_t cs = (size_t)nl * T.N, ds = 2 * cs;
    const size_t off = (size_t)i * T.N + idx;
    if ((pc.q >> 46) == 0 && ndiag <= kMaxDiagSmall) {      // uniform per CTA (one limb per blockIdx.z)
        constexpr u32 kLow = (1u << kCut) - 1;
        cols3 acc[NB][3][2] = {};
        const u64* __restrict__ db[NB];
#pragma unroll
        for (int g = 0; g < NB; g++) db[g] = dbs.p[min(g0 + g, nblocks - 1)] + off;
#pragma unroll 2
        for (int d = 0; d < ndiag; d++) {
            const u64* qa = rotq + (size_t)d * ds + off;
            ulonglong2 a0 = ldg128(qa), a1 = ldg128(qa + cs);
            ulonglong2 b0[NB], b1[NB];
#pragma unroll
            for (int g = 0; g < NB; g++) {
                b0[g] = ld_stream128(db[g] + (size_t)d * ds);
                b1[g] = ld_stream128(db[g] + (size_t)d * ds + cs);
            }
            const u32 a0x0 = (u32)a0.x & kLow, a0x1 = (u32)(a0.x >> kCut), a0y0 = (u32)a0.y & kLow, a0y1 = (u32)(a0.y >> kCut);
            const u32 a1x0 = (u32)a1.x & kLow, a1x1 = (u32)(a1.x >> kCut), a1y0 = (u32)a1.y & kLow, a1y1 = (u32)(a1.y >> kCut);
#pragma unroll
            for (int g = 0; g < NB; g++) {
                const u32 b0x0 = (u32)b0[g].x & kLow, b0x1 = (u32)(b0[g].x >> kCut), b0y0 = (u32)b0[g].y & kLow, b0y1 = (u32)(b0[g].y >> kCut);
                const u32 b1x0 = (u32)b1[g].x & kLow, b1x1 = (u32)(b1[g].x >> kCut), b1y0 = (u32)b1[g].y & kLow, b1y1 = (u32)(b1[g].y >> kCut);
                mac23(acc[g][0][0], a0x0, a0x1, b0x0, b0x1); mac23(acc[g][0][1], a0y0, a0y1, b0y0, b0y1);
                mac23(acc[g][1][0], a0x0, a0x1, b1x0, b1x1); mac23(acc[g][1][1], a0y0, a0y1, b1y0, b1y1);
                mac23(acc[g][1][0], a1x0, a1x1, b0x0, b0x1); mac23(acc[g][1][1], a1y0, a1y1, b0y0, b0y1);
                mac23(acc[g][2][0], a1x0, a1x1, b1x0, b1x1); mac23(acc[g][2][1], a1y0, a1y1, b1y0, b1y1);
            }
        }
#pragma unroll
        for (int g = 0; g < NB; g++) {
            if (g0 + g >= nblocks) break;
            u64* o = out + (size_t)(g0 + g) * 3 * cs + off;
#pragma unroll
            for (int c = 0; c < 3; c++) {
                ulonglong2 v;
                v.x = redc(fold128(fold23(acc[g][c][0]), pc.q, pc.mu64), pc.q, pc.qinv_neg);
                v.y = redc(fold128(fold23(acc[g][c][1]), pc.q, pc.mu64), pc.q, pc.qinv_neg);
                st128(o + (size_t)c * cs, v);
            }
        }
        return;
    }
    for (int g = g0; g < min(g0 + NB, nblocks); g++) {
        const u64* __restrict__ db = dbs.p[g];
        u64 r[3][2] = {{0, 0}, {0, 0}, {0, 0}};
        for (int d0 = 0; d0 < ndiag; d0 += kFold) {
            u128d acc[3][2] = {};
            int d1 = min(ndiag, d0 + kFold);
#pragma unroll 2
            for (int d = d0; d < d1; d++) {
                const u64* qa = rotq + (size_t)d * ds + off;
                const u64* da = db + (size_t)d * ds + off;
                ulonglong2 a0 = ldg128(qa), a1 = ldg128(qa + cs);
                ulonglong2 b0 = ld_stream128(da), b1 = ld_stream128(da + cs);
                acc_wide(acc[0][0], a0.x, b0.x); acc_wide(acc[0][1], a0.y, b0.y);
                acc_wide(acc[1][0], a0.x, b1.x); acc_wide(acc[1][1], a0.y, b1.y);
                acc_wide(acc[1][0], a1.x, b0.x); acc_wide(acc[1][1], a1.y, b0.y);
                acc_wide(acc[2][0], a1.x, b1.x); acc_wide(acc[2][1], a1.y, b1.y);
            }
#pragma unroll
            for (int c = 0; c < 3; c++)
#pragma unroll
                for (int e = 0; e < 2; e++)
                    r[c][e] = add_mod(r[c][e], redc(fold128(acc[c][e], pc.q, pc.mu64), pc.q, pc.qinv_neg), pc.q);
        }
        u64* o = out + (size_t)g * 3 * cs + off;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            ulonglong2 v = {r[c][0], r[c][1]};
            st128(o + (size_t)c * cs, v);
        }
    }
}
// The same sum with the operands fed by TMA (the kernel the engine uses): a producer thread keeps kMacStages diagonals (4 tiles of
// 4 KB: both components of the rotated query and of the database ciphertext) in flight per CTA on full/empty mbarriers, the 8
// consumer warps read them from shared memory. Loads in flight no longer depend on registers or on where the warps are in their
// multiply-accumulate code. Measured against the register-fed kernel above: 1 block 0.99 -> 0.93 ms (6.9 TB/s of DRAM reads),
// 8 blocks 5.9 -> 5.0 ms per launch.
constexpr int kMacStages = 6;
constexpr int kMacTileBytes = kThreads * 16;
constexpr int kMacSmemBytes = kMacStages * 4 * kMacTileBytes + 2 * kMacStages * 8;
__global__ void __launch_bounds__(kThreads + 32, 2) k_diag_mac_tma(DevTables T, const u64* __restrict__ rotq, const __grid_constant__ DbBlockPtrs dbs,
                                                                   u64* __restrict__ out, int ndiag, int nl, int first_limb) {
    pdl_wait();
    extern __shared__ __align__(128) unsigned char smraw[];
    u64* full = reinterpret_cast<u64*>(smraw + kMacStages * 4 * kMacTileBytes);
    u64* empty = full + kMacStages;
    const int t = threadIdx.x;
    const int i = first_limb + blockIdx.z;
    const int idx0 = blockIdx.y * 2 * kThreads;
    const size_t cs = (size_t)nl * T.N, ds = 2 * cs;
    const size_t off = (size_t)i * T.N + idx0;
    if (t == 0) {
        for (int s = 0; s < kMacStages; s++) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, kThreads / 32);
        }
    }
    __syncthreads();
    if (t >= kThreads) {
        if (t == kThreads) {
            const u64* qa = rotq + off;
            const u64* da = dbs.p[blockIdx.x] + off;
            for (int d = 0; d < ndiag; d++) {
                const int s = d % kMacStages;
                if (d >= kMacStages) mbar_wait(empty + s, ((d / kMacStages) - 1) & 1);
                mbar_expect_tx(full + s, 4 * kMacTileBytes);
                unsigned char* st = smraw + (size_t)s * 4 * kMacTileBytes;
                bulk_g2s(st, qa + (size_t)d * ds, kMacTileBytes, full + s);
                bulk_g2s(st + kMacTileBytes, qa + (size_t)d * ds + cs, kMacTileBytes, full + s);
                bulk_g2s(st + 2 * kMacTileBytes, da + (size_t)d * ds, kMacTileBytes, full + s);
                bulk_g2s(st + 3 * kMacTileBytes, da + (size_t)d * ds + cs, kMacTileBytes, full + s);
            }
        }
        return;
    }
    constexpr u32 kLow = (1u << kCut) - 1;
    const PrimeConst pc = T.pc[i];
    u64* o = out + (size_t)blockIdx.x * 3 * cs + off + 2 * t;
    if ((pc.q >> 46) != 0 || ndiag > kMaxDiagSmall) {      // the 60-bit limb: 128-bit accumulators folded every kFold diagonals
        u64 r[3][2] = {{0, 0}, {0, 0}, {0, 0}};
        u128d acc[3][2] = {};
        for (int d = 0; d < ndiag; d++) {
            const int s = d % kMacStages;
            mbar_wait(full + s, (d / kMacStages) & 1);
            const ulonglong2* st = reinterpret_cast<const ulonglong2*>(smraw + (size_t)s * 4 * kMacTileBytes);
            const ulonglong2 a0 = st[t], a1 = st[kThreads + t], b0 = st[2 * kThreads + t], b1 = st[3 * kThreads + t];
            __syncwarp();
            if ((t & 31) == 0) mbar_arrive(empty + s);
            acc_wide(acc[0][0], a0.x, b0.x); acc_wide(acc[0][1], a0.y, b0.y);
            acc_wide(acc[1][0], a0.x, b1.x); acc_wide(acc[1][1], a0.y, b1.y);
            acc_wide(acc[1][0], a1.x, b0.x); acc_wide(acc[1][1], a1.y, b0.y);
            acc_wide(acc[2][0], a1.x, b1.x); acc_wide(acc[2][1], a1.y, b1.y);
            if ((d + 1) % kFold == 0 || d + 1 == ndiag) {
#pragma unroll
                for (int c = 0; c < 3; c++)
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        r[c][e] = add_mod(r[c][e], redc(fold128(acc[c][e], pc.q, pc.mu64), pc.q, pc.qinv_neg), pc.q);
                        acc[c][e] = u128d{0, 0};
                    }
            }
        }
#pragma unroll
        for (int c = 0; c < 3; c++) {
            ulonglong2 v = {r[c][0], r[c][1]};
            st128(o + (size_t)c * cs, v);
        }
        return;
    }
    cols3 acc[3][2] = {};
    for (int d = 0; d < ndiag; d++) {
        const int s = d % kMacStages;
        mbar_wait(full + s, (d / kMacStages) & 1);
        const ulonglong2* st = reinterpret_cast<const ulonglong2*>(smraw + (size_t)s * 4 * kMacTileBytes);
        const ulonglong2 a0 = st[t], a1 = st[kThreads + t], b0 = st[2 * kThreads + t], b1 = st[3 * kThreads + t];
        __syncwarp();
        if ((t & 31) == 0) mbar_arrive(empty + s);
        const u32 a0x0 = (u32)a0.x & kLow, a0x1 = (u32)(a0.x >> kCut), a0y0 = (u32)a0.y & kLow, a0y1 = (u32)(a0.y >> kCut);
        const u32 a1x0 = (u32)a1.x & kLow, a1x1 = (u32)(a1.x >> kCut), a1y0 = (u32)a1.y & kLow, a1y1 = (u32)(a1.y >> kCut);
        const u32 b0x0 = (u32)b0.x & kLow, b0x1 = (u32)(b0.x >> kCut), b0y0 = (u32)b0.y & kLow, b0y1 = (u32)(b0.y >> kCut);
        const u32 b1x0 = (u32)b1.x & kLow, b1x1 = (u32)(b1.x >> kCut), b1y0 = (u32)b1.y & kLow, b1y1 = (u32)(b1.y >> kCut);
        mac23(acc[0][0], a0x0, a0x1, b0x0, b0x1); mac23(acc[0][1], a0y0, a0y1, b0y0, b0y1);
        mac23(acc[1][0], a0x0, a0x1, b1x0, b1x1); mac23(acc[1][1], a0y0, a0y1, b1y0, b1y1);
        mac23(acc[1][0], a1x0, a1x1, b0x0, b0x1); mac23(acc[1][1], a1y0, a1y1, b0y0, b0y1);
        mac23(acc[2][0], a1x0, a1x1, b1x0, b1x1); mac23(acc[2][1], a1y0, a1y1, b1y0, b1y1);
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
        ulonglong2 v;
        v.x = redc(fold128(fold23(acc[c][0]), pc.q, pc.mu64), pc.q, pc.qinv_neg);
        v.y = redc(fold128(fold23(acc[c][1]), pc.q, pc.mu64), pc.q, pc.qinv_neg);
        st128(o + (size_t)c * cs, v);
    }
}
void init_fused_kernels(int logN);   // hx_fused.cu
void init_device_kernels(int logN) {
    cudaError_t e = cudaFuncSetAttribute(k_diag_mac_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, kMacSmemBytes);
    if (e != cudaSuccess) throw std::runtime_error(std::string("cudaFuncSetAttribute(k_diag_mac_tma): ") + cudaGetErrorString(e));
    init_fused_kernels(logN);
}
void launch_diag_mac(const DevTables& T, const u64* rotq, const DbBlockPtrs& dbs, int nblocks, u64* out, int ndiag, int nl, cudaStream_t st) {
    if (T.N % (2 * kThreads) == 0) {
        launch_pdl(k_diag_mac_tma, dim3(nblocks, T.N / (2 * kThreads), nl), dim3(kThreads + 32), (size_t)kMacSmemBytes, st, T, rotq, dbs, out, ndiag,
                   nl, 0);
        return;
    }
    dim3 grid(nblocks, blocks_for(T.N), nl);      // rings smaller than one 512-coefficient tile: the register-fed kernel
    launch_pdl(k_diag_mac<1>, grid, dim3(kThreads), 0, st, T, rotq, dbs, out, ndiag, nl, nblocks);
}

}  // namespace hx
