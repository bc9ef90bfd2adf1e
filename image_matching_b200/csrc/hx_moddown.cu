// hx_moddown.cu -- ApproxModDown's coefficient-domain middle for rings of 2^15 / 2^16 coefficients:
// inverse column pass of the K P limbs, * (P/p_k)^-1 N^-1, approximate base conversion to the l Q limbs, forward column pass of
// each -- OpenFHE's ApproxModDown as reached from cc->EvalFastRotation (/root/reference/src/sender/sender_diag.cpp:25),
// RelinearizeInPlace (:79) and EvalRotate (src/openFHE_wrapper.cpp:124), between the two row-pass kernels of hx_fused.cu.
//
// Work decomposition (round 2; replaces the r1 launch pair k_inv_cols + k_conv_fwd_col for these rings): a CTA owns 8 adjacent columns of the
// 2^A x 2^B coefficient matrix of ONE (rotation, component); a WARP owns a task = one limb x a group of 512 >> A columns
// = 512 coefficients, 16 per thread, and runs the whole column transform of its task alone:
//   * 4 stages on the 16 register-held elements whose row index differs in its top 4 bits (their twiddles depend only on the
//     element's position in the thread, so they are warp-uniform loads),
//   * ONE exchange through a warp-private shared-memory buffer (__syncwarp only),
//   * the remaining A - 4 stages on 16 elements whose row index differs in its low bits, loaded from / stored to global memory
//     directly in that ownership (32-byte sectors fully used).
// No block-wide barrier inside a transform (the r1 kernel had ~40 per CTA), half the shared-memory exchange traffic, and the
// conversion's multiply-accumulate results feed the forward transform from registers.  The K scaled source limbs of the CTA
// are staged once in shared memory (30-bit split words) and read by every (target limb, column group) task; the only
// __syncthreads separates the K x groups inverse tasks from the l x groups forward tasks, both dealt round-robin to the 8 warps.
// All values are exact residues: bit-identical to the unfused kernels and to the CPU oracle.
#include "hx_kernels.cuh"
#include "hx_ntt_tile.cuh"
#include "hx_fpmod.cuh"
#include <algorithm>

#ifndef HX_MDW_MINB
#define HX_MDW_MINB 3
#endif
namespace hx {

template <int LOGN>
struct WGeo {
    static constexpr int A = LOGN / 2, B = LOGN - A;
    static constexpr int ROWS = 1 << A;
    static constexpr int NR = 1 << (A - 4);            // row-lows a thread holds in the low-bit round (8 / 16)
    static constexpr int LOGNC = 4 - (A - 4);           // log2 of the adjacent columns it holds there (1 / 0)
    static constexpr int NC = 1 << LOGNC;
    static constexpr int WCOLS = 512 >> A;              // columns per warp task (4 / 2)
    static constexpr int CTA_COLS = 8;
    static constexpr int GROUPS = CTA_COLS / WCOLS;     // tasks per limb (2 / 4)
    static constexpr int XS = 32 + (NC == 2 ? 4 : 2);   // words per top-bits block of the exchange buffer: padded, conflict free both ways
    static constexpr int XWORDS = 16 * XS;
    static constexpr int STAGE_WORDS = ROWS * CTA_COLS; // per source limb
    static constexpr int WARPS = 8;
    static_assert(A == 7 || A == 8, "warp-task column transform: rings of 2^15 / 2^16 coefficients");
};
template <int LOGN>
constexpr unsigned moddown_w_smem(int n_in) {
    return (unsigned)(sizeof(u64) * ((size_t)n_in * WGeo<LOGN>::STAGE_WORDS + (size_t)WGeo<LOGN>::WARPS * WGeo<LOGN>::XWORDS));
}

// ---- butterflies (same arithmetic as hx_ntt_tile.cuh)
__device__ __forceinline__ void bf_inv_int(u64& x, u64& y, const ulonglong2 w, u64 q, u64 q4) {   // Gentleman-Sande, values in [0,4q)
    const u64 a = x, b = y;
    u64 s = a + b;
    s = s >= q4 ? s - q4 : s;
    x = s;
    y = shoup_lazy4(a - b + q4, w.x, w.y, q);
}
__device__ __forceinline__ void bf_fwd_int(u64& x, u64& y, const ulonglong2 w, u64 q, u64 q4) {   // Cooley-Tukey, values in [0,8q)
    u64 a = x;
    a = a >= q4 ? a - q4 : a;
    const u64 m = shoup_lazy4(y, w.x, w.y, q);
    x = a + m;
    y = a - m + q4;
}
__device__ __forceinline__ void bf_fwd_fp(double& x, double& y, const double2 w, double q) {      // centred lazy doubles
    const double b = y;
    const double c = __dsub_rn(__fma_rn(b, w.y, kMagicRint), kMagicRint);   // rint(b * w/q)
    const double hi = __dmul_rn(b, w.x);
    const double lo = __fma_rn(b, w.x, -hi);
    const double m = __dadd_rn(__fma_rn(-c, q, hi), lo);
    const double a = x;
    x = __dadd_rn(a, m);
    y = __dsub_rn(a, m);
}
// Montgomery REDC without the final conditional subtraction: T < q 2^64 -> T 2^-64 mod q in [0, 2q)
__device__ __forceinline__ u64 redc_lazy(u128d t, u64 q, u64 qinv_neg) {
    const u64 m = t.lo * qinv_neg;
    return t.hi + __umul64hi(m, q) + (t.lo != 0 ? 1ull : 0ull);
}

__device__ __forceinline__ u64 as_word(u64 x) { return x; }
__device__ __forceinline__ u64 as_word(double x) { return fp_double_to_bits(x); }
template <class V> __device__ __forceinline__ V word_as(u64 w);
template <> __device__ __forceinline__ u64 word_as<u64>(u64 w) { return w; }
template <> __device__ __forceinline__ double word_as<double>(u64 w) { return fp_bits_to_double(w); }

// The two ownerships of a warp task (512 coefficients = ROWS x WCOLS, row r = rh * NR + rlo):
//   top:  lane = rlo * WCOLS + c holds rh = 0..15                        (stages 0..3 pair bits of rh)
//   low:  lane = rh * 2 + cq     holds rlo = 0..NR-1, c = cq*NC..+NC-1   (stages 4..A-1 pair bits of rlo); element index rlo*NC + cc
template <int LOGN, class V>
__device__ __forceinline__ void xch_top_to_low(u64* xch, int lane, const V (&v)[16], V (&u)[16]) {
    using W = WGeo<LOGN>;
    static_assert(sizeof(V) == 8, "64-bit words");
#pragma unroll
    for (int rh = 0; rh < 16; rh++) xch[rh * W::XS + lane] = as_word(v[rh]);
    __syncwarp();
    const u64* p = xch + (lane >> 1) * W::XS + (lane & 1) * W::NC;
#pragma unroll
    for (int rlo = 0; rlo < W::NR; rlo++) {
        if constexpr (W::NC == 2) {   // one 128-bit access per row: conflict free with the padded block stride
            const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(p + rlo * W::WCOLS);
            u[rlo * 2] = word_as<V>(x.x);
            u[rlo * 2 + 1] = word_as<V>(x.y);
        } else {
            u[rlo] = word_as<V>(p[rlo * W::WCOLS]);
        }
    }
    __syncwarp();
}
template <int LOGN, class V>
__device__ __forceinline__ void xch_low_to_top(u64* xch, int lane, const V (&u)[16], V (&v)[16]) {
    using W = WGeo<LOGN>;
    u64* p = xch + (lane >> 1) * W::XS + (lane & 1) * W::NC;
#pragma unroll
    for (int rlo = 0; rlo < W::NR; rlo++) {
        if constexpr (W::NC == 2) {
            ulonglong2 x = {as_word(u[rlo * 2]), as_word(u[rlo * 2 + 1])};
            *reinterpret_cast<ulonglong2*>(p + rlo * W::WCOLS) = x;
        } else {
            p[rlo * W::WCOLS] = as_word(u[rlo]);
        }
    }
    __syncwarp();
#pragma unroll
    for (int rh = 0; rh < 16; rh++) v[rh] = word_as<V>(xch[rh * W::XS + lane]);
    __syncwarp();
}

// forward stages 0..3 on the top ownership: twiddle index 2^s + (rh >> (4 - s)) -- position in the thread only
template <class V, class TW, class BF>
__device__ __forceinline__ void fwd_top(V (&v)[16], const TW* __restrict__ tw, BF bf) {
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const int hb = 3 - s;
#pragma unroll
        for (int h = 0; h < 8; h++) {
            const int i0 = ((h >> hb) << (hb + 1)) | (h & ((1 << hb) - 1)), i1 = i0 | (1 << hb);
            bf(v[i0], v[i1], __ldg(tw + (1 << s) + (i0 >> (hb + 1))));
        }
    }
}
// forward stages 4..A-1 on the low ownership: twiddle index 2^s + rh 2^(s-4) + (rlo >> (A - s))
template <int LOGN, class V, class TW, class BF>
__device__ __forceinline__ void fwd_low(V (&u)[16], int rh, const TW* __restrict__ tw, BF bf) {
    using W = WGeo<LOGN>;
#pragma unroll
    for (int s = 4; s < W::A; s++) {
        const int ib = (W::A - 1 - s) + W::LOGNC;      // element-index bit this stage pairs
        const TW* twp = tw + (1 << s) + (rh << (s - 4));
#pragma unroll
        for (int h = 0; h < 8; h++) {
            const int i0 = ((h >> ib) << (ib + 1)) | (h & ((1 << ib) - 1)), i1 = i0 | (1 << ib);
            bf(u[i0], u[i1], __ldg(twp + (i0 >> (ib + 1))));
        }
    }
}
// inverse: the same stages in the opposite order
template <int LOGN, class V, class TW, class BF>
__device__ __forceinline__ void inv_low(V (&u)[16], int rh, const TW* __restrict__ tw, BF bf) {
    using W = WGeo<LOGN>;
#pragma unroll
    for (int s = W::A - 1; s >= 4; s--) {
        const int ib = (W::A - 1 - s) + W::LOGNC;
        const TW* twp = tw + (1 << s) + (rh << (s - 4));
#pragma unroll
        for (int h = 0; h < 8; h++) {
            const int i0 = ((h >> ib) << (ib + 1)) | (h & ((1 << ib) - 1)), i1 = i0 | (1 << ib);
            bf(u[i0], u[i1], __ldg(twp + (i0 >> (ib + 1))));
        }
    }
}
template <class V, class TW, class BF>
__device__ __forceinline__ void inv_top(V (&v)[16], const TW* __restrict__ tw, BF bf) {
#pragma unroll
    for (int s = 3; s >= 0; s--) {
        const int hb = 3 - s;
#pragma unroll
        for (int h = 0; h < 8; h++) {
            const int i0 = ((h >> hb) << (hb + 1)) | (h & ((1 << hb) - 1)), i1 = i0 | (1 << hb);
            bf(v[i0], v[i1], __ldg(tw + (1 << s) + (i0 >> (hb + 1))));
        }
    }
}

// grid (2^B / 8 column tiles, nbatch) ; in: [nbatch][n_in][N] (K1's output: inverse row pass done, values in [0,4q));
// out: [nbatch][n_out][N], forward column pass done, pass-internal words (lazy doubles on FP64 limbs, [0,8q) on 60-bit limbs)
// One launch serves ModDown (one table, output limb j at position j) and ModUp (one table per digit: blockIdx.y = entry * ndig +
// digit, source limbs of digit d start at in_off[d], the digit's extended polynomial at out_off[d], output limb of prime p at
// its extended-basis position: Q limbs 0..ext_nl-1, then the P limbs).
struct MdwArgs {
    const BaseConvTable* tab[kMaxDigits];
    size_t in_off[kMaxDigits], out_off[kMaxDigits];
    int ndig;
    size_t in_bs, out_bs;      // words between batch entries
    int ext_nl, L;             // ext_nl < 0: ModDown positions
};
// NIN: the number of source limbs when every table of the launch has that many (the conversion's multiply-accumulates are then
// straight-line code the scheduler can interleave; with a run-time count each one sits behind its own branch), 0 = read it from the table.
template <int LOGN, int NIN>
__global__ void __launch_bounds__(256, LOGN <= 15 ? HX_MDW_MINB : 2) k_moddown_w(const __grid_constant__ DevTables T, const __grid_constant__ MdwArgs A,
                                                                       const u64* in, u64* out) {
    using W = WGeo<LOGN>;
    pdl_wait();
    extern __shared__ __align__(16) u64 smw[];
    const int dgt = blockIdx.y % A.ndig, entry = blockIdx.y / A.ndig;
    const BaseConvTable* __restrict__ tab = A.tab[dgt];
    const int n_in = NIN ? NIN : tab->n_in, n_out = tab->n_out;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64* const stage = smw;                                                   // [n_in][GROUPS][ROWS][WCOLS], 30-bit split words
    u64* const xch = smw + (size_t)n_in * W::STAGE_WORDS + warp * W::XWORDS;  // this warp's exchange buffer
    const int rhB = lane >> 1, cqB = lane & 1;                                // low ownership
    const size_t colbase = (size_t)blockIdx.x * W::CTA_COLS;
    in += (size_t)entry * A.in_bs + A.in_off[dgt];
    out += (size_t)entry * A.out_bs + A.out_off[dgt];

    // ---- inverse tasks: (source limb k, column group g)
    for (int task = warp; task < n_in * W::GROUPS; task += W::WARPS) {
        const int k = task / W::GROUPS, g = task % W::GROUPS;
        const int prime = tab->src[k];
        const u64 q = T.pc[prime].q, q4 = q << 2;
        const ulonglong2* __restrict__ twi = T.tw_inv + (size_t)prime * T.N;
        u64 u[16], v[16];
        const u64* src = in + (size_t)k * T.N + colbase + g * W::WCOLS + cqB * W::NC + ((size_t)(rhB * W::NR) << W::B);
#pragma unroll
        for (int rlo = 0; rlo < W::NR; rlo++) {
            if constexpr (W::NC == 2) {
                const ulonglong2 x = ldg128(src + ((size_t)rlo << W::B));
                u[rlo * 2] = x.x;
                u[rlo * 2 + 1] = x.y;
            } else {
                u[rlo] = ldg64(src + ((size_t)rlo << W::B));
            }
        }
        auto bf = [q, q4](u64& x, u64& y, const ulonglong2 w) { bf_inv_int(x, y, w, q, q4); };
        inv_low<LOGN>(u, rhB, twi, bf);
        xch_low_to_top<LOGN>(xch, lane, u, v);
        inv_top(v, twi, bf);
        const u64 h = tab->hatinv_n[k], hs = tab->hatinv_n_sh[k];
        u64* st = stage + ((size_t)(k * W::GROUPS + g) * W::ROWS) * W::WCOLS + lane;   // row rh*NR + rlo, column c: word (rh*NR)*WCOLS + lane
#pragma unroll
        for (int rh = 0; rh < 16; rh++) st[rh * W::NR * W::WCOLS] = split30(shoup_mul(v[rh], h, hs, q));
    }
    __syncthreads();

    // ---- forward tasks: (target limb j, column group g)
    for (int task = warp; task < n_out * W::GROUPS; task += W::WARPS) {
        const int j = task / W::GROUPS, g = task % W::GROUPS;
        const int prime = tab->dst[j];
        const PrimeConst pc = T.pc[prime];
        const u64 q = pc.q;
        constexpr int KM = NIN ? NIN : 7;
        u64 mk[KM];
#pragma unroll
        for (int k = 0; k < KM; k++) mk[k] = (NIN || k < n_in) ? tab->mat30[k * n_out + j] : 0ull;
        const u64* st = stage + ((size_t)g * W::ROWS) * W::WCOLS + lane;
        u64 r[16];
#pragma unroll
        for (int rh = 0; rh < 16; rh++) {
            acc3 acc = {0, 0, 0};
#pragma unroll
            for (int k = 0; k < KM; k++)
                if (NIN || k < n_in) mac30ss(acc, st[(size_t)k * W::STAGE_WORDS + rh * W::NR * W::WCOLS], mk[k]);
            r[rh] = redc_lazy(fold30(acc), q, pc.qinv_neg);       // in [0, 2q): sum of <= 7 products below 2^60 q
        }
        const int pos = A.ext_nl < 0 ? j : (prime < A.L ? prime : A.ext_nl + prime - A.L);
        u64* dst = out + (size_t)pos * T.N + colbase + g * W::WCOLS + cqB * W::NC + ((size_t)(rhB * W::NR) << W::B);
        if (fp_ok(q)) {
            const double qd = fp_from_u64(q);
            const double2* __restrict__ twd = T.tw_fwd_d + (size_t)prime * T.N;
            double v[16], u[16];
#pragma unroll
            for (int rh = 0; rh < 16; rh++) v[rh] = fp_from_u64(r[rh]);
            auto bf = [qd](double& x, double& y, const double2 w) { bf_fwd_fp(x, y, w, qd); };
            fwd_top(v, twd, bf);
            xch_top_to_low<LOGN>(xch, lane, v, u);
            fwd_low<LOGN>(u, rhB, twd, bf);
#pragma unroll
            for (int rlo = 0; rlo < W::NR; rlo++) {
                if constexpr (W::NC == 2) {
                    ulonglong2 x = {fp_double_to_bits(u[rlo * 2]), fp_double_to_bits(u[rlo * 2 + 1])};
                    st128(dst + ((size_t)rlo << W::B), x);
                } else {
                    dst[(size_t)rlo << W::B] = fp_double_to_bits(u[rlo]);
                }
            }
        } else {
            const u64 q4 = q << 2;
            const ulonglong2* __restrict__ twf = T.tw_fwd + (size_t)prime * T.N;
            u64 u[16];
            auto bf = [q, q4](u64& x, u64& y, const ulonglong2 w) { bf_fwd_int(x, y, w, q, q4); };
            fwd_top(r, twf, bf);
            xch_top_to_low<LOGN>(xch, lane, r, u);
            fwd_low<LOGN>(u, rhB, twf, bf);
#pragma unroll
            for (int rlo = 0; rlo < W::NR; rlo++) {
                if constexpr (W::NC == 2) {
                    ulonglong2 x = {u[rlo * 2], u[rlo * 2 + 1]};
                    st128(dst + ((size_t)rlo << W::B), x);
                } else {
                    dst[(size_t)rlo << W::B] = u[rlo];
                }
            }
        }
    }
}

bool moddown_w_supported(int logN, int n_in) { return (logN == 15 || logN == 16) && n_in >= 1 && n_in <= 7; }

template <int LOGN, int NIN>
static void init_tn() {
    cudaError_t e = cudaFuncSetAttribute(k_moddown_w<LOGN, NIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)moddown_w_smem<LOGN>(NIN ? NIN : 7));
    if (e != cudaSuccess) throw std::runtime_error(std::string("cudaFuncSetAttribute(k_moddown_w): ") + cudaGetErrorString(e));
}
template <int LOGN>
static void init_t() {
    init_tn<LOGN, 0>();
    init_tn<LOGN, 4>();
}
void init_moddown_kernels(int logN) {
    if (logN == 15) init_t<15>();
    if (logN == 16) init_t<16>();
}
// uniform_n_in: every table of the launch has max_n_in source limbs (the caller knows the digit sizes; the tables are device memory)
template <int LOGN, int NIN>
static void launch_tn(const DevTables& T, const MdwArgs& A, int max_n_in, int nbatch, const u64* in, u64* out, cudaStream_t st) {
    using W = WGeo<LOGN>;
    launch_pdl(k_moddown_w<LOGN, NIN>, dim3((1u << W::B) / W::CTA_COLS, nbatch * A.ndig), dim3(256), moddown_w_smem<LOGN>(max_n_in), st, T, A, in, out);
}
template <int LOGN>
static void launch_t(const DevTables& T, const MdwArgs& A, int max_n_in, bool uniform_n_in, int nbatch, const u64* in, u64* out, cudaStream_t st) {
    if (uniform_n_in && max_n_in == 4) launch_tn<LOGN, 4>(T, A, max_n_in, nbatch, in, out, st);
    else launch_tn<LOGN, 0>(T, A, max_n_in, nbatch, in, out, st);
}
static void launch_any(const DevTables& T, const MdwArgs& A, int max_n_in, bool uniform_n_in, int nbatch, const u64* in, u64* out, cudaStream_t st) {
    if (T.logN == 15) launch_t<15>(T, A, max_n_in, uniform_n_in, nbatch, in, out, st);
    else if (T.logN == 16) launch_t<16>(T, A, max_n_in, uniform_n_in, nbatch, in, out, st);
    else throw std::runtime_error("warp-task base conversion: unsupported ring dimension");
}
void launch_moddown_w(const DevTables& T, const BaseConvTable* tab, int n_in, int nbatch, const u64* in, size_t in_bs, u64* out, size_t out_bs, cudaStream_t st) {
    MdwArgs A{};
    A.tab[0] = tab; A.ndig = 1; A.in_bs = in_bs; A.out_bs = out_bs; A.ext_nl = -1; A.L = 0;
    launch_any(T, A, n_in, true, nbatch, in, out, st);
}
// ModUp's middle (EvalKeySwitchPrecomputeCore): scratch [nbatch][nl][N] (inverse row pass done) -> the converted limbs of every
// digit of digits [nbatch][ndig][nl+K][N] (forward column pass done; k_modup_rows finishes them and copies the digit's own limbs)
void launch_modup_w(const DevTables& T, const ModUpArgs& m, cudaStream_t st) {
    const int ext = m.nl + m.K;
    MdwArgs A{};
    A.ndig = m.ndig; A.in_bs = (size_t)m.nl * T.N; A.out_bs = (size_t)m.ndig * ext * T.N; A.ext_nl = m.nl; A.L = m.L;
    int max_in = 1, min_in = 1 << 30;
    for (int d = 0; d < m.ndig; d++) {
        const int lo = d * m.alpha, hi = (d == m.ndig - 1) ? m.nl : (d + 1) * m.alpha;
        A.tab[d] = m.tabs[d];
        A.in_off[d] = (size_t)lo * T.N;
        A.out_off[d] = (size_t)d * ext * T.N;
        max_in = std::max(max_in, hi - lo);
        min_in = std::min(min_in, hi - lo);
    }
    launch_any(T, A, max_in, min_in == max_in, m.nbatch, m.scratch, m.digits, st);
}

}  // namespace hx
