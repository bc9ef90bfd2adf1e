// hx_kernels.cuh -- launch interface of the sm_100a kernels (definitions in hx_kernels.cu).
// Polynomial data: uint64 limb-major [limb][N], EVALUATION form unless stated, canonical residues.
#pragma once
#include <stdexcept>
#include <string>
#include <utility>
#include "hx_modarith.cuh"

namespace hx {

constexpr int kMaxPrimes = 40;       // L + K upper bound (approach 2 needs 19 + 6)
constexpr int kMaxDigits = 4;

struct ScalarTable { u64 v[kMaxPrimes]; u64 sh[kMaxPrimes]; };   // per-limb constants with Shoup quotients

// prime index of the b-th polynomial of a batch = prime[b % period]
struct LimbMap {
    int period;
    short prime[kMaxPrimes];
};

struct DevTables {
    int N, logN;
    int nprimes;
    const PrimeConst* pc;        // [nprimes]
    const ulonglong2* tw_fwd;    // [nprimes][N]  {psi^brev(i), shoup}
    const ulonglong2* tw_inv;    // [nprimes][N]  {psi^-brev(i), shoup}
    const double2* tw_fwd_d;     // [nprimes][N]  {psi^brev(i), psi^brev(i)/q} as doubles (used for primes below 2^46)
};

// ---- NTT (row P2)
void launch_ntt(const DevTables& T, u64* data, int nbatch, const LimbMap& lm, bool inverse, cudaStream_t st);

// ---- base conversion (rows P4, P5): coefficient-domain ApproxSwitchCRTBasis
struct BaseConvTable {          // device-resident
    int n_in, n_out;
    short src[8];               // prime index of each source limb
    short dst[kMaxPrimes];      // prime index of each target limb
    u64 hatinv[8], hatinv_sh[8];           // (Q/q_i)^-1 mod q_i  (Shoup pair)
    u64 mat[8 * kMaxPrimes];               // (Q/q_i) * 2^64 mod p_j  (Montgomery form), [i][j]
    u64 hatinv_n[8], hatinv_n_sh[8];       // hatinv * N^-1 mod q_i: folds the inverse transform's scaling (fused path)
    u64 mat30[8 * kMaxPrimes];             // mat, pre-split into 30-bit halves (split30) for the carry-free MAC
};
// in: [nbatch][n_in][N] (stride_in limbs between batches), out: [nbatch][n_out][N]
void launch_baseconv(const DevTables& T, const BaseConvTable* tab, int n_in, int n_out, const u64* in, size_t in_batch_stride,
                     u64* out, size_t out_batch_stride, int nbatch, cudaStream_t st);

// ---- key-switch inner product (row P5): ct0/ct1[i] = sum_j digit_j[i] * keyB_j/keyA_j[i]
// digits: [ndig][nl+K][N]; keys (Montgomery form): [dnum][L+K][N]; out0,out1: [nl+K][N] per batch entry.
void launch_ks_inner(const DevTables& T, const u64* digits, int ndig, int nl, int L, int K,
                     const u64* const* keyA, const u64* const* keyB, u64* out, size_t out_batch_stride,
                     int nbatch, cudaStream_t st);

// ---- ModDown epilogue (+ optional add of c0, + optional automorphism gather) (rows P3, P5)
// out[i][j] = ( (ct[i][m] - conv[i][m]) * Pinv_i + (addend ? addend[i][m] : 0) ),  m = map ? map[j] : j
void launch_moddown_finish(const DevTables& T, const u64* ct, size_t ct_batch_stride, const u64* conv, size_t conv_batch_stride,
                           const u64* addend, int addend_comps, const u32* const* maps, u64* out, size_t out_batch_stride, int nl,
                           int nbatch, cudaStream_t st);

// ---- fused key switch (hx_fused.cu): rows P2-P5 in three launches per batch of keys
constexpr int kMaxKsBatch = 32;
struct FusedKsArgs {
    const u64* digits;              // [ndig][nl+K][N] ModUp output (EvalKeySwitchPrecomputeCore)
    int ndig, nl, L, K;
    const u64* keyA[kMaxKsBatch];   // per batch entry: evk a / b, Montgomery form, [dnum][L+K][N]
    const u64* keyB[kMaxKsBatch];
    u32 galois[kMaxKsBatch];        // Galois element applied to the result (1 = none)
    u64* tmpP;                      // scratch [nb][2][K][N]
    u64* conv;                      // scratch [nb][2][nl][N]
    const u64* add0;                // optional addend of component 0 / 1, [nl][N]
    const u64* add1;
    size_t digits_bs = 0;           // words between the digit sets of consecutive batch entries (0 = shared: hoisted rotations)
    size_t add_bs = 0;              // words between the addends of consecutive batch entries (0 = shared)
    u64* out;                       // [nb] x out_bs, each [2][nl][N]
    size_t out_bs;
    int moddown_w = 1;              // Engine::ks_moddown_w: warp-task ModDown middle (hx_moddown.cu) where the ring supports it
    unsigned long long fp_mask = 0; // bit i: Q limb i is an FP64 limb (q_i < 2^46); set by the engine
    short limbs[kMaxPrimes];        // filled per launch of the Q-limb kernel: the limbs grid.y walks (FP64 limbs / 60-bit limbs)
};
bool fused_ks_supported(int logN);
// per-device opt-ins (dynamic shared memory above 48 KB) of every kernel a ring of 2^logN coefficients launches; called by the
// Engine constructor with its device current
void init_device_kernels(int logN);
void launch_fused_ks(const DevTables& T, const FusedKsArgs& a, const BaseConvTable* moddown, int nb, cudaStream_t st);
// ModDown middle as warp tasks (hx_moddown.cu): in [nbatch][n_in][N] (inverse row pass done) -> out [nbatch][n_out][N] (forward column pass done)
bool moddown_w_supported(int logN, int n_in);
void init_moddown_kernels(int logN);
void launch_moddown_w(const DevTables& T, const BaseConvTable* tab, int n_in, int nbatch, const u64* in, size_t in_bs, u64* out, size_t out_bs, cudaStream_t st);

// fused ModUp: c [nbatch] x c_bs, each [nl][N]  ->  digits [nbatch][ndig][nl+K][N]
struct ModUpArgs {
    const u64* c; size_t c_bs; int nbatch;
    int nl, L, K, ndig, alpha;
    const BaseConvTable* tabs[kMaxDigits];   // per digit: own limbs -> complement (Q then P)
    u64* scratch;                            // [nbatch][nl][N]
    u64* digits;
    int warp_tasks = 1;                      // Engine::ks_moddown_w
};
void launch_fused_modup(const DevTables& T, const ModUpArgs& m, cudaStream_t st);
void launch_modup_w(const DevTables& T, const ModUpArgs& m, cudaStream_t st);   // hx_moddown.cu: the middle of ModUp as warp tasks
// fused rescale: c [ncomp] x c_bs, each [nl][N]  ->  out [ncomp] x out_bs, each [nl-1][N]
struct RescaleArgs {
    const u64* c; size_t c_bs; int ncomp, nl;
    const BaseConvTable* last_tab;           // 1-limb table of prime nl-1 (hatinv_n = N^-1)
    u64* last;                               // scratch [ncomp][N]
    u64* tmp;                                // scratch [ncomp][nl-1][N]
    u64* out; size_t out_bs;
    ScalarTable negqlinv, qlinv;
};
void launch_fused_rescale(const DevTables& T, const RescaleArgs& r, cudaStream_t st);

// ---- elementwise family (row P8)
enum EwOp { EW_ADD = 0, EW_SUB = 1, EW_MUL = 2 };
void launch_ew(const DevTables& T, EwOp op, u64* out, const u64* a, const u64* b, int nl, int ncomp, size_t comp_stride_out,
               size_t comp_stride_a, size_t comp_stride_b, cudaStream_t st);
void launch_mul_scalar(const DevTables& T, u64* out, const u64* a, const ScalarTable& s, int nl, int ncomp, size_t cs_out, size_t cs_a, cudaStream_t st);
// out = a with s added to component 0 of every ciphertext: ncomps_total = nbatch*ncomp contiguous components of nl limbs
void launch_add_scalar(const DevTables& T, u64* out, const u64* a, const ScalarTable& s, int nl, int ncomp, int ncomps_total, cudaStream_t st);
// EvalLinearWSum core: out = sum_t w_t * ct_t over n <= 8 ciphertexts of one shape (ncomps_total contiguous components of nl limbs)
constexpr int kMaxWsum = 8;
struct WsumArgs { const u64* p[kMaxWsum]; int n; u64 v[kMaxWsum][kMaxPrimes]; u64 sh[kMaxWsum][kMaxPrimes]; };
void launch_wsum(const DevTables& T, u64* out, const WsumArgs& a, int nl, int ncomps_total, cudaStream_t st);
// EvalMultCore: (a0*b0, a0*b1+a1*b0, a1*b1); a,b: [2][.][N] with component strides; out [3][nl][N]
// batch entry z: a + z*2*cs_a, b + z*2*cs_b (contiguous 2-component ciphertexts), out + z*3*nl*N
// bcast_a / bcast_b: that operand is ONE ciphertext used by every batch entry
void launch_tensor(const DevTables& T, u64* out, const u64* a, size_t cs_a, bool bcast_a, const u64* b, size_t cs_b, bool bcast_b, int nl, int nbatch, cudaStream_t st);
// EvalMultNoRelin of nbatch (query, database) pairs (row R7, sender_hers.cpp:61-98): query ciphertext z = a.p[z]
// ([2][nl_a][N], plain residues, component stride cs_a), database ciphertext z = db + z*2*cs_b (Montgomery form,
// as registered by hx_db_register); out [nbatch][3][nl][N].
struct CtPtrs { const u64* p[kMaxKsBatch]; };
void launch_tensor_db(const DevTables& T, u64* out, const CtPtrs& a, size_t cs_a, const u64* db, size_t cs_b, int nl, int nbatch, cudaStream_t st);
// acc[c][i][n] (+)= sum_z in[z][c][i][n] mod q_i over the nbatch entries of a batched ciphertext; ncl = ncomp*nl limbs
// per entry, limb y has prime index y % nl.  accumulate = 0 overwrites acc.
void launch_batch_sum(const DevTables& T, u64* acc, const u64* in, int nl, int ncomp, int nbatch, int accumulate, cudaStream_t st);
// in place x -> x mod q_limb for un-reduced 64-bit sums; data [ncomps][nl][N]
void launch_fold_mod(const DevTables& T, u64* data, int nl, int ncomps, cudaStream_t st);
void launch_automorph(const DevTables& T, u64* out, const u64* in, const u32* map, int nlimbs_total, cudaStream_t st);
void launch_automap(int N, int logN, u64 k, u32* map, cudaStream_t st);
// Launch with programmatic stream serialization (the kernel's first statement must be pdl_wait()): hides the launch and
// block-scheduling latency between the ~1000 small dependent kernels of a step.
// after every <<<>>> launch: a failed launch (bad configuration, sticky error) surfaces at the call, not at the next synchronisation
inline void check_launch(const char* kernel) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw std::runtime_error(std::string("kernel launch failed (") + kernel + "): " + cudaGetErrorString(e));
}
extern int g_pdl;   // 0: plain stream-ordered launches (hx_ctx_set_option "pdl"; diagnosis of launch-overlap hazards)
template <class... KArgs, class... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = g_pdl ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
    if (e != cudaSuccess) throw std::runtime_error(std::string("kernel launch failed: ") + cudaGetErrorString(e));
}
void launch_to_mont(const DevTables& T, u64* data, int nbatch, const LimbMap& lm, cudaStream_t st, bool inverse = false);
// Engine-internal evaluation-key form, in place on [dnum][L+K][N] plain residues (Engine::finish_key): Q limbs are multiplied by
// P^-1 mod q_i; limbs below 2^46 (the FP64 limbs of the key-switch kernels) become doubles holding that residue, all other limbs
// become Montgomery residues split into 30-bit halves.
void launch_key_finish(const DevTables& T, u64* key, int dnum, int L, int K, const ScalarTable& pinv, cudaStream_t st);
// the inverse: engine-internal key words -> plain residues (out of place), for key export (serialisation); pmod[i] = P mod q_i
void launch_key_unfinish(const DevTables& T, u64* out, const u64* key, int dnum, int L, int K, const ScalarTable& pmod, cudaStream_t st);
// in place: x -> split30(x) (engine-internal key form, see hx_modarith.cuh)
void launch_split30(u64* data, size_t words, cudaStream_t st);

// ---- rescale (row P7)
// last: [N] coefficient form of limb nl-1 ; tmp out: [nl-1][N] = lift(last)*(-q_l^-1) mod q_i
void launch_rescale_prep(const DevTables& T, const u64* last, size_t last_batch_stride, u64* tmp, size_t tmp_batch_stride,
                         int nl, const ScalarTable& negqlinv, int nbatch, cudaStream_t st);
// out_i = c_i * qlinv_i + tmp_i
void launch_rescale_finish(const DevTables& T, u64* out, size_t out_bs, const u64* c, size_t c_bs, const u64* tmp, size_t tmp_bs,
                           int nl, const ScalarTable& qlinv, int nbatch, cudaStream_t st);

// ---- fused diagonal multiply-accumulate (row P6)
// rotq: [ndiag][2][nl][N]; dbs.p[g] (Montgomery form): [ndiag][2][nl][N]; out: [nblocks][3][nl][N] (overwritten).
// All blocks of one launch share the rotated query: the block index is the fastest grid dimension, so the
// CTAs reading the same rotq tile are co-resident and every rotq line is fetched from HBM once per launch.
constexpr int kMaxDbBlocks = 64;
struct DbBlockPtrs { const u64* p[kMaxDbBlocks]; };
void launch_diag_mac(const DevTables& T, const u64* rotq, const DbBlockPtrs& dbs, int nblocks, u64* out, int ndiag, int nl, cudaStream_t st);

}  // namespace hx
