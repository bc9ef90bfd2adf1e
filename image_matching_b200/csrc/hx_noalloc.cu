// hx_noalloc.cu -- hx_rescale_noalloc: RescaleInPlace (DropLastElementAndScale) of one ciphertext without any allocation.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// Engine::rescale (hx_engine.cu) draws two scratch buffers and its output handle from the allocator and, the first time a level is
// rescaled, cudaMallocs the 1-limb conversion table of the dropped prime.  This variant launches the SAME four kernels
// (launch_fused_rescale: k_inv_rows, k_inv_cols, k_rescale_cols, k_rescale_rows) with all three of those in ONE file-scope
// __device__ array, and reads from / writes to device pointers the caller owns:
//
//     g_noalloc_scratch = [ 1-limb BaseConvTable | last: ncomp x N words | tmp: ncomp x (nl-1) x N words ]
//
// Limit: ncomp * nl * N <= HX_NOALLOC_MAX_WORDS (2^21 words = 16 MiB: 2 components x 16 limbs at N = 2^16, 2 x 32 at N = 2^15;
// bench.py's ciphertexts are 2 x 12 x 2^15).  Every GPU that loads the library holds these 16 MiB + 4 KiB until the process exits.
// The array is shared by every call and every context on a device and keeps the last call's contents: nothing in it has to start
// at zero (the table is rewritten by each call, `last` and `tmp` are fully overwritten before they are read), but two calls that
// overlap on the GPU -- different contexts, or one context's engine stream against another's -- race on it.  Calls of one context
// are ordered on its engine stream and are safe back to back.
#include <cstring>
#include <mutex>

#include "../../include/hydia_b200.h"
#include "hx_engine.h"
#include "hx_nt.h"

namespace hx {

constexpr size_t kNoallocTableWords = (sizeof(BaseConvTable) + 255) / 256 * 256 / sizeof(u64);   // table slot, 256-byte multiple
constexpr size_t kNoallocMaxWords = (size_t)HX_NOALLOC_MAX_WORDS;
static __device__ __align__(256) u64 g_noalloc_scratch[kNoallocTableWords + kNoallocMaxWords];

// Device code names the array directly: the table of the dropped prime (what Engine::last_limb_table uploads) is written in place.
// Only the fields the rescale kernels read are meaningful: one source limb, no targets, hatinv_n = N^-1 mod q.
__global__ void k_noalloc_last_table(int prime, u64 ninv, u64 ninv_sh) {
    BaseConvTable* t = reinterpret_cast<BaseConvTable*>(g_noalloc_scratch);
    t->n_in = 1;
    t->n_out = 0;
    t->src[0] = (short)prime;
    t->hatinv[0] = 1;
    t->hatinv_sh[0] = ~0ull;
    t->hatinv_n[0] = ninv;
    t->hatinv_n_sh[0] = ninv_sh;
}

// Host code needs the array's device address; its name in host code is only a stand-in.  Looked up once per device, with that
// device current (the caller holds the engine's Guard, which makes the engine's device current).
static u64* noalloc_scratch_ptr(int device) {
    static std::mutex mu;
    static u64* ptrs[64] = {};
    if (device < 0 || device >= 64) throw std::runtime_error("hx_rescale_noalloc: device index out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[device]) {
        u64* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) throw std::runtime_error(std::string("hx_rescale_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        ptrs[device] = p;
    }
    return ptrs[device];
}

// in: [ncomp][nl][N] evaluation-form residues at nl limbs; out: [ncomp][nl-1][N].  Both device memory of the engine's GPU, not overlapping.
void rescale_noalloc(Engine& e, const u64* in, int ncomp, int nl, u64* out) {
    Engine::Guard g(&e);
    if (!in || !out) throw std::runtime_error("hx_rescale_noalloc: null device pointer");
    if (ncomp < 1 || ncomp > 3) throw std::runtime_error("hx_rescale_noalloc: a ciphertext has 1..3 components");
    if (nl < 2 || nl > e.L) throw std::runtime_error("hx_rescale_noalloc: limb count must be 2..L (no limb left to drop, or more limbs than the context has)");
    if (!fused_ks_supported(e.logN)) throw std::runtime_error("hx_rescale_noalloc: ring dimension must be 2^12..2^16");
    const size_t need = (size_t)ncomp * nl * e.N;
    if (need > kNoallocMaxWords)
        throw std::runtime_error("hx_rescale_noalloc: ncomp * nlimbs * N = " + std::to_string(need) + " words exceeds the static scratch of " +
                                 std::to_string(kNoallocMaxWords) + " words (HX_NOALLOC_MAX_WORDS); use hx_rescale");
    u64* base = noalloc_scratch_ptr(e.device);
    const int N = e.N;
    const u64 ql = e.mods[nl - 1];
    RescaleArgs r;
    std::memset(&r.negqlinv, 0, sizeof(ScalarTable));
    std::memset(&r.qlinv, 0, sizeof(ScalarTable));
    for (int i = 0; i < nl - 1; i++) {   // q_l^-1 and -q_l^-1 mod q_i with their Shoup quotients (Engine::rescale / scalar_table)
        const u64 qi = e.mods[i], inv = invmod(ql % qi, qi), ng = (qi - inv) % qi;
        r.qlinv.v[i] = inv; r.qlinv.sh[i] = shoup(inv, qi);
        r.negqlinv.v[i] = ng; r.negqlinv.sh[i] = shoup(ng, qi);
    }
    const u64 ninv = invmod((u64)N % ql, ql);
    r.c = in; r.c_bs = (size_t)nl * N; r.ncomp = ncomp; r.nl = nl;
    r.last_tab = reinterpret_cast<const BaseConvTable*>(base);
    r.last = base + kNoallocTableWords;
    r.tmp = r.last + (size_t)ncomp * N;
    r.out = out; r.out_bs = (size_t)(nl - 1) * N;
    // launches are booked by hand: Engine::Scope may create CUDA events while the profiler is on
    e.launches += 5;
    e.prof_launches[PK_RESCALE] += 5;
    e.prof_bytes[PK_RESCALE] += 8.0 * N * ncomp * (nl + nl - 1);
    k_noalloc_last_table<<<1, 1, 0, e.st>>>(nl - 1, ninv, shoup(ninv, ql));
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) throw std::runtime_error(std::string("hx_rescale_noalloc: table kernel launch: ") + cudaGetErrorString(err));
    launch_fused_rescale(e.T, r, e.st);
}
size_t rescale_noalloc_max_words() { return kNoallocMaxWords; }

}  // namespace hx
