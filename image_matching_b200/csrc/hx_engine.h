// hx_engine.h -- host-side CKKS evaluation engine over the sm_100a kernels (product code).
// Mirrors the subset of lbcrypto::CryptoContext<DCRTPoly> the reference's Sender path calls
// (SURVEY §8(b) B2); each method cites the reference call site it serves.
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "hx_kernels.cuh"
#include "hx_nt.h"

namespace hx {

#define HX_CUDA(call)                                                                                     \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); \
    } while (0)

class Engine;

// Kernel families for launch accounting and the optional CUDA-event profiler (bench.py roofline).
enum ProfKind {
    PK_NTT_FWD = 0, PK_NTT_INV, PK_BASECONV, PK_KS_INNER, PK_MODDOWN, PK_EW, PK_SCALAR, PK_TENSOR, PK_AUTOMORPH,
    PK_RESCALE, PK_DIAG_MAC, PK_MISC, PK_KS_FUSED, PK_MODUP, PK_KS_HOISTED, PK_COUNT
};
const char* prof_kind_name(int kind);

// Ciphertext<DCRTPoly>: ncomp components x nl limbs x N, EVALUATION form, device resident.
// nb > 1: a BATCH of nb ciphertexts of identical shape/level/degree, contiguous [nb][ncomp][nl][N].  Every
// leveled op treats the batch entries independently with the same constants and keys in ONE launch per
// kernel; the HyDia path uses it to run the compare chain of all database blocks of a GPU together
// (the reference does the same chain per block in an OpenMP loop, sender_diag.cpp:40-43).
struct CtObj {
    Engine* eng = nullptr;
    std::shared_ptr<Engine> keep;   // a handle outliving hx_ctx_destroy keeps its engine alive
    u64* d = nullptr;
    int ncomp = 0, nl = 0;
    int nb = 1;         // batch entries
    int level = 0;      // GetLevel(): limbs dropped
    int deg = 1;        // GetNoiseScaleDeg()
    bool settled = false;   // contents complete on the device when the handle was returned (uploads synchronise) and never written again
    ~CtObj();
    u64* comp(int c) const;
    size_t comp_stride() const;
    size_t batch_stride() const { return (size_t)ncomp * comp_stride(); }
    int comps_total() const { return nb * ncomp; }
};
using Ct = std::shared_ptr<CtObj>;

struct EvalKey {        // device, Montgomery form, [dnum][L+K][N] each
    u64* a = nullptr;
    u64* b = nullptr;
};

class Engine : public std::enable_shared_from_this<Engine> {
public:
    Engine(int N, int L, int K, int dnum, const u64* q, const u64* p, const u64* psi_q, const u64* psi_p, int scale_bits, int device);
    ~Engine();

    int N, logN, L, K, dnum, alpha, scale_bits, device;
    std::vector<u64> mods, psis;     // Q then P
    DevTables T{};
    cudaStream_t st = nullptr;       // WORKING stream of the thread that holds `mu` (the hoisted rotations repoint it to their lanes)
    cudaStream_t st_main = nullptr;  // the engine stream; never changes.  Handles are freed on it (their destructors take `mu`)
    std::recursive_mutex mu;         // the reference enters from up to 48 OpenMP threads
    // Every entry point: take the engine lock and make the engine's device current for this host thread (contexts on several
    // devices in one process, worker threads whose current device is another one).
    struct Guard {
        std::unique_lock<std::recursive_mutex> lk;
        explicit Guard(Engine* e) : lk(e->mu) { cudaSetDevice(e->device); }
    };
    int ks_moddown_w = 1;            // ModDown's coefficient-domain middle as warp tasks (hx_moddown.cu) on rings of 2^15 / 2^16; 0 = the r1 launch sequence
    int rot_streams = 2;             // hoisted-rotation batches are issued round-robin on this many streams (1 = the engine stream only)
    cudaStream_t aux_streams[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t aux_events[4] = {nullptr, nullptr, nullptr, nullptr};
    // Rotation lane: the hoisted rotations of a SETTLED query (an uploaded ciphertext) depend on nothing the engine stream still
    // has queued, so the whole section (ModUp of the query + all batches) is issued on low-priority lane streams and the engine
    // stream (highest priority) joins at the end: the small dependent kernels of a previous call's compare chain, which leave
    // most SMs idle, run concurrently with the next call's rotations.  0 = everything ordered on the engine stream.
    int rot_lane = 1;
    cudaStream_t lane_streams[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t lane_events[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    int rot_chunk = 32;              // hoisted rotations per fused key-switch batch (intermediates stay L2 resident)
    bool force_unfused = false;      // tests: run the unfused reference kernels instead of hx_fused.cu
    int ks_kind = PK_KS_FUSED;       // profiler family of the next fused key switches (score_diag books its hoisted rotations separately)
    bool batch_blocks = true;        // HyDia path: relinearise/rescale/compare all blocks as one batch
    unsigned long long launches = 0; // kernels launched (bench "gpu_launches")

    // ---- launch accounting + optional CUDA-event profiler (events on `st`, the launching stream)
    struct ProfRec { int kind; cudaEvent_t a, b; };
    bool prof_no_events = false;             // inside a section timed as a whole (multi-stream hoisted rotations)
    unsigned prof_on = 0;                    // bitmask of ProfKind families being timed
    std::vector<ProfRec> prof_recs;
    std::vector<cudaEvent_t> prof_pool;
    unsigned long long prof_launches[PK_COUNT] = {};
    double prof_ms[PK_COUNT] = {};
    double prof_bytes[PK_COUNT] = {};        // algorithmic bytes (each operand read once, each result written once)
    struct Scope {                           // RAII around the launch(es) of one kernel family
        Engine* e; int kind; cudaEvent_t b = nullptr;
        Scope(Engine* e, int kind, int nlaunch, double bytes);
        ~Scope();
    };
    // ---- timeline marks (diagnostic; off by default): named CUDA events recorded on a stream, dumped as "name:ms;..." relative to
    // the mark that switched the trace on.  Used to name the limiter of a multi-GPU step without a system profiler.
    struct TraceRec { std::string name; cudaEvent_t ev; };
    bool trace_on = false;
    cudaEvent_t trace_origin = nullptr;
    std::vector<TraceRec> trace_recs;
    void trace_start();
    void trace_mark(const char* name, cudaStream_t s);   // no-op unless the trace is on
    std::string trace_dump();                            // synchronises the device, ends the trace
    void prof_enable(unsigned mask);
    void prof_reset();
    void prof_collect();                     // sync + fold event pairs into prof_ms
    // device timer for bench.py (events on `st`)
    cudaEvent_t timer_a = nullptr, timer_b = nullptr;
    void timer_start();
    double timer_stop_ms();

    // ---- memory
    u64* alloc_words(size_t words);
    void free_words(u64* p);
    Ct new_ct(int ncomp, int nl, int level, int deg, int nb = 1);
    Ct stack(const std::vector<Ct>& v);            // nb = v.size() batch from single ciphertexts (copies)
    std::vector<Ct> unstack(const Ct& a);          // batch -> single ciphertexts (copies)
    Ct concat(const std::vector<Ct>& v);           // batches of one shape -> one batch (entries in order; copies)
    std::vector<Ct> split(const Ct& a, int parts); // inverse of concat for equal-sized parts
    Ct upload(const u64* host, int ncomp, int nl, int level, int deg);
    void download(const Ct& ct, u64* host);
    Ct clone(const Ct& ct);
    void sync();

    // ---- keys (plain-form host arrays [dnum][L+K][N], evaluation form, OpenFHE limb order)
    void load_mult_key(const u64* a, const u64* b);
    void load_rot_key(u64 galois, const u64* a, const u64* b);
    void adopt_rot_key(u64 galois, EvalKey k);   // device-generated (client module)
    void adopt_mult_key(EvalKey k);
    void finish_key(EvalKey& k);                 // plain residues -> engine-internal key form (in place)
    bool has_rot_key(u64 galois) const { return rot_keys_.count(galois) != 0; }
    bool has_mult_key() const { return have_mult_key_; }
    std::vector<u64> rot_key_galois() const { std::vector<u64> v; for (auto& kv : rot_keys_) v.push_back(kv.first); return v; }
    // plain residues [dnum][L+K][N] of a loaded key back to the host (the inverse of finish_key); galois = 0: the mult key
    void export_key(u64 galois, u64* host_a, u64* host_b);
    u64 galois_for_rotation(int r) const;

    // ---- primitive evaluation ops
    Ct add(const Ct& a, const Ct& b);                       // EvalAdd / EvalAddInPlace (sender_diag.cpp:76)
    Ct sub(const Ct& a, const Ct& b);                       // EvalSub
    Ct add_const(const Ct& a, double c);                    // EvalAddInPlace(ct,double) (openFHE_wrapper.cpp:182)
    Ct mult_const(const Ct& a, double c);                   // EvalMult(ct,double)
    Ct mult_norelin(const Ct& a, const Ct& b);              // EvalMultNoRelin (sender_diag.cpp:93)
    Ct relinearize(const Ct& a);                            // RelinearizeInPlace (sender_diag.cpp:79)
    Ct mult(const Ct& a, const Ct& b);                      // EvalMult(ct,ct)
    Ct square(const Ct& a);                                 // EvalSquare
    // EvalMult(ct, Plaintext) (openFHE_wrapper.cpp:235,244,289): pt = encoded plaintext, host, [nlimbs][N], evaluation form
    Ct mult_plain(const Ct& a, const u64* host_pt, int nlimbs);
    Ct inner_product(const Ct& a, const Ct& b, int n);      // EvalInnerProduct (sender_base.cpp:93) = EvalSum(EvalMult(a,b), n)
    Ct rescale(const Ct& a);                                // RescaleInPlace / ModReduceInPlace (sender_diag.cpp:80)
    Ct level_reduce(const Ct& a, int levels);               // LevelReduceInPlace
    Ct rotate(const Ct& a, int r);                          // EvalRotate (openFHE_wrapper.cpp:124)
    Ct automorphism(const Ct& a, u64 galois);
    Ct eval_sum(const Ct& a, int n);                        // EvalSum (sender_diag.cpp:47)
    Ct add_many(const std::vector<Ct>& v);                  // EvalAddManyInPlace (sender_diag.cpp:46)
    // hoisted rotations (sender_diag.cpp:22-26)
    struct DbBlock { u64* d = nullptr; int ndiag = 0; int diag0 = 0; };   // diagonals diag0 .. diag0 + ndiag - 1 are resident
    // Diagonal range this context holds of EVERY database block (options "db_diag0" / "db_diag_count", set before registering):
    // the multi-GPU sender shards the 512 diagonals over the ranks, each rank computing its own rotations and the partial
    // sum over its diagonals for all blocks (sharded.py); a single GPU holds 0..511.
    int db_diag0 = 0, db_diag_count = 512;
    struct Digits { u64* d = nullptr; int ndig = 0, nl = 0; bool fused = false;   // fused ModUp stores 30-bit-split words
                    Engine* eng = nullptr; std::shared_ptr<Engine> keep; ~Digits(); };
    std::shared_ptr<Digits> fast_rotation_precompute(const Ct& a);
    Ct fast_rotation(const Ct& a, int r, const std::shared_ptr<Digits>& dg);

    // ---- PS evaluators (openFHE_wrapper.cpp:174,179)
    Ct chebyshev_series(const Ct& x, const std::vector<double>& coeffs, double a, double b);
    Ct eval_poly(const Ct& x, const std::vector<double>& coeffs);
    Ct chebyshev_compare(const Ct& x, double delta, int sign_depth);   // openFHE_wrapper.cpp:143-185
    Ct linear_wsum(const std::vector<Ct>& cts, const std::vector<double>& w);   // EvalLinearWSumMutable

    // ---- fused HyDia path (sender_diag.cpp:12-94)
    // database block g: 512 diagonal ciphertexts, kept HBM resident in Montgomery form
    void db_register(int block, int diag, const u64* host_ct /*[2][L][N]*/);
    void db_adopt(int block, u64* dev /*[ndiag][2][L][N], Montgomery form*/, int ndiag);
    void db_store_device(int block, int diag, const u64* dev_ct /*[2][L][N], plain residues, device*/);
    int db_blocks() const { return (int)db_.size(); }
    // rotations 0..ndiag-1 of the query (entry 0 = the query), hoisted and batched, as one [ndiag][2][L][N] buffer
    struct RotSet { u64* rotq = nullptr; int d0 = 0, ndiag = 0, nl = 0, level = 0, deg = 0; int slot = -1;   // rotations d0 .. d0 + ndiag - 1
                    Engine* eng = nullptr; std::shared_ptr<Engine> keep; ~RotSet(); };
    // The rotated-query buffers (ndiag ciphertexts, 3.2 GB at the reference's parameters) are written by the rotation lanes and
    // read by the diagonal MAC on the engine stream, and the NEXT query's rotations start while this one's MAC may still be queued.
    // They are engine-owned slots recycled under explicit events -- a released slot carries an event recorded on the engine stream
    // behind its last reader, and the next writer's stream waits for it -- instead of stream-ordered pool memory allocated on one
    // stream and freed on another: with pool recycling of these buffers ~1 % of the index scenarios on a 2^12 ring came out wrong
    // (tools/race_probe.py, profiles/r2_race_*.log), none with the buffer never recycled.
    struct RotSlot { u64* d = nullptr; size_t words = 0; bool busy = false; cudaEvent_t released = nullptr; bool has_release = false; };
    std::vector<RotSlot> rot_slots_;
    cudaEvent_t mac_gate_ = nullptr;                           // see mac_rotset
    int rot_slot_acquire(size_t words, cudaStream_t writer);   // returns the slot index; `writer` waits for the slot's last reader
    void rot_slot_release(int slot);                           // records the release event on the engine stream
    std::shared_ptr<RotSet> rotate_hoisted_all(const Ct& query, int ndiag) { return rotate_hoisted_range(query, 0, ndiag); }   // sender_diag.cpp:20-26
    std::shared_ptr<RotSet> rotate_hoisted_range(const Ct& query, int d0, int d1);   // rotations d0 .. d1 - 1 (rotation 0 = the query)
    // sum over the rotation set's diagonals of rotQ[d] (x) DB[g][d] for the named blocks: ONE batched 3-component ciphertext
    // [blocks][3][nl][N], not relinearised (sender_diag.cpp:70-77 restricted to a diagonal range)
    Ct mac_rotset(const RotSet& rs, const std::vector<int>& blocks);
    Ct diag_mac_partial(const Ct& query, int d0, int d1, const std::vector<int>& blocks);
    void fold_mod(const Ct& a);   // in place: every word (an un-reduced sum of residues below 2^63) -> its residue mod q_limb
    Ct alloc_like(int ncomp, int nl, int level, int deg, int nb) { Guard g(this); return new_ct(ncomp, nl, level, deg, nb); }
    Ct rotset_entry(const RotSet& rs, int i);                                 // one rotated query as a ciphertext (copy)
    // computeSimilarityMatrix (sender_diag.cpp:66-83) for the named blocks against one rotation set, as batches of <= 32
    std::vector<Ct> score_rotset(const RotSet& rs, const std::vector<int>& blocks);
    Ct db_entry(int block, int diag);                                         // plain-residue copy of a stored database ciphertext
    std::vector<Ct> score_diag(const Ct& query, int ndiag);            // computeSimilarity
    std::vector<Ct> score_diag_batched(const Ct& query, int ndiag);    // same, blocks kept as batches of <= 32
    Ct membership(const Ct& query, int ndiag, double delta, int depth); // membershipScenario
    std::vector<Ct> index(const Ct& query, int ndiag, double delta, int depth);   // indexScenario
    // ---- HERS (approach 4) scoring, sender_hers.cpp:13-98: query = one ciphertext per coordinate, database matrix m =
    // registered block m (index j = "diagonal" j).  One score ciphertext per matrix.
    std::vector<Ct> score_hers(const std::vector<Ct>& query, const std::vector<int>* matrices = nullptr /*null: all registered*/);
    // chebyshevCompare of many ciphertexts of one shape as batches (the OpenMP loops of sender_*.cpp over scoreCipher)
    std::vector<Ct> compare_many(const std::vector<Ct>& cts, double delta, int depth);

    // ---- building blocks exposed for parity tests
    void ntt_limbs(u64* data, int nbatch, const std::vector<int>& primes, bool inverse);
    // digits [nbatch][ndig][nl+K][N]; input polynomials c + b*c_bs
    u64* ks_precompute(const u64* c, int nl, int* ndig_out, int nbatch = 1, size_t c_bs = 0);
    // key-switch core on precomputed digits for `nbatch` keys: out[b] = [2][nl][N] (+ optional addend/maps)
    void ks_core_batch(const u64* digits, int ndig, int nl, const std::vector<const EvalKey*>& keys, const u64* addend, int addend_comps,
                       const std::vector<u64>& galois /*empty = no permutation*/, u64* out, size_t out_bs,
                       size_t digits_bs = 0 /*0: all entries share the digits*/, size_t addend_bs = 0);
    const EvalKey& rot_key(u64 galois) const;
    const u32* automap(u64 galois);
    LimbMap limb_map(const std::vector<int>& primes) const;
    ScalarTable scalar_table(const std::vector<u64>& v) const;   // v[i] mod q_i with Shoup quotients
    u64 P_mod(int i) const;     // (prod P) mod q_i

private:
    const BaseConvTable* modup_table(int nl, int digit);
    const BaseConvTable* moddown_table(int nl);
    const BaseConvTable* last_limb_table(int prime);
    BaseConvTable* make_table(const std::vector<int>& src, const std::vector<int>& dst, const std::vector<u64>* out_scale = nullptr);
    void align(Ct& a, Ct& b);
    std::vector<u64> const_for_mult(double c, int nl) const;
    std::vector<u64> const_for_add(double c, int nl, int deg) const;
    std::pair<Ct, Ct> two_products(Ct a0, Ct b0, Ct a1, Ct b1, bool twice);
    Ct inner_cheb(const std::vector<double>& coeffs, int k, int m, std::vector<Ct>& T, std::vector<Ct>& T2);
    Ct inner_poly(const std::vector<double>& coeffs, int k, int m, std::vector<Ct>& pw, std::vector<Ct>& pw2);
    // InnerEvalChebyshevPS / InnerEvalPolyPS of several coefficient vectors at the same recursion level m, evaluated level by
    // level: the independent ciphertext products of one tree level run as ONE batched mult + rescale (same operations on the
    // same operands as the recursion, hence the same bits)
    std::vector<Ct> inner_ps_many(const std::vector<std::vector<double>>& roots, int k, int m, std::vector<Ct>& basis,
                                  std::vector<Ct>& basis2, bool cheb);

    std::map<std::pair<int, int>, BaseConvTable*> modup_tabs_;
    std::map<int, BaseConvTable*> moddown_tabs_;
    std::map<int, BaseConvTable*> last_tabs_;
    std::map<u64, u32*> automaps_;
    std::map<u64, EvalKey> rot_keys_;
    EvalKey mult_key_;
    bool have_mult_key_ = false;
    std::map<int, DbBlock> db_;
    PrimeConst* d_pc_ = nullptr;
    ulonglong2 *d_twf_ = nullptr, *d_twi_ = nullptr;
    double2* d_twd_ = nullptr;
    std::vector<void*> owned_;   // device allocations released with the engine
};

}  // namespace hx
