// shim/openfhe.h -- the subset of OpenFHE v1.2.3's `lbcrypto` namespace that n7koirala/image_matching's
// sources name (every `cc->`, `Serial::`, `parameters.` and getter under src/),
// implemented over the C ABI of libhydia_b200.so (include/hydia_b200.h).  With this directory ahead of
// OpenFHE on the include path the reference's src/**/*.cpp compile UNMODIFIED and every evaluation call of the
// Sender runs as sm_100a kernels; shim/Makefile builds the reference's own `ImageMatching` that way.
//
// What maps where
//   CryptoContext<DCRTPoly>   -> hx_ctx   (+ the hx_client that stands for KeyGen / Encrypt / Decrypt)
//   Ciphertext<DCRTPoly>      -> shared_ptr to a mutable object owning one hx_ct ("...InPlace" swaps the handle
//                                inside the pointee, so aliases observe the update as they do with OpenFHE)
//   Plaintext                 -> the packed reals; encoded to RNS (hx_client_encode) when a ciphertext is multiplied by it
//   Serial::SerializeToFile / DeserializeFromFile for ciphertexts -> a path -> handle table in HBM
//       (the enrollers "write" serial/db_*/...bin, the senders "read" them inside their hot loops,
//       src/sender/sender_diag.cpp:87-91: no disk I/O, no parsing, the database stays device resident)
//   context / key serialisation -> accepted and ignored (keys live in HBM next to the context)
// Deferred execution: the DiagonalSender's call pattern -- EvalFastRotationPrecompute, 511 x
// EvalFastRotation, per block 512 x (DeserializeFromFile + EvalMultNoRelin), 511 x EvalAddInPlace, RelinearizeInPlace,
// RescaleInPlace (src/sender/sender_diag.cpp:20-33, 66-94) -- is recorded instead of executed, and when a block score is first
// read all recorded blocks run through the fused entries (hx_rot_fast_all + hx_score_blocks: batched hoisted rotations, one
// diagonal multiply-accumulate over all blocks, batched relinearise + rescale). Any other use of a recorded ciphertext
// executes its recipe call by call, so the results are the per-op results bit for bit either way.
// Errors: the C ABI's NULL / non-zero returns are re-thrown as lbcrypto::OpenFHEException, which is what the
// reference (never catching) would see from OpenFHE.  Thread safety: the library serialises callers per context,
// so the reference's "#pragma omp parallel for" loops are safe as they stand.
#pragma once

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "hydia_b200.h"
#include "hx_serial.h"

namespace lbcrypto {

class OpenFHEException : public std::runtime_error {
public:
    explicit OpenFHEException(const std::string& what) : std::runtime_error(what) {}
};

namespace shim {
inline void check(int rc, const char* where) {
    if (rc != 0) throw OpenFHEException(std::string(where) + ": " + hx_last_error());
}
template <class P>
inline P* check(P* p, const char* where) {
    if (!p) throw OpenFHEException(std::string(where) + ": " + hx_last_error());
    return p;
}
// The engine and the client that stands for the key holder; outlives every ciphertext that points at it.
struct Backend {
    hx_ctx* ctx = nullptr;
    hx_client* client = nullptr;
    int N = 0, L = 0, K = 0, dnum = 0, scale_bits = 0;
    bool sync_every_call = false;
    bool deferred = true;     // HYDIA_DEFERRED=0: every call is issued as it comes (pure per-op path)
    bool chains = true;       // HYDIA_DEFERRED=2 (default unless calls are synchronous): compare chains are recorded and batched too
    std::vector<uint64_t> pending_pkb, pending_pka;   // a deserialised public key waits here for its private key (hx_client_import)
    // HYDIA_SHIM_SERIAL_FILES=1: Serial:: and cc->Serialize...Key write and read real SerType::BINARY files (shim/hx_serial.h);
    // default: evaluation keys and database ciphertexts stay in HBM and the "files" are a path -> handle table
    static bool serial_files() {
        static const bool on = [] { const char* e = std::getenv("HYDIA_SHIM_SERIAL_FILES"); return e && std::atoi(e) != 0; }();
        return on;
    }
    // the context a ciphertext file is read into when no in-process handle is registered under its path (a fresh process that
    // deserialised serial/cryptocontext.bin): the most recently created or deserialised one
    static std::shared_ptr<Backend>& current() {
        static auto* c = new std::shared_ptr<Backend>();
        return *c;
    }
    ~Backend() {
        if (client) hx_client_destroy(client);
        if (ctx) hx_ctx_destroy(ctx);
    }
};
}  // namespace shim

// ---- enums and tags (src/main.cpp:170-179) ---------------------------------------------------------------------
enum SecurityLevel { HEStd_128_classic, HEStd_192_classic, HEStd_256_classic, HEStd_NotSet };
enum ScalingTechnique { FIXEDMANUAL, FIXEDAUTO, FLEXIBLEAUTO, FLEXIBLEAUTOEXT, NORESCALE };
enum KeySwitchTechnique { INVALID_KS_TECH, BV, HYBRID };
enum PKESchemeFeature { PKE = 0x01, KEYSWITCH = 0x02, PRE = 0x04, LEVELEDSHE = 0x08, ADVANCEDSHE = 0x10, MULTIPARTY = 0x20, FHE = 0x40 };
struct SerType {
    struct SERBINARY {};
    struct SERJSON {};
    static constexpr SERBINARY BINARY{};
    static constexpr SERJSON JSON{};
};

// Only ever named as the element type, and inside shared_ptr<vector<DCRTPoly>> (src/sender/sender_diag.cpp:22) where it
// carries the ModUp digits of EvalFastRotationPrecompute.
namespace shim {
struct Hoist;
}
class DCRTPoly {
public:
    std::shared_ptr<shim::Hoist> hoist;
};

// ---- CCParams<CryptoContextCKKSRNS> (src/main.cpp:169-173, src/openFHE_wrapper.cpp:50-58) ---------------------------
class CryptoContextCKKSRNS {};
template <class Scheme>
class CCParams;
template <>
class CCParams<CryptoContextCKKSRNS> {
public:
    void SetSecurityLevel(SecurityLevel s) { security_ = s; }
    void SetMultiplicativeDepth(uint32_t d) { depth_ = d; }
    void SetScalingModSize(uint32_t b) { scaling_bits_ = b; }
    void SetFirstModSize(uint32_t b) { first_bits_ = b; }
    void SetScalingTechnique(ScalingTechnique t) { technique_ = t; }
    void SetRingDim(uint32_t n) { ring_dim_ = n; }
    void SetBatchSize(uint32_t b) { batch_ = b; }
    void SetNumLargeDigits(uint32_t d) { num_large_digits_ = d; }
    void SetKeySwitchTechnique(KeySwitchTechnique k) { ks_ = k; }
    SecurityLevel GetSecurityLevel() const { return security_; }
    uint32_t GetMultiplicativeDepth() const { return depth_; }
    uint32_t GetScalingModSize() const { return scaling_bits_; }
    uint32_t GetFirstModSize() const { return first_bits_; }
    ScalingTechnique GetScalingTechnique() const { return technique_; }
    uint32_t GetRingDim() const { return ring_dim_; }
    uint32_t GetBatchSize() const { return batch_; }
    uint32_t GetNumLargeDigits() const { return num_large_digits_; }
    KeySwitchTechnique GetKeySwitchTechnique() const { return ks_; }
    double GetNoiseEstimate() const { return 0; }
    friend std::ostream& operator<<(std::ostream& os, const CCParams& p) {
        return os << "scheme: CKKSRNS; multiplicativeDepth: " << p.depth_ << "; scalingModSize: " << p.scaling_bits_
                  << "; firstModSize: " << p.first_bits_ << "; ringDim: " << p.ring_dim_ << "; scalTech: " << int(p.technique_)
                  << "; ksTech: HYBRID; numLargeDigits: " << p.num_large_digits_ << "; securityLevel: " << int(p.security_);
    }

private:
    SecurityLevel security_ = HEStd_128_classic;
    ScalingTechnique technique_ = FLEXIBLEAUTOEXT;
    KeySwitchTechnique ks_ = HYBRID;
    uint32_t depth_ = 1, scaling_bits_ = 50, first_bits_ = 60, ring_dim_ = 0, batch_ = 0, num_large_digits_ = 0;
};

class EncodingParamsImpl {
public:
    explicit EncodingParamsImpl(uint32_t batch = 0) : batch_(batch) {}
    uint32_t GetBatchSize() const { return batch_; }
    friend std::ostream& operator<<(std::ostream& os, const std::shared_ptr<EncodingParamsImpl>& p) {
        return p ? os << "[batch=" << p->batch_ << "]" : os << "[null]";
    }

private:
    uint32_t batch_;
};
using EncodingParams = std::shared_ptr<EncodingParamsImpl>;

// ---- keys: tokens, the material lives with the backend's client -----------------------------------------------------------
template <class Element>
struct PublicKeyImpl {
    std::shared_ptr<shim::Backend> be;
};
template <class Element>
struct PrivateKeyImpl {
    std::shared_ptr<shim::Backend> be;
};
template <class Element>
using PublicKey = std::shared_ptr<PublicKeyImpl<Element>>;
template <class Element>
using PrivateKey = std::shared_ptr<PrivateKeyImpl<Element>>;
template <class Element>
struct KeyPair {
    PublicKey<Element> publicKey;
    PrivateKey<Element> secretKey;
    bool good() const { return publicKey && secretKey; }
};

// ---- Plaintext ---------------------------------------------------------------------------------------------------------
class PlaintextImpl {
public:
    explicit PlaintextImpl(std::vector<double> v = {}) : values_(std::move(v)) {}
    const std::vector<double>& GetRealPackedValue() const { return values_; }
    std::vector<std::complex<double>> GetCKKSPackedValue() const { return {values_.begin(), values_.end()}; }
    void SetLength(size_t n) { values_.resize(n); }
    size_t GetLength() const { return values_.size(); }
    double GetLogPrecision() const { return 0; }
    friend std::ostream& operator<<(std::ostream& os, const std::shared_ptr<PlaintextImpl>& p) {
        os << "(";
        if (p)
            for (double v : p->values_) os << v << ", ";
        return os << ")";
    }
    // [L][N] evaluation-form residues of MakeCKKSPackedPlaintext, built on first use by EvalMult(ct, Plaintext)
    const std::vector<uint64_t>& rns(const shim::Backend& be) const {
        std::lock_guard<std::mutex> g(mu_);
        if (rns_.empty()) {
            rns_.resize(size_t(be.L) * be.N);
            shim::check(hx_client_encode(be.client, values_.data(), int(values_.size()), rns_.data()), "MakeCKKSPackedPlaintext");
        }
        return rns_;
    }

private:
    std::vector<double> values_;
    mutable std::vector<uint64_t> rns_;
    mutable std::mutex mu_;
};
using Plaintext = std::shared_ptr<PlaintextImpl>;
using ConstPlaintext = std::shared_ptr<const PlaintextImpl>;

// ---- the pointee of a Ciphertext: a device handle, or the recipe that produces one -------------------------------------
namespace shim {
inline std::recursive_mutex& lazy_mutex() {
    static auto* m = new std::recursive_mutex();
    return *m;
}
struct Node;
// One EvalFastRotationPrecompute: the digits, a private copy of the hoisted query, and the block scores waiting for it.
struct Hoist {
    std::shared_ptr<Backend> be;
    hx_digits* digits = nullptr;
    hx_ct* query = nullptr;
    const void* query_obj = nullptr;     // identity of the caller's ciphertext object at precompute time
    uint64_t query_version = 0;
    std::vector<std::weak_ptr<Node>> pending;
    ~Hoist() {
        hx_digits_free(digits);
        hx_ct_free(query);
    }
};
struct Node {
    enum Kind { READY, DB, ROT, PROD, SUM, SUM_RELIN, SCORE,      // DiagonalSender (approach 5)
                HPROD, HPROD_RELIN, HTERM, HSUM,                  // HersSender (approach 4)
                CHAIN };                                          // EvalChebyshevFunction / EvalPoly / + constant on one input
    struct UnaryOp {                   // one step of OpenFHEWrapper::chebyshevCompare (openFHE_wrapper.cpp:174-182)
        enum Type { CHEB, POLY, ADDC } type;
        std::vector<double> coeffs;
        double lo = 0, hi = 0, c = 0;
        bool operator==(const UnaryOp& o) const { return type == o.type && coeffs == o.coeffs && lo == o.lo && hi == o.hi && c == o.c; }
    };
    struct HersTerm {                  // Rescale(Relinearize(EvalMultNoRelin(query, database[block][diag])))
        std::shared_ptr<Node> query;
        uint64_t query_version;
        int diag;
    };
    Node(std::shared_ptr<Backend> b, hx_ct* handle) : be(std::move(b)), h(handle) {}
    Node(const Node&) = delete;
    Node& operator=(const Node&) = delete;
    virtual ~Node() { hx_ct_free(h.load()); }

    hx_ct* get() {                       // the value; a recorded recipe runs now
        hx_ct* p = h.load(std::memory_order_acquire);
        return p ? p : materialise();
    }
    void set(hx_ct* p) {                 // "...InPlace": the pointee changes, every alias sees it
        std::lock_guard<std::recursive_mutex> g(lazy_mutex());
        hx_ct_free(h.exchange(p));
        kind = READY;
        hoist.reset();
        terms.clear();
        hterms.clear();
        src.reset();
        ops.clear();
        db_block = db_diag = -1;
        db_layout = 0;
        version++;
        if (be->sync_every_call) check(hx_ctx_sync(be->ctx), "InPlace");
    }
    bool recorded() const { return h.load(std::memory_order_acquire) == nullptr; }

    std::shared_ptr<Backend> be;
    std::atomic<hx_ct*> h;
    Kind kind = READY;
    std::shared_ptr<Hoist> hoist;                  // ROT / PROD / SUM / SUM_RELIN / SCORE
    int rot = 0, block = -1, diag = -1, ndiag = 0; // ROT: rot; PROD: rot, block, diag; SUM..: block; SCORE: block, ndiag
    std::vector<std::pair<int, int>> terms;        // SUM / SUM_RELIN: (rotation, diagonal) of every product
    std::shared_ptr<Node> src;                     // CHAIN: the input, and the steps applied to it
    uint64_t src_version = 0;
    std::vector<UnaryOp> ops;
    static std::vector<std::weak_ptr<Node>>& chains() {      // every recorded chain (never destroyed, see Serial::detail::table)
        static auto* v = new std::vector<std::weak_ptr<Node>>();
        return *v;
    }
    std::vector<HersTerm> hterms;                  // HPROD..HSUM: the recorded products of one database matrix (block)
    int db_block = -1, db_diag = -1;               // database ciphertext: serial/db_diagonal/index{db_block*512+db_diag}.bin (layout 1)
    int db_layout = 0;                             //                   or serial/db_hers/matrix{db_block}/index{db_diag}.bin (layout 2)
    uint64_t version = 0;
    std::weak_ptr<Hoist> last_hoist;               // set on a query by EvalFastRotationPrecompute

private:
    hx_ct* product(int r, int d) const {
        hx_ct* a = r == 0 ? hoist->query : check(hx_rot_fast(hoist->query, r, hoist->digits), "EvalFastRotation");
        hx_ct* b = check(hx_db_get(be->ctx, block, d), "database ciphertext");
        hx_ct* p = hx_mult_norelin(a, b);
        if (r != 0) hx_ct_free(a);
        hx_ct_free(b);
        return check(p, "EvalMultNoRelin");
    }
    void score_pending() {               // every recorded block score of this hoist, through the fused entries
        std::shared_ptr<Hoist> H = hoist;
        std::vector<std::shared_ptr<Node>> nodes;
        std::vector<int> blocks;
        for (auto& w : H->pending)
            if (std::shared_ptr<Node> n = w.lock())
                if (n->kind == SCORE && n->recorded() && n->hoist == H && n->ndiag == ndiag) {
                    nodes.push_back(n);
                    blocks.push_back(n->block);
                }
        if (std::find_if(nodes.begin(), nodes.end(), [this](const std::shared_ptr<Node>& n) { return n.get() == this; }) == nodes.end())
            throw OpenFHEException("deferred score: ciphertext is not registered with its hoist");
        hx_rotset* rs = check(hx_rot_fast_all(H->query, ndiag), "EvalFastRotation (hoisted, batched)");
        std::vector<hx_ct*> out(nodes.size(), nullptr);
        int rc = hx_score_blocks(rs, blocks.data(), int(blocks.size()), out.data());
        hx_rotset_free(rs);
        check(rc, "EvalMultNoRelin / EvalAddInPlace / Relinearize / Rescale (fused)");
        for (size_t i = 0; i < nodes.size(); i++) {
            nodes[i]->h.store(out[i], std::memory_order_release);
            nodes[i]->kind = READY;
        }
        H->pending.clear();
    }
    // The compare loops of the senders ("#pragma omp parallel for" over scoreCipher, sender_diag.cpp:40-43, sender_hers.cpp:33-36)
    // apply one chain of steps to many ciphertexts of one shape. All recorded chains with the same steps run as ONE batched
    // ciphertext (hx_ct_stack, the steps, hx_ct_unstack): the same operations per entry, so the same bits.
    static hx_ct* apply(hx_ct* in, const std::vector<UnaryOp>& steps) {
        hx_ct* cur = in;
        for (const UnaryOp& op : steps) {
            hx_ct* next = op.type == UnaryOp::CHEB   ? hx_cheby_series(cur, op.coeffs.data(), int(op.coeffs.size()), op.lo, op.hi)
                          : op.type == UnaryOp::POLY ? hx_poly(cur, op.coeffs.data(), int(op.coeffs.size()))
                                                     : hx_add_const(cur, op.c);
            if (cur != in) hx_ct_free(cur);
            cur = check(next, op.type == UnaryOp::CHEB ? "EvalChebyshevFunction" : op.type == UnaryOp::POLY ? "EvalPoly" : "EvalAddInPlace");
        }
        return cur == in ? check(hx_ct_clone(in), "clone") : cur;
    }
    void chain() {
        constexpr size_t kMaxBatch = 32;
        std::vector<std::weak_ptr<Node>>& reg = chains();
        std::vector<std::shared_ptr<Node>> group;
        std::vector<std::weak_ptr<Node>> keep;
        for (auto& w : reg)
            if (std::shared_ptr<Node> n = w.lock())
                if (n->kind == CHAIN && n->recorded()) {
                    if (n.get() != this && n->be == be && n->ops == ops && group.size() + 1 < kMaxBatch) group.push_back(n);
                    else if (n.get() != this) keep.push_back(w);
                }
        reg.swap(keep);
        auto shape = [](hx_ct* c) {
            int v[4] = {0, 0, 0, 0};
            check(hx_ct_info(c, &v[0], &v[1], &v[2], &v[3]), "Ciphertext getter");
            return std::vector<int>(v, v + 4);
        };
        if (src->version != src_version) throw OpenFHEException("deferred chain: the input ciphertext changed after it was recorded");
        hx_ct* mine = src->get();
        const std::vector<int> want = shape(mine);
        std::vector<Node*> nodes{this};
        std::vector<hx_ct*> in{mine};
        std::vector<std::shared_ptr<Node>> alive;        // keeps the siblings alive while their results are produced
        for (auto& n : group) {
            if (n->src->version != n->src_version) { reg.push_back(n); continue; }
            hx_ct* x = n->src->get();
            if (shape(x) != want) { reg.push_back(n); continue; }
            nodes.push_back(n.get());
            in.push_back(x);
            alive.push_back(n);
        }
        std::vector<hx_ct*> out(nodes.size(), nullptr);
        if (nodes.size() == 1) {
            out[0] = apply(mine, ops);
        } else {
            hx_ct* batch = check(hx_ct_stack(in.data(), int(in.size())), "batch");
            hx_ct* res = apply(batch, ops);
            hx_ct_free(batch);
            int n = hx_ct_unstack(res, out.data(), int(out.size()));
            hx_ct_free(res);
            if (n != int(out.size())) throw OpenFHEException(std::string("batched chain: ") + hx_last_error());
        }
        for (size_t i = 0; i < nodes.size(); i++) {
            nodes[i]->h.store(out[i], std::memory_order_release);
            if (nodes[i] != this) {
                nodes[i]->kind = READY;
                nodes[i]->src.reset();
                nodes[i]->ops.clear();
            }
        }
    }
    // HersSender::computeSimilarityHelper (sender_hers.cpp:61-86). A complete matrix (column i of the query against
    // database column i, i = 0..n-1) goes through hx_score_hers_blocks; anything else runs term by term.
    hx_ct* hers() {
        for (const HersTerm& t : hterms)
            if (t.query->version != t.query_version) throw OpenFHEException("deferred HERS score: a query ciphertext changed after it was recorded");
        if (kind == HSUM) {
            std::vector<HersTerm> t = hterms;
            std::sort(t.begin(), t.end(), [](const HersTerm& a, const HersTerm& b) { return a.diag < b.diag; });
            bool complete = true;
            for (size_t i = 0; i < t.size() && complete; i++) complete = t[i].diag == int(i);
            if (complete) {
                std::vector<hx_ct*> q;
                for (const HersTerm& x : t) q.push_back(x.query->get());
                hx_ct* out = nullptr;
                if (hx_score_hers_blocks(be->ctx, q.data(), int(q.size()), &block, 1, &out) == 0) return out;
            }
        }
        hx_ct* acc = nullptr;
        for (const HersTerm& t : hterms) {
            hx_ct* b = check(hx_db_get(be->ctx, block, t.diag), "database ciphertext");
            hx_ct* p = hx_mult_norelin(t.query->get(), b);
            hx_ct_free(b);
            check(p, "EvalMultNoRelin");
            if (kind != HPROD) {
                hx_ct* r = hx_relin(p);
                hx_ct_free(p);
                p = check(r, "RelinearizeInPlace");
            }
            if (kind == HTERM || kind == HSUM) {
                hx_ct* r = hx_rescale(p);
                hx_ct_free(p);
                p = check(r, "RescaleInPlace");
            }
            if (!acc) {
                acc = p;
            } else {
                hx_ct* sum = hx_add(acc, p);
                hx_ct_free(acc);
                hx_ct_free(p);
                acc = check(sum, "EvalAddInPlace");
            }
        }
        return acc;
    }
    hx_ct* materialise() {
        std::lock_guard<std::recursive_mutex> g(lazy_mutex());
        if (hx_ct* p = h.load(std::memory_order_acquire)) return p;
        hx_ct* out = nullptr;
        switch (kind) {
            case DB: out = check(hx_db_get(be->ctx, db_block, db_diag), "database ciphertext"); break;
            case ROT: out = check(hx_rot_fast(hoist->query, rot, hoist->digits), "EvalFastRotation"); break;
            case PROD: out = product(rot, diag); break;
            case SUM:
            case SUM_RELIN:
                out = product(terms[0].first, terms[0].second);
                for (size_t t = 1; t < terms.size(); t++) {
                    hx_ct* p = product(terms[t].first, terms[t].second);
                    hx_ct* sum = hx_add(out, p);
                    hx_ct_free(p);
                    hx_ct_free(out);
                    out = check(sum, "EvalAddInPlace");
                }
                if (kind == SUM_RELIN) {
                    hx_ct* r = hx_relin(out);
                    hx_ct_free(out);
                    out = check(r, "RelinearizeInPlace");
                }
                break;
            case SCORE:
                score_pending();
                out = h.load(std::memory_order_acquire);
                break;
            case HPROD:
            case HPROD_RELIN:
            case HTERM:
            case HSUM: out = hers(); break;
            case CHAIN:
                chain();
                out = h.load(std::memory_order_acquire);
                break;
            default: throw OpenFHEException("ciphertext without a value");
        }
        h.store(out, std::memory_order_release);
        kind = READY;
        hoist.reset();
        terms.clear();
        hterms.clear();
        src.reset();
        ops.clear();
        return out;
    }
};
}  // namespace shim

// ---- Ciphertext --------------------------------------------------------------------------------------------------------
template <class Element>
class CiphertextImpl : public shim::Node {
public:
    CiphertextImpl(std::shared_ptr<shim::Backend> be, hx_ct* h) : shim::Node(std::move(be), h) {}
    hx_ct* handle() { return get(); }
    void reset(hx_ct* h) { set(h); }
    const std::shared_ptr<shim::Backend>& backend() const { return be; }
    // getters of src/openFHE_wrapper.cpp:65-68
    size_t GetSlots() const { return size_t(be->N / 2); }
    size_t GetLevel() { return size_t(info(2)); }
    size_t GetNoiseScaleDeg() { return size_t(info(3)); }
    double GetScalingFactor() { return std::pow(std::ldexp(1.0, be->scale_bits), double(info(3))); }
    EncodingParams GetEncodingParameters() const { return std::make_shared<EncodingParamsImpl>(uint32_t(be->N / 2)); }
    size_t NumberCiphertextElements() { return size_t(info(0)); }

private:
    int info(int which) {
        int v[4] = {0, 0, 0, 0};
        shim::check(hx_ct_info(get(), &v[0], &v[1], &v[2], &v[3]), "Ciphertext getter");
        return v[which];
    }
};
template <class Element>
using Ciphertext = std::shared_ptr<CiphertextImpl<Element>>;
template <class Element>
using ConstCiphertext = std::shared_ptr<CiphertextImpl<Element>>;

struct DecryptResult {
    bool isValid = true;
    size_t messageLength = 0;
};

// ---- CryptoContext -----------------------------------------------------------------------------------------------------
template <class Element>
class CryptoContextImpl {
    using CT = Ciphertext<Element>;

public:
    explicit CryptoContextImpl(std::shared_ptr<shim::Backend> be) : be_(std::move(be)) {}

    // called through a null CryptoContext at src/main.cpp:112-114: static, and there is no global key store to clear
    static void ClearEvalMultKeys() {}
    static void ClearEvalAutomorphismKeys() {}
    static void ClearEvalSumKeys() {}

    void Enable(PKESchemeFeature) {}
    void Enable(uint32_t) {}
    EncodingParams GetEncodingParams() const { return std::make_shared<EncodingParamsImpl>(uint32_t(be_->N / 2)); }
    uint32_t GetRingDimension() const { return uint32_t(be_->N); }
    uint32_t GetCyclotomicOrder() const { return uint32_t(2 * be_->N); }
    const std::shared_ptr<shim::Backend>& backend() const { return be_; }

    // -- client side (src/main.cpp:184-206, src/openFHE_wrapper.cpp:74-85): hx_client_*
    KeyPair<Element> KeyGen() {
        if (!be_->client) {
            const char* s = std::getenv("HYDIA_SEED");
            be_->client = shim::check(hx_client_create(be_->ctx, s ? std::strtoull(s, nullptr, 10) : 1), "KeyGen");
        }
        KeyPair<Element> kp;
        kp.publicKey = std::make_shared<PublicKeyImpl<Element>>(PublicKeyImpl<Element>{be_});
        kp.secretKey = std::make_shared<PrivateKeyImpl<Element>>(PrivateKeyImpl<Element>{be_});
        return kp;
    }
    void EvalMultKeyGen(const PrivateKey<Element>&) { shim::check(hx_client_gen_mult_key(client()), "EvalMultKeyGen"); }
    void EvalSumKeyGen(const PrivateKey<Element>&) { shim::check(hx_client_gen_sum_keys(client()), "EvalSumKeyGen"); }
    void EvalRotateKeyGen(const PrivateKey<Element>&, const std::vector<int32_t>& rotations) {
        shim::check(hx_client_gen_rot_keys(client(), rotations.data(), int(rotations.size())), "EvalRotateKeyGen");
    }
    Plaintext MakeCKKSPackedPlaintext(const std::vector<double>& values, size_t = 1, uint32_t = 0, const std::shared_ptr<void>& = nullptr,
                                      uint32_t = 0) const {
        if (values.size() > size_t(be_->N / 2)) throw OpenFHEException("MakeCKKSPackedPlaintext: more values than slots");
        return std::make_shared<PlaintextImpl>(values);
    }
    CT Encrypt(const PublicKey<Element>&, const Plaintext& pt) const {
        const std::vector<double>& v = pt->GetRealPackedValue();
        return wrap(hx_client_encrypt(client(), v.data(), int(v.size())), "Encrypt");
    }
    CT Encrypt(const Plaintext& pt, const PublicKey<Element>& pk) const { return Encrypt(pk, pt); }
    DecryptResult Decrypt(const PrivateKey<Element>&, const CT& ct, Plaintext* out) const {
        std::vector<double> v(size_t(be_->N / 2));
        shim::check(hx_client_decrypt(client(), ct->handle(), v.data()), "Decrypt");
        *out = std::make_shared<PlaintextImpl>(std::move(v));
        return DecryptResult{true, size_t(be_->N / 2)};
    }
    DecryptResult Decrypt(const CT& ct, const PrivateKey<Element>& sk, Plaintext* out) const { return Decrypt(sk, ct, out); }

    // -- evaluation (the Sender's calls); each is one C-ABI entry
    CT EvalAdd(const CT& a, const CT& b) const { return wrap(hx_add(a->handle(), b->handle()), "EvalAdd"); }
    CT EvalAdd(const CT& a, double c) const { return wrap(hx_add_const(a->handle(), c), "EvalAdd"); }
    void EvalAddInPlace(CT& a, const CT& b) const {
        {   // recorded: product (+ product)* of one hoist against one database block (sender_diag.cpp:75-77)
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            if (a->recorded() && b->recorded() && (a->kind == shim::Node::HTERM || a->kind == shim::Node::HSUM) && b->kind == shim::Node::HTERM &&
                a->block == b->block) {                       // sender_hers.cpp:75-77
                a->hterms.push_back(b->hterms[0]);
                a->kind = shim::Node::HSUM;
                a->version++;
                return;
            }
            if (a->recorded() && b->recorded() && (a->kind == shim::Node::PROD || a->kind == shim::Node::SUM) && b->kind == shim::Node::PROD &&
                a->hoist == b->hoist && a->block == b->block) {
                if (a->kind == shim::Node::PROD) {
                    a->terms.assign(1, std::make_pair(a->rot, a->diag));
                    a->kind = shim::Node::SUM;
                }
                a->terms.emplace_back(b->rot, b->diag);
                a->version++;
                return;
            }
        }
        a->reset(shim::check(hx_add(a->handle(), b->handle()), "EvalAddInPlace"));
    }
    void EvalAddInPlace(CT& a, double c) const {
        {
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            if (a->recorded() && a->kind == shim::Node::CHAIN) {
                shim::Node::UnaryOp op;
                op.type = shim::Node::UnaryOp::ADDC;
                op.c = c;
                a->ops.push_back(op);
                a->version++;
                return;
            }
        }
        a->reset(shim::check(hx_add_const(a->handle(), c), "EvalAddInPlace"));
    }
    CT EvalSub(const CT& a, const CT& b) const { return wrap(hx_sub(a->handle(), b->handle()), "EvalSub"); }
    CT EvalSub(const CT& a, double c) const { return wrap(hx_add_const(a->handle(), -c), "EvalSub"); }
    void EvalSubInPlace(CT& a, const CT& b) const { a->reset(shim::check(hx_sub(a->handle(), b->handle()), "EvalSubInPlace")); }
    CT EvalMult(const CT& a, const CT& b) const { return wrap(hx_mult(a->handle(), b->handle()), "EvalMult"); }
    CT EvalMult(const CT& a, double c) const { return wrap(hx_mult_const(a->handle(), c), "EvalMult"); }
    CT EvalMult(double c, const CT& a) const { return EvalMult(a, c); }
    CT EvalMult(const CT& a, const Plaintext& pt) const {
        return wrap(hx_mult_plain(a->handle(), pt->rns(*be_).data(), be_->L), "EvalMult(ct, Plaintext)");
    }
    CT EvalMult(const Plaintext& pt, const CT& a) const { return EvalMult(a, pt); }
    CT EvalMultNoRelin(const CT& a, const CT& b) const {
        if (be_->deferred && b->db_layout == 2 && !a->recorded()) {   // query column x database column (sender_hers.cpp:97)
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            CT out = std::make_shared<CiphertextImpl<Element>>(be_, nullptr);
            out->kind = shim::Node::HPROD;
            out->block = b->db_block;
            out->hterms.push_back(shim::Node::HersTerm{a, a->version, b->db_diag});
            return out;
        }
        if (be_->deferred && b->db_layout == 1) {   // rotation r of a hoisted query x database ciphertext (sender_diag.cpp:93)
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            std::shared_ptr<shim::Hoist> H;
            int rot = -1;
            if (a->recorded() && a->kind == shim::Node::ROT) {
                H = a->hoist;
                rot = a->rot;
            } else if (!a->recorded()) {           // rotatedQueryCipher[0] = queryCipher[0] (sender_diag.cpp:21)
                H = a->last_hoist.lock();
                if (H && H->query_obj == a.get() && H->query_version == a->version) rot = 0;
            }
            if (H && rot >= 0) {
                CT out = std::make_shared<CiphertextImpl<Element>>(be_, nullptr);
                out->kind = shim::Node::PROD;
                out->hoist = H;
                out->rot = rot;
                out->block = b->db_block;
                out->diag = b->db_diag;
                return out;
            }
        }
        return wrap(hx_mult_norelin(a->handle(), b->handle()), "EvalMultNoRelin");
    }
    CT Relinearize(const CT& a) const { return wrap(hx_relin(a->handle()), "Relinearize"); }
    void RelinearizeInPlace(CT& a) const {
        {
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            if (a->recorded() && (a->kind == shim::Node::SUM || a->kind == shim::Node::HPROD)) {
                a->kind = a->kind == shim::Node::SUM ? shim::Node::SUM_RELIN : shim::Node::HPROD_RELIN;
                a->version++;
                return;
            }
        }
        a->reset(shim::check(hx_relin(a->handle()), "RelinearizeInPlace"));
    }
    CT Rescale(const CT& a) const { return wrap(hx_rescale(a->handle()), "Rescale"); }
    void RescaleInPlace(CT& a) const {
        {   // a complete block: rotation i x diagonal i for i = 0..n-1, summed, relinearised (computeSimilarityMatrix, :66-83)
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            if (a->recorded() && a->kind == shim::Node::HPROD_RELIN) {
                a->kind = shim::Node::HTERM;
                a->version++;
                return;
            }
            if (a->recorded() && a->kind == shim::Node::SUM_RELIN) {
                std::vector<std::pair<int, int>> t = a->terms;
                std::sort(t.begin(), t.end());
                bool complete = !t.empty();
                for (size_t i = 0; i < t.size() && complete; i++) complete = t[i].first == int(i) && t[i].second == int(i);
                if (complete) {
                    a->kind = shim::Node::SCORE;
                    a->ndiag = int(t.size());
                    a->terms.clear();
                    a->version++;
                    a->hoist->pending.push_back(a);
                    return;
                }
            }
        }
        a->reset(shim::check(hx_rescale(a->handle()), "RescaleInPlace"));
    }
    CT ModReduce(const CT& a) const { return Rescale(a); }
    void ModReduceInPlace(CT& a) const { RescaleInPlace(a); }
    CT LevelReduce(const CT& a, const std::shared_ptr<void>&, size_t levels = 1) const {
        return wrap(hx_level_reduce(a->handle(), int(levels)), "LevelReduce");
    }
    void LevelReduceInPlace(CT& a, const std::shared_ptr<void>&, size_t levels = 1) const {
        a->reset(shim::check(hx_level_reduce(a->handle(), int(levels)), "LevelReduceInPlace"));
    }
    CT EvalSquare(const CT& a) const { return wrap(hx_square(a->handle()), "EvalSquare"); }
    void EvalSquareInPlace(CT& a) const { a = EvalSquare(a); }          // OpenFHE rebinds the pointer here
    CT EvalRotate(const CT& a, int32_t index) const { return wrap(hx_rotate(a->handle(), index), "EvalRotate"); }
    CT EvalAtIndex(const CT& a, int32_t index) const { return EvalRotate(a, index); }
    CT EvalSum(const CT& a, uint32_t batch) const { return wrap(hx_eval_sum(a->handle(), int(batch)), "EvalSum"); }
    CT EvalInnerProduct(const CT& a, const CT& b, uint32_t batch) const {
        return wrap(hx_inner_product(a->handle(), b->handle(), int(batch)), "EvalInnerProduct");
    }
    CT EvalAddMany(const std::vector<CT>& v) const {
        std::vector<hx_ct*> h;
        for (const CT& c : v) h.push_back(c->handle());
        return wrap(hx_add_many(h.data(), int(h.size())), "EvalAddMany");
    }
    CT EvalAddManyInPlace(std::vector<CT>& v) const {                   // the sum lands in v[0], as OpenFHE's tree leaves it
        std::vector<hx_ct*> h;
        for (const CT& c : v) h.push_back(c->handle());
        v[0]->reset(shim::check(hx_add_many(h.data(), int(h.size())), "EvalAddManyInPlace"));
        return v[0];
    }
    std::shared_ptr<std::vector<Element>> EvalFastRotationPrecompute(const CT& a) const {
        auto H = std::make_shared<shim::Hoist>();
        H->be = be_;
        H->digits = shim::check(hx_rot_precompute(a->handle()), "EvalFastRotationPrecompute");
        H->query = shim::check(hx_ct_clone(a->handle()), "EvalFastRotationPrecompute");
        {
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            H->query_obj = a.get();
            H->query_version = a->version;
            a->last_hoist = H;
        }
        auto out = std::make_shared<std::vector<Element>>(1);
        (*out)[0].hoist = H;
        return out;
    }
    CT EvalFastRotation(const CT& a, uint32_t index, uint32_t cyclotomic_order, const std::shared_ptr<std::vector<Element>>& digits) const {
        if (cyclotomic_order != uint32_t(2 * be_->N)) throw OpenFHEException("EvalFastRotation: cyclotomic order is not 2N");
        if (!digits || digits->empty() || !(*digits)[0].hoist) throw OpenFHEException("EvalFastRotation: no precomputed digits");
        const std::shared_ptr<shim::Hoist>& H = (*digits)[0].hoist;
        if (be_->deferred && index != 0) {
            std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
            if (H->query_obj == a.get() && H->query_version == a->version) {     // the hoisted ciphertext itself: record
                CT out = std::make_shared<CiphertextImpl<Element>>(be_, nullptr);
                out->kind = shim::Node::ROT;
                out->hoist = H;
                out->rot = int(index);
                return out;
            }
        }
        return wrap(hx_rot_fast(a->handle(), int(index), H->digits), "EvalFastRotation");
    }
    // EvalChebyshevCoefficients (interpolation at the Chebyshev nodes) then EvalChebyshevSeries (Paterson-Stockmeyer on the device)
    CT EvalChebyshevFunction(const std::function<double(double)>& func, const CT& a, double lo, double hi, uint32_t degree) const {
        return EvalChebyshevSeries(a, ChebyshevCoefficients(func, lo, hi, degree), lo, hi);
    }
    CT EvalChebyshevSeries(const CT& a, const std::vector<double>& coeffs, double lo, double hi) const {
        if (be_->chains) {
            shim::Node::UnaryOp op;
            op.type = shim::Node::UnaryOp::CHEB;
            op.coeffs = coeffs;
            op.lo = lo;
            op.hi = hi;
            return chain(a, op);
        }
        return wrap(hx_cheby_series(a->handle(), coeffs.data(), int(coeffs.size()), lo, hi), "EvalChebyshevSeries");
    }
    CT EvalPoly(const CT& a, const std::vector<double>& coeffs) const {
        if (be_->chains) {
            shim::Node::UnaryOp op;
            op.type = shim::Node::UnaryOp::POLY;
            op.coeffs = coeffs;
            return chain(a, op);
        }
        return wrap(hx_poly(a->handle(), coeffs.data(), int(coeffs.size())), "EvalPoly");
    }
    static std::vector<double> ChebyshevCoefficients(const std::function<double(double)>& func, double lo, double hi, uint32_t degree) {
        const size_t n = size_t(degree) + 1;
        const double half_span = 0.5 * (hi - lo), mid = 0.5 * (hi + lo), step = M_PI / double(n);
        std::vector<double> at_node(n), coeffs(n);
        for (size_t i = 0; i < n; i++) at_node[i] = func(std::cos(step * (double(i) + 0.5)) * half_span + mid);
        const double norm = 2.0 / double(n);
        for (size_t i = 0; i < n; i++) {
            double s = 0.0;
            for (size_t j = 0; j < n; j++) s += at_node[j] * std::cos(step * double(i) * (double(j) + 0.5));
            coeffs[i] = s * norm;
        }
        return coeffs;
    }

    // -- key (de)serialisation (src/main.cpp:124-165, 251-285): the keys never leave HBM
    // With HYDIA_SHIM_SERIAL_FILES=1 the streams carry the keys (shim/hx_serial.h: cereal BINARY conventions, layout unverified
    // against upstream); otherwise a marker line, because the keys stay resident.
    template <class ST> bool SerializeEvalMultKey(std::ostream& os, const ST&, const std::string& = "") const {
        if (!shim::Backend::serial_files()) return marker(os);
        try { return hxser::write_mult_key_stream(os, be_->ctx); } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; }
    }
    template <class ST> bool SerializeEvalSumKey(std::ostream& os, const ST&, const std::string& = "") const {
        if (!shim::Backend::serial_files()) return marker(os);
        const std::vector<uint32_t> sum = hxser::eval_sum_galois(be_->N);
        try { return hxser::write_auto_key_stream(os, be_->ctx, &sum); } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; }
    }
    template <class ST> bool SerializeEvalAutomorphismKey(std::ostream& os, const ST&, const std::string& = "") const {
        if (!shim::Backend::serial_files()) return marker(os);
        try { return hxser::write_auto_key_stream(os, be_->ctx, nullptr); } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; }
    }
    template <class ST> bool DeserializeEvalMultKey(std::istream& is, const ST&) const {
        if (!shim::Backend::serial_files()) return is.good();
        try { return hxser::read_mult_key_stream(is, be_->ctx); } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; }
    }
    template <class ST> bool DeserializeEvalSumKey(std::istream& is, const ST&) const {
        if (!shim::Backend::serial_files()) return is.good();
        try { return hxser::read_auto_key_stream(is, be_->ctx); } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; }
    }
    template <class ST> bool DeserializeEvalAutomorphismKey(std::istream& is, const ST&) const {
        if (!shim::Backend::serial_files()) return is.good();
        try { return hxser::read_auto_key_stream(is, be_->ctx); } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return false; }
    }

private:
    // a new recorded ciphertext = `a` followed by one more step (a's own steps are copied when `a` is still a recorded chain)
    CT chain(const CT& a, const shim::Node::UnaryOp& op) const {
        std::lock_guard<std::recursive_mutex> g(shim::lazy_mutex());
        CT out = std::make_shared<CiphertextImpl<Element>>(be_, nullptr);
        out->kind = shim::Node::CHAIN;
        if (a->recorded() && a->kind == shim::Node::CHAIN) {
            out->src = a->src;
            out->src_version = a->src_version;
            out->ops = a->ops;
        } else {
            out->src = a;
            out->src_version = a->version;
        }
        out->ops.push_back(op);
        shim::Node::chains().push_back(out);
        return out;
    }
    static bool marker(std::ostream& os) {
        os << "hydia-b200: evaluation keys are resident in HBM\n";
        return os.good();
    }
    hx_client* client() const {
        if (!be_->client) throw OpenFHEException("no key pair: call KeyGen first");
        return be_->client;
    }
    // Device work is queued asynchronously and completes in order; a result is waited for when it is read (Decrypt). With
    // HYDIA_SYNC_EVERY_CALL=1 every evaluation call returns only when its result is complete, as OpenFHE's do, so that the
    // reference's wall-clock timers around sender->...Scenario() (src/main.cpp:342-347, 361-366) see the device time.
    CT wrap(hx_ct* h, const char* where) const {
        CT out = std::make_shared<CiphertextImpl<Element>>(be_, shim::check(h, where));
        if (be_->sync_every_call) shim::check(hx_ctx_sync(be_->ctx), where);
        return out;
    }
    std::shared_ptr<shim::Backend> be_;
};
template <class Element>
using CryptoContext = std::shared_ptr<CryptoContextImpl<Element>>;

template <class Element>
class CryptoContextFactory {
public:
    static void ReleaseAllContexts() {}
};

// GenCryptoContext(HEStd_128_classic, depth, ScalingModSize, FIXEDMANUAL) (src/main.cpp:169-175): the parameter recipe is
// hx_params_generate; HYDIA_DEVICE picks the GPU.
inline CryptoContext<DCRTPoly> GenCryptoContext(const CCParams<CryptoContextCKKSRNS>& p) {
    if (p.GetScalingTechnique() != FIXEDMANUAL) throw OpenFHEException("GenCryptoContext: only FIXEDMANUAL is implemented");
    if (p.GetKeySwitchTechnique() != HYBRID) throw OpenFHEException("GenCryptoContext: only HYBRID key switching is implemented");
    auto be = std::make_shared<shim::Backend>();
    uint64_t q[40], pp[40], psi_q[40], psi_p[40];
    shim::check(hx_params_generate(int(p.GetMultiplicativeDepth()), int(p.GetScalingModSize()), int(p.GetFirstModSize()), int(p.GetRingDim()),
                                   int(p.GetNumLargeDigits()), &be->N, &be->L, &be->K, &be->dnum, q, pp, psi_q, psi_p),
                "GenCryptoContext");
    be->scale_bits = int(p.GetScalingModSize());
    const char* dev = std::getenv("HYDIA_DEVICE");
    be->ctx = shim::check(hx_ctx_create(be->N, be->L, be->K, be->dnum, q, pp, psi_q, psi_p, be->scale_bits, dev ? std::atoi(dev) : 0), "GenCryptoContext");
    const char* sync = std::getenv("HYDIA_SYNC_EVERY_CALL");
    be->sync_every_call = sync && std::atoi(sync) != 0;
    const char* deferred = std::getenv("HYDIA_DEFERRED");
    be->deferred = !deferred || std::atoi(deferred) != 0;
    // A recorded compare chain may leave the sender's scenario function unevaluated (indexScenario returns its outputs) and
    // complete when the receiver reads it; with synchronous calls that would move work out of the reference's timers, so
    // chains are then recorded only on request (HYDIA_DEFERRED=2).
    be->chains = be->deferred && (deferred ? std::atoi(deferred) >= 2 : !be->sync_every_call);
    shim::Backend::current() = be;
    return std::make_shared<CryptoContextImpl<DCRTPoly>>(be);
}
namespace shim {
// the same context from serialised parameters (Serial::DeserializeFromFile of serial/cryptocontext.bin, src/main.cpp:124)
inline std::shared_ptr<Backend> backend_from_params(const hxser::Params& P) {
    auto be = std::make_shared<Backend>();
    be->N = P.N; be->L = P.L; be->K = P.K; be->dnum = P.dnum; be->scale_bits = P.scale_bits;
    const char* dev = std::getenv("HYDIA_DEVICE");
    be->ctx = check(hx_ctx_create(P.N, P.L, P.K, P.dnum, P.q.data(), P.p.data(), P.psi_q.data(), P.psi_p.data(), P.scale_bits, dev ? std::atoi(dev) : 0),
                    "DeserializeFromFile(CryptoContext)");
    const char* sync = std::getenv("HYDIA_SYNC_EVERY_CALL");
    be->sync_every_call = sync && std::atoi(sync) != 0;
    const char* deferred = std::getenv("HYDIA_DEFERRED");
    be->deferred = !deferred || std::atoi(deferred) != 0;
    be->chains = be->deferred && (deferred ? std::atoi(deferred) >= 2 : !be->sync_every_call);
    Backend::current() = be;
    return be;
}
}  // namespace shim

// ---- Serial ------------------------------------------------------------------------------------------------------------
namespace Serial {
namespace detail {
// path -> ciphertext. Deliberately never destroyed: handles must not be released after the CUDA runtime has shut down.
inline std::map<std::string, Ciphertext<DCRTPoly>>& table() {
    static auto* t = new std::map<std::string, Ciphertext<DCRTPoly>>();
    return *t;
}
inline std::mutex& table_mutex() {
    static auto* m = new std::mutex();
    return *m;
}
// the ciphertext as a real SerType::BINARY file (HYDIA_SHIM_SERIAL_FILES=1), as the reference's enrollers produce
inline bool write_ct_file(const std::string& path, const Ciphertext<DCRTPoly>& ct) {
    if (!shim::Backend::serial_files()) return true;
    try {
        hxser::Session ses;
        ses.P = hxser::params_of(ct->backend()->ctx);
        ses.have_params = true;
        hxser::HostCt h;
        shim::check(hx_ct_info(ct->handle(), &h.ncomp, &h.nl, &h.level, &h.deg), "SerializeToFile");
        h.data.resize((size_t)h.ncomp * h.nl * ses.P.N);
        shim::check(hx_ct_download(ct->handle(), h.data.data()), "SerializeToFile");
        std::ofstream os(path, std::ios::out | std::ios::binary);
        if (!os.is_open()) return false;
        hxser::Writer w(os);
        hxser::write_ciphertext(w, ses, h);
        return w.good();
    } catch (const std::exception& e) {
        std::cerr << "SerializeToFile(" << path << "): " << e.what() << std::endl;
        return false;
    }
}
}  // namespace detail

// Enrollers (src/enroller/enroller_diag.cpp:158-164, enroller_hers.cpp:116-121, ...): the ciphertext becomes database entry
// `path`, resident in HBM. A private copy is kept so later "...InPlace" calls on the caller's object cannot touch it.
inline bool SerializeToFileResident(const std::string& path, const Ciphertext<DCRTPoly>& ct);
template <class ST>
bool SerializeToFile(const std::string& path, const Ciphertext<DCRTPoly>& ct, const ST&) {
    if (!ct) return false;
    if (!SerializeToFileResident(path, ct)) return false;
    return detail::write_ct_file(path, ct);
}
inline bool SerializeToFileResident(const std::string& path, const Ciphertext<DCRTPoly>& ct) {
    std::shared_ptr<CiphertextImpl<DCRTPoly>> entry;
    int index = -1, matrix = -1, consumed = 0;
    // serial/db_diagonal/index{block*512+diag}.bin (enroller_diag.cpp:159): goes into the engine's database store, which is what
    // the fused scoring reads; the table entry is a recipe for a plain copy, used only if a caller touches the ciphertext itself
    if (std::sscanf(path.c_str(), "serial/db_diagonal/index%d.bin%n", &index, &consumed) == 1 && size_t(consumed) == path.size() && index >= 0 &&
        hx_db_register_ct(ct->backend()->ctx, index / 512, index % 512, ct->handle()) == 0) {
        entry = std::make_shared<CiphertextImpl<DCRTPoly>>(ct->backend(), nullptr);
        entry->kind = shim::Node::DB;
        entry->db_block = index / 512;
        entry->db_diag = index % 512;
        entry->db_layout = 1;
    } else if (std::sscanf(path.c_str(), "serial/db_hers/matrix%d/index%d.bin%n", &matrix, &index, &consumed) == 2 &&
               size_t(consumed) == path.size() && matrix >= 0 && index >= 0 && index < 512 &&
               hx_db_register_ct(ct->backend()->ctx, matrix, index, ct->handle()) == 0) {     // enroller_hers.cpp:116-121
        entry = std::make_shared<CiphertextImpl<DCRTPoly>>(ct->backend(), nullptr);
        entry->kind = shim::Node::DB;
        entry->db_block = matrix;
        entry->db_diag = index;
        entry->db_layout = 2;
    } else {
        hx_ct* copy = hx_ct_clone(ct->handle());
        if (!copy) return false;
        entry = std::make_shared<CiphertextImpl<DCRTPoly>>(ct->backend(), copy);
    }
    {
        std::lock_guard<std::mutex> g(detail::table_mutex());
        detail::table()[path] = entry;
    }
    return true;
}
// Senders (src/sender/sender_diag.cpp:87-91 and the like): the handle registered under `path`; false when nothing was enrolled
// there, which the reference reports on cerr and carries on from.
template <class ST>
bool DeserializeFromFile(const std::string& path, Ciphertext<DCRTPoly>& ct, const ST&) {
    {
        std::lock_guard<std::mutex> g(detail::table_mutex());
        auto it = detail::table().find(path);
        if (it != detail::table().end()) { ct = it->second; return true; }
    }
    // no resident handle under this path: read the file into the current context (a process that started from serial/*.bin) and
    // keep it resident from then on -- the reference re-reads it on every query, the arithmetic that follows is the same
    std::shared_ptr<shim::Backend> be = shim::Backend::current();
    if (!shim::Backend::serial_files() || !be) return false;
    try {
        std::ifstream is(path, std::ios::in | std::ios::binary);
        if (!is.is_open()) return false;
        hxser::Session ses;
        ses.P = hxser::params_of(be->ctx);
        ses.have_params = true;
        hxser::Reader r(is);
        hxser::HostCt h = hxser::read_ciphertext(r, ses);
        Ciphertext<DCRTPoly> fresh = std::make_shared<CiphertextImpl<DCRTPoly>>(
            be, shim::check(hx_ct_upload(be->ctx, h.ncomp, h.nl, h.data.data(), h.level, h.deg), "DeserializeFromFile"));
        if (!SerializeToFileResident(path, fresh)) return false;
        std::lock_guard<std::mutex> g(detail::table_mutex());
        ct = detail::table()[path];
        return true;
    } catch (const std::exception& e) {
        std::cerr << "DeserializeFromFile(" << path << "): " << e.what() << std::endl;
        return false;
    }
}
// Context and key pair (src/main.cpp:122-134, 246-249).  Default: nothing to move, the objects are process resident, and reading
// them back fails as a missing file would.  HYDIA_SHIM_SERIAL_FILES=1: real files (shim/hx_serial.h) -- the context as its
// parameter block, the public key as CryptoObject + (b, a), the private key as CryptoObject + its N coefficients in {-1,0,1}
// (one byte each; upstream stores the key as a DCRTPoly: DEVIATION).
template <class ST> bool SerializeToFile(const std::string& path, const CryptoContext<DCRTPoly>& cc, const ST&) {
    if (!cc || !shim::Backend::serial_files()) return bool(cc);
    try {
        hxser::Session ses; ses.P = hxser::params_of(cc->backend()->ctx); ses.have_params = true;
        std::ofstream os(path, std::ios::out | std::ios::binary);
        hxser::Writer w(os);
        hxser::write_context(w, ses);
        return os.is_open() && w.good();
    } catch (const std::exception& e) { std::cerr << "SerializeToFile(" << path << "): " << e.what() << std::endl; return false; }
}
template <class ST> bool SerializeToFile(const std::string& path, const PublicKey<DCRTPoly>& k, const ST&) {
    if (!k || !shim::Backend::serial_files()) return bool(k);
    try {
        hxser::Session ses; ses.P = hxser::params_of(k->be->ctx); ses.have_params = true;
        const size_t words = (size_t)ses.P.L * ses.P.N;
        std::vector<uint64_t> ba(2 * words);
        shim::check(hx_client_export_public(k->be->client, ba.data(), ba.data() + words), "SerializeToFile(PublicKey)");
        std::ofstream os(path, std::ios::out | std::ios::binary);
        hxser::Writer w(os);
        w.shared_fresh();
        w.version(hxser::C_PUBKEY, hxser::kVerKey);
        hxser::write_object_base(w, ses);
        w.size(2);
        hxser::write_dcrt(w, ses, hxser::first_limbs(ses.P.L), ba.data());
        hxser::write_dcrt(w, ses, hxser::first_limbs(ses.P.L), ba.data() + words);
        return os.is_open() && w.good();
    } catch (const std::exception& e) { std::cerr << "SerializeToFile(" << path << "): " << e.what() << std::endl; return false; }
}
template <class ST> bool SerializeToFile(const std::string& path, const PrivateKey<DCRTPoly>& k, const ST&) {
    if (!k || !shim::Backend::serial_files()) return bool(k);
    try {
        hxser::Session ses; ses.P = hxser::params_of(k->be->ctx); ses.have_params = true;
        std::vector<int8_t> sk((size_t)ses.P.N);
        shim::check(hx_client_secret(k->be->client, sk.data()), "SerializeToFile(PrivateKey)");
        std::ofstream os(path, std::ios::out | std::ios::binary);
        hxser::Writer w(os);
        w.shared_fresh();
        w.version(hxser::C_PRIVKEY, hxser::kVerKey);
        hxser::write_object_base(w, ses);
        w.size(sk.size());
        w.raw(sk.data(), sk.size());
        return os.is_open() && w.good();
    } catch (const std::exception& e) { std::cerr << "SerializeToFile(" << path << "): " << e.what() << std::endl; return false; }
}
template <class ST> bool DeserializeFromFile(const std::string& path, CryptoContext<DCRTPoly>& cc, const ST&) {
    if (!shim::Backend::serial_files()) return false;
    try {
        std::ifstream is(path, std::ios::in | std::ios::binary);
        if (!is.is_open()) return false;
        hxser::Session ses;
        hxser::Reader r(is);
        hxser::read_context(r, ses);
        cc = std::make_shared<CryptoContextImpl<DCRTPoly>>(shim::backend_from_params(ses.P));
        return true;
    } catch (const std::exception& e) { std::cerr << "DeserializeFromFile(" << path << "): " << e.what() << std::endl; return false; }
}
template <class ST> bool DeserializeFromFile(const std::string& path, PublicKey<DCRTPoly>& k, const ST&) {
    std::shared_ptr<shim::Backend> be = shim::Backend::current();
    if (!shim::Backend::serial_files() || !be) return false;
    try {
        std::ifstream is(path, std::ios::in | std::ios::binary);
        if (!is.is_open()) return false;
        hxser::Session ses; ses.P = hxser::params_of(be->ctx); ses.have_params = true;
        hxser::Reader r(is);
        bool fresh;
        r.shared(&fresh);
        r.version(hxser::C_PUBKEY, hxser::kVerKey);
        hxser::read_object_base(r, ses);
        if (r.size(2) != 2) return false;
        std::vector<int> limbs;
        be->pending_pkb.clear(); be->pending_pka.clear();
        hxser::read_dcrt(r, ses, limbs, be->pending_pkb);
        hxser::read_dcrt(r, ses, limbs, be->pending_pka);
        k = std::make_shared<PublicKeyImpl<DCRTPoly>>(PublicKeyImpl<DCRTPoly>{be});
        return (int)limbs.size() == ses.P.L;
    } catch (const std::exception& e) { std::cerr << "DeserializeFromFile(" << path << "): " << e.what() << std::endl; return false; }
}
template <class ST> bool DeserializeFromFile(const std::string& path, PrivateKey<DCRTPoly>& k, const ST&) {
    std::shared_ptr<shim::Backend> be = shim::Backend::current();
    if (!shim::Backend::serial_files() || !be) return false;
    try {
        std::ifstream is(path, std::ios::in | std::ios::binary);
        if (!is.is_open()) return false;
        hxser::Session ses; ses.P = hxser::params_of(be->ctx); ses.have_params = true;
        hxser::Reader r(is);
        bool fresh;
        r.shared(&fresh);
        r.version(hxser::C_PRIVKEY, hxser::kVerKey);
        hxser::read_object_base(r, ses);
        std::vector<int8_t> sk(r.size(1 << 20));
        r.raw(sk.data(), sk.size());
        if ((int)sk.size() != ses.P.N || be->pending_pkb.empty()) return false;   // the reference reads the public key first (main.cpp:128-134)
        if (be->client) { hx_client_destroy(be->client); be->client = nullptr; }
        const char* seed = std::getenv("HYDIA_SEED");
        be->client = shim::check(hx_client_import(be->ctx, seed ? std::strtoull(seed, nullptr, 10) : 1, sk.data(), be->pending_pkb.data(), be->pending_pka.data()),
                                 "DeserializeFromFile(PrivateKey)");
        k = std::make_shared<PrivateKeyImpl<DCRTPoly>>(PrivateKeyImpl<DCRTPoly>{be});
        return true;
    } catch (const std::exception& e) { std::cerr << "DeserializeFromFile(" << path << "): " << e.what() << std::endl; return false; }
}
}  // namespace Serial

// OpenFHE prints vectors as "[ a b c ]" (src/main.cpp:388-389 streams the index results this way).
template <class T>
std::ostream& operator<<(std::ostream& os, const std::vector<T>& v) {
    os << "[";
    for (const T& x : v) os << " " << x;
    return os << " ]";
}

}  // namespace lbcrypto
