// shim/hx_serial.h -- SerType::BINARY files for the objects the reference serialises:
//   Ciphertext<DCRTPoly>                         src/enroller/*.cpp (serial/db_*/...bin), src/sender/sender_diag.cpp:87-91
//   CryptoContext, PublicKey, PrivateKey         src/main.cpp:124-134, 251-253
//   EvalMult / EvalAutomorphism / EvalSum keys   src/main.cpp:136-165, 255-285
//
// OpenFHE writes these with cereal's BinaryOutputArchive.  That archive has no header and no field names: a file is the raw
// little-endian bytes of the fields in the order the class's save() lists them, with cereal's conventions around them.  This
// header follows those CONVENTIONS exactly:
//   * arithmetic types: raw little-endian bytes of their size; bool = 1 byte; enums as their underlying 32-bit type
//   * std::string / std::vector / std::map: a 64-bit element count, then the elements (map: key, value)
//   * std::shared_ptr<T>, T polymorphic and of its static type: u32 0x40000000, then the pointer wrapper below
//   * pointer wrapper: u32 id; the first time an object is written the id has bit 31 set and the object's data follows, later
//     occurrences are the bare id (ILNativeParams of a prime, the CryptoContext of every CryptoObject)
//   * std::unique_ptr<T>: u8 valid, then the data
//   * a class with CEREAL_CLASS_VERSION: u32 version before the FIRST object of that class in the archive
// and, for the data-bearing classes, the FIELD ORDER of OpenFHE v1.2.3 as far as it is known here:
//   NativeVector   m (modulus, u64), l (length, u64), raw 8-byte words
//   ILNativeParams co (cyclotomic order, u32), rd (ring dimension, u32), 2n (u8), cm, ru, bm, br (u64 each)
//   NativePoly     v (unique_ptr<NativeVector>), f (format, u32: EVALUATION = 0), p (shared_ptr<ILNativeParams>)
//   ILDCRTParams   ElemParams fields with cm = 0 (the composite modulus does not fit a word; written as a vector of its limbs
//                  instead: DEVIATION, upstream writes a BigInteger), then p (vector<shared_ptr<ILNativeParams>>)
//   DCRTPoly       v (vector<NativePoly>), f (u32), p (shared_ptr<ILDCRTParams>)
//   CryptoObject   cc (shared_ptr<CryptoContext>), kt (key tag, string)
//   Ciphertext     CryptoObject, v (vector<DCRTPoly>), d (noiseScaleDeg, u32), l (level, u32), h (hops level, u32),
//                  m (metadata map, empty), s (scaling factor, f64), e (encoding type, u32: CKKS_PACKED_ENCODING = 3),
//                  si (scaling factor int, u64), sl (slots, u32)
//   EvalKeyRelin   CryptoObject, ak (vector<DCRTPoly>), bk (vector<DCRTPoly>)
//   PublicKey      CryptoObject, h (vector<DCRTPoly>: b, a)        PrivateKey   CryptoObject, s (DCRTPoly)
//   key files      SerializeEvalMultKey: map<keyTag, vector<EvalKey>>; SerializeEvalAutomorphismKey / SerializeEvalSumKey:
//                  map<keyTag, shared_ptr<map<u32 galois element, EvalKey>>>
// STATUS: byte layout UNVERIFIED against upstream -- no OpenFHE install and no file written by one exists in this environment.
// The CryptoContext block in particular is this repo's own field list (ring, limb moduli, roots, digits, scale, batch size) in
// cereal conventions: upstream's CryptoParametersCKKSRNS / scheme / encoding-parameter blocks are NOT reproduced.  What is
// checked (tests/cpp/shim_serial_roundtrip.cpp, tests/test_reference_unmodified.py): write -> read round trips are bit-identical
// for ciphertexts, evaluation keys and the context + key pair, and the reference's own main, rebuilt with
// READ_FROM_SERIAL = true (include/config.h:27), answers test/2_10.dat from the files a first run wrote.
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <istream>
#include <map>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "hydia_b200.h"

namespace lbcrypto {
namespace hxser {

constexpr uint32_t kMsb = 0x80000000u, kMsb2 = 0x40000000u;
constexpr uint32_t kVerVector = 1, kVerParams = 1, kVerPoly = 1, kVerDcrt = 1, kVerObject = 1, kVerCiphertext = 1, kVerEvalKey = 1, kVerKey = 1,
                   kVerContext = 1;
constexpr uint32_t kFormatEvaluation = 0, kEncodingCkksPacked = 3;

struct Params {   // what a context needs to be rebuilt (hx_ctx_create)
    int N = 0, L = 0, K = 0, dnum = 0, scale_bits = 0;
    std::vector<uint64_t> q, p, psi_q, psi_p;
    bool operator==(const Params& o) const {
        return N == o.N && L == o.L && K == o.K && dnum == o.dnum && scale_bits == o.scale_bits && q == o.q && p == o.p && psi_q == o.psi_q && psi_p == o.psi_p;
    }
    uint64_t mod(int prime) const { return prime < L ? q[prime] : p[prime - L]; }
    uint64_t root(int prime) const { return prime < L ? psi_q[prime] : psi_p[prime - L]; }
};
inline Params params_of(hx_ctx* ctx) {
    Params P;
    uint64_t q[64], p[64], sq[64], sp[64];
    if (hx_ctx_get_params(ctx, &P.N, &P.L, &P.K, &P.dnum, &P.scale_bits, q, p, sq, sp) != 0) throw std::runtime_error(hx_last_error());
    P.q.assign(q, q + P.L); P.psi_q.assign(sq, sq + P.L);
    P.p.assign(p, p + P.K); P.psi_p.assign(sp, sp + P.K);
    return P;
}

// ---- archive -----------------------------------------------------------------------------------------------------------
class Writer {
public:
    explicit Writer(std::ostream& os) : os_(os) {}
    template <class T> void put(T v) { os_.write(reinterpret_cast<const char*>(&v), sizeof(T)); }
    void size(uint64_t n) { put<uint64_t>(n); }
    void str(const std::string& s) { size(s.size()); os_.write(s.data(), (std::streamsize)s.size()); }
    void raw(const void* p, size_t bytes) { os_.write(reinterpret_cast<const char*>(p), (std::streamsize)bytes); }
    // class version: once per class per archive
    void version(int cls, uint32_t v) { if (!(seen_ & (1u << cls))) { seen_ |= 1u << cls; put<uint32_t>(v); } }
    // shared pointer of a polymorphic class of its static type; returns true when the object's data must follow
    bool shared(const void* key) {
        put<uint32_t>(kMsb2);
        auto it = ids_.find(key);
        if (it != ids_.end()) { put<uint32_t>(it->second); return false; }
        const uint32_t id = next_++;
        ids_[key] = id;
        put<uint32_t>(id | kMsb);
        return true;
    }
    // a pointer to an object that occurs once (a ciphertext, an evaluation key): always a fresh id, the data follows
    void shared_fresh() { put<uint32_t>(kMsb2); put<uint32_t>((next_++) | kMsb); }
    bool good() const { return os_.good(); }

private:
    std::ostream& os_;
    uint32_t seen_ = 0, next_ = 1;
    std::map<const void*, uint32_t> ids_;
};
class Reader {
public:
    explicit Reader(std::istream& is) : is_(is) {}
    template <class T> T get() {
        T v;
        is_.read(reinterpret_cast<char*>(&v), sizeof(T));
        if (!is_) throw std::runtime_error("hx_serial: unexpected end of file");
        return v;
    }
    uint64_t size(uint64_t limit) {
        const uint64_t n = get<uint64_t>();
        if (n > limit) throw std::runtime_error("hx_serial: implausible element count (not a file of this layout)");
        return n;
    }
    std::string str() { std::string s(size(1 << 20), '\0'); raw(&s[0], s.size()); return s; }
    void raw(void* p, size_t bytes) {
        is_.read(reinterpret_cast<char*>(p), (std::streamsize)bytes);
        if (!is_) throw std::runtime_error("hx_serial: unexpected end of file");
    }
    void version(int cls, uint32_t want) {
        if (seen_ & (1u << cls)) return;
        seen_ |= 1u << cls;
        if (get<uint32_t>() != want) throw std::runtime_error("hx_serial: class version mismatch");
    }
    // returns the id; *fresh = the object's data follows
    uint32_t shared(bool* fresh) {
        if (get<uint32_t>() != kMsb2) throw std::runtime_error("hx_serial: polymorphic pointer marker expected");
        const uint32_t id = get<uint32_t>();
        *fresh = (id & kMsb) != 0;
        return id & ~kMsb;
    }

private:
    std::istream& is_;
    uint32_t seen_ = 0;
};
enum Cls { C_VECTOR = 0, C_NPARAMS, C_POLY, C_DPARAMS, C_DCRT, C_OBJECT, C_CIPHERTEXT, C_EVALKEY, C_PUBKEY, C_PRIVKEY, C_CONTEXT };

// ---- objects -----------------------------------------------------------------------------------------------------------
struct HostCt {        // a ciphertext on the host: [ncomp][nl][N] residues, EVALUATION format
    int ncomp = 0, nl = 0, level = 0, deg = 1;
    std::vector<uint64_t> data;
};
struct HostKey {       // an evaluation key: a, b as [dnum][L+K][N] plain residues
    std::vector<uint64_t> a, b;
};
struct Session {       // per-archive bookkeeping of the shared objects (ids are per file, as in cereal)
    Params P;
    bool have_params = false;
    int ctx_key = 0, dparams_key[64] = {}, nparams_key[64] = {};       // addresses serve as pointer identities for the Writer
    std::map<uint32_t, int> nparams_id;                                // Reader: id -> prime index
    std::map<uint32_t, std::vector<int>> dparams_id;                   // Reader: id -> limb set
    uint32_t ctx_id = 0;
};

inline void write_context(Writer& w, Session& s) {
    if (!w.shared(&s.ctx_key)) return;
    const Params& P = s.P;
    w.version(C_CONTEXT, kVerContext);
    w.put<uint32_t>((uint32_t)P.N); w.put<uint32_t>((uint32_t)P.L); w.put<uint32_t>((uint32_t)P.K); w.put<uint32_t>((uint32_t)P.dnum);
    w.put<uint32_t>((uint32_t)P.scale_bits); w.put<uint32_t>((uint32_t)(P.N / 2));   // batch size
    for (const std::vector<uint64_t>* v : {&P.q, &P.p, &P.psi_q, &P.psi_p}) {
        w.size(v->size());
        w.raw(v->data(), v->size() * 8);
    }
}
inline void read_context(Reader& r, Session& s) {
    bool fresh;
    const uint32_t id = r.shared(&fresh);
    if (!fresh) {
        if (!s.have_params || id != s.ctx_id) throw std::runtime_error("hx_serial: reference to an unknown CryptoContext");
        return;
    }
    Params P;
    r.version(C_CONTEXT, kVerContext);
    P.N = (int)r.get<uint32_t>(); P.L = (int)r.get<uint32_t>(); P.K = (int)r.get<uint32_t>(); P.dnum = (int)r.get<uint32_t>();
    P.scale_bits = (int)r.get<uint32_t>();
    if (r.get<uint32_t>() != (uint32_t)(P.N / 2)) throw std::runtime_error("hx_serial: batch size is not N/2");
    for (std::vector<uint64_t>* v : {&P.q, &P.p, &P.psi_q, &P.psi_p}) {
        v->resize(r.size(64));
        r.raw(v->data(), v->size() * 8);
    }
    if ((int)P.q.size() != P.L || (int)P.p.size() != P.K || (int)P.psi_q.size() != P.L || (int)P.psi_p.size() != P.K)
        throw std::runtime_error("hx_serial: inconsistent CryptoContext block");
    if (s.have_params && !(s.P == P)) throw std::runtime_error("hx_serial: the file belongs to a different CryptoContext");
    s.P = P; s.have_params = true; s.ctx_id = id;
}
inline void write_object_base(Writer& w, Session& s) {     // CryptoObject
    w.version(C_OBJECT, kVerObject);
    write_context(w, s);
    w.str("");                                             // key tag
}
inline void read_object_base(Reader& r, Session& s) {
    r.version(C_OBJECT, kVerObject);
    read_context(r, s);
    r.str();
}
inline void write_nparams(Writer& w, Session& s, int prime) {
    if (!w.shared(&s.nparams_key[prime])) return;
    w.version(C_NPARAMS, kVerParams);
    w.put<uint32_t>((uint32_t)(2 * s.P.N)); w.put<uint32_t>((uint32_t)s.P.N); w.put<uint8_t>(1);
    w.put<uint64_t>(s.P.mod(prime)); w.put<uint64_t>(s.P.root(prime)); w.put<uint64_t>(0); w.put<uint64_t>(0);
}
inline int read_nparams(Reader& r, Session& s) {
    bool fresh;
    const uint32_t id = r.shared(&fresh);
    if (!fresh) {
        auto it = s.nparams_id.find(id);
        if (it == s.nparams_id.end()) throw std::runtime_error("hx_serial: reference to unknown ILNativeParams");
        return it->second;
    }
    r.version(C_NPARAMS, kVerParams);
    const uint32_t co = r.get<uint32_t>(), rd = r.get<uint32_t>();
    r.get<uint8_t>();
    const uint64_t cm = r.get<uint64_t>(), ru = r.get<uint64_t>();
    r.get<uint64_t>(); r.get<uint64_t>();
    if (!s.have_params || (int)rd != s.P.N || co != 2 * rd) throw std::runtime_error("hx_serial: ring dimension mismatch");
    for (int i = 0; i < s.P.L + s.P.K; i++)
        if (s.P.mod(i) == cm && s.P.root(i) == ru) { s.nparams_id[id] = i; return i; }
    throw std::runtime_error("hx_serial: a limb modulus / root of unity that the CryptoContext does not have");
}
// DCRTPoly over the limbs `primes`; data = [primes.size()][N]
inline void write_dcrt(Writer& w, Session& s, const std::vector<int>& primes, const uint64_t* data) {
    const size_t N = (size_t)s.P.N;
    w.version(C_DCRT, kVerDcrt);
    w.size(primes.size());
    for (size_t i = 0; i < primes.size(); i++) {           // NativePoly
        w.version(C_POLY, kVerPoly);
        w.put<uint8_t>(1);                                 // unique_ptr<NativeVector>: valid
        w.version(C_VECTOR, kVerVector);
        w.put<uint64_t>(s.P.mod(primes[i]));
        w.put<uint64_t>(N);
        w.raw(data + i * N, N * 8);
        w.put<uint32_t>(kFormatEvaluation);
        write_nparams(w, s, primes[i]);
    }
    w.put<uint32_t>(kFormatEvaluation);
    // ILDCRTParams: identified by its limb set (the sets that occur: 0..nl-1 and all L+K limbs)
    const int key = (primes.size() > (size_t)s.P.L) ? 63 : (int)primes.size();
    if (w.shared(&s.dparams_key[key])) {
        w.version(C_DPARAMS, kVerParams);
        w.put<uint32_t>((uint32_t)(2 * N)); w.put<uint32_t>((uint32_t)N); w.put<uint8_t>(1);
        w.size(primes.size());                             // composite modulus as its limbs (see the header comment)
        for (int p : primes) w.put<uint64_t>(s.P.mod(p));
        w.size(primes.size());
        for (int p : primes) write_nparams(w, s, p);
    }
}
inline void read_dcrt(Reader& r, Session& s, std::vector<int>& primes, std::vector<uint64_t>& data) {
    r.version(C_DCRT, kVerDcrt);
    const size_t n = r.size(64);
    primes.resize(n);
    for (size_t i = 0; i < n; i++) {
        r.version(C_POLY, kVerPoly);
        if (r.get<uint8_t>() != 1) throw std::runtime_error("hx_serial: empty NativePoly");
        r.version(C_VECTOR, kVerVector);
        const uint64_t m = r.get<uint64_t>(), len = r.get<uint64_t>();
        if (!s.have_params || len != (uint64_t)s.P.N) throw std::runtime_error("hx_serial: limb length is not the ring dimension");
        const size_t off = data.size();
        data.resize(off + len);
        r.raw(data.data() + off, len * 8);
        if (r.get<uint32_t>() != kFormatEvaluation) throw std::runtime_error("hx_serial: only EVALUATION format polynomials are supported");
        primes[i] = read_nparams(r, s);
        if (s.P.mod(primes[i]) != m) throw std::runtime_error("hx_serial: vector modulus differs from its parameters");
    }
    if (r.get<uint32_t>() != kFormatEvaluation) throw std::runtime_error("hx_serial: only EVALUATION format polynomials are supported");
    bool fresh;
    const uint32_t id = r.shared(&fresh);
    if (fresh) {
        r.version(C_DPARAMS, kVerParams);
        r.get<uint32_t>(); r.get<uint32_t>(); r.get<uint8_t>();
        const size_t k = r.size(64);
        for (size_t i = 0; i < k; i++) r.get<uint64_t>();
        const size_t k2 = r.size(64);
        std::vector<int> set(k2);
        for (size_t i = 0; i < k2; i++) set[i] = read_nparams(r, s);
        s.dparams_id[id] = set;
    }
    if (s.dparams_id[id] != primes) throw std::runtime_error("hx_serial: DCRTPoly limbs differ from its ILDCRTParams");
}

inline std::vector<int> first_limbs(int n) { std::vector<int> v(n); for (int i = 0; i < n; i++) v[i] = i; return v; }

// Ciphertext<DCRTPoly> as Serial::SerializeToFile(path, ct, SerType::BINARY) writes a shared_ptr<CiphertextImpl>
inline void write_ciphertext(Writer& w, Session& s, const HostCt& ct) {
    w.shared_fresh();
    w.version(C_CIPHERTEXT, kVerCiphertext);
    write_object_base(w, s);
    w.size((uint64_t)ct.ncomp);
    const std::vector<int> limbs = first_limbs(ct.nl);
    for (int c = 0; c < ct.ncomp; c++) write_dcrt(w, s, limbs, ct.data.data() + (size_t)c * ct.nl * s.P.N);
    w.put<uint32_t>((uint32_t)ct.deg);      // d: noiseScaleDeg
    w.put<uint32_t>((uint32_t)ct.level);    // l
    w.put<uint32_t>(0);                     // h: hops level
    w.size(0);                              // m: metadata map
    double sf = 1.0;
    for (int i = 0; i < ct.deg; i++) sf *= (double)(1ull << s.P.scale_bits);
    w.put<double>(sf);                      // s: scaling factor (FIXEDMANUAL: 2^(scale_bits * deg))
    w.put<uint32_t>(kEncodingCkksPacked);   // e
    w.put<uint64_t>(1);                     // si
    w.put<uint32_t>((uint32_t)(s.P.N / 2)); // sl: slots
}
inline HostCt read_ciphertext(Reader& r, Session& s) {
    bool fresh;
    r.shared(&fresh);
    if (!fresh) throw std::runtime_error("hx_serial: ciphertext body expected");
    r.version(C_CIPHERTEXT, kVerCiphertext);
    read_object_base(r, s);
    HostCt ct;
    ct.ncomp = (int)r.size(3);
    for (int c = 0; c < ct.ncomp; c++) {
        std::vector<int> limbs;
        read_dcrt(r, s, limbs, ct.data);
        if (c == 0) ct.nl = (int)limbs.size();
        if (limbs != first_limbs(ct.nl)) throw std::runtime_error("hx_serial: ciphertext components must hold limbs q_0..q_{l-1}");
    }
    ct.deg = (int)r.get<uint32_t>();
    ct.level = (int)r.get<uint32_t>();
    r.get<uint32_t>();
    if (r.size(1 << 16) != 0) throw std::runtime_error("hx_serial: ciphertext metadata is not supported");
    r.get<double>(); r.get<uint32_t>(); r.get<uint64_t>(); r.get<uint32_t>();
    return ct;
}
// EvalKeyRelin body (no pointer wrapper): CryptoObject, ak, bk -- dnum DCRTPolys over all L+K limbs each
inline void write_evalkey(Writer& w, Session& s, const HostKey& k) {
    w.shared_fresh();                        // every key is its own object
    w.version(C_EVALKEY, kVerEvalKey);
    write_object_base(w, s);
    const int np = s.P.L + s.P.K;
    const std::vector<int> all = first_limbs(np);
    for (const std::vector<uint64_t>* v : {&k.a, &k.b}) {
        w.size((uint64_t)s.P.dnum);
        for (int j = 0; j < s.P.dnum; j++) write_dcrt(w, s, all, v->data() + (size_t)j * np * s.P.N);
    }
}
inline HostKey read_evalkey(Reader& r, Session& s) {
    bool fresh;
    r.shared(&fresh);
    if (!fresh) throw std::runtime_error("hx_serial: evaluation key body expected");
    r.version(C_EVALKEY, kVerEvalKey);
    read_object_base(r, s);
    HostKey k;
    for (std::vector<uint64_t>* v : {&k.a, &k.b}) {
        const size_t n = r.size(8);
        if ((int)n != s.P.dnum) throw std::runtime_error("hx_serial: evaluation key with a different number of digits");
        for (size_t j = 0; j < n; j++) {
            std::vector<int> limbs;
            read_dcrt(r, s, limbs, *v);
            if (limbs != first_limbs(s.P.L + s.P.K)) throw std::runtime_error("hx_serial: evaluation key limbs must be Q then P");
        }
    }
    return k;
}

// ---- the files of src/main.cpp:251-285 from / into a live context ---------------------------------------------------------
inline HostKey export_key(hx_ctx* ctx, const Params& P, uint32_t galois) {
    HostKey k;
    const size_t words = (size_t)P.dnum * (P.L + P.K) * P.N;
    k.a.resize(words); k.b.resize(words);
    if (hx_key_export(ctx, galois, k.a.data(), k.b.data()) != 0) throw std::runtime_error(hx_last_error());
    return k;
}
// SerializeEvalMultKey: map<keyTag, vector<EvalKey>> with one tag and one key
inline bool write_mult_key_stream(std::ostream& os, hx_ctx* ctx) {
    Session s; s.P = params_of(ctx); s.have_params = true;
    Writer w(os);
    w.size(1); w.str("");
    w.size(1);
    write_evalkey(w, s, export_key(ctx, s.P, 0));
    return w.good();
}
inline bool read_mult_key_stream(std::istream& is, hx_ctx* ctx) {
    Session s; s.P = params_of(ctx); s.have_params = true;
    Reader r(is);
    if (r.size(16) != 1) return false;
    r.str();
    if (r.size(16) != 1) return false;
    HostKey k = read_evalkey(r, s);
    return hx_key_load_mult(ctx, k.a.data(), k.b.data()) == 0;
}
// SerializeEvalAutomorphismKey / SerializeEvalSumKey: map<keyTag, shared_ptr<map<u32, EvalKey>>>; `only` restricts the Galois
// elements (EvalSum keys = 5^(2^i) mod 2N)
inline bool write_auto_key_stream(std::ostream& os, hx_ctx* ctx, const std::vector<uint32_t>* only) {
    Session s; s.P = params_of(ctx); s.have_params = true;
    std::vector<uint32_t> g((size_t)hx_key_count(ctx, nullptr));
    if (hx_key_list(ctx, g.data(), (int)g.size()) < 0) return false;
    std::vector<uint32_t> sel;
    for (uint32_t x : g) {
        bool keep = !only;
        if (only) for (uint32_t y : *only) keep = keep || x == y;
        if (keep) sel.push_back(x);
    }
    Writer w(os);
    w.size(1); w.str("");
    w.shared_fresh();
    w.size(sel.size());
    for (uint32_t x : sel) {
        w.put<uint32_t>(x);
        HostKey k = export_key(ctx, s.P, x);
        write_evalkey(w, s, k);
    }
    return w.good();
}
inline bool read_auto_key_stream(std::istream& is, hx_ctx* ctx) {
    Session s; s.P = params_of(ctx); s.have_params = true;
    Reader r(is);
    if (r.size(16) != 1) return false;
    r.str();
    bool fresh;
    r.shared(&fresh);
    const size_t n = r.size(1 << 16);
    for (size_t i = 0; i < n; i++) {
        const uint32_t g = r.get<uint32_t>();
        HostKey k = read_evalkey(r, s);
        if (hx_key_load_auto(ctx, g, k.a.data(), k.b.data()) != 0) return false;
    }
    return true;
}
inline std::vector<uint32_t> eval_sum_galois(int N) {
    std::vector<uint32_t> v;
    uint64_t m = 2ull * N, g = 5;
    for (int i = 0; (1 << i) < N / 2; i++) { v.push_back((uint32_t)g); g = (g * g) % m; }
    return v;
}

}  // namespace hxser
}  // namespace lbcrypto
