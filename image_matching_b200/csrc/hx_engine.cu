// hx_engine.cu -- Engine: context tables, memory, keys, and the primitive evaluation ops.
#include <cstdlib>
#include <cstring>
#include "hx_engine.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace hx {

// ------------------------------------------------------------------------------------------ CtObj
// Handle destructors may run on any host thread at any time (hx_ct_free from another OpenMP thread while this engine is inside
// its hoisted rotations with `st` repointed): they take the engine lock and free on the engine stream, on which every handle's
// contents were produced or joined.
CtObj::~CtObj() {
    if (d && eng) { Engine::Guard g(eng); cudaFreeAsync(d, eng->st_main); }
}
u64* CtObj::comp(int c) const { return d + (size_t)c * comp_stride(); }
size_t CtObj::comp_stride() const { return (size_t)nl * eng->N; }
Engine::Digits::~Digits() {
    if (d && eng) { Engine::Guard g(eng); cudaFreeAsync(d, eng->st_main); }
}

// ------------------------------------------------------------------------------------------ setup
Engine::Engine(int N_, int L_, int K_, int dnum_, const u64* q, const u64* p, const u64* psi_q, const u64* psi_p, int scale_bits_,
               int device_)
    : N(N_), logN(ilog2(N_)), L(L_), K(K_), dnum(dnum_), scale_bits(scale_bits_), device(device_) {
    if ((1 << logN) != N || N < 1024 || N > 65536) throw std::runtime_error("ring dimension must be a power of two in [2^10, 2^16]");
    if (L + K > kMaxPrimes) throw std::runtime_error("too many RNS limbs");
    alpha = (L + dnum - 1) / dnum;
    if (alpha > 7 || K > 7) throw std::runtime_error("digit width or P basis above 7 limbs unsupported (carry-free MAC bound)");
    if (dnum < 1 || dnum > kMaxDigits) throw std::runtime_error("number of key-switching digits must be 1..4");
    for (int i = 0; i < L; i++) { mods.push_back(q[i]); psis.push_back(psi_q[i]); }
    for (int i = 0; i < K; i++) { mods.push_back(p[i]); psis.push_back(psi_p[i]); }
    for (u64 m : mods)
        if (m >> 60) throw std::runtime_error("moduli must be below 2^60");
    HX_CUDA(cudaSetDevice(device));
    {   // the engine stream outranks the rotation lanes (hx_hydia.cu): its short dependent kernels are scheduled first
        int least = 0, greatest = 0;
        HX_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        HX_CUDA(cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, greatest));
        st_main = st;
    }
    init_device_kernels(logN);   // > 48 KB dynamic shared memory opt-ins are per device
    // keep freed blocks in the stream-ordered pool: op outputs are allocated per call
    cudaMemPool_t pool;
    HX_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    unsigned long long thresh = ~0ull;
    HX_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));

    const int np = L + K;
    std::vector<PrimeConst> pcs(np);
    std::vector<ulonglong2> twf((size_t)np * N), twi((size_t)np * N);
    std::vector<double2> twd((size_t)np * N);
    for (int i = 0; i < np; i++) {
        u64 m = mods[i];
        PrimeConst& c = pcs[i];
        c.q = m;
        u64 inv = 1;                       // Newton iteration for m^-1 mod 2^64
        for (int it = 0; it < 6; it++) inv *= 2 - m * inv;
        c.qinv_neg = (u64)0 - inv;
        c.r1 = (u64)((((u128)1) << 64) % m);
        c.r2 = mulmod(c.r1, c.r1, m);
        c.mu64 = (u64)((((u128)1) << 64) / m);
        c.ninv = invmod((u64)N % m, m);
        c.ninv_sh = shoup(c.ninv, m);
        c.qinv_d = 1.0 / (double)m;
        if (powmod(psis[i], (u64)N, m) != m - 1) throw std::runtime_error("psi is not a primitive 2N-th root of unity");
        u64 psi_inv = invmod(psis[i], m);
        u64 x = 1, y = 1;
        for (int k = 0; k < N; k++) {
            size_t r = (size_t)i * N + bitrev((u64)k, logN);
            twf[r] = make_ulonglong2(x, shoup(x, m));
            twi[r] = make_ulonglong2(y, shoup(y, m));
            twd[r] = make_double2((double)x, (double)x / (double)m);   // exact / correctly rounded when m < 2^53
            x = mulmod(x, psis[i], m);
            y = mulmod(y, psi_inv, m);
        }
    }
    HX_CUDA(cudaMalloc(&d_pc_, sizeof(PrimeConst) * np));
    HX_CUDA(cudaMalloc(&d_twf_, sizeof(ulonglong2) * twf.size()));
    HX_CUDA(cudaMalloc(&d_twi_, sizeof(ulonglong2) * twi.size()));
    HX_CUDA(cudaMalloc(&d_twd_, sizeof(double2) * twd.size()));
    HX_CUDA(cudaMemcpy(d_twd_, twd.data(), sizeof(double2) * twd.size(), cudaMemcpyHostToDevice));
    HX_CUDA(cudaMemcpy(d_pc_, pcs.data(), sizeof(PrimeConst) * np, cudaMemcpyHostToDevice));
    HX_CUDA(cudaMemcpy(d_twf_, twf.data(), sizeof(ulonglong2) * twf.size(), cudaMemcpyHostToDevice));
    HX_CUDA(cudaMemcpy(d_twi_, twi.data(), sizeof(ulonglong2) * twi.size(), cudaMemcpyHostToDevice));
    T.N = N; T.logN = logN; T.nprimes = np; T.pc = d_pc_; T.tw_fwd = d_twf_; T.tw_inv = d_twi_; T.tw_fwd_d = d_twd_;
}

Engine::~Engine() {
    cudaSetDevice(device);
    cudaStreamSynchronize(st_main);
    for (auto& kv : modup_tabs_) cudaFree(kv.second);
    for (auto& kv : moddown_tabs_) cudaFree(kv.second);
    for (auto& kv : last_tabs_) cudaFree(kv.second);
    for (auto& kv : automaps_) cudaFree(kv.second);
    for (auto& kv : rot_keys_) { cudaFree(kv.second.a); cudaFree(kv.second.b); }
    if (have_mult_key_) { cudaFree(mult_key_.a); cudaFree(mult_key_.b); }
    for (auto& kv : db_) cudaFree(kv.second.d);
    for (auto& r : rot_slots_) { if (r.d) cudaFree(r.d); if (r.released) cudaEventDestroy(r.released); }
    for (void* p : owned_) cudaFree(p);
    for (auto& r : prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    for (auto ev : prof_pool) cudaEventDestroy(ev);
    if (timer_a) { cudaEventDestroy(timer_a); cudaEventDestroy(timer_b); }
    cudaFree(d_pc_); cudaFree(d_twf_); cudaFree(d_twi_); cudaFree(d_twd_);
    for (auto s : aux_streams) if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
    for (auto s : lane_streams) if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
    for (auto ev : aux_events) if (ev) cudaEventDestroy(ev);
    for (auto ev : lane_events) if (ev) cudaEventDestroy(ev);
    if (mac_gate_) cudaEventDestroy(mac_gate_);
    if (trace_origin) cudaEventDestroy(trace_origin);
    for (auto& r : trace_recs) cudaEventDestroy(r.ev);
    cudaStreamDestroy(st_main);
}

// ------------------------------------------------------------------------------------------ profiler
const char* prof_kind_name(int kind) {
    static const char* names[PK_COUNT] = {"ntt_fwd", "ntt_inv", "baseconv", "ks_inner", "moddown", "elementwise", "scalar", "tensor",
                                          "automorph", "rescale", "diag_mac", "misc", "ks_fused", "modup", "ks_hoisted"};
    return (kind >= 0 && kind < PK_COUNT) ? names[kind] : "?";
}
Engine::Scope::Scope(Engine* e_, int kind_, int nlaunch, double bytes) : e(e_), kind(kind_) {
    e->launches += nlaunch;
    e->prof_launches[kind] += nlaunch;
    e->prof_bytes[kind] += bytes;
    if (!((e->prof_on >> kind) & 1u) || e->prof_no_events) return;
    cudaEvent_t a;
    auto grab = [&]() {
        cudaEvent_t ev;
        if (!e->prof_pool.empty()) { ev = e->prof_pool.back(); e->prof_pool.pop_back(); }
        else cudaEventCreate(&ev);
        return ev;
    };
    a = grab(); b = grab();
    cudaEventRecord(a, e->st);
    e->prof_recs.push_back({kind, a, b});
}
Engine::Scope::~Scope() {
    if (b) cudaEventRecord(b, e->st);
}
void Engine::trace_start() {
    Guard g(this);
    if (!trace_origin) HX_CUDA(cudaEventCreate(&trace_origin));
    for (auto& r : trace_recs) cudaEventDestroy(r.ev);
    trace_recs.clear();
    HX_CUDA(cudaEventRecord(trace_origin, st_main));
    trace_on = true;
}
void Engine::trace_mark(const char* name, cudaStream_t s) {
    if (!trace_on) return;
    static const char* filter = std::getenv("HYDIA_TRACE_FILTER");   // diagnostic: comma list of the marks to keep
    if (filter && !std::strstr(filter, name)) return;
    Guard g(this);
    TraceRec r;
    r.name = name;
    HX_CUDA(cudaEventCreate(&r.ev));
    HX_CUDA(cudaEventRecord(r.ev, s ? s : st));
    trace_recs.push_back(std::move(r));
}
std::string Engine::trace_dump() {
    Guard g(this);
    std::string out;
    for (auto& r : trace_recs) {
        float ms = -1.f;
        HX_CUDA(cudaEventSynchronize(r.ev));   // each mark on its own stream: no device-wide synchronisation that would hide a stall
        cudaEventElapsedTime(&ms, trace_origin, r.ev);
        char buf[64];
        snprintf(buf, sizeof buf, ":%.3f;", ms);
        out += r.name + buf;
        cudaEventDestroy(r.ev);
    }
    trace_recs.clear();
    trace_on = false;
    return out;
}
void Engine::prof_enable(unsigned mask) {
    Guard g(this);
    prof_collect();
    prof_on = mask;
}
void Engine::prof_collect() {
    Guard g(this);
    if (prof_recs.empty()) return;
    HX_CUDA(cudaStreamSynchronize(st));
    for (auto& r : prof_recs) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) prof_ms[r.kind] += ms;
        prof_pool.push_back(r.a);
        prof_pool.push_back(r.b);
    }
    prof_recs.clear();
}
void Engine::prof_reset() {
    Guard g(this);
    prof_collect();
    for (int k = 0; k < PK_COUNT; k++) { prof_launches[k] = 0; prof_ms[k] = 0; prof_bytes[k] = 0; }
}
void Engine::timer_start() {
    Guard g(this);
    if (!timer_a) { HX_CUDA(cudaEventCreate(&timer_a)); HX_CUDA(cudaEventCreate(&timer_b)); }
    HX_CUDA(cudaEventRecord(timer_a, st));
}
double Engine::timer_stop_ms() {
    Guard g(this);
    if (!timer_a) throw std::runtime_error("timer_stop without timer_start");
    HX_CUDA(cudaEventRecord(timer_b, st));
    HX_CUDA(cudaEventSynchronize(timer_b));
    float ms = 0;
    HX_CUDA(cudaEventElapsedTime(&ms, timer_a, timer_b));
    return ms;
}

// ------------------------------------------------------------------------------------------ memory
u64* Engine::alloc_words(size_t words) {
    void* p = nullptr;
    HX_CUDA(cudaMallocAsync(&p, words * sizeof(u64), st));
    return (u64*)p;
}
void Engine::free_words(u64* p) { cudaFreeAsync(p, st); }
void Engine::sync() { HX_CUDA(cudaStreamSynchronize(st)); }

Ct Engine::new_ct(int ncomp, int nl, int level, int deg, int nb) {
    auto c = std::make_shared<CtObj>();
    c->eng = this; c->keep = weak_from_this().lock(); c->ncomp = ncomp; c->nl = nl; c->level = level; c->deg = deg; c->nb = nb;
    c->d = alloc_words((size_t)nb * ncomp * nl * N);
    return c;
}
Ct Engine::stack(const std::vector<Ct>& v) {
    Guard g(this);
    if (v.empty()) throw std::runtime_error("stack: empty input");
    const Ct& f = v[0];
    Ct out = new_ct(f->ncomp, f->nl, f->level, f->deg, (int)v.size());
    for (size_t i = 0; i < v.size(); i++) {
        if (v[i]->nb != 1 || v[i]->ncomp != f->ncomp || v[i]->nl != f->nl || v[i]->deg != f->deg)
            throw std::runtime_error("stack: ciphertexts differ in shape");
        HX_CUDA(cudaMemcpyAsync(out->d + i * out->batch_stride(), v[i]->d, sizeof(u64) * out->batch_stride(), cudaMemcpyDeviceToDevice, st));
    }
    return out;
}
std::vector<Ct> Engine::unstack(const Ct& a) {
    Guard g(this);
    std::vector<Ct> out;
    if (a->nb == 1) { out.push_back(a); return out; }
    for (int i = 0; i < a->nb; i++) {
        Ct c = new_ct(a->ncomp, a->nl, a->level, a->deg);
        HX_CUDA(cudaMemcpyAsync(c->d, a->d + (size_t)i * a->batch_stride(), sizeof(u64) * a->batch_stride(), cudaMemcpyDeviceToDevice, st));
        out.push_back(c);
    }
    return out;
}
Ct Engine::upload(const u64* host, int ncomp, int nl, int level, int deg) {
    Guard g(this);
    Ct c = new_ct(ncomp, nl, level, deg);
    HX_CUDA(cudaMemcpyAsync(c->d, host, sizeof(u64) * (size_t)ncomp * nl * N, cudaMemcpyHostToDevice, st));
    sync();
    c->settled = true;
    return c;
}
void Engine::download(const Ct& ct, u64* host) {
    Guard g(this);
    HX_CUDA(cudaMemcpyAsync(host, ct->d, sizeof(u64) * (size_t)ct->nb * ct->ncomp * ct->nl * N, cudaMemcpyDeviceToHost, st));
    sync();
}
Ct Engine::clone(const Ct& a) {
    Guard g(this);
    Ct c = new_ct(a->ncomp, a->nl, a->level, a->deg, a->nb);
    HX_CUDA(cudaMemcpyAsync(c->d, a->d, sizeof(u64) * (size_t)a->nb * a->ncomp * a->nl * N, cudaMemcpyDeviceToDevice, st));
    return c;
}

Ct Engine::concat(const std::vector<Ct>& v) {
    Guard g(this);
    if (v.empty()) throw std::runtime_error("concat: empty input");
    const Ct& f = v[0];
    int nb = 0;
    for (auto& c : v) {
        if (c->ncomp != f->ncomp || c->nl != f->nl || c->deg != f->deg) throw std::runtime_error("concat: ciphertexts differ in shape");
        nb += c->nb;
    }
    Ct out = new_ct(f->ncomp, f->nl, f->level, f->deg, nb);
    size_t off = 0;
    for (auto& c : v) {
        const size_t words = (size_t)c->nb * c->batch_stride();
        HX_CUDA(cudaMemcpyAsync(out->d + off, c->d, sizeof(u64) * words, cudaMemcpyDeviceToDevice, st));
        off += words;
    }
    return out;
}
std::vector<Ct> Engine::split(const Ct& a, int parts) {
    Guard g(this);
    if (parts <= 0 || a->nb % parts) throw std::runtime_error("split: batch size is not a multiple of the number of parts");
    std::vector<Ct> out;
    if (parts == 1) { out.push_back(a); return out; }
    const int nb = a->nb / parts;
    for (int i = 0; i < parts; i++) {
        Ct c = new_ct(a->ncomp, a->nl, a->level, a->deg, nb);
        const size_t words = (size_t)nb * a->batch_stride();
        HX_CUDA(cudaMemcpyAsync(c->d, a->d + (size_t)i * words, sizeof(u64) * words, cudaMemcpyDeviceToDevice, st));
        out.push_back(c);
    }
    return out;
}

// ------------------------------------------------------------------------------------------ helpers
LimbMap Engine::limb_map(const std::vector<int>& primes) const {
    LimbMap lm;
    lm.period = (int)primes.size();
    if (lm.period > kMaxPrimes) throw std::runtime_error("limb map too long");
    for (int i = 0; i < lm.period; i++) lm.prime[i] = (short)primes[i];
    return lm;
}
ScalarTable Engine::scalar_table(const std::vector<u64>& v) const {
    ScalarTable s;
    std::memset(&s, 0, sizeof(s));
    for (size_t i = 0; i < v.size(); i++) {
        s.v[i] = v[i] % mods[i];
        s.sh[i] = shoup(s.v[i], mods[i]);
    }
    return s;
}
u64 Engine::P_mod(int i) const {
    u64 r = 1;
    for (int k = 0; k < K; k++) r = mulmod(r, mods[L + k] % mods[i], mods[i]);
    return r;
}
void Engine::ntt_limbs(u64* data, int nbatch, const std::vector<int>& primes, bool inverse) {
    int active = 0;
    for (int b = 0; b < nbatch; b++) active += primes[b % primes.size()] >= 0;
    Scope sc(this, inverse ? PK_NTT_INV : PK_NTT_FWD, (logN > 11) ? 2 : 1, 16.0 * N * active);
    launch_ntt(T, data, nbatch, limb_map(primes), inverse, st);
}
u64 Engine::galois_for_rotation(int r) const {
    u64 m = 2ull * N;
    if (r == 0) return 1;
    u64 g = 5;
    if (r < 0) {   // 5 has order N/2 in (Z/2N)*, so 5^-1 = 5^(N/2 - 1)
        g = powmod(5, (u64)N / 2 - 1, m);
        r = -r;
    }
    return powmod(g, (u64)r, m);
}
const u32* Engine::automap(u64 galois) {
    auto it = automaps_.find(galois);
    if (it != automaps_.end()) return it->second;
    u32* m = nullptr;
    HX_CUDA(cudaMalloc(&m, sizeof(u32) * N));
    { Scope sc(this, PK_MISC, 1, 4.0 * N); launch_automap(N, logN, galois, m, st); }
    automaps_[galois] = m;
    return m;
}

// ApproxSwitchCRTBasis tables: src/dst are prime indices.
BaseConvTable* Engine::make_table(const std::vector<int>& src, const std::vector<int>& dst, const std::vector<u64>* out_scale) {
    BaseConvTable h;
    std::memset(&h, 0, sizeof(h));
    h.n_in = (int)src.size();
    h.n_out = (int)dst.size();
    for (int i = 0; i < h.n_in; i++) {
        u64 qi = mods[src[i]];
        h.src[i] = (short)src[i];
        u64 hat = 1;
        for (int k = 0; k < h.n_in; k++)
            if (k != i) hat = mulmod(hat, mods[src[k]] % qi, qi);
        h.hatinv[i] = invmod(hat, qi);
        h.hatinv_sh[i] = shoup(h.hatinv[i], qi);
        h.hatinv_n[i] = mulmod(h.hatinv[i], invmod((u64)N % qi, qi), qi);
        h.hatinv_n_sh[i] = shoup(h.hatinv_n[i], qi);
        for (int j = 0; j < h.n_out; j++) {
            u64 pj = mods[dst[j]];
            u64 hm = 1;
            for (int k = 0; k < h.n_in; k++)
                if (k != i) hm = mulmod(hm, mods[src[k]] % pj, pj);
            u64 r1 = (u64)((((u128)1) << 64) % pj);
            if (out_scale) hm = mulmod(hm, (*out_scale)[j] % pj, pj);
            h.mat[i * h.n_out + j] = mulmod(hm, r1, pj);   // Montgomery form
            h.mat30[i * h.n_out + j] = split30(h.mat[i * h.n_out + j]);
        }
    }
    for (int j = 0; j < h.n_out; j++) h.dst[j] = (short)dst[j];
    BaseConvTable* d = nullptr;
    HX_CUDA(cudaMalloc(&d, sizeof(BaseConvTable)));
    HX_CUDA(cudaMemcpy(d, &h, sizeof(BaseConvTable), cudaMemcpyHostToDevice));
    return d;
}
const BaseConvTable* Engine::modup_table(int nl, int digit) {
    auto key = std::make_pair(nl, digit);
    auto it = modup_tabs_.find(key);
    if (it != modup_tabs_.end()) return it->second;
    int ndig = std::min((nl + alpha - 1) / alpha, dnum);
    int lo = digit * alpha, hi = (digit == ndig - 1) ? nl : (digit + 1) * alpha;
    std::vector<int> src, dst;
    for (int i = lo; i < hi; i++) src.push_back(i);
    for (int i = 0; i < nl; i++)
        if (i < lo || i >= hi) dst.push_back(i);
    for (int k = 0; k < K; k++) dst.push_back(L + k);
    return modup_tabs_[key] = make_table(src, dst);
}
const BaseConvTable* Engine::last_limb_table(int prime) {
    auto it = last_tabs_.find(prime);
    if (it != last_tabs_.end()) return it->second;
    return last_tabs_[prime] = make_table({prime}, {});
}
const BaseConvTable* Engine::moddown_table(int nl) {
    auto it = moddown_tabs_.find(nl);
    if (it != moddown_tabs_.end()) return it->second;
    std::vector<int> src, dst;
    for (int k = 0; k < K; k++) src.push_back(L + k);
    for (int i = 0; i < nl; i++) dst.push_back(i);
    std::vector<u64> pinv(nl);
    for (int i = 0; i < nl; i++) pinv[i] = invmod(P_mod(i), mods[i]);
    return moddown_tabs_[nl] = make_table(src, dst, &pinv);   // ApproxModDown's final * P^-1 folded into the matrix
}

// ------------------------------------------------------------------------------------------ keys
static EvalKey upload_key(Engine& e, const u64* a, const u64* b) {
    EvalKey k;
    size_t words = (size_t)e.dnum * (e.L + e.K) * e.N;
    HX_CUDA(cudaMalloc(&k.a, words * sizeof(u64)));
    HX_CUDA(cudaMalloc(&k.b, words * sizeof(u64)));
    HX_CUDA(cudaMemcpyAsync(k.a, a, words * sizeof(u64), cudaMemcpyHostToDevice, e.st));
    HX_CUDA(cudaMemcpyAsync(k.b, b, words * sizeof(u64), cudaMemcpyHostToDevice, e.st));
    e.finish_key(k);
    e.sync();
    return k;
}
// Engine-internal key form (launch_key_finish): Q limbs pre-multiplied by P^-1 (so ModDown's last multiplication disappears from
// the key-switch kernels); limbs below 2^46 as doubles (FP64 inner product), the 60-bit limbs (q_0 and P) as Montgomery residues
// split into 30-bit halves (carry-free integer inner product).  Input: plain residues, [dnum][L+K][N].
void Engine::finish_key(EvalKey& k) {
    std::vector<u64> pinv(L);
    for (int i = 0; i < L; i++) pinv[i] = invmod(P_mod(i), mods[i]);
    ScalarTable ps = scalar_table(pinv);
    Scope sc(this, PK_MISC, 2, 32.0 * dnum * (L + K) * N);
    launch_key_finish(T, k.a, dnum, L, K, ps, st);
    launch_key_finish(T, k.b, dnum, L, K, ps, st);
}
void Engine::export_key(u64 galois, u64* host_a, u64* host_b) {
    Guard g(this);
    if (galois == 0 && !have_mult_key_) throw std::runtime_error("EvalMultKey is not generated");
    const EvalKey& k = galois ? rot_key(galois) : mult_key_;
    std::vector<u64> pm(L);
    for (int i = 0; i < L; i++) pm[i] = P_mod(i);
    ScalarTable ps = scalar_table(pm);
    const size_t words = (size_t)dnum * (L + K) * N;
    u64* tmp = alloc_words(words);
    const u64* src[2] = {k.a, k.b};
    u64* dst[2] = {host_a, host_b};
    for (int c = 0; c < 2; c++) {
        Scope sc(this, PK_MISC, 1, 16.0 * words);
        launch_key_unfinish(T, tmp, src[c], dnum, L, K, ps, st);
        HX_CUDA(cudaMemcpyAsync(dst[c], tmp, sizeof(u64) * words, cudaMemcpyDeviceToHost, st));
        sync();
    }
    free_words(tmp);
}
void Engine::load_mult_key(const u64* a, const u64* b) {
    Guard g(this);
    if (have_mult_key_) { cudaFree(mult_key_.a); cudaFree(mult_key_.b); }
    mult_key_ = upload_key(*this, a, b);
    have_mult_key_ = true;
}
void Engine::load_rot_key(u64 galois, const u64* a, const u64* b) {
    Guard g(this);
    auto it = rot_keys_.find(galois);
    if (it != rot_keys_.end()) { cudaFree(it->second.a); cudaFree(it->second.b); }
    rot_keys_[galois] = upload_key(*this, a, b);
}
void Engine::adopt_rot_key(u64 galois, EvalKey k) {
    Guard g(this);
    auto it = rot_keys_.find(galois);
    if (it != rot_keys_.end()) { cudaFree(it->second.a); cudaFree(it->second.b); }
    rot_keys_[galois] = k;
}
void Engine::adopt_mult_key(EvalKey k) {
    Guard g(this);
    if (have_mult_key_) { cudaFree(mult_key_.a); cudaFree(mult_key_.b); }
    mult_key_ = k;
    have_mult_key_ = true;
}
const EvalKey& Engine::rot_key(u64 galois) const {
    auto it = rot_keys_.find(galois);
    if (it == rot_keys_.end()) throw std::runtime_error("EvalKey for Galois index " + std::to_string(galois) + " is not generated");
    return it->second;
}

// ------------------------------------------------------------------------------------------ key switching
// EvalKeySwitchPrecomputeCore (upstream keyswitch-hybrid.cpp; sender_diag.cpp:22): digit split +
// ModUp. Returns [ndig][nl+K][N]; every digit holds its own limbs verbatim and the converted ones.
u64* Engine::ks_precompute(const u64* c, int nl, int* ndig_out, int nbatch, size_t c_bs) {
    int ndig = std::min((nl + alpha - 1) / alpha, dnum);
    int ext = nl + K;
    u64* digits = alloc_words((size_t)nbatch * ndig * ext * N);
    if (fused_ks_supported(logN) && !force_unfused) {
        ModUpArgs m;
        m.c = c; m.c_bs = c_bs; m.nbatch = nbatch;
        m.nl = nl; m.L = L; m.K = K; m.ndig = ndig; m.alpha = alpha;
        for (int j = 0; j < ndig; j++) m.tabs[j] = modup_table(nl, j);
        m.scratch = alloc_words((size_t)nbatch * nl * N);
        m.digits = digits;
        m.warp_tasks = ks_moddown_w;
        {   // algorithmic bytes: input polynomial once, digits once
            Scope sc(this, PK_MODUP, (ks_moddown_w && moddown_w_supported(logN, alpha)) ? 3 : 4, 8.0 * N * nbatch * (nl + (double)ndig * ext));
            launch_fused_modup(T, m, st);
        }
        free_words(m.scratch);
        *ndig_out = ndig;
        return digits;
    }
    if (nbatch != 1) throw std::runtime_error("batched key switching needs the fused pipeline");
    u64* scratch = alloc_words((size_t)alpha * N);
    for (int j = 0; j < ndig; j++) {
        int lo = j * alpha, hi = (j == ndig - 1) ? nl : (j + 1) * alpha, sz = hi - lo;
        u64* dj = digits + (size_t)j * ext * N;
        HX_CUDA(cudaMemcpyAsync(dj + (size_t)lo * N, c + (size_t)lo * N, sizeof(u64) * sz * N, cudaMemcpyDeviceToDevice, st));
        HX_CUDA(cudaMemcpyAsync(scratch, c + (size_t)lo * N, sizeof(u64) * sz * N, cudaMemcpyDeviceToDevice, st));
        std::vector<int> src;
        for (int i = lo; i < hi; i++) src.push_back(i);
        ntt_limbs(scratch, sz, src, true);
        // converted limbs: Q-limbs below lo, Q-limbs from hi, then P. Written in two pieces around the own limbs.
        const BaseConvTable* tab = modup_table(nl, j);
        int n_out = nl - sz + K;
        u64* compl_ = alloc_words((size_t)n_out * N);
        { Scope sc(this, PK_BASECONV, 1, 8.0 * N * (sz + n_out)); launch_baseconv(T, tab, sz, n_out, scratch, 0, compl_, 0, 1, st); }
        std::vector<int> dst;
        for (int i = 0; i < nl; i++)
            if (i < lo || i >= hi) dst.push_back(i);
        for (int k = 0; k < K; k++) dst.push_back(L + k);
        ntt_limbs(compl_, n_out, dst, false);
        if (lo > 0) HX_CUDA(cudaMemcpyAsync(dj, compl_, sizeof(u64) * (size_t)lo * N, cudaMemcpyDeviceToDevice, st));
        HX_CUDA(cudaMemcpyAsync(dj + (size_t)hi * N, compl_ + (size_t)lo * N, sizeof(u64) * (size_t)(n_out - lo) * N,
                                cudaMemcpyDeviceToDevice, st));
        free_words(compl_);
    }
    free_words(scratch);
    *ndig_out = ndig;
    return digits;
}

// EvalFastKeySwitchCoreExt + ApproxModDown for a batch of keys sharing one set of digits, fused
// with the "+ c0" and the Galois gather of EvalFastRotation (sender_diag.cpp:25).
void Engine::ks_core_batch(const u64* digits, int ndig, int nl, const std::vector<const EvalKey*>& keys, const u64* addend,
                           int addend_comps, const std::vector<u64>& galois, u64* out, size_t out_bs, size_t digits_bs, size_t addend_bs) {
    const int nb = (int)keys.size(), ext = nl + K;
    if (fused_ks_supported(logN) && !force_unfused) {
        // fused pipeline (hx_fused.cu): the evaluation keys are read once, nothing of size 2 x (l+K) is stored
        for (int b0 = 0; b0 < nb; b0 += kMaxKsBatch) {
            const int n = std::min(kMaxKsBatch, nb - b0);
            FusedKsArgs a;
            a.digits = digits + (size_t)b0 * digits_bs; a.ndig = ndig; a.nl = nl; a.L = L; a.K = K;
            a.digits_bs = digits_bs; a.add_bs = addend_bs;
            for (int i = 0; i < n; i++) {
                a.keyA[i] = keys[b0 + i]->a;
                a.keyB[i] = keys[b0 + i]->b;
                a.galois[i] = galois.empty() ? 1u : (u32)galois[b0 + i];
            }
            a.tmpP = alloc_words((size_t)n * 2 * K * N);
            a.conv = alloc_words((size_t)n * 2 * nl * N);
            const u64* ad = addend ? addend + (size_t)b0 * addend_bs : nullptr;
            a.add0 = (ad && addend_comps >= 1) ? ad : nullptr;
            a.add1 = (ad && addend_comps >= 2) ? ad + (size_t)nl * N : nullptr;
            a.out = out + (size_t)b0 * out_bs;
            a.out_bs = out_bs;
            {   // algorithmic bytes of one key switch after ModUp: evk once, digits once per launch,
                // addend once, result once
                a.moddown_w = ks_moddown_w;
                a.fp_mask = 0;
                for (int i = 0; i < nl; i++)
                    if ((mods[i] >> 46) == 0) a.fp_mask |= 1ull << i;
                const int n_fp = __builtin_popcountll(a.fp_mask);
                const int k3_launches = (n_fp > 0) + (n_fp < nl);   // the Q-limb kernel runs once for the FP64 limbs and once for the 60-bit ones
                const bool mw = ks_moddown_w && moddown_w_supported(logN, K);
                Scope sc(this, ks_kind, (mw ? 2 : 3) + k3_launches,   // launch_fused_tn: ModDown's middle in 1 or 2 launches
                         8.0 * N * ((double)n * 2 * ndig * ext + (double)(digits_bs ? n : 1) * ndig * ext + (double)(addend_bs ? n : 1) * addend_comps * nl + (double)n * 2 * nl));
                launch_fused_ks(T, a, moddown_table(nl), n, st);
            }
            free_words(a.conv);
            free_words(a.tmpP);
        }
        return;
    }
    if (digits_bs || addend_bs) throw std::runtime_error("batched key switching needs the fused pipeline");
    std::vector<const u64*> ha(nb), hb(nb);
    std::vector<const u32*> hm(nb, nullptr);
    for (int i = 0; i < nb; i++) { ha[i] = keys[i]->a; hb[i] = keys[i]->b; }
    if (!galois.empty())
        for (int i = 0; i < nb; i++) hm[i] = automap(galois[i]);
    // pointer tables (a | b | maps) in one device buffer
    u64* ptrs = alloc_words((size_t)3 * nb);
    HX_CUDA(cudaMemcpyAsync(ptrs, ha.data(), sizeof(u64) * nb, cudaMemcpyHostToDevice, st));
    HX_CUDA(cudaMemcpyAsync(ptrs + nb, hb.data(), sizeof(u64) * nb, cudaMemcpyHostToDevice, st));
    HX_CUDA(cudaMemcpyAsync(ptrs + 2 * nb, hm.data(), sizeof(u64) * nb, cudaMemcpyHostToDevice, st));
    HX_CUDA(cudaStreamSynchronize(st));   // host vectors go out of scope; tables are tiny
    size_t ct_bs = (size_t)2 * ext * N;
    u64* ct = alloc_words(ct_bs * nb);
    {   // algorithmic bytes: evk (a,b) streamed once per key, digits once per launch, outputs once
        Scope sc(this, PK_KS_INNER, 1, 8.0 * N * ext * ((double)nb * 2 * ndig + ndig + (double)nb * 2));
        launch_ks_inner(T, digits, ndig, nl, L, K, (const u64* const*)ptrs, (const u64* const*)(ptrs + nb), ct, ct_bs, nb, st);
    }
    // iNTT of the K P-limbs of both components, in place
    std::vector<int> pm(ext, -1);
    for (int k = 0; k < K; k++) pm[nl + k] = L + k;
    ntt_limbs(ct, nb * 2 * ext, pm, true);
    size_t conv_bs = (size_t)2 * nl * N;
    u64* conv = alloc_words(conv_bs * nb);
    {
        Scope sc(this, PK_BASECONV, 1, 8.0 * N * (K + nl) * nb * 2);
        launch_baseconv(T, moddown_table(nl), K, nl, ct + (size_t)nl * N, (size_t)ext * N, conv, (size_t)nl * N, nb * 2, st);
    }
    std::vector<int> qm;
    for (int i = 0; i < nl; i++) qm.push_back(i);
    ntt_limbs(conv, nb * 2 * nl, qm, false);
    {
        Scope sc(this, PK_MODDOWN, 1, 8.0 * N * nl * nb * (2 + 2 + 2 + (addend ? addend_comps : 0)));
        launch_moddown_finish(T, ct, ct_bs, conv, conv_bs, addend, addend_comps, galois.empty() ? nullptr : (const u32* const*)(ptrs + 2 * nb),
                              out, out_bs, nl, nb, st);
    }
    free_words(conv);
    free_words(ct);
    free_words(ptrs);
}

std::shared_ptr<Engine::Digits> Engine::fast_rotation_precompute(const Ct& a) {
    Guard g(this);
    auto dg = std::make_shared<Digits>();
    dg->eng = this;
    dg->keep = weak_from_this().lock();
    dg->nl = a->nl;
    dg->fused = fused_ks_supported(logN) && !force_unfused;
    dg->d = ks_precompute(a->comp(1), a->nl, &dg->ndig);
    return dg;
}
Ct Engine::fast_rotation(const Ct& a, int r, const std::shared_ptr<Digits>& dg) {
    Guard g(this);
    if (r == 0) return clone(a);
    if (a->ncomp != 2 || dg->nl != a->nl) throw std::runtime_error("EvalFastRotation: digits do not match the ciphertext");
    if (dg->fused != (fused_ks_supported(logN) && !force_unfused)) throw std::runtime_error("EvalFastRotation: digits were precomputed by the other key-switch pipeline");
    u64 k = galois_for_rotation(r);
    Ct out = new_ct(2, a->nl, a->level, a->deg);
    ks_core_batch(dg->d, dg->ndig, a->nl, {&rot_key(k)}, a->comp(0), 1, {k}, out->d, 0);
    return out;
}
Ct Engine::automorphism(const Ct& a, u64 k) {
    Guard g(this);
    if (a->ncomp != 2) throw std::runtime_error("EvalAutomorphism needs 2-component ciphertexts");
    int ndig;
    if (a->nb == 1) {
        u64* dg = ks_precompute(a->comp(1), a->nl, &ndig);
        Ct out = new_ct(2, a->nl, a->level, a->deg);
        ks_core_batch(dg, ndig, a->nl, {&rot_key(k)}, a->comp(0), 1, {k}, out->d, 0);
        free_words(dg);
        return out;
    }
    // a batch: every entry gets its own ModUp, all share the key and the Galois element
    u64* dg = ks_precompute(a->comp(1), a->nl, &ndig, a->nb, a->batch_stride());
    Ct out = new_ct(2, a->nl, a->level, a->deg, a->nb);
    std::vector<const EvalKey*> keys(a->nb, &rot_key(k));
    std::vector<u64> gal(a->nb, k);
    ks_core_batch(dg, ndig, a->nl, keys, a->comp(0), 1, gal, out->d, out->batch_stride(), (size_t)ndig * (a->nl + K) * N, a->batch_stride());
    free_words(dg);
    return out;
}
Ct Engine::rotate(const Ct& a, int r) {
    if (r == 0) return clone(a);
    return automorphism(a, galois_for_rotation(r));
}
Ct Engine::eval_sum(const Ct& a, int n) {
    // EvalSum2nComplex: g = 5; ceil(log2 n) times { ct += sigma_g(ct); g = g^2 mod 2N }
    u64 m = 2ull * N, gk = 5;
    Ct out = a;
    int rounds = 0;
    while ((1 << rounds) < n) rounds++;
    for (int i = 0; i < rounds; i++) {
        out = add(out, automorphism(out, gk));
        gk = mulmod(gk, gk, m);
    }
    return out;
}
Ct Engine::relinearize(const Ct& a) {
    Guard g(this);
    if (a->ncomp == 2) return clone(a);
    if (a->ncomp != 3) throw std::runtime_error("Relinearize supports 3-component ciphertexts");
    if (!have_mult_key_) throw std::runtime_error("EvalMultKey is not generated");
    int ndig;
    u64* dg = ks_precompute(a->comp(2), a->nl, &ndig, a->nb, a->batch_stride());
    Ct out = new_ct(2, a->nl, a->level, a->deg, a->nb);
    std::vector<const EvalKey*> keys(a->nb, &mult_key_);
    ks_core_batch(dg, ndig, a->nl, keys, a->d, 2, {}, out->d, out->batch_stride(), a->nb > 1 ? (size_t)ndig * (a->nl + K) * N : 0, a->nb > 1 ? a->batch_stride() : 0);
    free_words(dg);
    return out;
}

// ------------------------------------------------------------------------------------------ leveled ops
Ct Engine::level_reduce(const Ct& a, int levels) {
    Guard g(this);
    if (levels == 0) return a;
    if (levels < 0 || levels >= a->nl) throw std::runtime_error("LevelReduce: not enough limbs");
    int nl = a->nl - levels;
    Ct out = new_ct(a->ncomp, nl, a->level + levels, a->deg, a->nb);
    HX_CUDA(cudaMemcpy2DAsync(out->d, sizeof(u64) * out->comp_stride(), a->d, sizeof(u64) * a->comp_stride(), sizeof(u64) * (size_t)nl * N,
                              (size_t)a->comps_total(), cudaMemcpyDeviceToDevice, st));
    return out;
}
void Engine::align(Ct& a, Ct& b) {   // FIXEDMANUAL AdjustLevelsInPlace: drop limbs of the fresher operand
    if (a->nl > b->nl) a = level_reduce(a, a->nl - b->nl);
    else if (b->nl > a->nl) b = level_reduce(b, b->nl - a->nl);
}
Ct Engine::add(const Ct& a_, const Ct& b_) {
    Guard g(this);
    Ct a = a_, b = b_;
    align(a, b);
    if (a->deg != b->deg) throw std::runtime_error("EvalAdd: noiseScaleDeg mismatch");
    if (a->nb != b->nb) throw std::runtime_error("EvalAdd: batch size mismatch");
    if (a->ncomp < b->ncomp) std::swap(a, b);
    if (a->nb > 1 && a->ncomp != b->ncomp) throw std::runtime_error("EvalAdd: batched operands must have equal component counts");
    Ct out = new_ct(a->ncomp, a->nl, a->level, a->deg, a->nb);
    const int nc = a->nb > 1 ? a->comps_total() : b->ncomp;
    { Scope sc(this, PK_EW, 1, 24.0 * N * a->nl * nc); launch_ew(T, EW_ADD, out->d, a->d, b->d, a->nl, nc, out->comp_stride(), a->comp_stride(), b->comp_stride(), st); }
    for (int c = b->ncomp; c < a->ncomp; c++)
        HX_CUDA(cudaMemcpyAsync(out->comp(c), a->comp(c), sizeof(u64) * a->comp_stride(), cudaMemcpyDeviceToDevice, st));
    return out;
}
Ct Engine::sub(const Ct& a_, const Ct& b_) {
    Guard g(this);
    Ct a = a_, b = b_;
    align(a, b);
    if (a->deg != b->deg) throw std::runtime_error("EvalSub: noiseScaleDeg mismatch");
    if (a->ncomp != b->ncomp || a->nb != b->nb) throw std::runtime_error("EvalSub: component or batch count mismatch");
    Ct out = new_ct(a->ncomp, a->nl, a->level, a->deg, a->nb);
    { Scope sc(this, PK_EW, 1, 24.0 * N * a->nl * a->comps_total()); launch_ew(T, EW_SUB, out->d, a->d, b->d, a->nl, a->comps_total(), out->comp_stride(), a->comp_stride(), b->comp_stride(), st); }
    return out;
}
Ct Engine::add_many(const std::vector<Ct>& v_) {
    // EvalAddManyInPlace: for j = 1,2,4,..: for i = 0; i + j < n; i += 2j: v[i] += v[i+j]
    std::vector<Ct> v = v_;
    if (v.empty()) throw std::runtime_error("EvalAddMany: empty input");
    for (size_t j = 1; j < v.size(); j *= 2)
        for (size_t i = 0; i + j < v.size(); i += 2 * j) v[i] = add(v[i], v[i + j]);
    return v[0];
}
// GetElementForEvalMult: (int128)(c * 2^scale_bits + 0.5), truncating cast, per-limb residue with sign fix-up.
std::vector<u64> Engine::const_for_mult(double c, int nl) const {
    double sc = std::ldexp(1.0, scale_bits);
    __int128 large = (__int128)(c * sc + 0.5);
    std::vector<u64> out(nl);
    for (int i = 0; i < nl; i++) {
        __int128 r = large % (__int128)mods[i];
        if (r < 0) r += mods[i];
        out[i] = (u64)r;
    }
    return out;
}
// GetElementForEvalAddOrSub: (uint64)(c*2^scale_bits + 0.5) * ((uint64)(2^scale_bits + 0.5))^(deg-1), c >= 0.
std::vector<u64> Engine::const_for_add(double c, int nl, int deg) const {
    double sc = std::ldexp(1.0, scale_bits);
    u64 k = (u64)(c * sc + 0.5);
    u64 isf = (u64)(sc + 0.5);
    std::vector<u64> out(nl);
    for (int i = 0; i < nl; i++) {
        u64 v = k % mods[i];
        for (int d = 1; d < deg; d++) v = mulmod(v, isf % mods[i], mods[i]);
        out[i] = v;
    }
    return out;
}
Ct Engine::mult_const(const Ct& a, double c) {
    Guard g(this);
    Ct out = new_ct(a->ncomp, a->nl, a->level, a->deg + 1, a->nb);
    { Scope sc(this, PK_SCALAR, 1, 16.0 * N * a->nl * a->comps_total()); launch_mul_scalar(T, out->d, a->d, scalar_table(const_for_mult(c, a->nl)), a->nl, a->comps_total(), out->comp_stride(), a->comp_stride(), st); }
    return out;
}
Ct Engine::add_const(const Ct& a, double c) {
    Guard g(this);
    // the CryptoContext front end routes negative constants to EvalSub(ct, -c)
    std::vector<u64> k = const_for_add(std::fabs(c), a->nl, a->deg);
    if (c < 0)
        for (int i = 0; i < a->nl; i++) k[i] = (mods[i] - k[i]) % mods[i];
    Ct out = new_ct(a->ncomp, a->nl, a->level, a->deg, a->nb);
    { Scope sc(this, PK_SCALAR, 1, 16.0 * N * a->nl * a->comps_total()); launch_add_scalar(T, out->d, a->d, scalar_table(k), a->nl, a->ncomp, a->comps_total(), st); }
    return out;
}
Ct Engine::mult_norelin(const Ct& a_, const Ct& b_) {
    Guard g(this);
    Ct a = a_, b = b_;
    align(a, b);
    if (a->ncomp != 2 || b->ncomp != 2) throw std::runtime_error("EvalMultNoRelin supports 2-component inputs");
    if (a->nb != b->nb && a->nb != 1 && b->nb != 1) throw std::runtime_error("EvalMultNoRelin: batch size mismatch");
    const int nb = std::max(a->nb, b->nb);   // a single ciphertext is broadcast against a batch
    Ct out = new_ct(3, a->nl, a->level, a->deg + b->deg, nb);
    { Scope sc(this, PK_TENSOR, 1, 8.0 * N * a->nl * 7 * nb); launch_tensor(T, out->d, a->d, a->comp_stride(), a->nb == 1 && nb > 1, b->d, b->comp_stride(), b->nb == 1 && nb > 1, a->nl, nb, st); }
    return out;
}
Ct Engine::mult(const Ct& a, const Ct& b) { return relinearize(mult_norelin(a, b)); }
Ct Engine::mult_plain(const Ct& a, const u64* host_pt, int nlimbs) {
    Guard g(this);
    if (nlimbs < a->nl) throw std::runtime_error("EvalMult(ct, Plaintext): plaintext has fewer limbs than the ciphertext");
    // the plaintext was encoded at level 0; a ciphertext at a lower level uses its first nl limbs (same integer polynomial)
    u64* pt = alloc_words((size_t)a->nl * N);
    HX_CUDA(cudaMemcpyAsync(pt, host_pt, sizeof(u64) * (size_t)a->nl * N, cudaMemcpyHostToDevice, st));
    Ct out = new_ct(a->ncomp, a->nl, a->level, a->deg + 1, a->nb);
    { Scope sc(this, PK_EW, 1, 8.0 * N * a->nl * (2 * a->comps_total() + 1)); launch_ew(T, EW_MUL, out->d, a->d, pt, a->nl, a->comps_total(), out->comp_stride(), a->comp_stride(), 0, st); }
    HX_CUDA(cudaStreamSynchronize(st));   // the caller's host buffer may go away
    free_words(pt);
    return out;
}
Ct Engine::inner_product(const Ct& a, const Ct& b, int n) { return eval_sum(mult(a, b), n); }
Ct Engine::square(const Ct& a) { return mult(a, a); }

// DropLastElementAndScale (sender_diag.cpp:80).
Ct Engine::rescale(const Ct& a) {
    Guard g(this);
    int nl = a->nl;
    if (nl < 2) throw std::runtime_error("Rescale: no limb left to drop");
    int nc = a->comps_total();   // batch entries are independent components for DropLastElementAndScale
    u64 ql = mods[nl - 1];
    std::vector<u64> qlinv(nl - 1), neg(nl - 1);
    for (int i = 0; i < nl - 1; i++) {
        qlinv[i] = invmod(ql % mods[i], mods[i]);
        neg[i] = (mods[i] - qlinv[i]) % mods[i];
    }
    if (fused_ks_supported(logN) && !force_unfused) {
        RescaleArgs r;
        r.c = a->d; r.c_bs = a->comp_stride(); r.ncomp = nc; r.nl = nl;
        r.last_tab = last_limb_table(nl - 1);
        r.last = alloc_words((size_t)nc * N);
        r.tmp = alloc_words((size_t)nc * (nl - 1) * N);
        Ct out = new_ct(a->ncomp, nl - 1, a->level + 1, a->deg - 1, a->nb);
        r.out = out->d; r.out_bs = out->comp_stride();
        r.negqlinv = scalar_table(neg);
        r.qlinv = scalar_table(qlinv);
        {
            Scope sc(this, PK_RESCALE, 4, 8.0 * N * nc * (nl + nl - 1));
            launch_fused_rescale(T, r, st);
        }
        free_words(r.tmp);
        free_words(r.last);
        return out;
    }
    if (a->nb != 1) throw std::runtime_error("batched rescale needs the fused pipeline");
    u64* last = alloc_words((size_t)nc * N);
    for (int c = 0; c < nc; c++)
        HX_CUDA(cudaMemcpyAsync(last + (size_t)c * N, a->comp(c) + (size_t)(nl - 1) * N, sizeof(u64) * N, cudaMemcpyDeviceToDevice, st));
    ntt_limbs(last, nc, {nl - 1}, true);
    u64* tmp = alloc_words((size_t)nc * (nl - 1) * N);
    { Scope sc(this, PK_RESCALE, 1, 8.0 * N * nc * nl); launch_rescale_prep(T, last, N, tmp, (size_t)(nl - 1) * N, nl, scalar_table(neg), nc, st); }
    std::vector<int> qm;
    for (int i = 0; i < nl - 1; i++) qm.push_back(i);
    ntt_limbs(tmp, nc * (nl - 1), qm, false);
    Ct out = new_ct(nc, nl - 1, a->level + 1, a->deg - 1);
    { Scope sc(this, PK_RESCALE, 1, 24.0 * N * nc * (nl - 1)); launch_rescale_finish(T, out->d, out->comp_stride(), a->d, a->comp_stride(), tmp, (size_t)(nl - 1) * N, nl, scalar_table(qlinv), nc, st); }
    free_words(tmp);
    free_words(last);
    return out;
}

}  // namespace hx
