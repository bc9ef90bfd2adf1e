// hx_client.cu -- synthetic client side for smoke / bench / tests: key generation, encode + encrypt
// (Enroller / Receiver crypto) and decrypt + decode, running on the device next to the engine.
// Restates what the reference obtains from cc->KeyGen / EvalMultKeyGen / EvalRotateKeyGen /
// Encrypt / Decrypt (/root/reference/src/main.cpp:184-206, src/openFHE_wrapper.cpp:74-85).
// It is NOT on the Sender hot path: in deployment keys and ciphertexts arrive from the client
// through hx_key_load_* / hx_ct_upload.  Randomness is a counter-based hash PRNG (SplitMix64),
// adequate for synthetic benchmarks, not a CSPRNG.
#include <cmath>
#include <complex>

#include "hx_client.h"

namespace hx {

constexpr int kCT = 256;

__device__ __forceinline__ u64 splitmix(u64 x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__device__ __forceinline__ u64 rnd(u64 seed, u64 stream, u64 idx, u64 attempt) {
    return splitmix(splitmix(splitmix(seed ^ (stream * 0xD1342543DE82EF95ull)) + idx) + attempt * 0x2545F4914F6CDD1Dull);
}

// uniform residues in [0, q): rejection sampling below the largest multiple of q
__global__ void k_sample_uniform(DevTables T, u64* out, LimbMap lm, u64 seed, u64 stream) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T.N) return;
    size_t b = blockIdx.y;
    u64 q = T.pc[lm.prime[b % lm.period]].q;
    u64 lim = (~0ull / q) * q;
    u64 r, at = 0;
    do { r = rnd(seed, stream * 4096 + b, idx, at++); } while (r >= lim);
    out[b * T.N + idx] = r % q;
}
__global__ void k_sample_ternary(int N, int* out, u64 seed, u64 stream) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N) return;
    u64 r, at = 0;
    const u64 lim = (~0ull / 3) * 3;
    do { r = rnd(seed, stream, idx, at++); } while (r >= lim);
    out[idx] = (int)(r % 3) - 1;
}
// rounded continuous Gaussian, sigma = 3.19 (OpenFHE's default error distribution parameter)
__global__ void k_sample_gauss(int N, int* out, u64 seed, u64 stream) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N) return;
    u64 r1 = rnd(seed, stream, idx, 0), r2 = rnd(seed, stream, idx, 1);
    double u1 = ((double)(r1 >> 11) + 1.0) * (1.0 / 9007199254740993.0);
    double u2 = (double)(r2 >> 11) * (1.0 / 9007199254740992.0);
    double g = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2) * 3.19;
    out[idx] = (int)llrint(g);
}
// signed small coefficients -> residues for each limb of the map
__global__ void k_small_to_rns(DevTables T, u64* out, const long long* small, LimbMap lm) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T.N) return;
    u64 q = T.pc[lm.prime[blockIdx.y]].q;
    long long v = (long long)ldg64(reinterpret_cast<const u64*>(small + idx));
    long long r = v % (long long)q;
    out[(size_t)blockIdx.y * T.N + idx] = r < 0 ? (u64)(r + (long long)q) : (u64)r;
}
__global__ void k_widen(int N, long long* out, const int* in) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < N) out[idx] = in[idx];
}
// b = e - a*s_new + fac * s_old   (all EVALUATION form; KeySwitchGenInternal, keyswitch-hybrid.cpp)
__global__ void k_ksgen_b(DevTables T, u64* b, const u64* a, const u64* e, const u64* s_new, const u64* s_old, ScalarTable fac) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const PrimeConst pc = T.pc[i];
    size_t o = (size_t)i * T.N + idx;
    u64 v = sub_mod(ldg64(e + o), mulmod(ldg64(a + o), ldg64(s_new + o), pc), pc.q);
    if (s_old != nullptr && fac.v[i] != 0) v = add_mod(v, mulmod(fac.v[i], ldg64(s_old + o), pc), pc.q);
    b[o] = v;
}
// c0 = v*pkb + e0 + m ; c1 = v*pka + e1
__global__ void k_encrypt(DevTables T, u64* c, const u64* v, const u64* pkb, const u64* pka, const u64* e0, const u64* e1,
                          const u64* m, int nl) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const PrimeConst pc = T.pc[i];
    size_t o = (size_t)i * T.N + idx;
    c[o] = add_mod(add_mod(mulmod(ldg64(v + o), ldg64(pkb + o), pc), ldg64(e0 + o), pc.q), ldg64(m + o), pc.q);
    c[(size_t)nl * T.N + o] = add_mod(mulmod(ldg64(v + o), ldg64(pka + o), pc), ldg64(e1 + o), pc.q);
}
// acc = c0 + c1*s (+ c2*s^2)
__global__ void k_decrypt(DevTables T, u64* acc, const u64* c, int ncomp, size_t cs, const u64* s) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T.N) return;
    int i = blockIdx.y;
    const PrimeConst pc = T.pc[i];
    size_t o = (size_t)i * T.N + idx;
    u64 r = ldg64(c + o), sp = ldg64(s + o);
    for (int k = 1; k < ncomp; k++) {
        r = add_mod(r, mulmod(ldg64(c + k * cs + o), sp, pc), pc.q);
        sp = mulmod(sp, ldg64(s + o), pc);
    }
    acc[o] = r;
}

static dim3 grid1(int N, int y) { return dim3((N + kCT - 1) / kCT, y, 1); }

// ------------------------------------------------------------------------------------------ host FFT
static void fft_inplace(std::vector<std::complex<double>>& a, int sign) {
    size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; i++) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    for (size_t len = 2; len <= n; len <<= 1) {
        double ang = sign * 2.0 * M_PI / (double)len;
        std::vector<std::complex<double>> w(len / 2);
        for (size_t k = 0; k < len / 2; k++) w[k] = std::polar(1.0, ang * (double)k);
        for (size_t i = 0; i < n; i += len)
            for (size_t k = 0; k < len / 2; k++) {
                auto u = a[i + k], v = a[i + k + len / 2] * w[k];
                a[i + k] = u + v;
                a[i + k + len / 2] = u - v;
            }
    }
}

// ------------------------------------------------------------------------------------------ Client
Client::Client(std::shared_ptr<Engine> e, u64 seed_) : keep(e), eng(e.get()), seed(seed_) {
    Engine::Guard g(eng);
    const int N = eng->N, np = eng->L + eng->K;
    slot_idx.resize(N / 2);
    u64 m = 2ull * N, x = 1;
    for (int j = 0; j < N / 2; j++) { slot_idx[j] = (int)((x - 1) / 2); x = (x * 5) % m; }
    zeta.resize(N);
    for (int n = 0; n < N; n++) zeta[n] = std::polar(1.0, M_PI * (double)n / (double)N);
    // secret key: ternary, all L+K limbs, evaluation form
    HX_CUDA(cudaMalloc(&d_sk_small, sizeof(int) * N));
    k_sample_ternary<<<(N + kCT - 1) / kCT, kCT, 0, eng->st>>>(N, d_sk_small, seed, next_stream++);
    check_launch("k_sample_ternary");
    HX_CUDA(cudaMalloc(&d_sk, sizeof(u64) * (size_t)np * N));
    std::vector<int> all;
    for (int i = 0; i < np; i++) all.push_back(i);
    small_to_rns(d_sk_small, d_sk, all);
    // public key over Q: (b = -a*s + e, a)
    std::vector<int> qs;
    for (int i = 0; i < eng->L; i++) qs.push_back(i);
    HX_CUDA(cudaMalloc(&d_pka, sizeof(u64) * (size_t)eng->L * N));
    HX_CUDA(cudaMalloc(&d_pkb, sizeof(u64) * (size_t)eng->L * N));
    k_sample_uniform<<<grid1(N, eng->L), kCT, 0, eng->st>>>(eng->T, d_pka, eng->limb_map(qs), seed, next_stream++);
    check_launch("k_sample_uniform");
    u64* e_rns = gauss_rns(qs);
    ScalarTable zero = eng->scalar_table(std::vector<u64>(eng->L, 0));
    k_ksgen_b<<<grid1(N, eng->L), kCT, 0, eng->st>>>(eng->T, d_pkb, d_pka, e_rns, d_sk, nullptr, zero);
    check_launch("k_ksgen_b");
    eng->free_words(e_rns);
    eng->sync();
}
Client::Client(std::shared_ptr<Engine> e, u64 seed_, const int8_t* sk, const u64* pkb, const u64* pka) : keep(e), eng(e.get()), seed(seed_) {
    Engine::Guard g(eng);
    const int N = eng->N, np = eng->L + eng->K;
    slot_idx.resize(N / 2);
    u64 m = 2ull * N, x = 1;
    for (int j = 0; j < N / 2; j++) { slot_idx[j] = (int)((x - 1) / 2); x = (x * 5) % m; }
    zeta.resize(N);
    for (int n = 0; n < N; n++) zeta[n] = std::polar(1.0, M_PI * (double)n / (double)N);
    next_stream = 1u << 20;   // fresh randomness for encryptions made after the import: streams the generating client never used
    std::vector<int> small(N);
    for (int i = 0; i < N; i++) {
        if (sk[i] < -1 || sk[i] > 1) throw std::runtime_error("hx_client_import: secret key coefficients must be in {-1, 0, 1}");
        small[i] = sk[i];
    }
    HX_CUDA(cudaMalloc(&d_sk_small, sizeof(int) * N));
    HX_CUDA(cudaMemcpy(d_sk_small, small.data(), sizeof(int) * N, cudaMemcpyHostToDevice));
    HX_CUDA(cudaMalloc(&d_sk, sizeof(u64) * (size_t)np * N));
    std::vector<int> all;
    for (int i = 0; i < np; i++) all.push_back(i);
    small_to_rns(d_sk_small, d_sk, all);
    HX_CUDA(cudaMalloc(&d_pka, sizeof(u64) * (size_t)eng->L * N));
    HX_CUDA(cudaMalloc(&d_pkb, sizeof(u64) * (size_t)eng->L * N));
    HX_CUDA(cudaMemcpy(d_pka, pka, sizeof(u64) * (size_t)eng->L * N, cudaMemcpyHostToDevice));
    HX_CUDA(cudaMemcpy(d_pkb, pkb, sizeof(u64) * (size_t)eng->L * N, cudaMemcpyHostToDevice));
    eng->sync();
}
void Client::export_public(u64* pkb, u64* pka) {
    Engine::Guard g(eng);
    eng->sync();
    HX_CUDA(cudaMemcpy(pkb, d_pkb, sizeof(u64) * (size_t)eng->L * eng->N, cudaMemcpyDeviceToHost));
    HX_CUDA(cudaMemcpy(pka, d_pka, sizeof(u64) * (size_t)eng->L * eng->N, cudaMemcpyDeviceToHost));
}
Client::~Client() {
    cudaFree(d_sk_small); cudaFree(d_sk); cudaFree(d_pka); cudaFree(d_pkb);
}
void Client::small_to_rns(const int* d_small, u64* out, const std::vector<int>& primes) {
    const int N = eng->N;
    long long* wide = (long long*)eng->alloc_words(N);
    k_widen<<<(N + kCT - 1) / kCT, kCT, 0, eng->st>>>(N, wide, d_small);
    check_launch("k_widen");
    k_small_to_rns<<<grid1(N, (int)primes.size()), kCT, 0, eng->st>>>(eng->T, out, wide, eng->limb_map(primes));
    check_launch("k_small_to_rns");
    eng->ntt_limbs(out, (int)primes.size(), primes, false);
    eng->free_words((u64*)wide);
}
u64* Client::gauss_rns(const std::vector<int>& primes) {
    const int N = eng->N;
    int* small = (int*)eng->alloc_words(N / 2 + 1);
    k_sample_gauss<<<(N + kCT - 1) / kCT, kCT, 0, eng->st>>>(N, small, seed, next_stream++);
    check_launch("k_sample_gauss");
    u64* out = eng->alloc_words((size_t)primes.size() * N);
    small_to_rns(small, out, primes);
    eng->free_words((u64*)small);
    return out;
}
u64* Client::ternary_rns(const std::vector<int>& primes) {
    const int N = eng->N;
    int* small = (int*)eng->alloc_words(N / 2 + 1);
    k_sample_ternary<<<(N + kCT - 1) / kCT, kCT, 0, eng->st>>>(N, small, seed, next_stream++);
    check_launch("k_sample_ternary");
    u64* out = eng->alloc_words((size_t)primes.size() * N);
    small_to_rns(small, out, primes);
    eng->free_words((u64*)small);
    return out;
}

// KeySwitchGenInternal(old, new): digit j: a_j uniform, b_j = -a_j*s_new + e_j + [i in digit j] P*s_old.
EvalKey Client::ks_keygen(const u64* s_old, const u64* s_new) {
    const int N = eng->N, np = eng->L + eng->K;
    EvalKey k;
    size_t words = (size_t)eng->dnum * np * N;
    HX_CUDA(cudaMalloc(&k.a, sizeof(u64) * words));
    HX_CUDA(cudaMalloc(&k.b, sizeof(u64) * words));
    std::vector<int> all;
    for (int i = 0; i < np; i++) all.push_back(i);
    LimbMap lm = eng->limb_map(all);
    for (int j = 0; j < eng->dnum; j++) {
        u64* aj = k.a + (size_t)j * np * N;
        u64* bj = k.b + (size_t)j * np * N;
        k_sample_uniform<<<grid1(N, np), kCT, 0, eng->st>>>(eng->T, aj, lm, seed, next_stream++);
        check_launch("k_sample_uniform");
        u64* e = gauss_rns(all);
        std::vector<u64> fac(np, 0);
        for (int i = j * eng->alpha; i < eng->L && i < (j + 1) * eng->alpha; i++) fac[i] = eng->P_mod(i);
        k_ksgen_b<<<grid1(N, np), kCT, 0, eng->st>>>(eng->T, bj, aj, e, s_new, s_old, eng->scalar_table(fac));
        check_launch("k_ksgen_b");
        eng->free_words(e);
    }
    eng->finish_key(k);
    return k;
}
void Client::gen_mult_key() {
    Engine::Guard g(eng);
    const int N = eng->N, np = eng->L + eng->K;
    u64* s2 = eng->alloc_words((size_t)np * N);
    launch_ew(eng->T, EW_MUL, s2, d_sk, d_sk, np, 1, 0, 0, 0, eng->st);   // limb i <-> prime i for all L+K limbs
    eng->adopt_mult_key(ks_keygen(s2, d_sk));
    eng->free_words(s2);
    eng->sync();
}
void Client::gen_rot_key_galois(u64 k) {
    Engine::Guard g(eng);
    if (eng->has_rot_key(k)) return;
    const int N = eng->N, np = eng->L + eng->K;
    // EvalAutomorphismKeyGen: new key = sigma_{k^-1}(s)
    u64 m = 2ull * N, kinv = 1;
    {   // k^-1 mod 2N by exponentiation: the group has exponent N/2 * 2 = N
        u64 e = (u64)N - 1, b = k;
        while (e) { if (e & 1) kinv = (kinv * b) % m; b = (b * b) % m; e >>= 1; }
    }
    u64* sp = eng->alloc_words((size_t)np * N);
    launch_automorph(eng->T, sp, d_sk, eng->automap(kinv), np, eng->st);
    eng->adopt_rot_key(k, ks_keygen(d_sk, sp));
    eng->free_words(sp);
}
void Client::gen_rot_keys(const int* rot, int n) {
    for (int i = 0; i < n; i++) gen_rot_key_galois(eng->galois_for_rotation(rot[i]));
    eng->sync();
}
void Client::gen_sum_keys() {
    // EvalSumKeyGen: Galois elements 5^(2^i), i < log2(N/2)
    u64 m = 2ull * eng->N, g = 5;
    for (int i = 0; (1 << i) < eng->N / 2; i++) {
        gen_rot_key_galois(g);
        g = (g * g) % m;
    }
    eng->sync();
}

std::vector<long long> Client::encode(const double* values, int n) const {
    const int N = eng->N;
    std::vector<std::complex<double>> v(N, 0.0);
    for (int j = 0; j < n && j < N / 2; j++) {
        v[slot_idx[j]] = values[j];
        v[N - 1 - slot_idx[j]] = values[j];   // conjugate slot (real inputs)
    }
    fft_inplace(v, -1);
    double sc = std::ldexp(1.0, eng->scale_bits);
    std::vector<long long> out(N);
    for (int k = 0; k < N; k++) out[k] = std::llrint((v[k] * std::conj(zeta[k])).real() / (double)N * sc);
    return out;
}

void Client::encode_rns(const double* values, int n, u64* host_out) {
    Engine::Guard g(eng);
    const int N = eng->N, L = eng->L;
    std::vector<long long> coeffs = encode(values, n);
    long long* d_small = (long long*)eng->alloc_words(N);
    HX_CUDA(cudaMemcpyAsync(d_small, coeffs.data(), sizeof(long long) * N, cudaMemcpyHostToDevice, eng->st));
    std::vector<int> qs;
    for (int i = 0; i < L; i++) qs.push_back(i);
    u64* m = eng->alloc_words((size_t)L * N);
    k_small_to_rns<<<grid1(N, L), kCT, 0, eng->st>>>(eng->T, m, d_small, eng->limb_map(qs));
    check_launch("k_small_to_rns");
    eng->ntt_limbs(m, L, qs, false);
    HX_CUDA(cudaMemcpyAsync(host_out, m, sizeof(u64) * (size_t)L * N, cudaMemcpyDeviceToHost, eng->st));
    HX_CUDA(cudaStreamSynchronize(eng->st));
    eng->free_words(m); eng->free_words((u64*)d_small);
}

Ct Client::encrypt(const double* values, int n) {
    Engine::Guard g(eng);
    const int N = eng->N, L = eng->L;
    std::vector<long long> coeffs = encode(values, n);
    long long* d_small = (long long*)eng->alloc_words(N);
    HX_CUDA(cudaMemcpyAsync(d_small, coeffs.data(), sizeof(long long) * N, cudaMemcpyHostToDevice, eng->st));
    std::vector<int> qs;
    for (int i = 0; i < L; i++) qs.push_back(i);
    u64* m = eng->alloc_words((size_t)L * N);
    k_small_to_rns<<<grid1(N, L), kCT, 0, eng->st>>>(eng->T, m, d_small, eng->limb_map(qs));
    check_launch("k_small_to_rns");
    eng->ntt_limbs(m, L, qs, false);
    u64* v = ternary_rns(qs);
    u64* e0 = gauss_rns(qs);
    u64* e1 = gauss_rns(qs);
    Ct ct = eng->new_ct(2, L, 0, 1);
    k_encrypt<<<grid1(N, L), kCT, 0, eng->st>>>(eng->T, ct->d, v, d_pkb, d_pka, e0, e1, m, L);
    check_launch("k_encrypt");
    HX_CUDA(cudaStreamSynchronize(eng->st));   // coeffs (host) must outlive the copy
    eng->free_words(v); eng->free_words(e0); eng->free_words(e1); eng->free_words(m); eng->free_words((u64*)d_small);
    return ct;
}

void Client::enroll(int block, int diag, const double* values, int n) {
    Engine::Guard g(eng);
    Ct ct = encrypt(values, n);
    eng->db_store_device(block, diag, ct->d);
}

void Client::decrypt(const Ct& ct, double* out) {
    Engine::Guard g(eng);
    const int N = eng->N, nl = ct->nl;
    u64* acc = eng->alloc_words((size_t)nl * N);
    k_decrypt<<<grid1(N, nl), kCT, 0, eng->st>>>(eng->T, acc, ct->d, ct->ncomp, ct->comp_stride(), d_sk);
    check_launch("k_decrypt");
    std::vector<int> qs;
    for (int i = 0; i < nl; i++) qs.push_back(i);
    eng->ntt_limbs(acc, nl, qs, true);
    int use = nl >= 2 ? 2 : 1;
    std::vector<u64> h((size_t)use * N);
    HX_CUDA(cudaMemcpyAsync(h.data(), acc, sizeof(u64) * h.size(), cudaMemcpyDeviceToHost, eng->st));
    eng->sync();
    eng->free_words(acc);
    double sc = std::ldexp(1.0, eng->scale_bits * ct->deg);
    std::vector<std::complex<double>> x(N);
    u64 q0 = eng->mods[0];
    if (use == 1) {
        for (int k = 0; k < N; k++) {
            u64 r = h[k];
            double val = r > q0 / 2 ? -(double)(q0 - r) : (double)r;
            x[k] = zeta[k] * (val / sc);
        }
    } else {
        // two-limb CRT (Garner): x = r0 + q0 * ((r1 - r0) * q0^-1 mod q1), centred mod q0*q1
        u64 q1 = eng->mods[1];
        u64 q0inv = invmod(q0 % q1, q1);
        u128 Q = (u128)q0 * q1;
        for (int k = 0; k < N; k++) {
            u64 r0 = h[k], r1 = h[(size_t)N + k];
            u64 t = mulmod(submod(r1, r0 % q1, q1), q0inv, q1);
            u128 val = (u128)r0 + (u128)q0 * t;
            double dv = val > Q / 2 ? -(double)(Q - val) : (double)val;
            x[k] = zeta[k] * (dv / sc);
        }
    }
    fft_inplace(x, +1);
    for (int j = 0; j < N / 2; j++) out[j] = x[slot_idx[j]].real();
}

void Client::secret(int8_t* out) {
    Engine::Guard g(eng);
    std::vector<int> h(eng->N);
    HX_CUDA(cudaMemcpyAsync(h.data(), d_sk_small, sizeof(int) * eng->N, cudaMemcpyDeviceToHost, eng->st));
    eng->sync();
    for (int i = 0; i < eng->N; i++) out[i] = (int8_t)h[i];
}

}  // namespace hx
