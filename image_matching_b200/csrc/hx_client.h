// hx_client.h -- synthetic client side (keys, encrypt, decrypt) next to the engine; see hx_client.cu.
#pragma once
#include <complex>

#include "hx_engine.h"

namespace hx {

class Client {
public:
    Client(std::shared_ptr<Engine> e, u64 seed);
    // from a serialised key pair: secret as small coefficients in {-1,0,1}, public key (b, a) as plain residues [L][N] each
    Client(std::shared_ptr<Engine> e, u64 seed, const int8_t* sk, const u64* pkb, const u64* pka);
    void export_public(u64* pkb, u64* pka);
    ~Client();
    void gen_mult_key();                          // EvalMultKeyGen      (src/main.cpp:188)
    void gen_rot_keys(const int* rot, int n);     // EvalRotateKeyGen    (src/main.cpp:206)
    void gen_sum_keys();                          // EvalSumKeyGen       (src/main.cpp:192)
    Ct encrypt(const double* values, int n);      // encryptFromVector   (src/openFHE_wrapper.cpp:74)
    void enroll(int block, int diag, const double* values, int n);   // enroller_diag.cpp:158
    void decrypt(const Ct& ct, double* out);      // decryptToVector     (src/openFHE_wrapper.cpp:81)
    void secret(int8_t* out);
    std::vector<long long> encode(const double* values, int n) const;
    void encode_rns(const double* values, int n, u64* host_out);   // MakeCKKSPackedPlaintext: [L][N], evaluation form

    std::shared_ptr<Engine> keep;
    Engine* eng;
    u64 seed;

private:
    void gen_rot_key_galois(u64 k);
    EvalKey ks_keygen(const u64* s_old, const u64* s_new);
    void small_to_rns(const int* d_small, u64* out, const std::vector<int>& primes);
    u64* gauss_rns(const std::vector<int>& primes);
    u64* ternary_rns(const std::vector<int>& primes);
    u64 next_stream = 1;
    int* d_sk_small = nullptr;
    u64 *d_sk = nullptr, *d_pka = nullptr, *d_pkb = nullptr;
    std::vector<int> slot_idx;
    std::vector<std::complex<double>> zeta;
};

}  // namespace hx
