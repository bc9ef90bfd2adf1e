// SerType::BINARY files through the shim (shim/hx_serial.h): every object kind the reference serialises
// (src/main.cpp:124-165, 251-285; src/enroller/*.cpp; src/sender/sender_diag.cpp:87-91) is written by one context and read by a
// SECOND context built from the serialised parameter block, and must come back bit-identical:
//   ciphertexts (fresh, 3-component degree-2, rescaled), the EvalMult key, the automorphism keys, the EvalSum subset,
//   the key pair (a ciphertext encrypted under the re-imported public key decrypts under the re-imported secret).
// Prints "OK" and exits 0.  Built by shim/Makefile (needs no reference); run on the GPU by tests/test_reference_unmodified.py.
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "openfhe.h"

using namespace lbcrypto;
using CT = Ciphertext<DCRTPoly>;

static std::vector<uint64_t> bits(const CT& c) {
    int ncomp = 0, nl = 0, level = 0, deg = 0;
    shim::check(hx_ct_info(c->handle(), &ncomp, &nl, &level, &deg), "info");
    std::vector<uint64_t> v(size_t(ncomp) * nl * c->backend()->N + 4);
    shim::check(hx_ct_download(c->handle(), v.data() + 4), "download");
    v[0] = ncomp; v[1] = nl; v[2] = level; v[3] = deg;
    return v;
}
#define REQUIRE(x) do { if (!(x)) { std::fprintf(stderr, "FAILED: %s (line %d)\n", #x, __LINE__); return 1; } } while (0)

int main() {
    setenv("HYDIA_SHIM_SERIAL_FILES", "1", 1);
    char dir[] = "/tmp/hx_serial_XXXXXX";
    REQUIRE(mkdtemp(dir) != nullptr);
    REQUIRE(chdir(dir) == 0);
    mkdir("serial", 0755);
    mkdir("serial/db_diagonal", 0755);
    CCParams<CryptoContextCKKSRNS> params;
    params.SetSecurityLevel(HEStd_128_classic);
    params.SetMultiplicativeDepth(11);
    params.SetScalingModSize(45);
    params.SetScalingTechnique(FIXEDMANUAL);
    params.SetRingDim(4096);
    CryptoContext<DCRTPoly> cc = GenCryptoContext(params);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    cc->EvalSumKeyGen(kp.secretKey);
    cc->EvalRotateKeyGen(kp.secretKey, {1, 2, 5, -3, 512});
    const size_t slots = cc->GetEncodingParams()->GetBatchSize();
    std::vector<double> va(slots), vb(slots);
    for (size_t i = 0; i < slots; i++) { va[i] = std::sin(0.001 * i); vb[i] = std::cos(0.003 * i); }
    CT a = cc->Encrypt(kp.publicKey, cc->MakeCKKSPackedPlaintext(va)), b = cc->Encrypt(kp.publicKey, cc->MakeCKKSPackedPlaintext(vb));
    CT t = cc->EvalMultNoRelin(a, b);
    CT r = cc->EvalMult(a, b);
    cc->RescaleInPlace(r);
    // ---- write everything
    REQUIRE(Serial::SerializeToFile("serial/cryptocontext.bin", cc, SerType::BINARY));
    REQUIRE(Serial::SerializeToFile("serial/publickey.bin", kp.publicKey, SerType::BINARY));
    REQUIRE(Serial::SerializeToFile("serial/privatekey.bin", kp.secretKey, SerType::BINARY));
    { std::ofstream f("serial/multkey.bin", std::ios::binary); REQUIRE(cc->SerializeEvalMultKey(f, SerType::BINARY)); }
    { std::ofstream f("serial/rotkey.bin", std::ios::binary); REQUIRE(cc->SerializeEvalAutomorphismKey(f, SerType::BINARY)); }
    { std::ofstream f("serial/sumkey.bin", std::ios::binary); REQUIRE(cc->SerializeEvalSumKey(f, SerType::BINARY)); }
    REQUIRE(Serial::SerializeToFile("serial/ct_a.bin", a, SerType::BINARY));
    REQUIRE(Serial::SerializeToFile("serial/ct_t.bin", t, SerType::BINARY));
    REQUIRE(Serial::SerializeToFile("serial/ct_r.bin", r, SerType::BINARY));
    REQUIRE(Serial::SerializeToFile("serial/db_diagonal/index3.bin", b, SerType::BINARY));
    const hxser::Params P1 = hxser::params_of(cc->backend()->ctx);
    const hxser::HostKey m1 = hxser::export_key(cc->backend()->ctx, P1, 0);
    std::vector<uint32_t> g1((size_t)hx_key_count(cc->backend()->ctx, nullptr));
    REQUIRE(hx_key_list(cc->backend()->ctx, g1.data(), (int)g1.size()) == (int)g1.size());
    // ---- a second context from the files alone
    Serial::detail::table().clear();          // forget the resident handles: the files are all there is
    CryptoContext<DCRTPoly> cc2;
    REQUIRE(Serial::DeserializeFromFile("serial/cryptocontext.bin", cc2, SerType::BINARY));
    REQUIRE(cc2->backend()->ctx != cc->backend()->ctx);
    REQUIRE(hxser::params_of(cc2->backend()->ctx) == P1);
    PublicKey<DCRTPoly> pk2;
    PrivateKey<DCRTPoly> sk2;
    REQUIRE(Serial::DeserializeFromFile("serial/publickey.bin", pk2, SerType::BINARY));
    REQUIRE(Serial::DeserializeFromFile("serial/privatekey.bin", sk2, SerType::BINARY));
    { std::ifstream f("serial/multkey.bin", std::ios::binary); REQUIRE(cc2->DeserializeEvalMultKey(f, SerType::BINARY)); }
    { std::ifstream f("serial/sumkey.bin", std::ios::binary); REQUIRE(cc2->DeserializeEvalSumKey(f, SerType::BINARY)); }
    { std::ifstream f("serial/rotkey.bin", std::ios::binary); REQUIRE(cc2->DeserializeEvalAutomorphismKey(f, SerType::BINARY)); }
    // evaluation keys: bit-identical residues after write -> read -> load -> export
    const hxser::HostKey m2 = hxser::export_key(cc2->backend()->ctx, P1, 0);
    REQUIRE(m1.a == m2.a && m1.b == m2.b);
    std::vector<uint32_t> g2((size_t)hx_key_count(cc2->backend()->ctx, nullptr));
    REQUIRE(hx_key_list(cc2->backend()->ctx, g2.data(), (int)g2.size()) == (int)g2.size());
    REQUIRE(g1 == g2 && !g1.empty());
    for (uint32_t g : g1) {
        const hxser::HostKey k1 = hxser::export_key(cc->backend()->ctx, P1, g), k2 = hxser::export_key(cc2->backend()->ctx, P1, g);
        REQUIRE(k1.a == k2.a && k1.b == k2.b);
    }
    // ciphertexts: bit-identical, shape and level included
    const char* names[] = {"serial/ct_a.bin", "serial/ct_t.bin", "serial/ct_r.bin", "serial/db_diagonal/index3.bin"};
    const CT orig[] = {a, t, r, b};
    for (int i = 0; i < 4; i++) {
        CT back;
        REQUIRE(Serial::DeserializeFromFile(names[i], back, SerType::BINARY));
        REQUIRE(back->backend()->ctx == cc2->backend()->ctx);
        REQUIRE(bits(back) == bits(orig[i]));
    }
    // key pair: the public key's residues, and a fresh encryption under it decrypts under the re-imported secret; the same
    // evaluation on both contexts gives the same bits
    std::vector<uint64_t> p1(2 * (size_t)P1.L * P1.N), p2(p1.size());
    REQUIRE(hx_client_export_public(cc->backend()->client, p1.data(), p1.data() + p1.size() / 2) == 0);
    REQUIRE(hx_client_export_public(cc2->backend()->client, p2.data(), p2.data() + p2.size() / 2) == 0);
    REQUIRE(p1 == p2);
    CT a2;
    REQUIRE(Serial::DeserializeFromFile("serial/ct_a.bin", a2, SerType::BINARY));
    CT rot1 = cc->EvalRotate(a, 5), rot2 = cc2->EvalRotate(a2, 5);
    REQUIRE(bits(rot1) == bits(rot2));
    Plaintext out;
    cc2->Decrypt(sk2, cc2->Encrypt(pk2, cc2->MakeCKKSPackedPlaintext(va)), &out);
    double err = 0;
    for (size_t i = 0; i < slots; i++) err = std::max(err, std::fabs(out->GetRealPackedValue()[i] - va[i]));
    REQUIRE(err < 1e-7);
    // a truncated file is refused, not half-read
    { std::ifstream in("serial/ct_a.bin", std::ios::binary); std::string s((std::istreambuf_iterator<char>(in)), {}); std::ofstream o("serial/ct_cut.bin", std::ios::binary); o.write(s.data(), (std::streamsize)s.size() / 2); }
    CT cut;
    REQUIRE(!Serial::DeserializeFromFile("serial/ct_cut.bin", cut, SerType::BINARY));
    std::printf("OK\n");
    return 0;
}
