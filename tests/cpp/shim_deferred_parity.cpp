// The deferred-execution layer of shim/openfhe.h must not change a single bit: the DiagonalSender and HersSender call patterns
// (src/sender/sender_diag.cpp:20-33, 66-94; src/sender/sender_hers.cpp:61-98), issued through the shim's lbcrypto surface on a small
// ring, once with the layer on and once with HYDIA_DEFERRED=0 (every call executed as it comes), same seed => same keys and
// encryptions.  Also the fall-backs: an incomplete pattern, a recorded rotation read directly, a recorded sum read before the
// relinearisation.  Prints "OK" and exits 0 when every pair of outputs is identical.  Built by shim/Makefile (needs no reference).
#include <cstdio>
#include <cstdlib>
#include <random>

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

static std::vector<std::vector<uint64_t>> run(bool deferred) {
    setenv("HYDIA_DEFERRED", deferred ? "2" : "0", 1);
    const int ndiag = 8, nblocks = 2, ncols = 6;
    CCParams<CryptoContextCKKSRNS> params;
    params.SetSecurityLevel(HEStd_128_classic);
    params.SetMultiplicativeDepth(11);
    params.SetScalingModSize(45);
    params.SetScalingTechnique(FIXEDMANUAL);
    params.SetRingDim(4096);
    CryptoContext<DCRTPoly> cc = GenCryptoContext(params);
    cc->Enable(PKE);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    std::vector<int> rots;
    for (int i = 1; i < ndiag; i++) rots.push_back(i);
    cc->EvalRotateKeyGen(kp.secretKey, rots);
    const size_t slots = cc->GetEncodingParams()->GetBatchSize();
    std::mt19937_64 rng(7);
    std::uniform_real_distribution<double> u(-0.2, 0.2);
    auto fresh = [&]() {
        std::vector<double> v(slots);
        for (double& x : v) x = u(rng);
        return cc->Encrypt(kp.publicKey, cc->MakeCKKSPackedPlaintext(v));
    };
    for (int b = 0; b < nblocks; b++)
        for (int d = 0; d < ndiag; d++)
            if (!Serial::SerializeToFile("serial/db_diagonal/index" + std::to_string(b * 512 + d) + ".bin", fresh(), SerType::BINARY)) std::abort();
    for (int j = 0; j < ncols; j++)
        if (!Serial::SerializeToFile("serial/db_hers/matrix3/index" + std::to_string(j) + ".bin", fresh(), SerType::BINARY)) std::abort();
    auto load = [&](const std::string& path) {
        CT c;
        if (!Serial::DeserializeFromFile(path, c, SerType::BINARY)) std::abort();
        return c;
    };
    std::vector<std::vector<uint64_t>> out;
    CT q = fresh();

    // DiagonalSender::computeSimilarity, both blocks, then a consumer that reads block 1 first
    auto hoisted = [&](int skip) {
        auto pre = cc->EvalFastRotationPrecompute(q);
        std::vector<CT> rot(ndiag);
        rot[0] = q;
        for (int i = 1; i < ndiag; i++) rot[i] = cc->EvalFastRotation(q, i, 2 * cc->GetRingDimension(), pre);
        std::vector<CT> scores;
        for (int b = 0; b < nblocks; b++) {
            std::vector<CT> s;
            for (int i = 0; i < ndiag; i++)
                if (i != skip) s.push_back(cc->EvalMultNoRelin(rot[i], load("serial/db_diagonal/index" + std::to_string(b * 512 + i) + ".bin")));
            for (size_t i = 1; i < s.size(); i++) cc->EvalAddInPlace(s[0], s[i]);
            cc->RelinearizeInPlace(s[0]);
            cc->RescaleInPlace(s[0]);
            scores.push_back(s[0]);
        }
        if (skip < 0) {      // the compare loop of sender_diag.cpp:40-43 over both block scores (recorded chains -> one batch)
            std::vector<CT> cmp;
            for (const CT& sc : scores) {
                CT c = cc->EvalChebyshevFunction([](double x) -> double { return x >= 0.44 ? 1 : -1; }, sc, -1, 1, 59);
                c = cc->EvalPoly(c, {0.0, 315.0 / 128.0, 0.0, -420.0 / 128.0, 0.0, 378.0 / 128.0, 0.0, -180.0 / 128.0, 0.0, 35.0 / 128.0});
                cc->EvalAddInPlace(c, 1.0);
                cmp.push_back(c);
            }
            out.push_back(bits(cmp[1]));
            out.push_back(bits(cmp[0]));
            CT lone = cc->EvalPoly(scores[0], {0.5, 0.25, 0.125, 0.0, 0.03, 0.0, 0.01});      // a chain of its own kind: runs alone
            out.push_back(bits(lone));
        }
        out.push_back(bits(scores[1]));
        out.push_back(bits(scores[0]));
        out.push_back(bits(rot[2]));                                   // a recorded rotation read directly
    };
    hoisted(-1);      // complete pattern -> fused entries when the layer is on
    hoisted(3);       // diagonal 3 missing -> recipe executed call by call
    {                 // a recorded sum read before it is relinearised, then carried on with
        auto pre = cc->EvalFastRotationPrecompute(q);
        CT a = cc->EvalMultNoRelin(q, load("serial/db_diagonal/index0.bin"));
        CT b = cc->EvalMultNoRelin(cc->EvalFastRotation(q, 1, 2 * cc->GetRingDimension(), pre), load("serial/db_diagonal/index1.bin"));
        cc->EvalAddInPlace(a, b);
        out.push_back(bits(a));
        cc->RelinearizeInPlace(a);
        cc->RescaleInPlace(a);
        out.push_back(bits(a));
    }
    // HersSender::computeSimilarityHelper: one query ciphertext per column
    std::vector<CT> hq;
    for (int j = 0; j < ncols; j++) hq.push_back(fresh());
    auto hers = [&](int upto) {
        std::vector<CT> s(upto);
        for (int j = 0; j < upto; j++) {
            s[j] = cc->EvalMultNoRelin(hq[j], load("serial/db_hers/matrix3/index" + std::to_string(j) + ".bin"));
            cc->RelinearizeInPlace(s[j]);
            cc->RescaleInPlace(s[j]);
        }
        for (int j = 1; j < upto; j++) cc->EvalAddInPlace(s[0], s[j]);
        out.push_back(bits(s[0]));
    };
    hers(ncols);
    hers(1);          // a single term: never becomes a sum
    {                 // product read before its relinearisation
        CT p = cc->EvalMultNoRelin(hq[2], load("serial/db_hers/matrix3/index2.bin"));
        out.push_back(bits(p));
    }
    return out;
}

int main() {
    try {
        auto on = run(true), off = run(false);
        if (on.size() != off.size()) { std::printf("FAIL: %zu vs %zu outputs\n", on.size(), off.size()); return 1; }
        for (size_t i = 0; i < on.size(); i++)
            if (on[i] != off[i]) { std::printf("FAIL: output %zu differs between deferred and call-by-call execution\n", i); return 1; }
        std::printf("OK %zu outputs identical\n", on.size());
        return 0;
    } catch (const std::exception& e) {
        std::printf("FAIL: %s\n", e.what());
        return 2;
    }
}
