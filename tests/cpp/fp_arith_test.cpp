// CPU check of image_matching_b200/csrc/hx_fpmod.cuh against exact __int128 arithmetic: the FP64 modular product and
// normalisation the key-switch kernels use for the 45-bit primes (built with -ffp-contract=off by tests/test_host_cabi.py).
#include <cstdio>
#include <cstdlib>
#include <random>
#include "../../image_matching_b200/csrc/hx_fpmod.cuh"

typedef __int128 i128;
static long long modc(i128 x, long long q) { long long r = (long long)(x % q); return r < 0 ? r + q : r; }

int main() {
    // 45/46-bit primes = 1 mod 2^16 as the parameter recipe generates them, plus the extremes of the admitted range
    const long long primes[] = {35184372744193LL, 35184371138561LL, 35184373006337LL, 70368743489537LL /* < 2^46 */, 17592186028033LL /* 2^44 + */};
    std::mt19937_64 rng(12345);
    long long bad = 0, n = 0;
    double worst = 0;
    for (long long q : primes) {
        const double qd = (double)q, qinv = 1.0 / qd;
        for (int it = 0; it < 2000000; it++) {
            long long a, b = (long long)(rng() % (unsigned long long)q);
            switch (it & 7) {
                case 0: a = (long long)(rng() % (unsigned long long)q); break;                       // canonical
                case 1: a = (long long)(rng() % (1ull << 51)) - (1ll << 50); break;                // lazy, centred, |a| < 2^50
                case 2: a = (1ll << 51) - 1 - (long long)(rng() % 1024); break;                     // top of the admitted range
                case 3: a = -((1ll << 51) - 1) + (long long)(rng() % 1024); break;
                case 4: a = q - 1 - (it % 3); b = q - 1 - (it % 5); break;
                case 5: a = (long long)(rng() % 3); break;
                case 6: a = (long long)(rng() % (unsigned long long)q); b = (long long)(rng() % 3); break;
                default: a = -(long long)(rng() % (unsigned long long)q); break;
            }
            const double m = hx::fp_mulmod((double)a, (double)b, qd, qinv);
            const long long mi = (long long)m;
            n++;
            if ((double)mi != m || modc((i128)mi, q) != modc((i128)a * b, q)) { bad++; if (bad < 5) printf("mulmod mismatch q=%lld a=%lld b=%lld m=%.1f\n", q, a, b, m); }
            const double rel = (m < 0 ? -m : m) / qd;
            if (rel > worst) worst = rel;
            // sum of three products (one key-switch output), minus a lazy NTT word, plus a canonical addend, normalised
            const long long a2 = (long long)(rng() % (unsigned long long)q), b2 = (long long)(rng() % (unsigned long long)q);
            const long long a3 = (long long)(rng() % (1ull << 50)), b3 = (long long)(rng() % (unsigned long long)q);
            const long long conv = (long long)(rng() % (1ull << 50)) - (1ll << 49), add = (long long)(rng() % (unsigned long long)q);
            double acc = m;
            acc = acc + hx::fp_mulmod((double)a2, (double)b2, qd, qinv);
            acc = acc + hx::fp_mulmod((double)a3, (double)b3, qd, qinv);
            acc = acc - (double)conv + hx::fp_from_u64((uint64_t)add);
            const uint64_t r = hx::fp_to_u64(hx::fp_normalize(acc, qd, qinv));
            const long long want = modc((i128)a * b + (i128)a2 * b2 + (i128)a3 * b3 - conv + add, q);
            if ((long long)r != want) { bad++; if (bad < 5) printf("chain mismatch q=%lld got %llu want %lld\n", q, (unsigned long long)r, want); }
        }
    }
    printf("fp_arith_test: %lld cases, %lld mismatches, max |m|/q = %.4f\n", n, bad, worst);
    return bad ? 1 : 0;
}
