// hx_nt.cpp -- see hx_nt.h.
#include "hx_nt.h"
#include <stdexcept>

namespace hx {

u64 powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
    }
    return r;
}

bool is_prime(u64 n) {
    if (n < 2) return false;
    for (u64 b : {2ull, 3ull, 5ull, 7ull, 11ull, 13ull, 17ull, 19ull, 23ull, 29ull, 31ull, 37ull}) {
        if (n == b) return true;
        if (n % b == 0) return false;
    }
    u64 d = n - 1;
    int s = 0;
    while ((d & 1) == 0) { d >>= 1; ++s; }
    // this witness set is exact for every n < 2^64
    for (u64 b : {2ull, 3ull, 5ull, 7ull, 11ull, 13ull, 17ull, 19ull, 23ull, 29ull, 31ull, 37ull}) {
        u64 x = powmod(b, d, n);
        if (x == 1 || x == n - 1) continue;
        bool witness = true;
        for (int r = 1; r < s && witness; ++r) {
            x = mulmod(x, x, n);
            if (x == n - 1) witness = false;
        }
        if (witness) return false;
    }
    return true;
}

u64 first_prime_above(int bits, u64 step) {
    u64 base = (u64)1 << bits;
    u64 rem = base % step;
    u64 cand = base + (rem ? step - rem : 0) + 1;
    while (!is_prime(cand)) cand += step;
    return cand;
}
u64 next_prime(u64 q, u64 step) { do q += step; while (!is_prime(q)); return q; }
u64 prev_prime(u64 q, u64 step) { do q -= step; while (!is_prime(q)); return q; }

u64 min_primitive_root(u64 order, u64 q) {
    // order is a power of two: g primitive <=> g^(order/2) = -1; all primitive roots are odd powers of g.
    u64 cof = (q - 1) / order, g = 0;
    for (u64 x = 2;; ++x) {
        g = powmod(x, cof, q);
        if (powmod(g, order >> 1, q) == q - 1) break;
    }
    u64 gg = mulmod(g, g, q), cur = g, best = g;
    for (u64 k = 1; k < order; k += 2) {
        if (cur < best) best = cur;
        cur = mulmod(cur, gg, q);
    }
    return best;
}

static int bitlen_product(const std::vector<u64>& v) {
    // bit length of the product, via multi-word accumulation
    std::vector<u64> acc{1};
    for (u64 x : v) {
        u64 carry = 0;
        for (auto& w : acc) {
            u128 t = (u128)w * x + carry;
            w = (u64)t;
            carry = (u64)(t >> 64);
        }
        if (carry) acc.push_back(carry);
    }
    int top = 64 - __builtin_clzll(acc.back());
    return (int)(acc.size() - 1) * 64 + top;
}

ParamSet generate_params(int depth, int scale_bits, int first_mod_bits, int ring_dim, int num_large_digits) {
    static const int kStdBits[][2] = {{1024, 27}, {2048, 54}, {4096, 109}, {8192, 218}, {16384, 438},
                                      {32768, 881}, {65536, 1747}, {131072, 3523}};
    ParamSet P;
    P.scale_bits = scale_bits;
    P.L = depth + 1;
    int dnum = num_large_digits ? num_large_digits : (depth > 3 ? 3 : (depth > 0 ? 2 : 1));
    P.alpha = (P.L + dnum - 1) / dnum;
    P.dnum = (P.L + P.alpha - 1) / P.alpha;

    auto choose = [&](int N) {
        u64 m = 2ull * N;
        P.q.assign(P.L, 0);
        P.q[P.L - 1] = first_prime_above(scale_bits, m);
        u64 below = P.q[P.L - 1], above = below;
        bool down = true;
        for (int i = P.L - 2; i >= 1; --i, down = !down) {
            if (down) { below = prev_prime(below, m); P.q[i] = below; }
            else { above = next_prime(above, m); P.q[i] = above; }
        }
        P.q[0] = (first_mod_bits == scale_bits) ? prev_prime(below, m)
                                                : prev_prime(first_prime_above(first_mod_bits, m), m);
        int maxbits = 0;
        for (int j = 0; j < P.dnum; ++j) {
            std::vector<u64> part;
            for (int i = j * P.alpha; i < P.L && i < (j + 1) * P.alpha; ++i) part.push_back(P.q[i]);
            int b = bitlen_product(part);
            if (b > maxbits) maxbits = b;
        }
        P.K = (maxbits + 59) / 60;
        P.p.clear();
        u64 cur = first_prime_above(60, m);
        for (int i = 0; i < P.K; ++i) {
            bool clash;
            do {
                cur = prev_prime(cur, m);
                clash = false;
                for (u64 x : P.q) clash |= (x == cur);
            } while (clash);
            P.p.push_back(cur);
        }
    };

    if (ring_dim) {
        P.N = ring_dim;
    } else {
        // StdLatticeParm::FindRingDim on the NOMINAL budget qBound = firstModSize + (L-1)*scalingModSize + logP, with logP from
        // CryptoParametersRNS::EstimateLogP(..., addOne = true): (largest digit's nominal bits + 1) rounded up to whole 60-bit
        // auxiliary primes.  The estimate -- not the generated primes -- picks N: depth 12 gives 60 + 12*45 + 300 = 900 > 881,
        // hence N = 2^16 (pinned by /root/reference/tools/figures/15IndexTotals.csv:2,5 for approach 3).  The number of P limbs
        // (choose) still follows PrecomputeCRTTables: ceil(bits(largest digit product) / 60) on the generated primes.
        int maxbits = 0;
        for (int j = 0; j < P.dnum; ++j) {
            int b = 0;
            for (int i = j * P.alpha; i < P.L && i < (j + 1) * P.alpha; ++i) b += (i == 0 ? first_mod_bits : scale_bits);
            if (b > maxbits) maxbits = b;
        }
        const int size_p = (maxbits + 1 + 59) / 60;
        const int qbound = first_mod_bits + (P.L - 1) * scale_bits + size_p * 60;
        P.N = 0;
        for (auto& row : kStdBits)
            if (qbound <= row[1]) { P.N = row[0]; break; }
        if (!P.N) throw std::runtime_error("no ring dimension satisfies the security bound");
    }
    choose(P.N);
    for (u64 x : P.q) P.psi_q.push_back(min_primitive_root(2ull * P.N, x));
    for (u64 x : P.p) P.psi_p.push_back(min_primitive_root(2ull * P.N, x));
    return P;
}

}  // namespace hx
