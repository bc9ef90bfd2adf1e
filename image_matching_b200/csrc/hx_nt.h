// hx_nt.h -- host-side 64-bit number theory for the B200 CKKS engine (product code).
// Prime chain / root selection follows OpenFHE v1.2.3's FIXEDMANUAL + HYBRID recipe as used by
// /root/reference/src/main.cpp:169-175 (GenCryptoContext); see DESIGN.md "Parameter recipe".
#pragma once
#include <cstdint>
#include <vector>

namespace hx {
using u64 = unsigned long long;   // same width as uint64_t; matches CUDA 64-bit intrinsics
using u128 = unsigned __int128;

inline u64 mulmod(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }
inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
u64 powmod(u64 a, u64 e, u64 q);
inline u64 invmod(u64 a, u64 q) { return powmod(a % q, q - 2, q); }   // q prime
bool is_prime(u64 n);
u64 first_prime_above(int bits, u64 step);      // smallest prime > 2^bits, = 1 mod step
u64 next_prime(u64 q, u64 step);
u64 prev_prime(u64 q, u64 step);
u64 min_primitive_root(u64 order, u64 q);       // smallest primitive order-th root of unity mod q
inline u64 shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
inline int ilog2(u64 n) { int l = 0; while (((u64)1 << l) < n) l++; return l; }
inline u64 bitrev(u64 x, int bits) { u64 r = 0; for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; } return r; }

struct ParamSet {
    int N = 0, L = 0, K = 0, dnum = 0, alpha = 0, scale_bits = 45;
    std::vector<u64> q, p, psi_q, psi_p;
};
// depth -> L = depth+1 limbs; ring_dim = 0 selects the smallest 128-bit-classic ring.
ParamSet generate_params(int depth, int scale_bits, int first_mod_bits, int ring_dim, int num_large_digits);
}  // namespace hx
