/*
 * oracle/ckks_ref.c -- CPU restatement (plain C, unsigned __int128) of the RNS polynomial
 * arithmetic that the reference reaches through OpenFHE v1.2.3 on the HyDia Sender hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it.  The product path
 * (image_matching_b200/csrc) never links or calls anything in this directory.
 *
 * PARITY STATUS: "parity unpinned" at the polynomial level.  The arithmetic lives in the
 * un-vendored dependency openfheorg/openfhe-development @ v1.2.3 (/root/reference/dockerfile:7,
 * /root/reference/CMakeLists.txt:23); the reference holds no ciphertext-level golden vector.
 * The oracle is pinned at the decrypted level only (tests/test_oracle_*.py): signApprox.csv,
 * test/2_10.dat membership/index, |score - cosine| <= 1e-4.
 *
 * Each function names the upstream routine whose published algorithm it restates and the
 * reference call site (file:line under /root/reference) that reaches it.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef uint64_t u64;
typedef unsigned __int128 u128;

/* OpenMP team size of the CALLING thread for the limb-parallel loops below.  The timed CPU arm (oracle/cpu_path.py) restates the
 * reference's own parallel structure -- "#pragma omp parallel for num_threads(MAX_NUM_CORES)" over rotations / products / blocks
 * (sender_diag.cpp:23,40,57,70) -- with one host thread per loop iteration, each of which runs these primitives single-threaded. */
void ref_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n < 1 ? 1 : n);
#else
    (void)n;
#endif
}
int ref_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---------------------------------------------------------------- scalar number theory */

/* NativeInteger::ModMul -- exact a*b mod q (upstream: math/hal/intnat/ubintnat.h). */
u64 ref_mulmod(u64 a, u64 b, u64 q) { return (u64)(((u128)a * b) % q); }

u64 ref_powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    while (e) {
        if (e & 1) r = ref_mulmod(r, a, q);
        a = ref_mulmod(a, a, q);
        e >>= 1;
    }
    return r;
}

/* q prime => a^(q-2) */
u64 ref_invmod(u64 a, u64 q) { return ref_powmod(a, q - 2, q); }

/* MillerRabinPrimalityTest (upstream nbtheory.cpp); deterministic base set for 64-bit inputs. */
int ref_is_prime(u64 n) {
    static const u64 bases[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (n < 2) return 0;
    for (int i = 0; i < 12; i++) {
        if (n == bases[i]) return 1;
        if (n % bases[i] == 0) return 0;
    }
    u64 d = n - 1;
    int s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (int i = 0; i < 12; i++) {
        u64 x = ref_powmod(bases[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) {
            x = ref_mulmod(x, x, n);
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* FirstPrime(nBits, m): smallest prime q = 1 (mod m) with q > 2^nBits (upstream nbtheory-impl.h).
 * Reached from GenCryptoContext, /root/reference/src/main.cpp:175. */
u64 ref_first_prime(int bits, u64 m) {
    u64 q = ((u64)1 << bits);
    u64 r = q % m;
    q = q + (r ? (m - r) : 0) + 1;
    while (!ref_is_prime(q)) q += m;
    return q;
}
/* NextPrime / PreviousPrime(q, m): step by m until prime. */
u64 ref_next_prime(u64 q, u64 m) {
    do { q += m; } while (!ref_is_prime(q));
    return q;
}
u64 ref_prev_prime(u64 q, u64 m) {
    do { q -= m; } while (!ref_is_prime(q));
    return q;
}

/* RootOfUnity(m, q): the MINIMUM primitive m-th root of unity mod q (upstream nbtheory-impl.h
 * scans all coprime powers of some primitive root and keeps the smallest). m is a power of two
 * here, so the primitive m-th roots are g^k for odd k. */
u64 ref_min_root_of_unity(u64 m, u64 q) {
    u64 e = (q - 1) / m;
    u64 g = 0;
    for (u64 x = 2;; x++) {
        g = ref_powmod(x, e, q);
        /* primitive iff g^(m/2) == -1 */
        if (ref_powmod(g, m / 2, q) == q - 1) break;
    }
    u64 g2 = ref_mulmod(g, g, q);
    u64 cur = g, best = g;
    for (u64 k = 1; k < m; k += 2) {
        if (cur < best) best = cur;
        cur = ref_mulmod(cur, g2, q);
    }
    return best;
}

static inline u64 brev(u64 x, int bits) {
    u64 r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}
static inline int ilog2(u64 n) { int l = 0; while (((u64)1 << l) < n) l++; return l; }

/* ---------------------------------------------------------------- context: per-prime NTT tables */

typedef struct {
    u64 N;
    int logN;
    int nprimes;        /* Q primes first, then P primes (OpenFHE limb order) */
    u64* mod;           /* [nprimes] */
    u64* psi;           /* [nprimes] */
    u64* fwd;           /* [nprimes][N]  psi^brev(i)   (rootOfUnityReverse table) */
    u64* inv;           /* [nprimes][N]  psi^-brev(i)  */
    u64* fwdp;          /* Shoup quotients floor(w*2^64/q) (NativeInteger::PrepModMulConst) */
    u64* invp;
    u64* ninv;          /* [nprimes]     N^-1 */
} ref_ctx;

ref_ctx* ref_ctx_new(u64 N, int nprimes, const u64* mod, const u64* psi) {
    ref_ctx* c = (ref_ctx*)calloc(1, sizeof(ref_ctx));
    c->N = N; c->logN = ilog2(N); c->nprimes = nprimes;
    c->mod = (u64*)malloc(sizeof(u64) * nprimes);
    c->psi = (u64*)malloc(sizeof(u64) * nprimes);
    c->ninv = (u64*)malloc(sizeof(u64) * nprimes);
    c->fwd = (u64*)malloc(sizeof(u64) * nprimes * N);
    c->inv = (u64*)malloc(sizeof(u64) * nprimes * N);
    c->fwdp = (u64*)malloc(sizeof(u64) * nprimes * N);
    c->invp = (u64*)malloc(sizeof(u64) * nprimes * N);
    for (int p = 0; p < nprimes; p++) {
        u64 q = mod[p];
        c->mod[p] = q; c->psi[p] = psi[p];
        c->ninv[p] = ref_invmod(N % q, q);
        u64 psi_inv = ref_invmod(psi[p], q);
        u64 x = 1, y = 1;
        for (u64 i = 0; i < N; i++) {
            u64 r = brev(i, c->logN);
            c->fwd[(size_t)p * N + r] = x;
            c->inv[(size_t)p * N + r] = y;
            c->fwdp[(size_t)p * N + r] = (u64)(((u128)x << 64) / q);
            c->invp[(size_t)p * N + r] = (u64)(((u128)y << 64) / q);
            x = ref_mulmod(x, psi[p], q);
            y = ref_mulmod(y, psi_inv, q);
        }
    }
    return c;
}
void ref_ctx_free(ref_ctx* c) {
    if (!c) return;
    free(c->mod); free(c->psi); free(c->ninv); free(c->fwd); free(c->inv); free(c->fwdp); free(c->invp); free(c);
}

/* NativeInteger::ModMulFastConst: a*w mod q with the precomputed quotient wp = floor(w*2^64/q). */
static inline u64 mulmod_shoup(u64 a, u64 w, u64 wp, u64 q) {
    u64 hi = (u64)(((u128)a * wp) >> 64);
    u64 r = a * w - hi * q;
    return r >= q ? r - q : r;
}

/* ChineseRemainderTransformFTT::ForwardTransformToBitReverse (upstream transformnat-impl.h):
 * Cooley-Tukey, natural-order input, bit-reversed output; a[i] <- f(psi^(2*brev(i)+1)). */
static void ntt_fwd_one(u64* a, u64 N, u64 q, const u64* w, const u64* wp) {
    u64 t = N;
    for (u64 m = 1; m < N; m <<= 1) {
        t >>= 1;
        for (u64 i = 0; i < m; i++) {
            u64 j1 = 2 * i * t, j2 = j1 + t;
            u64 S = w[m + i], Sp = wp[m + i];
            for (u64 j = j1; j < j2; j++) {
                u64 U = a[j];
                u64 V = mulmod_shoup(a[j + t], S, Sp, q);
                u64 s = U + V; if (s >= q) s -= q;
                u64 d = U >= V ? U - V : U + q - V;
                a[j] = s; a[j + t] = d;
            }
        }
    }
}
/* InverseTransformFromBitReverse: Gentleman-Sande with psi^-1 table, then N^-1 scaling. */
static void ntt_inv_one(u64* a, u64 N, u64 q, const u64* w, const u64* wp, u64 ninv) {
    u64 t = 1;
    for (u64 m = N; m > 1; m >>= 1) {
        u64 h = m >> 1, j1 = 0;
        for (u64 i = 0; i < h; i++) {
            u64 j2 = j1 + t;
            u64 S = w[h + i], Sp = wp[h + i];
            for (u64 j = j1; j < j2; j++) {
                u64 U = a[j], V = a[j + t];
                u64 s = U + V; if (s >= q) s -= q;
                u64 d = U >= V ? U - V : U + q - V;
                a[j] = s; a[j + t] = mulmod_shoup(d, S, Sp, q);
            }
            j1 += 2 * t;
        }
        t <<= 1;
    }
    u64 np = (u64)(((u128)ninv << 64) / q);
    for (u64 j = 0; j < N; j++) a[j] = mulmod_shoup(a[j], ninv, np, q);
}

/* DCRTPoly::SetFormat over a set of limbs.  a: [nl][N]; idx[l] = prime index of limb l. */
void ref_ntt(const ref_ctx* c, u64* a, int nl, const int* idx, int inverse) {
    #pragma omp parallel for schedule(dynamic)
    for (int l = 0; l < nl; l++) {
        int p = idx[l];
        if (inverse) ntt_inv_one(a + (size_t)l * c->N, c->N, c->mod[p], c->inv + (size_t)p * c->N, c->invp + (size_t)p * c->N, c->ninv[p]);
        else ntt_fwd_one(a + (size_t)l * c->N, c->N, c->mod[p], c->fwd + (size_t)p * c->N, c->fwdp + (size_t)p * c->N);
    }
}

/* ---------------------------------------------------------------- limb-wise elementwise ops
 * All arrays are [nl][N], limb-major, canonical residues in [0, q_l). mod[l] = modulus of limb l. */

/* DCRTPoly::operator+ / EvalAddCoreInPlace (sender_diag.cpp:76 EvalAddInPlace). */
void ref_add(u64* out, const u64* a, const u64* b, int nl, u64 N, const u64* mod) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l];
        for (u64 j = 0; j < N; j++) {
            u64 s = a[l * N + j] + b[l * N + j];
            out[l * N + j] = s >= q ? s - q : s;
        }
    }
}
void ref_sub(u64* out, const u64* a, const u64* b, int nl, u64 N, const u64* mod) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l];
        for (u64 j = 0; j < N; j++) {
            u64 x = a[l * N + j], y = b[l * N + j];
            out[l * N + j] = x >= y ? x - y : x + q - y;
        }
    }
}
/* DCRTPoly::operator* in EVALUATION format (EvalMultCore, sender_diag.cpp:93). */
void ref_mul(u64* out, const u64* a, const u64* b, int nl, u64 N, const u64* mod) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l];
        for (u64 j = 0; j < N; j++) out[l * N + j] = (u64)(((u128)a[l * N + j] * b[l * N + j]) % q);
    }
}
/* out += a*b  (used for key-switch inner products, EvalFastKeySwitchCoreExt). */
void ref_mul_acc(u64* out, const u64* a, const u64* b, int nl, u64 N, const u64* mod) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l];
        for (u64 j = 0; j < N; j++) {
            u64 p = (u64)(((u128)a[l * N + j] * b[l * N + j]) % q);
            u64 s = out[l * N + j] + p;
            out[l * N + j] = s >= q ? s - q : s;
        }
    }
}
/* DCRTPoly * vector<NativeInteger>: per-limb scalar (EvalMultCoreInPlace(ct, double)). */
void ref_mul_scalar(u64* out, const u64* a, const u64* scal, int nl, u64 N, const u64* mod) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l], s = scal[l] % q;
        for (u64 j = 0; j < N; j++) out[l * N + j] = (u64)(((u128)a[l * N + j] * s) % q);
    }
}
/* DCRTPoly + vector<NativeInteger>: per-limb constant added to every evaluation point
 * (EvalAddInPlace(ct, double), openFHE_wrapper.cpp:182). */
void ref_add_scalar(u64* out, const u64* a, const u64* scal, int nl, u64 N, const u64* mod) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l], s = scal[l] % q;
        for (u64 j = 0; j < N; j++) {
            u64 v = a[l * N + j] + s;
            out[l * N + j] = v >= q ? v - q : v;
        }
    }
}

/* PrecomputeAutoMap (upstream utils/utilities / nbtheory2): evaluation-form index map of the
 * Galois automorphism X -> X^k.  Reached from EvalFastRotation (sender_diag.cpp:25). */
void ref_automap(u64 N, u64 k, uint32_t* map) {
    int logn = ilog2(N);
    u64 m = N << 1;
    for (u64 j = 0; j < N; j++) {
        u64 jt = (j << 1) + 1;
        u64 idx = ((jt * k) & (m - 1)) >> 1;
        map[brev(j, logn)] = (uint32_t)brev(idx, logn);
    }
}
/* DCRTPoly::AutomorphismTransform(k, map): out[j] = in[map[j]] on every limb. */
void ref_automorph(u64* out, const u64* in, int nl, u64 N, const uint32_t* map) {
    #pragma omp parallel for
    for (int l = 0; l < nl; l++)
        for (u64 j = 0; j < N; j++) out[l * N + j] = in[l * N + map[j]];
}

/* DCRTPoly::ApproxSwitchCRTBasis (upstream dcrtpoly-impl.h): coefficient-domain fast base
 * conversion without overflow correction:
 *   out_j = ( sum_i [x_i * hatinv_i]_{q_i} * (Qhat_i mod p_j) ) mod p_j
 * in: [n_in][N] with moduli qin; out: [n_out][N] with moduli pout;
 * hatinv[i] = (Q/q_i)^-1 mod q_i ; hatmod[i*n_out + j] = (Q/q_i) mod p_j.
 * Reached from EvalKeySwitchPrecomputeCore (ModUp) and ApproxModDown. */
void ref_baseconv(u64* out, const u64* in, int n_in, const u64* qin, int n_out, const u64* pout,
                  u64 N, const u64* hatinv, const u64* hatmod) {
    #pragma omp parallel for
    for (u64 r = 0; r < N; r++) {
        u64 y[64];
        for (int i = 0; i < n_in; i++) y[i] = (u64)(((u128)in[(size_t)i * N + r] * hatinv[i]) % qin[i]);
        for (int j = 0; j < n_out; j++) {
            u128 acc = 0;   /* n_in <= 8 products of < 2^121: no overflow */
            u64 pj = pout[j];
            u64 red = 0;
            for (int i = 0; i < n_in; i++) {
                acc += (u128)y[i] * hatmod[i * n_out + j];
                if ((i & 3) == 3) { red = (u64)((red + acc % pj) % pj); acc = 0; }
            }
            red = (u64)((red + acc % pj) % pj);
            out[(size_t)j * N + r] = red;
        }
    }
}

/* NativeVector::SwitchModulus: centred lift from q_old to q_new (used by DropLastElementAndScale,
 * i.e. RescaleInPlace, sender_diag.cpp:80). v > floor(q_old/2) is treated as v - q_old. */
void ref_switch_modulus(u64* out, const u64* in, u64 N, u64 q_old, u64 q_new) {
    u64 half = q_old >> 1;
    u64 qo_mod = q_old % q_new;
    for (u64 j = 0; j < N; j++) {
        u64 v = in[j];
        u64 r = v % q_new;
        if (v > half) r = r >= qo_mod ? r - qo_mod : r + q_new - qo_mod;
        out[j] = r;
    }
}

/* EvalMultCore on 2-component inputs + accumulate (sender_diag.cpp:93 then :75-77):
 *   acc0 += a0*b0 ; acc1 += a0*b1 + a1*b0 ; acc2 += a1*b1.
 * a, b: [2][nl][N]; acc: [3][nl][N]. */
void ref_tensor_acc(u64* acc, const u64* a, const u64* b, int nl, u64 N, const u64* mod) {
    size_t comp = (size_t)nl * N;
    #pragma omp parallel for
    for (int l = 0; l < nl; l++) {
        u64 q = mod[l];
        for (u64 j = 0; j < N; j++) {
            size_t o = (size_t)l * N + j;
            u64 a0 = a[o], a1 = a[comp + o], b0 = b[o], b1 = b[comp + o];
            u64 d0 = (u64)(((u128)a0 * b0) % q);
            u64 d1 = (u64)((((u128)a0 * b1) % q + ((u128)a1 * b0) % q) % q);
            u64 d2 = (u64)(((u128)a1 * b1) % q);
            acc[o] = (acc[o] + d0) % q;
            acc[comp + o] = (acc[comp + o] + d1) % q;
            acc[2 * comp + o] = (acc[2 * comp + o] + d2) % q;
        }
    }
}
