// microbench_tc.cu -- the number behind the "no tensor cores" decision for the base conversion of ModDown.
//
// Work: conv[j] = sum_{k<4} y_k * M[k][j] mod q_j for 12 target limbs, y_k < 2^60, for every coefficient of both components of a
// batch of 32 hoisted rotations (/root/reference/src/sender/sender_diag.cpp:25): 32 * 2 * 32768 coefficients, 12 outputs each.
//   A  "imad":     what the engine runs: operands pre-split in 30-bit halves, 4 IMAD.WIDE per product into three carry-free
//                  columns, one 128-bit fold and one Montgomery REDC per output (hx_modarith.cuh).
//   B  "tc-epi":   the int8-split GEMM formulation's EPILOGUE ONLY.  As a u8 x u8 -> s32 GEMM the conversion is
//                  [coefficients x 32 (4 sources x 8 bytes)] . [32 x 180 (12 targets x 15 byte-diagonals, Toeplitz in the bytes of
//                  M)]; every output then needs its 15 s32 partial sums recombined with weights 2^(8 s) into a 128-bit integer and
//                  reduced.  B times exactly that recombination + the same REDC, reading the 15 partial sums from shared memory --
//                  i.e. it charges NOTHING for the tensor-core MMAs, for splitting y into bytes or for moving accumulators out of
//                  TMEM / registers.  It is a lower bound for any tcgen05 / mma.sync version.
//   C  "tc-mma":   B plus the real thing on the legacy tensor path: mma.sync.m16n8k32 u8 x u8 -> s32 producing the 180 columns
//                  from byte-split operands in shared memory (fragment loads included).
// If B alone is not faster than A, no tensor-core version can win.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench_tc microbench_tc.cu && ./microbench_tc
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;

constexpr int kTargets = 12, kSources = 4;
constexpr size_t kCoefs = 32ull * 2 * 32768;   // one 32-rotation batch, both components
constexpr u64 kQ = 35184372744193ull;          // a 45-bit prime = 1 mod 2^16 (limb modulus stand-in)

__device__ __forceinline__ u64 madw(u32 a, u32 b, u64 c) { u64 d; asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c)); return d; }
struct acc3 { u64 lo, mid, hi; };
__device__ __forceinline__ void mac30ss(acc3& a, u64 xs, u64 ys) {
    const u32 x0 = (u32)xs, x1 = (u32)(xs >> 32), y0 = (u32)ys, y1 = (u32)(ys >> 32);
    a.lo = madw(x0, y0, a.lo); a.mid = madw(x0, y1, a.mid); a.mid = madw(x1, y0, a.mid); a.hi = madw(x1, y1, a.hi);
}
__device__ __forceinline__ void fold30(const acc3& a, u64& lo, u64& hi) {
    const u64 m_lo = a.mid << 30, m_hi = a.mid >> 34, h_lo = a.hi << 60, h_hi = a.hi >> 4;
    asm("add.cc.u64 %0, %2, %3;\n\taddc.u64 %1, %4, %5;" : "=l"(lo), "=l"(hi) : "l"(a.lo), "l"(m_lo), "l"(m_hi), "l"(h_hi));
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(lo), "+l"(hi) : "l"(h_lo));
}
__device__ __forceinline__ u64 redc_lazy(u64 lo, u64 hi, u64 q, u64 qinv_neg) {
    const u64 m = lo * qinv_neg;
    return hi + __umul64hi(m, q) + (lo != 0 ? 1ull : 0ull);
}

// A: one thread = one coefficient: 4 sources in registers, 12 outputs
__global__ void __launch_bounds__(256) k_imad(const u64* __restrict__ y, const u64* __restrict__ mat, u64* __restrict__ out, u64 q, u64 qinv_neg) {
    __shared__ u64 sm[kSources * kTargets];
    if (threadIdx.x < kSources * kTargets) sm[threadIdx.x] = mat[threadIdx.x];
    __syncthreads();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    u64 ys[kSources];
#pragma unroll
    for (int k = 0; k < kSources; k++) ys[k] = y[(size_t)k * kCoefs + i];
#pragma unroll
    for (int j = 0; j < kTargets; j++) {
        acc3 a = {0, 0, 0};
#pragma unroll
        for (int k = 0; k < kSources; k++) mac30ss(a, ys[k], sm[k * kTargets + j]);
        u64 lo, hi;
        fold30(a, lo, hi);
        out[(size_t)j * kCoefs + i] = redc_lazy(lo, hi, q, qinv_neg);
    }
}

// 15 byte-diagonal partial sums p[s] (each < 32 * 255^2 < 2^21) -> sum_s p[s] 2^(8 s) as (lo, hi)
__device__ __forceinline__ void recombine15(const int* p, u64& lo, u64& hi) {
    lo = 0; hi = 0;
#pragma unroll
    for (int s = 0; s < 15; s++) {
        const u64 v = (u64)(u32)p[s];
        if (8 * s < 64) {
            const u64 add_lo = v << (8 * s);
            const u64 add_hi = (8 * s == 0) ? 0ull : (v >> (64 - 8 * s));
            asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(lo), "+l"(hi) : "l"(add_lo), "l"(add_hi));
        } else {
            hi += v << (8 * s - 64);
        }
    }
}
// B: epilogue only.  The "accumulators" of a coefficient tile sit in shared memory as the MMA would leave them.
__global__ void __launch_bounds__(256) k_tc_epilogue(const int* __restrict__ partial_seed, u64* __restrict__ out, u64 q, u64 qinv_neg) {
    __shared__ int acc[256][16];   // one target's 15 diagonals per coefficient, refilled per target (cheap stand-in for LDTM / fragments)
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int seed = partial_seed[threadIdx.x];
#pragma unroll 1
    for (int j = 0; j < kTargets; j++) {
#pragma unroll
        for (int s = 0; s < 15; s++) acc[threadIdx.x][s] = (seed * (j + 3) + s * 7919) & 0x1FFFFF;
        __syncwarp();
        u64 lo, hi;
        recombine15(acc[threadIdx.x], lo, hi);
        out[(size_t)j * kCoefs + i] = redc_lazy(lo, hi, q, qinv_neg);
    }
}

// C: the GEMM on mma.sync m16n8k32 (u8 x u8 -> s32), then the same epilogue.  One warp = 16 coefficients x 184 columns (23 n-tiles of 8;
// 180 used).  A fragments from byte-split y in shared memory, B fragments (the Toeplitz matrix, 32 x 184 bytes) from shared memory.
__global__ void __launch_bounds__(256) k_tc_mma(const u64* __restrict__ y, const unsigned char* __restrict__ bmat, u64* __restrict__ out, u64 q, u64 qinv_neg) {
    __shared__ __align__(16) unsigned char sA[8][16][32];      // per warp: 16 coefficients x (4 sources x 8 bytes)
    __shared__ __align__(16) unsigned char sB[184][32];        // column-major: column n holds its 32 k-bytes contiguously
    __shared__ int sC[8][16][16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int t = threadIdx.x; t < 184 * 32 / 16; t += blockDim.x) reinterpret_cast<uint4*>(&sB[0][0])[t] = reinterpret_cast<const uint4*>(bmat)[t];
    const size_t c0 = ((size_t)blockIdx.x * 8 + warp) * 16;
    // byte-split the 16 coefficients' 4 sources: lanes 0..15 -> coefficient, two sources each half-warp pass
    for (int k = lane >> 4; k < kSources; k += 2) {
        const u64 v = y[(size_t)k * kCoefs + c0 + (lane & 15)];
        *reinterpret_cast<u64*>(&sA[warp][lane & 15][k * 8]) = v;
    }
    __syncthreads();
    // A fragment of m16n8k32 (.u8): a0 = (row g, k 4t..4t+3), a1 = (row g+8, same k), a2 = (row g, k 16+4t..), a3 = (row g+8, k 16+4t..)
    const int g = lane >> 2, t = lane & 3;
    u32 a[4];
    a[0] = *reinterpret_cast<const u32*>(&sA[warp][g][4 * t]);
    a[1] = *reinterpret_cast<const u32*>(&sA[warp][g + 8][4 * t]);
    a[2] = *reinterpret_cast<const u32*>(&sA[warp][g][16 + 4 * t]);
    a[3] = *reinterpret_cast<const u32*>(&sA[warp][g + 8][16 + 4 * t]);
#pragma unroll 1
    for (int j = 0; j < kTargets; j++) {
        // target j = columns 15 j .. 15 j + 14 -> two n-tiles of 8 (the second one overlaps the next target; harmless for timing)
#pragma unroll
        for (int nt = 0; nt < 2; nt++) {
            const int n0 = 15 * j + 8 * nt;
            u32 b[2];
            b[0] = *reinterpret_cast<const u32*>(&sB[n0 + g][4 * t]);
            b[1] = *reinterpret_cast<const u32*>(&sB[n0 + g][16 + 4 * t]);
            int c[4] = {0, 0, 0, 0};
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
            // C fragment: c0,c1 = (row g, cols 2t, 2t+1), c2,c3 = (row g+8, same cols)
            if (8 * nt + 2 * t < 16) {
                sC[warp][g][8 * nt + 2 * t] = c[0];
                if (8 * nt + 2 * t + 1 < 16) sC[warp][g][8 * nt + 2 * t + 1] = c[1];
                sC[warp][g + 8][8 * nt + 2 * t] = c[2];
                if (8 * nt + 2 * t + 1 < 16) sC[warp][g + 8][8 * nt + 2 * t + 1] = c[3];
            }
        }
        __syncwarp();
        if (lane < 16) {
            u64 lo, hi;
            recombine15(sC[warp][lane], lo, hi);
            out[(size_t)j * kCoefs + c0 + lane] = redc_lazy(lo, hi, q, qinv_neg);
        }
        __syncwarp();
    }
}

int main() {
    u64 *y, *mat, *out;
    int* seed;
    unsigned char* bmat;
    cudaMalloc(&y, sizeof(u64) * kSources * kCoefs);
    cudaMalloc(&out, sizeof(u64) * kTargets * kCoefs);
    cudaMalloc(&mat, sizeof(u64) * kSources * kTargets);
    cudaMalloc(&seed, sizeof(int) * 256);
    cudaMalloc(&bmat, 184 * 32);
    cudaMemset(y, 0x5A, sizeof(u64) * kSources * kCoefs);
    cudaMemset(mat, 0x33, sizeof(u64) * kSources * kTargets);
    cudaMemset(seed, 0x11, sizeof(int) * 256);
    cudaMemset(bmat, 0x27, 184 * 32);
    u64 inv = 1;
    for (int i = 0; i < 6; i++) inv *= 2 - kQ * inv;
    const u64 qinv_neg = 0 - inv;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const unsigned blocks = (unsigned)(kCoefs / 256);
    auto time = [&](const char* name, auto launch) {
        for (int i = 0; i < 3; i++) launch();
        cudaEventRecord(e0);
        for (int i = 0; i < 20; i++) launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        cudaError_t err = cudaGetLastError();
        printf("%-10s %8.1f us per 32-rotation batch (%s)\n", name, ms / 20 * 1e3, cudaGetErrorString(err));
    };
    time("A imad", [&] { k_imad<<<blocks, 256>>>(y, mat, out, kQ, qinv_neg); });
    time("B tc-epi", [&] { k_tc_epilogue<<<blocks, 256>>>(seed, out, kQ, qinv_neg); });
    time("C tc-mma", [&] { k_tc_mma<<<(unsigned)(kCoefs / 128), 256>>>(y, bmat, out, kQ, qinv_neg); });
    printf("outputs per batch: %zu; bytes written per kernel: %.1f MB (the store stream is common to all three)\n", kTargets * kCoefs, kTargets * kCoefs * 8 / 1e6);
    return 0;
}
