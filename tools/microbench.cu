// microbench.cu -- per-SM issue rates of the integer / FP64 instructions the modular arithmetic is
// built from (IMAD, IMAD.WIDE.U32, IMAD.HI, IADD3, DFMA), measured on the B200 this runs on.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;
constexpr int ITER = 4096, UNROLL = 8;

template <int OP>
__global__ void k(u64* out, u32 a0, u32 b0) {
    u32 a[UNROLL], b = b0 + threadIdx.x;
    u64 w[UNROLL];
    double d[UNROLL], dm = 1.0000001 + b0 * 1e-9, dc = 1e-7;
    for (int i = 0; i < UNROLL; i++) { a[i] = a0 + i + threadIdx.x; w[i] = a0 * 77ull + i; d[i] = 1.0 + i + threadIdx.x; }
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < UNROLL; i++) {
            if (OP == 0) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(b));
            if (OP == 1) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b));
            if (OP == 2) asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(b));
            if (OP == 3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 4) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(dm), "d"(dc));
            if (OP == 5) {   // mixed: one wide mad + one dfma per slot
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(dm), "d"(dc));
            }
            if (OP == 6) {   // mixed: mad.lo + add (fma pipe + alu pipe)
                asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(a[i]) : "r"(b));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(a[(i + 1) % UNROLL]) : "r"(b));
            }
            if (OP == 7) w[i] = __umul64hi(w[i], w[i] | 12345ull) + b;   // 64-bit mul.hi as the compiler emits it
            if (OP == 8) w[i] = w[i] * (w[i] | 12345ull) + b;            // 64-bit mul.lo
        }
    }
    u64 s = 0;
    for (int i = 0; i < UNROLL; i++) s += a[i] + w[i] + (u64)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, int per_slot) {
    int dev = 0, sms = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    u64* out;
    const int threads = 1024, blocks = sms * 2;
    cudaMalloc(&out, sizeof(u64) * threads * blocks);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k<OP><<<blocks, threads>>>(out, 3, 5);
    cudaEventRecord(a);
    k<OP><<<blocks, threads>>>(out, 3, 5);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    double ops = (double)blocks * threads * ITER * UNROLL * per_slot;
    double per_clk_sm = ops / (ms * 1e-3) / (khz * 1e3) / sms;
    printf("%-34s %8.3f ms  %7.1f thread-ops/clk/SM (at %d MHz nominal)\n", name, ms, per_clk_sm, khz / 1000);
    cudaFree(out);
}
int main() {
    run<0>("mad.lo.u32 (IMAD)", 1);
    run<1>("mad.wide.u32 (IMAD.WIDE)", 1);
    run<2>("mad.hi.u32 (IMAD.HI)", 1);
    run<3>("add.u32 (IADD3)", 1);
    run<4>("fma.rn.f64 (DFMA)", 1);
    run<5>("IMAD.WIDE + DFMA interleaved", 2);
    run<6>("IMAD + IADD interleaved", 2);
    run<7>("__umul64hi (64x64 high)", 1);
    run<8>("64-bit mul.lo", 1);
    return 0;
}
