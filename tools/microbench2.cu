// microbench2.cu -- compiler-generated vs hand-scheduled 64-bit multiply primitives (issue rates per SM).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;
constexpr int ITER = 2048, UNROLL = 8;

__device__ __forceinline__ u64 madwide(u32 a, u32 b, u64 c) {
    u64 r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u32 lo32(u64 x) { return (u32)x; }
__device__ __forceinline__ u32 hi32(u64 x) { return (u32)(x >> 32); }
// 64x64 -> 128 with four carry-free IMAD.WIDE (each a*b + c fits 64 bits)
__device__ __forceinline__ void mul128_hand(u64 a, u64 b, u64& lo, u64& hi) {
    const u32 a0 = lo32(a), a1 = hi32(a), b0 = lo32(b), b1 = hi32(b);
    const u64 p0 = madwide(a0, b0, 0);
    const u64 p1 = madwide(a0, b1, (u64)hi32(p0));
    const u64 p2 = madwide(a1, b0, (u64)lo32(p1));
    const u64 p3 = madwide(a1, b1, (u64)hi32(p1) + (u64)hi32(p2));
    lo = ((u64)lo32(p2) << 32) | lo32(p0);
    hi = p3;
}
__device__ __forceinline__ u64 umulhi_hand(u64 a, u64 b) {
    u64 lo, hi;
    mul128_hand(a, b, lo, hi);
    return hi;
}
__device__ __forceinline__ u64 mullo_hand(u64 a, u64 b) {
    const u32 a0 = lo32(a), a1 = hi32(a), b0 = lo32(b), b1 = hi32(b);
    const u64 p0 = madwide(a0, b0, 0);
    const u32 h = a0 * b1 + a1 * b0 + hi32(p0);
    return ((u64)h << 32) | lo32(p0);
}
__device__ __forceinline__ u64 shoup_c(u64 a, u64 w, u64 wsh, u64 q) { return a * w - __umul64hi(a, wsh) * q; }
__device__ __forceinline__ u64 shoup_h(u64 a, u64 w, u64 wsh, u64 q) { return mullo_hand(a, w) - mullo_hand(umulhi_hand(a, wsh), q); }

// approximate high product: drops a_lo*b_lo and the carries of the cross terms; result in {hi-2, hi-1, hi}
__device__ __forceinline__ u64 umulhi_approx(u64 a, u64 b) {
    const u32 a0 = lo32(a), a1 = hi32(a), b0 = lo32(b), b1 = hi32(b);
    const u64 s1 = madwide(a0, b1, 0);
    const u64 s2 = madwide(a1, b0, 0);
    return madwide(a1, b1, (u64)hi32(s1) + (u64)hi32(s2));
}
__device__ __forceinline__ u64 umulhi_approx2(u64 a, u64 b) {   // same, chained accumulate instead of the 33-bit add
    const u32 a0 = lo32(a), a1 = hi32(a), b0 = lo32(b), b1 = hi32(b);
    const u64 s1 = madwide(a0, b1, 0);
    const u64 s2 = madwide(a1, b0, (u64)hi32(s1));          // hi32(s2) = hi32(a1*b0 + hi32(s1)) : within 1 of the split sum
    return madwide(a1, b1, 0) + (u64)hi32(s2) + 0;            // but loses hi32(s1) weight: NOT equivalent, timing probe only
}
__device__ __forceinline__ u64 shoup_a(u64 a, u64 w, u64 wsh, u64 q) { return a * w - umulhi_approx(a, wsh) * q; }
// 128-bit MAC from one shared set of partial products, asm-chained
__device__ __forceinline__ void mac128_asm(u64& alo, u64& ahi, u64 a, u64 b) {
    const u32 a0 = lo32(a), a1 = hi32(a), b0 = lo32(b), b1 = hi32(b);
    u64 p0 = madwide(a0, b0, 0), p1 = madwide(a0, b1, 0), p2 = madwide(a1, b0, 0), p3 = madwide(a1, b1, 0);
    // acc += p0 + ((p1 + p2) << 32) + (p3 << 64)
    asm("{\n\t.reg .u32 m0, m1, m2, z;\n\t.reg .u32 l0, l1, h0, h1, x0, x1, y0, y1, r0, r1, r2, r3;\n\t"
        "mov.b64 {l0, l1}, %0;\n\tmov.b64 {h0, h1}, %1;\n\t"
        "mov.b64 {x0, x1}, %3;\n\tmov.b64 {y0, y1}, %4;\n\t"
        "add.cc.u32 m0, x0, y0;\n\taddc.cc.u32 m1, x1, y1;\n\taddc.u32 m2, 0, 0;\n\t"     // mid = p1 + p2 (65 bits)
        "mov.b64 {r0, r1}, %2;\n\tmov.b64 {r2, r3}, %5;\n\t"
        "add.cc.u32 r1, r1, m0;\n\taddc.cc.u32 r2, r2, m1;\n\taddc.u32 r3, r3, m2;\n\t"  // prod = p0 + mid<<32 + p3<<64
        "add.cc.u32 l0, l0, r0;\n\taddc.cc.u32 l1, l1, r1;\n\taddc.cc.u32 h0, h0, r2;\n\taddc.u32 h1, h1, r3;\n\t"
        "mov.b64 %0, {l0, l1};\n\tmov.b64 %1, {h0, h1};\n\t}"
        : "+l"(alo), "+l"(ahi) : "l"(p0), "l"(p1), "l"(p2), "l"(p3));
}

template <int OP>
__global__ void k(u64* out, u64 s0, u64 q) {
    u64 w[UNROLL], acc_lo[UNROLL], acc_hi[UNROLL];
    const u64 m = s0 * 0x9E3779B97F4A7C15ull | 1, msh = ~m;
    for (int i = 0; i < UNROLL; i++) { w[i] = s0 * 77ull + i + threadIdx.x; acc_lo[i] = acc_hi[i] = i; }
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < UNROLL; i++) {
            if (OP == 0) w[i] = __umul64hi(w[i], m) + 1;
            if (OP == 1) w[i] = umulhi_hand(w[i], m) + 1;
            if (OP == 2) w[i] = shoup_c(w[i], m, msh, q);
            if (OP == 3) w[i] = shoup_h(w[i], m, msh, q);
            if (OP == 4) {   // compiler 128-bit MAC
                u64 lo = w[i] * m, hi = __umul64hi(w[i], m);
                asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(acc_lo[i]), "+l"(acc_hi[i]) : "l"(lo), "l"(hi));
                w[i] += it;
            }
            if (OP == 5) {   // hand 128-bit MAC
                u64 lo, hi;
                mul128_hand(w[i], m, lo, hi);
                asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, %3;" : "+l"(acc_lo[i]), "+l"(acc_hi[i]) : "l"(lo), "l"(hi));
                w[i] += it;
            }
            if (OP == 8) w[i] = umulhi_approx(w[i], m) + 1;
            if (OP == 9) w[i] = shoup_a(w[i], m, msh, q);
            if (OP == 10) { mac128_asm(acc_lo[i], acc_hi[i], w[i], m); w[i] += it; }
            if (OP == 11) w[i] = umulhi_approx2(w[i], m) + 1;
            if (OP == 6) w[i] = w[i] * m + 1;
            if (OP == 7) w[i] = mullo_hand(w[i], m) + 1;
        }
    }
    u64 s = 0;
    for (int i = 0; i < UNROLL; i++) s += w[i] + acc_lo[i] + acc_hi[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
void run(const char* name) {
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    u64* out;
    const int threads = 512, blocks = sms * 4;
    cudaMalloc(&out, sizeof(u64) * threads * blocks);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k<OP><<<blocks, threads>>>(out, 3, (1ull << 59) + 12345);
    cudaEventRecord(a);
    k<OP><<<blocks, threads>>>(out, 3, (1ull << 59) + 12345);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    double ops = (double)blocks * threads * ITER * UNROLL;
    double per = ops / (ms * 1e-3) / (khz * 1e3) / sms;
    printf("%-36s %8.3f ms  %6.2f ops/clk/SM  = %5.1f IMAD-slots each (64/clk/SM)\n", name, ms, per, 64.0 / per);
    cudaFree(out);
}
int main() {
    run<0>("__umul64hi (compiler)");
    run<1>("umulhi (4 carry-free IMAD.WIDE)");
    run<2>("shoup_lazy (compiler)");
    run<3>("shoup_lazy (hand)");
    run<4>("128-bit MAC (compiler)");
    run<5>("128-bit MAC (hand)");
    run<6>("64-bit mul.lo (compiler)");
    run<7>("64-bit mul.lo (hand)");
    run<8>("umulhi approx (3 IMAD.WIDE)");
    run<9>("shoup_lazy with approx hi");
    run<10>("128-bit MAC (4 WIDE + asm carry chain)");
    run<11>("umulhi probe (chained)");
    return 0;
}
