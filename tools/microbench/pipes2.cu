// pipes2.cu -- second-round integer pipe microbenchmarks with simplification-proof chains.
// Each op works on 8 independent chains; operands depend on loop-carried state so ptxas cannot fold.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
#define ITERS 2048
#define CH 8
#define LOOP(BODY) for (int it = 0; it < ITERS; it++) { _Pragma("unroll") for (int i = 0; i < CH; i++) { BODY } }
template <int OP> __global__ void __launch_bounds__(256) k(u32* out, u32 seed) {
    u32 a[CH], b[CH], c[CH], d[CH];
    for (int i = 0; i < CH; i++) { a[i] = threadIdx.x * 7 + i + seed; b[i] = a[i] * 0x9E3779B9u + 1; c[i] = b[i] ^ 0x1234567; d[i] = c[i] + a[i]; }
    if (OP == 0) LOOP( asm volatile("{.reg .u64 w; mov.b64 w,{%0,%1}; mad.wide.u32 w,%0,%2,w; mov.b64 {%0,%1},w;}" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i])); )  // IMAD.WIDE
    if (OP == 1) LOOP( asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a[i]) : "r"(b[i]), "r"(c[i])); )                         // IMAD
    if (OP == 2) LOOP( asm volatile("add.u32 %0,%0,%1; add.u32 %1,%1,%0;" : "+r"(a[i]), "+r"(b[i])); )                      // 2 IADD
    if (OP == 3) LOOP( asm volatile("add.cc.u32 %0,%0,%2; addc.u32 %1,%1,%0;" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i])); )      // IADD3 P + IADD3.X
    if (OP == 4) LOOP( asm volatile("xor.b32 %0,%0,%1; and.b32 %1,%1,%0; or.b32 %1,%1,%2;" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i])); ) // LOP3
    if (OP == 5) LOOP( asm volatile("{.reg .pred p; setp.gt.u32 p,%0,%1; selp.u32 %0,%1,%2,p; add.u32 %1,%1,%0;}" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i])); ) // ISETP+SEL+IADD
    if (OP == 6) LOOP( asm volatile("{.reg .u64 w; mov.b64 w,{%0,%1}; mad.wide.u32 w,%0,%2,w; mov.b64 {%0,%1},w;}" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i]));
                       asm volatile("add.u32 %0,%0,%1; add.u32 %1,%1,%0;" : "+r"(c[i]), "+r"(d[i])); )                      // WIDE + 2 IADD
    if (OP == 7) LOOP( asm volatile("{.reg .u64 w; mov.b64 w,{%0,%1}; mad.wide.u32 w,%0,%2,w; mov.b64 {%0,%1},w;}" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i]));
                       asm volatile("add.cc.u32 %0,%0,%2; addc.u32 %1,%1,%0;" : "+r"(c[i]), "+r"(d[i]) : "r"(a[i])); )      // WIDE + carry pair
    if (OP == 8) LOOP( asm volatile("{.reg .u64 w; mov.b64 w,{%0,%1}; mad.wide.u32 w,%0,%2,w; mov.b64 {%0,%1},w;}" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i]));
                       asm volatile("xor.b32 %0,%0,%1; and.b32 %1,%1,%0;" : "+r"(c[i]), "+r"(d[i])); )                      // WIDE + 2 LOP3
    if (OP == 9) LOOP( asm volatile("add.cc.u32 %0,%0,%2; addc.cc.u32 %1,%1,%0; addc.u32 %2,%2,%1;" : "+r"(a[i]), "+r"(b[i]), "+r"(c[i])); ) // 3-long carry chain
    if (OP == 10) LOOP( asm volatile("{.reg .u64 w; mov.b64 w,{%0,%1}; mad.wide.u32 w,%0,%2,w; mov.b64 {%0,%1},w;}" : "+r"(a[i]), "+r"(b[i]) : "r"(c[i]));
                        asm volatile("add.cc.u32 %0,%0,%2; addc.cc.u32 %1,%1,%0; addc.u32 %0,%0,%1;" : "+r"(c[i]), "+r"(d[i]) : "r"(a[i])); ) // WIDE + 3 carry ops
    if (OP == 11) LOOP( asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a[i]) : "r"(b[i]), "r"(c[i]));
                        asm volatile("add.u32 %0,%0,%1; add.u32 %1,%1,%0;" : "+r"(c[i]), "+r"(d[i])); )                     // IMAD + 2 IADD
    u32 acc = 0;
    for (int i = 0; i < CH; i++) acc += a[i] + b[i] + c[i] + d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int OP> void run(const char* name, int per_iter, u32* out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = 148 * 8;
    k<OP><<<blocks, 256>>>(out, 1); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<OP><<<blocks, 256>>>(out, 1); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double wi = (double)blocks * 8 * ITERS * CH * per_iter;
    printf("%-26s %7.3f ms  %.3f src-instr/clk/SMSP (@1965MHz)  %.2f clk per chain-iter\n", name, ms,
           wi / (ms * 1e-3) / (148.0 * 4 * 1965e6), (ms * 1e-3) * 148.0 * 4 * 1965e6 / ((double)blocks * 8 * ITERS * CH));
}
int main() {
    u32* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("WIDE", 1, out); run<1>("IMAD", 1, out); run<2>("2xIADD", 2, out); run<3>("IADD.cc+addc", 2, out);
    run<4>("3xLOP3", 3, out); run<5>("ISETP+SEL+IADD", 3, out); run<6>("WIDE+2IADD", 3, out);
    run<7>("WIDE+cc pair", 3, out); run<8>("WIDE+2LOP", 3, out); run<9>("cc chain x3", 3, out);
    run<10>("WIDE+cc x3", 4, out); run<11>("IMAD+2IADD", 3, out);
    return 0;
}
