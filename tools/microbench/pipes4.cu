// pipes4.cu -- is the FP64 pipe of B200 a usable third integer-ish pipe?  DFMA / DADD issue cost alone and mixed
// with IMAD.WIDE and ALU work (same harness as pipes3.cu: clocks per chain iteration per scheduler).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
#define ITERS 2048
#define CH 8
#define LOOP(BODY) for (int it = 0; it < ITERS; it++) { _Pragma("unroll") for (int i = 0; i < CH; i++) { BODY } }
template <int OP> __global__ void __launch_bounds__(256) k(u32* out, u32 seed) {
    u64 w[CH]; u32 a[CH], b[CH], c[CH]; double d[CH], e[CH];
    for (int i = 0; i < CH; i++) { a[i] = threadIdx.x * 7 + i + seed; b[i] = a[i] * 0x9E3779B9u + 1; c[i] = b[i] ^ 0x1234567; w[i] = ((u64)a[i] << 32) | c[i];
        d[i] = (double)(a[i] & 0xFFFF); e[i] = (double)(b[i] & 0xFF) + 1.0; }
    if (OP == 0) LOOP( d[i] = fma(d[i], 3.0, e[i]); )                                              // DFMA
    if (OP == 1) LOOP( d[i] = d[i] + e[i]; )                                                       // DADD
    if (OP == 2) LOOP( d[i] = fma(d[i], 3.0, e[i]); w[i] = (u64)(u32)w[i] * c[i] + w[i]; )         // DFMA + WIDE
    if (OP == 3) LOOP( d[i] = fma(d[i], 3.0, e[i]); a[i] = (a[i] ^ b[i]) & (b[i] | c[i]); )        // DFMA + LOP3
    if (OP == 4) LOOP( d[i] = fma(d[i], 3.0, e[i]); w[i] = (u64)(u32)w[i] * c[i] + w[i]; a[i] = (a[i] ^ b[i]) & (b[i] | c[i]); b[i] ^= a[i]; ) // DFMA + WIDE + 2 LOP3
    if (OP == 5) LOOP( w[i] = (u64)(u32)w[i] * c[i] + w[i]; a[i] = (a[i] ^ b[i]) & (b[i] | c[i]); b[i] ^= a[i]; )                              // WIDE + 2 LOP3 (reference)
    if (OP == 6) LOOP( d[i] = fma(d[i], 3.0, e[i]); e[i] = fma(e[i], 0.5, d[i]); a[i] = a[i] * b[i] + c[i]; b[i] = (b[i] ^ a[i]) & (b[i] | c[i]); ) // 2 DFMA + IMAD + LOP3
    if (OP == 7) LOOP( a[i] = a[i] * b[i] + c[i]; b[i] = (b[i] ^ a[i]) & (b[i] | c[i]); )          // IMAD + LOP3 (reference)
    u32 acc = 0;
    for (int i = 0; i < CH; i++) acc += a[i] + b[i] + c[i] + (u32)w[i] + (u32)(w[i] >> 32) + (u32)(long long)d[i] + (u32)(long long)e[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int OP> void run(const char* name, u32* out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = 148 * 8;
    k<OP><<<blocks, 256>>>(out, 1); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<OP><<<blocks, 256>>>(out, 1); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s %7.3f ms  %.2f clk per chain-iter (@1965MHz)\n", name, ms, (ms * 1e-3) * 148.0 * 4 * 1965e6 / ((double)blocks * 8 * ITERS * CH));
}
int main() {
    u32* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("DFMA", out); run<1>("DADD", out); run<2>("DFMA+WIDE", out); run<3>("DFMA+LOP3", out); run<4>("DFMA+WIDE+2LOP3", out);
    run<5>("WIDE+2LOP3", out); run<6>("2DFMA+IMAD+LOP3", out); run<7>("IMAD+LOP3", out);
    return 0;
}
