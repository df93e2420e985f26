// pipes3.cu -- IMAD.WIDE issue cost on B200, alone and mixed with ALU-pipe work.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
#define ITERS 2048
#define CH 8
#define LOOP(BODY) for (int it = 0; it < ITERS; it++) { _Pragma("unroll") for (int i = 0; i < CH; i++) { BODY } }
template <int OP> __global__ void __launch_bounds__(256) k(u32* out, u32 seed) {
    u64 w[CH]; u32 a[CH], b[CH], c[CH];
    for (int i = 0; i < CH; i++) { a[i] = threadIdx.x * 7 + i + seed; b[i] = a[i] * 0x9E3779B9u + 1; c[i] = b[i] ^ 0x1234567; w[i] = ((u64)a[i] << 32) | c[i]; }
    if (OP == 0) LOOP( w[i] = (u64)(u32)w[i] * c[i] + w[i]; )                                   // WIDE acc (reg)
    if (OP == 1) LOOP( w[i] = (u64)(u32)w[i] * 41u + w[i]; )                                    // WIDE acc (imm)
    if (OP == 2) LOOP( w[i] = (u64)(u32)(w[i] >> 32) * c[i] + w[i]; )                           // WIDE acc, hi word as multiplicand
    if (OP == 3) LOOP( w[i] = (u64)(u32)w[i] * c[i] + w[i]; a[i] = (a[i] ^ b[i]) & (b[i] | c[i]); b[i] ^= a[i]; ) // WIDE + 2 LOP3
    if (OP == 4) LOOP( w[i] = (u64)(u32)w[i] * c[i] + w[i]; a[i] = a[i] * b[i] + c[i]; )        // WIDE + IMAD
    if (OP == 5) LOOP( a[i] = a[i] * b[i] + c[i]; b[i] = (b[i] ^ a[i]) & (b[i] | c[i]); )       // IMAD + LOP3
    if (OP == 6) LOOP( a[i] = a[i] * b[i] + c[i]; b[i] = (b[i] ^ a[i]) & (b[i] | c[i]); c[i] = __funnelshift_l(c[i], b[i], 7); ) // IMAD + LOP3 + SHF
    if (OP == 7) LOOP( w[i] = (u64)(u32)w[i] * c[i] + w[i]; w[i] = (u64)(u32)(w[i] >> 32) * b[i] + w[i]; a[i] = (a[i] ^ b[i]) & (a[i] | c[i]); ) // 2 WIDE + LOP
    if (OP == 8) LOOP( a[i] = __umulhi(a[i], b[i]) + c[i]; )                                    // IMAD.HI
    u32 acc = 0;
    for (int i = 0; i < CH; i++) acc += a[i] + b[i] + c[i] + (u32)w[i] + (u32)(w[i] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int OP> void run(const char* name, u32* out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = 148 * 8;
    k<OP><<<blocks, 256>>>(out, 1); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<OP><<<blocks, 256>>>(out, 1); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-26s %7.3f ms  %.2f clk per chain-iter (@1965MHz)\n", name, ms, (ms * 1e-3) * 148.0 * 4 * 1965e6 / ((double)blocks * 8 * ITERS * CH));
}
int main() {
    u32* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("WIDEacc reg", out); run<1>("WIDEacc imm", out); run<2>("WIDEacc hi-mult", out); run<3>("WIDE+2LOP", out);
    run<4>("WIDE+IMAD", out); run<5>("IMAD+LOP", out); run<6>("IMAD+LOP+SHF", out); run<7>("2WIDE+LOP", out); run<8>("IMAD.HI", out);
    return 0;
}
