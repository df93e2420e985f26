// pipes.cu -- integer-pipe throughput microbenchmarks on B200 (instr/clk/SM), to size the
// Goldilocks multiply / Poseidon MDS instruction mix.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;
#define ITERS 4096
#define CHAINS 8
template <int OP> __global__ void __launch_bounds__(256) k(u32* out, u32 seed, u32 mul) {
    u32 a[CHAINS], b[CHAINS]; u64 w[CHAINS];
    for (int i = 0; i < CHAINS; i++) { a[i] = threadIdx.x * 7 + i + seed; b[i] = a[i] ^ 0x1234567; w[i] = ((u64)a[i] << 32) | b[i]; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) {
            if (OP == 0) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(mul));       // IMAD.WIDE reg
            if (OP == 1) asm volatile("mad.wide.u32 %0, %1, 41, %0;" : "+l"(w[i]) : "r"(a[i]));                   // IMAD.WIDE imm
            if (OP == 2) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(mul), "r"(b[i]));           // IMAD
            if (OP == 3) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(mul), "r"(b[i]));           // IMAD.HI
            if (OP == 4) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b[i]));                            // IADD3
            if (OP == 5) asm volatile("{add.cc.u32 %0, %0, %2; addc.u32 %1, %1, %3;}" : "+r"(a[i]), "+r"(b[i]) : "r"(mul), "r"(seed)); // 64-bit add
            if (OP == 6) { asm volatile("mad.wide.u32 %0, %1, 41, %0;" : "+l"(w[i]) : "r"(a[i]));
                           asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(mul)); }                         // WIDE + IADD mix
            if (OP == 7) { asm volatile("mad.wide.u32 %0, %1, 41, %0;" : "+l"(w[i]) : "r"(a[i]));
                           asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(mul));
                           asm volatile("xor.b32 %0, %0, %1;" : "+r"(a[i]) : "r"(mul)); }                         // WIDE + 2 ALU
            if (OP == 8) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(mul), "r"(b[i]));       // LOP3
            if (OP == 9) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(b[i]));                  // SHF
        }
    }
    u32 acc = 0;
    for (int i = 0; i < CHAINS; i++) acc += a[i] + b[i] + (u32)w[i] + (u32)(w[i] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int OP> void run(const char* name, int per_iter, u32* out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = 148 * 8;
    k<OP><<<blocks, 256>>>(out, 1, 3); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<OP><<<blocks, 256>>>(out, 1, 3); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double warp_instr = (double)blocks * 8 * ITERS * CHAINS * per_iter;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-28s %8.3f ms  %7.2f Gwarp-instr/s  -> %.3f warp-instr/clk/SMSP at %d MHz nominal\n", name, ms,
           warp_instr / ms / 1e6, warp_instr / (ms * 1e-3) / (148.0 * 4 * clk * 1e3), clk / 1000);
}
int main() {
    u32* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("IMAD.WIDE.U32 reg", 1, out); run<1>("IMAD.WIDE.U32 imm", 1, out); run<2>("IMAD lo", 1, out);
    run<3>("IMAD.HI", 1, out); run<4>("IADD3", 1, out); run<5>("add.cc+addc (2 instr)", 2, out);
    run<6>("WIDE+IADD (2 instr)", 2, out); run<7>("WIDE+IADD+LOP (3 instr)", 3, out); run<8>("LOP3", 1, out);
    run<9>("SHF", 1, out);
    return 0;
}
