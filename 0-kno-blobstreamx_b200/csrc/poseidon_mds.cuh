// poseidon_mds.cuh -- the Poseidon-Goldilocks MDS layer without multiplications.
// Replaces plonky2 hash/poseidon.rs `mds_layer` / hash/poseidon_goldilocks.rs `mds_multiply_freq`
// for M = circ(17,15,41,16,2,28,13,13,39,18,34,20) + diag(8,0,...,0).
//
// The circulant part is a length-12 cyclic convolution.  Writing the index as 3a + b turns it into a
// product in Z[y]/(y^4 - 1)[x]/(x^3 - y); a 4-point DFT in y (roots 1, i, -1, -i) splits it into one
// cyclic 3-convolution (y = 1), one negacyclic (y = -1) and one complex 3-convolution (y = i), whose
// transformed kernel is tiny -- every entry is +-1 or +-2^k -- so the whole layer is adds and shifts:
//      y =  1 : K/4 = [16, 32, 16]
//      y = -1 : K/4 = [-1, -8, 2]
//      y =  i : K/2 = [2 + i, -4 - i, 16 - i]
// (derived and checked against the dense form by tools/gen_poseidon_constants.py `mds_fft_check`).
//
// B200 shape: the state is cut into three limbs of 22 bits held in 32-bit registers and every limb
// plane goes through the transform independently in wrapping u32 arithmetic.  All intermediate
// overflow is harmless because only ring operations are used and the true outputs are
// < (256 + 8) * 2^22 + 2^22 < 2^32.  No carry chains and no IMAD.WIDE: ~75 IADD3/LEA per plane instead
// of 100 quarter-rate IMAD.WIDE.U32, which is what bounded the leaf-hash kernel (profiles/).
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define BSX_HD __host__ __device__ __forceinline__
#else
#define BSX_HD static inline
#endif

#if defined(__CUDA_ARCH__)
#define BSX_UNROLL _Pragma("unroll")
#else
#define BSX_UNROLL
#endif

#define POSEIDON_LIMB_BITS 22
#define POSEIDON_LIMB_MASK 0x3FFFFFu

// o[r] = sum_i C[i] * x[(i + r) % 12] + 8 * x[0] * [r == 0] + k[r]   (mod 2^32)
BSX_HD void poseidon_mds_plane(const uint32_t x[12], const uint32_t *k, uint32_t o[12]) {
    uint32_t A[3], Bm[3], R[3], I[3];  // transforms at y = 1, -1, and (re, im) at y = i
    BSX_UNROLL
    for (int b = 0; b < 3; b++) {
        uint32_t x0 = x[b], x1 = x[3 + b], x2 = x[6 + b], x3 = x[9 + b];
        uint32_t e0 = x0 + x2, o0 = x1 + x3;
        A[b] = e0 + o0;
        Bm[b] = e0 - o0;
        R[b] = x0 - x2;
        I[b] = x1 - x3;
    }
    uint32_t T = A[0] + A[1] + A[2];
    uint32_t P1[3], Pm[3], Pr[3], Pi[3];
    P1[0] = (T + A[2]) << 4;
    P1[1] = (T + A[0]) << 4;
    P1[2] = (T + A[1]) << 4;
    Pm[0] = (Bm[2] << 3) - Bm[0] - (Bm[1] << 1);
    Pm[1] = 0u - (Bm[0] << 3) - Bm[1] - (Bm[2] << 1);
    Pm[2] = (Bm[0] << 1) - (Bm[1] << 3) - Bm[2];
    Pr[0] = (R[0] << 1) - I[0] + R[1] - (I[1] << 4) + R[2] + (I[2] << 2);
    Pr[1] = I[0] - (R[0] << 2) + (R[1] << 1) - I[1] + R[2] - (I[2] << 4);
    Pr[2] = (R[0] << 4) + I[0] - (R[1] << 2) + I[1] + (R[2] << 1) - I[2];
    Pi[0] = R[0] + (I[0] << 1) + (R[1] << 4) + I[1] - (R[2] << 2) + I[2];
    Pi[1] = R[1] - R[0] - (I[0] << 2) + (I[1] << 1) + (R[2] << 4) + I[2];
    Pi[2] = (I[0] << 4) - R[0] - R[1] - (I[1] << 2) + R[2] + (I[2] << 1);
    BSX_UNROLL
    for (int b = 0; b < 3; b++) {
        uint32_t p = P1[b] + Pm[b], q = P1[b] - Pm[b];
        o[b] = p + Pr[b] + k[b];
        o[3 + b] = q + Pi[b] + k[3 + b];
        o[6 + b] = p - Pr[b] + k[6 + b];
        o[9 + b] = q - Pi[b] + k[9 + b];
    }
    o[0] += x[0] << 3;
}
