// poseidon_v7_sparse.cuh -- the round-1 "v7" permutation kept OUT of the product for A/B measurements only
// (tools/microbench): partial rounds by the sparse factorisation CPU provers use (23 full-width multiplies per round
// plus a dense 11x11 pre-multiplication) and the MDS layer on 32-bit halves with IMAD.WIDE accumulation.  On B200 it
// costs 20.0 k instructions / 5.3 k IMAD.WIDE per permutation against 18.8 k / 2.4 k for the shipped limb-plane form
// (1.10 vs 1.39 G permutations/s).  Include after poseidon.cuh; defines poseidon_permute_lazy_v7.
#pragma once
#include "../../0-kno-blobstreamx_b200/csrc/poseidon.cuh"

__constant__ u64 POSEIDON_POST_LO[96] = {POSEIDON_POST_LO_LIST};
__constant__ u64 POSEIDON_POST_HI[96] = {POSEIDON_POST_HI_LIST};
__constant__ u32 POSEIDON_POST_L22[8 * 36] = {POSEIDON_POST_L22_LIST};
__constant__ u64 POSEIDON_FAST_INIT[121] = {POSEIDON_FAST_INIT_LIST};
__constant__ u64 POSEIDON_FAST_K[22] = {POSEIDON_FAST_K_LIST};
__constant__ u64 POSEIDON_FAST_V[242] = {POSEIDON_FAST_V_LIST};
__constant__ u64 POSEIDON_FAST_W[242] = {POSEIDON_FAST_W_LIST};

// MDS layer with the following constant vector fused: s <- M s + post, post given as 32-bit halves.
__device__ __forceinline__ void poseidon_mds_post(u64 s[12], const u64 *__restrict__ post_lo,
                                                  const u64 *__restrict__ post_hi) {
    const u32 C[12] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
    u32 lo[12], hi[12];
#pragma unroll
    for (int i = 0; i < 12; i++) {
        lo[i] = (u32)s[i];
        hi[i] = (u32)(s[i] >> 32);
    }
#pragma unroll
    for (int r = 0; r < 12; r++) {
        u64 al = post_lo[r], ah = post_hi[r];
#pragma unroll
        for (int i = 0; i < 12; i++) {
            al += (u64)C[i] * lo[(i + r) % 12];
            ah += (u64)C[i] * hi[(i + r) % 12];
        }
        if (r == 0) {
            al += (u64)8 * lo[0];
            ah += (u64)8 * hi[0];
        }
        s[r] = poseidon_fold96(al, ah);
    }
}


// lazy in, lazy out
__device__ __forceinline__ void poseidon_permute_lazy_v7(u64 s[12]) {
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = gl_add_lazy_canon(s[i], POSEIDON_RC[i]);
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
#pragma unroll 1
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 12; i++) s[i] = poseidon_sbox(s[i]);
            const int j = half * 4 + r;
            poseidon_mds_post(s, POSEIDON_POST_LO + 12 * j, POSEIDON_POST_HI + 12 * j);
        }
        if (half == 0) {
            // mds_partial_layer_init: x[1:] <- INIT x[1:] (dense 11x11, once per permutation).  Rows are
            // produced by a rolled loop into a small local array to keep the code short.
            u64 t[11];
#pragma unroll 1
            for (int r = 0; r < 11; r++) {
                gl_acc3 A = gl_acc3_zero();
#pragma unroll
                for (int c = 0; c < 11; c++) gl_acc3_mac(A, s[1 + c], POSEIDON_FAST_INIT[11 * r + c]);
                t[r] = gl_acc3_reduce(A);
            }
#pragma unroll
            for (int r = 0; r < 11; r++) s[1 + r] = t[r];
#pragma unroll 1
            for (int p = 0; p < 22; p++) {
                u64 s0 = gl_add_lazy_canon(poseidon_sbox(s[0]), POSEIDON_FAST_K[p]);
                // d = 25 s0 + V[p] . x[1:]
                gl_acc3 A = gl_acc3_zero();
                gl_acc3_mac(A, s0, 25);
#pragma unroll
                for (int i = 0; i < 11; i++) gl_acc3_mac(A, s[1 + i], POSEIDON_FAST_V[11 * p + i]);
#pragma unroll
                for (int i = 0; i < 11; i++) s[1 + i] = gl_mul_add_lazy(s0, POSEIDON_FAST_W[11 * p + i], s[1 + i]);
                s[0] = gl_acc3_reduce(A);
            }
#pragma unroll
            for (int i = 0; i < 12; i++) s[i] = gl_add_lazy_canon(s[i], POSEIDON_RC[12 * 26 + i]);
        }
    }
}

