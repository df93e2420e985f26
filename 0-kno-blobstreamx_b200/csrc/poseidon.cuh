// poseidon.cuh -- Poseidon-Goldilocks width-12 x^7 permutation and the sponge modes built on it.
// Replaces plonky2 hash/poseidon.rs (`Poseidon::poseidon`), hash/poseidon_goldilocks.rs (constants,
// MDS) and hash/hashing.rs (`hash_n_to_hash_no_pad`, `compress`).  Round structure 4 full +
// 22 partial + 4 full; MDS = circulant(17,15,41,16,2,28,13,13,39,18,34,20) + diag(8,0,...).
#pragma once
#include "gl.cuh"
#include "poseidon_rc.h"

__constant__ u64 POSEIDON_RC[POSEIDON_N_RC] = {POSEIDON_RC_LIST};

__device__ __forceinline__ u64 poseidon_sbox(u64 x) {
    u64 x2 = gl_sqr(x), x4 = gl_sqr(x2), x3 = gl_mul(x, x2);
    return gl_mul(x3, x4);
}

// MDS layer on 32-bit halves: each 12-term dot product with constants <= 41 stays below 2^42, so
// the low and high halves accumulate in plain u64 (IMAD.WIDE) and are recombined and reduced once.
__device__ __forceinline__ void poseidon_mds(u64 s[12]) {
    const u32 C[12] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
    u32 lo[12], hi[12];
#pragma unroll
    for (int i = 0; i < 12; i++) {
        lo[i] = (u32)s[i];
        hi[i] = (u32)(s[i] >> 32);
    }
#pragma unroll
    for (int r = 0; r < 12; r++) {
        u64 al = 0, ah = 0;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            al += (u64)C[i] * lo[(i + r) % 12];
            ah += (u64)C[i] * hi[(i + r) % 12];
        }
        if (r == 0) {
            al += (u64)8 * lo[0];
            ah += (u64)8 * hi[0];
        }
        // value = al + ah * 2^32 < 2^75: fold the part above 2^64 with 2^64 == EPS
        u64 low = al + (ah << 32);
        u64 top = (ah >> 32) + (low < al ? 1 : 0);  // < 2^11
        u64 t1 = (top << 32) - top;                  // top * EPS
        u64 v = low + t1;
        if (v < t1) v += GL_EPS;
        s[r] = gl_canon(v);
    }
}

// state canonical in, canonical out
__device__ __forceinline__ void poseidon_permute(u64 s[12]) {
    int rc = 0;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = poseidon_sbox(gl_add(s[i], POSEIDON_RC[rc + i]));
        poseidon_mds(s);
        rc += 12;
    }
#pragma unroll 1
    for (int r = 0; r < 22; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = gl_add(s[i], POSEIDON_RC[rc + i]);
        s[0] = poseidon_sbox(s[0]);
        poseidon_mds(s);
        rc += 12;
    }
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = poseidon_sbox(gl_add(s[i], POSEIDON_RC[rc + i]));
        poseidon_mds(s);
        rc += 12;
    }
}

// PoseidonHash::two_to_one: state = [l, r, 0, 0, 0, 0]; permute; first four words
__device__ __forceinline__ void poseidon_two_to_one(const u64 l[4], const u64 r[4], u64 out[4]) {
    u64 s[12] = {l[0], l[1], l[2], l[3], r[0], r[1], r[2], r[3], 0, 0, 0, 0};
    poseidon_permute(s);
    out[0] = s[0];
    out[1] = s[1];
    out[2] = s[2];
    out[3] = s[3];
}
