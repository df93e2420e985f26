// poseidon.cuh -- Poseidon-Goldilocks width-12 x^7 permutation and the sponge modes built on it.
// Replaces plonky2 hash/poseidon.rs (`Poseidon::poseidon`, `full_rounds`, `partial_rounds`,
// `mds_partial_layer_init/fast`), hash/poseidon_goldilocks.rs (constants, MDS) and hash/hashing.rs
// (`hash_n_to_hash_no_pad`, `compress`).  Round structure 4 full + 22 partial + 4 full; MDS =
// circulant(17,15,41,16,2,28,13,13,39,18,34,20) + diag(8,0,...).
//
// B200 shape of the computation (measured choices):
//  * state lives in registers as LAZY u64 (any representative), canonicalised only on output;
//  * full rounds: x^7 by two squarings and two multiplies (lazy); every MDS layer is multiplication-free on three
//    22-bit limb planes (poseidon_mds.cuh), the NEXT round's constants riding as third addends;
//  * partial rounds keep lanes 1..11 as limb planes from round to round (only lane 0 is merged for its S-box);
//    the CPU provers' sparse factorisation (23 full-width multiplies per round) is dearer on a machine whose only
//    multiplier is 32x32->64 at quarter rate -- that variant lives in tools/microbench/poseidon_v7_sparse.cuh;
//  * one shared full-round body and one partial-round body (no unrolling over rounds): the whole
//    permutation is ~34 KB of SASS and stays inside the instruction cache (v1 was 112 KB and stalled
//    on instruction fetch for 12.6 warp-cycles per issue).
#pragma once
#include "gl.cuh"
#include "poseidon_mds.cuh"
#include "poseidon_rc.h"

__constant__ __align__(16) u32 POSEIDON_POST30_L22[30 * 36] = {POSEIDON_POST30_L22_LIST};
// one round's 36 plane constants through nine 128-bit uniform loads (36 LDCU otherwise: issue slots)
__device__ __forceinline__ void poseidon_load_post(int round, u32 k[36]) {
    const uint4 *q = reinterpret_cast<const uint4 *>(POSEIDON_POST30_L22 + 36 * round);
#pragma unroll
    for (int i = 0; i < 9; i++) {
        uint4 v = q[i];
        k[4 * i] = v.x, k[4 * i + 1] = v.y, k[4 * i + 2] = v.z, k[4 * i + 3] = v.w;
    }
}
__constant__ u64 POSEIDON_RC[POSEIDON_N_RC] = {POSEIDON_RC_LIST};

__device__ __forceinline__ u64 poseidon_sbox(u64 x) {
    u64 x2 = gl_sqr_lazy(x), x4 = gl_sqr_lazy(x2), x3 = gl_mul_lazy(x, x2);
    return gl_mul_lazy(x3, x4);
}

// al + ah * 2^32 (al, ah < 2^43) -> lazy u64.   ah = hh 2^32 + hl:  == al + hh * EPS + hl 2^32
__device__ __forceinline__ u64 poseidon_fold96(u64 al, u64 ah) {
    u32 hl = (u32)ah, hh = (u32)(ah >> 32), r0, r1;
    u64 t = (u64)hh * 0xFFFFFFFFu + al;  // < 2^44, no overflow
    u32 t0 = (u32)t, t1 = (u32)(t >> 32);
    asm("{\n\t.reg .u32 c,m;\n\t"
        "add.cc.u32 %1, %3, %4;\n\t"   // high word += hl; a wrap is worth EPS, and the high word is
        "addc.u32 c, 0, 0;\n\t"         // then < 2^12 so adding EPS cannot wrap again
        "sub.u32 m, 0, c;\n\t"
        "add.cc.u32 %0, %2, m;\n\t"
        "addc.u32 %1, %1, 0;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(t0), "r"(t1), "r"(hl));
    return ((u64)r1 << 32) | r0;
}

// o0 + o1 2^22 + o2 2^44 (o_k < 2^32) -> lazy u64.  Three words w2:w1:w0 with w2 < 2^13, then
// w2 2^64 == w2 EPS folded in by one IMAD.WIDE with carry-out; a wrap is worth EPS once more and the
// wrapped value is < 2^45, so that add cannot wrap again.
__device__ __forceinline__ u64 poseidon_merge22(u32 o0, u32 o1, u32 o2) {
    u32 r0, r1;
    asm("{\n\t.reg .u32 w0,w1,w2,a,b,c,m;\n\t"
        "shl.b32 a, %3, 22;\n\t"
        "add.cc.u32 w0, %2, a;\n\t"
        "shr.b32 a, %3, 10;\n\t"
        "shl.b32 b, %4, 12;\n\t"
        "addc.cc.u32 w1, a, b;\n\t"
        "shr.b32 a, %4, 20;\n\t"
        "addc.u32 w2, a, 0;\n\t"
        "mad.lo.cc.u32 w0, w2, 0xFFFFFFFF, w0;\n\t"
        "madc.hi.cc.u32 w1, w2, 0xFFFFFFFF, w1;\n\t"
        "addc.u32 c, 0, 0;\n\t"
        "sub.u32 m, 0, c;\n\t"
        "add.cc.u32 %0, w0, m;\n\t"
        "addc.u32 %1, w1, 0;\n\t}"
        : "=r"(r0), "=r"(r1)
        : "r"(o0), "r"(o1), "r"(o2));
    return ((u64)r1 << 32) | r0;
}

// s <- M s + post_j on three 22-bit limb planes; lazy in, lazy out
__device__ __forceinline__ void poseidon_mds_post_fft(u64 s[12], const u32 *__restrict__ post) {
    u32 x0[12], x1[12], x2[12], y0[12], y1[12], y2[12];
#pragma unroll
    for (int i = 0; i < 12; i++) {
        u32 lo = (u32)s[i], hi = (u32)(s[i] >> 32);
        x0[i] = lo & POSEIDON_LIMB_MASK;
        x1[i] = __funnelshift_r(lo, hi, POSEIDON_LIMB_BITS) & POSEIDON_LIMB_MASK;
        x2[i] = hi >> 12;
    }
    poseidon_mds_plane(x0, post, y0);
    poseidon_mds_plane(x1, post + 12, y1);
    poseidon_mds_plane(x2, post + 24, y2);
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = poseidon_merge22(y0[i], y1[i], y2[i]);
}

// u64 -> three 22-bit limbs (22, 22, 20 bits)
__device__ __forceinline__ void poseidon_split22(u64 v, u32 &x0, u32 &x1, u32 &x2) {
    u32 lo = (u32)v, hi = (u32)(v >> 32);
    x0 = lo & POSEIDON_LIMB_MASK;
    x1 = __funnelshift_r(lo, hi, POSEIDON_LIMB_BITS) & POSEIDON_LIMB_MASK;
    x2 = hi >> 12;
}

// Limb normalisation between two plane-resident partial rounds.  In: plane outputs y_k in [0, 2^31).
// The bits of y2 above 2^20 are worth top * 2^64 == top * (2^32 - 1): + top << 10 on limb 1, - top on limb 0
// (which may go negative: the carries below use arithmetic shifts).  Out: y0, y1 < 2^22, y2 < 2^20 + 2^11,
// so the next layer's outputs stay below 264 * 2^22 + 2^22 < 2^31 (model and bounds checked in
// tools/gen_poseidon_constants.py `poseidon_planes`).  10 ALU instructions; merge + split is 18 + 1 IMAD.WIDE.
__device__ __forceinline__ void poseidon_norm22(u32 &y0, u32 &y1, u32 &y2) {
    u32 top = y2 >> 20;
    y2 &= 0xFFFFFu;
    y1 += top << 10;
    y0 -= top;
    int c0 = (int)y0 >> POSEIDON_LIMB_BITS;
    y0 &= POSEIDON_LIMB_MASK;
    y1 += (u32)c0;
    int c1 = (int)y1 >> POSEIDON_LIMB_BITS;
    y1 &= POSEIDON_LIMB_MASK;
    y2 += (u32)c1;
}

// lazy in, lazy out.  Round r: S-box, then s <- M s + RC[r + 1] (the constants ride as third addends).
__device__ __forceinline__ void poseidon_permute_lazy(u64 s[12]) {
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = gl_add_lazy_canon(s[i], POSEIDON_RC[i]);
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
#pragma unroll 1
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 12; i++) s[i] = poseidon_sbox(s[i]);
            u32 k[36];
            poseidon_load_post(half * 26 + r, k);
            poseidon_mds_post_fft(s, k);
        }
        if (half == 0) {
            u32 x0[12], x1[12], x2[12], a0, a1, a2;
#pragma unroll
            for (int i = 1; i < 12; i++) poseidon_split22(s[i], x0[i], x1[i], x2[i]);
            poseidon_split22(s[0], a0, a1, a2);
#pragma unroll 1
            for (int p = 4; p < 26; p++) {
                u32 post[36];
                poseidon_load_post(p, post);
                poseidon_split22(poseidon_sbox(poseidon_merge22(a0, a1, a2)), x0[0], x1[0], x2[0]);
                u32 y0[12], y1[12], y2[12];
                poseidon_mds_plane(x0, post, y0);
                poseidon_mds_plane(x1, post + 12, y1);
                poseidon_mds_plane(x2, post + 24, y2);
                a0 = y0[0], a1 = y1[0], a2 = y2[0];
#pragma unroll
                for (int i = 1; i < 12; i++) {
                    poseidon_norm22(y0[i], y1[i], y2[i]);
                    x0[i] = y0[i], x1[i] = y1[i], x2[i] = y2[i];
                }
            }
            s[0] = poseidon_merge22(a0, a1, a2);
#pragma unroll
            for (int i = 1; i < 12; i++) s[i] = poseidon_merge22(x0[i], x1[i], x2[i]);
        }
    }
}

// state canonical (or lazy) in, canonical out
__device__ __forceinline__ void poseidon_permute(u64 s[12]) {
    poseidon_permute_lazy(s);
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = gl_canon(s[i]);
}

// PoseidonHash::two_to_one: state = [l, r, 0, 0, 0, 0]; permute; first four words
__device__ __forceinline__ void poseidon_two_to_one(const u64 l[4], const u64 r[4], u64 out[4]) {
    u64 s[12] = {l[0], l[1], l[2], l[3], r[0], r[1], r[2], r[3], 0, 0, 0, 0};
    poseidon_permute_lazy(s);
    out[0] = gl_canon(s[0]);
    out[1] = gl_canon(s[1]);
    out[2] = gl_canon(s[2]);
    out[3] = gl_canon(s[3]);
}

// Where permutations depend on one another (the Fiat-Shamir transcript, the narrow top levels of a Merkle tree) what
// counts is the LATENCY of one permutation: a single thread takes ~39 us (30 rounds x 12 S-boxes back to back).  Here lane i < 12 of one warp
// owns state word i: the twelve S-boxes of a full round run side by side and the MDS row of every lane is
// gathered with shuffles (Poseidon::poseidon in its naive form, hash/poseidon.rs `poseidon_naive`; all words
// canonical).  ~5 us per permutation.
__device__ __forceinline__ u64 poseidon_permute_warp(u64 x, int lane) {
    const u32 C[12] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
    const int me = lane < 12 ? lane : 0;
    for (int r = 0; r < 30; r++) {
        x = gl_add(x, POSEIDON_RC[12 * r + me]);
        const bool full = r < 4 || r >= 26;
        if (full || lane == 0) {
            u64 x2 = gl_mul_lazy(x, x), x4 = gl_mul_lazy(x2, x2);  // lazy words are fine for the halves below
            x = gl_mul_lazy(gl_mul_lazy(x, x2), x4);
        }
        // y[lane] = sum_i C[i] x[(lane + i) % 12] (+ 8 x[0] for lane 0), on 32-bit halves
        const u32 lo = (u32)x, hi = (u32)(x >> 32);
        u64 al = 0, ah = 0;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            int src = me + i;
            src = src >= 12 ? src - 12 : src;
            al += (u64)C[i] * __shfl_sync(0xFFFFFFFFu, lo, src);
            ah += (u64)C[i] * __shfl_sync(0xFFFFFFFFu, hi, src);
        }
        if (lane == 0) {
            al += (u64)8 * lo;
            ah += (u64)8 * hi;
        }
        x = gl_canon(poseidon_fold96(al, ah));
    }
    return x;
}

