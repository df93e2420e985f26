// gl.cuh -- Goldilocks field F_p, p = 2^64 - 2^32 + 1, device primitives (kernel K1 of SURVEY.md).
// Replaces plonky2 field/src/goldilocks_field.rs (Add/Sub/Mul, reduce128) and
// field/src/extension/quadratic.rs (W = 7).  Storage is canonical u64 (< p) everywhere a value
// crosses a kernel boundary; inside kernels a few helpers accept lazily reduced inputs and say so.
#pragma once
#include <cstdint>

typedef unsigned long long u64;
typedef unsigned int u32;

#define GL_P 0xFFFFFFFF00000001ULL
#define GL_EPS 0xFFFFFFFFULL  // 2^64 mod p

__host__ __device__ __forceinline__ u64 gl_canon(u64 x) { return x >= GL_P ? x - GL_P : x; }

// a, b canonical -> canonical.
__device__ __forceinline__ u64 gl_add(u64 a, u64 b) {
    u64 s = a + b;
    // a + b < 2p: subtract p when the 64-bit add wrapped or landed in [p, 2^64)
    return (s < a || s >= GL_P) ? s - GL_P : s;
}
__device__ __forceinline__ u64 gl_sub(u64 a, u64 b) {
    u64 d = a - b;
    return (a < b) ? d + GL_P : d;
}
__device__ __forceinline__ u64 gl_neg(u64 a) { return a ? GL_P - a : 0; }
__device__ __forceinline__ u64 gl_double(u64 a) { return gl_add(a, a); }

// 128-bit (hi:lo) -> canonical.  2^64 == EPS, 2^96 == -1 (mod p).
__device__ __forceinline__ u64 gl_reduce128(u64 hi, u64 lo) {
    u32 hh = (u32)(hi >> 32), hl = (u32)hi;
    u64 t0 = lo - (u64)hh;
    if (lo < (u64)hh) t0 -= GL_EPS;               // borrowed 2^64 == EPS
    u64 t1 = ((u64)hl << 32) - (u64)hl;           // hl * EPS
    u64 r = t0 + t1;
    if (r < t1) r += GL_EPS;                      // wrapped 2^64 == EPS
    return gl_canon(r);
}

// 64x64 -> 128 from four 32x32->64 IMAD.WIDE products with 64-bit accumulation.
__device__ __forceinline__ void gl_mul_wide(u64 a, u64 b, u64 &hi, u64 &lo) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32);
    u64 p00 = (u64)a0 * b0;
    u64 t = (u64)a0 * b1 + (p00 >> 32);           // < 2^64: (2^32-1)^2 + 2^32 - 1
    u64 t2 = (u64)a1 * b0 + (u64)(u32)t;          // < 2^64
    hi = (u64)a1 * b1 + (t >> 32) + (t2 >> 32);   // < 2^64
    lo = (t2 << 32) | (u64)(u32)p00;
}
__device__ __forceinline__ u64 gl_mul(u64 a, u64 b) {
    u64 hi, lo;
    gl_mul_wide(a, b, hi, lo);
    return gl_reduce128(hi, lo);
}
__device__ __forceinline__ u64 gl_sqr(u64 a) { return gl_mul(a, a); }

// small-constant multiply: c < 2^32, a canonical
__device__ __forceinline__ u64 gl_mul_small(u64 a, u32 c) {
    u64 l = (u64)(u32)a * c;
    u64 h = (u64)(u32)(a >> 32) * c + (l >> 32);  // < 2^64
    // value = h * 2^32 + (u32)l ; h = hh*2^32 + hl -> hh*2^64 + hl*2^32 + l32
    u64 lo = (h << 32) | (u64)(u32)l;
    u64 hi = h >> 32;                              // < 2^32
    u64 t1 = (hi << 32) - hi;                      // hi * EPS
    u64 r = lo + t1;
    if (r < t1) r += GL_EPS;
    return gl_canon(r);
}


__device__ __forceinline__ u64 gl_pow(u64 a, u64 e) {
    u64 r = 1;
    while (e) {
        if (e & 1) r = gl_mul(r, a);
        a = gl_sqr(a);
        e >>= 1;
    }
    return r;
}
__device__ __forceinline__ u64 gl_inv(u64 a) { return gl_pow(a, GL_P - 2); }

// ---- quadratic extension F_p[x]/(x^2 - 7), element (c0, c1) ----
struct gl2 {
    u64 c0, c1;
};
__device__ __forceinline__ gl2 gl2_add(gl2 a, gl2 b) { return {gl_add(a.c0, b.c0), gl_add(a.c1, b.c1)}; }
__device__ __forceinline__ gl2 gl2_sub(gl2 a, gl2 b) { return {gl_sub(a.c0, b.c0), gl_sub(a.c1, b.c1)}; }
__device__ __forceinline__ gl2 gl2_mul(gl2 a, gl2 b) {
    u64 a0b0 = gl_mul(a.c0, b.c0), a1b1 = gl_mul(a.c1, b.c1);
    return {gl_add(a0b0, gl_mul_small(a1b1, 7)), gl_add(gl_mul(a.c0, b.c1), gl_mul(a.c1, b.c0))};
}
__device__ __forceinline__ gl2 gl2_mul_base(gl2 a, u64 b) { return {gl_mul(a.c0, b), gl_mul(a.c1, b)}; }
__device__ __forceinline__ gl2 gl2_pow(gl2 a, u64 e) {
    gl2 r = {1, 0};
    while (e) {
        if (e & 1) r = gl2_mul(r, a);
        a = gl2_mul(a, a);
        e >>= 1;
    }
    return r;
}
__device__ __forceinline__ gl2 gl2_inv(gl2 a) {
    u64 n = gl_sub(gl_sqr(a.c0), gl_mul_small(gl_sqr(a.c1), 7));
    u64 ni = gl_inv(n);
    return {gl_mul(a.c0, ni), gl_mul(gl_neg(a.c1), ni)};
}

// ---- host-side scalar helpers (table seeds, shifts); not a compute path ----
static inline u64 gl_host_mul(u64 a, u64 b) {
    unsigned __int128 x = (unsigned __int128)a * b;
    u64 lo = (u64)x, hi = (u64)(x >> 64), hh = hi >> 32, hl = hi & GL_EPS;
    u64 t0 = lo - hh;
    if (lo < hh) t0 -= GL_EPS;
    u64 t1 = hl * GL_EPS, r = t0 + t1;
    if (r < t1) r += GL_EPS;
    return r >= GL_P ? r - GL_P : r;
}
static inline u64 gl_host_pow(u64 a, u64 e) {
    u64 r = 1;
    while (e) {
        if (e & 1) r = gl_host_mul(r, a);
        a = gl_host_mul(a, a);
        e >>= 1;
    }
    return r;
}
static inline u64 gl_host_inv(u64 a) { return gl_host_pow(a, GL_P - 2); }
// primitive 2^k-th root of unity: POWER_OF_TWO_GENERATOR^(2^(32-k))
static inline u64 gl_host_root(int k) {
    u64 w = 1753635133440165772ULL;
    for (int i = k; i < 32; i++) w = gl_host_mul(w, w);
    return w;
}
