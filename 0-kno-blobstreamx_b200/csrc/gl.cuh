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

// ---------------------------------------------------------------------------------------------
// Lazy arithmetic for the hot kernels.  "Lazy" values are any u64 (a representative of the
// residue, not necessarily < p).  Everything below is branch-free inline PTX on 32-bit limbs so
// that ptxas emits IMAD.WIDE.U32 (+carry-out / .X carry-in) for the products and IADD3/IMAD.X
// carry chains for the reduction -- no IMAD.HI (quarter rate on B200, tools/microbench/pipes.cu)
// and no compare+select sequences.  16 SASS instructions per multiply instead of ~36.

// EPS as a constant-bank operand: ptxas then keeps `x2 * EPS + (x1:x0)` as ONE IMAD.WIDE.U32 with carry-out
// (with the literal 0xFFFFFFFF it strength-reduces the low half to a subtraction and adds an IMAD.HI).
static __constant__ u32 GL_EPS32_C = 0xFFFFFFFFu;

// Tail shared by the multiplies: (x3:x2:x1:x0) -> lazy u64.
//   r = (x1:x0) + x2 * EPS - x3  in 96-bit two's complement: the IMAD.WIDE's carry and the borrow of the
//   subtraction meet in r2 in {-1, 0, 1}; one signed correction r2 * EPS = (r2 >> 31 : -r2) finishes.  It
//   cannot wrap again: r2 = 1 means the low 64 bits are < 2^64 - 2^32, r2 = -1 means they are >= 2^64 - 2^32
//   (x3 < 2^32).  11 SASS instructions after the four products (the two-fix form took 15).
#define GL_REDUCE_TAIL_WIDE(EPS_OPERAND, SUB_HI)             \
    "mad.lo.cc.u32 x0, x2, " EPS_OPERAND ", x0;\n\t"         \
    "madc.hi.cc.u32 x1, x2, " EPS_OPERAND ", x1;\n\t"        \
    "addc.u32 r2, 0, 0;\n\t"                                 \
    "sub.cc.u32 x0, x0, x3;\n\t"                             \
    "subc.cc.u32 x1, x1, " SUB_HI ";\n\t"                    \
    "subc.u32 r2, r2, 0;\n\t"                                \
    "sub.u32 n, 0, r2;\n\t"                                  \
    "shr.s32 eh, r2, 31;\n\t"                                \
    "add.cc.u32 %0, x0, n;\n\t"                              \
    "addc.u32 %1, x1, eh;\n\t}"
// Same tail with x2 * EPS = (x2 << 32) - x2 formed by two ALU subtractions instead of the fifth IMAD.WIDE:
// +3 ALU instructions, -1 quarter-rate FMA-heavy instruction.  Pays when the FMA-heavy pipe is the busier one
// (Poseidon after the MDS/accumulator rewrites: 86 % vs 57 % ALU).  Measured: 1.10 -> 1.02 Gperm/s, because the
// kernel then issues 14 % more instructions at an unchanged ~0.6 IPC -- kept only as a documented dead end.
#define GL_REDUCE_TAIL_ALU(EPS_OPERAND, SUB_HI)              \
    "sub.cc.u32 n, 0, x2;\n\t"                               \
    "subc.u32 eh, x2, 0;\n\t"                                \
    "add.cc.u32 x0, x0, n;\n\t"                              \
    "addc.cc.u32 x1, x1, eh;\n\t"                            \
    "addc.u32 r2, 0, 0;\n\t"                                 \
    "sub.cc.u32 x0, x0, x3;\n\t"                             \
    "subc.cc.u32 x1, x1, " SUB_HI ";\n\t"                    \
    "subc.u32 r2, r2, 0;\n\t"                                \
    "sub.u32 n, 0, r2;\n\t"                                  \
    "shr.s32 eh, r2, 31;\n\t"                                \
    "add.cc.u32 %0, x0, n;\n\t"                              \
    "addc.u32 %1, x1, eh;\n\t}"
#ifndef GL_MUL_TAIL_ALU
#define GL_MUL_TAIL_ALU 0
#endif
#if GL_MUL_TAIL_ALU
#define GL_REDUCE_TAIL(E, S) GL_REDUCE_TAIL_ALU(E, S)
#else
#define GL_REDUCE_TAIL(E, S) GL_REDUCE_TAIL_WIDE(E, S)
#endif

// 128-bit product + 64-bit addend, reduced:  a * b + c  (mod p), all lazy.  c is added to the partial
// products a0*b0 and a0*b1 (neither sum can overflow 64 bits), so the addend costs no carry handling later.
//   x = x0 + x1 2^32 + x2 2^64 + x3 2^96 == (x1:x0) - x3 + x2 * EPS   (2^64 == EPS, 2^96 == -1)
__device__ __forceinline__ u64 gl_mul_add_lazy(u64 a, u64 b, u64 c) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32);
    u32 c0 = (u32)c, c1 = (u32)(c >> 32), r0, r1;
    asm("{\n\t.reg .u32 x0,x1,x2,x3,t0,t1,t2,r2,n,eh;\n\t.reg .u64 w0,w3,t,cx0,cx1;\n\t"
        "cvt.u64.u32 cx0, %7;\n\t"
        "cvt.u64.u32 cx1, %8;\n\t"
        "mad.wide.u32 w0, %2, %4, cx0;\n\t"
        "mul.wide.u32 w3, %3, %5;\n\t"
        "mad.wide.u32 t, %2, %5, cx1;\n\t"
        "mov.b64 {t0,t1}, t;\n\t"
        "mad.lo.cc.u32 t0, %3, %4, t0;\n\t"
        "madc.hi.cc.u32 t1, %3, %4, t1;\n\t"
        "addc.u32 t2, 0, 0;\n\t"
        "mov.b64 {x0,x1}, w0;\n\t"
        "mov.b64 {x2,x3}, w3;\n\t"
        "add.cc.u32 x1, x1, t0;\n\t"
        "addc.cc.u32 x2, x2, t1;\n\t"
        "addc.u32 x3, x3, t2;\n\t"
        GL_REDUCE_TAIL("%6", "0")
        : "=r"(r0), "=r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(GL_EPS32_C), "r"(c0), "r"(c1));
    return ((u64)r1 << 32) | r0;
}

// a * b (mod p), lazy in, lazy out: 5 IMAD.WIDE + 11 other instructions
__device__ __forceinline__ u64 gl_mul_lazy(u64 a, u64 b) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32), r0, r1;
    asm("{\n\t.reg .u32 x0,x1,x2,x3,t0,t1,t2,r2,n,eh;\n\t.reg .u64 w0,w3,t;\n\t"
        "mul.wide.u32 w0, %2, %4;\n\t"
        "mul.wide.u32 w3, %3, %5;\n\t"
        "mul.wide.u32 t, %2, %5;\n\t"
        "mov.b64 {t0,t1}, t;\n\t"
        "mad.lo.cc.u32 t0, %3, %4, t0;\n\t"
        "madc.hi.cc.u32 t1, %3, %4, t1;\n\t"
        "addc.u32 t2, 0, 0;\n\t"
        "mov.b64 {x0,x1}, w0;\n\t"
        "mov.b64 {x2,x3}, w3;\n\t"
        "add.cc.u32 x1, x1, t0;\n\t"
        "addc.cc.u32 x2, x2, t1;\n\t"
        "addc.u32 x3, x3, t2;\n\t"
        GL_REDUCE_TAIL("%6", "0")
        : "=r"(r0), "=r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(GL_EPS32_C));
    return ((u64)r1 << 32) | r0;
}

// a * a (mod p), lazy in, lazy out: the cross product once, doubled on the ALU (GL_SQR_MODE 1: by shifts,
// 2: added twice) -- three IMAD.WIDE for the product instead of four.
#ifndef GL_SQR_MODE
#define GL_SQR_MODE 1
#endif
#ifndef GL_SQR_TAIL_ALU
#define GL_SQR_TAIL_ALU 0
#endif
__device__ __forceinline__ u64 gl_sqr_lazy(u64 a) {
#if GL_SQR_MODE == 0
    return gl_mul_lazy(a, a);
#else
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), r0, r1;
    asm("{\n\t.reg .u32 x0,x1,x2,x3,t0,t1,d0,d1,d2,r2,n,eh;\n\t.reg .u64 w0,w3,t;\n\t"
        "mul.wide.u32 w0, %2, %2;\n\t"
        "mul.wide.u32 w3, %3, %3;\n\t"
        "mul.wide.u32 t, %2, %3;\n\t"
        "mov.b64 {t0,t1}, t;\n\t"
        "mov.b64 {x0,x1}, w0;\n\t"
        "mov.b64 {x2,x3}, w3;\n\t"
#if GL_SQR_MODE == 1
        "shf.l.wrap.b32 d1, t0, t1, 1;\n\t"
        "shl.b32 d0, t0, 1;\n\t"
        "shr.b32 d2, t1, 31;\n\t"
        "add.cc.u32 x1, x1, d0;\n\t"
        "addc.cc.u32 x2, x2, d1;\n\t"
        "addc.u32 x3, x3, d2;\n\t"
#else
        "add.cc.u32 x1, x1, t0;\n\t"
        "addc.cc.u32 x2, x2, t1;\n\t"
        "addc.u32 x3, x3, 0;\n\t"
        "add.cc.u32 x1, x1, t0;\n\t"
        "addc.cc.u32 x2, x2, t1;\n\t"
        "addc.u32 x3, x3, 0;\n\t"
#endif
#if GL_SQR_TAIL_ALU
        GL_REDUCE_TAIL_ALU("%4", "0")
#else
        GL_REDUCE_TAIL("%4", "0")
#endif
        : "=r"(r0), "=r"(r1)
        : "r"(a0), "r"(a1), "r"(GL_EPS32_C));
    return ((u64)r1 << 32) | r0;
#endif
}

// ---- canonical carry-chain forms for the NTT butterflies (inputs < p, outputs < p) ----
// a - b: 5 SASS instructions (the compiler's compare+select form of gl_sub is 8)
__device__ __forceinline__ u64 gl_sub_c(u64 a, u64 b) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32), r0, r1;
    asm("{\n\t.reg .u32 m;\n\t"
        "sub.cc.u32 %0, %2, %4;\n\t"
        "subc.cc.u32 %1, %3, %5;\n\t"
        "subc.u32 m, 0, 0;\n\t"       // 0xFFFFFFFF when a < b: add p == subtract EPS
        "sub.cc.u32 %0, %0, m;\n\t"
        "subc.u32 %1, %1, 0;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
    return ((u64)r1 << 32) | r0;
}
// p - b in (0, p]: 2 instructions.  p itself (b == 0) is a valid SUBTRAHEND for gl_sub_c.
__device__ __forceinline__ u64 gl_neg_p(u64 b) {
    u32 b0 = (u32)b, b1 = (u32)(b >> 32), r0, r1;
    asm("sub.cc.u32 %0, 1, %2;\n\tsubc.u32 %1, 0xFFFFFFFF, %3;" : "=r"(r0), "=r"(r1) : "r"(b0), "r"(b1));
    return ((u64)r1 << 32) | r0;
}
// (a + b, a - b) in 12 instructions: a + b = a - (p - b)
__device__ __forceinline__ void gl_butterfly_c(u64 &a, u64 &b) {
    u64 nb = gl_neg_p(b);
    u64 d = gl_sub_c(a, b);
    a = gl_sub_c(a, nb);
    b = d;
}
// lazy -> canonical: x >= p  <=>  hi == 0xFFFFFFFF and lo != 0, and then x - p = lo - 1
__device__ __forceinline__ u64 gl_canon_fast(u64 x) {
    u32 lo = (u32)x, hi = (u32)(x >> 32);
    bool ge = (hi == 0xFFFFFFFFu) && (lo != 0);
    return ge ? (u64)(lo - 1) : x;
}

// ---- pipe-balanced forms for the transforms (ntt_kernels.cu) ----
// The NTT kernels are bound by the ALU pipe (76 % busy) while the FMA-heavy pipe idles at ~40 %: carry-OUT adds exist
// only as IADD3 on the ALU, and ptxas also puts most of the carry-IN-only adds there (IADD3.X 710 : IMAD.X 269 in
// pass LO).  An add written as a multiply-add by a run-time 1 (constant bank, opaque to ptxas) can only be an IMAD /
// IMAD.X, i.e. it moves to the FMA pipe at the same instruction count.  GL_FMA_* select which sites are moved.
// Measured on B200 (profiles/r02_ntt_variants.txt): LDE 25.89 -> 24.03 ms, iNTT 3.27 -> 3.03 ms with both on.  Only
// add-type carries are handed to madc: after sub.cc ptxas passes the HARDWARE carry (no borrow = 1) to a following
// madc, not PTX's documented borrow, so a subtraction rewritten that way computes the wrong value (tried, rejected).
static __constant__ u32 GL_ONE_C = 1u, GL_ZERO_C = 0u, GL_MINUS1_C = 0xFFFFFFFFu;
#ifndef GL_FMA_MUL
#define GL_FMA_MUL 0
#endif
#ifndef GL_FMA_CANON
#define GL_FMA_CANON 0
#endif

// a * b -> CANONICAL.  Same product and reduction as gl_mul_lazy; with GL_FMA_MUL the five carry-in-only adds are
// IMAD.X, with GL_FMA_CANON the final canonicalisation is 2 ALU + 3 FMA instructions instead of 2 ISETP + 2 SEL:
//   r >= p  <=>  r + EPS carries out of 64 bits; then r - p = (r1 + 1 : r0 - 1) (r1 = 0xFFFFFFFF wraps to 0, r0 >= 1).
__device__ __forceinline__ u64 gl_mul_canon(u64 a, u64 b) {
#if !GL_FMA_MUL && !GL_FMA_CANON
    return gl_canon_fast(gl_mul_lazy(a, b));
#else
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32), r0, r1;
    asm("{\n\t.reg .u32 x0,x1,x2,x3,t0,t1,t2,r2,n,eh,c0,c1,g;\n\t.reg .u64 w0,w3,t;\n\t"
        "mul.wide.u32 w0, %2, %4;\n\t"
        "mul.wide.u32 w3, %3, %5;\n\t"
        "mul.wide.u32 t, %2, %5;\n\t"
        "mov.b64 {t0,t1}, t;\n\t"
        "mad.lo.cc.u32 t0, %3, %4, t0;\n\t"
        "madc.hi.cc.u32 t1, %3, %4, t1;\n\t"
#if GL_FMA_MUL
        "madc.lo.u32 t2, %8, %7, 0;\n\t"         // 0 * 1 + 0 + carry
#else
        "addc.u32 t2, 0, 0;\n\t"
#endif
        "mov.b64 {x0,x1}, w0;\n\t"
        "mov.b64 {x2,x3}, w3;\n\t"
        "add.cc.u32 x1, x1, t0;\n\t"
        "addc.cc.u32 x2, x2, t1;\n\t"
#if GL_FMA_MUL
        "madc.lo.u32 x3, t2, %7, x3;\n\t"
#else
        "addc.u32 x3, x3, t2;\n\t"
#endif
        "mad.lo.cc.u32 x0, x2, %6, x0;\n\t"
        "madc.hi.cc.u32 x1, x2, %6, x1;\n\t"
#if GL_FMA_MUL
        "madc.lo.u32 r2, %8, %7, 0;\n\t"
#else
        "addc.u32 r2, 0, 0;\n\t"
#endif
        "sub.cc.u32 x0, x0, x3;\n\t"
        "subc.cc.u32 x1, x1, 0;\n\t"
        "subc.u32 r2, r2, 0;\n\t"
#if GL_FMA_MUL
        "mul.lo.u32 n, r2, %9;\n\t"               // -r2
#else
        "sub.u32 n, 0, r2;\n\t"
#endif
        "shr.s32 eh, r2, 31;\n\t"
        "add.cc.u32 x0, x0, n;\n\t"
#if GL_FMA_MUL
        "madc.lo.u32 x1, eh, %7, x1;\n\t"
#else
        "addc.u32 x1, x1, eh;\n\t"
#endif
#if GL_FMA_CANON
        "add.cc.u32 c0, x0, 0xFFFFFFFF;\n\t"     // carries iff x0 != 0
        "addc.cc.u32 c1, x1, 0;\n\t"             // carries iff x1 == 0xFFFFFFFF and x0 != 0, i.e. x >= p
        "madc.lo.u32 g, %8, %7, 0;\n\t"          // g = 1 iff x >= p
        "mad.lo.u32 %0, g, %9, x0;\n\t"          // x0 - g
        "mad.lo.u32 %1, g, %7, x1;\n\t}"         // x1 + g (0xFFFFFFFF wraps to 0)
#else
        "mov.u32 %0, x0;\n\t"
        "mov.u32 %1, x1;\n\t}"
#endif
        : "=r"(r0), "=r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1), "r"(GL_EPS32_C), "r"(GL_ONE_C), "r"(GL_ZERO_C), "r"(GL_MINUS1_C));
    u64 r = ((u64)r1 << 32) | r0;
#if GL_FMA_CANON
    return r;
#else
    return gl_canon_fast(r);
#endif
#endif
}

// ---- multiplication by a power of two (twiddles of the small roots of unity) ----
// In Goldilocks 2 has order 192 (2^96 == -1) and plonky2's omega_16 = omega_{2^32}^(2^28) is 2^156 = -2^60, so every
// power of omega_16 (and of omega_8, omega_4) is +-2^k, k a multiple of 12: a shift into four 32-bit limbs and one
// reduction instead of a 64 x 64 product (1 IMAD.WIDE instead of 5, no table read).
// (x3:x2:x1:x0) - sub_hi 2^32 -> canonical:  (x1:x0) + x2 EPS - (sub_hi:x3); same single signed correction as the multiply
// (sub_hi < 2^31 here, x3 < 2^32), canonicalised the way gl_mul_canon does.
__device__ __forceinline__ u64 gl_reduce4_canon(u32 l0, u32 l1, u32 l2, u32 l3, u32 sub_hi) {
    u32 r0, r1;
    asm("{\n\t.reg .u32 x0,x1,r2,n,eh,c0,c1,g;\n\t"
        "mad.lo.cc.u32 x0, %4, %7, %2;\n\t"
        "madc.hi.cc.u32 x1, %4, %7, %3;\n\t"
        "madc.lo.u32 r2, %9, %8, 0;\n\t"          // carry
        "sub.cc.u32 x0, x0, %5;\n\t"
        "subc.cc.u32 x1, x1, %6;\n\t"
        "subc.u32 r2, r2, 0;\n\t"
        "mul.lo.u32 n, r2, %10;\n\t"               // -r2
        "shr.s32 eh, r2, 31;\n\t"
        "add.cc.u32 x0, x0, n;\n\t"
        "madc.lo.u32 x1, eh, %8, x1;\n\t"
        "add.cc.u32 c0, x0, 0xFFFFFFFF;\n\t"      // carries iff x0 != 0
        "addc.cc.u32 c1, x1, 0;\n\t"              // carries iff x >= p
        "madc.lo.u32 g, %9, %8, 0;\n\t"
        "mad.lo.u32 %0, g, %10, x0;\n\t"
        "mad.lo.u32 %1, g, %8, x1;\n\t}"
        : "=r"(r0), "=r"(r1)
        : "r"(l0), "r"(l1), "r"(l2), "r"(l3), "r"(sub_hi), "r"(GL_EPS32_C), "r"(GL_ONE_C), "r"(GL_ZERO_C), "r"(GL_MINUS1_C));
    return ((u64)r1 << 32) | r0;
}
// x * 2^E (mod p), 0 < E < 96; x canonical -> canonical.  E must be a compile-time constant after unrolling: the limb
// placement below is then resolved by the compiler and the zero limbs fold away.
__device__ __forceinline__ u64 gl_mul_pow2_canon(u64 x, int E) {
    const u32 x0 = (u32)x, x1 = (u32)(x >> 32);
    const int s = E & 31;
    const u32 z0 = x0 << s, z1 = __funnelshift_l(x0, x1, s), z2 = s ? (x1 >> (32 - s)) : 0u;  // x << s, three limbs
    if (E < 32) return gl_reduce4_canon(z0, z1, z2, 0u, 0u);
    if (E < 64) return gl_reduce4_canon(0u, z0, z1, z2, 0u);
    return gl_reduce4_canon(0u, 0u, z0, z1, z2);  // 2^128 == -2^32: the top limb is subtracted one word up
}

// a + b (mod p): a lazy, b CANONICAL (< p) -> lazy.  One wrap fix suffices:
// a + b - 2^64 + EPS <= (2^64-1) + (p-1) - 2^64 + EPS < 2^64.
__device__ __forceinline__ u64 gl_add_lazy_canon(u64 a, u64 b) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32), r0, r1;
    asm("{\n\t.reg .u32 c,m;\n\t"
        "add.cc.u32 %0, %2, %4;\n\t"
        "addc.cc.u32 %1, %3, %5;\n\t"
        "addc.u32 c, 0, 0;\n\t"
        "sub.u32 m, 0, c;\n\t"
        "add.cc.u32 %0, %0, m;\n\t"
        "addc.u32 %1, %1, 0;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
    return ((u64)r1 << 32) | r0;
}

// Five-limb accumulator for sums of products (up to 2^12, see gl_acc_reduce): acc += a * b.
struct gl_acc160 {
    u32 x0, x1, x2, x3, x4;
};
__device__ __forceinline__ void gl_acc_mac(gl_acc160 &A, u64 a, u64 b) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32);
    asm("{\n\t.reg .u32 y0,y1,y2,y3,t0,t1,t2;\n\t.reg .u64 w0,w3,t;\n\t"
        "mul.wide.u32 w0, %5, %7;\n\t"
        "mul.wide.u32 w3, %6, %8;\n\t"
        "mul.wide.u32 t, %5, %8;\n\t"
        "mov.b64 {t0,t1}, t;\n\t"
        "mad.lo.cc.u32 t0, %6, %7, t0;\n\t"
        "madc.hi.cc.u32 t1, %6, %7, t1;\n\t"
        "addc.u32 t2, 0, 0;\n\t"
        "mov.b64 {y0,y1}, w0;\n\t"
        "mov.b64 {y2,y3}, w3;\n\t"
        "add.cc.u32 y1, y1, t0;\n\t"
        "addc.cc.u32 y2, y2, t1;\n\t"
        "addc.u32 y3, y3, t2;\n\t"
        "add.cc.u32 %0, %0, y0;\n\t"
        "addc.cc.u32 %1, %1, y1;\n\t"
        "addc.cc.u32 %2, %2, y2;\n\t"
        "addc.cc.u32 %3, %3, y3;\n\t"
        "addc.u32 %4, %4, 0;\n\t}"
        : "+r"(A.x0), "+r"(A.x1), "+r"(A.x2), "+r"(A.x3), "+r"(A.x4)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
}
// x0 + x1 2^32 + x2 2^64 + x3 2^96 + x4 2^128 == (x1:x0) - (x4:x3) + x2 * EPS   (2^128 == -2^32).
// x4 counts the products (each < 2^128): callers keep a sum below 2^12 products (x4 < 2^12, checked where the
// term count comes from outside: api_fri.cu final_poly_device), so (x4:x3) < 2^44 and the single signed
// correction is still exact: r2 in {-1, 0, 1} only needs the subtrahend's high word below 2^32 - 2.
__device__ __forceinline__ u64 gl_acc_reduce(const gl_acc160 &A) {
    u32 r0, r1;
    asm("{\n\t.reg .u32 x0,x1,x2,x3,r2,n,eh;\n\t"
        "mov.u32 x0, %2;\n\t"
        "mov.u32 x1, %3;\n\t"
        "mov.u32 x2, %4;\n\t"
        "mov.u32 x3, %5;\n\t"
        GL_REDUCE_TAIL_WIDE("%7", "%6")
        : "=r"(r0), "=r"(r1)
        : "r"(A.x0), "r"(A.x1), "r"(A.x2), "r"(A.x3), "r"(A.x4), "r"(GL_EPS32_C));
    return ((u64)r1 << 32) | r0;
}

// Column accumulators for dot products: the partial products of weight 1, 2^32 and 2^64 are summed in
// three separate 64-bit accumulators by the IMAD.WIDE addend itself, and each carry-out is counted in a
// third word (one IADD3.X).  A term costs 4 IMAD.WIDE + 4 carry counts = 8 instructions (the five-limb
// form above: 13); the columns are lined up once per dot product in gl_acc3_reduce.
struct gl_acc3 {
    u32 l0, h0, c0, l1, h1, c1, l2, h2, c2;
};
__device__ __forceinline__ gl_acc3 gl_acc3_zero() { return {0, 0, 0, 0, 0, 0, 0, 0, 0}; }
__device__ __forceinline__ void gl_acc3_mac(gl_acc3 &A, u64 a, u64 b) {
    u32 a0 = (u32)a, a1 = (u32)(a >> 32), b0 = (u32)b, b1 = (u32)(b >> 32);
    asm("mad.lo.cc.u32 %0, %9, %11, %0;\n\t"
        "madc.hi.cc.u32 %1, %9, %11, %1;\n\t"
        "addc.u32 %2, %2, 0;\n\t"
        "mad.lo.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.hi.cc.u32 %4, %9, %12, %4;\n\t"
        "addc.u32 %5, %5, 0;\n\t"
        "mad.lo.cc.u32 %3, %10, %11, %3;\n\t"
        "madc.hi.cc.u32 %4, %10, %11, %4;\n\t"
        "addc.u32 %5, %5, 0;\n\t"
        "mad.lo.cc.u32 %6, %10, %12, %6;\n\t"
        "madc.hi.cc.u32 %7, %10, %12, %7;\n\t"
        "addc.u32 %8, %8, 0;"
        : "+r"(A.l0), "+r"(A.h0), "+r"(A.c0), "+r"(A.l1), "+r"(A.h1), "+r"(A.c1), "+r"(A.l2), "+r"(A.h2), "+r"(A.c2)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
}
// (c0:h0:l0) + (c1:h1:l1) 2^32 + (c2:h2:l2) 2^64 -> five limbs -> lazy u64 (at most 2^12 terms, see gl_acc_reduce)
__device__ __forceinline__ u64 gl_acc3_reduce(const gl_acc3 &A) {
    gl_acc160 X;
    asm("add.cc.u32 %0, %5, %7;\n\t"          // x1 = h0 + l1
        "addc.cc.u32 %1, %6, %8;\n\t"         // x2 = c0 + h1 + carry
        "addc.cc.u32 %2, %9, 0;\n\t"          // x3 = c1 + carry
        "addc.u32 %3, 0, 0;\n\t"              // x4 = carry
        "add.cc.u32 %1, %1, %10;\n\t"         // x2 += l2
        "addc.cc.u32 %2, %2, %11;\n\t"        // x3 += h2
        "addc.u32 %3, %3, %12;"                // x4 += c2
        : "=&r"(X.x1), "=&r"(X.x2), "=&r"(X.x3), "=&r"(X.x4)
        : "r"(0), "r"(A.h0), "r"(A.c0), "r"(A.l1), "r"(A.h1), "r"(A.c1), "r"(A.l2), "r"(A.h2), "r"(A.c2));
    X.x0 = A.l0;
    return gl_acc_reduce(X);
}

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
