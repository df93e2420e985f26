// api_quotient.cu -- the vanishing polynomial on the quotient coset, on the device.
// Replaces plonky2 plonk/prover.rs `compute_quotient_polys` and plonk/vanishing_poly.rs
// `eval_vanishing_poly_base_batch` / `evaluate_gate_constraints_base_batch` (with plonk/plonk_common.rs
// `ZeroPolyOnCoset`, `reduce_with_powers_multi`, plonk/vanishing_poly.rs `check_partial_products`, gates/gate.rs
// `compute_filter`) for circuits made of NoopGate, ConstantGate, PublicInputGate, ArithmeticGate, BaseSumGate<2>,
// ArithmeticExtensionGate<2>, MulExtensionGate<2>, ExponentiationGate, ReducingGate<2>, ReducingExtensionGate<2>,
// RandomAccessGate<2> (up to 4 bits), PoseidonMdsGate<2>, CosetInterpolationGate<2> (this kernel) and PoseidonGate (a second
// kernel that adds its share).
//
// Upstream walks the quotient domain in batches of 32 points and gathers, per point i, the rows
// get_lde_values(i, step) of the constants+sigmas, wires and Zs commitments -- on the CPU that means the whole LDE
// has to be in host memory (9 GB over PCIe at the headline shape).  Here the LDEs never leave HBM: they sit
// column-major in LEAF order, and the rows the quotient needs, reverse_bits(i * step), are exactly the first
// 2^(degree_bits + q) leaves, so thread l reads leaf l of every column (coalesced) and handles the domain point
// i = reverse_bits(l).  Only Z(g x) lives in another row (2 gathered words of ~240).
//
// Field arithmetic is exact, so the order of operations is free: every vanishing term is folded straight into one
// three-column accumulator per alpha with its precomputed power (reduce_with_powers_multi = sum_j alpha^j term_j),
// products run in lazy form, and L_0(x) = Z_H(x) / (n (x - 1)) costs one Fermat inversion per point.
#include "api_internal.h"
#include "poseidon.cuh"

#define QV_MAX_CHALLENGES 4
#define QV_MAX_TERMS 4096  // gl_acc3 is exact for sums of up to 2^12 products
#ifndef QV_BATCH
#define QV_BATCH 4  // wires loaded ahead of the multiplies that consume them (1: 13.2 ms, 2: 11.8, 4: 11.2, 8: 16.4 at 2^20)
#endif

struct QvParams {
    const u64 *cs, *wires, *zs;  // LDE buffers, column-major, leaf order
    size_t stride;               // words per column (2^(degree_bits + rate_bits)), the same for the three batches
    int log_q, qdb;              // quotient domain = 2^log_q points, log_q = degree_bits + qdb
    int nc, n_routed, max_degree, n_chunks, num_constants, num_selectors, n_gates, many_selector;
    int n_terms, gate_base;
    const bsx_gate *gates;
    const u64 *k_is;
    const u64 *apow;  // [nc][n_terms]: alpha_a^j
    const u64 *zh;    // [2^qdb] Z_H on the coset, then [2^qdb] its inverses
    const u64 *ci_tab;  // CosetInterpolationGate: the subgroup of order 2^bits, then its barycentric weights
    u64 betas[QV_MAX_CHALLENGES], gammas[QV_MAX_CHALLENGES], pi_hash[4];
    u64 n_field;      // F::from_canonical_usize(n)
    PowTable tw;      // powers of omega_{2^log_q}
    u64 *out;         // [nc][2^log_q], natural order
};

__device__ __forceinline__ u64 qv_inv(u64 a) {  // a^(p - 2), p - 2 = 2^64 - 2^32 - 1: bits 63..33 and 31..0 set
    u64 t = a;                                   // t_k = a^(2^k - 1)
    u64 t2 = gl_mul_lazy(gl_sqr_lazy(t), a);     // 2 bits
    u64 t4 = gl_mul_lazy(gl_sqr_lazy(gl_sqr_lazy(t2)), t2);
    u64 t8 = t4;
    for (int i = 0; i < 4; i++) t8 = gl_sqr_lazy(t8);
    t8 = gl_mul_lazy(t8, t4);
    u64 t16 = t8;
    for (int i = 0; i < 8; i++) t16 = gl_sqr_lazy(t16);
    t16 = gl_mul_lazy(t16, t8);
    u64 t24 = t16;
    for (int i = 0; i < 8; i++) t24 = gl_sqr_lazy(t24);
    t24 = gl_mul_lazy(t24, t8);
    u64 t28 = t24;
    for (int i = 0; i < 4; i++) t28 = gl_sqr_lazy(t28);
    t28 = gl_mul_lazy(t28, t4);
    u64 t30 = gl_mul_lazy(gl_sqr_lazy(gl_sqr_lazy(t28)), t2);
    u64 t31 = gl_mul_lazy(gl_sqr_lazy(t30), a);
    u64 t32 = gl_mul_lazy(gl_sqr_lazy(t31), a);
    u64 r = t31;  // a^(2^31 - 1) -> shifted up by 33 bits, then the low 32 ones
    for (int i = 0; i < 33; i++) r = gl_sqr_lazy(r);
    return gl_mul_lazy(r, t32);
}

// term (lazy) times alpha_a^idx into every accumulator
template <int NC>
__device__ __forceinline__ void qv_fold(gl_acc3 *res, const QvParams &P, u64 term, int idx) {
#pragma unroll
    for (int a = 0; a < NC; a++) gl_acc3_mac(res[a], term, __ldg(P.apow + (size_t)a * P.n_terms + idx));
}

// gates/gate.rs compute_filter(row, group_range, s, many_selector)
__device__ __forceinline__ u64 qv_filter(const QvParams &P, const bsx_gate &g, const u64 *crow) {
    const u64 s = __ldg(crow + (size_t)g.selector_index * P.stride);
    u64 f = 1;
    for (int r = g.group_start; r < g.group_end; r++)
        if (r != g.row) f = gl_mul_lazy(f, gl_sub_c((u64)r, s));
    if (P.many_selector) f = gl_mul_lazy(f, gl_sub_c(0xFFFFFFFFULL, s));  // UNUSED_SELECTOR
    return f;
}

// Templated on the number of challenges: accumulators, running products and betas are per challenge.
template <int NC>
__global__ void __launch_bounds__(256, 2) quotient_values_kernel(const QvParams P) {
    const size_t l = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // leaf = row of the LDE buffers
    const size_t q_size = (size_t)1 << P.log_q;
    if (l >= q_size) return;
    const u32 i = __brev((u32)l) >> (32 - P.log_q);                   // domain index, x = omega^i
    const u32 i_next = (i + (1u << P.qdb)) & (u32)(q_size - 1);
    const size_t l_next = __brev(i_next) >> (32 - P.log_q);
    constexpr int nc = NC;
    u64 x = __ldg(P.tw.lo + (i & 4095));
    if (i >> 12) x = gl_mul_lazy(x, __ldg(P.tw.hi + (i >> 12)));
    const u64 sx = gl_canon_fast(gl_mul_lazy(x, 7));  // F::coset_shift() * x
    gl_acc3 res[NC];
#pragma unroll
    for (int a = 0; a < NC; a++) res[a] = gl_acc3_zero();

    // ---- vanishing_z_1_terms: L_0(x) (Z(x) - 1), L_0(x) = Z_H(x) / (n (x - 1))
    const u64 zh = __ldg(P.zh + (i & ((1u << P.qdb) - 1)));
    const u64 l0 = gl_mul_lazy(zh, qv_inv(gl_mul_lazy(P.n_field, gl_sub_c(sx, 1))));
    const u64 *zrow = P.zs + l;
#pragma unroll
    for (int k = 0; k < NC; k++)
        if (k < nc) qv_fold<NC>(res, P, gl_mul_lazy(l0, gl_sub_c(__ldg(zrow + (size_t)k * P.stride), 1)), k);

    // ---- vanishing_partial_products_terms: check_partial_products for every challenge, one pass over the wires
    u64 bx[NC];
#pragma unroll
    for (int k = 0; k < NC; k++) bx[k] = k < nc ? gl_mul_lazy(P.betas[k], sx) : 0;
    const u64 *wrow = P.wires + l, *srow = P.cs + (size_t)P.num_constants * P.stride + l;
    for (int m = 0; m < P.n_chunks; m++) {
        u64 num[NC], den[NC];
#pragma unroll
        for (int k = 0; k < NC; k++) num[k] = den[k] = 1;
        const int j0 = m * P.max_degree, j1 = min(P.n_routed, j0 + P.max_degree);
        // QV_BATCH wires at a time: their loads are issued before the first multiply needs one of them
        for (int jb = j0; jb < j1; jb += QV_BATCH) {
            u64 wv[QV_BATCH], sv[QV_BATCH], kv[QV_BATCH];
#pragma unroll
            for (int t = 0; t < QV_BATCH; t++) {
                const int j = min(jb + t, j1 - 1);
                wv[t] = __ldg(wrow + (size_t)j * P.stride);
                sv[t] = __ldg(srow + (size_t)j * P.stride);
                kv[t] = __ldg(P.k_is + j);
            }
#pragma unroll
            for (int t = 0; t < QV_BATCH; t++) {
                if (jb + t >= j1) break;
#pragma unroll
                for (int k = 0; k < NC; k++)
                    if (k < nc) {
                        const u64 wg = gl_add_lazy_canon(wv[t], P.gammas[k]);                 // w + gamma (w < p: no second wrap)
                        num[k] = gl_mul_lazy(num[k], gl_mul_add_lazy(bx[k], kv[t], wg));       // w + beta k_j x + gamma
                        den[k] = gl_mul_lazy(den[k], gl_mul_add_lazy(P.betas[k], sv[t], wg));  // w + beta sigma_j + gamma
                    }
            }
        }
#pragma unroll
        for (int k = 0; k < NC; k++)
            if (k < nc) {
                // product_accs = z_x, partial products of challenge k, z_gx; window (m, m + 1)
                const u64 *pp = zrow + (size_t)(nc + k * (P.n_chunks - 1)) * P.stride;
                const u64 prev = m == 0 ? __ldg(zrow + (size_t)k * P.stride) : __ldg(pp + (size_t)(m - 1) * P.stride);
                const u64 next = m == P.n_chunks - 1 ? __ldg(P.zs + (size_t)k * P.stride + l_next) : __ldg(pp + (size_t)m * P.stride);
                const u64 t = gl_sub_c(gl_canon_fast(gl_mul_lazy(prev, num[k])), gl_canon_fast(gl_mul_lazy(next, den[k])));
                qv_fold<NC>(res, P, t, nc + k * P.n_chunks + m);
            }
    }

    // ---- evaluate_gate_constraints_base_batch: filter * constraint, slot j of every gate shares alpha^(gate_base + j)
    const u64 *crow = P.cs + l;                                          // local_constants (selectors first)
    const u64 *gconst = P.cs + (size_t)P.num_selectors * P.stride + l;   // vars.remove_prefix(num_selectors)
    for (int gi = 0; gi < P.n_gates; gi++) {
        const bsx_gate g = P.gates[gi];
        if (g.kind == BSX_GATE_NOOP) continue;
        if (g.kind == BSX_GATE_POSEIDON) continue;  // quotient_poseidon_kernel
        // sum_j alpha^(gate_base + j) filter c_j = filter * sum_j alpha^(gate_base + j) c_j: the constraints of one gate go
        // into accumulators of their own and the filter multiplies the folded sum once
        gl_acc3 ga[NC];
#pragma unroll
        for (int a = 0; a < NC; a++) ga[a] = gl_acc3_zero();
        auto gfold = [&](u64 c, int slot_abs) {
#pragma unroll
            for (int a = 0; a < NC; a++) gl_acc3_mac(ga[a], c, __ldg(P.apow + (size_t)a * P.n_terms + slot_abs));
        };
        if (g.kind == BSX_GATE_CONSTANT) {
            for (int t = 0; t < g.param; t++) {
                const u64 c = gl_sub_c(__ldg(gconst + (size_t)t * P.stride), __ldg(wrow + (size_t)t * P.stride));
                gfold(c, P.gate_base + t);
            }
        } else if (g.kind == BSX_GATE_PUBLIC_INPUT) {
            for (int t = 0; t < 4; t++)
                gfold(gl_sub_c(__ldg(wrow + (size_t)t * P.stride), P.pi_hash[t]), P.gate_base + t);
        } else if (g.kind == BSX_GATE_ARITHMETIC) {  // output - (m0 m1 const_0 + addend const_1)
            const u64 c0 = __ldg(gconst), c1 = __ldg(gconst + P.stride);
            for (int t = 0; t < g.param; t++) {
                const u64 *wq = wrow + (size_t)(4 * t) * P.stride;
                const u64 m0 = __ldg(wq), m1 = __ldg(wq + P.stride), ad = __ldg(wq + 2 * P.stride), o = __ldg(wq + 3 * P.stride);
                const u64 computed = gl_mul_add_lazy(gl_mul_lazy(m0, m1), c0, gl_mul_lazy(ad, c1));
                gfold(gl_sub_c(o, gl_canon_fast(computed)), P.gate_base + t);
            }
        } else if (g.kind == BSX_GATE_ARITHMETIC_EXTENSION || g.kind == BSX_GATE_MUL_EXTENSION) {
            // output - ((m0 m1).scalar_mul(const_0) [+ addend.scalar_mul(const_1)]) in F_p[X] / (X^2 - 7), two constraints
            // per operation: (a0 + a1 X)(b0 + b1 X) = a0 b0 + 7 a1 b1 + (a0 b1 + a1 b0) X
            const bool arith = g.kind == BSX_GATE_ARITHMETIC_EXTENSION;
            const int per = arith ? 8 : 6;
            const u64 c0 = __ldg(gconst), c1 = arith ? __ldg(gconst + P.stride) : 0;
            for (int t = 0; t < g.param; t++) {
                const u64 *wq = wrow + (size_t)(per * t) * P.stride;
                const u64 a0 = __ldg(wq), a1 = __ldg(wq + P.stride), b0 = __ldg(wq + 2 * P.stride), b1 = __ldg(wq + 3 * P.stride);
                const u64 pr0 = gl_mul_add_lazy(a0, b0, gl_mul_lazy(gl_mul_lazy(a1, b1), 7));
                const u64 pr1 = gl_mul_add_lazy(a0, b1, gl_mul_lazy(a1, b0));
                u64 e0 = gl_mul_lazy(pr0, c0), e1 = gl_mul_lazy(pr1, c0);
                if (arith) {
                    e0 = gl_mul_add_lazy(__ldg(wq + 4 * P.stride), c1, e0);
                    e1 = gl_mul_add_lazy(__ldg(wq + 5 * P.stride), c1, e1);
                }
                const u64 o0 = __ldg(wq + (size_t)(per - 2) * P.stride), o1 = __ldg(wq + (size_t)(per - 1) * P.stride);
                gfold(gl_sub_c(o0, gl_canon_fast(e0)), P.gate_base + 2 * t);
                gfold(gl_sub_c(o1, gl_canon_fast(e1)), P.gate_base + 2 * t + 1);
            }
        } else if (g.kind == BSX_GATE_EXPONENTIATION) {
            // wires: base 0, power bits 1.. (little-endian), output 1 + n, intermediate values 2 + n..;
            // prev^2 (bit base + 1 - bit) - intermediate_i with the bits taken from the top, then output - last
            const int n = g.param;
            const u64 base = __ldg(wrow);
            u64 prev = 1;
            for (int t = 0; t < n; t++) {
                const u64 bit = __ldg(wrow + (size_t)(n - t) * P.stride), iv = __ldg(wrow + (size_t)(2 + n + t) * P.stride);
                const u64 sel = gl_canon_fast(gl_mul_add_lazy(bit, gl_sub_c(base, 1), 1));  // bit base + (1 - bit)
                gfold(gl_sub_c(gl_canon_fast(gl_mul_lazy(prev, sel)), iv), P.gate_base + t);
                prev = gl_sqr_lazy(iv);
            }
            const u64 last = __ldg(wrow + (size_t)(2 + n + n - 1) * P.stride);
            gfold(gl_sub_c(__ldg(wrow + (size_t)(1 + n) * P.stride), last), P.gate_base + n);
        } else if (g.kind == BSX_GATE_REDUCING || g.kind == BSX_GATE_REDUCING_EXTENSION) {
            // wires: output 0..2, alpha 2..4, old_acc 4..6, coeffs from 6 (one or two words each), then the accumulators (the
            // last one is the output): acc alpha + coeff_i - acc_i in F_p[X] / (X^2 - 7)
            const int n = g.param, cw = g.kind == BSX_GATE_REDUCING_EXTENSION ? 2 : 1, start_accs = 6 + n * cw;
            const u64 al0 = __ldg(wrow + 2 * P.stride), al1 = __ldg(wrow + 3 * P.stride);
            u64 a0 = __ldg(wrow + 4 * P.stride), a1 = __ldg(wrow + 5 * P.stride);
            for (int t = 0; t < n; t++) {
                const u64 *acc_t = wrow + (size_t)(t == n - 1 ? 0 : start_accs + 2 * t) * P.stride;
                const u64 n0 = __ldg(acc_t), n1 = __ldg(acc_t + P.stride);
                const u64 co0 = __ldg(wrow + (size_t)(6 + cw * t) * P.stride);
                u64 t0 = gl_mul_add_lazy(a0, al0, gl_mul_lazy(gl_mul_lazy(a1, al1), 7));
                u64 t1 = gl_mul_add_lazy(a0, al1, gl_mul_lazy(a1, al0));
                t0 = gl_add_lazy_canon(t0, co0);
                if (cw == 2) t1 = gl_add_lazy_canon(t1, __ldg(wrow + (size_t)(6 + 2 * t + 1) * P.stride));
                gfold(gl_sub_c(gl_canon_fast(t0), n0), P.gate_base + 2 * t);
                gfold(gl_sub_c(gl_canon_fast(t1), n1), P.gate_base + 2 * t + 1);
                a0 = n0, a1 = n1;
            }
        } else if (g.kind == BSX_GATE_RANDOM_ACCESS) {
            // per copy: access_index, claimed_element, 2^bits list items; the index bits sit after the routed wires.
            // bit (bit - 1) per bit, bits recomposed - index, the list folded by the bits - claimed element; then the extra
            // constants.  bits <= 4 here: the list lives in 16 registers.
            const int bits = g.param & 0xFF, copies = (g.param >> 8) & 0xFF, extra = g.param >> 16, vec = 1 << bits;
            const int routed = (2 + vec) * copies + extra;
            int slot = P.gate_base;
            for (int cp = 0; cp < copies; cp++) {
                const u64 *q = wrow + (size_t)((2 + vec) * cp) * P.stride, *bw = wrow + (size_t)(routed + cp * bits) * P.stride;
                u64 b[4], list[16];
                u64 rec = 0;
#pragma unroll
                for (int t = 3; t >= 0; t--) {
                    b[t] = t < bits ? __ldg(bw + (size_t)t * P.stride) : 0;
                    if (t < bits) rec = gl_add_lazy_canon(gl_add_lazy_canon(rec, gl_canon_fast(rec)), b[t]);
                }
#pragma unroll
                for (int t = 0; t < 4; t++)
                    if (t < bits) gfold(gl_mul_lazy(b[t], gl_sub_c(b[t], 1)), slot++);
                gfold(gl_sub_c(gl_canon_fast(rec), __ldg(q)), slot++);
#pragma unroll
                for (int j = 0; j < 16; j++) list[j] = j < vec ? __ldg(q + (size_t)(2 + j) * P.stride) : 0;
#pragma unroll
                for (int t = 0; t < 4; t++) {
                    if (t < bits) {
#pragma unroll
                        for (int j = 0; j < (8 >> t); j++)  // x + b (y - x), canonical
                            list[j] = gl_canon_fast(gl_mul_add_lazy(b[t], gl_sub_c(list[2 * j + 1], list[2 * j]), list[2 * j]));
                    }
                }
                gfold(gl_sub_c(list[0], __ldg(q + P.stride)), slot++);
            }
            for (int t = 0; t < extra; t++)
                gfold(gl_sub_c(__ldg(gconst + (size_t)t * P.stride),
                                                        __ldg(wrow + (size_t)((2 + vec) * copies + t) * P.stride)), slot++);
        } else if (g.kind == BSX_GATE_COSET_INTERPOLATION) {
            // wires: shift 0, values 1 + 2 q, evaluation point, evaluation value, intermediate evals, intermediate prods,
            // shifted evaluation point.  evaluation_point - shifted_point * shift; the barycentric sum over the subgroup in
            // chunks (first `degree` points, then degree - 1 each), (eval, prod) <- (eval (x - x_q) + v_q w_q prod,
            // prod (x - x_q)) in F_p[X]/(X^2 - 7), pinned to wires after every chunk; evaluation_value - the last eval
            const int bits = g.param & 0xFF, degree = g.param >> 8, np = 1 << bits, ni = (np - 2) / (degree - 1);
            const int ep = 1 + 2 * np, ev = ep + 2, im = ev + 2, sp = im + 4 * ni;
            const u64 *dom = P.ci_tab, *wts = P.ci_tab + np;
            const u64 shift = __ldg(wrow), x0 = __ldg(wrow + (size_t)sp * P.stride), x1 = __ldg(wrow + (size_t)(sp + 1) * P.stride);
            int slot = P.gate_base;
            gfold(gl_sub_c(__ldg(wrow + (size_t)ep * P.stride), gl_canon_fast(gl_mul_lazy(x0, shift))), slot++);
            gfold(gl_sub_c(__ldg(wrow + (size_t)(ep + 1) * P.stride), gl_canon_fast(gl_mul_lazy(x1, shift))), slot++);
            u64 e0 = 0, e1 = 0, p0 = 1, p1 = 0;  // canonical between the steps
            int first = 0, last = degree;
            for (int i = 0; i <= ni; i++) {
                for (int q = first; q < last; q++) {
                    const u64 w = __ldg(wts + q);
                    const u64 v0 = gl_mul_lazy(__ldg(wrow + (size_t)(1 + 2 * q) * P.stride), w);
                    const u64 v1 = gl_mul_lazy(__ldg(wrow + (size_t)(2 + 2 * q) * P.stride), w);
                    const u64 t0 = gl_sub_c(x0, __ldg(dom + q)), t1 = x1;
                    // (a0 + a1 X)(b0 + b1 X) = a0 b0 + 7 a1 b1 + (a0 b1 + a1 b0) X, for eval * term + val * prod and prod * term
                    const u64 n0 = gl_mul_add_lazy(e0, t0, gl_mul_add_lazy(v0, p0, gl_mul_lazy(gl_add_lazy_canon(gl_mul_lazy(e1, t1), gl_canon_fast(gl_mul_lazy(v1, p1))), 7)));
                    const u64 n1 = gl_mul_add_lazy(e0, t1, gl_mul_add_lazy(e1, t0, gl_mul_add_lazy(v0, p1, gl_mul_lazy(v1, p0))));
                    const u64 q0 = gl_mul_add_lazy(p0, t0, gl_mul_lazy(gl_mul_lazy(p1, t1), 7));
                    const u64 q1 = gl_mul_add_lazy(p0, t1, gl_mul_lazy(p1, t0));
                    e0 = gl_canon_fast(n0), e1 = gl_canon_fast(n1), p0 = gl_canon_fast(q0), p1 = gl_canon_fast(q1);
                }
                if (i == ni) break;
                const u64 *ie = wrow + (size_t)(im + 2 * i) * P.stride, *ip = wrow + (size_t)(im + 2 * (ni + i)) * P.stride;
                const u64 ie0 = __ldg(ie), ie1 = __ldg(ie + P.stride), ip0 = __ldg(ip), ip1 = __ldg(ip + P.stride);
                gfold(gl_sub_c(ie0, e0), slot++);
                gfold(gl_sub_c(ie1, e1), slot++);
                gfold(gl_sub_c(ip0, p0), slot++);
                gfold(gl_sub_c(ip1, p1), slot++);
                e0 = ie0, e1 = ie1, p0 = ip0, p1 = ip1;
                first = 1 + (degree - 1) * (i + 1);
                last = min(first + degree - 1, np);
            }
            gfold(gl_sub_c(__ldg(wrow + (size_t)ev * P.stride), e0), slot++);
            gfold(gl_sub_c(__ldg(wrow + (size_t)(ev + 1) * P.stride), e1), slot++);
        } else if (g.kind == BSX_GATE_POSEIDON_MDS) {
            // output_r - (sum_i CIRC[i] input_{(i + r) % 12} + 8 input_0 [r == 0]), the two coordinates of the algebra
            // elements separately: the multiplication-free layer of the permutation with zero constants
            u32 zero[36];
#pragma unroll
            for (int t = 0; t < 36; t++) zero[t] = 0;
            for (int d = 0; d < 2; d++) {
                u64 lane[12];
#pragma unroll
                for (int t = 0; t < 12; t++) lane[t] = __ldg(wrow + (size_t)(2 * t + d) * P.stride);
                poseidon_mds_post_fft(lane, zero);
#pragma unroll
                for (int t = 0; t < 12; t++)
                    gfold(gl_sub_c(__ldg(wrow + (size_t)(24 + 2 * t + d) * P.stride), gl_canon_fast(lane[t])), P.gate_base + 2 * t + d);
            }
        } else if (g.kind == BSX_GATE_BASE_SUM) {  // reduce_with_powers(limbs, 2) - sum; limb (limb - 1) per limb
            u64 acc = 0;
            for (int t = g.param; t >= 1; t--) {
                const u64 limb = __ldg(wrow + (size_t)t * P.stride);
                acc = gl_mul_add_lazy(acc, 2, limb);
                gfold(gl_mul_lazy(limb, gl_sub_c(limb, 1)), P.gate_base + t);
            }
            gfold(gl_sub_c(gl_canon_fast(acc), __ldg(wrow)), P.gate_base);
        }
        const u64 f = qv_filter(P, g, crow);
#pragma unroll
        for (int a = 0; a < NC; a++) gl_acc3_mac(res[a], f, gl_acc3_reduce(ga[a]));
    }

    // ---- * z_h_on_coset.eval_inverse(i)
    const u64 zh_inv = __ldg(P.zh + (1u << P.qdb) + (i & ((1u << P.qdb) - 1)));
#pragma unroll
    for (int a = 0; a < NC; a++)
        if (a < nc) P.out[(size_t)a * q_size + i] = gl_canon_fast(gl_mul_lazy(gl_acc3_reduce(res[a]), zh_inv));
}

// gates/poseidon.rs PoseidonGate wire map
#define PG_OUT 12
#define PG_SWAP 24
#define PG_DELTA 25
#define PG_FULL0 29
#define PG_PARTIAL 65
#define PG_FULL1 87
#define PG_WIRES 135
#define PG_CONSTRAINTS 123

// The PoseidonGate's share of the vanishing polynomial: out[a][i] += filter * sum_j alpha_a^(gate_base + j) c_j / Z_H.
// PoseidonGate::eval_unfiltered_base_one walks the permutation from the (possibly swapped) inputs; wherever the S-box
// inputs of a round are wires, the running state is compared with them (one constraint each) and continues from the
// wires.  One thread per point, the state never leaves registers, and the schedule is poseidon_permute_lazy's
// (poseidon.cuh): S-box, then the multiplication-free MDS layer with the next round's constants; lanes 1..11 stay in
// limb planes through the partial rounds.  Those are taken in the naive form -- upstream's factored form
// (mds_partial_layer_init / _fast) yields the same lane-0 values for any wires, it is the same affine maps regrouped
// (tests/test_gates.py).  The filter multiplies the folded sum once instead of each of the 123 constraints.
// Templated on the number of challenges: the accumulators are 9 registers each.
template <int NC>
__global__ void __launch_bounds__(128, 4) quotient_poseidon_kernel(const QvParams P, const bsx_gate g) {
    const size_t l = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t q_size = (size_t)1 << P.log_q;
    if (l >= q_size) return;
    const u32 i = __brev((u32)l) >> (32 - P.log_q);
    const u64 *w = P.wires + l;
#define PG_W(j) __ldg(w + (size_t)(j) * P.stride)
    gl_acc3 res[NC];
#pragma unroll
    for (int a = 0; a < NC; a++) res[a] = gl_acc3_zero();
    const u64 *apow = P.apow + P.gate_base;
    int slot = 0;
    auto fold = [&](u64 c) {
#pragma unroll
        for (int a = 0; a < NC; a++) gl_acc3_mac(res[a], c, __ldg(apow + (size_t)a * P.n_terms + slot));
        slot++;
    };
    u64 s[12];
    {
        const u64 swap = PG_W(PG_SWAP);
        fold(gl_mul_lazy(swap, gl_sub_c(swap, 1)));
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const u64 lhs = PG_W(k), rhs = PG_W(k + 4), d = PG_W(PG_DELTA + k);
            fold(gl_sub_c(gl_canon_fast(gl_mul_lazy(swap, gl_sub_c(rhs, lhs))), d));
            s[k] = gl_add_lazy_canon(gl_add_lazy_canon(lhs, d), POSEIDON_RC[k]);
            s[k + 4] = gl_add_lazy_canon(gl_sub_c(rhs, d), POSEIDON_RC[k + 4]);
        }
#pragma unroll
        for (int k = 8; k < 12; k++) s[k] = gl_add_lazy_canon(PG_W(k), POSEIDON_RC[k]);
    }
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
#pragma unroll 1
        for (int r = 0; r < 4; r++) {
            if (half + r) {  // s = the state after this round's constant layer: the wires say what it has to be
                const int w0 = half ? PG_FULL1 + 12 * r : PG_FULL0 + 12 * (r - 1);
#pragma unroll
                for (int k = 0; k < 12; k++) {
                    const u64 win = PG_W(w0 + k);
                    fold(gl_sub_c(gl_canon_fast(s[k]), win));
                    s[k] = win;
                }
            }
#pragma unroll
            for (int k = 0; k < 12; k++) s[k] = poseidon_sbox(s[k]);
            u32 post[36];
            poseidon_load_post(half * 26 + r, post);
            poseidon_mds_post_fft(s, post);
        }
        if (half == 0) {
            u32 x0[12], x1[12], x2[12], a0, a1, a2;
#pragma unroll
            for (int k = 1; k < 12; k++) poseidon_split22(s[k], x0[k], x1[k], x2[k]);
            poseidon_split22(s[0], a0, a1, a2);
#pragma unroll 1
            for (int p = 4; p < 26; p++) {
                const u64 win = PG_W(PG_PARTIAL + p - 4);
                fold(gl_sub_c(gl_canon_fast(poseidon_merge22(a0, a1, a2)), win));
                u32 post[36];
                poseidon_load_post(p, post);
                poseidon_split22(poseidon_sbox(win), x0[0], x1[0], x2[0]);
                u32 y0[12], y1[12], y2[12];
                poseidon_mds_plane(x0, post, y0);
                poseidon_mds_plane(x1, post + 12, y1);
                poseidon_mds_plane(x2, post + 24, y2);
                a0 = y0[0], a1 = y1[0], a2 = y2[0];
#pragma unroll
                for (int k = 1; k < 12; k++) {
                    poseidon_norm22(y0[k], y1[k], y2[k]);
                    x0[k] = y0[k], x1[k] = y1[k], x2[k] = y2[k];
                }
            }
            s[0] = poseidon_merge22(a0, a1, a2);
#pragma unroll
            for (int k = 1; k < 12; k++) s[k] = poseidon_merge22(x0[k], x1[k], x2[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < 12; k++) fold(gl_sub_c(gl_canon_fast(s[k]), PG_W(PG_OUT + k)));
#undef PG_W
    const u64 f = qv_filter(P, g, P.cs + l);
    const u64 scale = gl_mul_lazy(f, __ldg(P.zh + (1u << P.qdb) + (i & ((1u << P.qdb) - 1))));
#pragma unroll
    for (int a = 0; a < NC; a++) {
        u64 *o = P.out + (size_t)a * q_size + i;
        *o = gl_canon_fast(gl_mul_add_lazy(gl_acc3_reduce(res[a]), scale, *o));
    }
}

// gates/poseidon.rs PoseidonGenerator::run_once on row-major rows of 135 wires: inputs and swap are read, the deltas, the
// S-box input wires and the outputs are written (canonical).  One thread per row.
__global__ void __launch_bounds__(128) poseidon_gate_witness_kernel(u64 *rows, size_t n) {
    const size_t row = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    u64 *w = rows + row * PG_WIRES;
    u64 s[12];
    const u64 swap = gl_canon(w[PG_SWAP]);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const u64 lhs = gl_canon(w[k]), rhs = gl_canon(w[k + 4]);
        const u64 d = gl_canon_fast(gl_mul_lazy(swap, gl_sub_c(rhs, lhs)));
        w[PG_DELTA + k] = d;
        s[k] = gl_add_lazy_canon(gl_add_lazy_canon(lhs, d), POSEIDON_RC[k]);
        s[k + 4] = gl_add_lazy_canon(gl_sub_c(rhs, d), POSEIDON_RC[k + 4]);
    }
#pragma unroll
    for (int k = 8; k < 12; k++) s[k] = gl_add_lazy_canon(gl_canon(w[k]), POSEIDON_RC[k]);
#pragma unroll 1
    for (int r = 0; r < 30; r++) {
        const bool full = r < 4 || r >= 26;
        if (r >= 1) {
            const int w0 = r < 4 ? PG_FULL0 + 12 * (r - 1) : r < 26 ? PG_PARTIAL + (r - 4) : PG_FULL1 + 12 * (r - 26);
            if (full) {
#pragma unroll
                for (int k = 0; k < 12; k++) w[w0 + k] = s[k] = gl_canon(s[k]);
            } else {
                w[w0] = s[0] = gl_canon(s[0]);
            }
        }
        if (full) {
#pragma unroll
            for (int k = 0; k < 12; k++) s[k] = poseidon_sbox(s[k]);
        } else {
            s[0] = poseidon_sbox(s[0]);
        }
        u32 post[36];
        poseidon_load_post(r, post);
        poseidon_mds_post_fft(s, post);
    }
#pragma unroll
    for (int k = 0; k < 12; k++) w[PG_OUT + k] = gl_canon(s[k]);
}

extern "C" int32_t bsx_poseidon_gate_witness(bsx_ctx *ctx, uint64_t *rows, int32_t mem, size_t n) {
    if (!ctx || (!rows && n)) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (n == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *d = (u64 *)rows;
    if (mem != BSX_DEVICE) {
        TRY(bsx_alloc(ctx, &d, n * PG_WIRES));
        TRY(bsx_copy_in(ctx, d, (const u64 *)rows, n * PG_WIRES, mem));
    }
    BSX_COUNT_LAUNCH();
    poseidon_gate_witness_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(d, n);
    CU(ctx, cudaGetLastError());
    if (mem != BSX_DEVICE) {
        TRY(bsx_copy_out(ctx, (u64 *)rows, d, n * PG_WIRES, mem));
        bsx_free(ctx, d);
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

static int gate_constraints(const bsx_gate &g) {
    switch (g.kind) {
        case BSX_GATE_NOOP: return 0;
        case BSX_GATE_CONSTANT: return g.param;
        case BSX_GATE_PUBLIC_INPUT: return 4;
        case BSX_GATE_ARITHMETIC: return g.param;
        case BSX_GATE_POSEIDON: return PG_CONSTRAINTS;
        case BSX_GATE_ARITHMETIC_EXTENSION: return g.param < 0 ? -1 : 2 * g.param;
        case BSX_GATE_MUL_EXTENSION: return g.param < 0 ? -1 : 2 * g.param;
        case BSX_GATE_BASE_SUM: return g.param < 1 ? -1 : 1 + g.param;
        case BSX_GATE_EXPONENTIATION: return g.param < 1 ? -1 : g.param + 1;
        case BSX_GATE_REDUCING: return g.param < 1 ? -1 : 2 * g.param;
        case BSX_GATE_REDUCING_EXTENSION: return g.param < 1 ? -1 : 2 * g.param;
        case BSX_GATE_RANDOM_ACCESS: {
            const int bits = g.param & 0xFF, copies = (g.param >> 8) & 0xFF, extra = g.param >> 16;
            return bits < 1 || bits > 4 || copies < 1 || extra < 0 ? -1 : copies * (bits + 2) + extra;
        }
        case BSX_GATE_POSEIDON_MDS: return 24;
        case BSX_GATE_COSET_INTERPOLATION: {
            const int bits = g.param & 0xFF, degree = g.param >> 8;
            return bits < 1 || bits > 8 || degree < 2 || degree > (1 << bits) ? -1 : 2 * (2 + 2 * (((1 << bits) - 2) / (degree - 1)));
        }
        default: return -1;
    }
}

static int log2_ceil(int x) {
    int b = 0;
    while ((1 << b) < x) b++;
    return b;
}

// values on the device: out_dev[a][i]
static int quotient_values_device(bsx_ctx *ctx, const bsx_circuit *c, bsx_batch *cs, bsx_batch *wires, bsx_batch *zs,
                                  const uint64_t *pi_hash, const uint64_t *betas, const uint64_t *gammas,
                                  const uint64_t *alphas, u64 *out_dev) {
    const int nc = c->num_challenges, max_degree = c->quotient_degree_factor;
    if (nc < 1 || nc > QV_MAX_CHALLENGES)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "num_challenges=%d outside [1, %d]", nc, QV_MAX_CHALLENGES);
    if (max_degree < 2) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "quotient_degree_factor=%d: check_partial_products needs > 1", max_degree);
    const int qdb = log2_ceil(max_degree);
    if (qdb > c->rate_bits)  // upstream assert
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "Having constraints of degree higher than the rate is not supported yet.");
    if (c->degree_bits < 0 || c->degree_bits + qdb > 24 || c->degree_bits + qdb < 1)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "degree_bits=%d + quotient_degree_bits=%d outside [1, 24]", c->degree_bits, qdb);
    if (c->num_routed_wires < 1 || c->num_routed_wires > c->num_wires || !c->k_is || (c->n_gates > 0 && !c->gates))
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad circuit description");
    const int n_chunks = (c->num_routed_wires + max_degree - 1) / max_degree;
    if (c->num_partial_products != n_chunks - 1)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "num_partial_products=%d, expected %d", c->num_partial_products, n_chunks - 1);
    bsx_batch *bs[3] = {cs, wires, zs};
    const int want_cols[3] = {c->num_constants + c->num_routed_wires, c->num_wires, nc * n_chunks};
    for (int k = 0; k < 3; k++) {
        if (!bs[k] || bs[k]->ctx != ctx) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "commitment %d belongs to another context", k);
        if (bs[k]->log_n != c->degree_bits || bs[k]->rate_bits != c->rate_bits || bs[k]->n_cols != want_cols[k])
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "commitment %d has shape 2^%d x %d rate_bits %d, the circuit wants 2^%d x %d rate_bits %d",
                            k, bs[k]->log_n, bs[k]->n_cols, bs[k]->rate_bits, c->degree_bits, want_cols[k], c->rate_bits);
    }
    if (c->num_selectors < 0 || c->num_selectors > c->num_constants) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad selector count");
    const int n_gate_consts = c->num_constants - c->num_selectors;
    int ci_gate = -1;
    for (int g = 0; g < c->n_gates; g++) {
        const bsx_gate &G = c->gates[g];
        const int k = gate_constraints(G);
        bool ok = k >= 0 && k <= c->num_gate_constraints && G.selector_index >= 0 && G.selector_index < c->num_selectors &&
                  G.group_start <= G.row && G.row < G.group_end && G.group_end - G.group_start <= 64;
        if (ok && G.kind == BSX_GATE_CONSTANT) ok = G.param <= n_gate_consts && G.param <= c->num_wires;
        if (ok && G.kind == BSX_GATE_PUBLIC_INPUT) ok = c->num_wires >= 4;
        if (ok && G.kind == BSX_GATE_ARITHMETIC) ok = 4 * G.param <= c->num_wires && n_gate_consts >= 2;
        if (ok && G.kind == BSX_GATE_POSEIDON) ok = c->num_wires >= PG_WIRES;
        if (ok && G.kind == BSX_GATE_ARITHMETIC_EXTENSION) ok = 8 * G.param <= c->num_wires && n_gate_consts >= 2;
        if (ok && G.kind == BSX_GATE_MUL_EXTENSION) ok = 6 * G.param <= c->num_wires && n_gate_consts >= 1;
        if (ok && G.kind == BSX_GATE_BASE_SUM) ok = G.param <= 63 && 1 + G.param <= c->num_wires;
        if (ok && G.kind == BSX_GATE_EXPONENTIATION) ok = 2 + 2 * G.param <= c->num_wires;
        if (ok && G.kind == BSX_GATE_REDUCING) ok = 6 + G.param + 2 * (G.param - 1) <= c->num_wires;
        if (ok && G.kind == BSX_GATE_REDUCING_EXTENSION) ok = 6 + 2 * G.param + 2 * (G.param - 1) <= c->num_wires;
        if (ok && G.kind == BSX_GATE_RANDOM_ACCESS) {
            const int bits = G.param & 0xFF, copies = (G.param >> 8) & 0xFF, extra = G.param >> 16;
            ok = (2 + (1 << bits)) * copies + extra + copies * bits <= c->num_wires && extra <= n_gate_consts;
        }
        if (ok && G.kind == BSX_GATE_POSEIDON_MDS) ok = c->num_wires >= 48;
        if (ok && G.kind == BSX_GATE_COSET_INTERPOLATION) {
            const int bits = G.param & 0xFF, degree = G.param >> 8, np = 1 << bits;
            ok = 1 + 2 * np + 6 + 4 * ((np - 2) / (degree - 1)) <= c->num_wires && ci_gate < 0;  // one such gate per circuit
            ci_gate = g;
        }
        if (!ok) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "gate %d (kind %d, param %d) does not fit the circuit", g, G.kind, G.param);
    }
    QvParams P;
    P.nc = nc;
    P.n_terms = nc + nc * n_chunks + c->num_gate_constraints;
    P.gate_base = nc + nc * n_chunks;
    if (P.n_terms > QV_MAX_TERMS) return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "%d vanishing terms (at most %d)", P.n_terms, QV_MAX_TERMS);
    CU(ctx, cudaSetDevice(ctx->device));
    // host-side tables: alpha powers, Z_H on the coset and its inverses (ZeroPolyOnCoset::new), k_is, gates
    const int rate = 1 << qdb;
    const int ci_np = ci_gate >= 0 ? 1 << (c->gates[ci_gate].param & 0xFF) : 0;
    std::vector<u64> h((size_t)nc * P.n_terms + 2 * rate + c->num_routed_wires + 2 * (size_t)ci_np);
    for (int a = 0; a < nc; a++) {
        u64 pw = 1, al = alphas[a] % GL_P;
        for (int j = 0; j < P.n_terms; j++) {
            h[(size_t)a * P.n_terms + j] = pw;
            pw = gl_host_mul(pw, al);
        }
    }
    u64 g_pow_n = 7;
    for (int k = 0; k < c->degree_bits; k++) g_pow_n = gl_host_mul(g_pow_n, g_pow_n);
    const u64 v = gl_host_root(qdb);
    u64 xv = 1;
    u64 *zh = h.data() + (size_t)nc * P.n_terms;
    for (int k = 0; k < rate; k++) {
        const u64 e = gl_host_mul(g_pow_n, xv);
        zh[k] = e ? e - 1 : GL_P - 1;
        if (zh[k] == 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "Z_H vanishes on the coset");
        zh[rate + k] = gl_host_inv(zh[k]);
        xv = gl_host_mul(xv, v);
    }
    for (int j = 0; j < c->num_routed_wires; j++) zh[2 * rate + j] = c->k_is[j] % GL_P;
    if (ci_np) {  // F::two_adic_subgroup(bits) and field/interpolation.rs barycentric_weights on it
        u64 *dom = zh + 2 * rate + c->num_routed_wires, *wts = dom + ci_np;
        const u64 gen = gl_host_root(c->gates[ci_gate].param & 0xFF);
        u64 xg = 1;
        for (int k = 0; k < ci_np; k++, xg = gl_host_mul(xg, gen)) dom[k] = xg;
        for (int k = 0; k < ci_np; k++) {
            u64 pr = 1;
            for (int j = 0; j < ci_np; j++)
                if (j != k) pr = gl_host_mul(pr, dom[k] >= dom[j] ? dom[k] - dom[j] : dom[k] + (GL_P - dom[j]));
            wts[k] = gl_host_inv(pr);
        }
    }
    u64 *tab;
    bsx_gate *gates_dev = nullptr;
    TRY(bsx_alloc(ctx, &tab, h.size()));
    TRY(bsx_copy_in(ctx, tab, h.data(), h.size(), BSX_HOST));
    if (c->n_gates > 0) {
        const size_t gw = ((size_t)c->n_gates * sizeof(bsx_gate) + 7) / 8;
        TRY(bsx_alloc(ctx, (u64 **)&gates_dev, gw));
        CU(ctx, cudaMemcpyAsync(gates_dev, c->gates, (size_t)c->n_gates * sizeof(bsx_gate), cudaMemcpyHostToDevice, ctx->stream));
    }
    P.cs = cs->lde;
    P.wires = wires->lde;
    P.zs = zs->lde;
    P.stride = (size_t)1 << (c->degree_bits + c->rate_bits);
    P.log_q = c->degree_bits + qdb;
    P.qdb = qdb;
    P.n_routed = c->num_routed_wires;
    P.max_degree = max_degree;
    P.n_chunks = n_chunks;
    P.num_constants = c->num_constants;
    P.num_selectors = c->num_selectors;
    P.n_gates = c->n_gates;
    P.many_selector = c->num_selectors > 1;
    P.gates = gates_dev;
    P.apow = tab;
    P.zh = tab + (size_t)nc * P.n_terms;
    P.k_is = P.zh + 2 * rate;
    P.ci_tab = P.k_is + c->num_routed_wires;
    for (int k = 0; k < QV_MAX_CHALLENGES; k++) {
        P.betas[k] = k < nc ? betas[k] % GL_P : 0;
        P.gammas[k] = k < nc ? gammas[k] % GL_P : 0;
    }
    for (int k = 0; k < 4; k++) P.pi_hash[k] = pi_hash ? pi_hash[k] % GL_P : 0;
    P.n_field = ((u64)1 << c->degree_bits) % GL_P;
    TRY(bsx_pow_table(ctx, P.log_q, 0, &P.tw));
    P.out = out_dev;
    const size_t q_size = (size_t)1 << P.log_q;
    BSX_COUNT_LAUNCH();
    {
        const unsigned blocks = (unsigned)((q_size + 255) / 256);
        switch (nc) {
            case 1: quotient_values_kernel<1><<<blocks, 256, 0, ctx->stream>>>(P); break;
            case 2: quotient_values_kernel<2><<<blocks, 256, 0, ctx->stream>>>(P); break;
            case 3: quotient_values_kernel<3><<<blocks, 256, 0, ctx->stream>>>(P); break;
            default: quotient_values_kernel<4><<<blocks, 256, 0, ctx->stream>>>(P); break;
        }
    }
    CU(ctx, cudaGetLastError());
    for (int g = 0; g < c->n_gates; g++)
        if (c->gates[g].kind == BSX_GATE_POSEIDON) {
            BSX_COUNT_LAUNCH();
            const unsigned blocks = (unsigned)((q_size + 127) / 128);
            switch (nc) {
                case 1: quotient_poseidon_kernel<1><<<blocks, 128, 0, ctx->stream>>>(P, c->gates[g]); break;
                case 2: quotient_poseidon_kernel<2><<<blocks, 128, 0, ctx->stream>>>(P, c->gates[g]); break;
                case 3: quotient_poseidon_kernel<3><<<blocks, 128, 0, ctx->stream>>>(P, c->gates[g]); break;
                default: quotient_poseidon_kernel<4><<<blocks, 128, 0, ctx->stream>>>(P, c->gates[g]); break;
            }
            CU(ctx, cudaGetLastError());
        }
    bsx_free(ctx, tab);
    bsx_free(ctx, (u64 *)gates_dev);
    return BSX_OK;
}

extern "C" int32_t bsx_quotient_values(bsx_ctx *ctx, const bsx_circuit *circuit, bsx_batch *constants_sigmas,
                                       bsx_batch *wires, bsx_batch *zs_partial_products,
                                       const uint64_t *public_inputs_hash, const uint64_t *betas, const uint64_t *gammas,
                                       const uint64_t *alphas, uint64_t *out, int32_t out_mem) {
    if (!ctx || !circuit || !betas || !gammas || !alphas || !out) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    const int qdb = log2_ceil(circuit->quotient_degree_factor < 1 ? 1 : circuit->quotient_degree_factor);
    if (circuit->num_challenges < 1 || circuit->degree_bits < 0 || circuit->degree_bits + qdb > 24)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "bad circuit shape");
    const size_t words = ((size_t)circuit->num_challenges << (circuit->degree_bits + qdb));
    u64 *dev = (u64 *)out;
    if (out_mem != BSX_DEVICE) TRY(bsx_alloc(ctx, &dev, words));
    TRY(quotient_values_device(ctx, circuit, constants_sigmas, wires, zs_partial_products, public_inputs_hash, betas,
                               gammas, alphas, dev));
    if (out_mem != BSX_DEVICE) {
        TRY(bsx_copy_out(ctx, (u64 *)out, dev, words, out_mem));
        bsx_free(ctx, dev);
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

extern "C" int32_t bsx_compute_quotient_polys(bsx_ctx *ctx, const bsx_circuit *circuit, bsx_batch *constants_sigmas,
                                              bsx_batch *wires, bsx_batch *zs_partial_products,
                                              const uint64_t *public_inputs_hash, const uint64_t *betas,
                                              const uint64_t *gammas, const uint64_t *alphas, int32_t cap_height,
                                              int32_t blinding, bsx_batch **out, bsx_timing *timing, float *eval_ms) {
    if (!ctx || !circuit || !betas || !gammas || !alphas || !out) return BSX_ERR_INVALID_ARG;
    *out = nullptr;
    BsxScope scope(ctx);
    const int qdb = log2_ceil(circuit->quotient_degree_factor < 1 ? 1 : circuit->quotient_degree_factor);
    if (circuit->num_challenges < 1 || circuit->degree_bits < 0 || circuit->degree_bits + qdb > 24)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "bad circuit shape");
    const size_t words = ((size_t)circuit->num_challenges << (circuit->degree_bits + qdb));
    u64 *dev;
    TRY(bsx_alloc(ctx, &dev, words));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (eval_ms) {
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, ctx->stream);
    }
    int rc = quotient_values_device(ctx, circuit, constants_sigmas, wires, zs_partial_products, public_inputs_hash, betas,
                                    gammas, alphas, dev);
    if (eval_ms) cudaEventRecord(e1, ctx->stream);
    // the tail of compute_quotient_polys (coset_ifft) and, in prove(), trim_to_len / chunks / from_coeffs
    if (rc == BSX_OK)
        rc = bsx_batch_from_quotient_values(ctx, (const uint64_t *)dev, BSX_DEVICE, circuit->degree_bits, qdb,
                                            circuit->quotient_degree_factor, circuit->num_challenges, circuit->rate_bits,
                                            cap_height, blinding, nullptr, out, timing);
    if (eval_ms) {
        *eval_ms = 0;
        if (rc == BSX_OK && cudaEventSynchronize(e1) == cudaSuccess) cudaEventElapsedTime(eval_ms, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    bsx_free(ctx, dev);
    return rc;
}
