/* plonky2_plonk_oracle.c -- CPU restatement of the prover steps either side of the commitments:
 *   plonk/prover.rs `compute_quotient_polys` with plonk/vanishing_poly.rs `eval_vanishing_poly_base_batch`
 *   (the vanishing polynomial on the quotient coset) for circuits built from NoopGate, ConstantGate,
 *   PublicInputGate, ArithmeticGate, PoseidonGate, ArithmeticExtensionGate, MulExtensionGate, BaseSumGate<2>,
 *   ExponentiationGate, ReducingGate, ReducingExtensionGate, RandomAccessGate, PoseidonMdsGate and
 *   CosetInterpolationGate
 *   (gates/{noop,constant,public_input,arithmetic_base,poseidon,arithmetic_extension,multiplication_extension,
 *   base_sum,exponentiation,reducing,reducing_extension,random_access,poseidon_mds,coset_interpolation}.rs), and
 *   util/serialization/mod.rs `write_proof_with_public_inputs` (`ProofWithPublicInputs::to_bytes`).
 * TEST INFRASTRUCTURE ONLY; "parity unpinned" against the real prover -- see plonky2_oracle.h. */
#include "plonky2_oracle.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define P ORC_P

/* plonk/plonk_common.rs ZeroPolyOnCoset::new(n_log, rate_bits): evals[i] = g^n * v^i - 1, v a primitive
 * rate-th root of unity; eval(i) = evals[i % rate] */
typedef struct {
    uint64_t n;
    size_t rate;
    uint64_t *evals, *inverses;
} zero_poly_on_coset;
static void zpoc_new(zero_poly_on_coset *z, int n_log, int rate_bits) {
    z->n = orc_pow(2, (uint64_t)n_log) % P; /* F::from_canonical_usize(1 << n_log) */
    z->rate = (size_t)1 << rate_bits;
    z->evals = (uint64_t *)malloc(sizeof(uint64_t) * z->rate);
    z->inverses = (uint64_t *)malloc(sizeof(uint64_t) * z->rate);
    uint64_t g_pow_n = 7; /* F::coset_shift().exp_power_of_2(n_log) */
    for (int i = 0; i < n_log; i++) g_pow_n = orc_mul(g_pow_n, g_pow_n);
    uint64_t v = orc_root_of_unity(rate_bits), x = 1;
    for (size_t i = 0; i < z->rate; i++) {
        z->evals[i] = orc_sub(orc_mul(g_pow_n, x), 1);
        z->inverses[i] = orc_inv(z->evals[i]);
        x = orc_mul(x, v);
    }
}
static void zpoc_free(zero_poly_on_coset *z) {
    free(z->evals);
    free(z->inverses);
}
/* eval_l_0(index, x) = eval(index) * (n * (x - 1)).inverse() */
static uint64_t zpoc_eval_l_0(const zero_poly_on_coset *z, size_t index, uint64_t x) {
    return orc_mul(z->evals[index % z->rate], orc_inv(orc_mul(z->n, orc_sub(x, 1))));
}

/* fri/oracle.rs get_lde_values(index, step): leaf reverse_bits(index * step, lde_bits), salt stripped */
static const uint64_t *lde_row(const orc_batch_view *b, size_t index, size_t step) {
    int bits = b->log_n + b->rate_bits;
    return b->leaves + orc_reverse_bits(index * step, bits) * (size_t)(b->n_cols + b->n_salt);
}

/* gates/gate.rs compute_filter(row, group_range, s, many_selector) */
static uint64_t compute_filter(int row, int g0, int g1, uint64_t s, int many_selector) {
    uint64_t f = 1;
    for (int i = g0; i < g1; i++)
        if (i != row) f = orc_mul(f, orc_sub((uint64_t)i, s));
    if (many_selector) f = orc_mul(f, orc_sub(0xFFFFFFFFULL /* UNUSED_SELECTOR = u32::MAX */, s));
    return f;
}

/* gates/poseidon.rs PoseidonGate::eval_unfiltered_base_one.  Wires: input i, output 12 + i, swap 24, delta 25 + i (i < 4),
 * S-box inputs of the first full rounds 29 + 12 (r - 1) + i (r = 1..3), of the partial rounds 65 + r, of the second full
 * rounds 87 + 12 r + i.  Constraints in upstream's order: swap binary, the four deltas, then for every round whose
 * S-box inputs are wires `state - wire` (the state continues from the wire), and the outputs: 1 + 4 + 36 + 22 + 48 + 12.
 * The partial rounds are written in the naive form (constants, S-box on lane 0, MDS); upstream's factored form
 * (mds_partial_layer_init / mds_partial_layer_fast) gives the same lane-0 values for ANY wire values, satisfying or not,
 * because it is the same affine maps regrouped (checked numerically against tools/gen_poseidon_constants.py's
 * factorisation in tests/test_gates.py). */
#define PG_SWAP 24
#define PG_DELTA 25
#define PG_FULL0 29
#define PG_PARTIAL 65
#define PG_FULL1 87
static int poseidon_gate_eval(const uint64_t *w, uint64_t *out) {
    const uint64_t *rc = orc_round_constants();
    int t = 0;
    const uint64_t swap = w[PG_SWAP];
    out[t++] = orc_mul(swap, orc_sub(swap, 1));
    for (int i = 0; i < 4; i++) out[t++] = orc_sub(orc_mul(swap, orc_sub(w[i + 4], w[i])), w[PG_DELTA + i]);
    uint64_t s[12];
    for (int i = 0; i < 4; i++) {
        s[i] = orc_add(w[i], w[PG_DELTA + i]);
        s[i + 4] = orc_sub(w[i + 4], w[PG_DELTA + i]);
    }
    for (int i = 8; i < 12; i++) s[i] = w[i];
    for (int r = 0; r < 30; r++) {
        for (int i = 0; i < 12; i++) s[i] = orc_add(s[i], rc[12 * r + i]);
        const int full = r < 4 || r >= 26;
        if (r >= 1) {
            const uint64_t *win = r < 4 ? w + PG_FULL0 + 12 * (r - 1) : r < 26 ? w + PG_PARTIAL + (r - 4) : w + PG_FULL1 + 12 * (r - 26);
            for (int i = 0; i < (full ? 12 : 1); i++) {
                out[t++] = orc_sub(s[i], win[i]);
                s[i] = win[i];
            }
        }
        for (int i = 0; i < (full ? 12 : 1); i++) s[i] = orc_poseidon_sbox(s[i]);
        orc_poseidon_mds(s);
    }
    for (int i = 0; i < 12; i++) out[t++] = orc_sub(s[i], w[12 + i]);
    return t;
}

/* gates/poseidon.rs PoseidonGenerator::run_once for n rows of 135 wires (row-major): inputs 0..11 and swap (0 or 1) are
 * read, the deltas, every S-box input wire and the outputs are written. */
void orc_poseidon_gate_fill(uint64_t *rows, size_t n) {
    const uint64_t *rc = orc_round_constants();
    for (size_t row = 0; row < n; row++) {
        uint64_t *w = rows + 135 * row, s[12];
        const uint64_t swap = w[PG_SWAP] % P;
        for (int i = 0; i < 4; i++) {
            const uint64_t d = orc_mul(swap, orc_sub(w[i + 4] % P, w[i] % P));
            w[PG_DELTA + i] = d;
            s[i] = orc_add(w[i] % P, d);
            s[i + 4] = orc_sub(w[i + 4] % P, d);
        }
        for (int i = 8; i < 12; i++) s[i] = w[i] % P;
        for (int r = 0; r < 30; r++) {
            for (int i = 0; i < 12; i++) s[i] = orc_add(s[i], rc[12 * r + i]);
            const int full = r < 4 || r >= 26;
            if (r >= 1) {
                uint64_t *win = r < 4 ? w + PG_FULL0 + 12 * (r - 1) : r < 26 ? w + PG_PARTIAL + (r - 4) : w + PG_FULL1 + 12 * (r - 26);
                for (int i = 0; i < (full ? 12 : 1); i++) win[i] = s[i];
            }
            for (int i = 0; i < (full ? 12 : 1); i++) s[i] = orc_poseidon_sbox(s[i]);
            orc_poseidon_mds(s);
        }
        for (int i = 0; i < 12; i++) w[12 + i] = s[i];
    }
}

/* F_p[X] / (X^2 - 7) on base-field pairs: vars.get_local_ext(range) in the base-field evaluation */
static void x2_mul(const uint64_t a[2], const uint64_t b[2], uint64_t o[2]) { orc_ext_mul(a, b, o); }

/* gates/coset_interpolation.rs: the subgroup of order 2^bits and field/interpolation.rs barycentric_weights on it,
 * w_i = 1 / prod_{j != i} (x_i - x_j) */
static void ci_tables(int bits, uint64_t *domain, uint64_t *weights) {
    const int n = 1 << bits;
    const uint64_t g = orc_root_of_unity(bits);
    uint64_t x = 1;
    for (int i = 0; i < n; i++, x = orc_mul(x, g)) domain[i] = x;
    for (int i = 0; i < n; i++) {
        uint64_t pr = 1;
        for (int j = 0; j < n; j++)
            if (j != i) pr = orc_mul(pr, orc_sub(domain[i], domain[j]));
        weights[i] = orc_inv(pr);
    }
}
/* gates/coset_interpolation.rs partial_interpolate over F_p[X]/(X^2 - 7): for the points [first, last) fold
 * (eval, prod) <- (eval (x - x_i) + value_i w_i prod, prod (x - x_i)); value i sits at wires 1 + 2 i */
static void ci_partial(const uint64_t *domain, const uint64_t *weights, const uint64_t *wires, int first, int last,
                       const uint64_t x[2], uint64_t eval[2], uint64_t prod[2]) {
    for (int i = first; i < last; i++) {
        const uint64_t val[2] = {orc_mul(wires[1 + 2 * i], weights[i]), orc_mul(wires[2 + 2 * i], weights[i])};
        const uint64_t term[2] = {orc_sub(x[0], domain[i]), x[1]};
        uint64_t a[2], b[2];
        x2_mul(eval, term, a);
        x2_mul(val, prod, b);
        eval[0] = orc_add(a[0], b[0]);
        eval[1] = orc_add(a[1], b[1]);
        x2_mul(prod, term, a);
        prod[0] = a[0];
        prod[1] = a[1];
    }
}

/* eval_unfiltered of the gates; out[j] for j < num_constraints(gate); returns the constraint count.
 * consts = local_constants with the selector prefix removed (vars.remove_prefix(num_selectors)). */
static int gate_eval_unfiltered(const orc_gate *g, const uint64_t *consts, const uint64_t *wires,
                                const uint64_t pi_hash[4], uint64_t *out) {
    switch (g->kind) {
        case ORC_GATE_NOOP:
            return 0;
        case ORC_GATE_CONSTANT: /* constants[const_input(i)] - wires[wire_output(i)], both = i */
            for (int i = 0; i < g->param; i++) out[i] = orc_sub(consts[i], wires[i]);
            return g->param;
        case ORC_GATE_PUBLIC_INPUT: /* wires[i] - public_inputs_hash.elements[i] */
            for (int i = 0; i < 4; i++) out[i] = orc_sub(wires[i], pi_hash[i]);
            return 4;
        case ORC_GATE_ARITHMETIC: /* output - (m0 * m1 * const_0 + addend * const_1), wires 4i .. 4i+3 */
            for (int i = 0; i < g->param; i++) {
                uint64_t m0 = wires[4 * i], m1 = wires[4 * i + 1], ad = wires[4 * i + 2], o = wires[4 * i + 3];
                uint64_t computed = orc_add(orc_mul(orc_mul(m0, m1), consts[0]), orc_mul(ad, consts[1]));
                out[i] = orc_sub(o, computed);
            }
            return g->param;
        case ORC_GATE_POSEIDON:
            return poseidon_gate_eval(wires, out);
        case ORC_GATE_ARITHMETIC_EXTENSION: /* output - ((m0 * m1).scalar_mul(const_0) + addend.scalar_mul(const_1)), D = 2 */
            for (int i = 0; i < g->param; i++) {
                const uint64_t *q = wires + 8 * i;
                uint64_t pr[2];
                x2_mul(q, q + 2, pr);
                for (int d = 0; d < 2; d++)
                    out[2 * i + d] = orc_sub(q[6 + d], orc_add(orc_mul(pr[d], consts[0]), orc_mul(q[4 + d], consts[1])));
            }
            return 2 * g->param;
        case ORC_GATE_MUL_EXTENSION: /* output - (m0 * m1).scalar_mul(const_0) */
            for (int i = 0; i < g->param; i++) {
                const uint64_t *q = wires + 6 * i;
                uint64_t pr[2];
                x2_mul(q, q + 2, pr);
                for (int d = 0; d < 2; d++) out[2 * i + d] = orc_sub(q[4 + d], orc_mul(pr[d], consts[0]));
            }
            return 2 * g->param;
        case ORC_GATE_EXPONENTIATION: { /* wires: base 0, power bits 1.. (little-endian), output 1 + n, intermediate values 2 + n.. */
            const int n = g->param;
            const uint64_t *iv = wires + 2 + n;
            for (int i = 0; i < n; i++) {
                const uint64_t prev = i == 0 ? 1 : orc_mul(iv[i - 1], iv[i - 1]);
                const uint64_t bit = wires[1 + (n - 1 - i)]; /* accumulated in big-endian order */
                out[i] = orc_sub(orc_mul(prev, orc_add(orc_mul(bit, wires[0]), orc_sub(1, bit))), iv[i]);
            }
            out[n] = orc_sub(wires[1 + n], iv[n - 1]);
            return n + 1;
        }
        case ORC_GATE_REDUCING:
        case ORC_GATE_REDUCING_EXTENSION: { /* wires: output 0..2, alpha 2..4, old_acc 4..6, coeffs from 6, then the accumulators
                                             * (the last one is the output): acc alpha + coeff_i - acc_i */
            const int n = g->param, cw = g->kind == ORC_GATE_REDUCING_EXTENSION ? 2 : 1, start_accs = 6 + n * cw;
            const uint64_t *acc = wires + 4;
            for (int i = 0; i < n; i++) {
                const uint64_t *acc_i = i == n - 1 ? wires : wires + start_accs + 2 * i;
                uint64_t t[2];
                x2_mul(acc, wires + 2, t);
                t[0] = orc_add(t[0], wires[6 + cw * i]);
                if (cw == 2) t[1] = orc_add(t[1], wires[6 + 2 * i + 1]);
                out[2 * i] = orc_sub(t[0], acc_i[0]);
                out[2 * i + 1] = orc_sub(t[1], acc_i[1]);
                acc = acc_i;
            }
            return 2 * n;
        }
        case ORC_GATE_RANDOM_ACCESS: { /* per copy: access_index, claimed_element, 2^bits list items; bits after the routed wires */
            const int bits = g->param & 0xFF, copies = (g->param >> 8) & 0xFF, extra = g->param >> 16, vec = 1 << bits;
            const int routed = (2 + vec) * copies + extra;
            int t = 0;
            uint64_t list[256];
            for (int c = 0; c < copies; c++) {
                const uint64_t *q = wires + (2 + vec) * c, *b = wires + routed + c * bits;
                for (int i = 0; i < bits; i++) out[t++] = orc_mul(b[i], orc_sub(b[i], 1));
                uint64_t rec = 0;
                for (int i = bits - 1; i >= 0; i--) rec = orc_add(orc_add(rec, rec), b[i]);
                out[t++] = orc_sub(rec, q[0]);
                for (int j = 0; j < vec; j++) list[j] = q[2 + j];
                for (int i = 0, len = vec; i < bits; i++, len >>= 1)
                    for (int j = 0; j < len / 2; j++)
                        list[j] = orc_add(list[2 * j], orc_mul(b[i], orc_sub(list[2 * j + 1], list[2 * j])));
                out[t++] = orc_sub(list[0], q[1]);
            }
            for (int i = 0; i < extra; i++) out[t++] = orc_sub(consts[i], wires[(2 + vec) * copies + i]);
            return t;
        }
        case ORC_GATE_COSET_INTERPOLATION: {
            /* wires: shift 0, values 1 + 2 i, evaluation point, evaluation value, then intermediate evals, intermediate
             * prods and the shifted evaluation point.  evaluation_point - shifted_point * shift; the barycentric sum in
             * chunks (the first of `degree` points, then degree - 1 each) with the running (eval, prod) pinned to wires
             * after every chunk; evaluation_value - the last eval */
            const int bits = g->param & 0xFF, degree = g->param >> 8, np = 1 << bits, ni = (np - 2) / (degree - 1);
            const int ep = 1 + 2 * np, ev = ep + 2, im = ev + 2, sp = im + 4 * ni;
            uint64_t domain[256], weights[256];
            ci_tables(bits, domain, weights);
            int t = 0;
            out[t++] = orc_sub(wires[ep], orc_mul(wires[sp], wires[0]));
            out[t++] = orc_sub(wires[ep + 1], orc_mul(wires[sp + 1], wires[0]));
            uint64_t eval[2] = {0, 0}, prod[2] = {1, 0};
            ci_partial(domain, weights, wires, 0, degree, wires + sp, eval, prod);
            for (int i = 0; i < ni; i++) {
                const uint64_t *ie = wires + im + 2 * i, *ip = wires + im + 2 * (ni + i);
                out[t++] = orc_sub(ie[0], eval[0]);
                out[t++] = orc_sub(ie[1], eval[1]);
                out[t++] = orc_sub(ip[0], prod[0]);
                out[t++] = orc_sub(ip[1], prod[1]);
                const int start = 1 + (degree - 1) * (i + 1), end = start + degree - 1 < np ? start + degree - 1 : np;
                eval[0] = ie[0], eval[1] = ie[1], prod[0] = ip[0], prod[1] = ip[1];
                ci_partial(domain, weights, wires, start, end, wires + sp, eval, prod);
            }
            out[t++] = orc_sub(wires[ev], eval[0]);
            out[t++] = orc_sub(wires[ev + 1], eval[1]);
            return t;
        }
        case ORC_GATE_POSEIDON_MDS: { /* output_r - (sum_i CIRC[i] input_{(i + r) % 12} + DIAG[r] input_r), component by component */
            for (int d = 0; d < 2; d++) {
                uint64_t lane[12];
                for (int i = 0; i < 12; i++) lane[i] = wires[2 * i + d];
                orc_poseidon_mds(lane);
                for (int r = 0; r < 12; r++) out[2 * r + d] = orc_sub(wires[24 + 2 * r + d], lane[r]);
            }
            return 24;
        }
        case ORC_GATE_BASE_SUM: { /* reduce_with_powers(limbs, 2) - sum, then limb (limb - 1) per limb; sum = wire 0 */
            uint64_t acc = 0;
            for (int i = g->param; i >= 1; i--) acc = orc_add(orc_mul(acc, 2), wires[i]);
            out[0] = orc_sub(acc, wires[0]);
            for (int i = 1; i <= g->param; i++) out[i] = orc_mul(wires[i], orc_sub(wires[i], 1));
            return 1 + g->param;
        }
        default:
            return -1;
    }
}

int orc_gate_num_constraints(const orc_gate *g) {
    switch (g->kind) {
        case ORC_GATE_NOOP: return 0;
        case ORC_GATE_CONSTANT: return g->param;
        case ORC_GATE_PUBLIC_INPUT: return 4;
        case ORC_GATE_ARITHMETIC: return g->param;
        case ORC_GATE_POSEIDON: return 123;
        case ORC_GATE_ARITHMETIC_EXTENSION: return 2 * g->param;
        case ORC_GATE_MUL_EXTENSION: return 2 * g->param;
        case ORC_GATE_BASE_SUM: return 1 + g->param;
        case ORC_GATE_EXPONENTIATION: return g->param < 1 ? -1 : g->param + 1;
        case ORC_GATE_REDUCING: return g->param < 1 ? -1 : 2 * g->param;
        case ORC_GATE_REDUCING_EXTENSION: return g->param < 1 ? -1 : 2 * g->param;
        case ORC_GATE_RANDOM_ACCESS: {
            const int bits = g->param & 0xFF, copies = (g->param >> 8) & 0xFF, extra = g->param >> 16;
            return bits < 1 || bits > 8 || copies < 1 ? -1 : copies * (bits + 2) + extra;
        }
        case ORC_GATE_POSEIDON_MDS: return 24;
        case ORC_GATE_COSET_INTERPOLATION: {
            const int bits = g->param & 0xFF, degree = g->param >> 8;
            return bits < 1 || bits > 8 || degree < 2 || degree > (1 << bits) ? -1 : 2 * (2 + 2 * (((1 << bits) - 2) / (degree - 1)));
        }
        default: return -1;
    }
}

/* plonk/vanishing_poly.rs eval_vanishing_poly_base_batch for ONE point: the vanishing terms in upstream's order
 * (vanishing_z_1_terms, vanishing_partial_products_terms, gate constraint terms) reduced with the powers of every
 * alpha (plonk_common.rs reduce_with_powers_multi).  l_0_x = L_0 at the point.  terms / gate_out: scratch of
 * n_terms + 1 and num_gate_constraints + 4 words.  Returns -1 on a gate that does not fit the description. */
static int vanishing_at_point(const orc_circuit *c, uint64_t x, uint64_t l_0_x, const uint64_t *local_constants,
                              const uint64_t *s_sigmas, const uint64_t *local_wires, const uint64_t *local_zs_pp,
                              const uint64_t *next_zs, const uint64_t pi_hash[4], const uint64_t *betas,
                              const uint64_t *gammas, const uint64_t *alphas, uint64_t *terms, uint64_t *gate_out,
                              uint64_t *res) {
    const int nc = c->num_challenges, max_degree = c->quotient_degree_factor, num_prods = c->num_partial_products;
    const int n_chunks = num_prods + 1;
    const uint64_t *partial_products = local_zs_pp + nc; /* partial_products_range */
    int t = 0, bad = 0;
    /* vanishing_z_1_terms */
    for (int k = 0; k < nc; k++) terms[t++] = orc_mul(l_0_x, orc_sub(local_zs_pp[k], 1));
    /* vanishing_partial_products_terms: check_partial_products per challenge */
    for (int k = 0; k < nc; k++) {
        const uint64_t z_x = local_zs_pp[k], z_gx = next_zs[k];
        const uint64_t *pp = partial_products + (size_t)k * num_prods;
        for (int m = 0; m < n_chunks; m++) {
            uint64_t num = 1, den = 1;
            for (int j = m * max_degree; j < c->num_routed_wires && j < (m + 1) * max_degree; j++) {
                const uint64_t wv = local_wires[j], s_id = orc_mul(c->k_is[j] % P, x);
                num = orc_mul(num, orc_add(orc_add(wv, orc_mul(betas[k], s_id)), gammas[k]));
                den = orc_mul(den, orc_add(orc_add(wv, orc_mul(betas[k], s_sigmas[j])), gammas[k]));
            }
            const uint64_t prev_acc = m == 0 ? z_x : pp[m - 1], next_acc = m == n_chunks - 1 ? z_gx : pp[m];
            terms[t++] = orc_sub(orc_mul(prev_acc, num), orc_mul(next_acc, den));
        }
    }
    /* evaluate_gate_constraints_base_batch: filtered constraints of every gate summed per slot */
    uint64_t *cons = terms + t;
    for (int j = 0; j < c->num_gate_constraints; j++) cons[j] = 0;
    for (int gi = 0; gi < c->n_gates; gi++) {
        const orc_gate *g = &c->gates[gi];
        const uint64_t filter = compute_filter(g->row, g->group_start, g->group_end, local_constants[g->selector_index],
                                               c->num_selectors > 1);
        int k = orc_gate_num_constraints(g); /* checked before the evaluation writes gate_out */
        if (k < 0 || k > c->num_gate_constraints) { bad = 1; continue; }
        k = gate_eval_unfiltered(g, local_constants + c->num_selectors, local_wires, pi_hash, gate_out);
        for (int j = 0; j < k; j++) cons[j] = orc_add(cons[j], orc_mul(filter, gate_out[j]));
    }
    t += c->num_gate_constraints;
    for (int a = 0; a < nc; a++) { /* reduce_with_powers_multi */
        uint64_t acc = 0;
        for (int j = t - 1; j >= 0; j--) acc = orc_add(terms[j], orc_mul(acc, alphas[a]));
        res[a] = acc;
    }
    return bad ? -1 : 0;
}

static int circuit_terms(const orc_circuit *c) {
    return c->num_challenges * (2 + c->num_partial_products) + c->num_gate_constraints;
}

/* The vanishing polynomial at an arbitrary base-field point x from the polynomials' values there (what the verifier
 * does with the opening set, plonk/vanishing_poly.rs eval_vanishing_poly): Z_H(x) = x^n - 1,
 * L_0(x) = (x^n - 1) / (n (x - 1)).  rows: constants then sigmas | wires | zs then partial products | zs at g x. */
int orc_vanishing_at(const orc_circuit *c, uint64_t x, const uint64_t *constants_sigmas_row, const uint64_t *wires_row,
                     const uint64_t *zs_pp_row, const uint64_t *next_zs_row, const uint64_t pi_hash[4],
                     const uint64_t *betas, const uint64_t *gammas, const uint64_t *alphas, uint64_t *res) {
    uint64_t xn = x % P;
    for (int i = 0; i < c->degree_bits; i++) xn = orc_mul(xn, xn);
    const uint64_t z_h = orc_sub(xn, 1), n = orc_pow(2, (uint64_t)c->degree_bits);
    const uint64_t l_0_x = orc_mul(z_h, orc_inv(orc_mul(n, orc_sub(x % P, 1))));
    uint64_t *terms = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(circuit_terms(c) + 1));
    uint64_t *gate_out = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(c->num_gate_constraints + 4));
    int rc = vanishing_at_point(c, x % P, l_0_x, constants_sigmas_row, constants_sigmas_row + c->num_constants, wires_row,
                                zs_pp_row, next_zs_row, pi_hash, betas, gammas, alphas, terms, gate_out, res);
    free(terms);
    free(gate_out);
    return rc;
}

/* ------------------------------------------------------------ verifier side, extension field
 * plonk/vanishing_poly.rs eval_vanishing_poly at an EXTENSION point from the opening set, as plonk/verifier.rs
 * verify_with_challenges uses it.  Written separately from the base-field prover path above (different arithmetic,
 * no shared helper beyond the gate table), so that it checks that path rather than repeats it. */
typedef struct { uint64_t c[2]; } e2;
static e2 e2_of(uint64_t x) { e2 r = {{x % P, 0}}; return r; }
static e2 e2_add(e2 a, e2 b) { e2 r = {{orc_add(a.c[0], b.c[0]), orc_add(a.c[1], b.c[1])}}; return r; }
static e2 e2_sub(e2 a, e2 b) { e2 r = {{orc_sub(a.c[0], b.c[0]), orc_sub(a.c[1], b.c[1])}}; return r; }
static e2 e2_mul(e2 a, e2 b) { e2 r; orc_ext_mul(a.c, b.c, r.c); return r; }
static e2 e2_inv(e2 a) { e2 r; orc_ext_inv(a.c, r.c); return r; }
static e2 e2_scale(e2 a, uint64_t k) { e2 r = {{orc_mul(a.c[0], k % P), orc_mul(a.c[1], k % P)}}; return r; }

/* openings: constants | plonk_sigmas | wires | plonk_zs | plonk_zs_next | partial_products, each [n][2].
 * res[num_challenges][2] = reduce_with_powers_multi(vanishing terms, alphas) at zeta. */
int orc_vanishing_at_ext(const orc_circuit *c, const uint64_t zeta[2], const uint64_t *constants, const uint64_t *sigmas,
                         const uint64_t *wires, const uint64_t *zs, const uint64_t *zs_next, const uint64_t *partial_products,
                         const uint64_t pi_hash[4], const uint64_t *betas, const uint64_t *gammas, const uint64_t *alphas,
                         uint64_t *res) {
    const int nc = c->num_challenges, max_degree = c->quotient_degree_factor, num_prods = c->num_partial_products;
    const int n_chunks = num_prods + 1, n_terms = nc * (1 + n_chunks) + c->num_gate_constraints;
    if (n_chunks != (c->num_routed_wires + max_degree - 1) / max_degree) return -1;
    e2 *terms = (e2 *)malloc(sizeof(e2) * (size_t)(n_terms + 1));
    const e2 x = {{zeta[0] % P, zeta[1] % P}}, one = e2_of(1);
#define EXT_AT(arr, i) (*(const e2 *)((arr) + 2 * (size_t)(i)))
    /* plonk_common.rs eval_l_0(n, x) = (x^n - 1) / (n (x - 1)) */
    e2 xn = x;
    for (int i = 0; i < c->degree_bits; i++) xn = e2_mul(xn, xn);
    const e2 l_0_x = e2_mul(e2_sub(xn, one), e2_inv(e2_scale(e2_sub(x, one), orc_pow(2, (uint64_t)c->degree_bits))));
    int t = 0;
    for (int k = 0; k < nc; k++) terms[t++] = e2_mul(l_0_x, e2_sub(EXT_AT(zs, k), one)); /* vanishing_z_1_terms */
    for (int k = 0; k < nc; k++) { /* check_partial_products */
        for (int m = 0; m < n_chunks; m++) {
            e2 num = one, den = one;
            for (int j = m * max_degree; j < c->num_routed_wires && j < (m + 1) * max_degree; j++) {
                const e2 w = EXT_AT(wires, j), s_id = e2_scale(x, c->k_is[j]);
                num = e2_mul(num, e2_add(e2_add(w, e2_scale(s_id, betas[k])), e2_of(gammas[k])));
                den = e2_mul(den, e2_add(e2_add(w, e2_scale(EXT_AT(sigmas, j), betas[k])), e2_of(gammas[k])));
            }
            const e2 prev = m == 0 ? EXT_AT(zs, k) : EXT_AT(partial_products, k * num_prods + m - 1);
            const e2 next = m == n_chunks - 1 ? EXT_AT(zs_next, k) : EXT_AT(partial_products, k * num_prods + m);
            terms[t++] = e2_sub(e2_mul(prev, num), e2_mul(next, den));
        }
    }
    /* evaluate_gate_constraints: every gate's filtered constraints summed per slot */
    for (int j = 0; j < c->num_gate_constraints; j++) terms[t + j] = e2_of(0);
    int bad = 0;
    for (int gi = 0; gi < c->n_gates; gi++) {
        const orc_gate *g = &c->gates[gi];
        const e2 s = EXT_AT(constants, g->selector_index);
        e2 f = one; /* compute_filter */
        for (int i = g->group_start; i < g->group_end; i++)
            if (i != g->row) f = e2_mul(f, e2_sub(e2_of((uint64_t)i), s));
        if (c->num_selectors > 1) f = e2_mul(f, e2_sub(e2_of(0xFFFFFFFFULL), s));
        const uint64_t *gc = constants + 2 * (size_t)c->num_selectors; /* constants after the selector prefix */
        if (orc_gate_num_constraints(g) < 0 || orc_gate_num_constraints(g) > c->num_gate_constraints) { bad = 1; continue; }
        switch (g->kind) {
            case ORC_GATE_NOOP: break;
            case ORC_GATE_CONSTANT:
                for (int i = 0; i < g->param; i++) terms[t + i] = e2_add(terms[t + i], e2_mul(f, e2_sub(EXT_AT(gc, i), EXT_AT(wires, i))));
                break;
            case ORC_GATE_PUBLIC_INPUT:
                for (int i = 0; i < 4; i++) terms[t + i] = e2_add(terms[t + i], e2_mul(f, e2_sub(EXT_AT(wires, i), e2_of(pi_hash[i]))));
                break;
            case ORC_GATE_ARITHMETIC:
                for (int i = 0; i < g->param; i++) {
                    const e2 prod = e2_mul(e2_mul(EXT_AT(wires, 4 * i), EXT_AT(wires, 4 * i + 1)), EXT_AT(gc, 0));
                    const e2 computed = e2_add(prod, e2_mul(EXT_AT(wires, 4 * i + 2), EXT_AT(gc, 1)));
                    terms[t + i] = e2_add(terms[t + i], e2_mul(f, e2_sub(EXT_AT(wires, 4 * i + 3), computed)));
                }
                break;
            case ORC_GATE_POSEIDON: { /* PoseidonGate::eval_unfiltered: the same schedule over F_p^2 (constants and MDS are base-field) */
                const uint64_t *rc = orc_round_constants();
                int k = 0;
#define PG_EMIT(v) do { terms[t + k] = e2_add(terms[t + k], e2_mul(f, (v))); k++; } while (0)
                const e2 swap = EXT_AT(wires, PG_SWAP);
                PG_EMIT(e2_mul(swap, e2_sub(swap, one)));
                for (int i = 0; i < 4; i++)
                    PG_EMIT(e2_sub(e2_mul(swap, e2_sub(EXT_AT(wires, i + 4), EXT_AT(wires, i))), EXT_AT(wires, PG_DELTA + i)));
                e2 st[12];
                for (int i = 0; i < 4; i++) {
                    st[i] = e2_add(EXT_AT(wires, i), EXT_AT(wires, PG_DELTA + i));
                    st[i + 4] = e2_sub(EXT_AT(wires, i + 4), EXT_AT(wires, PG_DELTA + i));
                }
                for (int i = 8; i < 12; i++) st[i] = EXT_AT(wires, i);
                for (int r = 0; r < 30; r++) {
                    for (int i = 0; i < 12; i++) st[i] = e2_add(st[i], e2_of(rc[12 * r + i]));
                    const int full = r < 4 || r >= 26;
                    if (r >= 1) {
                        const int w0 = r < 4 ? PG_FULL0 + 12 * (r - 1) : r < 26 ? PG_PARTIAL + (r - 4) : PG_FULL1 + 12 * (r - 26);
                        for (int i = 0; i < (full ? 12 : 1); i++) {
                            PG_EMIT(e2_sub(st[i], EXT_AT(wires, w0 + i)));
                            st[i] = EXT_AT(wires, w0 + i);
                        }
                    }
                    for (int i = 0; i < (full ? 12 : 1); i++) { /* x^7 */
                        const e2 x2 = e2_mul(st[i], st[i]), x4 = e2_mul(x2, x2), x3 = e2_mul(x2, st[i]);
                        st[i] = e2_mul(x3, x4);
                    }
                    uint64_t lane[2][12]; /* the MDS matrix is over the base field: component by component */
                    for (int i = 0; i < 12; i++) lane[0][i] = st[i].c[0], lane[1][i] = st[i].c[1];
                    orc_poseidon_mds(lane[0]);
                    orc_poseidon_mds(lane[1]);
                    for (int i = 0; i < 12; i++) st[i].c[0] = lane[0][i], st[i].c[1] = lane[1][i];
                }
                for (int i = 0; i < 12; i++) PG_EMIT(e2_sub(st[i], EXT_AT(wires, 12 + i)));
#undef PG_EMIT
                break;
            }
            case ORC_GATE_ARITHMETIC_EXTENSION:
            case ORC_GATE_MUL_EXTENSION: {
                /* vars.get_local_ext_algebra: pairs of F_p^2 openings as elements of F_p^2[X] / (X^2 - 7)
                 * (ExtensionAlgebra): (a0 + a1 X)(b0 + b1 X) = a0 b0 + 7 a1 b1 + (a0 b1 + a1 b0) X */
                const int arith = g->kind == ORC_GATE_ARITHMETIC_EXTENSION, per = arith ? 8 : 6;
                for (int i = 0; i < g->param; i++) {
                    const int q = per * i;
                    const e2 a0 = EXT_AT(wires, q), a1 = EXT_AT(wires, q + 1), b0 = EXT_AT(wires, q + 2), b1 = EXT_AT(wires, q + 3);
                    e2 pr[2];
                    pr[0] = e2_add(e2_mul(a0, b0), e2_scale(e2_mul(a1, b1), 7));
                    pr[1] = e2_add(e2_mul(a0, b1), e2_mul(a1, b0));
                    for (int d = 0; d < 2; d++) {
                        e2 computed = e2_mul(pr[d], EXT_AT(gc, 0));
                        if (arith) computed = e2_add(computed, e2_mul(EXT_AT(wires, q + 4 + d), EXT_AT(gc, 1)));
                        const e2 o = EXT_AT(wires, q + (arith ? 6 : 4) + d);
                        terms[t + 2 * i + d] = e2_add(terms[t + 2 * i + d], e2_mul(f, e2_sub(o, computed)));
                    }
                }
                break;
            }
            case ORC_GATE_EXPONENTIATION: {
                const int n = g->param;
                for (int i = 0; i < n; i++) {
                    const e2 prev = i == 0 ? one : e2_mul(EXT_AT(wires, 2 + n + i - 1), EXT_AT(wires, 2 + n + i - 1));
                    const e2 bit = EXT_AT(wires, 1 + (n - 1 - i));
                    const e2 computed = e2_mul(prev, e2_add(e2_mul(bit, EXT_AT(wires, 0)), e2_sub(one, bit)));
                    terms[t + i] = e2_add(terms[t + i], e2_mul(f, e2_sub(computed, EXT_AT(wires, 2 + n + i))));
                }
                terms[t + n] = e2_add(terms[t + n], e2_mul(f, e2_sub(EXT_AT(wires, 1 + n), EXT_AT(wires, 2 + n + n - 1))));
                break;
            }
            case ORC_GATE_REDUCING:
            case ORC_GATE_REDUCING_EXTENSION: { /* ExtensionAlgebra: acc * alpha + coeff - acc_i, two F_p^2 constraints each */
                const int n = g->param, cw = g->kind == ORC_GATE_REDUCING_EXTENSION ? 2 : 1, start_accs = 6 + n * cw;
                const e2 al0 = EXT_AT(wires, 2), al1 = EXT_AT(wires, 3);
                int acc = 4;
                for (int i = 0; i < n; i++) {
                    const int acc_i = i == n - 1 ? 0 : start_accs + 2 * i;
                    const e2 a0 = EXT_AT(wires, acc), a1 = EXT_AT(wires, acc + 1);
                    e2 t0 = e2_add(e2_mul(a0, al0), e2_scale(e2_mul(a1, al1), 7)), t1 = e2_add(e2_mul(a0, al1), e2_mul(a1, al0));
                    t0 = e2_add(t0, EXT_AT(wires, 6 + cw * i));
                    if (cw == 2) t1 = e2_add(t1, EXT_AT(wires, 6 + 2 * i + 1));
                    terms[t + 2 * i] = e2_add(terms[t + 2 * i], e2_mul(f, e2_sub(t0, EXT_AT(wires, acc_i))));
                    terms[t + 2 * i + 1] = e2_add(terms[t + 2 * i + 1], e2_mul(f, e2_sub(t1, EXT_AT(wires, acc_i + 1))));
                    acc = acc_i;
                }
                break;
            }
            case ORC_GATE_RANDOM_ACCESS: {
                const int bits = g->param & 0xFF, copies = (g->param >> 8) & 0xFF, extra = g->param >> 16, vec = 1 << bits;
                const int routed = (2 + vec) * copies + extra;
                int k = 0;
                e2 list[256];
                for (int cp = 0; cp < copies; cp++) {
                    const int q = (2 + vec) * cp, b = routed + cp * bits;
                    for (int i = 0; i < bits; i++, k++)
                        terms[t + k] = e2_add(terms[t + k], e2_mul(f, e2_mul(EXT_AT(wires, b + i), e2_sub(EXT_AT(wires, b + i), one))));
                    e2 rec = e2_of(0);
                    for (int i = bits - 1; i >= 0; i--) rec = e2_add(e2_add(rec, rec), EXT_AT(wires, b + i));
                    terms[t + k] = e2_add(terms[t + k], e2_mul(f, e2_sub(rec, EXT_AT(wires, q))));
                    k++;
                    for (int j = 0; j < vec; j++) list[j] = EXT_AT(wires, q + 2 + j);
                    for (int i = 0, len = vec; i < bits; i++, len >>= 1)
                        for (int j = 0; j < len / 2; j++)
                            list[j] = e2_add(list[2 * j], e2_mul(EXT_AT(wires, b + i), e2_sub(list[2 * j + 1], list[2 * j])));
                    terms[t + k] = e2_add(terms[t + k], e2_mul(f, e2_sub(list[0], EXT_AT(wires, q + 1))));
                    k++;
                }
                for (int i = 0; i < extra; i++, k++)
                    terms[t + k] = e2_add(terms[t + k], e2_mul(f, e2_sub(EXT_AT(gc, i), EXT_AT(wires, (2 + vec) * copies + i))));
                break;
            }
            case ORC_GATE_COSET_INTERPOLATION: { /* the same over the ExtensionAlgebra: elements (A0, A1) of F_p^2 pairs */
                const int bits = g->param & 0xFF, degree = g->param >> 8, np = 1 << bits, ni = (np - 2) / (degree - 1);
                const int ep = 1 + 2 * np, ev = ep + 2, im = ev + 2, sp = im + 4 * ni;
                uint64_t domain[256], weights[256];
                ci_tables(bits, domain, weights);
                int k = 0;
#define CI_EMIT(v) do { terms[t + k] = e2_add(terms[t + k], e2_mul(f, (v))); k++; } while (0)
                const e2 shift = EXT_AT(wires, 0), x0 = EXT_AT(wires, sp), x1 = EXT_AT(wires, sp + 1);
                CI_EMIT(e2_sub(EXT_AT(wires, ep), e2_mul(x0, shift)));
                CI_EMIT(e2_sub(EXT_AT(wires, ep + 1), e2_mul(x1, shift)));
                e2 ev0 = e2_of(0), ev1 = e2_of(0), pr0 = one, pr1 = e2_of(0);
                int first = 0, last = degree;
                for (int i = 0; i <= ni; i++) {
                    for (int q = first; q < last; q++) { /* partial_interpolate */
                        const e2 v0 = e2_scale(EXT_AT(wires, 1 + 2 * q), weights[q]), v1 = e2_scale(EXT_AT(wires, 2 + 2 * q), weights[q]);
                        const e2 t0 = e2_sub(x0, e2_of(domain[q])), t1 = x1;
                        /* (a0 + a1 X)(b0 + b1 X) = a0 b0 + 7 a1 b1 + (a0 b1 + a1 b0) X */
                        const e2 n0 = e2_add(e2_add(e2_mul(ev0, t0), e2_scale(e2_mul(ev1, t1), 7)), e2_add(e2_mul(v0, pr0), e2_scale(e2_mul(v1, pr1), 7)));
                        const e2 n1 = e2_add(e2_add(e2_mul(ev0, t1), e2_mul(ev1, t0)), e2_add(e2_mul(v0, pr1), e2_mul(v1, pr0)));
                        const e2 p0 = e2_add(e2_mul(pr0, t0), e2_scale(e2_mul(pr1, t1), 7)), p1 = e2_add(e2_mul(pr0, t1), e2_mul(pr1, t0));
                        ev0 = n0, ev1 = n1, pr0 = p0, pr1 = p1;
                    }
                    if (i == ni) break;
                    const int ie = im + 2 * i, ip = im + 2 * (ni + i);
                    CI_EMIT(e2_sub(EXT_AT(wires, ie), ev0));
                    CI_EMIT(e2_sub(EXT_AT(wires, ie + 1), ev1));
                    CI_EMIT(e2_sub(EXT_AT(wires, ip), pr0));
                    CI_EMIT(e2_sub(EXT_AT(wires, ip + 1), pr1));
                    ev0 = EXT_AT(wires, ie), ev1 = EXT_AT(wires, ie + 1), pr0 = EXT_AT(wires, ip), pr1 = EXT_AT(wires, ip + 1);
                    first = 1 + (degree - 1) * (i + 1);
                    last = first + degree - 1 < np ? first + degree - 1 : np;
                }
                CI_EMIT(e2_sub(EXT_AT(wires, ev), ev0));
                CI_EMIT(e2_sub(EXT_AT(wires, ev + 1), ev1));
#undef CI_EMIT
                break;
            }
            case ORC_GATE_POSEIDON_MDS: { /* mds_layer_algebra: the matrix is over the base field, so every one of the four
                                           * base-field coordinates of the 12 algebra elements goes through it on its own */
                for (int d = 0; d < 2; d++)
                    for (int cc = 0; cc < 2; cc++) {
                        uint64_t lane[12];
                        for (int i = 0; i < 12; i++) lane[i] = EXT_AT(wires, 2 * i + d).c[cc];
                        orc_poseidon_mds(lane);
                        for (int r = 0; r < 12; r++) {
                            e2 diff = e2_of(0);
                            diff.c[cc] = orc_sub(EXT_AT(wires, 24 + 2 * r + d).c[cc], lane[r]);
                            terms[t + 2 * r + d] = e2_add(terms[t + 2 * r + d], e2_mul(f, diff));
                        }
                    }
                break;
            }
            case ORC_GATE_BASE_SUM: {
                e2 acc = e2_of(0);
                for (int i = g->param; i >= 1; i--) acc = e2_add(e2_scale(acc, 2), EXT_AT(wires, i));
                terms[t] = e2_add(terms[t], e2_mul(f, e2_sub(acc, EXT_AT(wires, 0))));
                for (int i = 1; i <= g->param; i++)
                    terms[t + i] = e2_add(terms[t + i], e2_mul(f, e2_mul(EXT_AT(wires, i), e2_sub(EXT_AT(wires, i), one))));
                break;
            }
            default: bad = 1;
        }
    }
    t += c->num_gate_constraints;
    for (int a = 0; a < nc; a++) {
        e2 acc = e2_of(0);
        for (int j = t - 1; j >= 0; j--) acc = e2_add(terms[j], e2_scale(acc, alphas[a]));
        res[2 * a] = acc.c[0];
        res[2 * a + 1] = acc.c[1];
    }
#undef EXT_AT
    free(terms);
    return bad ? -1 : 0;
}

/* plonk/prover.rs compute_quotient_polys up to (not including) the final coset_ifft:
 * out[a * lde_size + i] = quotient value of challenge alpha_a at point i of the coset 7 * <omega_{n 2^q}>. */
int orc_quotient_values(const orc_circuit *c, const orc_batch_view *constants_sigmas, const orc_batch_view *wires_b,
                        const orc_batch_view *zs_b, const uint64_t public_inputs_hash[4], const uint64_t *betas,
                        const uint64_t *gammas, const uint64_t *alphas, uint64_t *out) {
    const int nc = c->num_challenges, max_degree = c->quotient_degree_factor, num_prods = c->num_partial_products;
    int qdb = 0;
    while ((1 << qdb) < c->quotient_degree_factor) qdb++; /* log2_ceil */
    if (qdb > c->rate_bits) return -1;                     /* upstream assert */
    if (max_degree < 2 || nc < 1) return -1;
    const int n_chunks = (c->num_routed_wires + max_degree - 1) / max_degree;
    if (num_prods != n_chunks - 1) return -1;
    if (constants_sigmas->n_cols != c->num_constants + c->num_routed_wires || wires_b->n_cols != c->num_wires ||
        zs_b->n_cols != nc * (1 + num_prods))
        return -1;
    const size_t step = (size_t)1 << (c->rate_bits - qdb), next_step = (size_t)1 << qdb;
    const size_t lde_size = (size_t)1 << (c->degree_bits + qdb);
    zero_poly_on_coset zh;
    zpoc_new(&zh, c->degree_bits, qdb);
    const uint64_t w = orc_root_of_unity(c->degree_bits + qdb);
    int bad = 0;
#pragma omp parallel
    {
        uint64_t *terms = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(circuit_terms(c) + 1));
        uint64_t *gate_out = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(c->num_gate_constraints + 4));
        uint64_t res[64];
#pragma omp for schedule(static)
        for (size_t i = 0; i < lde_size; i++) {
            const uint64_t x = orc_pow(w, i), shifted_x = orc_mul(7, x);
            const size_t i_next = (i + next_step) % lde_size;
            const uint64_t *local_cs = lde_row(constants_sigmas, i, step);
            const uint64_t *local_wires = lde_row(wires_b, i, step);
            const uint64_t *local_zs_pp = lde_row(zs_b, i, step), *next_zs_pp = lde_row(zs_b, i_next, step);
            const uint64_t l_0_x = zpoc_eval_l_0(&zh, i, shifted_x);
            if (nc > 64 || vanishing_at_point(c, shifted_x, l_0_x, local_cs, local_cs + c->num_constants, local_wires,
                                              local_zs_pp, next_zs_pp, public_inputs_hash, betas, gammas, alphas, terms,
                                              gate_out, res) != 0) {
                bad = 1;
                continue;
            }
            /* quotient_values *= z_h_on_coset.eval_inverse(i) */
            const uint64_t denominator_inv = zh.inverses[i % zh.rate];
            for (int a = 0; a < nc; a++) out[(size_t)a * lde_size + i] = orc_mul(res[a], denominator_inv);
        }
        free(terms);
        free(gate_out);
    }
    zpoc_free(&zh);
    return bad ? -1 : 0;
}

/* ------------------------------------------------------------ ProofWithPublicInputs::to_bytes */
typedef struct {
    uint8_t *buf;
    size_t pos;
} wr;
static void w_u64s(wr *w, const uint64_t *v, size_t n) { /* field elements / hashes: u64 little-endian */
    for (size_t i = 0; i < n; i++) {
        if (w->buf)
            for (int b = 0; b < 8; b++) w->buf[w->pos + b] = (uint8_t)(v[i] >> (8 * b));
        w->pos += 8;
    }
}
static const uint64_t *w_merkle_proof(wr *w, const uint64_t *p, int n_sib) { /* write_merkle_proof */
    if (w->buf) w->buf[w->pos] = (uint8_t)n_sib;
    w->pos += 1;
    w_u64s(w, p, 4 * (size_t)n_sib);
    return p + 4 * (size_t)n_sib;
}
/* write_fri_proof over the flat stream of orc_fri_prove_openings (field order of FriProof) */
static const uint64_t *w_fri_proof(wr *w, const uint64_t *p, const orc_fri_params *fp, const int *widths, int n_oracles) {
    const int log_nb = fp->degree_bits + fp->rate_bits;
    for (int r = 0; r < fp->n_arity; r++) { /* commit_phase_merkle_caps */
        w_u64s(w, p, (size_t)4 << fp->cap_height);
        p += (size_t)4 << fp->cap_height;
    }
    for (int q = 0; q < fp->num_query_rounds; q++) { /* write_fri_query_round */
        for (int o = 0; o < n_oracles; o++) {        /* write_fri_initial_proof */
            w_u64s(w, p, (size_t)widths[o]);
            p = w_merkle_proof(w, p + widths[o], log_nb - fp->cap_height);
        }
        int lg = log_nb;
        for (int r = 0; r < fp->n_arity; r++) { /* write_fri_query_step */
            lg -= fp->arity_bits[r];
            w_u64s(w, p, (size_t)2 << fp->arity_bits[r]);
            p = w_merkle_proof(w, p + ((size_t)2 << fp->arity_bits[r]), lg - fp->cap_height);
        }
    }
    int tot = 0;
    for (int r = 0; r < fp->n_arity; r++) tot += fp->arity_bits[r];
    w_u64s(w, p, ((size_t)2 << (fp->degree_bits - tot)) + 1); /* final_poly.coeffs, pow_witness */
    return p + ((size_t)2 << (fp->degree_bits - tot)) + 1;
}
/* returns the byte count; out == NULL only counts */
size_t orc_write_proof_with_public_inputs(const uint64_t *wires_cap, const uint64_t *zs_cap, const uint64_t *quotient_cap,
                                          const uint64_t *const openings[7], const int n_openings[7],
                                          const uint64_t *fri_words, const orc_fri_params *fp, const int *widths,
                                          int n_oracles, const uint64_t *public_inputs, int n_public_inputs, uint8_t *out) {
    wr w = {out, 0};
    const size_t cap_words = (size_t)4 << fp->cap_height;
    w_u64s(&w, wires_cap, cap_words); /* write_proof */
    w_u64s(&w, zs_cap, cap_words);
    w_u64s(&w, quotient_cap, cap_words);
    for (int k = 0; k < 7; k++) w_u64s(&w, openings[k], 2 * (size_t)n_openings[k]); /* write_opening_set */
    w_fri_proof(&w, fri_words, fp, widths, n_oracles);
    w_u64s(&w, public_inputs, (size_t)n_public_inputs); /* write_field_vec(public_inputs) */
    return w.pos;
}
