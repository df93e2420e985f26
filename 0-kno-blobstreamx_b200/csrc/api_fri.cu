// api_fri.cu -- Challenger and FRI entry points of the C ABI.
// Upstream: plonky2 iop/challenger.rs, fri/oracle.rs `PolynomialBatch::prove_openings`,
// fri/prover.rs `fri_proof`, `fri_committed_trees`, `fri_proof_of_work`, `fri_prover_query_rounds`.
// The whole proof is produced on the device (challenger included, so no beta ever crosses PCIe);
// the host only sequences launches and makes one small read for the proof-of-work result.
#include "api_internal.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

static_assert(sizeof(bsx_challenger) == 28 * 8 + 8, "bsx_challenger layout must match DevChallenger");

struct FriLayout {
    int log_n, rb, log_nb, cap, n_rounds;
    size_t caps_words, per_q, final_len, total;
    size_t off_oracle[64], off_round[16];
    int round_log_leaves[16], round_shift[16];
};

static int fri_layout(bsx_ctx *ctx, bsx_batch *const *oracles, int n_oracles, const bsx_fri_params *p, FriLayout *L) {
    if (!p || n_oracles < 0 || n_oracles > 64 || (n_oracles > 0 && !oracles))
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad oracle list");
    if (p->n_arity < 0 || p->n_arity > 16 || p->degree_bits < 0 || p->rate_bits < 0 || p->cap_height < 0 ||
        p->num_query_rounds < 0 || p->proof_of_work_bits < 0 || p->proof_of_work_bits > 40)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad FRI parameters");
    L->log_n = p->degree_bits;
    L->rb = p->rate_bits;
    L->log_nb = p->degree_bits + p->rate_bits;
    L->cap = p->cap_height;
    L->n_rounds = p->n_arity;
    if (L->log_nb - L->rb > 24 || L->log_nb > 32) return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "degree_bits too large");
    L->caps_words = (size_t)p->n_arity * ((size_t)4 << L->cap);
    size_t per_q = 0;
    for (int o = 0; o < n_oracles; o++) {
        const bsx_batch *b = oracles[o];
        if (!b || b->log_n != L->log_n || b->rate_bits != L->rb || b->tree.cap_height != L->cap)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "oracle %d does not match the FRI parameters", o);
        L->off_oracle[o] = per_q;
        per_q += (size_t)(b->n_cols + b->n_salt) + 4 * (size_t)(L->log_nb - L->cap);
    }
    int lg = L->log_nb, tot = 0;
    for (int r = 0; r < p->n_arity; r++) {
        int ab = p->arity_bits[r];
        if (ab < 1 || ab > 5) return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "arity_bits must be in [1, 5]");
        lg -= ab;
        tot += ab;
        if (lg < L->cap || tot > L->log_n)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "reduction_arity_bits leave fewer leaves than the cap");
        L->off_round[r] = per_q;
        L->round_log_leaves[r] = lg;
        L->round_shift[r] = tot;
        per_q += ((size_t)2 << ab) + 4 * (size_t)(lg - L->cap);
    }
    L->per_q = per_q;
    L->final_len = (size_t)1 << (L->log_n - tot);
    L->total = L->caps_words + per_q * (size_t)p->num_query_rounds + 2 * L->final_len + 1;
    return BSX_OK;
}

extern "C" int32_t bsx_fri_proof_words(bsx_batch *const *oracles, int32_t n_oracles, const bsx_fri_params *params,
                                       size_t *out) {
    if (!out) return BSX_ERR_INVALID_ARG;
    FriLayout L;
    bsx_ctx *ctx = (n_oracles > 0 && oracles && oracles[0]) ? oracles[0]->ctx : nullptr;
    TRY(fri_layout(ctx, oracles, n_oracles, params, &L));
    *out = L.total;
    return BSX_OK;
}

// ------------------------------------------------------------------------------------ challenger
extern "C" void bsx_challenger_init(bsx_challenger *ch) {
    if (ch) memset(ch, 0, sizeof(*ch));
}

static int chal_upload(bsx_ctx *ctx, const bsx_challenger *ch, u64 **dev) {
    if (ch->n_input < 0 || ch->n_input > 7 || ch->n_output < 0 || ch->n_output > 8)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "corrupt challenger buffers");
    TRY(bsx_alloc(ctx, dev, sizeof(bsx_challenger) / 8));
    CU(ctx, cudaMemcpyAsync(*dev, ch, sizeof(bsx_challenger), cudaMemcpyHostToDevice, ctx->stream));
    return BSX_OK;
}
static int chal_download(bsx_ctx *ctx, bsx_challenger *ch, u64 *dev) {
    CU(ctx, cudaMemcpyAsync(ch, dev, sizeof(bsx_challenger), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    bsx_free(ctx, dev);
    return BSX_OK;
}

extern "C" int32_t bsx_challenger_observe(bsx_ctx *ctx, bsx_challenger *ch, const uint64_t *elems_host, size_t n) {
    if (!ctx || !ch || (!elems_host && n)) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (n == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *dev, *data;
    TRY(chal_upload(ctx, ch, &dev));
    TRY(bsx_alloc(ctx, &data, n));
    CU(ctx, cudaMemcpyAsync(data, elems_host, n * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    launch_chal_observe(dev, data, n, 0, 0, ctx->stream);
    bsx_free(ctx, data);
    return chal_download(ctx, ch, dev);
}

extern "C" int32_t bsx_challenger_get(bsx_ctx *ctx, bsx_challenger *ch, uint64_t *out_host, int32_t n) {
    if (!ctx || !ch || n < 0 || (!out_host && n)) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (n == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *dev, *out;
    TRY(chal_upload(ctx, ch, &dev));
    TRY(bsx_alloc(ctx, &out, (size_t)n));
    launch_chal_get(dev, out, n, 0, ctx->stream);
    CU(ctx, cudaMemcpyAsync(out_host, out, (size_t)n * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    bsx_free(ctx, out);
    return chal_download(ctx, ch, dev);
}

// ------------------------------------------------------------------------------------ final polynomial
// fin: planar [2][n], natural order, multiplied by X.  alpha_dev: 2 words on device.
static int final_poly_device(bsx_ctx *ctx, bsx_batch *const *oracles, int n_oracles, const bsx_fri_instance *inst,
                             const u64 *alpha_dev, int log_n, u64 *fin) {
    if (!inst || inst->n_batches < 0 || inst->n_batches > 16) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad instance");
    const size_t n = (size_t)1 << log_n;
    int kmax = 0, total = 0;
    for (int b = 0; b < inst->n_batches; b++) {
        if (inst->n_polys[b] < 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad instance");
        // combine_kernel's column accumulators are exact for sums of up to 2^12 products (gl_acc_reduce)
        if (inst->n_polys[b] > 4096)
            return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "opening batch %d holds %d polynomials; at most 4096 are supported", b,
                            inst->n_polys[b]);
        if (inst->n_polys[b] > kmax) kmax = inst->n_polys[b];
        total += inst->n_polys[b];
    }
    for (int i = 0; i < total; i++) {
        int o = inst->poly_oracle[i];
        if (o == -1) continue;  // held by another rank (sharded prove): keeps its place in the alpha powers
        if (o < 0 || o >= n_oracles || inst->poly_index[i] < 0 || inst->poly_index[i] >= oracles[o]->n_cols)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "instance names polynomial %d of oracle %d, which does not exist",
                            inst->poly_index[i], o);
    }
    CU(ctx, cudaMemsetAsync(fin, 0, 2 * n * sizeof(u64), ctx->stream));
    if (inst->n_batches == 0) return BSX_OK;
    u64 *pw, *pts, *sh, *comp, *tmp, *work, *ptr_dev;
    int *idx_dev;
    TRY(bsx_alloc(ctx, &pw, 2 * (size_t)(kmax > 0 ? kmax : 1)));
    TRY(bsx_alloc(ctx, &pts, 2 * (size_t)inst->n_batches));
    TRY(bsx_alloc(ctx, &sh, 2));
    TRY(bsx_alloc(ctx, &comp, 2 * n));
    TRY(bsx_alloc(ctx, &tmp, 2 * n));
    TRY(bsx_alloc(ctx, &work, divide_by_linear_workspace_words(n)));
    TRY(bsx_alloc(ctx, &ptr_dev, (size_t)(kmax > 0 ? kmax : 1)));
    TRY(bsx_alloc(ctx, (u64 **)&idx_dev, (size_t)(kmax / 2 + 1)));
    CU(ctx, cudaMemcpyAsync(pts, inst->points, 2 * (size_t)inst->n_batches * sizeof(u64), cudaMemcpyHostToDevice,
                            ctx->stream));
    launch_ext_powers(alpha_dev, pw, kmax, ctx->stream);
    int pos = 0;
    std::vector<const u64 *> ptrs;
    std::vector<int> pidx;
    for (int b = 0; b < inst->n_batches; b++) {
        const int k = inst->n_polys[b];
        bool have = false;
        // storage order 1 = bit-reversed (from_values), 0 = natural (from_coeffs); bit-reversed group
        // is combined in storage order (coalesced) and permuted once
        for (int order = 1; order >= 0; order--) {
            ptrs.clear();
            pidx.clear();
            for (int j = 0; j < k; j++) {
                if (inst->poly_oracle[pos + j] < 0) continue;
                const bsx_batch *o = oracles[inst->poly_oracle[pos + j]];
                if ((o->coeffs_bitrev ? 1 : 0) != order) continue;
                ptrs.push_back(o->coeffs + (size_t)inst->poly_index[pos + j] * n);
                pidx.push_back(j);
            }
            if (ptrs.empty()) continue;
            // the pinned-free small uploads are ordered on the stream; sync before reusing the host vectors
            CU(ctx, cudaMemcpyAsync(ptr_dev, ptrs.data(), ptrs.size() * sizeof(u64 *), cudaMemcpyHostToDevice, ctx->stream));
            CU(ctx, cudaMemcpyAsync(idx_dev, pidx.data(), pidx.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
            CU(ctx, cudaStreamSynchronize(ctx->stream));
            if (order == 1) {
                launch_combine((const u64 *const *)ptr_dev, idx_dev, (int)ptrs.size(), pw, n, tmp, tmp + n, 0, ctx->stream);
                launch_bitrev_copy(tmp, n, comp, n, log_n, 2, ctx->stream);
            } else {
                launch_combine((const u64 *const *)ptr_dev, idx_dev, (int)ptrs.size(), pw, n, comp, comp + n, have ? 1 : 0,
                               ctx->stream);
            }
            have = true;
        }
        if (!have) CU(ctx, cudaMemsetAsync(comp, 0, 2 * n * sizeof(u64), ctx->stream));
        launch_ext_pow(alpha_dev, (u64)k, sh, ctx->stream);
        launch_divide_by_linear_accumulate(comp, comp + n, n, pts + 2 * b, sh, fin, fin + n, work, ctx->stream);
        pos += k;
    }
    CU(ctx, cudaGetLastError());
    bsx_free(ctx, pw);
    bsx_free(ctx, pts);
    bsx_free(ctx, sh);
    bsx_free(ctx, comp);
    bsx_free(ctx, tmp);
    bsx_free(ctx, work);
    bsx_free(ctx, ptr_dev);
    bsx_free(ctx, (u64 *)idx_dev);
    return BSX_OK;
}

extern "C" int32_t bsx_fri_final_poly(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                      const bsx_fri_instance *instance, const uint64_t alpha[2], uint64_t *out_host) {
    if (!ctx || !oracles || n_oracles <= 0 || !alpha || !out_host) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    CU(ctx, cudaSetDevice(ctx->device));
    const int log_n = oracles[0]->log_n;
    const size_t n = (size_t)1 << log_n;
    for (int o = 0; o < n_oracles; o++)
        if (oracles[o]->log_n != log_n) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "oracles differ in degree");
    u64 *a, *fin, *inter;
    TRY(bsx_alloc(ctx, &a, 2));
    TRY(bsx_alloc(ctx, &fin, 2 * n));
    TRY(bsx_alloc(ctx, &inter, 2 * n));
    CU(ctx, cudaMemcpyAsync(a, alpha, 2 * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    TRY(final_poly_device(ctx, oracles, n_oracles, instance, a, log_n, fin));
    launch_ext_interleave(fin, n, n, inter, ctx->stream);
    CU(ctx, cudaMemcpyAsync(out_host, inter, 2 * n * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    bsx_free(ctx, a);
    bsx_free(ctx, fin);
    bsx_free(ctx, inter);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// ------------------------------------------------------------------------------------ fri_proof
// values <- coset_fft(coeffs, shift) for a planar extension polynomial of length 2^lg whose
// coefficients beyond 2^(lg - rb) are zero: run as 2^rb coset transforms of size 2^(lg - rb).
static void dbg_chal(bsx_ctx *ctx, const u64 *chal, const char *tag) {
    if (!getenv("BSX_DEBUG_FRI")) return;
    bsx_challenger h;
    cudaMemcpyAsync(&h, chal, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    fprintf(stderr, "[fri] %-14s st0=%016llx st1=%016llx n_in=%d n_out=%d in0=%016llx\n", tag,
            (unsigned long long)h.sponge_state[0], (unsigned long long)h.sponge_state[1], h.n_input, h.n_output,
            (unsigned long long)h.input_buffer[0]);
}

static int ext_coset_ntt(bsx_ctx *ctx, const u64 *coeffs, u64 *values, int lg, int rb, u64 shift) {
    NttPlan plan;
    plan.n_dest = 0;
    plan.log_n = lg - rb;
    bsx_split(plan.log_n, &plan.a, &plan.b);
    plan.inverse = 0;
    plan.src_bitrev = 0;
    plan.n_cosets = 1 << rb;
    TRY(bsx_coset_tables(ctx, plan.log_n, rb, shift, &plan.pre_hi, &plan.pre_lo));
    TRY(bsx_pow_table(ctx, plan.log_n, 0, &plan.tw));
    plan.scale = 1;
    plan.post_full = nullptr;  // two columns only: the two-level twiddle tables are cheaper than a full table
    const size_t len = (size_t)1 << lg;
    launch_ntt(ctx->tables, plan, coeffs, len, values, len, 2, ctx->stream);
    CU(ctx, cudaGetLastError());
    return BSX_OK;
}

// coeffs: planar [2][2^log_nb] natural order on device (consumed).  chal: device challenger.
static int fri_proof_device(bsx_ctx *ctx, bsx_batch *const *oracles, int n_oracles, u64 *coeffs, u64 *chal,
                            const bsx_fri_params *p, const FriLayout &L, u64 *proof, u64 *idx_out_host) {
    const int rb = L.rb, cap = L.cap;
    int lg = L.log_nb;
    size_t len = (size_t)1 << lg;
    u64 shift = 7;
    u64 *values = nullptr, *beta = nullptr;
    TRY(bsx_alloc(ctx, &values, 2 * len));
    TRY(bsx_alloc(ctx, &beta, 2));
    TRY(ext_coset_ntt(ctx, coeffs, values, lg, rb, shift));
    std::vector<bsx_tree *> trees;
    int rc = BSX_OK;
    for (int r = 0; r < L.n_rounds && rc == BSX_OK; r++) {
        const int ab = p->arity_bits[r];
        // fri_committed_trees: values are already bit-reversed (DIF order); leaf = 2^ab consecutive ones
        u64 *leaves = nullptr;
        if ((rc = bsx_alloc(ctx, &leaves, 2 * len))) break;
        launch_fri_leaves(values, values + len, len, ab, leaves, ctx->stream);
        bsx_tree *t = new bsx_tree();
        t->ctx = ctx;
        t->log_leaves = lg - ab;
        t->leaf_len = 2 << ab;
        t->cap_height = cap;
        t->leaves = leaves;
        t->col_stride = len >> ab;
        t->owns_leaves = true;
        t->embedded = false;
        trees.push_back(t);
        if ((rc = bsx_tree_build(t, nullptr))) break;
        const size_t cap_words = (size_t)4 << cap;
        cudaMemcpyAsync(proof + (size_t)r * cap_words, t->cap_dev, cap_words * sizeof(u64), cudaMemcpyDeviceToDevice,
                        ctx->stream);
        launch_chal_observe(chal, t->cap_dev, cap_words, 0, 0, ctx->stream);  // challenger.observe_cap
        launch_chal_get(chal, beta, 2, 0, ctx->stream);                       // get_extension_challenge
        // coeffs <- fold by powers of beta; shift <- shift^arity; values <- coset_fft
        const size_t nl = len >> ab;
        u64 *nc = nullptr;
        if ((rc = bsx_alloc(ctx, &nc, 2 * nl))) break;
        launch_fri_fold(coeffs, coeffs + len, nl, ab, beta, nc, nc + nl, ctx->stream);
        bsx_free(ctx, coeffs);
        coeffs = nc;
        shift = gl_host_pow(shift, (u64)1 << ab);
        len = nl;
        lg -= ab;
        if ((rc = ext_coset_ntt(ctx, coeffs, values, lg, rb, shift))) break;
    }
    u64 *idx = nullptr, *best = nullptr;
    if (rc == BSX_OK) do {
        // coeffs.truncate(len >> rate_bits); challenger.observe_extension_elements(coeffs)
        const size_t final_len = len >> rb;
        dbg_chal(ctx, chal, "before final");
        launch_chal_observe(chal, coeffs, 2 * final_len, 1, len, ctx->stream);
        dbg_chal(ctx, chal, "after final");
        u64 *fin_out = proof + L.caps_words + L.per_q * (size_t)p->num_query_rounds;
        launch_ext_interleave(coeffs, len, final_len, fin_out, ctx->stream);
        // fri_proof_of_work
        if ((rc = bsx_alloc(ctx, &best, 2))) break;
        cudaMemsetAsync(best, 0xFF, sizeof(u64), ctx->stream);
        u64 found = ~0ULL;
        // candidates are tried in growing spans from 0, the smallest valid one of the first successful span wins
        // (= the smallest valid witness overall).  First span = 4 x the expected waiting time 2^pow_bits (98 % hit
        // rate) instead of a fixed 2^20: 0.25 ms instead of 1 ms at 16 bits
        u64 span = (u64)1 << (p->proof_of_work_bits + 2 < 14 ? 14 : p->proof_of_work_bits + 2 > 22 ? 22 : p->proof_of_work_bits + 2);
        for (u64 base = 0; found == ~0ULL; base += span, span = span < ((u64)1 << 22) ? span << 1 : span) {
            launch_pow(chal, base, span, p->proof_of_work_bits, best, ctx->stream);
            if (cudaMemcpyAsync(&found, best, sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
                cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
                rc = bsx_fail(ctx, BSX_ERR_CUDA, "proof-of-work search failed: %s", cudaGetErrorString(cudaGetLastError()));
                break;
            }
            if (base > ((u64)1 << 44)) {
                rc = bsx_fail(ctx, BSX_ERR_CUDA, "Proof of work failed. This is highly unlikely!");
                break;
            }
        }
        if (rc) break;
        cudaMemcpyAsync(fin_out + 2 * final_len, best, sizeof(u64), cudaMemcpyDeviceToDevice, ctx->stream);
        dbg_chal(ctx, chal, "before finish");
        launch_chal_pow_finish(chal, best, best + 1, ctx->stream);
        dbg_chal(ctx, chal, "after finish");
        // fri_prover_query_rounds
        const int nq = p->num_query_rounds;
        if (nq > 0) {
            if ((rc = bsx_alloc(ctx, &idx, (size_t)nq))) break;
            launch_chal_get(chal, idx, nq, (u64)1 << L.log_nb, ctx->stream);
            if (idx_out_host)  // read after the caller's final synchronisation
                cudaMemcpyAsync(idx_out_host, idx, (size_t)nq * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream);
            u64 *qbase = proof + L.caps_words;
            for (int o = 0; o < n_oracles; o++) {
                const bsx_batch *b = oracles[o];
                launch_fri_query_gather(idx, nq, 0, b->lde, (size_t)1 << L.log_nb, b->n_cols + b->n_salt, b->tree.digests,
                                        L.log_nb, L.log_nb - cap, qbase, L.per_q, L.off_oracle[o], ctx->stream);
            }
            for (int r = 0; r < L.n_rounds; r++) {
                const bsx_tree *t = trees[r];
                launch_fri_query_gather(idx, nq, L.round_shift[r], t->leaves, t->col_stride, t->leaf_len, t->digests,
                                        t->log_leaves, t->log_leaves - cap, qbase, L.per_q, L.off_round[r], ctx->stream);
            }
        }
        if (cudaGetLastError() != cudaSuccess) rc = bsx_fail(ctx, BSX_ERR_CUDA, "FRI launch failed");
    } while (0);
    for (bsx_tree *t : trees) {
        bsx_tree_release(t);
        delete t;
    }
    bsx_free(ctx, idx);
    bsx_free(ctx, best);
    bsx_free(ctx, values);
    bsx_free(ctx, beta);
    bsx_free(ctx, coeffs);
    return rc;
}

static int fri_run(bsx_ctx *ctx, bsx_batch *const *oracles, int n_oracles, const bsx_fri_instance *inst,
                   const u64 *lde_coeffs_host, bsx_challenger *ch, const bsx_fri_params *p, u64 *proof_host,
                   const u64 *final_poly = nullptr, int final_mem = BSX_HOST, u64 *idx_out_host = nullptr) {
    if (!ctx || !ch || !p || !proof_host) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    FriLayout L;
    TRY(fri_layout(ctx, oracles, n_oracles, p, &L));
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << L.log_n, nb = (size_t)1 << L.log_nb;
    u64 *chal, *proof, *coeffs;
    TRY(chal_upload(ctx, ch, &chal));
    TRY(bsx_alloc(ctx, &proof, L.total));
    TRY(bsx_alloc(ctx, &coeffs, 2 * nb));
    int rc = BSX_OK;
    if (inst) {
        // prove_openings: alpha, final polynomial, zero-padded to the LDE length
        u64 *alpha, *fin;
        TRY(bsx_alloc(ctx, &alpha, 2));
        TRY(bsx_alloc(ctx, &fin, 2 * n));
        launch_chal_get(chal, alpha, 2, 0, ctx->stream);
        rc = final_poly_device(ctx, oracles, n_oracles, inst, alpha, L.log_n, fin);
        if (rc == BSX_OK) {
            cudaMemsetAsync(coeffs, 0, 2 * nb * sizeof(u64), ctx->stream);
            cudaMemcpyAsync(coeffs, fin, n * sizeof(u64), cudaMemcpyDeviceToDevice, ctx->stream);
            cudaMemcpyAsync(coeffs + nb, fin + n, n * sizeof(u64), cudaMemcpyDeviceToDevice, ctx->stream);
        }
        bsx_free(ctx, alpha);
        bsx_free(ctx, fin);
    } else if (final_poly) {
        // the final polynomial itself (2^degree_bits interleaved extension coefficients), zero-padded here
        u64 *tmp;
        TRY(bsx_alloc(ctx, &tmp, 2 * n));
        TRY(bsx_copy_in(ctx, tmp, final_poly, 2 * n, final_mem));
        CU(ctx, cudaMemsetAsync(coeffs, 0, 2 * nb * sizeof(u64), ctx->stream));
        launch_ext_deinterleave(tmp, n, coeffs, nb, ctx->stream);
        bsx_free(ctx, tmp);
    } else {
        u64 *tmp;
        TRY(bsx_alloc(ctx, &tmp, 2 * nb));
        CU(ctx, cudaMemcpyAsync(tmp, lde_coeffs_host, 2 * nb * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
        launch_ext_deinterleave(tmp, nb, coeffs, nb, ctx->stream);
        bsx_free(ctx, tmp);
    }
    if (rc == BSX_OK) rc = fri_proof_device(ctx, oracles, n_oracles, coeffs, chal, p, L, proof, idx_out_host);
    else bsx_free(ctx, coeffs);
    if (rc == BSX_OK) {
        cudaError_t e = cudaMemcpyAsync(proof_host, proof, L.total * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream);
        if (e != cudaSuccess) rc = bsx_fail(ctx, BSX_ERR_CUDA, "proof copy failed: %s", cudaGetErrorString(e));
    }
    bsx_free(ctx, proof);
    int rc2 = chal_download(ctx, ch, chal);
    return rc != BSX_OK ? rc : rc2;
}

extern "C" int32_t bsx_fri_prove_openings(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                          const bsx_fri_instance *instance, bsx_challenger *challenger,
                                          const bsx_fri_params *params, uint64_t *proof_host) {
    if (!instance) return BSX_ERR_INVALID_ARG;
    return fri_run(ctx, oracles, n_oracles, instance, nullptr, challenger, params, (u64 *)proof_host);
}

extern "C" int32_t bsx_fri_proof(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                 const uint64_t *lde_coeffs_host, bsx_challenger *challenger,
                                 const bsx_fri_params *params, uint64_t *proof_host) {
    if (!lde_coeffs_host) return BSX_ERR_INVALID_ARG;
    return fri_run(ctx, oracles, n_oracles, nullptr, (const u64 *)lde_coeffs_host, challenger, params, (u64 *)proof_host);
}

// fri_proof from the final polynomial of prove_openings (2^degree_bits extension coefficients, `mem` side) instead of
// its zero-padded LDE form, with the query indices returned: the building block of a prove over row-sharded
// oracles, where the initial-tree openings come from the owning ranks (n_oracles may be 0).
extern "C" int32_t bsx_fri_proof_from_final_poly(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                                 const uint64_t *final_poly, int32_t mem, bsx_challenger *challenger,
                                                 const bsx_fri_params *params, uint64_t *proof_host,
                                                 uint64_t *query_indices_host) {
    if (!final_poly) return BSX_ERR_INVALID_ARG;
    return fri_run(ctx, oracles, n_oracles, nullptr, nullptr, challenger, params, (u64 *)proof_host,
                   (const u64 *)final_poly, mem, (u64 *)query_indices_host);
}

// out[i] = sum_k parts[k * words + i] (mod p): the partial final polynomials of the ranks added up
extern "C" int32_t bsx_field_sum(bsx_ctx *ctx, const uint64_t *parts, int32_t mem, int32_t k, size_t words, uint64_t *out,
                                 int32_t mem_out) {
    if (!ctx || k < 0 || (words && k && (!parts || !out))) return BSX_ERR_INVALID_ARG;
    if (words == 0) return BSX_OK;
    BsxScope scope(ctx);
    CU(ctx, cudaSetDevice(ctx->device));
    const u64 *src = (const u64 *)parts;
    u64 *tmp = nullptr, *dst = (u64 *)out;
    if (mem != BSX_DEVICE && k > 0) {
        TRY(bsx_alloc(ctx, &tmp, (size_t)k * words));
        TRY(bsx_copy_in(ctx, tmp, (const u64 *)parts, (size_t)k * words, mem));
        src = tmp;
    }
    if (mem_out != BSX_DEVICE) TRY(bsx_alloc(ctx, &dst, words));
    launch_field_sum(src, k, words, dst, ctx->stream);
    CU(ctx, cudaGetLastError());
    if (mem_out != BSX_DEVICE) TRY(bsx_copy_out(ctx, (u64 *)out, dst, words, mem_out));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// The part of prove_openings' final polynomial that the listed polynomials contribute: `instance` names every
// polynomial of the opening batches, with poly_oracle == -1 for those not in `oracles`; they keep their place in
// the powers of alpha.  Division by (X - z), the shift between batches and the multiplication by X are linear, so the
// sum of the parts over any partition of the polynomials is the final polynomial.  out = 2^degree_bits x [c0, c1].
extern "C" int32_t bsx_fri_final_poly_part(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                           const bsx_fri_instance *instance, const uint64_t alpha[2], int32_t degree_bits,
                                           uint64_t *out, int32_t mem_out) {
    if (!ctx || !instance || !alpha || !out || degree_bits < 0 || degree_bits > 24 || n_oracles < 0 ||
        (n_oracles > 0 && !oracles))
        return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << degree_bits;
    for (int o = 0; o < n_oracles; o++)
        if (!oracles[o] || oracles[o]->log_n != degree_bits) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "oracles differ in degree");
    u64 *a, *fin, *inter = (u64 *)out;
    TRY(bsx_alloc(ctx, &a, 2));
    TRY(bsx_alloc(ctx, &fin, 2 * n));
    if (mem_out != BSX_DEVICE) TRY(bsx_alloc(ctx, &inter, 2 * n));
    CU(ctx, cudaMemcpyAsync(a, alpha, 2 * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    TRY(final_poly_device(ctx, oracles, n_oracles, instance, a, degree_bits, fin));
    launch_ext_interleave(fin, n, n, inter, ctx->stream);
    if (mem_out != BSX_DEVICE) TRY(bsx_copy_out(ctx, (u64 *)out, inter, 2 * n, mem_out));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// ------------------------------------------------------------------------------------ opening set
extern "C" int32_t bsx_batch_eval_ext(bsx_batch *b, int32_t col_begin, int32_t n_cols, const uint64_t *points_host,
                                      int32_t n_points, uint64_t *out_host) {
    if (!b) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = b->ctx;
    BsxScope scope(ctx);
    if (col_begin < 0 || n_cols < 0 || col_begin + n_cols > b->n_cols)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "polynomial range %d..%d outside the batch (%d polynomials)", col_begin,
                        col_begin + n_cols, b->n_cols);
    if (n_points < 0 || (n_points > 0 && !points_host) || (n_points > 0 && n_cols > 0 && !out_host))
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "null points / output");
    if (n_points == 0 || n_cols == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << b->log_n, n_out = 2 * (size_t)n_points * n_cols;
    u64 *pts, *work, *out;
    TRY(bsx_alloc(ctx, &pts, 2 * (size_t)n_points));
    TRY(bsx_alloc(ctx, &work, eval_workspace_words(b->log_n, n_cols, n_points)));
    TRY(bsx_alloc(ctx, &out, n_out));
    TRY(bsx_copy_in(ctx, pts, (const u64 *)points_host, 2 * (size_t)n_points, BSX_HOST));
    launch_eval_ext(b->coeffs + (size_t)col_begin * n, n, b->log_n, b->coeffs_bitrev, n_cols, pts, n_points, work, out,
                    ctx->stream);
    CU(ctx, cudaGetLastError());
    TRY(bsx_copy_out(ctx, (u64 *)out_host, out, n_out, BSX_HOST));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    bsx_free(ctx, pts);
    bsx_free(ctx, work);
    bsx_free(ctx, out);
    return BSX_OK;
}
